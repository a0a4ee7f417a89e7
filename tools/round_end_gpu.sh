# Round-end evidence on one B200 (under gpurun): GPU tests, smoke, both bench arms, the ncu launch list of the bench
# command and one ncu --set full capture of the roofline kernels.  TAG names the outputs (gpurun_out/${TAG}_*).
TAG=${1:-r02}
set -x
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/${TAG}_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.txt 2>&1
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_ref.err
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-rooflines --no-train --no-varying --no-sweep > gpurun_out/${TAG}_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${TAG}_prof python tools/profile_kernels.py > gpurun_out/${TAG}_prof.log 2>&1
tail -3 gpurun_out/${TAG}_pytest.txt; cat gpurun_out/${TAG}_smoke.txt | tail -2; head -c 600 gpurun_out/${TAG}_bench.json
