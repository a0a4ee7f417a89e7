set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r01f_pytest.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r01f_smoke.txt 2>&1
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r01f_bench_reference.json 2> gpurun_out/r01f_ref.err
python bench.py > gpurun_out/r01f_bench.json 2> gpurun_out/r01f_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01f_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-rooflines --no-train > gpurun_out/r01f_ncu.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/r01f_prof python tools/profile_kernels.py > gpurun_out/r01f_prof.log 2>&1
tail -3 gpurun_out/r01f_pytest.txt; cat gpurun_out/r01f_smoke.txt | tail -2; cat gpurun_out/r01f_bench.json
