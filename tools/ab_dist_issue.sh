# A/B runs of the data-parallel training leg at N = 2 (under gpurun --gpus 2)
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-rooflines --no-cpu-baseline --no-sweep --no-varying"
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["training"]["ms_per_step"], d["training"]["ms_per_step_rank0_min_median_max"], d["training"]["slow_steps_rank0"])'
for i in 1 2 3; do
echo -n "20 "; $T 2>/dev/null | python -c "$P"
echo -n "60 "; $T --train-steps 60 2>/dev/null | python -c "$P"
done
