"""Data-parallel training step check (run under torchrun on >= 2 GPUs):  the gradients the overlapped, staged all-reduce of
train._GradAllReduce leaves in p.grad equal the rank mean of the single-rank gradients.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_train_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from skghoi_b200 import synthetic, train
from skghoi_b200.head import build_interaction_head

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
otv = synthetic.hico_object_to_verb()


def grads(distributed: bool):
    ih = build_interaction_head(otv, seed=42, distributed=False)     # local loss normalisers in both runs: exact comparison
    ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
    ih = ih.to(dev).train(); ih.box_pair_head.transh_head.transh.to(dev)
    ih.box_pair_head.transh_head.negatives_on_device = True
    batch = synthetic.make_batch([(5, 12), (3, 7), (8, 9)], seed=100 + rank, with_targets=True)
    d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev), boxes=[b.to(dev) for b in batch["box_coords"]],
             labels=[l.to(dev) for l in batch["box_labels"]], scores=[s.to(dev) for s in batch["box_scores"]],
             targets=[{k: v.to(dev) for k, v in t.items()} for t in batch["targets"]])
    if distributed:
        train.enable_gradient_allreduce(ih)
    out = []
    for _ in range(2):                                   # twice: the second backward re-uses events / tables
        ih.zero_grad(set_to_none=False) if out else None
        res = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"], d["targets"])
        ls = res[-1]
        # un-normalised sum so that the distributed normaliser averaging (head.py:167-228) does not enter the comparison
        (ls["hoi_loss"] + ls["interactiveness_loss"]).backward()
        torch.cuda.synchronize()
        out = [p.grad.detach().clone() for p in ih.parameters()]
    return out


g_dist = grads(True)
g_single = grads(False)
worst = 0.0
names = [k for k, _ in build_interaction_head(otv, seed=42).named_parameters()]
bad = []
for k, a, b in zip(names, g_dist, g_single):
    m = b.clone()
    dist.all_reduce(m)
    m /= world
    den = float(m.norm()) + 1e-7            # adjacency.bias has a mathematically zero gradient (softmax shift invariance): noise / noise
    e = float((a - m).norm()) / den
    if e > 2e-2:
        bad.append((k, e, float(a.norm()), den))
    worst = max(worst, e)
if rank == 0:
    print("tensors off: %d of %d" % (len(bad), len(names)))
    for x in bad[:12]:
        print("   %-60s rel %.3e |got| %.3e |want| %.3e" % x)
t = torch.tensor([worst], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print("dist_train_check: world %d, worst per-tensor relative L2 between staged all-reduce and the rank mean: %.3e" % (world, float(t)))
    # the two runs differ by the split-K summation order only
    print("OK" if float(t) < 2e-2 else "FAIL")
dist.destroy_process_group()
