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


def train_run(sync_mode: bool):
    """Six optimiser steps alternating between two batches of the SAME geometry (so the forward / backward graphs replay)
    with DIFFERENT contents: a stage event that fired early (stale record) would all-reduce half-written gradients."""
    from skghoi_b200 import shard
    if sync_mode:
        os.environ["SKG_DIST_SYNC"] = "1"
    else:
        os.environ.pop("SKG_DIST_SYNC", None)
    ih = build_interaction_head(otv, seed=42, distributed=True)
    ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
    ih = ih.to(dev).train(); ih.box_pair_head.transh_head.transh.to(dev)
    ih.box_pair_head.transh_head.negatives_on_device = True
    params = [p for p in ih.parameters() if p.requires_grad]
    buckets = shard.GradBuckets(params).attach(flatten_params=True, module=ih)
    opt = torch.optim.SGD(buckets.flat_params, lr=1e-2, momentum=0.9, fused=True)
    train.enable_gradient_allreduce(ih)
    data = []
    for k in range(2):
        batch = synthetic.make_batch([(6, 12), (4, 9), (8, 10)], seed=500 + 10 * k + rank, with_targets=True)
        n_per = [int(b.shape[0]) for b in batch["box_coords"]]
        data.append((batch, dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev),
                                 boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per)),
                                 labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per)),
                                 scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per)),
                                 targets=[{k2: v.to(dev) for k2, v in t.items()} for t in batch["targets"]])))
    # persistent input buffers (stable addresses -> the graphs replay), refilled with the step's batch
    buf = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in data[0][1].items() if k in ("bf",)}
    f3 = data[0][1]["f"]["3"].clone()
    bx = torch.cat(data[0][1]["boxes"]).clone(); lb = torch.cat(data[0][1]["labels"]).clone(); sc = torch.cat(data[0][1]["scores"]).clone()
    n_per = [int(b.shape[0]) for b in data[0][0]["box_coords"]]
    losses = []
    for step in range(8):
        batch, d = data[step % 2]
        f3.copy_(d["f"]["3"]); buf["bf"].copy_(d["bf"]); bx.copy_(torch.cat(d["boxes"])); lb.copy_(torch.cat(d["labels"])); sc.copy_(torch.cat(d["scores"]))
        buckets.zero_grad()
        res = ih.forward_from_pooled({"3": f3}, batch["image_shapes"], buf["bf"], list(bx.split(n_per)), list(lb.split(n_per)),
                                     list(sc.split(n_per)), d["targets"])
        ls = res[-1]
        (ls["hoi_loss"] + ls["interactiveness_loss"]).backward()
        opt.step()
        losses.append(float((ls["hoi_loss"] + ls["interactiveness_loss"]).detach()))
    torch.cuda.synchronize()
    import ctypes
    from skghoi_b200 import _lib
    return torch.cat([p.detach().flatten() for p in ih.parameters()]).clone(), losses


w_overlap, l_overlap = train_run(False)
w_sync, l_sync = train_run(True)
os.environ.pop("SKG_DIST_SYNC", None)
rel = float((w_overlap - w_sync).norm() / w_sync.norm())
tr = torch.tensor([rel], device=dev)
dist.all_reduce(tr, op=dist.ReduceOp.MAX)
if rank == 0:
    print("8 optimiser steps, alternating batch contents under graph replay: overlapped vs synchronised all-reduce, "
          "relative L2 of the final weights %.3e; losses %s vs %s" % (float(tr), ["%.4f" % x for x in l_overlap], ["%.4f" % x for x in l_sync]))
    print("TRAIN_OK" if float(tr) < 1e-3 else "TRAIN_FAIL")

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
