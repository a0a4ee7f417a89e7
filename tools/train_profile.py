"""Where a training step (bench.py's `training` leg: batch 16 x (16, 32), GradBuckets, SGD momentum) spends its time:
host enqueue per phase, device->host synchronisations (torch sync debug mode), GPU time per kernel (torch.profiler).
python tools/train_profile.py  (on the GPU box)."""
import os, sys, time, warnings
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import engine, synthetic, shard          # noqa: E402
from skghoi_b200.head import build_interaction_head       # noqa: E402

dev = torch.device("cuda:0")
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).train(); ih.box_pair_head.transh_head.transh.to(dev)
ih.box_pair_head.transh_head.negatives_on_device = True
batch = synthetic.make_batch([(16, 32)] * 16, seed=1042, with_targets=True)
n_per = [int(b.shape[0]) for b in batch["box_coords"]]
d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev), boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per)),
         labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per)), scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per)),
         targets=[{k: v.to(dev) for k, v in t.items()} for t in batch["targets"]])
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
params = [p for p in ih.parameters() if p.requires_grad]
buckets = shard.GradBuckets(params).attach(flatten_params=True, module=ih)
opt = torch.optim.SGD(buckets.flat_params, lr=1e-4, momentum=0.9, fused=True)
T = {}


def step(sync=False):
    def mark(k, t0):
        if sync:
            torch.cuda.synchronize()
        T[k] = T.get(k, 0.0) + time.perf_counter() - t0
        return time.perf_counter()
    t0 = time.perf_counter()
    buckets.zero_grad()
    t0 = mark("zero_grad", t0)
    res = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"], d["targets"], meta=meta)
    t0 = mark("forward + losses", t0)
    ls = res[-1]
    (ls["hoi_loss"] + ls["interactiveness_loss"] + ls["transH_loss"].sum()).backward()
    t0 = mark("backward", t0)
    opt.step()
    mark("optimizer", t0)


for _ in range(3):
    step()
torch.cuda.synchronize()
for sync in (False, True):
    T.clear()
    t0 = time.perf_counter()
    for _ in range(10):
        step(sync)
    torch.cuda.synchronize()
    print("%s: %.2f ms/step | " % ("phases with a synchronise after each" if sync else "host enqueue per phase (no synchronise)",
                                     (time.perf_counter() - t0) / 10 * 1e3), {k: "%.2f ms" % (v / 10 * 1e3) for k, v in T.items()})
import ctypes
from skghoi_b200 import _lib
a, b_, c = ctypes.c_longlong(), ctypes.c_longlong(), ctypes.c_longlong()
_lib.lib().skg_head_backward_graph_stats(ctypes.byref(a), ctypes.byref(b_), ctypes.byref(c))
print("backward graph cache (replays, captures, plain):", a.value, b_.value, c.value, "| forward:", _lib.graph_stats())
print("---- device->host synchronisations of ONE step:")
torch.cuda.set_sync_debug_mode("warn")
with warnings.catch_warnings(record=True) as w:
    warnings.simplefilter("always")
    step()
torch.cuda.set_sync_debug_mode("default")
import traceback
for x in w:
    print("  sync:", x.filename.replace(ROOT, "."), x.lineno)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=90))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(10):
    step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr, stream=sys.stdout).sort_stats("cumulative").print_stats(60)
