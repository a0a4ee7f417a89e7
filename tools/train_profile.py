"""Where a training step spends its time (torch.profiler): python tools/train_profile.py  (on the GPU box)."""
import os, sys, time
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
d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev), boxes=[b.to(dev) for b in batch["box_coords"]],
         labels=[l.to(dev) for l in batch["box_labels"]], scores=[s.to(dev) for s in batch["box_scores"]],
         targets=[{k: v.to(dev) for k, v in t.items()} for t in batch["targets"]])
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
params = [p for p in ih.parameters() if p.requires_grad]
opt = torch.optim.SGD(params, lr=1e-4, momentum=0.9)
T = {}


def step(timed=False):
    def mark(k, t0):
        if timed:
            torch.cuda.synchronize(); T[k] = T.get(k, 0.0) + time.perf_counter() - t0
        return time.perf_counter()
    t0 = time.perf_counter()
    opt.zero_grad(set_to_none=True)
    res = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"], d["targets"], meta=meta)
    t0 = mark("forward (native + taped)", t0)
    ls = res[-1]
    (ls["hoi_loss"] + ls["interactiveness_loss"] + ls["transH_loss"].sum()).backward()
    t0 = mark("backward", t0)
    opt.step()
    mark("optimizer", t0)


for _ in range(3):
    step()
torch.cuda.synchronize()
for _ in range(5):
    step(True)
print({k: "%.2f ms" % (v / 5 * 1e3) for k, v in T.items()})
# split the forward
import skghoi_b200.train as TR
orig = engine.interaction_head_from_pooled
acc = {"native": 0.0}
def wrapped(*a, **k):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = orig(*a, **k); torch.cuda.synchronize(); acc["native"] += time.perf_counter() - t0; return r
engine.interaction_head_from_pooled = wrapped
for _ in range(5):
    step()
print("native train-mode forward inside the step: %.2f ms" % (acc["native"] / 5 * 1e3))
engine.interaction_head_from_pooled = orig
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=18, max_name_column_width=50))
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=14, max_name_column_width=70))
