"""Host-vs-device breakdown of one bench step (C2): enqueue time, device time, cProfile of the host side.
    python tools/step_profile.py            (on the GPU box)"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import _lib, engine, synthetic            # noqa: E402
from skghoi_b200.head import build_interaction_head        # noqa: E402

dev = torch.device("cuda:0")
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).eval()
ih.box_pair_head.transh_head.transh.to(dev)
batch = synthetic.make_batch([(16, 32)] * 8, seed=42)
d = dict(f3=batch["features"]["3"].to(dev), bf=batch["box_features"].to(dev), boxes=[b.to(dev) for b in batch["box_coords"]],
         labels=[l.to(dev) for l in batch["box_labels"]], scores=[s.to(dev) for s in batch["box_scores"]])
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)


def step():
    with torch.no_grad():
        return engine.interaction_head_from_pooled(ih, {"3": d["f3"]}, batch["image_shapes"], d["bf"], d["boxes"],
                                                   d["labels"], d["scores"], meta=meta)


def core():
    boxes_cat = torch.cat(d["boxes"])
    with torch.no_grad():
        return engine.head_forward(ih, d["f3"], batch["image_shapes"], d["bf"], boxes_cat, meta)


for fn, name in ((step, "interaction_head_from_pooled"), (core, "head_forward only")):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    N = 50
    host = dev_ms = 0.0
    for _ in range(N):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter(); s.record(); fn(); e.record(); host += time.perf_counter() - t0
        torch.cuda.synchronize(); dev_ms += s.elapsed_time(e)
    # back-to-back (GPU never idle if host is faster)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(N):
        fn()
    torch.cuda.synchronize(); b2b = (time.perf_counter() - t0) / N
    print("%-32s host enqueue %.3f ms | events (idle start) %.3f ms | back-to-back %.3f ms/step" % (name, host / N * 1e3, dev_ms / N, b2b * 1e3))

print("graph cache (replays, captures, plain):", _lib.graph_stats())
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(45)

# per-launch device timeline of skg_head_forward inside the real pipeline (events after each launch; warm caches)
boxes_cat = torch.cat(d["boxes"])
acc = {}
REP = 20
for _ in range(REP):
    with torch.no_grad():
        run = engine.head_forward(ih, d["f3"], batch["image_shapes"], d["bf"], boxes_cat, meta, profile=True)
    for i, (nm, sid, us) in enumerate(run["profile"]):
        acc.setdefault((i, nm, sid), []).append(us)
tot = [0.0] * 4
print("\nper-launch timeline (median of %d, us since previous event on the same stream; stream 0 = caller, 1 = side)" % REP)
for (i, nm, sid), v in sorted(acc.items()):
    v.sort(); m = v[len(v) // 2]; tot[sid] += m
    print("  s%d %-32s %8.2f" % (sid, nm, m))
print("  sums per stream (us):", ["%.1f" % t for t in tot])
