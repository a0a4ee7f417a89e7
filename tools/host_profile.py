"""Host cost of one inference step, split: python tools/host_profile.py  (on the GPU box)."""
import os, sys, time, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from skghoi_b200 import _lib, engine, synthetic
from skghoi_b200.head import build_interaction_head

dev = torch.device("cuda:0")
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).eval(); ih.box_pair_head.transh_head.transh.to(dev)
shapes = [(int(os.environ.get("NH", 16)), int(os.environ.get("N", 32)))] * 8
batch = synthetic.make_batch(shapes, seed=42)
n_per = [int(b.shape[0]) for b in batch["box_coords"]]
d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev),
         boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per)), labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per)),
         scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per)))
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
l = _lib.lib()
T = {}
for name in ("skg_session_prepare", "skg_session_run"):
    fn = getattr(l, name)
    def wrap(*a, _fn=fn, _n=name):
        t0 = time.perf_counter_ns(); r = _fn(*a); T[_n] = T.get(_n, 0) + time.perf_counter_ns() - t0; return r
    setattr(l, name, wrap)
orig_sf = engine._session_forward
def sf(*a, **k):
    t0 = time.perf_counter_ns(); r = orig_sf(*a, **k); T["_session_forward"] = T.get("_session_forward", 0) + time.perf_counter_ns() - t0; return r
engine._session_forward = sf
orig_pi = engine._prep_inputs
def pi(*a, **k):
    t0 = time.perf_counter_ns(); r = orig_pi(*a, **k); T["_prep_inputs"] = T.get("_prep_inputs", 0) + time.perf_counter_ns() - t0; return r
engine._prep_inputs = pi

def step(fresh_meta):
    m = dict(n=meta["n"], n_h=meta["n_h"], labels=meta["labels"], scores=meta["scores"]) if fresh_meta else meta
    with torch.no_grad():
        return engine.interaction_head_from_pooled(ih, d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"], meta=m)

for fresh in (False, True):
    for _ in range(20):
        step(fresh)
    torch.cuda.synchronize(); T.clear()
    N = 200
    t0 = time.perf_counter_ns()
    for _ in range(N):
        step(fresh)
    tot = time.perf_counter_ns() - t0
    torch.cuda.synchronize()
    print("fresh meta per step: %s | whole step host %.1f us | %s" % (fresh, tot / N / 1e3, {k: "%.1f us" % (v / N / 1e3) for k, v in T.items()}))
    # GPU-side duration of the same loop (events), to see whether the host or the GPU is the bound
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N):
        step(fresh)
    e1.record(); torch.cuda.synchronize()
    print("   events: %.1f us/step" % (e0.elapsed_time(e1) / N * 1e3))
# raw graph launch cost: replay the session step's C call only
a, b = ctypes.c_longlong(), ctypes.c_longlong()
l.skg_tensor_map_cache_stats(ctypes.byref(a), ctypes.byref(b))
print("tensor map cache (hits, encodes):", a.value, b.value, " graph stats:", _lib.graph_stats())
