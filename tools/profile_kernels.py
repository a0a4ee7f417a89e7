"""Launches each roofline-relevant kernel a few times inside a cudaProfilerStart/Stop window, for
    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/prof python tools/profile_kernels.py
Kernels: the fused fc_2 GEMM (dominant), the box_head GEMM, pairs_spatial at 2^20 cells, TransH grid
forward in score-only and API-faithful modes.  Not a benchmark: numbers printed under ncu are never reported."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from skghoi_b200 import _lib, engine, synthetic            # noqa: E402
from skghoi_b200.head import build_interaction_head        # noqa: E402

which = set(sys.argv[1:]) or {"gemm", "spatial", "transh"}
dev = torch.device("cuda:0")
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).eval()
ih.box_pair_head.transh_head.transh.to(dev)
batch = synthetic.make_batch([(16, 32)] * 8, seed=42)
with torch.no_grad():
    for _ in range(2):
        ih.forward_from_pooled({"3": batch["features"]["3"].to(dev)}, batch["image_shapes"], batch["box_features"].to(dev),
                               [b.to(dev) for b in batch["box_coords"]], [l.to(dev) for l in batch["box_labels"]],
                               [s.to(dev) for s in batch["box_scores"]])
run = ih.__dict__["_skg_last_run"]
g = engine.dominant_gemm_closure(ih, run)
p = _lib.ptr
B = 16384        # 2^23 cells: the shape of bench.py's spatial roofline leg
gen = torch.Generator().manual_seed(1)
bx, _, _ = synthetic.make_detections([(16, 32)], [(800, 1216)], gen)
boxes = bx[0].repeat(B, 1).to(dev)
seg = engine.Segments([16] * B, [32] * B, [(800, 1216)] * B, dev)
C, Pn = seg.total_cells, seg.total_pairs
enc = torch.empty(C, 46, dtype=torch.float32, device=dev)
xk = torch.empty(Pn, dtype=torch.int64, device=dev); yk = torch.empty(Pn, dtype=torch.int64, device=dev)
nanf = torch.zeros(B, dtype=torch.int32, device=dev)
tr = ih.box_pair_head.transh_head.transh
e, r, w = tr.ent_embeddings.weight.detach(), tr.rel_embeddings.weight.detach(), tr.norm_vector.weight.detach()
cells = 1 << 21
tails = (torch.arange(cells, device=dev) % 32).to(torch.int64)
score = torch.empty(cells * 117, dtype=torch.float32, device=dev)
cells2 = 16384
outs = [torch.empty(cells2 * 117, 50, dtype=torch.float32, device=dev) for _ in range(4)]
st = _lib.stream_ptr(dev)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(2):
    if "gemm" in which:
        g["launch"]()
    if "spatial" in which:
        _lib.call("skg_pairs_spatial_encode", p(boxes), p(seg.d_box_off), p(seg.d_n_h), p(seg.d_cell_off), p(seg.d_pair_off),
                  p(seg.d_hum_off), p(seg.d_hw), B, C, 23, 1e-10, 0, p(enc), None, None, p(xk), p(yk), None, None, None,
                  None, None, p(nanf), st)
    if "transh" in which:
        _lib.call("skg_transh_forward_grid", p(e), p(r), p(w), 80, 117, 50, 49, p(tails), cells, 2, 1, None, None, None, None,
                  p(score), st)
        _lib.call("skg_transh_forward_grid", p(e), p(r), p(w), 80, 117, 50, 49, p(tails), cells2, 2, 1, p(outs[0]), p(outs[1]),
                  p(outs[2]), p(outs[3]), p(score), st)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled kernels launched")
