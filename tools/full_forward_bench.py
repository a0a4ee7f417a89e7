"""InteractionHead.forward (detections + FPN pyramid resident on the GPU -> preprocess -> RoIAlign -> pair path) on the
C2 shapes, native RoIAlign (bf16 operand rows) against the stock torchvision pooler, plus the two poolers
alone.      python tools/full_forward_bench.py            (on the GPU box)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import synthetic                         # noqa: E402
from skghoi_b200.head import build_interaction_head      # noqa: E402
from skghoi_b200.roi import MultiScaleRoIAlign           # noqa: E402

dev = torch.device("cuda:0")
ih = build_interaction_head(synthetic.hico_object_to_verb(), seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).eval()
ih.box_pair_head.transh_head.transh.to(dev)
batch = synthetic.make_batch([(16, 32)] * 8, seed=42, levels=("0", "1", "2", "3"))
feats = {k: v.to(dev) for k, v in batch["features"].items()}
dets = [dict(boxes=b.to(dev), labels=l.to(dev), scores=s.to(dev))
        for b, l, s in zip(batch["box_coords"], batch["box_labels"], batch["box_scores"])]
shapes = batch["image_shapes"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / n


def full():
    with torch.no_grad():
        return ih(feats, dets, shapes)


pairs = 0
print("| path | ms/step | note |")
print("|---|---:|---|")
for native in (True, False):
    ih.native_roi_align = native
    res = full()
    pairs = sum(int(r["boxes_h"].shape[0]) for r in res)
    ms = timed(full)
    print("| InteractionHead.forward, %s RoIAlign | %.3f | %d pairs kept after NMS/caps: %.2f M pairs/s |"
          % ("native (bf16 operand rows)" if native else "stock torchvision", ms, pairs, pairs / ms / 1e3))
coords = [d["boxes"] for d in dets]
mine = MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
rows = torch.empty(sum(c.shape[0] for c in coords), 12544, dtype=torch.bfloat16, device=dev)
print("| pooler alone: native, fp32 [256,256,7,7] | %.3f | 256 boxes, 4 levels |" % timed(lambda: mine(feats, coords, shapes)))
print("| pooler alone: native, bf16 rows | %.3f | |" % timed(lambda: mine.pool(feats, coords, shapes, want_f32=False, bf16_out=rows)))
print("| pooler alone: torchvision CUDA | %.3f | |" % timed(lambda: ih.box_roi_pool(feats, coords, shapes)))
