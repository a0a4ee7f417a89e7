"""Logit-error statistics of the bf16 pipeline against the fp32 oracle for several configurations (diagnostics; used to
calibrate helpers.assert_logits_parity):  python tools/parity_stats.py  (on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from helpers import build_head, head_params, synthetic
from oracle import head_oracle as O
from skghoi_b200 import engine
from skghoi_b200.head import build_interaction_head

DEV = "cuda:0"


def stats(name, got, ref):
    got = got.double().cpu(); ref = ref.double()
    d = (got - ref).abs()
    rms = ref.pow(2).mean().sqrt()
    rowmax = ref.abs().amax(1, keepdim=True)
    band_b = float((d / (ref.abs() + rms)).max())
    band_r = float((d / (ref.abs() + torch.maximum(rowmax, rms))).max())
    viol_b = float((d > 1e-2 * ref.abs() + 1e-2 * rms).double().mean())
    viol_r = float((d > 1e-2 * ref.abs() + 1e-2 * torch.maximum(rowmax, rms)).double().mean())
    print("%-28s n=%7d max_rel %.2e rel_l2 %.2e | batch-rms band: viol %.2e worst %.2e | row band: viol %.2e worst %.2e | max|ref|/rms %.1f"
          % (name, ref.numel(), float(d.max() / ref.abs().max()), float((got - ref).norm() / ref.norm()), viol_b, band_b, viol_r, band_r,
             float(ref.abs().max() / rms)))


def run(name, shapes, seed, ih, P, tables, otv, nc, hidx):
    batch = synthetic.make_batch(shapes, seed=seed, human_idx=hidx)
    with torch.no_grad():
        res, lp, ls = O.interaction_head_forward(P, tables, batch["features"]["3"], batch["image_shapes"], batch["box_features"],
                                                 batch["box_coords"], batch["box_labels"], batch["box_scores"], otv, num_cls=nc,
                                                 human_idx=hidx, row_pick="slice")
        meta = engine._host_meta(batch["box_labels"], batch["box_scores"], hidx)
        for precise in (True, False):
            out = engine.head_forward(ih, batch["features"]["3"].to(DEV), batch["image_shapes"], batch["box_features"].to(DEV),
                                      torch.cat(batch["box_coords"]).to(DEV), meta, precise_logits=precise, graph=False)
            lg = out["logits"].cpu()
            tag = "%s %s" % (name, "precise" if precise else "plain")
            stats(tag + " p", lg[:, :nc], lp); stats(tag + " s", lg[:, nc:], ls); stats(tag + " p|s", lg, torch.cat([lp, ls], 1))


otv = synthetic.hico_object_to_verb()
ih = build_head(42); tables = synthetic.make_transh_tables(42)
ih.box_pair_head.transh_head.set_tables(*tables); ih = ih.to(DEV).eval(); ih.box_pair_head.transh_head.transh.to(DEV)
for name, shapes, seed in (("c2_8x16x32", [(16, 32)] * 8, 42), ("c5_3x32x64", [(32, 64)] * 3, 43), ("c5_2x40x80", [(40, 80)] * 2, 44),
                           ("c5_2x64x64", [(64, 64)] * 2, 45), ("ragged", [(4, 8), (8, 16), (15, 30), (1, 2), (40, 80), (2, 2)], 46)):
    run(name, shapes, seed, ih, head_params(), tables, otv, 117, 49)
# V-COCO parameterisation, several seeds: how lucky is the single suppressor column?
g = torch.Generator().manual_seed(3)
votv = []
for _ in range(80):
    k = int(torch.randint(1, 7, (1,), generator=g)); votv.append(sorted(set(torch.randint(0, 24, (k,), generator=g).tolist())))
for seed in (7, 8, 9):
    vh = build_interaction_head(votv, human_idx=1, num_classes=24, seed=seed)
    VP = {k: v.detach().clone() for k, v in vh.state_dict().items()}
    vt = synthetic.make_transh_tables(11, 80, 24, 50)
    vh.box_pair_head.transh_head.set_tables(*vt); vh = vh.to(DEV).eval(); vh.box_pair_head.transh_head.transh.to(DEV)
    run("vcoco seed %d" % seed, [(3, 9), (6, 6), (2, 14)], 21, vh, VP, vt, votv, 24, 1)
