"""Stage-by-stage error attribution of the bf16 pipeline against the fp32 oracle on the benched batch (diagnostics)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import build_head, head_params, synthetic
from oracle import head_oracle as O
from skghoi_b200 import engine

DEV = "cuda:0"
shapes = [(16, 32)] * int(os.environ.get("IMAGES", "8"))
batch = synthetic.make_batch(shapes, seed=42)
P = head_params()
tables = synthetic.make_transh_tables(42)
taps = {}
with torch.no_grad():
    g = O.graph_head_forward(P, tables, batch["features"]["3"], batch["image_shapes"], batch["box_features"], batch["box_coords"],
                             batch["box_labels"], batch["box_scores"], synthetic.hico_object_to_verb(), row_pick="slice", taps=taps)
feat_ref = torch.cat(g["feat"])
lp_ref = torch.nn.functional.linear(feat_ref, P["box_pair_predictor.weight"], P["box_pair_predictor.bias"])
ih = build_head(42)
ih.box_pair_head.transh_head.set_tables(*tables)
ih = ih.to(DEV).eval(); ih.box_pair_head.transh_head.transh.to(DEV)
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
with torch.no_grad():
    out = engine.head_forward(ih, batch["features"]["3"].to(DEV), batch["image_shapes"], batch["box_features"].to(DEV),
                              torch.cat(batch["box_coords"]).to(DEV), meta, want_taps=True, precise_logits=True, graph=False)
torch.cuda.synchronize()
NH = out["NH"]; R = 1024


def rep(name, got, ref):
    got = got.double().cpu(); ref = ref.double()
    d = (got - ref).abs()
    rms = ref.pow(2).mean().sqrt()
    print("%-18s rel_l2 %.3e  max|d|/rms %.3e  max|d|/max|ref| %.3e  frac(|d| > 1e-2|ref| + 1e-2 rms) %.2e"
          % (name, float((got - ref).norm() / ref.norm()), float(d.max() / rms), float(d.max() / ref.abs().max()),
             float((d > 1e-2 * ref.abs() + 1e-2 * rms).double().mean())))


rep("sp", out["sp"].float(), torch.cat(taps["spatial1024"]))
rep("hN", out["hN32"], torch.cat(taps["graph_human_node"]))
rep("oN", out["oN32"], torch.cat(taps["graph_object_node"]))
rep("adjacency", out["adjacency"], torch.cat([a.flatten() for a in taps["adjacency"]]))
rep("h_enc", out["enc32"][:NH], torch.cat(taps["h_node_enc"]))
rep("o_enc", out["enc32"][NH:], torch.cat(taps["node_enc"]))
feat = out["feat"].float().cpu()
hi, lo = feat[:, :2 * R], feat[:, 2 * R:]
rep("feat(hi)", hi, feat_ref); rep("feat(hi+lo)", hi + lo, feat_ref)
rep("feat att1", (hi + lo)[:, :R], feat_ref[:, :R]); rep("feat att2 (g)", (hi + lo)[:, R:], feat_ref[:, R:])
lg = out["logits"].cpu()
rep("logits_p", lg[:, :117], lp_ref)
# predictor alone: fp32 predictor on the GPU features vs the GPU logits
rep("pred-only", lg[:, :117], torch.nn.functional.linear((hi + lo), P["box_pair_predictor.weight"], P["box_pair_predictor.bias"]))
# how do the violations distribute over pairs?
d = (lg[:, :117].double() - lp_ref.double()).abs(); rms = lp_ref.double().pow(2).mean().sqrt()
bad = (d > 1e-2 * lp_ref.abs().double() + 1e-2 * rms)
print("violating elements %d of %d in %d pairs; feat-norm of violating pairs vs all: %.3f vs %.3f"
      % (int(bad.sum()), bad.numel(), int(bad.any(1).sum()), float(feat_ref[bad.any(1)].norm(dim=1).mean()) if bad.any() else 0.0,
         float(feat_ref.norm(dim=1).mean())))
fe = (hi + lo - feat_ref).norm(dim=1) / feat_ref.norm(dim=1)
print("per-pair feat rel err: median %.3e  p99 %.3e  max %.3e" % (float(fe.median()), float(fe.quantile(0.99)), float(fe.max())))
