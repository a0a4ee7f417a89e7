"""Diagnostics: the SKG_HEAD_TRAIN forward layout against the inference layout, buffer by buffer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import build_head, synthetic
from skghoi_b200 import engine, train

DEV = "cuda:0"
shapes = [(3, 7), (2, 5)] if len(sys.argv) < 2 else [(16, 32)] * int(sys.argv[1])
batch = synthetic.make_batch(shapes, seed=77, with_targets=True)
ih = build_head(42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(DEV).train(); ih.box_pair_head.transh_head.transh.to(DEV)
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
f3 = batch["features"]["3"].to(DEV); bf = batch["box_features"].to(DEV); bc = torch.cat(batch["box_coords"]).to(DEV)
ts = train.train_state(ih)
with torch.no_grad():
    ev = engine.head_forward(ih, f3, batch["image_shapes"], bf, bc, meta, precise_logits=False, graph=False)
    ev = {k: ev[k].clone() for k in ("logits", "feat", "H_att", "S_raw", "H_os", "H_so", "H_g", "adjacency", "H2", "sp", "hproj", "oproj", "U2h", "U2o", "pair2cell", "enc16")}
    tr = engine.head_forward(ih, f3, batch["image_shapes"], bf, bc, meta, precise_logits=False, graph=False, train=True, pk=ts.pk)
torch.cuda.synchronize()
p2c = ev["pair2cell"].long()
R = 1024
def cmp(name, a, b):
    a = a.float(); b = b.float()
    print("%-12s max|d| %.3e  max|ref| %.3e" % (name, float((a - b).abs().max()), float(b.abs().max())))
for k in ("sp", "hproj", "oproj", "H_att", "H_os", "H_so", "H_g", "adjacency", "enc16", "U2h", "U2o", "H2"):
    cmp(k, tr[k], ev[k])
cmp("S_raw(att)", tr["S_raw"][:, :R], ev["S_raw"])
cmp("feat", tr["feat"][p2c], ev["feat"][:, :2 * R])
cmp("logits", tr["logits"][p2c], ev["logits"])
pk = ts.pk
ref_wts = torch.relu(ev["H_att"].float() @ pk.W3att.float().t() + pk.b3att)
cmp("wts", tr["wts"], ref_wts)
ref_adj = ref_wts @ pk.w_adj + pk.b_adj
cmp("adj(eval)", ev["adjacency"], ref_adj)
cmp("adj(train)", tr["adjacency"], ref_adj)
dp = tr["dot_part"]
print("dot_part train: sum over parts vs ref", float((dp.sum(1) + pk.b_adj - ref_adj).abs().max()))
print(dp[:2])
print("ref  ", ref_adj[:8].tolist())
print("eval ", ev["adjacency"][:8].tolist())
print("train", tr["adjacency"][:8].tolist())
with torch.no_grad():
    ev2 = engine.head_forward(ih, f3, batch["image_shapes"], bf, bc, meta, precise_logits=False, graph=False)
    print("eval2", ev2["adjacency"][:8].tolist())
    print("eval2 dot_part", ev2["dot_part"][:2])
    ev3 = engine.head_forward(ih, f3, batch["image_shapes"], bf, bc, meta, precise_logits=False, graph=False, want_taps=True)
    print("eval3", ev3["adjacency"][:8].tolist())
