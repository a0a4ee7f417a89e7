"""Design check (CPU, torch): emulates the planned bf16 kernel pipeline -- same algebraic
restructuring and the same bf16 rounding points as the CUDA engine -- and compares its logits
with the reference golden.  Not part of the product or the tests."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import torch, torch.nn.functional as F
from helpers import load_golden, build_head, synthetic, logits_close
from oracle import head_oracle as O

def r(x, site=None): return x.bfloat16().float()

def run(case, round_fn=r):
    gold = load_golden(case)
    ih = build_head(42); G = ih.box_pair_head
    ent, rel, nrm = synthetic.make_transh_tables(42)
    batch = synthetic.make_batch(gold["shapes_hn"], seed=gold["seed"], image_shapes=gold["image_shapes"])
    rb = round_fn
    with torch.no_grad():
        lin = lambda x, m, site='lin': rb(x, site) @ rb(m.weight, site+'_w').t() + m.bias
        bf = batch["box_features"].flatten(1)
        x1 = F.relu(lin(bf, G.box_head[1], 'box')); node_all = F.relu(lin(x1, G.box_head[3], 'box'))
        glob = batch["features"]["3"].mean(dim=(2, 3))
        att = G.attention_head.packed(); os_ = G.obj_to_sub.packed(); so = G.sub_to_obj.packed(); gg = G.attention_head_g.packed()
        feats = []; c = 0
        for b, (nh, n) in enumerate(gold["shapes_hn"]):
            coords = batch["box_coords"][b]; node = node_all[c:c+n]; c += n
            x, y = torch.meshgrid(torch.arange(nh), torch.arange(n), indexing="ij"); xk, yk = torch.nonzero(x != y).unbind(1)
            x = x.flatten(); y = y.flatten()
            enc = O.compute_spatial_ratio_encodings([coords[x]], [coords[y]], [batch["image_shapes"][b]])
            s1 = F.relu(enc @ G.spatial_head[0].weight.t() + G.spatial_head[0].bias)   # fp32 layer 1
            s2 = F.relu(lin(s1, G.spatial_head[2], 'spmlp')); sp = F.relu(lin(s2, G.spatial_head[4], 'spmlp'))
            hN = F.relu(lin(torch.cat([node[:nh], ent[49].expand(nh, -1)], 1), G.fc_head[0], 'fc'))
            oN = F.relu(lin(torch.cat([node, ent[:n]], 1), G.fc_tail[0], 'fc'))
            W1, b1, W2, b2, W3, b3 = att
            def node_proj(hx, ox):
                Uh = rb(hx,'nodeproj') @ rb(W1[:, :1024],'nodeproj_w').t(); Uo = rb(ox,'nodeproj') @ rb(W1[:, 1024:],'nodeproj_w').t() + b1
                return Uh, Uo
            Uh, Uo = node_proj(hN, oN)
            Vos = rb(oN,'nodeproj') @ rb(os_[0],'nodeproj_w').t() + os_[1]; Vso = rb(hN,'nodeproj') @ rb(so[0],'nodeproj_w').t() + so[1]
            Gb = rb(glob[b:b+1],'nodeproj') @ rb(gg[0],'nodeproj_w').t() + gg[1]
            S_att = rb(sp,'sp') @ rb(W2,'W2').t() + b2
            H_att = F.relu(S_att * (Uh[x] + Uo[y]))
            H_os = F.relu((rb(sp,'sp') @ rb(os_[2],'W2').t() + os_[3]) * Vos[y])
            H_so = F.relu((rb(sp,'sp') @ rb(so[2],'W2').t() + so[3]) * Vso[x])
            H_g = F.relu((rb(sp,'sp') @ rb(gg[2],'W2').t() + gg[3]) * Gb)
            wts = F.relu(rb(H_att,'H_att') @ rb(W3,'W3att_adj').t() + b3)
            adj = (wts @ G.adjacency.weight.t() + G.adjacency.bias).reshape(nh, n)
            Zh = (adj.softmax(1)[..., None] * rb(H_os,'H_msg').reshape(nh, n, -1)).sum(1)
            Zo = (adj.t().softmax(1)[..., None] * rb(H_so,'H_msg').reshape(nh, n, -1).permute(1, 0, 2)).sum(1)
            msg_h = F.relu(rb(Zh,'Z') @ rb(os_[4],'W3msg').t() + os_[5]); msg_o = F.relu(rb(Zo,'Z') @ rb(so[4],'W3msg').t() + so[5])
            h_enc = G.norm_h(hN + msg_h); enc_o = G.norm_o(oN + msg_o)
            Uh2, Uo2 = node_proj(h_enc, enc_o)
            cell = xk * n + yk
            H2 = F.relu(rb(S_att,'S_raw')[cell] * (Uh2[xk] + Uo2[yk]))
            f1 = F.relu(rb(H2,'H2') @ rb(W3,'W3fin').t() + b3); f2 = F.relu(rb(H_g,'H2')[cell] @ rb(gg[4],'W3fin').t() + gg[5])
            feats.append(torch.cat([f1, f2], 1))
        feat = torch.cat(feats)
        lp = lin(feat, ih.box_pair_predictor,'pred'); ls = lin(feat, ih.box_pair_suppressor,'pred')
    for name, got, ref in (("logits_p", lp, gold["logits_p"]), ("logits_s", ls, gold["logits_s"])):
        ok, st = logits_close(got, ref)
        cos = F.cosine_similarity(got.flatten(), ref.flatten(), dim=0)
        print(case, name, "ok" if ok else "FAIL", st, "cos", float(cos))

if __name__ == "__main__":
    for case in ("head_small2_eval", "head_c1_eval"):
        run(case, round_fn=lambda x, site=None: x)    # restructuring only (fp32): must be ~exact
        run(case)                          # bf16 rounding points
