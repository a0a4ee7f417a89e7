"""Generates the golden fixtures under tests/golden/ by running the UNMODIFIED reference head
(imported from /root/reference through oracle/ref_harness.py) on seeded synthetic inputs.

Run in the build container only (the reference checkout does not exist on the GPU box):
    python tests/golden/make_golden.py
The fixtures it writes are committed; tests/test_oracle_vs_golden.py pins oracle/head_oracle.py
to them and the GPU parity tests compare the CUDA path with both.

Weights are NOT stored (29.6 M parameters): every consumer rebuilds them with the PyTorch default
initialisers under ``torch.manual_seed(42)`` in the reference's construction order
(models.py:164-191); ``weights_checksum.pt`` holds per-tensor sums to prove the rebuild is identical.
"""
import os
import sys
from contextlib import contextmanager

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from oracle import ref_harness as R          # noqa: E402
from skghoi_b200 import synthetic as S       # noqa: E402

torch.set_num_threads(8)


@contextmanager
def cuda_literal_to_cpu():
    """MarginLoss.py:33 / head.py:232 call ``.to('cuda')``; map the literal to CPU here."""
    orig = torch.Tensor.to

    def to(self, *a, **k):
        a = tuple("cpu" if (isinstance(x, str) and x == "cuda") else x for x in a)
        return orig(self, *a, **k)
    torch.Tensor.to = to
    orig_mod = torch.nn.Module.to

    def mto(self, *a, **k):
        a = tuple("cpu" if (isinstance(x, str) and x == "cuda") else x for x in a)
        return orig_mod(self, *a, **k)
    torch.nn.Module.to = mto
    try:
        yield
    finally:
        torch.Tensor.to = orig
        torch.nn.Module.to = orig_mod


def run_case(name, shapes_hn, image_shapes=None, seed=42, train=False, taps=True):
    otv = S.hico_object_to_verb()
    graph, ih = R.build_heads(otv, seed=42)
    tables = S.make_transh_tables(42)
    batch = S.make_batch(shapes_hn, seed=seed, image_shapes=image_shapes, with_targets=train)
    ih.train(train)
    head_mod = R.load()[0]
    rec = {}
    # tap intermediates by wrapping sub-modules (no reference file is edited)
    hooks = []
    if taps:
        def tap(key):
            def fn(mod, inp, out):
                rec.setdefault(key, []).append(out.detach().clone())
            return fn
        hooks.append(graph.spatial_head.register_forward_hook(tap("spatial1024")))
        hooks.append(graph.norm_h.register_forward_hook(tap("h_node_enc_all")))
        hooks.append(graph.norm_o.register_forward_hook(tap("node_enc_all")))
        hooks.append(graph.adjacency.register_forward_hook(tap("adjacency_all")))
        hooks.append(graph.fc_head.register_forward_hook(tap("graph_human_node_full")))
        hooks.append(graph.fc_tail.register_forward_hook(tap("graph_object_node_full")))
        orig_sp = head_mod.compute_spatial_ratio_encodings

        def sp_tap(*a, **k):
            out = orig_sp(*a, **k)
            rec.setdefault("spatial46", []).append(out.detach().clone())
            return out
        head_mod.compute_spatial_ratio_encodings = sp_tap
    with torch.no_grad(), R.fixed_transh(tables), cuda_literal_to_cpu():
        torch.manual_seed(1234)         # fixes torch.randperm at head.py:939
        outs = graph(batch["features"], batch["image_shapes"], batch["box_features"],
                     batch["box_coords"], batch["box_labels"], batch["box_scores"], batch["targets"])
        feat = torch.cat(outs[0])
        logits_p = ih.box_pair_predictor(feat)
        logits_s = ih.box_pair_suppressor(feat)
        results = ih.postprocess(logits_p, logits_s, outs[5], outs[1], outs[2], outs[3], outs[4])
        gold = dict(shapes_hn=list(shapes_hn), image_shapes=batch["image_shapes"], seed=seed, train=train,
                    logits_p=logits_p, logits_s=logits_s,
                    boxes_h=outs[1], boxes_o=outs[2], object=outs[3],
                    results=[{k: v for k, v in r.items() if k not in ("boxes_h", "boxes_o")} for r in results],
                    feat_rowsum=feat.double().sum(1).float(), feat_head=feat[:8].clone())
        if train:
            gold["labels"] = outs[4]
            gold["pos"] = outs[6]; gold["neg"] = outs[7]
            gold["hoi_loss"] = ih.compute_interaction_classification_loss(results)
            gold["interactiveness_loss"] = ih.compute_interactiveness_loss(results)
            # intended call of head.py:229-233 (SURVEY defect 2): the class's own 1-argument contract
            _, _, _, ML, NS = R.load()
            ns = NS(loss=ML(margin=1), batch_size=256)
            n_p = len(torch.nonzero(torch.cat([r["unary_labels"] for r in results])))
            flat = torch.cat([torch.cat(outs[6]), torch.cat(outs[7])])
            gold["transH_loss"] = ns(flat) / n_p
    for h in hooks:
        h.remove()
    if taps:
        head_mod.compute_spatial_ratio_encodings = orig_sp
        num_iter = 2
        gold["spatial46"] = rec["spatial46"]
        gold["spatial1024_rowsum"] = [t.double().sum(1).float() for t in rec["spatial1024"]]
        # hooks fire once per message-passing iteration; iterations are identical (SURVEY defect 5)
        gold["h_node_enc"] = rec["h_node_enc_all"][num_iter - 1::num_iter]
        gold["node_enc"] = rec["node_enc_all"][num_iter - 1::num_iter]
        adj = rec["adjacency_all"][num_iter - 1::num_iter]
        gold["adjacency"] = [a.reshape(nh, n) for a, (nh, n) in zip(adj, shapes_hn)]
        gold["graph_human_node"] = [t[::n].clone() for t, (nh, n) in zip(rec["graph_human_node_full"], shapes_hn)]
        gold["graph_object_node"] = [t[:n].clone() for t, (nh, n) in zip(rec["graph_object_node_full"], shapes_hn)]
    # pair indices exactly as head.py:847-860 builds them
    gold["x"], gold["y"], gold["x_keep"], gold["y_keep"] = [], [], [], []
    for nh, n in shapes_hn:
        x, y = torch.meshgrid(torch.arange(nh), torch.arange(n))
        xk, yk = torch.nonzero(x != y).unbind(1)
        gold["x"].append(x.flatten()); gold["y"].append(y.flatten())
        gold["x_keep"].append(xk); gold["y_keep"].append(yk)
    torch.save(gold, os.path.join(HERE, name + ".pt"))
    print("wrote", name, "P=", feat.shape[0], "L=", sum(len(r["scores"]) for r in results),
          {k: float(gold[k]) for k in ("hoi_loss", "interactiveness_loss", "transH_loss") if k in gold})
    return graph, ih


def main():
    otv = S.hico_object_to_verb()
    head_mod, ops, TransH, MarginLoss, NegativeSampling = R.load()

    # ---- weights checksum (rebuild-from-seed proof)
    graph, ih = R.build_heads(otv, seed=42)
    sd = {k: v for k, v in ih.state_dict().items()}
    chk = {k: torch.stack([v.double().sum(), v.double().abs().sum()]) for k, v in sd.items()}
    torch.save(dict(checksum=chk, shapes={k: tuple(v.shape) for k, v in sd.items()},
                    num_params=sum(v.numel() for v in sd.values())), os.path.join(HERE, "weights_checksum.pt"))
    print("state_dict tensors:", len(sd), "params:", sum(v.numel() for v in sd.values()))

    # ---- spatial encodings: edge cases + random, both variants (ops.py:20-83, 85-157)
    b1, b2, shape = S.edge_case_boxes()
    gen = torch.Generator().manual_seed(7)
    rb, _, _ = S.make_detections([(8, 64)], [shape], gen)
    r1 = rb[0][torch.randint(0, 64, (200,), generator=gen)]; r2 = rb[0][torch.randint(0, 64, (200,), generator=gen)]
    sp = dict(shape=shape, b1=b1, b2=b2, r1=r1, r2=r2,
              edge36=ops.compute_spatial_encodings([b1], [b2], [shape]),
              edge46=ops.compute_spatial_ratio_encodings([b1], [b2], [shape]),
              rand36=ops.compute_spatial_encodings([r1], [r2], [shape]),
              rand46=ops.compute_spatial_ratio_encodings([r1], [r2], [shape]),
              rand46_two_images=ops.compute_spatial_ratio_encodings([r1[:50], r1[50:]], [r2[:50], r2[50:]],
                                                                    [shape, (480, 640)]))
    x = torch.rand(300, generator=gen).clamp(1e-4, 1 - 1e-4); yb = (torch.rand(300, generator=gen) > 0.8).float()
    sp["focal_x"] = x; sp["focal_y"] = yb
    sp["focal_g02_sum"] = ops.binary_focal_loss(x, yb, reduction="sum", gamma=0.2)
    sp["focal_g20_sum"] = ops.binary_focal_loss(x, yb, reduction="sum", gamma=2.0)
    sp["focal_g20_mean"] = ops.binary_focal_loss(x, yb)
    sp["focal_g20_none"] = ops.binary_focal_loss(x, yb, reduction="none")
    torch.save(sp, os.path.join(HERE, "spatial_ops.pt"))
    print("wrote spatial_ops  nan count edge46:", int(torch.isnan(sp["edge46"]).sum()))

    # ---- TransH / MarginLoss / NegativeSampling standalone (TransH.py:87-106, MarginLoss.py:28-36)
    ent, rel, nrm = S.make_transh_tables(42)
    T = 4096
    hd = torch.randint(0, 80, (T,), generator=gen); rl = torch.randint(0, 117, (T,), generator=gen)
    tl = torch.randint(0, 80, (T,), generator=gen)
    kg = dict(head=hd, relation=rl, tail=tl)
    with cuda_literal_to_cpu(), torch.no_grad():
        for p in (1, 2):
            for nf in (True, False):
                m = TransH(80, 117, dim=50, p_norm=p, norm_flag=nf)
                m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
                h_, r, w, t_, score = m(hd, rl, tl)
                kg[f"score_p{p}_norm{int(nf)}"] = score
                if p == 2 and nf:
                    kg["h_"], kg["r"], kg["r_norm"], kg["t_"] = h_[:64].clone(), r[:64].clone(), w[:64].clone(), t_[:64].clone()
                    kg["regularization"] = m.regularization(hd, rl, tl)
        m = TransH(80, 117, dim=50, p_norm=1, norm_flag=True, margin=6.0, epsilon=2.0)
        kg["margin_mode_tables"] = (m.ent_embeddings.weight.clone(), m.rel_embeddings.weight.clone(), m.norm_vector.weight.clone())
        kg["margin_mode_score"] = m(hd, rl, tl)
        s2 = kg["score_p2_norm1"]
        kg["margin1"] = MarginLoss(margin=1)(s2[:2048, None], s2[2048:, None])
        kg["margin6"] = MarginLoss(margin=6.0)(s2[:2048, None], s2[2048:, None])
        kg["margin_adv"] = MarginLoss(adv_temperature=1.5, margin=6.0)(s2[:1024].view(-1, 1), s2[1024:].view(-1, 1024).permute(1, 0))
        kg["negsamp_margin1"] = NegativeSampling(loss=MarginLoss(margin=1), batch_size=256)(s2)
    # gradients of the margin loss wrt the three tables (autograd through the reference classes)
    with cuda_literal_to_cpu():
        m = TransH(80, 117, dim=50, p_norm=2, norm_flag=True)
        with torch.no_grad():
            m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
        score = m(hd, rl, tl)[4]
        loss = NegativeSampling(loss=MarginLoss(margin=1), batch_size=256)(score)
        loss.backward()
        kg["grad_ent"] = m.ent_embeddings.weight.grad.clone()
        kg["grad_rel"] = m.rel_embeddings.weight.grad.clone()
        kg["grad_norm"] = m.norm_vector.weight.grad.clone()
        kg["grad_loss"] = loss.detach()
    torch.save(kg, os.path.join(HERE, "transh_ops.pt"))
    print("wrote transh_ops")

    # ---- head cases
    run_case("head_small2_eval", [(3, 7), (2, 5)], image_shapes=[(480, 640), (600, 800)], seed=3)
    run_case("head_small2_train", [(3, 7), (2, 5)], image_shapes=[(480, 640), (600, 800)], seed=3, train=True)
    run_case("head_c1_eval", [(16, 32)], seed=42)


if __name__ == "__main__":
    main()
