"""Pins oracle/head_oracle.py (the CPU restatement) to the golden vectors produced by the
UNMODIFIED reference (tests/golden/make_golden.py).  CPU only."""
import pytest
import torch

from helpers import load_golden, head_params, synthetic
from oracle import head_oracle as O


def test_weights_rebuild_matches_reference_checksum():
    chk = load_golden("weights_checksum")
    P = head_params()
    assert set(P) == set(chk["checksum"])
    assert sum(v.numel() for v in P.values()) == chk["num_params"] == 29565175
    for k, v in P.items():
        assert tuple(v.shape) == chk["shapes"][k]
        got = torch.stack([v.double().sum(), v.double().abs().sum()])
        assert torch.equal(got, chk["checksum"][k]), k


def _same_nan_inf(a, b):
    return torch.equal(torch.isnan(a), torch.isnan(b)) and torch.equal(torch.isinf(a), torch.isinf(b))


@pytest.mark.parametrize("key,fn,boxes", [
    ("edge36", O.compute_spatial_encodings, ("b1", "b2")),
    ("edge46", O.compute_spatial_ratio_encodings, ("b1", "b2")),
    ("rand36", O.compute_spatial_encodings, ("r1", "r2")),
    ("rand46", O.compute_spatial_ratio_encodings, ("r1", "r2")),
])
def test_spatial_oracle(key, fn, boxes):
    g = load_golden("spatial_ops")
    out = fn([g[boxes[0]]], [g[boxes[1]]], [g["shape"]])
    assert _same_nan_inf(out, g[key])
    fin = torch.isfinite(g[key])
    torch.testing.assert_close(out[fin], g[key][fin], rtol=1e-6, atol=1e-6)


def test_spatial_oracle_two_images_and_focal():
    g = load_golden("spatial_ops")
    out = O.compute_spatial_ratio_encodings([g["r1"][:50], g["r1"][50:]], [g["r2"][:50], g["r2"][50:]],
                                            [g["shape"], (480, 640)])
    torch.testing.assert_close(out, g["rand46_two_images"], rtol=1e-6, atol=1e-6)
    x, y = g["focal_x"], g["focal_y"]
    torch.testing.assert_close(O.binary_focal_loss(x, y, reduction="sum", gamma=0.2), g["focal_g02_sum"])
    torch.testing.assert_close(O.binary_focal_loss(x, y, reduction="sum", gamma=2.0), g["focal_g20_sum"])
    torch.testing.assert_close(O.binary_focal_loss(x, y), g["focal_g20_mean"])
    torch.testing.assert_close(O.binary_focal_loss(x, y, reduction="none"), g["focal_g20_none"])


def test_transh_oracle():
    g = load_golden("transh_ops")
    ent, rel, nrm = synthetic.make_transh_tables(42)
    for p in (1, 2):
        for nf in (True, False):
            out = O.transh_forward(ent, rel, nrm, g["head"], g["relation"], g["tail"], p, nf)
            torch.testing.assert_close(out[4], g[f"score_p{p}_norm{int(nf)}"], rtol=1e-5, atol=1e-6)
    out = O.transh_forward(ent, rel, nrm, g["head"], g["relation"], g["tail"], 2, True)
    assert torch.equal(out[0][:64], g["h_"]) and torch.equal(out[1][:64], g["r"])
    assert torch.equal(out[2][:64], g["r_norm"]) and torch.equal(out[3][:64], g["t_"])
    s2 = g["score_p2_norm1"]
    torch.testing.assert_close(O.margin_loss(s2[:2048, None], s2[2048:, None], 1.0), g["margin1"])
    torch.testing.assert_close(O.margin_loss(s2[:2048, None], s2[2048:, None], 6.0), g["margin6"])
    torch.testing.assert_close(O.margin_loss(s2[:1024].view(-1, 1), s2[1024:].view(-1, 1024).permute(1, 0), 6.0, 1.5),
                               g["margin_adv"])
    torch.testing.assert_close(O.negative_sampling_loss(s2, 1.0), g["negsamp_margin1"])
    me, mr, mn = g["margin_mode_tables"]
    out = O.transh_forward(me, mr, mn, g["head"], g["relation"], g["tail"], 1, True)
    torch.testing.assert_close(6.0 - out[4], g["margin_mode_score"], rtol=1e-5, atol=1e-5)


def _run_oracle(gold, taps=None, row_pick="slice"):
    P = head_params()
    tables = synthetic.make_transh_tables(42)
    batch = synthetic.make_batch(gold["shapes_hn"], seed=gold["seed"], image_shapes=gold["image_shapes"],
                                 with_targets=gold["train"])
    torch.manual_seed(1234)
    return O.interaction_head_forward(P, tables, batch["features"]["3"], batch["image_shapes"], batch["box_features"],
                                      batch["box_coords"], batch["box_labels"], batch["box_scores"],
                                      synthetic.hico_object_to_verb(), batch["targets"], training=gold["train"],
                                      row_pick=row_pick, taps=taps)


@pytest.mark.parametrize("case,row_pick", [("head_small2_eval", "loop"), ("head_small2_train", "slice"),
                                           ("head_c1_eval", "slice")])
def test_head_oracle_matches_reference(case, row_pick):
    gold = load_golden(case)
    taps = {}
    with torch.no_grad():
        results, logits_p, logits_s = _run_oracle(gold, taps, row_pick)
    for k in ("x", "y", "x_keep", "y_keep"):
        for a, b in zip(taps[k], gold[k]):
            assert torch.equal(a, b), k
    for a, b in zip(taps["spatial46"], gold["spatial46"]):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-6)
    for k in ("adjacency", "h_node_enc", "node_enc", "graph_human_node", "graph_object_node"):
        for a, b in zip(taps[k], gold[k]):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(logits_p, gold["logits_p"], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(logits_s, gold["logits_s"], rtol=1e-4, atol=1e-4)
    n_img = len(gold["shapes_hn"])
    for r, gr in zip(results[:n_img], gold["results"]):
        assert torch.equal(r["index"], gr["index"]) and torch.equal(r["prediction"], gr["prediction"])
        assert torch.equal(r["object"], gr["object"])
        torch.testing.assert_close(r["scores"], gr["scores"], rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(r["prior"], gr["prior"], rtol=1e-6, atol=1e-7)
        torch.testing.assert_close(r["weights"], gr["weights"], rtol=1e-4, atol=1e-6)
        if gold["train"]:
            assert torch.equal(r["labels"], gr["labels"]) and torch.equal(r["unary_labels"], gr["unary_labels"])
    if gold["train"]:
        losses = results[-1]
        for k in ("hoi_loss", "interactiveness_loss", "transH_loss"):
            torch.testing.assert_close(losses[k].reshape(()), gold[k].reshape(()), rtol=1e-4, atol=1e-6)
