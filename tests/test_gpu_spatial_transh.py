"""GPU parity: pair enumeration + spatial encoding, TransH, margin / focal loss -- CUDA library
(through the C-ABI) against the CPU oracle and the golden vectors of the reference."""
import pytest
import torch

from helpers import load_golden, synthetic
from oracle import head_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _same_nan_inf(a, b):
    return torch.equal(torch.isnan(a), torch.isnan(b)) and torch.equal(torch.isinf(a), torch.isinf(b))


def _close_finite(got, ref, rtol=1e-5, atol=1e-5):
    assert _same_nan_inf(got, ref)
    fin = torch.isfinite(ref)
    torch.testing.assert_close(got[fin], ref[fin], rtol=rtol, atol=atol)


@pytest.mark.parametrize("key,name,boxes", [
    ("edge36", "compute_spatial_encodings", ("b1", "b2")),
    ("edge46", "compute_spatial_ratio_encodings", ("b1", "b2")),
    ("rand36", "compute_spatial_encodings", ("r1", "r2")),
    ("rand46", "compute_spatial_ratio_encodings", ("r1", "r2")),
])
def test_spatial_rows_vs_golden(key, name, boxes):
    from skghoi_b200 import ops
    g = load_golden("spatial_ops")
    out = getattr(ops, name)([g[boxes[0]].to(DEV)], [g[boxes[1]].to(DEV)], [g["shape"]]).cpu()
    _close_finite(out, g[key])


def test_spatial_rows_two_images_and_errors():
    from skghoi_b200 import ops
    g = load_golden("spatial_ops")
    out = ops.compute_spatial_ratio_encodings([g["r1"][:50].to(DEV), g["r1"][50:].to(DEV)],
                                              [g["r2"][:50].to(DEV), g["r2"][50:].to(DEV)],
                                              [g["shape"], (480, 640)]).cpu()
    _close_finite(out, g["rand46_two_images"])
    with pytest.raises(AssertionError):       # ops.py:43-44
        ops.compute_spatial_encodings([g["r1"].to(DEV)], [g["r2"].to(DEV)], [(0, 10)])
    empty = ops.compute_spatial_ratio_encodings([torch.zeros(0, 4, device=DEV)], [torch.zeros(0, 4, device=DEV)],
                                                [(10, 10)])
    assert empty.shape == (0, 46)


@pytest.mark.parametrize("shapes", [[(16, 32)], [(3, 7), (2, 5)], [(1, 2)], [(4, 4), (15, 30), (2, 80), (40, 80)]])
def test_pair_enumeration_bit_exact_and_encoding(shapes):
    from skghoi_b200.engine import enumerate_pairs
    gen = torch.Generator().manual_seed(5)
    img_shapes = [(480 + 16 * k, 640 + 32 * k) for k in range(len(shapes))]
    boxes, labels, scores = synthetic.make_detections(shapes, img_shapes, gen)
    out = enumerate_pairs([b.to(DEV) for b in boxes], [s[0] for s in shapes], img_shapes)
    co, po = 0, 0
    for b, (nh, n) in enumerate(shapes):
        x, y = torch.meshgrid(torch.arange(nh), torch.arange(n), indexing="ij")
        xk, yk = torch.nonzero(x != y).unbind(1)
        N2, Pn = nh * n, len(xk)
        assert torch.equal(out["x"][co:co + N2].cpu(), x.flatten())
        assert torch.equal(out["y"][co:co + N2].cpu(), y.flatten())
        assert torch.equal(out["x_keep"][po:po + Pn].cpu(), xk)
        assert torch.equal(out["y_keep"][po:po + Pn].cpu(), yk)
        mask = (x != y).flatten()
        c2p = out["cell2pair"][co:co + N2].cpu()
        assert torch.equal(c2p >= 0, mask)
        assert torch.equal(c2p[mask].long(), torch.arange(Pn) + po)
        assert torch.equal(out["pair2cell"][po:po + Pn].cpu().long(), torch.nonzero(mask).squeeze(1) + co)
        ref = O.compute_spatial_ratio_encodings([boxes[b][x.flatten()]], [boxes[b][y.flatten()]], [img_shapes[b]])
        if torch.isnan(ref).sum() > 0:
            ref = torch.nan_to_num(ref)
        got = out["enc"][co:co + N2].cpu()
        torch.testing.assert_close(got, ref, rtol=1e-5, atol=1e-5)
        co += N2; po += Pn
    assert out["enc"].shape[0] == co and out["x_keep"].shape[0] == po


def test_pair_encoding_matches_reference_golden_and_nan_guard():
    from skghoi_b200.engine import enumerate_pairs
    gold = load_golden("head_c1_eval")
    batch = synthetic.make_batch(gold["shapes_hn"], seed=gold["seed"], image_shapes=gold["image_shapes"])
    out = enumerate_pairs([b.to(DEV) for b in batch["box_coords"]], [s[0] for s in gold["shapes_hn"]],
                          batch["image_shapes"])
    torch.testing.assert_close(out["enc"].cpu(), gold["spatial46"][0], rtol=1e-5, atol=1e-5)
    assert torch.equal(out["x_keep"].cpu(), gold["x_keep"][0]) and torch.equal(out["y_keep"].cpu(), gold["y_keep"][0])
    # NaN guard (head.py:866-868): image 0 has two zero-area boxes -> NaN -> nan_to_num on that image only
    b0 = torch.tensor([[5., 5., 5., 5.], [9., 9., 9., 9.], [1., 2., 30., 40.]])
    b1 = torch.tensor([[1., 1., 20., 30.], [0., 0., 0., 10.], [4., 4., 9., 9.]])
    shapes = [(100, 200), (100, 200)]
    out = enumerate_pairs([b0.to(DEV), b1.to(DEV)], [2, 1], shapes)
    for b, (bx, nh) in enumerate(((b0, 2), (b1, 1))):
        x, y = torch.meshgrid(torch.arange(nh), torch.arange(3), indexing="ij")
        ref = O.compute_spatial_ratio_encodings([bx[x.flatten()]], [bx[y.flatten()]], [shapes[b]])
        had_nan = bool(torch.isnan(ref).any())
        assert had_nan == (b == 0)
        if had_nan:
            ref = torch.nan_to_num(ref)
        got = out["enc"][b * 6: b * 6 + nh * 3].cpu()
        assert _same_nan_inf(got, ref)
        torch.testing.assert_close(got, ref, rtol=1e-5, atol=1e-5, equal_nan=True)
        assert int(out["nan_flag"][b]) == int(had_nan)


def _tables(dev=DEV):
    return [t.to(dev) for t in synthetic.make_transh_tables(42)]


@pytest.mark.parametrize("p_norm", [1, 2])
@pytest.mark.parametrize("norm_flag", [True, False])
def test_transh_scores_vs_golden(p_norm, norm_flag):
    from skghoi_b200.kge import TransH
    g = load_golden("transh_ops")
    ent, rel, nrm = _tables()
    m = TransH(80, 117, dim=50, p_norm=p_norm, norm_flag=norm_flag).to(DEV)
    with torch.no_grad():
        m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
        h_, r, w, t_, score = m(g["head"].to(DEV), g["relation"].to(DEV), g["tail"].to(DEV))
    ref = g[f"score_p{p_norm}_norm{int(norm_flag)}"]
    assert float((score.cpu() - ref).abs().max()) <= 1e-4          # north_star: TransH scores within 1e-4
    assert torch.equal(h_[:64].cpu(), g["h_"]) and torch.equal(r[:64].cpu(), g["r"])
    assert torch.equal(w[:64].cpu(), g["r_norm"]) and torch.equal(t_[:64].cpu(), g["t_"])


def test_transh_margin_mode_and_regularization():
    from skghoi_b200.kge import TransH
    g = load_golden("transh_ops")
    m = TransH(80, 117, dim=50, p_norm=1, norm_flag=True, margin=6.0, epsilon=2.0).to(DEV)
    me, mr, mn = g["margin_mode_tables"]
    with torch.no_grad():
        m.ent_embeddings.weight.copy_(me); m.rel_embeddings.weight.copy_(mr); m.norm_vector.weight.copy_(mn)
        out = m(g["head"].to(DEV), g["relation"].to(DEV), g["tail"].to(DEV))
    assert float((out.cpu() - g["margin_mode_score"]).abs().max()) <= 1e-4
    ent, rel, nrm = _tables()
    m2 = TransH(80, 117, dim=50, p_norm=2).to(DEV)
    with torch.no_grad():
        m2.ent_embeddings.weight.copy_(ent); m2.rel_embeddings.weight.copy_(rel); m2.norm_vector.weight.copy_(nrm)
        reg = m2.regularization(g["head"].to(DEV), g["relation"].to(DEV), g["tail"].to(DEV))
    torch.testing.assert_close(reg.cpu(), g["regularization"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("dim,ent_tot,rel_tot", [(50, 80, 117), (8, 5, 3), (100, 300, 40), (200, 1000, 50)])
def test_transh_generic_dims_vs_oracle(dim, ent_tot, rel_tot):
    """Other table shapes, incl. one that does not fit in shared memory (1000x200 + 2x50x200 fp32)."""
    from skghoi_b200.kge import TransH
    gen = torch.Generator().manual_seed(dim)
    ent = torch.randn(ent_tot, dim, generator=gen) * 0.3
    rel = torch.randn(rel_tot, dim, generator=gen) * 0.3
    nrm = torch.randn(rel_tot, dim, generator=gen) * 0.3
    T = 3001
    hd = torch.randint(0, ent_tot, (T,), generator=gen); rl = torch.randint(0, rel_tot, (T,), generator=gen)
    tl = torch.randint(0, ent_tot, (T,), generator=gen)
    for p in (1, 2):
        m = TransH(ent_tot, rel_tot, dim=dim, p_norm=p).to(DEV)
        with torch.no_grad():
            m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
            got = m(hd.to(DEV), rl.to(DEV), tl.to(DEV))
        ref = O.transh_forward(ent, rel, nrm, hd, rl, tl, p, True)
        assert float((got[4].cpu() - ref[4]).abs().max()) <= 1e-4 * max(1.0, float(ref[4].abs().max()))
        assert torch.equal(got[0].cpu(), ref[0]) and torch.equal(got[3].cpu(), ref[3])


@pytest.mark.parametrize("n_h,n", [(3, 7), (1, 2), (1, 3), (16, 32), (15, 23)])
def test_transh_head_grid_form(n_h, n):
    """transH_head.forward (head.py:558-582): implicit (cell, verb) grid == explicit index vectors.  Odd and even cell
    counts, fewer and more cell pairs than SMs (the rows go out through the bulk-copy kernel, pair by pair)."""
    from skghoi_b200.head import transH_head
    ent, rel, nrm = synthetic.make_transh_tables(42)
    th = transH_head(transh_dim=50, transh_p_norm=2, transh_norm_flag=True, human_idx=49, num_object=80, num_cls=117)
    th.set_tables(ent, rel, nrm, device=DEV)
    x, y = torch.meshgrid(torch.arange(n_h), torch.arange(n), indexing="ij")
    x, y = x.flatten(), y.flatten()
    got = th(x.to(DEV), y.to(DEV), n_h, n, None, None)
    N2 = n_h * n
    relations = torch.arange(117).repeat(N2); heads = torch.full((N2 * 117,), 49); tails = y.repeat_interleave(117)
    ref = O.transh_forward(ent, rel, nrm, heads, relations, tails, 2, True)
    for a, b in zip(got[:4], ref[:4]):
        assert torch.equal(a.cpu(), b)
    assert float((got[4].cpu() - ref[4]).abs().max()) <= 1e-4


def test_margin_loss_and_negative_sampling_vs_golden():
    from skghoi_b200.kge import MarginLoss, NegativeSampling
    g = load_golden("transh_ops")
    s2 = g["score_p2_norm1"].to(DEV)
    torch.testing.assert_close(MarginLoss(margin=1).to(DEV)(s2[:2048, None], s2[2048:, None]).cpu(), g["margin1"],
                               rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(MarginLoss(margin=6.0).to(DEV)(s2[:2048, None], s2[2048:, None]).cpu(), g["margin6"],
                               rtol=1e-5, atol=1e-6)
    adv = MarginLoss(adv_temperature=1.5, margin=6.0).to(DEV)(s2[:1024].view(-1, 1),
                                                               s2[1024:].view(-1, 1024).permute(1, 0).contiguous())
    torch.testing.assert_close(adv.cpu(), g["margin_adv"], rtol=1e-5, atol=1e-6)
    ns = NegativeSampling(loss=MarginLoss(margin=1), batch_size=256).to(DEV)
    torch.testing.assert_close(ns(s2).cpu(), g["negsamp_margin1"], rtol=1e-5, atol=1e-6)
    # six-argument call-site form of head.py:233
    six = ns([s2[:1000], s2[1000:2048]], [s2[2048:3000], s2[3000:]], None, None, None, None)
    torch.testing.assert_close(six.cpu(), g["negsamp_margin1"], rtol=1e-5, atol=1e-6)


def test_transh_margin_gradients_vs_reference_autograd():
    """d loss / d tables through TransH -> NegativeSampling(MarginLoss(1)) vs the reference's autograd."""
    from skghoi_b200.kge import TransH, MarginLoss, NegativeSampling
    g = load_golden("transh_ops")
    ent, rel, nrm = _tables()
    m = TransH(80, 117, dim=50, p_norm=2, norm_flag=True).to(DEV)
    with torch.no_grad():
        m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
    score = m(g["head"].to(DEV), g["relation"].to(DEV), g["tail"].to(DEV))[4]
    loss = NegativeSampling(loss=MarginLoss(margin=1), batch_size=256).to(DEV)(score)
    loss.backward()
    torch.testing.assert_close(loss.detach().cpu(), g["grad_loss"], rtol=1e-5, atol=1e-6)
    for got, key in ((m.ent_embeddings.weight.grad, "grad_ent"), (m.rel_embeddings.weight.grad, "grad_rel"),
                     (m.norm_vector.weight.grad, "grad_norm")):
        ref = g[key]
        assert float((got.cpu() - ref).abs().max()) <= 1e-4 * max(1.0, float(ref.abs().max())), key


def test_focal_loss_forward_backward():
    from skghoi_b200 import ops
    g = load_golden("spatial_ops")
    x, y = g["focal_x"].to(DEV), g["focal_y"].to(DEV)
    torch.testing.assert_close(ops.binary_focal_loss(x, y, reduction="sum", gamma=0.2).cpu(), g["focal_g02_sum"],
                               rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(ops.binary_focal_loss(x, y, reduction="sum", gamma=2.0).cpu(), g["focal_g20_sum"],
                               rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(ops.binary_focal_loss(x, y).cpu(), g["focal_g20_mean"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ops.binary_focal_loss(x, y, reduction="none").cpu(), g["focal_g20_none"],
                               rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        ops.binary_focal_loss(x, y, reduction="bogus")
    for gamma in (0.2, 2.0):
        xr = g["focal_x"].clone().requires_grad_(True)
        O.binary_focal_loss(xr, g["focal_y"], reduction="sum", gamma=gamma).backward()
        xg = g["focal_x"].to(DEV).requires_grad_(True)
        ops.binary_focal_loss(xg, y, reduction="sum", gamma=gamma).backward()
        torch.testing.assert_close(xg.grad.cpu(), xr.grad, rtol=1e-4, atol=1e-5)


def test_spatial_large_roundtrip_properties():
    """Full-size property check (2^17 cells): identical boxes give IoU 1 and zero offsets; the
    log half equals log(first half + eps) recomputed on the host."""
    from skghoi_b200.engine import enumerate_pairs
    shapes = [(16, 32)] * 256
    gen = torch.Generator().manual_seed(11)
    img_shapes = [(800, 1216)] * len(shapes)
    boxes, _, _ = synthetic.make_detections(shapes, img_shapes, gen)
    out = enumerate_pairs([b.to(DEV) for b in boxes], [16] * len(shapes), img_shapes)
    enc = out["enc"].cpu()
    assert enc.shape == (256 * 512, 46)
    diag = (out["x"] == out["y"]).cpu()
    assert int(diag.sum()) == 256 * 16
    assert torch.allclose(enc[diag][:, 18], torch.ones(int(diag.sum())))
    assert float(enc[diag][:, 19:23].abs().max()) == 0.0
    torch.testing.assert_close(enc[:, 23:], torch.log(enc[:, :23] + 1e-10), rtol=1e-5, atol=1e-5)
