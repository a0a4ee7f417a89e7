"""Kernel-level parity of the backward-pass kernels (csrc/trainops.cu) and the ReLU-gate epilogue of the GEMM, each
through its C-ABI entry point against plain PyTorch autograd of the formula it implements (reference: autograd through
heads/adamixer_transH_spatial_r50_head.py:431-530, 896-925 and ops.py:159-211)."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

from helpers import synthetic
from oracle import head_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _lib():
    from skghoi_b200 import _lib
    return _lib


def _segments(shapes):
    from skghoi_b200 import engine
    return engine.Segments([h for h, _ in shapes], [n for _, n in shapes], [(800, 1216)] * len(shapes), torch.device(DEV))


def _maps(seg):
    """hrow / orow / img of every cell (row-major human x box grid per image)."""
    hr, orow, img = [], [], []
    for b in range(seg.B):
        nh, n = int(seg.n_h[b]), int(seg.n[b])
        for i in range(nh):
            for j in range(n):
                hr.append(int(seg.hum_off[b]) + i); orow.append(int(seg.box_off[b]) + j); img.append(b)
    t = lambda v: torch.tensor(v, dtype=torch.int64, device=DEV)   # noqa: E731
    return t(hr), t(orow), t(img)


@pytest.mark.parametrize("rows,cols,use_map", [(200, 130, False), (64, 64, False), (1000, 1024, True), (7, 8, False)])
def test_transpose_bf16_with_colsum_and_row_map(rows, cols, use_map):
    L = _lib()
    g = torch.Generator(device=DEV).manual_seed(1)
    src_rows = rows + 13
    x = torch.randn(src_rows, cols + 6, device=DEV, generator=g).to(torch.bfloat16)
    xin = x[:, :cols]
    rm = None
    if use_map:
        rm = torch.randint(0, src_rows, (rows,), device=DEV, generator=g).to(torch.int32)
        rm[::17] = -1
    ld_out = (rows + 63) // 64 * 64
    out = torch.full((cols, ld_out), 7.0, dtype=torch.bfloat16, device=DEV)
    cs = torch.zeros(cols, dtype=torch.float32, device=DEV)
    L.call("skg_transpose_bf16", L.ptr(xin), xin.stride(0), rows, cols, L.ptr(rm), L.ptr(out), ld_out, L.ptr(cs), L.stream_ptr(torch.device(DEV)))
    torch.cuda.synchronize()
    if use_map:
        ref = torch.where((rm >= 0)[:, None], xin[rm.clamp(min=0).long()], torch.zeros((), dtype=torch.bfloat16, device=DEV))
    else:
        ref = xin[:rows]
    assert torch.equal(out[:, :rows], ref.t())
    torch.testing.assert_close(cs, ref.float().sum(0), rtol=1e-4, atol=1e-3)


def test_gemm_relu_gate_epilogue_with_residual_and_scatter():
    """dX = gate(G W + residual) with the gate indexed by the GEMM row and the rows scattered through out_row."""
    L = _lib()
    g = torch.Generator(device=DEV).manual_seed(2)
    M, N, K = 300, 256, 512                      # contraction K, output width N
    G = torch.randn(M, K, device=DEV, generator=g).to(torch.bfloat16)
    WT = (torch.randn(N, K, device=DEV, generator=g) / K ** 0.5).to(torch.bfloat16)
    mask = torch.randn(M, N, device=DEV, generator=g).to(torch.bfloat16)
    mask[::3] = 0
    res = torch.randn(M, N, device=DEV, generator=g)
    perm = torch.randperm(M + 20, device=DEV, generator=g)[:M].to(torch.int32)
    out = torch.zeros(M + 20, N, dtype=torch.bfloat16, device=DEV)
    L.gemm_bf16(G, WT, [L.epilogue_group(mask_bf16=mask, residual=res, out_bf16=out, out_row=perm)])
    out2 = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    L.gemm_bf16(G, WT, [L.epilogue_group(mask_bf16=mask, out_bf16=out2)])
    torch.cuda.synchronize()
    prod = G.float() @ WT.float().t()
    ref = torch.where(mask.float() > 0, prod + res, torch.zeros((), device=DEV))
    torch.testing.assert_close(out[perm.long()].float(), ref, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(out2.float(), torch.where(mask.float() > 0, prod, torch.zeros((), device=DEV)), rtol=2e-2, atol=2e-2)
    assert bool((out2.float()[mask.float() <= 0] == 0).all())


def test_hadamard_backward_both_phases_vs_autograd():
    from skghoi_b200 import train
    L = _lib()
    shapes = [(3, 7), (2, 5), (4, 4)]
    seg = _segments(shapes)
    hr, orow, img = _maps(seg)
    C, NH, NB, B, R = seg.total_cells, seg.total_humans, seg.total_boxes, seg.B, 1024
    g = torch.Generator(device=DEV).manual_seed(3)
    rn = lambda *s: torch.randn(*s, device=DEV, generator=g)   # noqa: E731
    S = rn(C, 4 * R).to(torch.bfloat16)
    dP = [rn(C, R).to(torch.bfloat16) for _ in range(5)]       # att, fin, os, so, g
    hproj, oproj, U2h, U2o, GT = rn(NH, 2 * R), rn(NB, 2 * R), rn(NH, R), rn(NB, R), rn(B, R)
    # autograd reference
    leaves = [t.clone().requires_grad_(True) for t in (hproj, oproj, U2h, U2o, GT)]
    Sf = S.float().clone().requires_grad_(True)
    hp, op, u2h, u2o, gt = leaves
    S_att, S_os, S_so, S_g = Sf.split(R, 1)
    pre = [S_att * (hp[hr, :R] + op[orow, :R]), S_att * (u2h[hr] + u2o[orow]), S_os * op[orow, R:], S_so * hp[hr, R:], S_g * gt[img]]
    sum((p * d.float()).sum() for p, d in zip(pre, dP)).backward()
    a = train._HadamardBwd()
    dS = torch.zeros(C, 4 * R, dtype=torch.bfloat16, device=DEV)
    dh, do_, du2h, du2o, dgt = (torch.zeros_like(t) for t in (hproj, oproj, U2h, U2o, GT))
    a.S, a.ld_s = S.data_ptr(), 4 * R
    a.dP_att, a.dP_fin, a.dP_os, a.dP_so, a.dP_g = (t.data_ptr() for t in dP)
    a.ld_dp = R
    a.hproj, a.oproj, a.U2h, a.U2o, a.GT = (t.data_ptr() for t in (hproj, oproj, U2h, U2o, GT))
    a.dS, a.ld_ds = dS.data_ptr(), 4 * R
    a.dhproj, a.doproj, a.dU2h, a.dU2o, a.dGT = (t.data_ptr() for t in (dh, do_, du2h, du2o, dgt))
    a.box_off, a.hum_off, a.cell_off = seg.d_box_off.data_ptr(), seg.d_hum_off.data_ptr(), seg.d_cell_off.data_ptr()
    a.num_images, a.R = B, R
    for phase in (1, 2):
        L.call("skg_hadamard_backward", ctypes.byref(a), NH, phase, L.stream_ptr(torch.device(DEV)))
    torch.cuda.synchronize()
    tol = dict(rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(dS.float(), Sf.grad, **tol)
    for got, leaf, name in ((dh, hp, "dhproj"), (do_, op, "doproj"), (du2h, u2h, "dU2h"), (du2o, u2o, "dU2o"), (dgt, gt, "dGT")):
        torch.testing.assert_close(got, leaf.grad, rtol=1e-3, atol=1e-2, msg=lambda m, n=name: n + ": " + m)


def test_message_aggregate_backward_vs_autograd():
    L = _lib()
    shapes = [(3, 7), (2, 5), (4, 4), (1, 2)]
    seg = _segments(shapes)
    hr, orow, img = _maps(seg)
    C, NH, NB, B, R = seg.total_cells, seg.total_humans, seg.total_boxes, seg.B, 1024
    g = torch.Generator(device=DEV).manual_seed(4)
    rn = lambda *s: torch.randn(*s, device=DEV, generator=g)   # noqa: E731
    adj = rn(C).requires_grad_(True)
    H_os = rn(C, R).relu().to(torch.bfloat16); H_so = rn(C, R).relu().to(torch.bfloat16)
    dZh = rn(NH, R).to(torch.bfloat16); dZo = rn(NB, R).to(torch.bfloat16)
    from train_ref import segment_softmax
    Hos_f = H_os.float().clone().requires_grad_(True); Hso_f = H_so.float().clone().requires_grad_(True)
    Zh = torch.zeros(NH, R, device=DEV).index_add(0, hr, segment_softmax(adj, hr, NH)[:, None] * Hos_f)
    Zo = torch.zeros(NB, R, device=DEV).index_add(0, orow, segment_softmax(adj, orow, NB)[:, None] * Hso_f)
    ((Zh * dZh.float()).sum() + (Zo * dZo.float()).sum()).backward()
    dP_os = torch.zeros(C, R, dtype=torch.bfloat16, device=DEV); dP_so = torch.zeros_like(dP_os)
    dah = torch.zeros(C, device=DEV); dao = torch.zeros(C, device=DEV)
    p = L.ptr
    L.call("skg_message_aggregate_backward", p(adj.detach()), p(dZh), p(dZo), p(H_os), p(H_so), R, p(seg.d_box_off), p(seg.d_hum_off),
           p(seg.d_cell_off), B, NB, NH, int(seg.max_n), R, p(dP_os), p(dP_so), R, p(dah), p(dao), L.stream_ptr(torch.device(DEV)))
    torch.cuda.synchronize()
    torch.testing.assert_close(dah + dao, adj.grad, rtol=1e-3, atol=1e-3)
    # the kernel hands back the gradient of the PRE-activation: gated by H > 0
    torch.testing.assert_close(dP_os.float(), Hos_f.grad * (H_os.float() > 0), rtol=2e-2, atol=1e-2)
    torch.testing.assert_close(dP_so.float(), Hso_f.grad * (H_so.float() > 0), rtol=2e-2, atol=1e-2)


def test_adjacency_layernorm_spatial_fc1_backward_vs_autograd():
    L = _lib()
    dev = torch.device(DEV)
    g = torch.Generator(device=DEV).manual_seed(5)
    rn = lambda *s: torch.randn(*s, device=DEV, generator=g)   # noqa: E731
    p = L.ptr
    # adjacency: adj = wts . w + b
    C, R = 333, 1024
    wts = rn(C, R).relu().to(torch.bfloat16)
    w = rn(R).requires_grad_(True)
    dah, dao = rn(C), rn(C)
    wf = wts.float().clone().requires_grad_(True)
    ((wf @ w) * (dah + dao)).sum().backward()
    dwts = torch.zeros(C, R, dtype=torch.bfloat16, device=DEV); dw = torch.zeros(R, device=DEV); db = torch.zeros(4, device=DEV)
    L.call("skg_adjacency_backward", p(dah), p(dao), p(wts), R, p(w.detach()), C, R, p(dwts), R, p(dw), p(db), L.stream_ptr(dev))
    torch.cuda.synchronize()
    torch.testing.assert_close(dwts.float(), wf.grad * (wts.float() > 0), rtol=2e-2, atol=1e-2)
    torch.testing.assert_close(dw, w.grad, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(db[0], (dah + dao).sum(), rtol=1e-4, atol=1e-3)
    # layernorm on base + relu(msg)
    NH, NB = 9, 21
    base = rn(NH + NB, R); msg = rn(NH + NB, R)
    x = (base + msg.relu()).detach()
    wa, wb = (rn(R) * 0.1 + 1).requires_grad_(True), (rn(R) * 0.1 + 1).requires_grad_(True)
    ba, bb = rn(R).requires_grad_(True), rn(R).requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    y = torch.cat([F.layer_norm(xr[:NH], (R,), wa, ba, 1e-5), F.layer_norm(xr[NH:], (R,), wb, bb, 1e-5)])
    dy = rn(NH + NB, R)
    (y * dy).sum().backward()
    dx = torch.zeros_like(x); dmsg = torch.zeros(NH + NB, R, dtype=torch.bfloat16, device=DEV)
    gw = [torch.zeros(R, device=DEV) for _ in range(4)]
    L.call("skg_layernorm_backward", p(x), p(dy), p(base[:NH].contiguous()), p(base[NH:].contiguous()), NH + NB, NH, R, p(wa.detach()),
           p(wb.detach()), 1e-5, p(dx), p(dmsg), p(gw[0]), p(gw[1]), p(gw[2]), p(gw[3]), L.stream_ptr(dev))
    torch.cuda.synchronize()
    torch.testing.assert_close(dx, xr.grad, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(dmsg.float(), xr.grad * (msg > 0), rtol=2e-2, atol=1e-2)
    for got, ref in zip(gw, (wa.grad, ba.grad, wb.grad, bb.grad)):
        torch.testing.assert_close(got, ref, rtol=1e-3, atol=1e-3)
    # spatial fc1
    cells = 1000
    enc = rn(cells, 46) * 5
    gg = rn(cells, 128).to(torch.bfloat16)
    dW = torch.zeros(128, 46, device=DEV); dbb = torch.zeros(128, device=DEV)
    L.call("skg_spatial_fc1_backward", p(enc), p(gg), cells, 46, 128, p(dW), p(dbb), L.stream_ptr(dev))
    torch.cuda.synchronize()
    torch.testing.assert_close(dW, gg.float().t() @ enc, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dbb, gg.float().sum(0), rtol=1e-3, atol=1e-2)


def test_head_loss_backward_vs_autograd():
    """d(hoi + 0.5 * inter)/d logits from the kernel against autograd through sigmoid / priors / binary_focal_loss."""
    from skghoi_b200 import engine
    L = _lib()
    dev = torch.device(DEV)
    shapes = [(3, 7), (2, 5)]
    batch = synthetic.make_batch(shapes, seed=3, with_targets=True)
    seg = _segments(shapes)
    boxes = torch.cat(batch["box_coords"]).to(DEV)
    pe = engine.enumerate_pairs(None, [h for h, _ in shapes], batch["image_shapes"], seg=seg, boxes_cat=boxes)
    Pn, C, nc = seg.total_pairs, seg.total_cells, 117
    labels_dev = torch.cat(batch["box_labels"]).to(DEV); scores_dev = torch.cat(batch["box_scores"]).to(DEV)
    otv = synthetic.hico_object_to_verb()
    vt = engine.VerbTable(otv, dev)
    t = L.PostTables()
    t.verb_off, t.verb_idx = vt.d_off.data_ptr(), vt.d_idx.data_ptr()
    t.pair2cell, t.cell_hrow, t.cell_orow, t.cell_img = (pe[k].data_ptr() for k in ("pair2cell", "cell_hrow", "cell_orow", "cell_img"))
    t.hum_off, t.box_off, t.pair_off = seg.d_hum_off.data_ptr(), seg.d_box_off.data_ptr(), seg.d_pair_off.data_ptr()
    t.labels, t.scores = labels_dev.data_ptr(), scores_dev.data_ptr()
    g = torch.Generator(device=DEV).manual_seed(6)
    logits = torch.randn(C, nc + 1, device=DEV, generator=g)
    gt = (torch.rand(Pn, nc, device=DEV, generator=g) < 0.02).float()
    n_p = torch.tensor([7.0, 3.0], device=DEV)
    g_hoi = torch.tensor([1.0], device=DEV); g_int = torch.tensor([0.5], device=DEV)
    dl = torch.zeros(C, 128, dtype=torch.bfloat16, device=DEV)
    L.call("skg_head_loss_backward", L.ptr(logits), nc + 1, nc, ctypes.byref(t), L.ptr(pe["pair2cell"]), Pn, 1.0, L.ptr(gt), L.ptr(n_p),
           L.ptr(g_hoi), L.ptr(g_int), L.ptr(dl), 128, L.stream_ptr(dev))
    torch.cuda.synchronize()
    # autograd reference on the pair rows
    p2c = pe["pair2cell"].long()
    lg = logits.clone().requires_grad_(True)
    lp = lg[p2c]
    w = torch.sigmoid(lp[:, nc]); s = torch.sigmoid(lp[:, :nc])
    xk, yk = pe["x_keep"], pe["y_keep"]
    pair_img = pe["cell_img"].long()[p2c]
    boff = seg.d_box_off.long()[pair_img]
    sh = scores_dev[boff + xk]; so = scores_dev[boff + yk]
    valid = torch.zeros(80, nc, device=DEV)
    for o, verbs in enumerate(otv):
        valid[o, verbs] = 1
    m = valid[labels_dev[boff + yk]] > 0
    x = (s * (sh * so)[:, None] * w.detach()[:, None])[m]
    hoi = O.binary_focal_loss(x, gt[m], reduction="sum", gamma=0.2) / n_p[0]
    inter = O.binary_focal_loss(w, gt.sum(1).clamp(max=1), reduction="sum", gamma=2.0) / n_p[1]
    (hoi + 0.5 * inter).backward()
    ref = lg.grad
    got = dl.float()[:, :nc + 1]
    scale = float(ref.abs().max())
    assert float((got - ref).abs().max()) <= 1e-2 * scale, (float((got - ref).abs().max()), scale)
    diag = torch.ones(C, dtype=torch.bool, device=DEV); diag[p2c] = False
    assert float(got[diag].abs().max()) == 0.0          # cells i == j carry no gradient


def test_select_negatives_cuda_counts_validity_and_randomness():
    """skg_select_count / skg_select_fill: positives in flat order image by image, per image as many DISTINCT zero-label
    entries of the same image, a new draw per call (head.py:926-963 in its device form)."""
    from skghoi_b200 import engine
    torch.manual_seed(0)
    nc = 117
    per_img = [10, 25, 0, 5, 300]
    P = sum(per_img)
    pair_off = torch.tensor([0] + list(torch.tensor(per_img).cumsum(0)), dtype=torch.int64, device=DEV)
    labels = (torch.rand(P, nc) > 0.97).float()
    labels[35:40] = 0; labels[36, 5] = 1
    lab = labels.to(DEV)
    pos, neg = engine.select_negatives_cuda(lab, pair_off, len(per_img))
    flat = labels.view(-1)
    assert torch.equal(pos.cpu(), torch.nonzero(flat).squeeze(1))
    negc = neg.cpu()
    assert bool((flat[negc] == 0).all()) and negc.unique().numel() == negc.numel()
    img_of = torch.repeat_interleave(torch.arange(len(per_img)), torch.tensor(per_img))
    assert torch.equal(img_of[pos.cpu() // nc], img_of[negc // nc])            # negative i comes from the image of positive i
    _, neg2 = engine.select_negatives_cuda(lab, pair_off, len(per_img))
    assert neg2.numel() == neg.numel() and not torch.equal(neg2.cpu().sort().values, negc.sort().values)
    # uniformity over the zero labels of a small image: 2 positives among 12 entries, many draws
    small = torch.zeros(2, 6); small[0, 1] = 1; small[1, 4] = 1
    sm = small.to(DEV); off = torch.tensor([0, 2], dtype=torch.int64, device=DEV)
    hits = torch.zeros(12)
    for _ in range(600):
        _, n = engine.select_negatives_cuda(sm, off, 1)
        hits[n.cpu()] += 1
    assert hits[1] == 0 and hits[10] == 0
    z = hits[[0, 2, 3, 4, 5, 6, 7, 8, 9, 11]]
    assert float(z.min()) > 0.6 * float(z.mean()) and float(z.max()) < 1.4 * float(z.mean()), z
    empty = torch.zeros(6, nc, device=DEV)
    p0, n0 = engine.select_negatives_cuda(empty, torch.tensor([0, 6], dtype=torch.int64, device=DEV), 1)
    assert p0.numel() == 0 and n0.numel() == 0
