"""Differentiable (training) pass of the interaction head: BASELINE configs[2], SURVEY 8(b) "fwd/bwd", 8(e).

Every dense contraction of the head -- forward, input gradient and weight gradient -- runs on the hand-written
tcgen05 GEMM (``skg_gemm_bf16``, csrc/gemm.cu) through ``LinearFn``; pair enumeration + spatial encoding, GT
association, priors, post-processing indices, the focal losses and the TransH scores / margin loss are the same
CUDA kernels the inference path uses.  The element-wise glue between the GEMMs (Hadamard products, ReLU masks,
segment softmax, LayerNorm) is expressed with ATen ops so that autograd provides its backward: PyTorch is the
plumbing here (device memory + autograd tape), the contractions are ours.  The forward uses the same algebraic
restructuring as the inference engine: 16-branch MBFs concatenated, ``attention_head.fc_1``
split per node, one fused ``fc_2`` GEMM for the four MBFs, ``fc_3`` of the messages applied after the weighted sum,
the two identical message-passing iterations computed once (SURVEY defect 5: only the last one reaches the loss).

Reference: GraphHead.forward head.py:769-993, InteractionHead.forward head.py:389-427, losses head.py:153-235.
"""
from typing import List, Optional

import torch
import torch.nn.functional as F
from torch import Tensor

from . import _lib


def _ceil(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def _operand(t: Tensor) -> Tensor:
    """[rows, K] -> bf16, K contiguous, row pitch a multiple of 8 elements (TMA: 16-byte pitch), 16-byte aligned."""
    rows, k = t.shape
    kp = _ceil(k, 8)
    if kp == k:
        return t.to(torch.bfloat16).contiguous()
    out = torch.zeros(rows, kp, dtype=torch.bfloat16, device=t.device)
    out[:, :k] = t
    return out


def _operand_t(t: Tensor) -> Tensor:
    """[rows, K] -> its transpose [K, rows] as a bf16 operand (rows become the contraction dimension)."""
    rows, k = t.shape
    rp = _ceil(rows, 8)
    if rp == rows:
        return t.to(torch.bfloat16).t().contiguous()
    out = torch.zeros(k, rp, dtype=torch.bfloat16, device=t.device)
    out[:, :rows] = t.t()
    return out


def gemm_nt(A: Tensor, B: Tensor, K: int, bias: Optional[Tensor] = None, relu: bool = False) -> Tensor:
    """fp32 [M, N] = A[M, :K] @ B[N, :K]^T (+ bias) (ReLU) on skg_gemm_bf16; A, B bf16 operands from ``_operand``."""
    M, N = A.shape[0], B.shape[0]
    np_ = _ceil(N, 8)
    out = torch.empty(M, np_, dtype=torch.float32, device=A.device)
    if M == 0:
        return out[:, :N]
    bp = None
    if bias is not None:
        bp = torch.zeros(_ceil(N, 256), dtype=torch.float32, device=A.device)
        bp[:N] = bias
    g = _lib.epilogue_group(relu=relu, bias=bp, out_f32=out)
    _lib.gemm_bf16(A, B, [g], M=M, N=N, K=K)
    return out if np_ == N else out[:, :N]


class LinearFn(torch.autograd.Function):
    """y = x W^T + b (optionally ReLU) with all three GEMMs (y, dx, dW) on the tcgen05 kernel; bf16 operands,
    fp32 accumulation, fp32 results; the bias gradient is a column sum."""

    @staticmethod
    def forward(ctx, x: Tensor, W: Tensor, b: Optional[Tensor], relu: bool):
        _lib.require_cuda(x, W)
        K = x.shape[1]
        y = gemm_nt(_operand(x), _operand(W), K, b, relu)
        ctx.save_for_backward(x, W, y if relu else None)
        ctx.has_bias = b is not None
        ctx.relu = relu
        return y

    @staticmethod
    def backward(ctx, g: Tensor):
        x, W, y = ctx.saved_tensors
        if ctx.relu:
            g = g * (y > 0)
        M, K = x.shape
        N = W.shape[0]
        gx = gW = gb = None
        if ctx.needs_input_grad[0]:
            gx = gemm_nt(_operand(g), _operand_t(W), N)                 # [M, K] = g[M, N] @ W[N, K]
            gx = gx if gx.is_contiguous() else gx.contiguous()
        if ctx.needs_input_grad[1]:
            gW = gemm_nt(_operand_t(g), _operand_t(x), M)               # [N, K] = g^T[N, M] @ x[M, K]
            gW = gW if gW.is_contiguous() else gW.contiguous()
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g.sum(0)
        return gx, gW, gb, None


def linear(x: Tensor, W: Tensor, b: Optional[Tensor] = None, relu: bool = False) -> Tensor:
    return LinearFn.apply(x, W, b, relu)


def segment_softmax(v: Tensor, seg: Tensor, num_segments: int) -> Tensor:
    """softmax of ``v`` inside each segment id (head.py:903, 914: adjacency rows / columns of one image)."""
    m = torch.full((num_segments,), float("-inf"), dtype=v.dtype, device=v.device).scatter_reduce(0, seg, v.detach(), "amax")
    e = torch.exp(v - m[seg])
    s = torch.zeros(num_segments, dtype=v.dtype, device=v.device).index_add(0, seg, e)
    return e / s[seg]


def head_logits(ih, features3: Tensor, box_features: Tensor, pe: dict) -> Tensor:
    """Differentiable logits [pairs, num_cls + 1] (predictor | suppressor) from the pooled box features.
    ``pe`` = engine.enumerate_pairs(...) of the batch (encodings + row maps, no gradient)."""
    G = ih.box_pair_head
    seg = pe["seg"]
    dev = box_features.device
    R = G.representation_size
    hrow, orow, img = pe["cell_hrow"].long(), pe["cell_orow"].long(), pe["cell_img"].long()
    p2c = pe["pair2cell"].long()
    NH, NB = seg.total_humans, seg.total_boxes
    # node rows: box_head on every box, entity rows appended (head.py:812, 875-885)
    node = linear(linear(box_features.flatten(1), G.box_head[1].weight, G.box_head[1].bias, True),
                  G.box_head[3].weight, G.box_head[3].bias, True)
    ent = G.transh_head.transh.ent_embeddings.weight.detach()
    box_local = torch.cat([torch.arange(int(k)) for k in seg.n]).to(dev)                   # box index inside its image
    human_rows = torch.cat([torch.arange(int(o), int(o) + int(k)) for o, k in zip(seg.box_off[:-1], seg.n_h)]).to(dev)
    hN = linear(torch.cat([node[human_rows], ent[G.human_idx].expand(NH, -1)], 1), G.fc_head[0].weight, G.fc_head[0].bias, True)
    oN = linear(torch.cat([node, ent[box_local]], 1), G.fc_tail[0].weight, G.fc_tail[0].bias, True)
    # spatial head (head.py:662-669); the 46 -> 128 layer stays fp32 (inputs span [-23, 1e3]) like the inference kernel
    sh = G.spatial_head
    sp = F.relu(F.linear(pe["enc"], sh[0].weight, sh[0].bias))
    sp = linear(linear(sp, sh[2].weight, sh[2].bias, True), sh[4].weight, sh[4].bias, True)
    a_W1, a_b1, a_W2, a_b2, a_W3, a_b3 = G.attention_head.packed()
    o_W1, o_b1, o_W2, o_b2, o_W3, o_b3 = G.obj_to_sub.packed()
    s_W1, s_b1, s_W2, s_b2, s_W3, s_b3 = G.sub_to_obj.packed()
    g_W1, g_b1, g_W2, g_b2, g_W3, g_b3 = G.attention_head_g.packed()
    # one fc_2 GEMM for the four MBFs that read the spatial features
    S = linear(sp, torch.cat([a_W2, o_W2, s_W2, g_W2]), torch.cat([a_b2, o_b2, s_b2, g_b2]))
    S_att, S_os, S_so, S_g = S.split(R, dim=1)
    W1h, W1o = a_W1[:, :R], a_W1[:, R:]

    def att_tables(h, o):                                                                  # fc_1 of cat(h[x], o[y]) per node
        return linear(h, W1h), linear(o, W1o, a_b1)
    Uh, Uo = att_tables(hN, oN)
    wts = linear(F.relu(S_att * (Uh[hrow] + Uo[orow])), a_W3, a_b3, True)                  # head.py:896
    adj = F.linear(wts, G.adjacency.weight, G.adjacency.bias).squeeze(1)                   # head.py:897
    # messages (head.py:900-925): fc_3 commutes with the softmax-weighted sum (weights sum to 1)
    H_os = F.relu(S_os * linear(oN, o_W1, o_b1)[orow])
    H_so = F.relu(S_so * linear(hN, s_W1, s_b1)[hrow])
    Zh = torch.zeros(NH, R, device=dev).index_add(0, hrow, segment_softmax(adj, hrow, NH)[:, None] * H_os)
    Zo = torch.zeros(NB, R, device=dev).index_add(0, orow, segment_softmax(adj, orow, NB)[:, None] * H_so)
    h_enc = F.layer_norm(hN + linear(Zh, o_W3, o_b3, True), (R,), G.norm_h.weight, G.norm_h.bias, G.norm_h.eps)
    o_enc = F.layer_norm(oN + linear(Zo, s_W3, s_b3, True), (R,), G.norm_o.weight, G.norm_o.bias, G.norm_o.eps)
    # final fusion on the valid pairs (head.py:966-973)
    Uh2, Uo2 = att_tables(h_enc, o_enc)
    f1 = linear(F.relu(S_att[p2c] * (Uh2[hrow[p2c]] + Uo2[orow[p2c]])), a_W3, a_b3, True)
    glob = features3.mean(dim=(2, 3))
    f2 = linear(F.relu(S_g[p2c] * linear(glob, g_W1, g_b1)[img[p2c]]), g_W3, g_b3, True)
    feat = torch.cat([f1, f2], 1)
    Wl = torch.cat([ih.box_pair_predictor.weight, ih.box_pair_suppressor.weight])
    bl = torch.cat([ih.box_pair_predictor.bias, ih.box_pair_suppressor.bias])
    return linear(feat, Wl, bl)                                                            # head.py:410-411


def training_forward(ih, features, image_shapes, box_features: Tensor, box_coords: List[Tensor], box_labels, box_scores,
                     targets, meta=None) -> List[dict]:
    """InteractionHead.forward in training mode from the pooled box features on (head.py:389-427): the per-image
    result dicts plus the loss dict, with ``hoi_loss`` / ``interactiveness_loss`` differentiable w.r.t. every head
    parameter and ``transH_loss`` w.r.t. the TransH tables."""
    from . import engine, ops
    assert targets is not None, "Targets should be passed during training"                 # head.py:380-381
    G = ih.box_pair_head
    meta, boxes_cat, labels_dev, scores_dev = engine._prep_inputs(G, features, box_features, box_coords, box_labels,
                                                                  box_scores, meta)
    n, n_h = meta["n"], meta["n_h"]
    if any(k > G.transh_head.num_object for k in n):
        raise IndexError("index out of range in self")                                     # head.py:572
    n_h_eff = [0 if (h == 0 or k <= 1) else h for h, k in zip(n_h, n)]
    dev = box_features.device
    # pair enumeration + spatial encoding + row maps: the same kernel as inference, no tape (boxes carry no gradient)
    pe = engine.enumerate_pairs(None, n_h_eff, image_shapes, seg=engine.Segments(n_h_eff, n, image_shapes, dev),
                                boxes_cat=boxes_cat)
    seg = pe["seg"]
    with torch.enable_grad():
        logits = head_logits(ih, features["3"], box_features, pe)
    # post-processing indices, priors, GT labels and the TransH loss from the native kernels, fed with the taped logits
    lg = logits.detach()
    run = dict(pe)
    run.update(logits=lg if lg.is_contiguous() else lg.contiguous(),
               post_i32=torch.empty(2 * max(seg.total_boxes, 1), dtype=torch.int32, device=dev),
               post_i64=torch.empty(max(seg.total_humans, 1) + 2 * seg.B + 1, dtype=torch.int64, device=dev))
    was_training = ih.training
    ih.train()
    try:
        with torch.enable_grad():           # the TransH loss is differentiable w.r.t. the tables (kge autograd Functions)
            results = engine.interaction_head_from_pooled(ih, features, image_shapes, box_features.detach(), box_coords,
                                                          box_labels, box_scores, targets=targets, meta=meta, run=run)
    finally:
        ih.train(was_training)
    with torch.enable_grad():
        native_losses = results.pop()
        nc = G.num_cls
        dev = logits.device
        gx = torch.cat([r["index"] + int(o) for r, o in zip(results, seg.pair_off[:-1])])
        gy = torch.cat([r["prediction"] for r in results])
        prior = torch.cat([r["prior"] for r in results], 1)
        lab = torch.cat([r["labels"] for r in results]); unary = torch.cat([r["unary_labels"] for r in results])
        weights = torch.sigmoid(logits[:, nc])
        scores = torch.sigmoid(logits[:, :nc])[gx, gy] * prior.prod(0) * weights[gx].detach()   # head.py:316
        n_p = torch.stack([(lab != 0).sum(), (unary != 0).sum()]).float()
        if ih.distributed:
            from .shard import average_normalisers
            n_p = average_normalisers(n_p)
        hoi = ops.binary_focal_loss(scores, lab, reduction="sum", gamma=0.2) / n_p[0]      # head.py:153-177
        inter = ops.binary_focal_loss(weights, unary, reduction="sum", gamma=2.0) / n_p[1]  # head.py:180-205
    ih.__dict__["_skg_last_train"] = dict(logits=logits, scores=scores, weights=weights)
    results.append(dict(hoi_loss=hoi, interactiveness_loss=inter, transH_loss=native_losses["transH_loss"]))
    return results
