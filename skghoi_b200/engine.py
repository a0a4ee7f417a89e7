"""Host-side orchestration of the interaction-head hot path on libskghoi_b200.so.

``Segments`` is the per-batch segment table (image -> box / human / cell / pair offsets) every
batched kernel consumes; ``enumerate_pairs`` is the pair-enumeration + spatial-encoding call
(reference head.py:847-868).  The GEMM pipeline (``HeadEngine``) lives further down.
CUDA only -- nothing here computes on the CPU beyond building the few-hundred-byte segment table.
"""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch import Tensor

from . import _lib


class Segments:
    """Per-batch segment table.  For image b with n boxes (n_h humans first):
    cells N2 = n_h*n (row-major human x box grid), valid pairs P = n_h*(n-1)."""

    def __init__(self, n_h: Sequence[int], n: Sequence[int], image_shapes: Sequence[Tuple[int, int]], device):
        B = len(n)
        assert len(n_h) == B and len(image_shapes) == B
        nh = np.asarray(n_h, dtype=np.int64); nn_ = np.asarray(n, dtype=np.int64)
        if B and (np.any(nh > nn_) or np.any(nh < 0)):
            raise ValueError("n_h must satisfy 0 <= n_h <= n")
        self.B = B
        self.n_h = nh; self.n = nn_
        self.box_off = np.concatenate([[0], np.cumsum(nn_)]).astype(np.int64)
        self.hum_off = np.concatenate([[0], np.cumsum(nh)]).astype(np.int64)
        self.cell_off = np.concatenate([[0], np.cumsum(nh * nn_)]).astype(np.int64)
        self.pair_off = np.concatenate([[0], np.cumsum(nh * np.maximum(nn_ - 1, 0))]).astype(np.int64)
        self.total_boxes = int(self.box_off[-1]); self.total_humans = int(self.hum_off[-1])
        self.total_cells = int(self.cell_off[-1]); self.total_pairs = int(self.pair_off[-1])
        if self.total_cells >= 2 ** 31:
            raise ValueError("too many grid cells in one batch")
        hw = np.asarray([[float(h), float(w), float(h * w)] for h, w in image_shapes], dtype=np.float32).reshape(B, 3)
        # one packed host buffer -> one H2D copy
        i32 = np.concatenate([self.box_off, nh, self.hum_off]).astype(np.int32)
        i64 = np.concatenate([self.cell_off, self.pair_off]).astype(np.int64)
        raw = np.concatenate([i64.view(np.uint8), i32.view(np.uint8), hw.reshape(-1).view(np.uint8)])
        pad = (-len(raw)) % 16
        raw = np.concatenate([raw, np.zeros(pad, dtype=np.uint8)])
        self.host = torch.from_numpy(raw)
        self.dev = self.host.to(device, non_blocking=False)
        o = 0
        self.d_cell_off = self.dev[o:o + 8 * (B + 1)].view(torch.int64); o += 8 * (B + 1)
        self.d_pair_off = self.dev[o:o + 8 * (B + 1)].view(torch.int64); o += 8 * (B + 1)
        self.d_box_off = self.dev[o:o + 4 * (B + 1)].view(torch.int32); o += 4 * (B + 1)
        self.d_n_h = self.dev[o:o + 4 * B].view(torch.int32); o += 4 * B
        self.d_hum_off = self.dev[o:o + 4 * (B + 1)].view(torch.int32); o += 4 * (B + 1)
        self.d_hw = self.dev[o:o + 12 * B].view(torch.float32)
        self.device = self.dev.device


def enumerate_pairs(box_coords: List[Tensor], n_h: Sequence[int], image_shapes: Sequence[Tuple[int, int]],
                    num_features: int = 23, eps: float = 1e-10, nan_guard: bool = True,
                    want_indices: bool = True, seg: Optional[Segments] = None,
                    boxes_cat: Optional[Tensor] = None) -> Dict[str, Tensor]:
    """Pair enumeration + spatial encoding for a whole batch in one launch (head.py:847-868).

    Returns enc [cells, 2F] fp32, x / y [cells] int64, x_keep / y_keep [pairs] int64 (per-image local
    indices, concatenated over images in order), cell2pair, pair2cell, cell_hrow, cell_orow,
    cell_img (int32 row maps for the fused GEMM epilogues) and nan_flag [B]."""
    dev = box_coords[0].device if boxes_cat is None else boxes_cat.device
    _lib.require_cuda(*(box_coords if boxes_cat is None else [boxes_cat]))
    if seg is None:
        seg = Segments(n_h, [len(b) for b in box_coords], image_shapes, dev)
    if boxes_cat is None:
        boxes_cat = torch.cat(list(box_coords)).to(torch.float32).contiguous()
    C, Pn, F = seg.total_cells, seg.total_pairs, num_features
    out = dict(enc=torch.empty(C, 2 * F, dtype=torch.float32, device=dev),
               nan_flag=torch.zeros(max(seg.B, 1), dtype=torch.int32, device=dev), seg=seg)
    if want_indices:
        out.update(x=torch.empty(C, dtype=torch.int64, device=dev), y=torch.empty(C, dtype=torch.int64, device=dev),
                   x_keep=torch.empty(Pn, dtype=torch.int64, device=dev),
                   y_keep=torch.empty(Pn, dtype=torch.int64, device=dev),
                   cell2pair=torch.empty(C, dtype=torch.int32, device=dev),
                   pair2cell=torch.empty(Pn, dtype=torch.int32, device=dev),
                   cell_hrow=torch.empty(C, dtype=torch.int32, device=dev),
                   cell_orow=torch.empty(C, dtype=torch.int32, device=dev),
                   cell_img=torch.empty(C, dtype=torch.int32, device=dev))
    g = lambda k: _lib.ptr(out.get(k))  # noqa: E731
    with torch.cuda.device(dev):
        _lib.call("skg_pairs_spatial_encode", _lib.ptr(boxes_cat), _lib.ptr(seg.d_box_off), _lib.ptr(seg.d_n_h),
                  _lib.ptr(seg.d_cell_off), _lib.ptr(seg.d_pair_off), _lib.ptr(seg.d_hum_off), _lib.ptr(seg.d_hw),
                  seg.B, C, F, float(eps), int(nan_guard), g("enc"), g("x"), g("y"), g("x_keep"), g("y_keep"),
                  g("cell2pair"), g("cell_hrow"), g("cell_orow"), g("cell_img"), g("pair2cell"), g("nan_flag"),
                  _lib.stream_ptr(dev))
    return out


# =============================================================================================
# packed weights
# =============================================================================================
def _bf16(t: Tensor, pad_k: int = 0, pad_rows: int = 0) -> Tensor:
    """fp32 [rows, K] parameter -> contiguous bf16 [rows(+pad), K(+pad)] operand."""
    t = t.detach().float()
    rows, k = t.shape
    kp = k if pad_k == 0 else (k + pad_k - 1) // pad_k * pad_k
    rp = rows if pad_rows == 0 else (rows + pad_rows - 1) // pad_rows * pad_rows
    if kp != k or rp != rows:
        out = torch.zeros(rp, kp, dtype=torch.bfloat16, device=t.device)
        out[:rows, :k] = t.to(torch.bfloat16)
        return out
    return t.to(torch.bfloat16).contiguous()


def _f32(t: Tensor, pad: int = 0) -> Tensor:
    t = t.detach().float().contiguous()
    if pad and t.numel() % pad:
        out = torch.zeros((t.numel() + pad - 1) // pad * pad, dtype=torch.float32, device=t.device)
        out[:t.numel()] = t
        return out
    return t


def _versions(module) -> tuple:
    return tuple((p.data_ptr(), p._version) for p in module.parameters())


class PackedMBF:
    """Concatenated-branch bf16 operands of one MultiBranchFusion / MessageMBF (head.py:446-474)."""

    def __init__(self, mbf):
        W1, b1, W2, b2, W3, b3 = mbf.packed()
        self.W1, self.b1 = _bf16(W1, pad_k=8), _f32(b1)
        self.W2, self.b2 = _bf16(W2, pad_k=8), _f32(b2)
        self.W3, self.b3 = _bf16(W3), _f32(b3)
        self.R = W3.shape[0]
        self.key = _versions(mbf)


def _packed_mbf(mbf) -> PackedMBF:
    pk = mbf.__dict__.get("_skg_packed")
    if pk is None or pk.key != _versions(mbf):
        pk = PackedMBF(mbf)
        mbf.__dict__["_skg_packed"] = pk
    return pk


def _to_bf16_operand(x: Tensor) -> Tensor:
    """Activation [M, K] (fp32 or bf16) -> bf16 [M, K8] operand through the cast kernel."""
    M, K = x.shape
    Kp = (K + 7) // 8 * 8
    if x.dtype == torch.bfloat16 and Kp == K and x.stride(1) == 1 and x.stride(0) % 8 == 0:
        return x
    xf = x.float().contiguous()
    out = torch.empty(M, Kp, dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.call("skg_cast_pad_bf16", _lib.ptr(xf), xf.stride(0), None, 0, M, K, _lib.ptr(out), Kp, Kp,
                  _lib.stream_ptr(x.device))
    return out


def _mbf_apply(mbf, appearance: Tensor, spatial: Tensor, row_p: Optional[Tensor], rows: int) -> Tensor:
    """relu?(W3 relu((W1 app + b1)[row_p] * (W2 sp + b2)) + b3) for ``rows`` spatial rows."""
    _lib.require_cuda(appearance, spatial)
    pk = _packed_mbf(mbf)
    dev = spatial.device
    R = pk.R
    if R % 128 or pk.W2.shape[1] % 8:
        raise _lib.SkgError("representation_size must be a multiple of 128 for the tensor-core path")
    a16 = _to_bf16_operand(appearance)
    s16 = _to_bf16_operand(spatial)
    a1 = torch.empty(a16.shape[0], R, dtype=torch.float32, device=dev)
    _lib.gemm_bf16(a16, pk.W1, [_lib.epilogue_group(bias=pk.b1, out_f32=a1)], K=pk.W1.shape[1])
    H = torch.empty(rows, R, dtype=torch.bfloat16, device=dev)
    _lib.gemm_bf16(s16, pk.W2, [_lib.epilogue_group(relu=True, bias=pk.b2, mul_p=a1, row_p=row_p, out_bf16=H)],
                   K=pk.W2.shape[1])
    out = torch.empty(rows, R, dtype=torch.float32, device=dev)
    _lib.gemm_bf16(H, pk.W3, [_lib.epilogue_group(relu=mbf._outer_relu, bias=pk.b3, out_f32=out)])
    return out


def mbf_forward(mbf, appearance: Tensor, spatial: Tensor) -> Tensor:
    """MultiBranchFusion.forward (head.py:469-474); appearance may have 1 row (broadcast, head.py:971)."""
    rows = spatial.shape[0]
    if appearance.shape[0] == rows:
        row_p = None
    elif appearance.shape[0] == 1:
        row_p = torch.zeros(rows, dtype=torch.int32, device=spatial.device)
    else:
        raise RuntimeError("The size of tensor a (%d) must match the size of tensor b (%d) at non-singleton dimension 0"
                           % (appearance.shape[0], rows))
    return _mbf_apply(mbf, appearance, spatial, row_p, rows)


def message_mbf_forward(mbf, appearance: Tensor, spatial: Tensor) -> Tensor:
    """MessageMBF.forward (head.py:509-530): spatial [n_h, n, S]; 'object' -> [n_h, n, R], 'human' -> [n, n_h, R]."""
    n_h, n = spatial.shape[:2]
    dev = spatial.device
    flat = spatial.reshape(n_h * n, -1)
    ar = torch.arange(n_h * n, device=dev, dtype=torch.int32)
    if mbf.node_type == "human":
        assert len(appearance) == n_h, "Incorrect size of dim0 for appearance features"     # head.py:511
        out = _mbf_apply(mbf, appearance, flat, (ar // n).int(), n_h * n)
        return out.reshape(n_h, n, -1).permute(1, 0, 2).contiguous()
    assert len(appearance) == n, "Incorrect size of dim0 for appearance features"           # head.py:521
    out = _mbf_apply(mbf, appearance, flat, (ar % n).int(), n_h * n)
    return out.reshape(n_h, n, -1)


class PackedHead:
    """All operands of the interaction head, packed once per parameter version."""

    def __init__(self, ih):
        G = ih.box_pair_head
        dev = next(ih.parameters()).device
        if dev.type != "cuda":
            raise _lib.SkgError("the interaction head must live on a CUDA device (no CPU fallback)")
        self.device = dev
        self.key = _versions(ih)
        R = G.representation_size; E = G.node_encoding_size
        self.R, self.E = R, E
        self.Wb1, self.bb1 = _bf16(G.box_head[1].weight, pad_k=64), _f32(G.box_head[1].bias)
        self.Wb2, self.bb2 = _bf16(G.box_head[3].weight), _f32(G.box_head[3].bias)
        self.Wsp0, self.bsp0 = _f32(G.spatial_head[0].weight), _f32(G.spatial_head[0].bias)
        self.Wsp1, self.bsp1 = _bf16(G.spatial_head[2].weight), _f32(G.spatial_head[2].bias)
        self.Wsp2, self.bsp2 = _bf16(G.spatial_head[4].weight), _f32(G.spatial_head[4].bias)
        self.Wfh, self.bfh = _bf16(G.fc_head[0].weight, pad_k=64), _f32(G.fc_head[0].bias)
        self.Wft, self.bft = _bf16(G.fc_tail[0].weight, pad_k=64), _f32(G.fc_tail[0].bias)
        self.fc_k = self.Wfh.shape[1]
        att = G.attention_head.packed(); os_ = G.obj_to_sub.packed(); so = G.sub_to_obj.packed()
        gg = G.attention_head_g.packed()
        z = torch.zeros(R, device=dev)
        # human nodes -> [U_h | V_so], object nodes -> [U_o | V_os]   (W1 of attention_head split as [W1h | W1o])
        self.W_hproj = _bf16(torch.cat([att[0][:, :E], so[0]], 0)); self.b_hproj = _f32(torch.cat([z, so[1]]))
        self.W_oproj = _bf16(torch.cat([att[0][:, E:], os_[0]], 0)); self.b_oproj = _f32(torch.cat([att[1], os_[1]]))
        self.Wg1, self.bg1 = _bf16(gg[0], pad_k=64), _f32(gg[1])
        self.W2cat = _bf16(torch.cat([att[2], os_[2], so[2], gg[2]], 0))
        self.b2cat = _f32(torch.cat([att[3], os_[3], so[3], gg[3]]))
        self.W3att, self.b3att = _bf16(att[4]), _f32(att[5])
        self.W3os, self.b3os = _bf16(os_[4]), _f32(os_[5])
        self.W3so, self.b3so = _bf16(so[4]), _f32(so[5])
        self.W3g, self.b3g = _bf16(gg[4]), _f32(gg[5])
        self.w_adj = _f32(G.adjacency.weight.reshape(-1)); self.b_adj = float(G.adjacency.bias.detach().float().item())
        self.ln = [_f32(t) for t in (G.norm_h.weight, G.norm_h.bias, G.norm_o.weight, G.norm_o.bias)]
        self.ln_eps = float(G.norm_h.eps)
        nc = ih.box_pair_predictor.weight.shape[0]
        self.num_cls = nc
        Wp = torch.cat([ih.box_pair_predictor.weight, ih.box_pair_suppressor.weight], 0).detach().float()
        bp = torch.cat([ih.box_pair_predictor.bias, ih.box_pair_suppressor.bias], 0)
        self.n_logit = nc + 1
        self.Wpred = _bf16(Wp, pad_rows=128)
        self.bpred = _f32(bp, pad=128)
        # split-bf16 predictor: W = W_hi + W_lo, used with A = [hi | lo]  (near-fp32 product, 3 bf16 MMAs)
        hi = Wp.to(torch.bfloat16); lo = (Wp - hi.float()).to(torch.bfloat16)
        self.Wpred_hh = _bf16(torch.cat([hi.float(), hi.float()], 1), pad_rows=128)     # [128, 2*2R] against [hi | lo]
        self.Wpred_lo = _bf16(lo.float(), pad_rows=128)                                 # [128, 2R]   against hi


def packed_head(ih) -> PackedHead:
    pk = ih.__dict__.get("_skg_packed")
    if pk is None or pk.key != _versions(ih):
        pk = PackedHead(ih)
        ih.__dict__["_skg_packed"] = pk
    return pk


# =============================================================================================
# the batched forward pipeline (GraphHead.forward head.py:769-993 + predictor head.py:408-411)
# =============================================================================================
def _g(**kw):
    return _lib.epilogue_group(**kw)


def _host_meta(box_labels: List[Tensor], box_scores: Optional[List[Tensor]], human_idx: int, meta=None):
    """n, n_h, labels and scores on the host.  The reference syncs per image (``.item()``, head.py:826);
    here it is ONE device->host copy per batch, or none when the caller supplies ``meta`` (host copies)."""
    if meta is not None:
        return meta
    n = [int(l.shape[0]) for l in box_labels]
    if len(box_labels) == 0 or sum(n) == 0:
        return dict(n=n, n_h=[0] * len(n), labels=np.zeros(0, np.int64), scores=np.zeros(0, np.float32))
    lab = torch.cat(list(box_labels))
    if box_scores is not None:
        packed = torch.stack([lab.to(torch.float64), torch.cat(list(box_scores)).to(torch.float64)]).cpu().numpy()
        labels = packed[0].astype(np.int64); scores = packed[1].astype(np.float32)
    else:
        labels = lab.cpu().numpy().astype(np.int64); scores = None
    off = np.concatenate([[0], np.cumsum(n)])
    n_h = [int((labels[off[b]:off[b + 1]] == human_idx).sum()) for b in range(len(n))]
    for b in range(len(n)):
        if n_h[b] and not np.all(labels[off[b]:off[b] + n_h[b]] == human_idx):
            raise ValueError("Human detections are not permuted to the top")            # head.py:840-841
    return dict(n=n, n_h=n_h, labels=labels, scores=scores)


def head_forward(ih, features3: Tensor, image_shapes, box_features: Tensor, boxes_cat: Tensor, meta: dict,
                 want_taps: bool = False, precise_logits: bool = True) -> dict:
    """Runs the whole pair path for a batch.  Inputs: features['3'] [B,256,h,w] fp32, pooled box
    features [sum n, 256, 7, 7] fp32, concatenated boxes [sum n, 4] fp32 (all CUDA), host ``meta``.
    Returns a dict of device tensors: logits [P,118] fp32 (117 verb logits + suppressor logit),
    feat (bf16 [P, 2*R] hi part, and lo part when precise_logits), pair indices, encodings, etc."""
    pk = packed_head(ih)
    G = ih.box_pair_head
    dev = pk.device
    R, E = pk.R, pk.E
    n, n_h = meta["n"], meta["n_h"]
    if any(k > G.transh_head.num_object for k in n):
        raise IndexError("index out of range in self")   # head.py:572: box index addresses the 80-row entity table
    # images the reference skips (head.py:829-839) contribute no cells: treat them as n_h = 0
    n_h_eff = [0 if (h == 0 or k <= 1) else h for h, k in zip(n_h, n)]
    seg = Segments(n_h_eff, n, image_shapes, dev)
    B, NB, NH, C, Pn = seg.B, seg.total_boxes, seg.total_humans, seg.total_cells, seg.total_pairs
    out = dict(seg=seg)
    f32 = dict(dtype=torch.float32, device=dev); b16 = dict(dtype=torch.bfloat16, device=dev)
    st = _lib.stream_ptr(dev)
    with torch.cuda.device(dev):
        pe = enumerate_pairs(None, n_h_eff, image_shapes, seg=seg, boxes_cat=boxes_cat)
        out.update(pe)
        if C == 0:
            out.update(logits=torch.zeros(0, pk.n_logit, **f32), feat=torch.zeros(0, 2 * R, **b16))
            return out
        # ---- spatial head (head.py:662-669, 888): fp32 layer 0, then two tensor-core layers
        s1 = torch.empty(C, 128, **b16)
        _lib.call("skg_spatial_fc1", _lib.ptr(pe["enc"]), _lib.ptr(pk.Wsp0), _lib.ptr(pk.bsp0), C, 46, 128, _lib.ptr(s1), st)
        s2 = torch.empty(C, 256, **b16)
        _lib.gemm_bf16(s1, pk.Wsp1, [_g(relu=True, bias=pk.bsp1, out_bf16=s2)])
        sp = torch.empty(C, R, **b16)
        _lib.gemm_bf16(s2, pk.Wsp2, [_g(relu=True, bias=pk.bsp2, out_bf16=sp)])
        # ---- box head (head.py:635-641, 812)
        K0 = pk.Wb1.shape[1]
        bf = box_features.reshape(NB, -1)
        xb = torch.empty(NB, K0, **b16)
        _lib.call("skg_cast_pad_bf16", _lib.ptr(bf), bf.stride(0), None, 0, NB, bf.shape[1], _lib.ptr(xb), K0, K0, st)
        x1 = torch.empty(NB, E, **b16)
        _lib.gemm_bf16(xb, pk.Wb1, [_g(relu=True, bias=pk.bb1, out_bf16=x1)])
        node = torch.empty(NB, E, **b16)
        _lib.gemm_bf16(x1, pk.Wb2, [_g(relu=True, bias=pk.bb2, out_bf16=node)])
        # ---- fc_head / fc_tail on the n_h (resp. n) DISTINCT rows (head.py:875-885)
        tr = G.transh_head.transh
        if tr.ent_embeddings.weight.device != dev:
            tr.to(dev)
        ent = tr.ent_embeddings.weight.detach()
        h_in = torch.empty(NH, pk.fc_k, **b16); o_in = torch.empty(NB, pk.fc_k, **b16)
        _lib.call("skg_node_inputs", _lib.ptr(node), E, _lib.ptr(ent), ent.shape[1], ent.shape[0], G.human_idx,
                  _lib.ptr(seg.d_box_off), _lib.ptr(seg.d_hum_off), B, NB, NH, _lib.ptr(h_in), _lib.ptr(o_in), pk.fc_k, st)
        hN32 = torch.empty(NH, R, **f32); hN16 = torch.empty(NH, R, **b16)
        oN32 = torch.empty(NB, R, **f32); oN16 = torch.empty(NB, R, **b16)
        _lib.gemm_bf16(h_in, pk.Wfh, [_g(relu=True, bias=pk.bfh, out_f32=hN32, out_bf16=hN16)])
        _lib.gemm_bf16(o_in, pk.Wft, [_g(relu=True, bias=pk.bft, out_f32=oN32, out_bf16=oN16)])
        # ---- per-node halves of the MBF appearance terms
        hproj = torch.empty(NH, 2 * R, **f32); oproj = torch.empty(NB, 2 * R, **f32)
        _lib.gemm_bf16(hN16, pk.W_hproj, [_g(bias=pk.b_hproj, out_f32=hproj)])
        _lib.gemm_bf16(oN16, pk.W_oproj, [_g(bias=pk.b_oproj, out_f32=oproj)])
        # ---- global feature term of attention_head_g (head.py:811, 971)
        Cg = features3.shape[1]
        f3 = features3.contiguous()
        Kg = pk.Wg1.shape[1]
        gfeat = torch.zeros(B, Kg, **b16)
        if Kg == Cg:
            _lib.call("skg_global_pool", _lib.ptr(f3), B * Cg, f3.shape[2] * f3.shape[3], _lib.ptr(gfeat), None, st)
        else:
            g32 = torch.empty(B, Cg, **f32)
            _lib.call("skg_global_pool", _lib.ptr(f3), B * Cg, f3.shape[2] * f3.shape[3], None, _lib.ptr(g32), st)
            _lib.call("skg_cast_pad_bf16", _lib.ptr(g32), Cg, None, 0, B, Cg, _lib.ptr(gfeat), Kg, Kg, st)
        Gt = torch.empty(B, R, **f32)
        _lib.gemm_bf16(gfeat, pk.Wg1, [_g(bias=pk.bg1, out_f32=Gt)])
        # ---- fused fc_2 GEMM of the four MBFs that read the spatial features (N = 4R)
        H_att = torch.empty(C, R, **b16); S_raw = torch.empty(C, R, **b16)
        H_os = torch.empty(C, R, **b16); H_so = torch.empty(C, R, **b16); H_g = torch.empty(C, R, **b16)
        hrow, orow, img = pe["cell_hrow"], pe["cell_orow"], pe["cell_img"]
        _lib.gemm_bf16(sp, pk.W2cat, [
            _g(n_begin=0, relu=True, bias=pk.b2cat, mul_p=hproj[:, :R], row_p=hrow, mul_q=oproj[:, :R], row_q=orow,
               out_bf16=H_att, raw_bf16=S_raw),
            _g(n_begin=R, relu=True, bias=pk.b2cat, mul_p=oproj[:, R:], row_p=orow, out_bf16=H_os),
            _g(n_begin=2 * R, relu=True, bias=pk.b2cat, mul_p=hproj[:, R:], row_p=hrow, out_bf16=H_so),
            _g(n_begin=3 * R, relu=True, bias=pk.b2cat, mul_p=Gt, row_p=img, out_bf16=H_g)])
        # ---- attention weights -> adjacency (head.py:896-897); the two iterations of head.py:892 are identical
        nt = R // 256
        dot_part = torch.empty(C, nt, **f32)
        _lib.gemm_bf16(H_att, pk.W3att, [_g(relu=True, bias=pk.b3att, dot_w=pk.w_adj, dot_out=dot_part)], block_n=256)
        # ---- messages (head.py:907-925)
        Z = torch.empty(NH + NB, R, **b16)
        adjacency = torch.empty(C, **f32)
        _lib.call("skg_message_aggregate", _lib.ptr(dot_part), nt, pk.b_adj, _lib.ptr(H_os), _lib.ptr(H_so), R,
                  _lib.ptr(seg.d_box_off), _lib.ptr(seg.d_hum_off), _lib.ptr(seg.d_cell_off), B, NB, NH, int(max(n)), R,
                  _lib.ptr(Z[:NH]), _lib.ptr(Z[NH:]), _lib.ptr(adjacency), st)
        pre = torch.empty(NH + NB, R, **f32)
        _lib.gemm_bf16(Z[:NH], pk.W3os, [_g(relu=True, bias=pk.b3os, residual=hN32, out_f32=pre[:NH])])
        _lib.gemm_bf16(Z[NH:], pk.W3so, [_g(relu=True, bias=pk.b3so, residual=oN32, out_f32=pre[NH:])])
        enc16 = torch.empty(NH + NB, R, **b16)
        enc32 = torch.empty(NH + NB, R, **f32) if want_taps else None
        _lib.call("skg_layernorm", _lib.ptr(pre), NH + NB, NH, R, _lib.ptr(pk.ln[0]), _lib.ptr(pk.ln[1]), _lib.ptr(pk.ln[2]),
                  _lib.ptr(pk.ln[3]), pk.ln_eps, _lib.ptr(enc32), _lib.ptr(enc16), st)
        # ---- final attention heads on the valid pairs (head.py:966-973)
        U2h = torch.empty(NH, R, **f32); U2o = torch.empty(NB, R, **f32)
        _lib.gemm_bf16(enc16[:NH], pk.W_hproj, [_g(out_f32=U2h)], N=R)
        _lib.gemm_bf16(enc16[NH:], pk.W_oproj, [_g(bias=pk.b_oproj, out_f32=U2o)], N=R)
        H2 = torch.empty(C, R, **b16)
        _lib.call("skg_hadamard_cells", _lib.ptr(S_raw), R, _lib.ptr(U2h), R, _lib.ptr(U2o), R, _lib.ptr(hrow), _lib.ptr(orow),
                  C, R, _lib.ptr(H2), R, st)
        # feat rows are the valid pairs (cell2pair scatter); columns [hi(att1) hi(att2) | lo(att1) lo(att2)]
        feat = torch.empty(Pn, 4 * R if precise_logits else 2 * R, **b16)
        c2p = pe["cell2pair"]
        lo1 = feat[:, 2 * R:3 * R] if precise_logits else None
        lo2 = feat[:, 3 * R:] if precise_logits else None
        _lib.gemm_bf16(H2, pk.W3att, [_g(relu=True, bias=pk.b3att, out_bf16=feat[:, :R], out_row=c2p, lo_bf16=lo1)])
        _lib.gemm_bf16(H_g, pk.W3g, [_g(relu=True, bias=pk.b3g, out_bf16=feat[:, R:2 * R], out_row=c2p, lo_bf16=lo2)])
        # ---- predictor + suppressor (head.py:410-411) as one N = num_cls + 1 GEMM
        logits = torch.empty(Pn, pk.n_logit, **f32)
        if precise_logits:
            tmp = torch.empty(Pn, pk.Wpred.shape[0], **f32)       # pitch padded to 128 (read back as residual)
            _lib.gemm_bf16(feat, pk.Wpred_hh, [_g(bias=pk.bpred, out_f32=tmp)], N=pk.n_logit)
            _lib.gemm_bf16(feat[:, :2 * R], pk.Wpred_lo, [_g(residual=tmp, out_f32=logits)], N=pk.n_logit, K=2 * R)
        else:
            _lib.gemm_bf16(feat, pk.Wpred, [_g(bias=pk.bpred, out_f32=logits)], N=pk.n_logit)
    out.update(logits=logits, feat=feat, adjacency=adjacency, node16=node, hN32=hN32, oN32=oN32, enc32=enc32,
               enc16=enc16, sp=sp, NH=NH, NB=NB)
    return out
