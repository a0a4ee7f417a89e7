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
