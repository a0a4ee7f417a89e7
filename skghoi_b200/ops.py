"""Reference-facing operator surface of ``ops.py`` (reference ops.py:20-211) on the CUDA library.

Same names, argument meaning and error behaviour as the reference functions; the arithmetic runs
in libskghoi_b200.so (csrc/spatial.cu, csrc/losses.cu).  CUDA tensors only -- there is no CPU path.
"""
from typing import List, Sequence, Tuple

import torch
from torch import Tensor

from . import _lib


def _encode_rows(boxes_1: Sequence[Tensor], boxes_2: Sequence[Tensor], shapes: Sequence[Tuple[int, int]],
                 eps: float, num_features: int, check_shapes: bool) -> Tensor:
    if not (len(boxes_1) == len(boxes_2) == len(shapes)):
        # zip() in the reference silently truncates; mismatched lists are a caller bug we surface
        n = min(len(boxes_1), len(boxes_2), len(shapes))
        boxes_1, boxes_2, shapes = boxes_1[:n], boxes_2[:n], shapes[:n]
    if len(boxes_1) == 0:
        raise RuntimeError("torch.cat(): expected a non-empty list of Tensors")   # ops.py:157
    _lib.require_cuda(*boxes_1, *boxes_2)
    dev = boxes_1[0].device
    hw, off = [], [0]
    for b1, b2, (h, w) in zip(boxes_1, boxes_2, shapes):
        if check_shapes:
            assert h > 0        # ops.py:43
            assert w > 0        # ops.py:44
        if b1.shape != b2.shape or b1.dim() != 2 or b1.shape[1] != 4:
            raise RuntimeError("boxes_1 / boxes_2 entries must both be [M, 4], got %s and %s"
                               % (tuple(b1.shape), tuple(b2.shape)))
        hw.append([float(h), float(w), float(h * w)])
        off.append(off[-1] + b1.shape[0])
    rows = off[-1]
    b1 = torch.cat(list(boxes_1)).to(torch.float32).contiguous()
    b2 = torch.cat(list(boxes_2)).to(torch.float32).contiguous()
    out = torch.empty(rows, 2 * num_features, dtype=torch.float32, device=dev)
    if rows == 0:
        return out
    hw_t = torch.tensor(hw, dtype=torch.float32).to(dev)
    off_t = torch.tensor(off, dtype=torch.int64).to(dev)
    with torch.cuda.device(dev):
        _lib.call("skg_spatial_encode_rows", _lib.ptr(b1), _lib.ptr(b2), _lib.ptr(off_t), _lib.ptr(hw_t),
                  len(hw), rows, num_features, float(eps), _lib.ptr(out), None, _lib.stream_ptr(dev))
    return out


def compute_spatial_encodings(boxes_1: List[Tensor], boxes_2: List[Tensor],
                              shapes: List[Tuple[int, int]], eps: float = 1e-10) -> Tensor:
    """Reference ops.py:20-83: 18 features + their logs -> Tensor[N, 36].

    The reference's ``features = torch.cat(features)`` sits inside its loop (ops.py:81) and raises
    for lists longer than one image; the evident intent -- concatenate over images -- is
    implemented, which is identical to the reference for 1-image lists."""
    return _encode_rows(boxes_1, boxes_2, shapes, eps, 18, check_shapes=True)


def compute_spatial_ratio_encodings(boxes_1: List[Tensor], boxes_2: List[Tensor],
                                    shapes: List[Tuple[int, int]], eps: float = 1e-10) -> Tensor:
    """Reference ops.py:85-157: 23 features + their logs -> Tensor[N, 46] (the variant the head
    calls at head.py:863)."""
    return _encode_rows(boxes_1, boxes_2, shapes, eps, 23, check_shapes=False)


class _FocalLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, alpha, gamma, eps):
        _lib.require_cuda(x, y)
        xc = x.contiguous().float(); yc = y.contiguous().float()
        out = torch.empty_like(xc)
        with torch.cuda.device(x.device):
            _lib.call("skg_focal_loss", _lib.ptr(xc), _lib.ptr(yc), xc.numel(), float(alpha), float(gamma),
                      float(eps), _lib.ptr(out), None, None, _lib.stream_ptr(x.device))
        ctx.save_for_backward(xc, yc)
        ctx.hp = (alpha, gamma, eps)
        return out

    @staticmethod
    def backward(ctx, g):
        xc, yc = ctx.saved_tensors
        alpha, gamma, eps = ctx.hp
        gc = g.contiguous().float()
        dx = torch.empty_like(xc)
        with torch.cuda.device(xc.device):
            _lib.call("skg_focal_loss", _lib.ptr(xc), _lib.ptr(yc), xc.numel(), float(alpha), float(gamma),
                      float(eps), None, _lib.ptr(gc), _lib.ptr(dx), _lib.stream_ptr(xc.device))
        return dx, None, None, None, None


def binary_focal_loss(x: Tensor, y: Tensor, alpha: float = 0.5, gamma: float = 2.0,
                      reduction: str = "mean", eps: float = 1e-6) -> Tensor:
    """Reference ops.py:159-211: ``|1-y-alpha| * (|y-x|+eps)^gamma * BCE(x, y)``."""
    if reduction not in ("mean", "sum", "none"):
        raise ValueError("Unsupported reduction method {}".format(reduction))   # ops.py:211
    if x.shape != y.shape:
        raise ValueError("Using a target size ({}) that is different to the input size ({}) is deprecated."
                         " Please ensure they have the same size.".format(y.shape, x.shape))
    loss = _FocalLoss.apply(x, y, alpha, gamma, eps).view(x.shape)
    if reduction == "mean":
        return loss.mean()
    if reduction == "sum":
        return loss.sum()
    return loss
