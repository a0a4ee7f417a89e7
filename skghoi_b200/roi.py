"""``box_roi_pool`` of the interaction head on the native kernel.

``MultiScaleRoIAlign`` mirrors ``torchvision.ops.MultiScaleRoIAlign`` as the reference builds and calls it
(models.py:158-162: ``MultiScaleRoIAlign(['0', '1', '2', '3'], 7, 2)``; head.py:387:
``box_roi_pool(features, box_coords, image_shapes)``): same constructor arguments, same call, same ``[sum n, C, 7, 7]``
fp32 result -- one launch for the whole batch instead of one ``roi_align`` + ``where`` + indexed copy per pyramid level.
CUDA only; there is no CPU fallback.
"""
import ctypes
import math
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch
from torch import Tensor

from . import _lib


def infer_scales(features: Sequence[Tensor], image_shapes: Sequence[Tuple[int, int]]) -> List[float]:
    """torchvision poolers._setup_scales / _infer_scale: the scale of a level is the power of two nearest to
    feature height / largest image height of the batch."""
    if not image_shapes:
        raise ValueError("images list should not be empty")
    max_h = max(int(s[0]) for s in image_shapes)
    return [2.0 ** float(round(math.log2(float(f.shape[-2]) / float(max_h)))) for f in features]


class MultiScaleRoIAlign(torch.nn.Module):
    def __init__(self, featmap_names: List[str], output_size: Union[int, Tuple[int, int], List[int]], sampling_ratio: int,
                 *, canonical_scale: int = 224, canonical_level: int = 4):
        super().__init__()
        if isinstance(output_size, int):
            output_size = (output_size, output_size)
        if output_size[0] != output_size[1]:
            raise ValueError("the native RoIAlign pools square outputs (the reference uses 7 x 7)")
        if sampling_ratio <= 0:
            raise ValueError("adaptive sampling (sampling_ratio <= 0) is not supported (the reference uses 2)")
        self.featmap_names = list(featmap_names)
        self.output_size = tuple(output_size)
        self.sampling_ratio = int(sampling_ratio)
        self.canonical_scale, self.canonical_level = canonical_scale, canonical_level
        self.scales: Optional[List[float]] = None

    def _levels(self, x: Dict[str, Tensor]) -> List[Tensor]:
        return [v for k, v in x.items() if k in self.featmap_names]           # poolers._filter_input

    def pool(self, x: Dict[str, Tensor], boxes: List[Tensor], image_shapes: List[Tuple[int, int]], want_f32: bool = True,
             bf16_out: Optional[Tensor] = None) -> Optional[Tensor]:
        """The batch's pooled features: fp32 ``[sum n, C, P, P]`` (``want_f32``) and / or rounded to bf16 into the rows of
        ``bf16_out`` ``[sum n, >= C*P*P]`` (the K-major operand of box_head's first GEMM)."""
        feats = self._levels(x)
        if not feats:
            raise ValueError("none of the feature maps %s is in the input" % self.featmap_names)
        _lib.require_cuda(*feats, *boxes)
        dev = feats[0].device
        feats = [f if (f.dtype == torch.float32 and f.is_contiguous()) else f.float().contiguous() for f in feats]
        if self.scales is None or len(self.scales) != len(feats):
            self.scales = infer_scales(feats, image_shapes)                   # set once, like torchvision
        k_min = int(-math.log2(self.scales[0])); k_max = int(-math.log2(self.scales[-1]))
        n = [int(b.shape[0]) for b in boxes]
        total = sum(n)
        C, P = int(feats[0].shape[1]), self.output_size[0]
        out = torch.empty(total, C, P, P, dtype=torch.float32, device=dev) if want_f32 else None
        if total == 0:
            return out
        from .engine import _cat_rows
        boxes_cat = _cat_rows(boxes, torch.float32, dev, 4)
        if boxes_cat.data_ptr() % 16:
            boxes_cat = boxes_cat.clone()
        off = np.zeros(len(n) + 1, dtype=np.int32)
        np.cumsum(n, out=off[1:])
        d_off = torch.from_numpy(off).to(dev, non_blocking=True)
        L = len(feats)
        ptrs = (ctypes.c_void_p * L)(*[f.data_ptr() for f in feats])
        hs = (ctypes.c_int32 * L)(*[int(f.shape[-2]) for f in feats])
        ws = (ctypes.c_int32 * L)(*[int(f.shape[-1]) for f in feats])
        sc = (ctypes.c_float * L)(*self.scales)
        if bf16_out is not None:
            assert bf16_out.dtype == torch.bfloat16 and bf16_out.is_cuda and bf16_out.stride(1) == 1 and bf16_out.shape[0] >= total
        with torch.cuda.device(dev):
            _lib.call("skg_multiscale_roi_align", ptrs, hs, ws, sc, L, C, len(n), _lib.ptr(boxes_cat), _lib.ptr(d_off), total,
                      P, self.sampling_ratio, k_min, k_max, float(self.canonical_scale), int(self.canonical_level),
                      _lib.ptr(out), _lib.ptr(bf16_out), int(bf16_out.stride(0)) if bf16_out is not None else 0,
                      _lib.stream_ptr(dev))
        return out

    def channels_times_bins(self, x: Dict[str, Tensor]) -> int:
        return int(self._levels(x)[0].shape[1]) * self.output_size[0] * self.output_size[1]

    def forward(self, x: Dict[str, Tensor], boxes: List[Tensor], image_shapes: List[Tuple[int, int]]) -> Tensor:
        return self.pool(x, boxes, image_shapes)
