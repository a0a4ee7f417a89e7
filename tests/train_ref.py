"""Plain-PyTorch pieces used as autograd references by the GPU tests of the training kernels (test infrastructure)."""
import torch


def segment_softmax(v: torch.Tensor, seg: torch.Tensor, num_segments: int) -> torch.Tensor:
    """softmax of ``v`` inside each segment id (head.py:903, 914: adjacency rows / columns of one image)."""
    m = torch.full((num_segments,), float("-inf"), dtype=v.dtype, device=v.device).scatter_reduce(0, seg, v.detach(), "amax")
    e = torch.exp(v - m[seg])
    s = torch.zeros(num_segments, dtype=v.dtype, device=v.device).index_add(0, seg, e)
    return e / s[seg]
