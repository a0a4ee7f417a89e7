"""box_head.1-shaped split-K GEMM (M = boxes, N = 1024, K = 12544) under different tile widths / split depths
(SKG_SPLITK_MIN_KB): python tools/splitk_bench.py   (on the GPU box)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import _lib
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for M in (256, 512):
    N, K = 1024, 12544
    A = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16); B = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device=dev); out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    scratch = torch.zeros(int(_lib.lib().skg_gemm_splitk_scratch_bytes(0, 0)), dtype=torch.uint8, device=dev)
    g = [_lib.epilogue_group(relu=True, bias=bias, out_bf16=out)]
    ref = torch.relu(A.float() @ B.float().t() + bias)
    for bn in (64, 128, 256):
        fn = lambda: _lib.gemm_bf16(A, B, g, block_n=bn, scratch=scratch)
        for _ in range(3): fn()
        torch.cuda.synchronize()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        tot = 0.0
        for _ in range(20):
            flush.zero_(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize(); tot += s.elapsed_time(e)
        # warm (operands in L2, as inside the step after the cast kernel)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(20): fn()
        e.record(); torch.cuda.synchronize()
        print("M=%d bn=%3d min_kb=%s : flushed %.2f us  warm back-to-back %.2f us  max rel err %.1e" % (M, bn, os.environ.get("SKG_SPLITK_MIN_KB", "16"), tot / 20 * 1e3, s.elapsed_time(e) / 20 * 1e3, err))
