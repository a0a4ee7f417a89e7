"""Microbenchmark of skg_gemm_bf16: mainloop ceiling vs epilogue cost (CUDA events, L2 flushed between launches).
    python tools/gemm_bench.py            (on the GPU box)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=20, do_flush=True):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        if do_flush:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / reps * 1e-3


def run(M, N, K, kind, bn=0):
    A = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16)
    B = (torch.randn(N, K, device=dev) * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device=dev)
    out16 = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    if kind == "plain":
        groups = [_lib.epilogue_group(relu=True, bias=bias, out_bf16=out16)]
    elif kind == "f32":
        out32 = torch.empty(M, N, device=dev)
        groups = [_lib.epilogue_group(bias=bias, out_f32=out32)]
    elif kind == "fused4":       # the head's fused fc_2 epilogue: 4 groups, Hadamard with gathered tables
        R = N // 4
        nh, nb = M // 32, M // 16
        hproj = torch.randn(nh, 2 * R, device=dev); oproj = torch.randn(nb, 2 * R, device=dev); gt = torch.randn(8, R, device=dev)
        ar = torch.arange(M, device=dev)
        hrow = (ar // 32).int(); orow = ((ar // 512) * 32 + ar % 32).int(); img = (ar // 512).int() % 8
        outs = [torch.empty(M, R, dtype=torch.bfloat16, device=dev) for _ in range(5)]
        g = _lib.epilogue_group
        groups = [g(n_begin=0, relu=True, bias=bias, mul_p=hproj[:, :R], row_p=hrow, mul_q=oproj[:, :R], row_q=orow, out_bf16=outs[0], raw_bf16=outs[1]),
                  g(n_begin=R, relu=True, bias=bias, mul_p=oproj[:, R:], row_p=orow, out_bf16=outs[2]),
                  g(n_begin=2 * R, relu=True, bias=bias, mul_p=hproj[:, R:], row_p=hrow, out_bf16=outs[3]),
                  g(n_begin=3 * R, relu=True, bias=bias, mul_p=gt, row_p=img, out_bf16=outs[4])]
    elif kind in ("P_zero", "P_orow", "P_ident"):   # one group over all columns: Hadamard with ONE gathered table
        ar = torch.arange(M, device=dev)
        if kind == "P_zero":
            tab = torch.randn(8, N, device=dev); rows = torch.zeros(M, dtype=torch.int32, device=dev)
        elif kind == "P_orow":
            tab = torch.randn(M // 16, N, device=dev); rows = ((ar // 512) * 32 + ar % 32).int()
        else:
            tab = torch.randn(M, N, device=dev); rows = ar.int()
        groups = [_lib.epilogue_group(relu=True, bias=bias, mul_p=tab, row_p=rows, out_bf16=out16)]
    sec = timeit(lambda: _lib.gemm_bf16(A, B, groups, block_n=bn))
    tf = 2.0 * M * N * K / sec / 1e12
    print("%-7s M=%5d N=%5d K=%5d bn=%3d : %8.2f us  %7.1f TFLOP/s" % (kind, M, N, K, bn, sec * 1e6, tf))
    if kind == "plain" and M * N * K <= 4096 ** 3:
        sec = timeit(lambda: torch.matmul(A, B.t()))
        print("        cuBLAS same shape            : %8.2f us  %7.1f TFLOP/s" % (sec * 1e6, 2.0 * M * N * K / sec / 1e12))


def anatomy():
    """Per-CTA timeline T = setup + ML + (t-1)*max(EP, ML) + EP from launches with t tiles per CTA (N = 256, M = t*148*128)
    at K = 64 (ML ~ 0 -> EP) and K = 1024, for the plain and the fused (P+Q+RAW+OB, P+OB) epilogues."""
    for kind in ("plain", "fusedA", "fusedP"):
        for K in (64, 1024, 2048):
            for t in (1, 2, 4):
                M, N = t * 148 * 128, 256
                A = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16)
                B = (torch.randn(N, K, device=dev) * 0.1).to(torch.bfloat16)
                bias = torch.randn(N, device=dev)
                o0 = torch.empty(M, N, dtype=torch.bfloat16, device=dev); o1 = torch.empty_like(o0)
                ar = torch.arange(M, device=dev)
                hrow = (ar // 32).int(); orow = ((ar // 512) * 32 + ar % 32).int()
                hp = torch.randn(M // 32, N, device=dev); op = torch.randn(M // 16, N, device=dev)
                g = _lib.epilogue_group
                if kind == "plain":
                    groups = [g(relu=True, bias=bias, out_bf16=o0)]
                elif kind == "fusedA":
                    groups = [g(relu=True, bias=bias, mul_p=hp, row_p=hrow, mul_q=op, row_q=orow, out_bf16=o0, raw_bf16=o1)]
                else:
                    groups = [g(relu=True, bias=bias, mul_p=op, row_p=orow, out_bf16=o0)]
                sec = timeit(lambda: _lib.gemm_bf16(A, B, groups, block_n=256))
                print("anatomy %-6s K=%4d tiles/CTA=%d : %7.2f us" % (kind, K, t, sec * 1e6))


if "--anatomy" in sys.argv:
    anatomy()
    sys.exit(0)
if "--epi" in sys.argv:
    run(4096, 4096, 4096, "plain")
    for kind in ("plain", "P_zero", "P_orow", "P_ident", "fused4", "plain"):
        run(4096, 4096, 1024, kind)
    sys.exit(0)
if "--fused-only" in sys.argv:       # A/B protocol: same pre-conditioning every time, plain GEMM of the same shape as control
    run(4096, 4096, 4096, "plain")
    for _ in range(3):
        run(4096, 4096, 1024, "plain")
        run(4096, 4096, 1024, "fused4")
    sys.exit(0)
run(8192, 8192, 8192, "plain")
run(4096, 4096, 4096, "plain")
run(4096, 4096, 1024, "plain")
run(4096, 4096, 1024, "plain", 128)
run(4096, 4096, 1024, "fused4")
run(4096, 1024, 1024, "plain")
run(4096, 1024, 1024, "plain", 128)
run(4096, 1024, 256, "plain")
run(3968, 128, 4096, "f32")
run(256, 1024, 1024, "plain")
run(256, 2048, 1024, "f32")
