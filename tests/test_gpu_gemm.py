"""GPU parity of the tcgen05 bf16 GEMM and its fused epilogues against a plain fp32 torch
reference computed from the same bf16-rounded operands."""
import pytest
import torch

from skghoi_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops(M, N, K, seed=0):
    g = torch.Generator().manual_seed(seed)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    B = (torch.randn(N, K, generator=g) * 0.1).to(torch.bfloat16).to(DEV)
    return A, B


@pytest.mark.parametrize("M,N,K,bn", [(128, 256, 64, 0), (128, 128, 64, 128), (256, 512, 1024, 0), (300, 384, 200, 0),
                                      (4096, 1024, 1024, 0), (31, 128, 2048, 0), (1000, 1024, 256, 128),
                                      (520, 4096, 1024, 256)])
def test_gemm_bias_relu(M, N, K, bn):
    A, B = _ops(M, N, K, seed=M + N + K)
    Kp = (K + 7) // 8 * 8
    if Kp != K:
        A = torch.nn.functional.pad(A, (0, Kp - K)); B = torch.nn.functional.pad(B, (0, Kp - K))
    bias = torch.randn(N, device=DEV)
    out32 = torch.full((M, N), float("nan"), device=DEV)
    out16 = torch.zeros(M, N, dtype=torch.bfloat16, device=DEV)
    g = _lib.epilogue_group(relu=True, bias=bias, out_f32=out32, out_bf16=out16)
    _lib.gemm_bf16(A, B, [g], K=Kp, block_n=bn)
    ref = torch.relu(A.float() @ B.float().t() + bias)
    torch.testing.assert_close(out32, ref, rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(out16.float(), ref, rtol=1e-2, atol=1e-2)


def test_gemm_unpadded_n_and_rowmap():
    """N = 118 (predictor + suppressor) with fp32 output masked at N; row scatter with skipped rows."""
    M, N, K = 500, 118, 2048
    A, B = _ops(M, 128, K, seed=3)
    B[N:] = 0
    bias = torch.zeros(128, device=DEV); bias[:N] = torch.randn(N, device=DEV)
    out = torch.full((M, N), 7.0, device=DEV)
    g = _lib.epilogue_group(bias=bias, out_f32=out)
    _lib.gemm_bf16(A, B, [g], N=N)
    ref = A.float() @ B[:N].float().t() + bias[:N]
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-3)
    # scatter: even rows -> compact positions, odd rows skipped
    rowmap = torch.full((M,), -1, dtype=torch.int32, device=DEV)
    rowmap[::2] = torch.arange((M + 1) // 2, dtype=torch.int32, device=DEV)
    out2 = torch.full(((M + 1) // 2, N), 7.0, device=DEV)
    _lib.gemm_bf16(A, B, [_lib.epilogue_group(bias=bias, out_f32=out2, out_row=rowmap)], N=N)
    torch.testing.assert_close(out2, ref[::2], rtol=2e-3, atol=2e-3)


def test_gemm_hadamard_groups_dot_residual():
    """Four column groups with different epilogues in one launch (the fused fc_2 GEMM of the head)."""
    M, K, R = 700, 1024, 256
    N = 4 * R
    A, B = _ops(M, N, K, seed=9)
    gen = torch.Generator().manual_seed(1)
    bias = torch.randn(N, generator=gen).to(DEV)
    nP, nQ = 13, 29
    P = torch.randn(nP, R, generator=gen).to(DEV); Q = torch.randn(nQ, R, generator=gen).to(DEV)
    rp = torch.randint(0, nP, (M,), generator=gen).int().to(DEV); rq = torch.randint(0, nQ, (M,), generator=gen).int().to(DEV)
    res = torch.randn(M, R, generator=gen).to(DEV)
    w = torch.randn(R, generator=gen).to(DEV)
    o0 = torch.zeros(M, R, dtype=torch.bfloat16, device=DEV); raw0 = torch.zeros(M, R, dtype=torch.bfloat16, device=DEV)
    o1 = torch.zeros(M, R, dtype=torch.bfloat16, device=DEV)
    o2 = torch.zeros(M, R, device=DEV)
    dot = torch.zeros(M, 4 * (R // 256), device=DEV)      # four partials per 256-column tile
    o3 = torch.zeros(M, R, device=DEV)
    groups = [
        _lib.epilogue_group(0, relu=True, bias=bias, mul_p=P, row_p=rp, mul_q=Q, row_q=rq, out_bf16=o0, raw_bf16=raw0),
        _lib.epilogue_group(R, relu=True, bias=bias, mul_p=Q, row_p=rq, out_bf16=o1),
        _lib.epilogue_group(2 * R, relu=True, bias=bias, residual=res, out_f32=o2),
        _lib.epilogue_group(3 * R, relu=True, bias=bias, dot_w=w, dot_out=dot, out_f32=o3),
    ]
    _lib.gemm_bf16(A, B, groups)
    C = A.float() @ B.float().t() + bias
    c0, c1, c2, c3 = C.split(R, dim=1)
    torch.testing.assert_close(raw0.float(), c0, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(o0.float(), torch.relu(c0 * (P[rp.long()] + Q[rq.long()])), rtol=1e-2, atol=2e-2)
    torch.testing.assert_close(o1.float(), torch.relu(c1 * Q[rq.long()]), rtol=1e-2, atol=2e-2)
    torch.testing.assert_close(o2, torch.relu(c2) + res, rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(o3, torch.relu(c3), rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(dot.sum(1), torch.relu(c3) @ w, rtol=2e-3, atol=2e-2)


def test_gemm_rejects_bad_arguments():
    A, B = _ops(64, 128, 64)
    with pytest.raises(_lib.SkgError):
        _lib.gemm_bf16(A, B, [_lib.epilogue_group(n_begin=64)])
    with pytest.raises(_lib.SkgError):
        _lib.gemm_bf16(A[:, :60].contiguous(), B[:, :60].contiguous(), [_lib.epilogue_group()])   # pitch 60 % 8 != 0


@pytest.mark.parametrize("M,N,K,kind", [(256, 1024, 12544, "ob"), (128, 1024, 1088, "obof"), (130, 2048, 1024, "of"),
                                        (256, 1024, 1024, "res"), (3968, 118, 4096, "of")])
def test_gemm_split_k(M, N, K, kind):
    """Few-tile shapes (node-row GEMMs of the head) run split-K through a zeroed scratch region that the kernel
    must leave zeroed; results equal the unsplit product up to fp32 summation order."""
    Np = (N + 127) // 128 * 128
    A, B = _ops(M, Np, K, seed=M + N)
    B[N:] = 0
    bias = torch.zeros(Np, device=DEV); bias[:N] = torch.randn(N, device=DEV)
    nbytes = int(_lib.lib().skg_gemm_splitk_scratch_bytes(0, 0))
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    ref = A.float() @ B[:N].float().t() + bias[:N]
    out32 = torch.full((M, N), float("nan"), device=DEV)
    out16 = torch.zeros(M, Np, dtype=torch.bfloat16, device=DEV)
    res = torch.randn(M, Np, device=DEV)
    for rep in range(2):                     # second launch re-uses the self-cleaned scratch
        out32.fill_(float("nan"))
        if kind == "ob":
            g = _lib.epilogue_group(relu=True, bias=bias, out_bf16=out16)
            _lib.gemm_bf16(A, B, [g], N=N, scratch=scratch)
            torch.testing.assert_close(out16[:, :N].float(), torch.relu(ref), rtol=1e-2, atol=1e-2)
        elif kind == "obof":
            g = _lib.epilogue_group(relu=True, bias=bias, out_bf16=out16, out_f32=out32)
            _lib.gemm_bf16(A, B, [g], N=N, scratch=scratch)
            torch.testing.assert_close(out32, torch.relu(ref), rtol=2e-3, atol=2e-3)
            torch.testing.assert_close(out16[:, :N].float(), torch.relu(ref), rtol=1e-2, atol=1e-2)
        elif kind == "of":
            g = _lib.epilogue_group(bias=bias, out_f32=out32)
            _lib.gemm_bf16(A, B, [g], N=N, scratch=scratch)
            torch.testing.assert_close(out32, ref, rtol=2e-3, atol=3e-3)
        else:
            g = _lib.epilogue_group(relu=True, bias=bias, residual=res, out_f32=out32)
            _lib.gemm_bf16(A, B, [g], N=N, scratch=scratch)
            torch.testing.assert_close(out32, torch.relu(ref) + res[:, :N], rtol=2e-3, atol=3e-3)
        assert int(scratch.count_nonzero()) == 0, "split-K scratch not left zeroed"
