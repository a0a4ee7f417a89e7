"""skg_gemm_bf16_splitk_static: the allocation-free split-K GEMM entry point (csrc/noalloc.cu; no counterpart in the
reference).  Called through the C-ABI and checked against the existing
skg_gemm_bf16_splitk (caller-owned scratch) on the same operands, two calls in a row (the second starts from whatever the
first left in the shared static array), within the fp32 tolerance of tests/test_gpu_gemm.py::test_gemm_split_k.

Collected last on purpose (test_zz_*)."""

import pytest
import torch

from skghoi_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops(M, N, K, seed):
    g = torch.Generator().manual_seed(seed)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    B = (torch.randn(N, K, generator=g) * 0.1).to(torch.bfloat16).to(DEV)
    return A, B


def _static(A, B, group, M, N, K, block_n=0):
    arr = (_lib.EpilogueGroup * 1)(group)
    dev = torch.device(DEV)
    with _lib.on_device(dev):
        _lib.call("skg_gemm_bf16_splitk_static", _lib.ptr(A), A.stride(0), _lib.ptr(B), B.stride(0), M, N, K, arr, 1, block_n,
                  _lib.stream_ptr(dev))


@pytest.mark.parametrize("M,N,K,block_n", [(256, 1024, 12544, 0), (1000, 1024, 8192, 256), (130, 512, 4096, 128)])
def test_static_scratch_gemm_matches_caller_scratch_gemm(M, N, K, block_n):
    assert int(_lib.lib().skg_gemm_static_scratch_bytes()) == 4 * 1024 * 1024 + 4096
    A, B = _ops(M, N, K, seed=M + N + K)
    bias = torch.randn(N, device=DEV)
    nbytes = int(_lib.lib().skg_gemm_splitk_scratch_bytes(0, 0))
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    want32 = torch.full((M, N), float("nan"), device=DEV)
    want16 = torch.zeros(M, N, dtype=torch.bfloat16, device=DEV)
    _lib.gemm_bf16(A, B, [_lib.epilogue_group(relu=True, bias=bias, out_f32=want32, out_bf16=want16)], block_n=block_n, scratch=scratch)
    ref = torch.relu(A.float() @ B.float().t() + bias)
    torch.testing.assert_close(want32, ref, rtol=2e-3, atol=3e-3)       # the existing entry point itself is right
    got32 = torch.empty((M, N), device=DEV)
    got16 = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    group = _lib.epilogue_group(relu=True, bias=bias, out_f32=got32, out_bf16=got16)
    torch.cuda.synchronize()
    for call in range(2):                                               # two calls in a row on the shared array
        got32.fill_(float("nan")); got16.zero_()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        torch_before = torch.cuda.memory_reserved(0)
        _static(A, B, group, M, N, K, block_n)
        torch.cuda.synchronize()
        if call == 1:                                                   # (the first call may load the module lazily)
            drop = free_before - torch.cuda.mem_get_info(0)[0]
            if drop > 0:                                                # someone else on the device?  a call that allocates does so every time
                drops = []
                for _ in range(3):
                    before = torch.cuda.mem_get_info(0)[0]
                    _static(A, B, group, M, N, K, block_n)
                    torch.cuda.synchronize()
                    drops.append(before - torch.cuda.mem_get_info(0)[0])
                assert min(drops) <= 0, "the call took device memory: %s" % (drops,)
            assert torch.cuda.memory_reserved(0) == torch_before
        torch.testing.assert_close(got32, want32, rtol=2e-3, atol=3e-3)
        torch.testing.assert_close(got16.float(), want16.float(), rtol=1e-2, atol=1e-2)


def test_static_scratch_gemm_after_a_different_shape():
    """The array keeps the previous call's footprint: a smaller, differently tiled problem right after a larger one."""
    outs = []
    for M, N, K in [(1024, 1024, 4096), (200, 768, 8192), (1024, 1024, 4096)]:
        A, B = _ops(M, N, K, seed=K + M)
        out = torch.full((M, N), float("nan"), device=DEV)
        _static(A, B, _lib.epilogue_group(out_f32=out), M, N, K)
        torch.testing.assert_close(out, A.float() @ B.float().t(), rtol=2e-3, atol=3e-3)
        outs.append(out)
    torch.testing.assert_close(outs[0], outs[2], rtol=2e-3, atol=3e-3)


def test_static_scratch_gemm_rejects_shapes_beyond_the_array():
    A, B = _ops(128, 128, 64, seed=1)                                    # never read: the size check comes first
    out = torch.zeros(1, device=DEV)
    for M, N in [(2048, 1024), (1025, 1024), (1024, 1025), (128, 16384)]:
        with pytest.raises(_lib.SkgError, match="static array holds"):
            _static(A, B, _lib.epilogue_group(out_f32=out), M, N, 64)
    with pytest.raises(_lib.SkgError, match="block_n"):
        _static(A, B, _lib.epilogue_group(out_f32=out), 128, 128, 64, block_n=64)
