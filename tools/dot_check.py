import sys; sys.path.insert(0, '/root/repo')
import torch
from skghoi_b200 import _lib
DEV='cuda:0'
g = torch.Generator(device=DEV).manual_seed(2)
M, R = 300, 1024
A = torch.randn(M, R, device=DEV, generator=g).to(torch.bfloat16)
W = (torch.randn(R, R, device=DEV, generator=g) / 32).to(torch.bfloat16)
bias = torch.randn(R, device=DEV, generator=g)
w = torch.randn(R, device=DEV, generator=g)
ref = torch.relu(A.float() @ W.float().t() + bias) @ w
for variant in ("dot", "dot+ob", "dot+of"):
    dot = torch.full((M, 16), 9.0, device=DEV)
    kw = {}
    if variant == "dot+ob": kw["out_bf16"] = torch.empty(M, R, dtype=torch.bfloat16, device=DEV)
    if variant == "dot+of": kw["out_f32"] = torch.empty(M, R, device=DEV)
    _lib.gemm_bf16(A, W, [_lib.epilogue_group(relu=True, bias=bias, dot_w=w, dot_out=dot, **kw)], block_n=256)
    torch.cuda.synchronize()
    print(variant, "max err", float((dot.sum(1) - ref).abs().max()), "ref max", float(ref.abs().max()), dot[0, :4].tolist())
