"""A few launches of the large plain product of the layer (M = 40960, N = 1024, K = 512, bf16: CTA-pair form, TMA-store epilogue)
for ncu.  Usage: python scripts/gemm_prof.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib
L = _lib.lib(); dev = "cuda:0"
M, N, K = 40960, 1024, 512
x = torch.randn(M, K, device=dev).to(torch.bfloat16); w = (torch.randn(N, K, device=dev) / K ** 0.5).to(torch.bfloat16)
b = torch.randn(N, device=dev); y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    _lib.check(L.bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), None, _lib.ptr(y), M, N, K, 0, _lib.BF16, _lib.BF16, _lib.BF16,
                                 _lib.stream_ptr()), "linear")
torch.cuda.synchronize()
ref = x.float() @ w.float().t() + b
print("ok err %.3e" % float((y.float() - ref).abs().max() / ref.abs().max()))
