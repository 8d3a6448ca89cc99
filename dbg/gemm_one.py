import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib
L = _lib.lib(); dev = "cuda:0"
M, N, K = 40960, 1024, 512
x = torch.randn(M, K, device=dev).to(torch.bfloat16); w = torch.randn(N, K, device=dev).to(torch.bfloat16)
y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    _lib.check(L.bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), None, None, _lib.ptr(y), M, N, K, 0, _lib.BF16, _lib.BF16, _lib.BF16, _lib.stream_ptr()), "linear")
torch.cuda.synchronize()
