import ctypes, os, sys
sys.path.insert(0, "/root/repo")
import torch
from bist_b200 import _lib
L = _lib.lib(); dev = "cuda:0"
M, N, K = 2048, 512, 512
x = torch.randn(M, K, device=dev).to(torch.bfloat16); w = torch.randn(N, K, device=dev).to(torch.bfloat16); b = torch.randn(N, device=dev)
y = torch.empty(M, N, device=dev, dtype=torch.float32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(20):
    _lib.check(L.bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), None, _lib.ptr(y), M, N, K, 0, _lib.BF16, _lib.BF16, _lib.F32, _lib.stream_ptr()), "linear")
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (3 * 64))()
L.bist_debug_gemm_trace.argtypes = [ctypes.c_void_p]
assert L.bist_debug_gemm_trace(buf) == 0
tr = [[buf[r * 64 + i] for i in range(64)] for r in range(3)]
t0 = tr[2][60]
print("start 0, setup done %d, prev done %d" % (tr[2][61] - t0, tr[2][62] - t0))
print("tma issued:", [v - t0 for v in tr[0][:8]])
print("mma committed:", [v - t0 for v in tr[1][:8]])
print("epilogue start/end:", tr[2][0] - t0, tr[2][1] - t0)
