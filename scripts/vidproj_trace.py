"""Timeline of CTA 0 of the fused projection kernel over its first 64 slabs (needs a library built with -DBIST_GEMM_TRACE):
clock64 stamps of: feature slab requested (producer), slab landed (worker warp 2), slab converted, UMMAs of the slab committed."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bist_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
N, D, d = 40960, 2048, 512
fts = torch.rand(N, D, device=dev)
W = torch.randn(d, D, device=dev) / D ** 0.5
bias, a2, b2 = torch.zeros(d, device=dev), torch.ones(d, device=dev), torch.zeros(d, device=dev)
out = torch.empty(N, d, dtype=torch.bfloat16, device=dev)
ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, _lib.BF16), dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(30):
    flush.zero_()
    _lib.check(L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D, d, 1e-6,
                                  _lib.ptr(ws), ws.numel(), _lib.BF16, _lib.BF16, _lib.stream_ptr()), "vidproj")
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (4 * 64))()
L.bist_debug_vp_trace.argtypes = [ctypes.c_void_p]
assert L.bist_debug_vp_trace(buf) == 0
tr = [[buf[r * 64 + i] for i in range(64)] for r in range(4)]
t0 = min(x for r in tr for x in r if x > 0)
print("slab  requested  landed  converted  committed   (clk since first stamp; per-slab deltas of 'committed' at the end)")
for i in range(64):
    print("%4d  %9d %7d %10d %10d" % (i, tr[0][i] - t0, tr[2][i] - t0, tr[3][i] - t0, tr[1][i] - t0))
c = [tr[1][i] for i in range(64) if tr[1][i] > 0]
print("committed deltas:", " ".join(str(c[i + 1] - c[i]) for i in range(len(c) - 1)))
