"""Timeline of CTA 0 of gemm_tc_kernel (needs a library whose gemm_tc.cu was compiled with -DBIST_GEMM_TRACE): clock64 stamps of the
first 64 TMA requests and UMMA commits, and the start / end of warp 2's epilogue of each tile.
Usage: python scripts/gemm_trace.py [M N K [bf16|f32]]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib
L = _lib.lib(); dev = "cuda:0"
M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (2048, 512, 512)
odt = torch.bfloat16 if len(sys.argv) >= 5 and sys.argv[4] == "bf16" else torch.float32
x = torch.randn(M, K, device=dev).to(torch.bfloat16); w = torch.randn(N, K, device=dev).to(torch.bfloat16); b = torch.randn(N, device=dev)
y = torch.empty(M, N, device=dev, dtype=odt)
for _ in range(20):
    _lib.check(L.bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), None if "nobias" in sys.argv else _lib.ptr(b), None, _lib.ptr(y), M, N, K, 0, _lib.BF16, _lib.BF16, _lib.dtype_code(odt), _lib.stream_ptr()), "linear")
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (3 * 64))()
L.bist_debug_gemm_trace.argtypes = [ctypes.c_void_p]
assert L.bist_debug_gemm_trace(buf) == 0
tr = [[buf[r * 64 + i] for i in range(64)] for r in range(3)]
t0 = tr[2][60]
print("M=%d N=%d K=%d out=%s" % (M, N, K, odt))
print("start 0, setup done %d, prev done %d" % (tr[2][61] - t0, tr[2][62] - t0))
print("tma issued:", [v - t0 for v in tr[0][:40]])
print("mma committed:", [v - t0 for v in tr[1][:40]])
print("epilogue (start, end) per tile:", [(tr[2][2 * i] - t0, tr[2][2 * i + 1] - t0) for i in range(12) if tr[2][2 * i + 1] > t0])
print("third tile, warp 2, per 32-column chunk (top, tmem data in registers, staged + syncwarp, smem loads issued, chunk end) relative to the tile's epilogue start:")
for c in range(4):
    print("  chunk %d:" % c, [tr[2][24 + 5 * c + i] - tr[2][4] for i in range(5)])
