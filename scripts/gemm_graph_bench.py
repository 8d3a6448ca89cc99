"""Time of one small product inside a CUDA graph: a chain of N identical bist_linear_fwd launches (each reads the previous one's
output) is captured once and replayed; per-launch time = replay time / N.  Shows the fixed cost of a launch of gemm_tc_kernel
(set-up, PDL hand-over, first TMA round trip, epilogue, tear-down) that the eager-launch numbers of gemm_bench.py hide.
Usage: python scripts/gemm_graph_bench.py [M N K]..."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib
L = _lib.lib(); dev = "cuda:0"
shapes = [(2048, 512, 512), (2048, 2048, 512), (2048, 512, 2048), (768, 512, 512), (4096, 512, 512), (1024, 1536, 512)]
if len(sys.argv) >= 4:
    shapes = [tuple(int(v) for v in sys.argv[1:4])]
NCH = 64
for (M, N, K) in shapes:
    x = [torch.randn(M, K, device=dev).to(torch.bfloat16) for _ in range(2)]
    w = torch.randn(N, K, device=dev).to(torch.bfloat16) / K ** 0.5
    b = torch.randn(N, device=dev)
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    def call():
        _lib.check(L.bist_linear_fwd(_lib.ptr(x[0]), _lib.ptr(w), _lib.ptr(b), None, _lib.ptr(y), M, N, K, 0, _lib.BF16, _lib.BF16,
                                     _lib.BF16, _lib.stream_ptr()), "linear")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(NCH):
                call()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(10):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (10 * NCH)
    print("M=%5d N=%5d K=%5d  %.2f us per launch in a graph chain  (%.0f TFLOP/s)" % (M, N, K, us, 2.0 * M * N * K / us / 1e6), flush=True)
