"""Micro-benchmark of bist_linear_fwd (tcgen05 GEMM) on the shapes the BiST layer uses.  GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
shapes = [(40960, 1024, 512), (40960, 512, 512), (16384, 512, 512), (2048, 512, 512), (2048, 1536, 512),
          (2048, 2048, 512), (2048, 512, 2048), (8192, 512, 512), (1024, 512, 64)]
for (M, N, K) in shapes:
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    b = torch.randn(N, device=dev)
    for odt in (torch.bfloat16, torch.float32):
        y = torch.empty(M, N, device=dev, dtype=odt)
        def run():
            _lib.check(L.bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), None, _lib.ptr(y), M, N, K, 0,
                                         _lib.BF16, _lib.BF16, _lib.dtype_code(odt), _lib.stream_ptr()), "linear")
        for _ in range(5):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 50
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / n
        ref = (x.float() @ w.float().t() + b)
        err = float((y.float() - ref).abs().max() / ref.abs().max())
        print("M=%6d N=%5d K=%5d out=%-8s %8.1f us  %7.1f TFLOP/s  err=%.2e" % (M, N, K, str(odt)[6:], us, 2.0 * M * N * K / us / 1e6, err), flush=True)
