"""bist_linear_bwd (dw/db) vs torch on shapes that take the tcgen05 weight-gradient kernel."""
import sys, torch
sys.path.insert(0, ".")
from bist_b200 import _lib
L = _lib.lib()
dev = "cuda:0"
torch.manual_seed(0)
for (M, N, K) in [(256, 128, 256), (1024, 512, 512), (4096, 64, 512), (40960, 512, 2048), (16384, 1536, 512), (640, 128, 512)]:
    dy = torch.randn(M, N, device=dev).to(torch.bfloat16)
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    dw = torch.zeros(N, K, device=dev); db = torch.zeros(N, device=dev)
    _lib.check(L.bist_linear_bwd(_lib.ptr(dy), _lib.ptr(x), None, None, _lib.ptr(dw), _lib.ptr(db), M, N, K, 1, 1, 1, 1, _lib.stream_ptr()), "lin_bwd")
    torch.cuda.synchronize()
    ref = dy.float().t() @ x.float()
    rb = dy.float().sum(0)
    e = float((dw - ref).abs().max() / ref.abs().max()); eb = float((db - rb).abs().max() / rb.abs().max())
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(10):
        L.bist_linear_bwd(_lib.ptr(dy), _lib.ptr(x), None, None, _lib.ptr(dw), _lib.ptr(db), M, N, K, 1, 1, 1, 1, _lib.stream_ptr())
    ev[1].record(); torch.cuda.synchronize()
    us = ev[0].elapsed_time(ev[1]) * 100
    print("M=%6d N=%5d K=%5d  dw err %.2e  db err %.2e  %8.1f us  %7.1f TFLOP/s" % (M, N, K, e, eb, us, 2.0 * M * N * K / us / 1e6))
