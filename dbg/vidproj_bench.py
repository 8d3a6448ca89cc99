"""Time bist_vidproj_fwd (fused fp32-in projection + LayerNorm) at config 2 and check it against torch.  GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
for (N, D, d) in ((40960, 2048, 512), (40960, 1024, 512), (40960, 512, 512), (40960, 2048, 256), (40960, 1024, 256), (18944, 2048, 512), (18944, 1024, 512)):
    g = torch.Generator(device="cpu").manual_seed(1)
    fts = torch.rand(N, D, generator=g).to(dev)
    W = ((torch.rand(d, D, generator=g) * 2 - 1) * (6.0 / (d + D)) ** 0.5).to(dev)
    bias = (0.02 * torch.randn(d, generator=g)).to(dev)
    a2 = (1 + 0.1 * torch.randn(d, generator=g)).to(dev)
    b2 = (0.1 * torch.randn(d, generator=g)).to(dev)
    out = torch.empty((N, d), dtype=torch.bfloat16, device=dev)
    ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, _lib.BF16), dtype=torch.uint8, device=dev)

    def run():
        _lib.check(L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out),
                                      N, D, d, 1e-6, _lib.ptr(ws), ws.numel(), _lib.BF16, _lib.dtype_code(torch.bfloat16),
                                      _lib.stream_ptr()), "bist_vidproj_fwd")
    for _ in range(5):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 30
    e0.record()
    for _ in range(n):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    act = torch.relu(fts[:2048].to(torch.bfloat16).float() @ W.to(torch.bfloat16).float().t() + bias)
    mu, sd = act.mean(-1, keepdim=True), act.std(-1, keepdim=True)
    ref = a2 * (act - mu) / (sd + 1e-6) + b2
    err = float((out[:2048].float() - ref).abs().max() / ref.abs().max())
    print("N=%d D=%d d=%d  %7.1f us (incl. weight pack)  %6.1f TFLOP/s  %6.1f GB/s in  err=%.2e" %
          (N, D, d, us, 2.0 * N * D * d / us / 1e6, N * D * 4 / us / 1e3, err), flush=True)
