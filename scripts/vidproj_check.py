"""Check + timing of the fused projection kernel (bist_vidproj_fwd in bf16 mode: csrc/vidproj_tc.cu) against a torch
restatement of VidEncoder8's video branch (/root/reference/model/encoder.py:72-80, LayerNorm of modules.py:20-31).
GPU box only.  Usage: python scripts/vidproj_check.py [--time]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bist_b200 import _lib

L = _lib.lib()
dev = "cuda:0"


def ref(fts, W, bias, a2, b2, eps):
    x = torch.relu(fts.to(torch.bfloat16).float() @ W.to(torch.bfloat16).float().t() + bias)
    mean, std = x.mean(-1, keepdim=True), x.std(-1, keepdim=True)
    return a2 * (x - mean) / (std + eps) + b2


def run(N, D, d, time=False):
    g = torch.Generator().manual_seed(N + D + d)
    fts = torch.rand(N, D, generator=g).to(dev)
    W = (torch.randn(d, D, generator=g) / D ** 0.5).to(dev)
    bias, a2, b2 = (torch.randn(d, generator=g) * 0.1).to(dev), (1 + 0.1 * torch.randn(d, generator=g)).to(dev), (0.1 * torch.randn(d, generator=g)).to(dev)
    out = torch.full((N, d), float("nan"), dtype=torch.bfloat16, device=dev)
    ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, _lib.BF16), dtype=torch.uint8, device=dev)

    def call():
        _lib.check(L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out),
                                      N, D, d, 1e-6, _lib.ptr(ws), ws.numel(), _lib.BF16, _lib.BF16, _lib.stream_ptr()), "bist_vidproj_fwd")
    call()
    torch.cuda.synchronize()
    r = ref(fts, W, bias, a2, b2, 1e-6)
    err = float((out.float() - r).abs().max() / r.abs().max()) if torch.isfinite(out.float()).all() else float("inf")
    msg = "N=%6d D=%5d d=%4d  err %.3e" % (N, D, d, err)
    if time:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for _ in range(20):
            call()
        L.bist_prof_enable(1)
        for _ in range(10):
            flush.zero_()               # evict the features from L2 between launches
            call()
        torch.cuda.synchronize()
        buf = ctypes.create_string_buffer(1 << 16)
        L.bist_prof_collect(buf, len(buf))
        L.bist_prof_enable(0)
        for ln in buf.value.decode().strip().splitlines():
            tag, n, tot = ln.split(",")
            us = 1e3 * float(tot) / int(n)
            msg += "   %s %.1f us" % (tag, us)
            if tag.startswith("vidproj"):
                msg += " = %.0f TFLOP/s, %.0f GB/s of fp32 features" % (2.0 * N * D * d / us / 1e6, N * D * 4 / us / 1e3)
    print(msg, flush=True)
    return err


if __name__ == "__main__":
    bad = 0
    for (N, D, d) in [(300, 128, 64), (128, 256, 128), (1000, 512, 256), (2049, 2048, 512), (40960, 2048, 512), (768, 128, 512), (33, 384, 320)]:
        bad += run(N, D, d, "--time" in sys.argv and N >= 40960) > 2e-2
    print("FAIL" if bad else "OK")
    sys.exit(1 if bad else 0)
