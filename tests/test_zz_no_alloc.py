"""bist_vidproj_fwd_noalloc (include/bist_b200.h): the bf16-mode projection with its scratch in a static device array inside
the library, for callers that cannot allocate device memory.  Checked through the C ABI against the existing entry point
bist_vidproj_fwd on the same inputs, on two calls in a row (the scratch keeps the previous call's contents), plus the size
limit.  This file is named so that pytest collects it last."""
import ctypes

import pytest
import torch

from bist_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-2          # bf16 mode, normalised max error (tests/test_gpu_parity.py)


def _inputs(seed, N, D, d):
    g = torch.Generator().manual_seed(seed)
    fts = torch.rand(N, D, generator=g).to(DEV)
    W = (torch.randn(d, D, generator=g) / D ** 0.5).to(DEV)
    bias = (torch.randn(d, generator=g) * 0.1).to(DEV)
    a2 = (1 + 0.1 * torch.randn(d, generator=g)).to(DEV)
    b2 = (0.1 * torch.randn(d, generator=g)).to(DEV)
    return fts, W, bias, a2, b2


def _existing(L, fts, W, bias, a2, b2, N, D, d, odt):
    out = torch.full((N, d), float("nan"), dtype=odt, device=DEV)
    ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, _lib.BF16), dtype=torch.uint8, device=DEV)
    _lib.check(L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D, d,
                                  1e-6, _lib.ptr(ws), ws.numel(), _lib.BF16, _lib.dtype_code(odt), _lib.stream_ptr()), "bist_vidproj_fwd")
    return out


def _nerr(a, b):
    a, b = a.float(), b.float()
    assert torch.isfinite(a).all()
    return float((a - b).abs().max() / b.abs().max())


def test_noalloc_matches_workspace_entry_point_twice():
    L = _lib.lib()
    # two calls in a row with different weights AND different shapes: the second must not see the first one's scratch
    for seed, (N, D, d), odt in ((1, (1000, 2048, 512), torch.bfloat16), (2, (333, 256, 128), torch.float32)):
        fts, W, bias, a2, b2 = _inputs(seed, N, D, d)
        want = _existing(L, fts, W, bias, a2, b2, N, D, d, odt)
        out = torch.full((N, d), float("nan"), dtype=odt, device=DEV)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = L.bist_vidproj_fwd_noalloc(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D, d,
                                        1e-6, _lib.dtype_code(odt), _lib.stream_ptr())
        _lib.check(rc, "bist_vidproj_fwd_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert _nerr(out, want) <= TOL
        assert torch.equal(out, want)       # same kernels on the same operands: bit-equal, not just within tolerance


def test_noalloc_rejects_sizes_beyond_the_static_array():
    L = _lib.lib()
    N, D, d = 128, 4096, 512                # d*D*2 = 4 MiB > the 2 MiB array
    fts, W, bias, a2, b2 = _inputs(3, N, D, d)
    out = torch.zeros(N, d, dtype=torch.bfloat16, device=DEV)
    rc = L.bist_vidproj_fwd_noalloc(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D, d,
                                    1e-6, _lib.BF16, _lib.stream_ptr())
    assert rc == 3                          # BIST_ERR_WORKSPACE
    assert b"exceed" in L.bist_last_error()
    # a shape the fused kernel does not take (d % 64 != 0) needs the [N, d] workspace of the two-kernel path
    N, D, d = 64, 128, 96
    fts, W, bias, a2, b2 = _inputs(4, N, D, d)
    out = torch.zeros(N, d, dtype=torch.bfloat16, device=DEV)
    rc = L.bist_vidproj_fwd_noalloc(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D, d,
                                    1e-6, _lib.BF16, _lib.stream_ptr())
    assert rc == 2                          # BIST_ERR_UNSUPPORTED
    torch.cuda.synchronize()
