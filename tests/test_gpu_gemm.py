"""GPU tests of the linear-layer entry point (bist_linear_fwd) on both GEMM paths: the tcgen05/TMEM/TMA
kernel (bf16 operands, K % 64 == 0; fp32-as-tf32 inside bist_vidproj_fwd) and the CUDA-core kernel,
against torch matmul in fp32 on the same (rounded) operands."""
import pytest
import torch

from bist_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _linear(x, w, bias=None, res=None, relu=0, out_dtype=None):
    out_dtype = out_dtype or x.dtype
    M, K = x.shape
    N = w.shape[0]
    y = torch.empty((M, N), dtype=out_dtype, device=x.device)
    if res is not None:
        assert res.dtype == out_dtype
    _lib.check(_lib.lib().bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(res), _lib.ptr(y), M, N, K,
                                          relu, _lib.dtype_code(x.dtype), _lib.dtype_code(w.dtype),
                                          _lib.dtype_code(out_dtype), _lib.stream_ptr()), "bist_linear_fwd")
    return y


def _ref(x, w, bias, res, relu):
    y = x.float() @ w.float().t()
    if bias is not None:
        y = y + bias
    if relu:
        y = torch.relu(y)
    if res is not None:
        y = y + res.float()
    return y


def _err(a, b):
    return float((a.float() - b).abs().max() / b.abs().max())


@pytest.mark.parametrize("M,N,K", [
    (128, 64, 64), (128, 128, 64), (128, 256, 128), (1024, 512, 512), (1000, 512, 512), (77, 640, 512),
    (4096, 1024, 512), (20000, 512, 2048), (130, 72, 192), (1024, 2048, 512), (1, 64, 64), (300, 1536, 512),
])
def test_tcgen05_bf16_linear(M, N, K):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    x = torch.randn(M, K, generator=g).to(DEV).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(DEV).to(torch.bfloat16)
    bias = torch.randn(N, generator=g).to(DEV)
    res = torch.randn(M, N, generator=g).to(DEV)
    # plain product, fp32 output
    assert _err(_linear(x, w, out_dtype=torch.float32), _ref(x, w, None, None, 0)) < 2e-5
    # bias + relu + residual, fp32 output
    y = _linear(x, w, bias, res, relu=1, out_dtype=torch.float32)
    assert _err(y, _ref(x, w, bias, res, 1)) < 2e-5
    # bf16 output (rounding of the result only)
    yb = _linear(x, w, bias, None, relu=0, out_dtype=torch.bfloat16)
    assert _err(yb, _ref(x, w, bias, None, 0)) < 6e-3


def test_paths_agree_and_launch_counts_move():
    """Same bf16 problem through the tensor-core path (K % 64 == 0) and, shifted by a K that is not
    a multiple of 64, through the CUDA-core path."""
    g = torch.Generator(device="cpu").manual_seed(5)
    for K in (512, 520):
        x = torch.randn(333, K, generator=g).to(DEV).to(torch.bfloat16)
        w = (torch.randn(256, K, generator=g) / K ** 0.5).to(DEV).to(torch.bfloat16)
        n0 = _lib.lib().bist_launch_count()
        y = _linear(x, w, out_dtype=torch.float32)
        assert _lib.lib().bist_launch_count() == n0 + 1
        assert _err(y, _ref(x, w, None, None, 0)) < 2e-5


@pytest.mark.parametrize("N,D,d", [(4096, 2048, 512), (1000, 256, 128), (640, 128, 512)])
def test_vidproj_tf32_and_fp32(N, D, d):
    g = torch.Generator(device="cpu").manual_seed(N + D + d)
    fts = torch.rand(N, D, generator=g).to(DEV)
    W = ((torch.rand(d, D, generator=g) * 2 - 1) * (6.0 / (d + D)) ** 0.5).to(DEV)
    bias = (0.02 * torch.randn(d, generator=g)).to(DEV)
    a2 = (1 + 0.1 * torch.randn(d, generator=g)).to(DEV)
    b2 = (0.1 * torch.randn(d, generator=g)).to(DEV)
    act = torch.relu(fts.double() @ W.double().t() + bias.double())
    mu, sd = act.mean(-1, keepdim=True), act.std(-1, keepdim=True)
    ref = (a2.double() * (act - mu) / (sd + 1e-6) + b2.double()).float()
    L = _lib.lib()
    for cdt, odt, tol in ((_lib.F32, torch.float32, 1e-5), (_lib.BF16, torch.bfloat16, 8e-3)):
        out = torch.empty((N, d), dtype=odt, device=DEV)
        ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, cdt), dtype=torch.uint8, device=DEV)
        _lib.check(L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out),
                                      N, D, d, 1e-6, _lib.ptr(ws), ws.numel(), cdt, _lib.dtype_code(odt),
                                      _lib.stream_ptr()), "bist_vidproj_fwd")
        assert _err(out, ref) < tol, (cdt, _err(out, ref))
