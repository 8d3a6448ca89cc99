"""Drop-in mirrors of /root/reference/model/modules.py over libbist_b200.so.

Same constructor signatures, forward signatures and parameter names as the reference
(``LayerNorm`` :20-31, ``SublayerConnection`` :33-44, ``MultiHeadedAttention`` :66-100,
``PositionwiseFeedForward`` :102-113, ``clones`` :10-12), so a reference ``state_dict``
loads with ``strict=True`` and ``model/mtn.py`` can build its model from these classes.
The forward bodies call the C ABI (include/bist_b200.h) on torch's current CUDA stream.

``LayerNorm`` (and the video modules of encoder.py) carry autograd through the library's backward
entry points; ``MultiHeadedAttention`` / ``PositionwiseFeedForward`` / ``SublayerConnection`` used on
their own are still forward-only (eval mode).  Inputs may be float32 or bfloat16 ("bf16
mode": bf16 storage, fp32 accumulation / statistics); parameters stay fp32 masters.
"""
import copy
import ctypes

import torch
import torch.nn as nn

from . import _lib


def clones(module, N):
    "Produce N identical layers (modules.py:10-12)."
    return nn.ModuleList([copy.deepcopy(module) for _ in range(N)])


def _eval_only(m):
    if m.training:
        raise NotImplementedError(
            "%s: this module of the bist_b200 CUDA path is forward-only this round; call .eval()" % type(m).__name__)


def _no_dropout(m, *ps):
    """Training mode is supported by the CUDA path only with dropout probability 0 (dropout kernels land next)."""
    if m.training and any(p > 0 for p in ps):
        raise NotImplementedError(
            "%s: train() with dropout p > 0 is not built yet on the bist_b200 CUDA path; use dropout=0 or .eval() "
            "(gradients flow in both modes)" % type(m).__name__)


def _needs_grad(*tensors):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


class _LayerNormFn(torch.autograd.Function):
    """LayerNorm forward/backward through the C ABI (modules.py:28-31)."""

    @staticmethod
    def forward(ctx, x, a_2, b_2, eps):
        x = x.contiguous()
        d = x.shape[-1]
        y = torch.empty_like(x)
        dt = _lib.dtype_code(x.dtype)
        _lib.check(_lib.lib().bist_ln_fwd(_lib.ptr(x), _lib.ptr(a_2), _lib.ptr(b_2), _lib.ptr(y), x.numel() // d, d,
                                          float(eps), dt, dt, _lib.stream_ptr()), "bist_ln_fwd")
        ctx.save_for_backward(x, a_2)
        ctx.eps = float(eps)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, a_2 = ctx.saved_tensors
        gy = gy.contiguous()
        d = x.shape[-1]
        dx = torch.empty_like(x)
        da = torch.zeros_like(a_2, dtype=torch.float32)
        db = torch.zeros_like(a_2, dtype=torch.float32)
        _lib.check(_lib.lib().bist_ln_bwd(_lib.ptr(gy), _lib.ptr(x), _lib.ptr(a_2), _lib.ptr(dx), _lib.ptr(da),
                                          _lib.ptr(db), x.numel() // d, d, ctx.eps, _lib.dtype_code(gy.dtype),
                                          _lib.dtype_code(x.dtype), _lib.dtype_code(dx.dtype), _lib.stream_ptr()),
                   "bist_ln_bwd")
        return dx, da, db, None


def _ws(nbytes, device):
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def _mask_u8(mask, shape):
    """bool/int mask broadcast to `shape`, as a contiguous uint8 tensor (0 = masked)."""
    m = mask.expand(shape) if tuple(mask.shape) != tuple(shape) else mask
    m = (m != 0) if m.dtype != torch.bool else m
    return m.contiguous().view(torch.uint8)


# Packed-weight caches are keyed on the parameters' version counters, which fused optimizers (torch._fused_adam_)
# do not bump; a global optimizer-step hook therefore advances an epoch that is part of every cache key.
_WEIGHT_EPOCH = [0]


def _on_optimizer_step(*_args, **_kwargs):
    _WEIGHT_EPOCH[0] += 1


try:
    from torch.optim.optimizer import register_optimizer_step_post_hook as _reg_hook
    _reg_hook(_on_optimizer_step)
except ImportError:                                   # very old torch: fall back to invalidating after every backward
    _reg_hook = None


def weights_updated():
    """Call after changing parameters in a way torch cannot see (custom in-place kernels): drops packed caches."""
    _WEIGHT_EPOCH[0] += 1


class _BF16Cache:
    """bf16 copies of fp32 parameters, refreshed when a parameter is updated in place."""

    def __init__(self):
        self._store = {}

    def get(self, p):
        key = id(p)
        hit = self._store.get(key)
        ver = (p._version, _WEIGHT_EPOCH[0])
        if hit is None or hit[0] != ver or hit[1].device != p.device:
            hit = (ver, p.detach().to(torch.bfloat16).contiguous())
            self._store[key] = hit
        return hit[1]


class LayerNorm(nn.Module):
    "modules.py:20-31 — a_2 * (x - mean) / (std + eps) + b_2, unbiased std, eps added to the std."

    def __init__(self, features, eps=1e-6):
        super(LayerNorm, self).__init__()
        self.a_2 = nn.Parameter(torch.ones(features))
        self.b_2 = nn.Parameter(torch.zeros(features))
        self.eps = eps

    def forward(self, x):
        _lib.require_cuda(x, self.a_2)
        if _needs_grad(x, self.a_2, self.b_2):
            return _LayerNormFn.apply(x, self.a_2, self.b_2, self.eps)
        x = x.detach().contiguous()
        d = x.shape[-1]
        y = torch.empty_like(x)
        dt = _lib.dtype_code(x.dtype)
        _lib.check(_lib.lib().bist_ln_fwd(_lib.ptr(x), _lib.ptr(self.a_2), _lib.ptr(self.b_2), _lib.ptr(y),
                                          x.numel() // d, d, float(self.eps), dt, dt, _lib.stream_ptr()),
                   "bist_ln_fwd")
        return y


class SublayerConnection(nn.Module):
    "modules.py:33-44 — pre-norm residual: x + dropout(sublayer(norm(x)))."

    def __init__(self, size, dropout):
        super(SublayerConnection, self).__init__()
        self.norm = LayerNorm(size)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x, sublayer):
        _eval_only(self)
        return x + sublayer(self.norm(x))


class MultiHeadedAttention(nn.Module):
    "modules.py:66-100."

    def __init__(self, h, d_model, d_in=-1, dropout=0.1):
        super(MultiHeadedAttention, self).__init__()
        assert d_model % h == 0
        self.d_k = d_model // h
        self.h = h
        if d_in < 0:
            d_in = d_model
        self.linears = clones(nn.Linear(d_in, d_model), 3)
        self.linears.append(nn.Linear(d_model, d_in))
        self.attn = None
        self.dropout = nn.Dropout(p=dropout)
        self.store_attn = True      # the reference always keeps p_attn (generator.py:62-63 reads it)
        self._bf16 = _BF16Cache()

    def __deepcopy__(self, memo):
        new = type(self).__new__(type(self))
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = _BF16Cache() if k == "_bf16" else copy.deepcopy(v, memo)
        return new

    def forward(self, query, key, value, mask=None):
        _eval_only(self)
        _lib.require_cuda(query, key, value, self.linears[0].weight)
        dt = _lib.dtype_code(query.dtype)
        query, key, value = (t.detach().contiguous() for t in (query, key, value))
        Bp, Lq, d_in = query.shape
        Lk = key.shape[1]
        d = self.h * self.d_k
        P = _lib.MhaParams()
        keep = []
        for i, lin in enumerate(self.linears):
            w = lin.weight.detach() if dt == _lib.F32 else self._bf16.get(lin.weight)
            keep.append(w)
            P.w[i] = w.data_ptr()
            P.b[i] = lin.bias.data_ptr()
        mask_mode, m8 = _lib.MASK_NONE, None
        if mask is not None:
            if mask.dim() == 3 and mask.shape[1] == 1:
                mask_mode, m8 = _lib.MASK_KEY, _mask_u8(mask, (Bp, 1, Lk))
            else:
                mask_mode, m8 = _lib.MASK_FULL, _mask_u8(mask, (Bp, Lq, Lk))
        out = torch.empty((Bp, Lq, d_in), dtype=query.dtype, device=query.device)
        attn = torch.empty((Bp, self.h, Lq, Lk), dtype=torch.float32, device=query.device) if self.store_attn else None
        L = _lib.lib()
        ws = _ws(L.bist_mha_workspace_bytes(Bp, Lq, Lk, d, d_in, self.h), query.device)
        _lib.check(L.bist_mha_fwd(_lib.ptr(query), _lib.ptr(key), _lib.ptr(value), _lib.ptr(m8), mask_mode,
                                  ctypes.byref(P), _lib.ptr(out), _lib.ptr(attn), Bp, Lq, Lk, d, d_in, self.h,
                                  _lib.ptr(ws), ws.numel(), dt, _lib.stream_ptr()), "bist_mha_fwd")
        self.attn = attn if attn is None or dt == _lib.F32 else attn.to(query.dtype)
        return out


class PositionwiseFeedForward(nn.Module):
    "modules.py:102-113 — w_2(dropout(relu(w_1 x)))."

    def __init__(self, d_model, d_ff, dropout=0.1, d_out=-1):
        super(PositionwiseFeedForward, self).__init__()
        self.w_1 = nn.Linear(d_model, d_ff)
        if d_out < 0:
            d_out = d_model
        self.w_2 = nn.Linear(d_ff, d_out)
        self.dropout = nn.Dropout(dropout)
        self._bf16 = _BF16Cache()

    __deepcopy__ = MultiHeadedAttention.__deepcopy__

    def forward(self, x):
        _eval_only(self)
        _lib.require_cuda(x, self.w_1.weight)
        dt = _lib.dtype_code(x.dtype)
        x = x.detach().contiguous()
        d, d_ff, d_out = self.w_1.in_features, self.w_1.out_features, self.w_2.out_features
        rows = x.numel() // d
        w1 = self.w_1.weight.detach() if dt == _lib.F32 else self._bf16.get(self.w_1.weight)
        w2 = self.w_2.weight.detach() if dt == _lib.F32 else self._bf16.get(self.w_2.weight)
        y = torch.empty(x.shape[:-1] + (d_out,), dtype=x.dtype, device=x.device)
        ws = _ws(rows * d_ff * x.element_size() + 256, x.device)
        _lib.check(_lib.lib().bist_ffn_fwd(_lib.ptr(x), _lib.ptr(w1), _lib.ptr(self.w_1.bias), _lib.ptr(w2),
                                           _lib.ptr(self.w_2.bias), _lib.ptr(y), rows, d, d_ff, d_out,
                                           _lib.ptr(ws), ws.numel(), dt, _lib.stream_ptr()), "bist_ffn_fwd")
        return y
