"""Drop-in mirrors of /root/reference/model/modules.py over libbist_b200.so.

Same constructor signatures, forward signatures and parameter names as the reference
(``LayerNorm`` :20-31, ``SublayerConnection`` :33-44, ``MultiHeadedAttention`` :66-100,
``PositionwiseFeedForward`` :102-113, ``clones`` :10-12), so a reference ``state_dict``
loads with ``strict=True`` and ``model/mtn.py`` can build its model from these classes.
The forward bodies call the C ABI (include/bist_b200.h) on torch's current CUDA stream.

Every module carries autograd through the library's backward entry points (``bist_ln_bwd``, ``bist_mha_bwd``,
``bist_ffn_bwd``, ``bist_dropout_fwd``) and honours ``train()`` mode: dropout at the reference's three sites
(modules.py:44,62-63,113) with counter-based masks, a fresh seed per call drawn from torch's CPU generator
(``torch.manual_seed`` reproduces a run; ``module.last_dropout`` = (p, seed, site) lets a test rebuild the mask
with ``bist_dropout_mask``).  Inputs may be float32 or bfloat16 ("bf16 mode": bf16 storage, fp32 accumulation /
statistics); parameters stay fp32 masters and their gradients are fp32.
"""
import copy
import ctypes
import weakref

import torch
import torch.nn as nn

from . import _lib
from . import arena as _arena


def clones(module, N):
    "Produce N identical layers (modules.py:10-12)."
    return nn.ModuleList([copy.deepcopy(module) for _ in range(N)])


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
        dab = _arena.zeros((2,) + tuple(a_2.shape), a_2.device)
        da, db = dab[0], dab[1]
        _lib.check(_lib.lib().bist_ln_bwd(_lib.ptr(gy), _lib.ptr(x), _lib.ptr(a_2), _lib.ptr(dx), _lib.ptr(da),
                                          _lib.ptr(db), x.numel() // d, d, ctx.eps, _lib.dtype_code(gy.dtype),
                                          _lib.dtype_code(x.dtype), _lib.dtype_code(dx.dtype), _lib.stream_ptr()),
                   "bist_ln_bwd")
        return dx, da, db, None


def _bucket_block(block, params):
    """Hand a module's gradient block to the bucketed all-reduce (dist.grad_block_ready) when autograd is going to ADOPT its slices
    as the parameters' .grad (all of them still None); a module applied twice in one step accumulates instead and is left to the
    final all-reduce."""
    from . import dist as _dist
    if _dist._pending is not None and all(p.grad is None for p in params):
        _dist.grad_block_ready(block, bucket=True)


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


class _CopyArena:
    """Compute-dtype (and transposed) copies of fp32 weight matrices for one (device, dtype), kept in ONE buffer and
    refreshed TOGETHER — ceil(n / 72) launches of ``bist_pack_params`` after an optimizer step instead of one or two
    torch kernels per copy (a full model asks for ~200 copies per step).  A copy requested for the first time in the
    middle of a step is made on the spot and joins the arena at the next refresh."""

    def __init__(self, device, dtype):
        self.device, self.dtype = device, dtype
        self.slots = {}              # (id(param), transposed) -> [weakref, offset, rows, cols, version, view]
        self.size, self.buf = 0, None
        self.refreshed_epoch = None
        self.stale_layout = True
        self.extra = {}

    def get(self, p, transposed):
        key = (id(p), transposed)
        slot = self.slots.get(key)
        if slot is None or slot[0]() is not p:
            rows, cols = (p.shape[1], p.shape[0]) if transposed else (p.shape[0], p.shape[1])
            slot = [weakref.ref(p), 0, rows, cols, None, None]
            self.slots[key] = slot
            self.stale_layout = True
            self.extra.pop(key, None)        # ids are recycled: never hand out a dead parameter's copy
        cur = _WEIGHT_EPOCH[0]
        if self.refreshed_epoch == cur and slot[5] is not None and slot[4] == p._version:
            return slot[5]
        if self.refreshed_epoch == cur and slot[5] is None:          # new this step: one-off copy, arena next time
            e = self.extra.get(key)
            if e is None or e[0] != p._version:
                w = p.detach().t() if transposed else p.detach()
                e = (p._version, w.to(self.dtype).contiguous())
                self.extra[key] = e
            return e[1]
        self.refresh()
        return self.slots[key][5]

    def refresh(self):
        live = {k: s for k, s in self.slots.items() if s[0]() is not None}
        if len(live) != len(self.slots):
            self.slots, self.stale_layout = live, True
        if not self.slots:
            self.refreshed_epoch = _WEIGHT_EPOCH[0]
            return
        if self.stale_layout or self.buf is None:
            off = 0
            for s in self.slots.values():
                s[1] = off
                off += (s[2] * s[3] + 63) // 64 * 64                # 128-byte aligned slots
            self.size = off
            self.buf = torch.empty(max(off, 64), dtype=self.dtype, device=self.device)
            self.stale_layout = False
        items = (_lib.PackItem * len(self.slots))()
        for it, (key, s) in zip(items, self.slots.items()):
            p = s[0]()
            it.src, it.dst_off, it.rows, it.cols, it.transpose = p.data_ptr(), s[1], s[2], s[3], int(key[1])
            s[4] = p._version
            s[5] = self.buf[s[1]:s[1] + s[2] * s[3]].view(s[2], s[3])
        _lib.check(_lib.lib().bist_pack_params(items, len(self.slots), _lib.ptr(self.buf), _lib.dtype_code(self.dtype),
                                               _lib.stream_ptr()), "bist_pack_params")
        self.refreshed_epoch = _WEIGHT_EPOCH[0]
        self.extra.clear()


_ARENAS = {}


def refresh_weight_copies(device):
    """Bring the shared weight-copy arenas of `device` up to date on the CURRENT stream.  Callers that are about to run
    modules on several streams do this first, so that no stream reads a slot another stream is still writing."""
    for (dev, _dtype), arena in _ARENAS.items():
        if dev != device or not arena.slots:
            continue
        stale = arena.refreshed_epoch != _WEIGHT_EPOCH[0]
        if not stale:                                # a parameter changed in place without an optimizer step
            for s in arena.slots.values():
                p = s[0]()
                if p is not None and s[5] is not None and s[4] != p._version:
                    stale = True
                    break
        if stale:
            arena.refresh()


class _BF16Cache:
    """Copies of fp32 parameters in the compute dtype (bf16 copies; transposed copies for the backward GEMMs),
    refreshed when a parameter is updated in place.  Weight matrices on a CUDA device live in the shared ``_CopyArena``
    of their (device, dtype); anything else is copied on its own."""

    def __init__(self):
        self._store = {}

    def get(self, p, dtype=torch.bfloat16, transposed=False):
        if p.is_cuda and p.dim() == 2 and p.dtype == torch.float32 and p.is_contiguous():
            ak = (p.device, dtype)
            arena = _ARENAS.get(ak)
            if arena is None:
                arena = _ARENAS[ak] = _CopyArena(p.device, dtype)
            return arena.get(p, transposed)
        key = (id(p), dtype, transposed)
        hit = self._store.get(key)
        ver = (p._version, _WEIGHT_EPOCH[0])
        if hit is None or hit[0] != ver or hit[1].device != p.device:
            w = p.detach().t() if transposed else p.detach()
            hit = (ver, w.to(dtype).contiguous())
            self._store[key] = hit
        return hit[1]


# dropout sites of the stand-alone modules (the video path's sites 0-10 are listed in include/bist_b200.h)
SITE_MHA, SITE_FFN, SITE_SUBLAYER = 100, 101, 102


def _draw_seed():
    return int(torch.randint(0, 2 ** 62, (1,)).item())


class _DropoutFn(torch.autograd.Function):
    """y = x o m / (1 - p) with the library's counter-based mask; the backward is the same call on dy."""

    @staticmethod
    def forward(ctx, x, p, seed, site):
        x = x.contiguous()
        d = x.shape[-1]
        y = torch.empty_like(x)
        _lib.check(_lib.lib().bist_dropout_fwd(_lib.ptr(x), _lib.ptr(y), x.numel() // d, d, float(p), seed, site,
                                               _lib.dtype_code(x.dtype), _lib.stream_ptr()), "bist_dropout_fwd")
        ctx.cfg = (float(p), seed, site)
        return y

    @staticmethod
    def backward(ctx, gy):
        gy = gy.contiguous()
        d = gy.shape[-1]
        gx = torch.empty_like(gy)
        p, seed, site = ctx.cfg
        _lib.check(_lib.lib().bist_dropout_fwd(_lib.ptr(gy), _lib.ptr(gx), gy.numel() // d, d, p, seed, site,
                                               _lib.dtype_code(gy.dtype), _lib.stream_ptr()), "bist_dropout_fwd")
        return gx, None, None, None


class _DropoutAddFn(torch.autograd.Function):
    """x + dropout(y) in one kernel (modules.py:44); backward: dx = g, dy = g o m."""

    @staticmethod
    def forward(ctx, y, x, p, seed, site):
        y, x = y.contiguous(), x.contiguous()
        d = x.shape[-1]
        out = torch.empty_like(x)
        _lib.check(_lib.lib().bist_dropout_add_fwd(_lib.ptr(y), _lib.ptr(x), _lib.ptr(out), x.numel() // d, d, float(p), seed,
                                                   site, _lib.dtype_code(x.dtype), _lib.stream_ptr()), "bist_dropout_add_fwd")
        ctx.cfg = (float(p), seed, site)
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        d = g.shape[-1]
        gy = torch.empty_like(g)
        p, seed, site = ctx.cfg
        _lib.check(_lib.lib().bist_dropout_fwd(_lib.ptr(g), _lib.ptr(gy), g.numel() // d, d, p, seed, site,
                                               _lib.dtype_code(g.dtype), _lib.stream_ptr()), "bist_dropout_fwd")
        return gy, g, None, None, None


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


class _MhaFn(torch.autograd.Function):
    """MultiHeadedAttention forward/backward (modules.py:81-100) — bist_mha_train_fwd / bist_mha_bwd."""

    @staticmethod
    def forward(ctx, mod, query, key, value, m8, mask_mode, p_drop, seed, *params):
        dt = _lib.dtype_code(query.dtype)
        query, key, value = query.contiguous(), key.contiguous(), value.contiguous()
        Bp, Lq, d_in = query.shape
        Lk = key.shape[1]
        d = mod.h * mod.d_k
        P, keep = mod._param_struct(dt, transposed=False)
        out = torch.empty((Bp, Lq, d_in), dtype=query.dtype, device=query.device)
        attn = torch.empty((Bp, mod.h, Lq, Lk), dtype=torch.float32, device=query.device) if mod.store_attn else None
        L = _lib.lib()
        tape = _ws(L.bist_mha_workspace_bytes(Bp, Lq, Lk, d, d_in, mod.h), query.device)
        _lib.check(L.bist_mha_train_fwd(_lib.ptr(query), _lib.ptr(key), _lib.ptr(value), _lib.ptr(m8), mask_mode,
                                        ctypes.byref(P), _lib.ptr(out), _lib.ptr(attn), Bp, Lq, Lk, d, d_in, mod.h,
                                        _lib.ptr(tape), tape.numel(), dt, float(p_drop), seed, SITE_MHA,
                                        _lib.stream_ptr()), "bist_mha_train_fwd")
        ctx.save_for_backward(query, key, value, m8, tape)
        ctx.mod, ctx.cfg = mod, (Bp, Lq, Lk, d, d_in, mask_mode, float(p_drop), seed, dt)
        if attn is None:
            attn = torch.empty(0, device=query.device)
        ctx.mark_non_differentiable(attn)
        return out, attn

    @staticmethod
    def backward(ctx, g_out, _g_attn):
        query, key, value, m8, tape = ctx.saved_tensors
        mod = ctx.mod
        Bp, Lq, Lk, d, d_in, mask_mode, p_drop, seed, dt = ctx.cfg
        g_out = g_out.to(query.dtype).contiguous()
        PT, keep = mod._param_struct(dt, transposed=True)
        # one flat fp32 block for the eight parameter gradients (one all-reduce per module, dist.allreduce_grads)
        nw, nb = d * d_in, (d, d, d, d_in)
        block = _arena.zeros(4 * nw + sum(nb), query.device)
        dW = [block[i * nw:(i + 1) * nw] for i in range(4)]
        db, off = [], 4 * nw
        for n in nb:
            db.append(block[off:off + n])
            off += n
        d_in_t = [torch.empty_like(t) if need else None for t, need in zip((query, key, value), ctx.needs_input_grad[1:4])]
        L = _lib.lib()
        ws = _ws(L.bist_mha_bwd_workspace_bytes(Bp, Lq, Lk, d, dt), query.device)
        _lib.check(L.bist_mha_bwd(_lib.ptr(g_out), _lib.ptr(query), _lib.ptr(key), _lib.ptr(value), _lib.ptr(m8),
                                  mask_mode, ctypes.byref(PT), _lib.ptr(tape), tape.numel(), _lib.ptr(d_in_t[0]),
                                  _lib.ptr(d_in_t[1]), _lib.ptr(d_in_t[2]), _ptr_array(dW), _ptr_array(db), Bp, Lq, Lk,
                                  d, d_in, mod.h, _lib.ptr(ws), ws.numel(), dt, p_drop, seed, SITE_MHA,
                                  _lib.stream_ptr()), "bist_mha_bwd")
        grads = []
        for i, lin in enumerate(mod.linears):
            grads += [dW[i].view(lin.weight.shape), db[i]]
        _bucket_block(block, mod.parameters())
        return (None, d_in_t[0], d_in_t[1], d_in_t[2], None, None, None, None) + tuple(grads)


class _FfnFn(torch.autograd.Function):
    """PositionwiseFeedForward forward/backward (modules.py:112-113) — bist_ffn_train_fwd / bist_ffn_bwd."""

    @staticmethod
    def forward(ctx, mod, x, p_drop, seed, w1, b1, w2, b2):
        dt = _lib.dtype_code(x.dtype)
        x = x.contiguous()
        d, d_ff, d_out = w1.shape[1], w1.shape[0], w2.shape[0]
        rows = x.numel() // d
        cw1, cw2 = (w1.detach(), w2.detach()) if dt == _lib.F32 else (mod._bf16.get(w1), mod._bf16.get(w2))
        y = torch.empty(x.shape[:-1] + (d_out,), dtype=x.dtype, device=x.device)
        hidden = torch.empty((rows, d_ff), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().bist_ffn_train_fwd(_lib.ptr(x), _lib.ptr(cw1), _lib.ptr(b1), _lib.ptr(cw2), _lib.ptr(b2),
                                                 _lib.ptr(y), _lib.ptr(hidden), rows, d, d_ff, d_out, dt, float(p_drop),
                                                 seed, SITE_FFN, _lib.stream_ptr()), "bist_ffn_train_fwd")
        ctx.save_for_backward(x, hidden, w1, w2)
        ctx.mod, ctx.cfg = mod, (rows, d, d_ff, d_out, float(p_drop), dt)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, hidden, w1, w2 = ctx.saved_tensors
        rows, d, d_ff, d_out, p_drop, dt = ctx.cfg
        gy = gy.to(x.dtype).contiguous()
        w1T = ctx.mod._bf16.get(ctx.mod.w_1.weight, x.dtype, transposed=True)
        w2T = ctx.mod._bf16.get(ctx.mod.w_2.weight, x.dtype, transposed=True)
        o1, o2, o3 = d_ff * d, d_ff * d + d_ff, d_ff * d + d_ff + d_out * d_ff
        block = _arena.zeros(o3 + d_out, x.device)     # one block: one all-reduce per module
        dW1, db1, dW2, db2 = block[:o1], block[o1:o2], block[o2:o3], block[o3:]
        dx = torch.empty_like(x) if ctx.needs_input_grad[1] else None
        ws = _ws(rows * d_ff * x.element_size() + 256, x.device)
        _lib.check(_lib.lib().bist_ffn_bwd(_lib.ptr(gy), _lib.ptr(x), _lib.ptr(w1T), _lib.ptr(w2T), _lib.ptr(hidden),
                                           _lib.ptr(dx), _lib.ptr(dW1), _lib.ptr(db1), _lib.ptr(dW2), _lib.ptr(db2), rows,
                                           d, d_ff, d_out, _lib.ptr(ws), ws.numel(), dt, p_drop, _lib.stream_ptr()),
                   "bist_ffn_bwd")
        _bucket_block(block, ctx.mod.parameters())
        return None, dx, None, None, dW1.view(w1.shape), db1, dW2.view(w2.shape), db2


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
        y = sublayer(self.norm(x))
        self.last_dropout = None
        if self.training and self.dropout.p > 0:                       # modules.py:44
            _lib.require_cuda(y)
            seed = _draw_seed()
            self.last_dropout = (self.dropout.p, seed, SITE_SUBLAYER)
            if y.dtype == x.dtype and y.shape == x.shape:
                return _DropoutAddFn.apply(y, x, self.dropout.p, seed, SITE_SUBLAYER)
            y = _DropoutFn.apply(y, self.dropout.p, seed, SITE_SUBLAYER)
        return x + y


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

    def _param_struct(self, dt, transposed):
        """bist_mha_params_t over the compute-dtype (optionally transposed) copies of the four Linear weights."""
        P, keep = _lib.MhaParams(), []
        tdt = torch.float32 if dt == _lib.F32 else torch.bfloat16
        for i, lin in enumerate(self.linears):
            if transposed or dt != _lib.F32:
                w = self._bf16.get(lin.weight, tdt, transposed)
            else:
                w = lin.weight.detach()
            keep.append(w)
            P.w[i] = w.data_ptr()
            P.b[i] = lin.bias.data_ptr()
        return P, keep

    def forward(self, query, key, value, mask=None):
        _lib.require_cuda(query, key, value, self.linears[0].weight)
        dt = _lib.dtype_code(query.dtype)
        if key.dtype != query.dtype:                 # e.g. fp32 video streams attended by a bf16 decoder state
            key = key.to(query.dtype)
        if value.dtype != query.dtype:
            value = value.to(query.dtype)
        Bp, Lq, d_in = query.shape
        Lk = key.shape[1]
        d = self.h * self.d_k
        mask_mode, m8 = _lib.MASK_NONE, None
        if mask is not None:
            if mask.dim() == 3 and mask.shape[1] == 1:
                mask_mode, m8 = _lib.MASK_KEY, _mask_u8(mask, (Bp, 1, Lk))
            else:
                mask_mode, m8 = _lib.MASK_FULL, _mask_u8(mask, (Bp, Lq, Lk))
        p_drop = self.dropout.p if self.training else 0.0
        self.last_dropout = None
        params = [t for lin in self.linears for t in (lin.weight, lin.bias)]
        if p_drop > 0 or _needs_grad(query, key, value, *params):
            seed = _draw_seed() if p_drop > 0 else 0
            if p_drop > 0:
                self.last_dropout = (p_drop, seed, SITE_MHA)
            out, attn = _MhaFn.apply(self, query, key, value, m8, mask_mode, p_drop, seed, *params)
            attn = attn if self.store_attn else None
        else:
            query, key, value = (t.detach().contiguous() for t in (query, key, value))
            P, keep = self._param_struct(dt, transposed=False)
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
        _lib.require_cuda(x, self.w_1.weight)
        dt = _lib.dtype_code(x.dtype)
        p_drop = self.dropout.p if self.training else 0.0
        self.last_dropout = None
        params = (self.w_1.weight, self.w_1.bias, self.w_2.weight, self.w_2.bias)
        if p_drop > 0 or _needs_grad(x, *params):
            seed = _draw_seed() if p_drop > 0 else 0
            if p_drop > 0:
                self.last_dropout = (p_drop, seed, SITE_FFN)
            return _FfnFn.apply(self, x, p_drop, seed, *params)
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


class _LinearFn(torch.autograd.Function):
    """nn.Linear through the library (bist_linear_fwd / bist_linear_bwd): y = x W^T + b in x's dtype."""

    @staticmethod
    def forward(ctx, x, weight, bias, cache):
        x = x.contiguous()
        dt = _lib.dtype_code(x.dtype)
        N, K = weight.shape
        M = x.numel() // K
        w = weight.detach() if dt == _lib.F32 else cache.get(weight)
        y = torch.empty(x.shape[:-1] + (N,), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().bist_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), None, _lib.ptr(y), M, N, K, 0,
                                              dt, dt, dt, _lib.stream_ptr()), "bist_linear_fwd")
        ctx.save_for_backward(x, weight)
        ctx.cache = cache
        return y

    @staticmethod
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        dt = _lib.dtype_code(x.dtype)
        N, K = weight.shape
        M = x.numel() // K
        gy = gy.to(x.dtype).contiguous()
        wT = ctx.cache.get(weight, x.dtype, transposed=True)
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        block = _arena.zeros(N * K + N, x.device)
        _lib.check(_lib.lib().bist_linear_bwd(_lib.ptr(gy), _lib.ptr(x), _lib.ptr(wT), _lib.ptr(dx), _lib.ptr(block),
                                              _lib.ptr(block[N * K:]), M, N, K, dt, dt, dt, dt, _lib.stream_ptr()),
                   "bist_linear_bwd")
        return dx, block[:N * K].view(N, K), block[N * K:], None


class _GateFn(torch.autograd.Function):
    """out = sum_k softmax(logits)_k * src_k (decoder.py:155-167) — bist_gate_combine_fwd / bwd."""

    @staticmethod
    def forward(ctx, logits, *srcs):
        logits = logits.float().contiguous()
        srcs = tuple(s.contiguous() for s in srcs)
        d = srcs[0].shape[-1]
        R = srcs[0].numel() // d
        out = torch.empty_like(srcs[0])
        _lib.check(_lib.lib().bist_gate_combine_fwd(_ptr_array(srcs), _lib.ptr(logits), _lib.ptr(out), R, d, len(srcs),
                                                    _lib.dtype_code(out.dtype), _lib.stream_ptr()), "bist_gate_combine_fwd")
        ctx.save_for_backward(logits, *srcs)
        return out

    @staticmethod
    def backward(ctx, g):
        logits, *srcs = ctx.saved_tensors
        g = g.to(srcs[0].dtype).contiguous()
        d = srcs[0].shape[-1]
        R = srcs[0].numel() // d
        dsrc = [torch.empty_like(s) for s in srcs]
        dlog = torch.empty_like(logits)
        _lib.check(_lib.lib().bist_gate_combine_bwd(_lib.ptr(g), _ptr_array(srcs), _lib.ptr(logits), _ptr_array(dsrc),
                                                    _lib.ptr(dlog), R, d, len(srcs), _lib.dtype_code(g.dtype),
                                                    _lib.stream_ptr()), "bist_gate_combine_bwd")
        return (dlog,) + tuple(dsrc)


def linear(x, lin, cache):
    """``lin(x)`` for an nn.Linear through the library, with autograd."""
    _lib.require_cuda(x, lin.weight)
    return _LinearFn.apply(x, lin.weight, lin.bias, cache)


def gate_combine(logits, srcs):
    """softmax(logits) over the last dim, then the weighted sum of the K streams (each [..., d])."""
    _lib.require_cuda(logits, *srcs)
    return _GateFn.apply(logits, *srcs)


# ---------------------------------------------------------------------------------------------- text side (modules.py:115-141)
SITE_EMBED = 103


class _EmbedFn(torch.autograd.Function):
    """dropout(lut(ids) * sqrt(d) + pe) in one kernel — bist_embed_fwd / bist_embed_bwd."""

    @staticmethod
    def forward(ctx, ids, lut, pe, p_drop, seed, dtype):
        ids = ids.contiguous()
        B, L = ids.shape
        V, d = lut.shape
        out = torch.empty((B, L, d), dtype=dtype, device=lut.device)
        _lib.check(_lib.lib().bist_embed_fwd(_lib.ptr(ids), _lib.ptr(lut), _lib.ptr(pe), _lib.ptr(out), B * L, L, d, V,
                                             float(p_drop), seed, SITE_EMBED, _lib.dtype_code(dtype), _lib.stream_ptr()),
                   "bist_embed_fwd")
        ctx.save_for_backward(ids)
        ctx.cfg = (V, d, float(p_drop), seed)
        return out

    @staticmethod
    def backward(ctx, gy):
        (ids,) = ctx.saved_tensors
        V, d, p_drop, seed = ctx.cfg
        gy = gy.contiguous()
        dlut = _arena.zeros((V, d), gy.device)
        _lib.check(_lib.lib().bist_embed_bwd(_lib.ptr(ids), _lib.ptr(gy), _lib.ptr(dlut), ids.numel(), d, p_drop, seed,
                                             SITE_EMBED, _lib.dtype_code(gy.dtype), _lib.stream_ptr()), "bist_embed_bwd")
        return None, dlut, None, None, None, None


class Embeddings(nn.Module):
    "modules.py:115-123 — lut(x) * sqrt(d_model).  Used through ``EmbedSequence`` (fused with the positional encoding)."

    def __init__(self, d_model, vocab):
        super(Embeddings, self).__init__()
        self.lut = nn.Embedding(vocab, d_model)
        self.d_model = d_model

    def forward(self, x):
        raise NotImplementedError("Embeddings runs fused with PositionalEncoding: call the EmbedSequence that holds both")


class PositionalEncoding(nn.Module):
    "modules.py:125-141 — the sinusoidal table ``pe`` (buffer, same construction) and the dropout after the sum."

    def __init__(self, d_model, dropout, max_len=5000):
        super(PositionalEncoding, self).__init__()
        self.dropout = nn.Dropout(p=dropout)
        import math
        pe = torch.zeros(max_len, d_model)
        position = torch.arange(0., max_len).unsqueeze(1)
        div_term = torch.exp(torch.arange(0., d_model, 2) * -(math.log(10000.0) / d_model))
        pe[:, 0::2] = torch.sin(position * div_term)
        pe[:, 1::2] = torch.cos(position * div_term)
        self.register_buffer('pe', pe.unsqueeze(0))

    def forward(self, x):
        raise NotImplementedError("PositionalEncoding runs fused with Embeddings: call the EmbedSequence that holds both")


class EmbedSequence(nn.Sequential):
    """``nn.Sequential(Embeddings, PositionalEncoding)`` of mtn.py:78-80 (same children, same state_dict keys
    ``0.lut.weight`` / ``1.pe``) whose forward is ONE kernel: lookup, scale, add the table, dropout."""
    precision = "fp32"

    def forward(self, ids):
        if ids is None:
            return None
        emb, pos = self[0], self[1]
        _lib.require_cuda(ids, emb.lut.weight)
        p_drop = pos.dropout.p if self.training else 0.0
        seed = _draw_seed() if p_drop > 0 else 0
        self.last_dropout = (p_drop, seed, SITE_EMBED) if p_drop > 0 else None
        dtype = torch.float32 if self.precision == "fp32" else torch.bfloat16
        return _EmbedFn.apply(ids, emb.lut.weight, pos.pe[0], p_drop, seed, dtype)
