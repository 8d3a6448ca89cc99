"""Drop-in mirrors of /root/reference/model/generator.py and label_smoothing.py over libbist_b200.so
(SURVEY.md §8(f) row 2).

``Generator`` (generator.py:11-27), ``PointerGenerator`` (:29-75) and ``MultiPointerGenerator`` (:77-127) keep the
reference's constructor signatures, attribute names (``vocab_gen``, ``pointer_gen_W``, ``pointer_attn``, ``proj``) and
``forward(ft, batch, args)`` returning log-probabilities ``[B, Lt, V]``.  For training they add ``loss_rows``: the
label-smoothed KL of every target position, computed by ONE kernel pass per vocabulary row that also produces the
gradient with respect to the logits, the pointer probabilities and the switch (``bist_genloss_fwd_bwd``) — the
``[B*Lt, V]`` probabilities and log-probabilities are never written to memory.  ``LabelSmoothing`` mirrors
label_smoothing.py for callers that hold log-probabilities already.
"""
import ctypes

import torch
import torch.nn as nn

from . import _lib
from . import arena as _arena
from .modules import _BF16Cache, _LinearFn, _ptr_array


class _PtrAttnFn(torch.autograd.Function):
    """Pointer probabilities and the text vector (generator.py:62-69) — bist_ptr_attn_fwd / bwd."""

    @staticmethod
    def forward(ctx, q, k, v, mask_u8):
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        B, Lt, d = q.shape
        Ls = k.shape[1]
        P = torch.empty((B, Lt, Ls), dtype=torch.float32, device=q.device)
        tvec = torch.empty_like(q)
        _lib.check(_lib.lib().bist_ptr_attn_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(mask_u8), _lib.ptr(P),
                                                _lib.ptr(tvec), B, Lt, Ls, d, _lib.dtype_code(q.dtype), _lib.stream_ptr()),
                   "bist_ptr_attn_fwd")
        ctx.save_for_backward(q, k, v, P)
        return P, tvec

    @staticmethod
    def backward(ctx, dP, dtvec):
        q, k, v, P = ctx.saved_tensors
        B, Lt, d = q.shape
        Ls = k.shape[1]
        dP = torch.zeros_like(P) if dP is None else dP.float().contiguous()
        dtvec = torch.zeros_like(q) if dtvec is None else dtvec.to(q.dtype).contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _lib.check(_lib.lib().bist_ptr_attn_bwd(_lib.ptr(dP), _lib.ptr(dtvec), _lib.ptr(q), _lib.ptr(k), _lib.ptr(v),
                                                _lib.ptr(P), _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv), B, Lt, Ls, d,
                                                _lib.dtype_code(q.dtype), _lib.stream_ptr()), "bist_ptr_attn_bwd")
        return dq, dk, dv, None


def _i64_array(vals):
    arr = (ctypes.c_int64 * len(vals))()
    for i, v in enumerate(vals):
        arr[i] = int(v)
    return arr


def _genloss(Z, logp, Ps, dPs, texts, sw, dsw, target, row_loss, Lt, pad, smoothing, want_grad):
    R, V = Z.shape
    n = len(Ps)
    _lib.check(_lib.lib().bist_genloss_fwd_bwd(
        _lib.ptr(Z), _lib.ptr(logp), n, _ptr_array(Ps) if n else None, _ptr_array(dPs) if n and dPs else None,
        _ptr_array(texts) if n else None, _i64_array([p.shape[-1] for p in Ps]) if n else None, _lib.ptr(sw),
        _lib.ptr(dsw), _lib.ptr(target), _lib.ptr(row_loss), R, V, Lt, int(pad), float(smoothing), int(want_grad),
        _lib.dtype_code(Z.dtype), _lib.stream_ptr()), "bist_genloss_fwd_bwd")


class _GenLossFn(torch.autograd.Function):
    """row_loss[r] of the (pointer) generator + label smoothing; the forward already holds every gradient."""

    @staticmethod
    def forward(ctx, x, E, cache, target, pad, smoothing, texts, sw, *Ps):
        x = x.contiguous()
        B, Lt, d = x.shape
        V = E.shape[0]
        R = B * Lt
        dt = _lib.dtype_code(x.dtype)
        Ec = E.detach() if dt == _lib.F32 else cache.get(E)
        Z = torch.empty((R, V), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().bist_linear_fwd(_lib.ptr(x), _lib.ptr(Ec), None, None, _lib.ptr(Z), R, V, d, 0, dt, dt, dt,
                                              _lib.stream_ptr()), "bist_linear_fwd")           # generator.py:82
        Ps = [p.float().contiguous().view(R, -1) for p in Ps]
        sw = sw.float().contiguous().view(R, -1) if sw is not None else None
        dPs = [torch.empty_like(p) for p in Ps]
        dsw = torch.empty_like(sw) if sw is not None else None
        row_loss = torch.empty(R, dtype=torch.float32, device=x.device)
        target = target.contiguous().view(-1)
        _genloss(Z, None, Ps, dPs, texts, sw, dsw, target, row_loss, Lt, pad, smoothing, True)
        ctx.save_for_backward(x, E, Z, dsw, *dPs)          # Z now holds d(sum row_loss)/dZ
        ctx.cache, ctx.has_sw, ctx.shapes = cache, sw is not None, (B, Lt, d, V)
        return row_loss

    @staticmethod
    def backward(ctx, g_rows):
        x, E, dZ, dsw, *dPs = ctx.saved_tensors
        B, Lt, d, V = ctx.shapes
        R = B * Lt
        dt = _lib.dtype_code(x.dtype)
        g = g_rows.float().contiguous().view(R, 1)
        dZ = (dZ * g.to(dZ.dtype)).contiguous()
        ET = ctx.cache.get(E, x.dtype, transposed=True)                                          # [d, V]
        dx = torch.empty_like(x)
        dE = _arena.zeros(tuple(E.shape), E.device)
        _lib.check(_lib.lib().bist_linear_bwd(_lib.ptr(dZ), _lib.ptr(x), _lib.ptr(ET), _lib.ptr(dx), _lib.ptr(dE), None, R,
                                              V, d, dt, dt, dt, dt, _lib.stream_ptr()), "bist_linear_bwd")
        d_sw = (dsw * g).view(B, Lt, -1) if ctx.has_sw else None
        d_Ps = tuple((p * g).view(B, Lt, -1) for p in dPs)
        return (dx, dE, None, None, None, None, None, d_sw) + d_Ps


def _lin(x, lin, cache):
    return _LinearFn.apply(x, lin.weight, lin.bias, cache)


def _source(name, batch, ft):
    if name == 'query':
        return batch.query, ft['encoded_query'], batch.query_mask
    if name == 'his':
        return batch.his, ft['encoded_his'], batch.his_mask
    if name == 'cap':
        return batch.cap, ft['encoded_cap'], batch.cap_mask
    raise NotImplementedError("pointer source %r (query, his, cap are built)" % name)


class _PointerBase(nn.Module):
    """Shared body of the two pointer generators: pointer attentions -> switch logits -> fused mix/loss kernel."""

    def _sources(self, args):
        return args.ptr_ft.split(',')

    def _attn_module(self, i):
        raise NotImplementedError

    def _switch_logits(self, x, encoded_tgt, tvecs):
        raise NotImplementedError

    def _prepare(self, ft, batch, args):
        x = ft['decoded_text']
        _lib.require_cuda(x, self.vocab_gen)
        Ps, tvecs, texts = [], [], []
        for i, name in enumerate(self._sources(args)):
            text, enc, mask = _source(name, batch, ft)
            if args.mask_unk:
                mask = mask & (text != 0).unsqueeze(-2)                      # generator.py:59-60
            att = self._attn_module(i)
            enc = enc.to(x.dtype)
            q = _lin(x, att.linears[0], self._cache)
            k = _lin(enc, att.linears[1], self._cache)
            m8 = mask.reshape(mask.shape[0], -1).contiguous().view(torch.uint8)
            P, tvec = _PtrAttnFn.apply(q, k, enc, m8)
            Ps.append(P)
            tvecs.append(tvec)
            texts.append(text.contiguous())
        sw = self._switch_logits(x, ft['encoded_tgt'].to(x.dtype), tvecs).float()
        return x, Ps, texts, sw

    def forward(self, ft, batch, args):
        "log p_out [B, Lt, V] (generator.py:75 / :127)."
        with torch.no_grad():
            x, Ps, texts, sw = self._prepare(ft, batch, args)
            B, Lt, d = x.shape
            V = self.vocab_gen.shape[0]
            dt = _lib.dtype_code(x.dtype)
            Ec = self.vocab_gen.detach() if dt == _lib.F32 else self._cache.get(self.vocab_gen)
            Z = torch.empty((B * Lt, V), dtype=x.dtype, device=x.device)
            _lib.check(_lib.lib().bist_linear_fwd(_lib.ptr(x.contiguous()), _lib.ptr(Ec), None, None, _lib.ptr(Z), B * Lt, V,
                                                  d, 0, dt, dt, dt, _lib.stream_ptr()), "bist_linear_fwd")
            logp = torch.empty((B * Lt, V), dtype=torch.float32, device=x.device)
            Ps = [p.contiguous().view(B * Lt, -1) for p in Ps]
            _genloss(Z, logp, Ps, None, texts, sw.contiguous().view(B * Lt, -1), None, None, None, Lt, 0, 0.0, False)
        return logp.view(B, Lt, V)

    def loss_rows(self, ft, batch, args, target, pad, smoothing):
        "Un-normalised label-smoothed KL per target position [B*Lt] (label_smoothing.py:20-30 on log p_out)."
        x, Ps, texts, sw = self._prepare(ft, batch, args)
        return _GenLossFn.apply(x, self.vocab_gen, self._cache, target, pad, smoothing, texts, sw, *Ps)


class PointerGenerator(_PointerBase):
    "generator.py:29-75 — one pointer source, sigmoid switch."

    def __init__(self, d_model, vocab_gen, pointer_attn):
        super(PointerGenerator, self).__init__()
        self.vocab_gen = vocab_gen
        self.pointer_gen_W = nn.Linear(d_model * 3, 1)
        self.pointer_attn = pointer_attn
        self._cache = _BF16Cache()

    def _sources(self, args):
        if args.ptr_ft == 'query+cap':
            raise NotImplementedError("ptr_ft='query+cap' (concatenated sources) is not built; use 'query,cap'")
        return [args.ptr_ft]

    def _attn_module(self, i):
        return self.pointer_attn

    def _switch_logits(self, x, encoded_tgt, tvecs):
        w = _lin(torch.cat([x, tvecs[0], encoded_tgt], -1), self.pointer_gen_W, self._cache)     # generator.py:71-72
        return torch.cat([torch.zeros_like(w), w], -1)      # softmax([0, w]) = [1 - sigmoid(w), sigmoid(w)]  (:73)


class MultiPointerGenerator(_PointerBase):
    "generator.py:77-127 — several pointer sources, softmax switch (vocabulary weight last)."

    def __init__(self, d_model, vocab_gen, pointer_attn, nb_pointer_ft):
        super(MultiPointerGenerator, self).__init__()
        self.vocab_gen = vocab_gen
        self.pointer_gen_W = nn.Linear(d_model * (nb_pointer_ft + 2), nb_pointer_ft + 1)
        self.pointer_attn = pointer_attn
        self._cache = _BF16Cache()

    def _attn_module(self, i):
        return self.pointer_attn[i]

    def _switch_logits(self, x, encoded_tgt, tvecs):
        return _lin(torch.cat([x, encoded_tgt] + tvecs, -1), self.pointer_gen_W, self._cache)    # generator.py:120-121


class Generator(nn.Module):
    "generator.py:11-27 — log_softmax(x W^T (+ b)); W is the tied embedding when given."

    def __init__(self, d_model, vocab, W=None):
        super(Generator, self).__init__()
        if W is not None:
            self.proj = W
            self.shared_W = True
        else:
            self.proj = nn.Linear(d_model, vocab)
            self.shared_W = False
        self._cache = _BF16Cache()

    def _weight(self):
        if not self.shared_W:
            raise NotImplementedError("Generator with its own Linear (separate_out_linear) is not built; pass the tied W")
        return self.proj

    def forward(self, ft, batch, args, ft_key='decoded_text'):
        x = ft[ft_key]
        E = self._weight()
        _lib.require_cuda(x, E)
        with torch.no_grad():
            x = x.contiguous()
            B, L, d = x.shape
            dt = _lib.dtype_code(x.dtype)
            Ec = E.detach() if dt == _lib.F32 else self._cache.get(E)
            Z = torch.empty((B * L, E.shape[0]), dtype=x.dtype, device=x.device)
            _lib.check(_lib.lib().bist_linear_fwd(_lib.ptr(x), _lib.ptr(Ec), None, None, _lib.ptr(Z), B * L, E.shape[0], d, 0,
                                                  dt, dt, dt, _lib.stream_ptr()), "bist_linear_fwd")
            logp = torch.empty(Z.shape, dtype=torch.float32, device=x.device)
            _genloss(Z, logp, [], None, [], None, None, None, None, L, 0, 0.0, False)
        return logp.view(B, L, -1)

    def loss_rows(self, ft, batch, args, target, pad, smoothing, ft_key='decoded_text'):
        x = ft[ft_key]
        if 'decoded_text' in ft and x.dtype != ft['decoded_text'].dtype:
            x = x.to(ft['decoded_text'].dtype)           # the video streams leave VidEncoderLayer4 in fp32
        _lib.require_cuda(x, self._weight())
        return _GenLossFn.apply(x, self._weight(), self._cache, target, pad, smoothing, [], None)


for _cls in (PointerGenerator, MultiPointerGenerator, Generator):
    def _deepcopy(self, memo, _cls=_cls):
        import copy
        new = type(self).__new__(type(self))
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = _BF16Cache() if k == "_cache" else copy.deepcopy(v, memo)
        return new
    _cls.__deepcopy__ = _deepcopy


class LabelSmoothing(nn.Module):
    """label_smoothing.py:9-30 for callers that already hold log-probabilities ``x`` [R, V]: summed KL against the
    smoothed target distribution, through the same kernel (the log-probabilities are treated as logits of an
    already-normalised row: softmax(log p) = p)."""

    def __init__(self, size, padding_idx, smoothing=0.0):
        super(LabelSmoothing, self).__init__()
        self.padding_idx = padding_idx
        self.confidence = 1.0 - smoothing
        self.smoothing = smoothing
        self.size = size

    def forward(self, x, target):
        assert x.size(1) == self.size
        _lib.require_cuda(x)
        return _LogpLossFn.apply(x, target, self.padding_idx, self.smoothing).sum()


class _LogpLossFn(torch.autograd.Function):
    """Forward through the fused kernel; the gradient with respect to log-probabilities is just -t (built densely
    here: this is the compatibility path, the training step uses ``generator.loss_rows`` and never holds [R, V])."""

    @staticmethod
    def forward(ctx, logp, target, pad, smoothing):
        Z = logp.detach().float().clone().contiguous()
        target = target.contiguous().view(-1)
        row_loss = torch.empty(Z.shape[0], dtype=torch.float32, device=Z.device)
        _genloss(Z, None, [], None, [], None, None, target, row_loss, 1, pad, smoothing, False)
        ctx.save_for_backward(target)
        ctx.cfg = (Z.shape[1], pad, smoothing, logp.dtype)
        return row_loss

    @staticmethod
    def backward(ctx, g_rows):
        (target,) = ctx.saved_tensors
        V, pad, smoothing, dtype = ctx.cfg
        t = torch.full((target.numel(), V), smoothing / (V - 2), dtype=torch.float32, device=target.device)
        t.scatter_(1, target.unsqueeze(1), 1.0 - smoothing)
        t[:, pad] = 0
        t[target == pad] = 0
        return (-t * g_rows.float().unsqueeze(1)).to(dtype), None, None, None
