"""Drop-in mirrors of the video hot path of /root/reference/model/encoder.py over libbist_b200.so.

``VidEncoder8``       encoder.py:55-93   (2048 -> d_model projection + ReLU + LayerNorm)
``VidEncoderLayer4``  encoder.py:95-201  (query self-attention, temporal2spatial, spatial2temporal)
``VideoPath``         the slice of ``MTN.forward`` that is the hot path (mtn.py:49-51 and
                      decoder.py:107-129), with the reference's parameter names, used by
                      bench.py and the parity tests.

Constructor and forward signatures, the ``b`` / ``ft`` / ``in_ft`` dict protocol and the
state_dict key names are the reference's.  ``precision`` ("fp32" | "bf16") selects the storage
type of the projected video tensor and of the weights inside the kernels; parameters stay
fp32 masters and the [B,Lq,d] query streams stay fp32 in both modes.

Autograd: when a gradient is required (or the module is in ``train()`` mode with dropout) the
modules run the library's training forward (``bist_vidproj_train_fwd`` /
``bist_st_layer_train_fwd``: tape + dropout) and ``loss.backward()`` runs its hand-written
backward (``bist_st_layer_bwd`` / ``bist_vidproj_bwd``); under ``torch.no_grad()`` in eval mode
they run the inference forward (folded formulation, CUDA-graph capturable).
"""
import copy
import ctypes

import torch
import torch.nn as nn

from . import _lib
from . import arena as _arena
from . import dist as _dist
from .modules import (LayerNorm, SublayerConnection, clones, _needs_grad, _ws, _WEIGHT_EPOCH,
                      _reg_hook, weights_updated)

_TORCH_DTYPE = {"fp32": torch.float32, "bf16": torch.bfloat16}


def set_precision(module, precision):
    """Set the compute/storage precision of every bist_b200 video module under `module`."""
    if precision not in _TORCH_DTYPE:
        raise ValueError("precision must be 'fp32' or 'bf16'")
    for m in module.modules():
        if isinstance(m, (VidEncoder8, VidEncoderLayer4)):
            m.precision = precision
    return module


def _mask_bytes(mask, B, L):
    m = mask.reshape(B, L)
    m = m if m.dtype == torch.bool else (m != 0)
    return m.contiguous().view(torch.uint8)


class _GradBox:
    """fp32 accumulator for d(ft['spatiotemporal_ft']) shared by the projection and the layers that read it.

    The layers' backward adds into ``acc`` in place (fp32, one buffer for all blocks and directions) instead of
    returning bf16 gradients for autograd to sum; the projection's backward consumes it.
    """

    def __init__(self):
        self.acc = None

    def get(self, like):
        if self.acc is None:
            self.acc = torch.zeros(like.shape, dtype=torch.float32, device=like.device)
        return self.acc


class _VidProjFn(torch.autograd.Function):
    """LayerNorm(relu(W fts + b)) with a backward (encoder.py:75,80) — bist_vidproj_train_fwd / bist_vidproj_bwd."""

    @staticmethod
    def forward(ctx, fts, W, bias, a_2, b_2, eps, precision, box):
        D, d = W.shape[1], W.shape[0]
        N = fts.numel() // D
        cdt = _lib.BF16 if precision == "bf16" else _lib.F32
        out = torch.empty(fts.shape[:-1] + (d,), dtype=_TORCH_DTYPE[precision], device=fts.device)
        L = _lib.lib()
        tape = _ws(L.bist_vidproj_tape_bytes(N, D, d, cdt), fts.device)
        _lib.check(L.bist_vidproj_train_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a_2), _lib.ptr(b_2),
                                            _lib.ptr(out), _lib.ptr(tape), tape.numel(), N, D, d, float(eps), cdt, cdt,
                                            _lib.stream_ptr()), "bist_vidproj_train_fwd")
        handle = torch.zeros(1, dtype=torch.float32, device=fts.device)
        ctx.save_for_backward(fts, tape, a_2, W)
        ctx.N = N
        ctx.box, ctx.eps, ctx.cdt = box, float(eps), cdt
        ctx.set_materialize_grads(False)
        return out, handle

    @staticmethod
    def backward(ctx, g_out, g_handle):
        fts, tape, a_2, W = ctx.saved_tensors
        d, D = W.shape
        N = ctx.N
        acc = ctx.box.acc
        ctx.box.acc = None
        if g_out is not None:                       # some other consumer differentiated the tensor directly
            acc = g_out.float().reshape(N, d) if acc is None else acc.add_(g_out.float().reshape(acc.shape))
        if acc is None:
            return (None,) * 8
        dW = _arena.zeros(tuple(W.shape), W.device)
        vec = _arena.zeros((3, d), W.device)
        L = _lib.lib()
        ws = _ws(L.bist_vidproj_bwd_workspace_bytes(N, D, d, ctx.cdt), W.device)
        _lib.check(L.bist_vidproj_bwd(_lib.ptr(acc), _lib.ptr(fts), _lib.ptr(tape), tape.numel(), _lib.ptr(a_2), _lib.ptr(dW),
                                      _lib.ptr(vec[0]), _lib.ptr(vec[1]), _lib.ptr(vec[2]), N, D, d, ctx.eps,
                                      _lib.ptr(ws), ws.numel(), ctx.cdt, _lib.stream_ptr()), "bist_vidproj_bwd")
        return None, dW, vec[0], vec[1], vec[2], None, None, None


def _layer_grad_offsets(layer, d, d_ff, n_dirs):
    """(offset, numel) of every parameter of a VidEncoderLayer4 inside the library's gradient block
    (layout documented in include/bist_b200.h), in ``layer.named_parameters()`` order."""
    dd = d * d
    attn_stride = 4 * dd + 4 * d
    ff = 3 * attn_stride
    ln = ff + 2 * d * d_ff + d_ff + d
    gdir = ln + 8 * d
    out = []
    for name, p in layer.named_parameters():
        parts = name.split(".")
        if parts[0] == "attn":
            i, j = int(parts[1]), int(parts[3])
            base = (i // 3) * gdir + (i % 3) * attn_stride
            off = base + (j * dd if parts[4] == "weight" else 4 * dd + j * d)
        elif parts[0] == "ff":
            base = int(parts[1]) * gdir + ff
            off = base + {"w_1.weight": 0, "w_1.bias": d_ff * d, "w_2.weight": d_ff * d + d_ff,
                          "w_2.bias": 2 * d_ff * d + d_ff}[parts[2] + "." + parts[3]]
        elif parts[0] == "sublayer":
            i = int(parts[1])
            off = (i // 4) * gdir + ln + (i % 4) * 2 * d + (0 if parts[3] == "a_2" else d)
        else:
            raise ValueError("unexpected VidEncoderLayer4 parameter %r" % name)
        out.append((off, p.numel()))
    assert max(o + n for o, n in out) <= n_dirs * gdir
    return out, n_dirs * gdir


class _StLayerFn(torch.autograd.Function):
    """VidEncoderLayer4.forward with a backward — bist_st_layer_train_fwd / bist_st_layer_bwd."""

    @staticmethod
    def forward(ctx, layer, vft, handle, box, qm, tm, dir_mask, dropout, x_t2s, x_s2t, *params):
        B, T, S, d = vft.shape
        d_model, d_ff, h = layer._dims()
        nd = layer._n_dirs()
        run_t2s, run_s2t = bool(dir_mask & 1), bool(dir_mask & 2)
        Lq = (x_t2s if run_t2s else x_s2t).shape[1]
        vdt = _lib.dtype_code(vft.dtype)
        o_t2s = torch.empty_like(x_t2s) if run_t2s else None
        o_s2t = torch.empty_like(x_s2t) if run_s2t else None
        packed = layer.train_packed_weights()
        L = _lib.lib()
        tape = _ws(L.bist_st_layer_tape_bytes(B, T, S, Lq, d, d_ff, h, nd, vdt), vft.device)
        eps = float(layer.sublayer[0].norm.eps)
        _lib.check(L.bist_st_layer_train_fwd(_lib.ptr(x_t2s), _lib.ptr(x_s2t), _lib.ptr(vft), _lib.ptr(qm), _lib.ptr(tm),
                                             _lib.ptr(packed), _lib.ptr(o_t2s), _lib.ptr(o_s2t), B, T, S, Lq, d, d_ff, h,
                                             nd, dir_mask, eps, _lib.ptr(tape), tape.numel(), vdt,
                                             ctypes.byref(dropout) if dropout is not None else None, _lib.stream_ptr()),
                   "bist_st_layer_train_fwd")
        ctx.dropout = dropout
        ctx.save_for_backward(vft, qm, tm, packed, tape, x_t2s, x_s2t)
        ctx.layer, ctx.box, ctx.dims = layer, box, (B, T, S, Lq, d, d_ff, h, nd, dir_mask, eps, vdt)
        ctx.vft_needs_grad = vft.requires_grad
        return tuple(o for o in (o_t2s, o_s2t) if o is not None)

    @staticmethod
    def backward(ctx, *g_outs):
        vft, qm, tm, packed, tape, x_t2s, x_s2t = ctx.saved_tensors
        B, T, S, Lq, d, d_ff, h, nd, dir_mask, eps, vdt = ctx.dims
        run_t2s, run_s2t = bool(dir_mask & 1), bool(dir_mask & 2)
        g_outs = list(g_outs)
        g_t2s = g_outs.pop(0).float().contiguous() if run_t2s else None
        g_s2t = g_outs.pop(0).float().contiguous() if run_s2t else None
        if run_t2s and run_s2t and g_t2s.data_ptr() == g_s2t.data_ptr():
            g_s2t = g_s2t.clone()
        d_t2s = torch.empty_like(x_t2s) if run_t2s else None
        d_s2t = torch.empty_like(x_s2t) if run_s2t else None
        shared = ctx.box is not None
        acc = ctx.box.get(vft) if shared else torch.zeros(vft.shape, dtype=torch.float32, device=vft.device)
        offs, total = _layer_grad_offsets(ctx.layer, d, d_ff, nd)
        L = _lib.lib()
        if total != L.bist_st_layer_grad_elems(d, d_ff, nd):
            raise RuntimeError("gradient block layout mismatch between bist_b200/encoder.py and libbist_b200.so")
        grads = _arena.zeros(total, vft.device)
        ws = _ws(L.bist_st_layer_bwd_workspace_bytes(B, T, S, Lq, d, d_ff, h, nd, vdt), vft.device)
        _lib.check(L.bist_st_layer_bwd(_lib.ptr(g_t2s), _lib.ptr(g_s2t), _lib.ptr(x_t2s), _lib.ptr(x_s2t), _lib.ptr(vft),
                                       _lib.ptr(qm), _lib.ptr(tm), _lib.ptr(packed), _lib.ptr(tape), tape.numel(),
                                       _lib.ptr(d_t2s), _lib.ptr(d_s2t), _lib.ptr(acc), _lib.ptr(grads), B, T, S, Lq, d,
                                       d_ff, h, nd, dir_mask, eps, _lib.ptr(ws), ws.numel(), vdt,
                                       ctypes.byref(ctx.dropout) if ctx.dropout is not None else None, _lib.stream_ptr()),
                   "bist_st_layer_bwd")
        if _reg_hook is None:
            weights_updated()
        if all(p.grad is None for p in ctx.layer.parameters()):     # autograd will adopt the slices: the block IS .grad
            _dist.grad_block_ready(grads)
        pg = tuple(grads[o:o + n].view(p.shape) for (o, n), p in zip(offs, ctx.layer.parameters()))
        g_vft = None if shared or not ctx.vft_needs_grad else acc.to(vft.dtype)
        g_handle = torch.zeros(1, dtype=torch.float32, device=vft.device) if shared else None
        return (None, g_vft, g_handle, None, None, None, None, None, d_t2s, d_s2t) + pg


class VidEncoder8(nn.Module):
    "encoder.py:55-93."

    def __init__(self, W, a_W, vid_position, v_N, a_N, size, args):
        super(VidEncoder8, self).__init__()
        self.v_N = v_N
        self.a_N = a_N
        self.args = args
        self.precision = "fp32"
        if self.v_N > 0:
            self.W = W
            self.vid_position = vid_position
            self.in_norm = LayerNorm(size)
        if self.a_N > 0:
            self.a_W = a_W
            self.vid_position = vid_position
            self.a_in_norm = LayerNorm(size)

    def _project(self, x, lin, norm):
        """LayerNorm(relu(lin(x))) as one C-ABI call (encoder.py:75,80 / :87,90)."""
        _lib.require_cuda(x, lin.weight)
        x = x.detach()
        needs_grad = _needs_grad(lin.weight, lin.bias, norm.a_2, norm.b_2)
        if x.dtype == torch.bfloat16 and self.precision == "bf16" and not needs_grad:
            # features held in bf16 by the ingest cache (ingest.py): the projection reads 2 bytes per value
            from .modules import _BF16Cache
            cache = self.__dict__.setdefault("_wcache", _BF16Cache())
            x = x.contiguous()
            D, d = lin.in_features, lin.out_features
            N = x.numel() // D
            out = torch.empty(x.shape[:-1] + (d,), dtype=torch.bfloat16, device=x.device)
            ws = _ws(N * d * 4 + 256, x.device)
            _lib.check(_lib.lib().bist_vidproj_fwd_bf16in(_lib.ptr(x), _lib.ptr(cache.get(lin.weight)), _lib.ptr(lin.bias),
                                                          _lib.ptr(norm.a_2), _lib.ptr(norm.b_2), _lib.ptr(out), N, D, d,
                                                          float(norm.eps), _lib.ptr(ws), ws.numel(), _lib.BF16,
                                                          _lib.stream_ptr()), "bist_vidproj_fwd_bf16in")
            return out
        if x.dtype != torch.float32:
            x = x.float()
        x = x.contiguous()
        if needs_grad:
            box = _GradBox()
            out, handle = _VidProjFn.apply(x, lin.weight, lin.bias, norm.a_2, norm.b_2, norm.eps, self.precision, box)
            out._bist_box, out._bist_handle = box, handle      # picked up by VidEncoderLayer4.forward
            return out
        D, d = lin.in_features, lin.out_features
        N = x.numel() // D
        cdt = _lib.BF16 if self.precision == "bf16" else _lib.F32
        out = torch.empty(x.shape[:-1] + (d,), dtype=_TORCH_DTYPE[self.precision], device=x.device)
        L = _lib.lib()
        ws = _ws(L.bist_vidproj_workspace_bytes(N, D, d, cdt), x.device)
        _lib.check(L.bist_vidproj_fwd(_lib.ptr(x), _lib.ptr(lin.weight), _lib.ptr(lin.bias), _lib.ptr(norm.a_2),
                                      _lib.ptr(norm.b_2), _lib.ptr(out), N, D, d, float(norm.eps), _lib.ptr(ws),
                                      ws.numel(), cdt, cdt, _lib.stream_ptr()), "bist_vidproj_fwd")
        return out

    def forward(self, b, ft):
        if self.v_N > 0:
            if self.vid_position is not None:      # always None in the reference (mtn.py:108)
                raise NotImplementedError("vid_position is never set by make_model (mtn.py:108)")
            ft['spatiotemporal_ft'] = self._project(b.fts, self.W, self.in_norm)
        if self.a_N > 0:
            if getattr(self.args, 'noW_venc', 0):  # encoder.py:84 (attribute absent from the configs)
                ft['encoded_audio'] = self.a_in_norm(b.audio_fts)
            else:
                ft['encoded_audio'] = self._project(b.audio_fts, self.a_W, self.a_in_norm).float()
        return ft


class VidEncoderLayer4(nn.Module):
    "encoder.py:95-201 with enc_st_combine == 'none'."

    def __init__(self, size, attn, nb_attn, ff, nb_ff, dropout, args):
        super(VidEncoderLayer4, self).__init__()
        self.size = size
        self.attn = clones(attn, nb_attn)
        self.ff = clones(ff, nb_ff)
        self.sublayer = clones(SublayerConnection(size, dropout), nb_attn + nb_ff)
        self.args = args
        self.precision = "fp32"
        if getattr(self.args, 'enc_st_combine', 'none') in ['early_sum', 'early_dyn']:
            raise NotImplementedError(
                "enc_st_combine=%r: the early_* modes only run with nb_blocks=1 in the reference "
                "(SURVEY.md §8 a2) and are not built" % self.args.enc_st_combine)
        self._packed = None          # (key, tensor)

    def __deepcopy__(self, memo):
        new = type(self).__new__(type(self))
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = None if k in ("_packed", "_train_packed") else copy.deepcopy(v, memo)
        return new

    # -- parameter hand-over ------------------------------------------------------------------
    def _dims(self):
        lin = self.attn[0].linears[0]
        return lin.out_features, self.ff[0].w_1.out_features, self.attn[0].h

    def _n_dirs(self):
        if len(self.attn) == 6 and len(self.ff) == 2:
            return 2
        if len(self.attn) == 3 and len(self.ff) == 1:
            return 1
        raise ValueError("VidEncoderLayer4 needs 6 attn / 2 ff or 3 attn / 1 ff clones (mtn.py:123-128)")

    def _param_struct(self, d):
        P = _lib.StLayerParams()
        for i, a in enumerate(self.attn):
            for j, lin in enumerate(a.linears):
                if lin.in_features != d or lin.out_features != d:
                    raise ValueError("VidEncoderLayer4 kernels need d_in == d_model attention clones")
                P.attn_w[i][j] = lin.weight.data_ptr()
                P.attn_b[i][j] = lin.bias.data_ptr()
        for i, f in enumerate(self.ff):
            P.ff_w1[i], P.ff_b1[i] = f.w_1.weight.data_ptr(), f.w_1.bias.data_ptr()
            P.ff_w2[i], P.ff_b2[i] = f.w_2.weight.data_ptr(), f.w_2.bias.data_ptr()
        for i, s in enumerate(self.sublayer):
            P.ln_a[i], P.ln_b[i] = s.norm.a_2.data_ptr(), s.norm.b_2.data_ptr()
        return P

    def packed_weights(self):
        """Pack (and cache until a parameter changes) the layer weights for the kernels."""
        params = list(self.parameters())
        key = (self.precision, params[0].device, _WEIGHT_EPOCH[0], tuple(p._version for p in params),
               tuple(p.data_ptr() for p in params))
        if self._packed is not None and self._packed[0] == key:
            return self._packed[1]
        d, d_ff, h = self._dims()
        nd = self._n_dirs()
        dt = _lib.BF16 if self.precision == "bf16" else _lib.F32
        P = self._param_struct(d)
        L = _lib.lib()
        nbytes = L.bist_st_layer_packed_bytes(d, d_ff, nd, dt)
        buf = torch.empty(nbytes, dtype=torch.uint8, device=params[0].device)
        _lib.check(L.bist_st_layer_pack(ctypes.byref(P), _lib.ptr(buf), nbytes, d, d_ff, h, nd, dt,
                                        _lib.stream_ptr()), "bist_st_layer_pack")
        self._packed = (key, buf)
        return buf

    def _dropout_cfg(self):
        """bist_dropout_t of this call: the three probabilities of the reference's nn.Dropout members (modules.py:40,79,
        109) and a fresh seed from torch's CPU generator (so torch.manual_seed reproduces a run); None in eval mode."""
        self.last_dropout = None
        if not self.training:
            return None
        p_sub, p_attn, p_ffn = self.sublayer[0].dropout.p, self.attn[0].dropout.p, self.ff[0].dropout.p
        if p_sub == 0 and p_attn == 0 and p_ffn == 0:
            return None
        cfg = _lib.Dropout(float(p_attn), float(p_sub), float(p_ffn), int(torch.randint(0, 2 ** 62, (1,)).item()))
        self.last_dropout = cfg            # tests read the seed back to rebuild the masks
        return cfg

    def train_packed_weights(self):
        """Packed weights of the training formulation (both operand orientations), cached per weight version."""
        params = list(self.parameters())
        key = (self.precision, params[0].device, _WEIGHT_EPOCH[0], tuple(p._version for p in params),
               tuple(p.data_ptr() for p in params))
        hit = getattr(self, "_train_packed", None)
        if hit is not None and hit[0] == key:
            return hit[1]
        d, d_ff, h = self._dims()
        nd = self._n_dirs()
        dt = _lib.BF16 if self.precision == "bf16" else _lib.F32
        P = self._param_struct(d)
        L = _lib.lib()
        nbytes = L.bist_st_layer_train_packed_bytes(d, d_ff, nd, dt)
        buf = torch.empty(nbytes, dtype=torch.uint8, device=params[0].device)
        _lib.check(L.bist_st_layer_train_pack(ctypes.byref(P), _lib.ptr(buf), nbytes, d, d_ff, h, nd, dt,
                                              _lib.stream_ptr()), "bist_st_layer_train_pack")
        self._train_packed = (key, buf)
        return buf

    # -- forward ------------------------------------------------------------------------------
    def forward(self, in_ft, ft, b):
        vft = ft['spatiotemporal_ft']
        _lib.require_cuda(vft, self.attn[0].linears[0].weight)
        run_t2s = (not hasattr(self.args, 't2s')) or bool(self.args.t2s)      # encoder.py:174
        run_s2t = (not hasattr(self.args, 's2t')) or bool(self.args.s2t)      # encoder.py:182
        dir_mask = (1 if run_t2s else 0) | (2 if run_s2t else 0)
        if dir_mask == 0:
            return in_ft
        vdt = _lib.dtype_code(vft.dtype)
        want = _lib.BF16 if self.precision == "bf16" else _lib.F32
        if vdt != want:
            raise TypeError("ft['spatiotemporal_ft'] is %s but this layer's precision is %s" % (vft.dtype, self.precision))
        B, T, S, d = vft.shape
        d_model, d_ff, h = self._dims()
        if d != d_model:
            raise ValueError("video feature width %d != d_model %d" % (d, d_model))
        nd = self._n_dirs()
        params = list(self.parameters())
        dropout = self._dropout_cfg()
        if dropout is not None or _needs_grad(vft, in_ft.get('t2s') if run_t2s else None,
                                              in_ft.get('s2t') if run_s2t else None, *params):
            x_t2s = in_ft['t2s'].float().contiguous() if run_t2s else None
            x_s2t = in_ft['s2t'].float().contiguous() if run_s2t else None
            Lq = (x_t2s if run_t2s else x_s2t).shape[1]
            qm = _mask_bytes(b.query_mask, B, Lq)
            tm = _mask_bytes(b.temporal_mask, B, T)
            box, handle = getattr(vft, "_bist_box", None), getattr(vft, "_bist_handle", None)
            outs = list(_StLayerFn.apply(self, vft.contiguous(), handle, box, qm, tm, dir_mask, dropout,
                                         x_t2s, x_s2t, *params))
            if run_t2s:
                in_ft['t2s'] = outs.pop(0)
            if run_s2t:
                in_ft['s2t'] = outs.pop(0)
            return in_ft
        vft = vft.detach().contiguous()
        x_t2s = in_ft['t2s'].detach().float().contiguous() if run_t2s else None
        x_s2t = in_ft['s2t'].detach().float().contiguous() if run_s2t else None
        Lq = (x_t2s if run_t2s else x_s2t).shape[1]
        qm = _mask_bytes(b.query_mask, B, Lq)
        tm = _mask_bytes(b.temporal_mask, B, T)
        o_t2s = torch.empty_like(x_t2s) if run_t2s else None
        o_s2t = torch.empty_like(x_s2t) if run_s2t else None
        packed = self.packed_weights()
        L = _lib.lib()
        ws = _ws(L.bist_st_layer_workspace_bytes(B, T, S, Lq, d, d_ff, h, nd, vdt), vft.device)
        _lib.check(L.bist_st_layer_fwd(_lib.ptr(x_t2s), _lib.ptr(x_s2t), _lib.ptr(vft), _lib.ptr(qm), _lib.ptr(tm),
                                       _lib.ptr(packed), _lib.ptr(o_t2s), _lib.ptr(o_s2t), B, T, S, Lq, d, d_ff, h,
                                       nd, dir_mask, float(self.sublayer[0].norm.eps), _lib.ptr(ws), ws.numel(),
                                       vdt, _lib.stream_ptr()), "bist_st_layer_fwd")
        if run_t2s:
            in_ft['t2s'] = o_t2s
        if run_s2t:
            in_ft['s2t'] = o_s2t
        return in_ft


class _DecoderSlice(nn.Module):
    """The members of MultimodalDecoder8 that belong to the video path (decoder.py:73-77)."""

    def __init__(self, v_layer, v_N, size):
        super(_DecoderSlice, self).__init__()
        self.v_layers = clones(v_layer, v_N)
        self.spatial_out_norm = LayerNorm(size)
        self.temporal_out_norm = LayerNorm(size)


class VideoPath(nn.Module):
    """vid_encoder + v_layers + output norms, keyed like the reference ``MTN.state_dict()``.

    forward(b, encoded_query) follows mtn.py:49-51 and decoder.py:107-129 and returns ``ft`` with
    'spatiotemporal_ft', 'temporal_ft' (if s2t) and 'spatial_ft' (if t2s), plus the raw streams
    under 't2s' / 's2t'.
    """

    def __init__(self, d_model=512, h=8, n_blocks=3, ft_size=2048, t2s=1, s2t=1, d_ff=None, dropout=0.1,
                 precision="fp32"):
        super(VideoPath, self).__init__()
        from argparse import Namespace
        from .modules import MultiHeadedAttention, PositionwiseFeedForward
        args = Namespace(t2s=t2s, s2t=s2t, enc_st_combine='none')
        self.args = args
        attn = MultiHeadedAttention(h, d_model)                              # mtn.py:77
        ff = PositionwiseFeedForward(d_model, d_ff or 4 * d_model, dropout)   # mtn.py:70,78
        nb_attn, nb_ff = (6, 2) if (t2s and s2t) else (3, 1)                 # mtn.py:123-128
        self.vid_encoder = VidEncoder8(nn.Linear(ft_size, d_model), None, None, n_blocks, 0, d_model, args)
        v_layer = VidEncoderLayer4(d_model, attn, nb_attn, ff, nb_ff, dropout, args)
        self.mutlimodal_decoder = _DecoderSlice(v_layer, n_blocks, d_model)   # (sic) the reference's spelling
        set_precision(self, precision)
        self.eval()

    def capture(self, b, encoded_query, warmup=3):
        """Capture one forward into a CUDA graph and return a ``GraphedVideoPath``.

        The kernels are stream-ordered, never synchronise and take every buffer from torch's caching
        allocator, so the whole path (projection + all blocks + output norms, ~80 launches) records into
        one graph.  ``graphed(b, encoded_query)`` copies the new inputs into the static input buffers and
        replays; ``graphed.replay()`` replays on whatever the static buffers hold.
        """
        return GraphedVideoPath(self, b, encoded_query, warmup)

    def forward(self, b, encoded_query):
        ft = {'encoded_query': encoded_query}
        ft = self.vid_encoder(b, ft)                                          # mtn.py:49-51
        dec = self.mutlimodal_decoder
        in_ft = {'t2s': encoded_query, 's2t': encoded_query}                  # decoder.py:109-110
        for layer in dec.v_layers:
            in_ft = layer(in_ft, ft, b)                                       # decoder.py:116
        if self.args.s2t:
            ft['temporal_ft'] = dec.temporal_out_norm(in_ft['s2t'])           # decoder.py:127
            ft['s2t'] = in_ft['s2t']
        if self.args.t2s:
            ft['spatial_ft'] = dec.spatial_out_norm(in_ft['t2s'])             # decoder.py:129
            ft['t2s'] = in_ft['t2s']
        return ft


class GraphedVideoPath:
    """CUDA-graph replay of ``VideoPath.forward`` for a fixed batch shape (launch overhead removed)."""

    def __init__(self, model, b, encoded_query, warmup=3):
        from types import SimpleNamespace
        _lib.require_cuda(b.fts, encoded_query)
        self.model = model
        self.static_b = SimpleNamespace(fts=b.fts.clone(), query_mask=b.query_mask.clone(),
                                        temporal_mask=b.temporal_mask.clone())
        self.static_eq = encoded_query.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):           # packs weights, sets kernel attributes, warms the allocator
                model(self.static_b, self.static_eq)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = model(self.static_b, self.static_eq)

    def load(self, b, encoded_query, non_blocking=True):
        "Copy a new batch (host or device tensors of the captured shapes) into the static input buffers."
        self.static_b.fts.copy_(b.fts, non_blocking=non_blocking)
        self.static_b.query_mask.copy_(b.query_mask, non_blocking=non_blocking)
        self.static_b.temporal_mask.copy_(b.temporal_mask, non_blocking=non_blocking)
        self.static_eq.copy_(encoded_query, non_blocking=non_blocking)

    def replay(self):
        self.graph.replay()
        return self.out

    def __call__(self, b, encoded_query):
        self.load(b, encoded_query)
        return self.replay()


class PipelinedVideoPath:
    """Host -> device -> host pipeline over ``depth`` captured copies of the forward: while one copy's graph runs, the
    next batch's pinned-host inputs are already crossing PCIe into the other copy's static buffers on a copy stream
    (H2D, compute and D2H use different engines).  Every submitted batch is still copied in, computed and copied out;
    only the waiting is overlapped, so the steady-state step time is max(H2D, compute) instead of their sum.

        pipe = PipelinedVideoPath(model, device_batch, device_eq)
        for host_batch, host_eq in loader:
            pipe.submit(host_batch, host_eq, out_temporal_pinned, out_spatial_pinned)
        pipe.drain()
    """

    def __init__(self, model, b, encoded_query, depth=2, warmup=3):
        self.slots = [GraphedVideoPath(model, b, encoded_query, warmup) for _ in range(depth)]
        self.copy_stream = torch.cuda.Stream()
        self.loaded = [torch.cuda.Event() for _ in range(depth)]
        self.free = [torch.cuda.Event() for _ in range(depth)]
        for e in self.free:
            e.record()
        self.n = 0

    def submit(self, b, encoded_query, out_temporal=None, out_spatial=None, fill=None):
        """``fill(slot)``: optional callable that writes the slot's static inputs itself (e.g. ``FeatureCache.collate``
        into ``slot.static_b.fts`` / ``.temporal_mask``); it runs on the copy stream.  Otherwise ``b`` and
        ``encoded_query`` (host or device tensors of the captured shapes) are copied in."""
        k = self.n % len(self.slots)
        g = self.slots[k]
        cur = torch.cuda.current_stream()
        self.copy_stream.wait_event(self.free[k])          # slot k's previous replay and read-back are done
        with torch.cuda.stream(self.copy_stream):
            if fill is not None:
                fill(g)
            else:
                g.load(b, encoded_query, non_blocking=True)
            self.loaded[k].record(self.copy_stream)
        cur.wait_event(self.loaded[k])
        out = g.replay()
        if out_temporal is not None:
            out_temporal.copy_(out['temporal_ft'], non_blocking=True)
        if out_spatial is not None:
            out_spatial.copy_(out['spatial_ft'], non_blocking=True)
        self.free[k].record(cur)
        self.n += 1
        return out

    def drain(self):
        torch.cuda.current_stream().synchronize()


class _MemoryEncoderLayer(nn.Module):
    """Self attention over the query-length stream, attention over one encoded memory, FFN — the common body of
    CapEncoderLayer and AudioEncoderLayer (encoder.py:203-235); compositions of the library-backed blocks."""
    stream, memory, memory_mask = None, None, None

    def __init__(self, size, attn, nb_attn, ff, dropout):
        super(_MemoryEncoderLayer, self).__init__()
        self.size = size
        self.attn = clones(attn, nb_attn)
        self.ff = ff
        self.sublayer = clones(SublayerConnection(size, dropout), nb_attn + 1)

    def forward(self, in_ft, ft, b):
        x = in_ft[self.stream]
        mem, mem_mask = ft[self.memory], getattr(b, self.memory_mask)
        x = self.sublayer[0](x, lambda y: self.attn[0](y, y, y, b.query_mask))
        x = self.sublayer[1](x, lambda y: self.attn[1](y, mem, mem, mem_mask))
        in_ft[self.stream] = self.sublayer[2](x, self.ff)
        return in_ft


class CapEncoderLayer(_MemoryEncoderLayer):
    "encoder.py:203-218 — in_ft['cap'] attends over ft['encoded_cap'] (b.cap_mask)."
    stream, memory, memory_mask = 'cap', 'encoded_cap', 'cap_mask'


class AudioEncoderLayer(_MemoryEncoderLayer):
    "encoder.py:220-235 — in_ft['audio'] attends over ft['encoded_audio'] (b.audio_mask)."
    stream, memory, memory_mask = 'audio', 'encoded_audio', 'audio_mask'


class Encoder(nn.Module):
    """The text "encoder" of encoder.py:11-40: one LayerNorm per sequence, consumed in order (lists of sequences take
    one norm per element, None passes through)."""

    def __init__(self, size, nb_layers):
        super(Encoder, self).__init__()
        self.norm = nn.ModuleList(LayerNorm(size) for _ in range(nb_layers))
        self.nb_layers = nb_layers

    def forward(self, *seqs):
        norms = iter(self.norm)
        out = []
        for s in seqs:
            if isinstance(s, list):
                out.append([next(norms)(t) for t in s])
            else:
                out.append(None if s is None else next(norms)(s))
        return out
