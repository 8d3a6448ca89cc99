"""Drop-in mirrors of /root/reference/model/decoder.py over libbist_b200.so.

``MultimodalDecoderLayer12`` (decoder.py:11-60) and ``MultimodalDecoder8`` (decoder.py:62-186) keep the reference's
constructor signatures, forward signatures, attribute names and therefore ``state_dict`` keys, so
``model/mtn.py:make_model`` can build its decoder from these classes and a reference checkpoint loads with
``strict=True``.  They are compositions of the library-backed blocks of modules.py (every matmul, softmax, LayerNorm,
dropout and the softmax gate run in libbist_b200.so, forward and backward); the video blocks inside are the fused
``VidEncoderLayer4`` of encoder.py.

Scope: ``enc_st_combine == 'none'`` (train_configs.py:40 and exec.sh).  The other values make the reference overwrite
``ft['spatiotemporal_ft']`` with a ``[B,Lq,d]`` tensor and only run with a single block; they raise
NotImplementedError here.  The control flow is organised as a table of attention memories (``_memory_plan``) instead
of the reference's nested branches; ``oracle/bist_oracle.py:decoder_memories`` restates the same table and is pinned
against the unmodified reference by tests/golden/decoder_*.npz.
"""
import torch
import torch.nn as nn

from . import _lib
import contextlib
import os

from .modules import LayerNorm, SublayerConnection, clones, gate_combine, linear, refresh_weight_copies, _BF16Cache

_BRANCH = {}


def _branch_streams_enabled():
    on = os.environ.get("BIST_BRANCH_STREAMS", "1") != "0"
    if on and not _BRANCH.get("quiet"):
        # parameters shared by modules on different branches (the tied embedding, the text norms) accumulate gradients
        # that arrive from several streams: intended here, autograd synchronises them
        quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
        if quiet is not None:
            quiet(False)
        _BRANCH["quiet"] = True
    return on


def _decoder_stream(device):
    """The stream of the decoder-layer chain (``MultimodalDecoder8.forward``)."""
    key = ("dec", torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device())
    if key not in _BRANCH:
        _BRANCH[key] = torch.cuda.Stream(device=device, priority=int(os.environ.get("BIST_BRANCH_PRIO", "0")))
    return _BRANCH[key]


def _loss_streams(device, n):
    """Streams for the auto-encoder heads of ``SimpleLossCompute.loss``."""
    key = ("loss", torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device())
    have = _BRANCH.setdefault(key, [])
    while len(have) < n:
        have.append(torch.cuda.Stream(device=device))
    return have[:n]


def _branch_streams(device):
    """Two long-lived streams per device for the caption and audio layers of a block (``_block_memories``)."""
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if key not in _BRANCH:
        prio = int(os.environ.get("BIST_BRANCH_PRIO", "0"))
        _BRANCH[key] = (torch.cuda.Stream(device=device, priority=prio), torch.cuda.Stream(device=device, priority=prio))
    return _BRANCH[key]


def _opt(args, name, default):
    return getattr(args, name, default)


def _require_plain_st(args):
    if _opt(args, "enc_st_combine", "none") != "none":
        raise NotImplementedError(
            "bist_b200 builds the decoder side for enc_st_combine='none' only (got %r)" % args.enc_st_combine)


def _memory_plan(args):
    """[(ft key, mask attribute of the batch)] attended after the self / history / query attentions, and whether
    temporal_ft and spatial_ft are attended from the same input and summed (decoder.py:25-57)."""
    v = args.nb_venc_blocks > 0
    c = args.nb_cenc_blocks > 0
    a = _opt(args, "nb_aenc_blocks", 0) > 0
    if v and c and _opt(args, "enc_vc_combine", "none") != "none":
        return [("encoded_ft", "query_mask")], False
    plan, parallel = [], False
    if args.include_caption != "none":
        plan.append(("cap_ft", "query_mask") if c else ("encoded_cap", "cap_mask"))
    if v:
        if args.dec_st_combine == "seq":
            plan += [(k, "query_mask") for k, on in (("temporal_ft", _opt(args, "s2t", 1)), ("spatial_ft", _opt(args, "t2s", 1))) if on]
        else:
            plan += [("temporal_ft", "query_mask"), ("spatial_ft", "query_mask")]
            parallel = True
    if a:
        plan.append(("audio_ft", "query_mask"))
    return plan, parallel


class MultimodalDecoderLayer12(nn.Module):
    "decoder.py:11-60 — self, history and query attention, then the multimodal memories, then the FFN."

    def __init__(self, size, attn, nb_attn, ff, dropout, args):
        super(MultimodalDecoderLayer12, self).__init__()
        self.size = size
        self.attn = clones(attn, nb_attn)
        self.ff = ff
        self.sublayer = clones(SublayerConnection(size, dropout), nb_attn + 1)
        self.args = args

    def _attend(self, i, x, mem, mask):
        att = self.attn[i]
        if mem is None:
            return self.sublayer[i](x, lambda y: att(y, y, y, mask))
        return self.sublayer[i](x, lambda y: att(y, mem, mem, mask))

    def forward(self, b, ft, x):
        _require_plain_st(self.args)
        x = self._attend(0, x, None, b.trg_mask)
        x = self._attend(1, x, ft['encoded_his'], b.his_mask)
        x = self._attend(2, x, ft['encoded_query'], b.query_mask)
        plan, parallel = _memory_plan(self.args)
        i = 3
        for key, mname in plan:
            if parallel and key == "spatial_ft":
                continue                                    # consumed together with temporal_ft below
            if parallel and key == "temporal_ft":           # both streams attend from the same input (decoder.py:47-52)
                x = (self._attend(i, x, ft['temporal_ft'], b.query_mask)
                     + self._attend(i + 1, x, ft['spatial_ft'], b.query_mask))
                i += 2
                continue
            x = self._attend(i, x, ft[key], getattr(b, mname))
            i += 1
        return self.sublayer[i](x, self.ff)


class MultimodalDecoder8(nn.Module):
    "decoder.py:62-186."

    def __init__(self, v_layer, c_layer, a_layer, layer, venc_N, cenc_N, aenc_N, N, args):
        super(MultimodalDecoder8, self).__init__()
        self.layers = clones(layer, N)
        self.N = N
        self.v_N = venc_N
        self.c_N = cenc_N
        self.a_N = aenc_N
        self.norm = LayerNorm(layer.size)
        self.args = args
        _require_plain_st(args)
        s2t, t2s = _opt(args, "s2t", 1), _opt(args, "t2s", 1)
        if self.v_N > 0:
            self.v_layers = clones(v_layer, self.v_N)
            self.spatial_out_norm = LayerNorm(v_layer.size)
            self.temporal_out_norm = LayerNorm(v_layer.size)
        if self.c_N > 0:
            self.c_layers = clones(c_layer, self.c_N)
            self.cap_out_norm = LayerNorm(c_layer.size)
        if self.a_N > 0:
            self.a_layers = clones(a_layer, self.a_N)
            self.a_out_norm = LayerNorm(a_layer.size)
        if self.v_N > 0 and _opt(args, "enc_vc_combine", "none") == 'dyn':
            # gate input = [encoded_query, cap?, spatial?, temporal?, audio?]; one logit per stream but the query
            n_in = 1 + int(args.include_caption != 'none') + int(bool(t2s)) + int(bool(s2t)) + int(self.a_N > 0)
            self.vc_combine_W = nn.Linear(v_layer.size * n_in, n_in - 1)
        self._cache = _BF16Cache()

    def __deepcopy__(self, memo):
        import copy
        new = type(self).__new__(type(self))
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = _BF16Cache() if k == "_cache" else copy.deepcopy(v, memo)
        return new

    def _combine(self, ft):
        """ft['encoded_ft'] (decoder.py:137-167): plain sum, or the softmax gate over the encoder streams."""
        args = self.args
        s2t, t2s = bool(_opt(args, "s2t", 1)), bool(_opt(args, "t2s", 1))
        if args.enc_vc_combine == 'sum':
            return ft['temporal_ft'] + ft['spatial_ft'] + ft['cap_ft']
        gate_in = [ft['encoded_query'], ft['cap_ft']]
        if t2s:
            gate_in.append(ft['spatial_ft'])
        if s2t:
            gate_in.append(ft['temporal_ft'])
        if self.a_N > 0:
            gate_in.append(ft['audio_ft'])
        if s2t and t2s:
            streams = [ft['temporal_ft'], ft['spatial_ft'], ft['cap_ft']]
        elif s2t:
            streams = [ft['temporal_ft'], ft['cap_ft']]
        else:
            streams = [ft['spatial_ft'], ft['cap_ft']]
        if self.a_N > 0:
            if len(streams) != 3:
                raise NotImplementedError("the audio stream takes gate column 3 (decoder.py:167): needs both s2t and t2s")
            streams.append(ft['audio_ft'])
        dt = ft['encoded_query'].dtype                       # the video streams come back fp32 from VidEncoderLayer4
        gate_in = [t.to(dt) for t in gate_in]
        streams = [t.to(dt) for t in streams]
        logits = linear(torch.cat(gate_in, dim=-1), self.vc_combine_W, self._cache)
        if logits.shape[-1] != len(streams):
            raise RuntimeError("vc_combine_W has %d outputs for %d streams" % (logits.shape[-1], len(streams)))
        return gate_combine(logits, streams)

    def _block_memories(self, count, in_ft, ft, b):
        """Block `count` of the query-conditioned encoders (decoder.py:115-167): video, caption and audio layers, their
        output norms and the gate.  None of it depends on the target sequence.

        The three layers of a block read different streams of ``in_ft`` and never each other's outputs (they meet in the
        gate), so on a CUDA device the caption and audio layers are enqueued on two branch streams beside the video layer
        — their ~20 small launches each hide behind the video layer's large kernels, in eager mode and as parallel
        branches of a captured CUDA graph alike (autograd runs each backward on its forward's stream).  The branches are
        joined before the gate.  ``BIST_BRANCH_STREAMS=0`` keeps everything on the current stream."""
        args = self.args
        eq = ft['encoded_query']
        side = (self.v_N > 0 and (self.c_N > 0 or self.a_N > 0) and eq.is_cuda and _branch_streams_enabled())
        if side:
            main = torch.cuda.current_stream(eq.device)
            s_cap, s_aud = _branch_streams(eq.device)
            refresh_weight_copies(eq.device)        # on the main stream, before any branch reads the shared arena
            for on, s in ((self.c_N > 0, s_cap), (self.a_N > 0, s_aud)):
                if on:
                    s.wait_stream(main)
        cap_stream = torch.cuda.stream(s_cap) if side else contextlib.nullcontext()
        aud_stream = torch.cuda.stream(s_aud) if side else contextlib.nullcontext()
        made = []
        if self.c_N > 0:
            with cap_stream:
                in_ft = self.c_layers[count](in_ft, ft, b)
                ft['cap_ft'] = self.cap_out_norm(in_ft['cap'])
            made += [in_ft['cap'], ft['cap_ft']]
        if self.a_N > 0:
            with aud_stream:
                in_ft = self.a_layers[count](in_ft, ft, b)
                ft['audio_ft'] = self.a_out_norm(in_ft['audio'])
            made += [in_ft['audio'], ft['audio_ft']]
        # the video layer last: its large kernels fill in around the branches' chains of small ones (enqueued the other
        # way round the chains queue up behind them, in a captured graph too: 12.5 vs 11.8 ms for the full training step)
        if self.v_N > 0:
            in_ft = self.v_layers[count](in_ft, ft, b)
            if _opt(args, "s2t", 1):
                ft['temporal_ft'] = self.temporal_out_norm(in_ft['s2t'])
            if _opt(args, "t2s", 1):
                ft['spatial_ft'] = self.spatial_out_norm(in_ft['t2s'])
        if side:
            if self.c_N > 0:
                main.wait_stream(s_cap)
            if self.a_N > 0:
                main.wait_stream(s_aud)
            for t in made:                          # allocated on a branch stream, read on the main one from here on
                t.record_stream(main)
        if self.v_N > 0 and self.c_N > 0 and _opt(args, "enc_vc_combine", "none") in ('sum', 'dyn'):
            ft['encoded_ft'] = self._combine(ft)
        return in_ft

    def forward(self, b, ft, x):
        """decoder.py:107-186.  Decoder layer k needs the memories of block k, but block k+1's video / caption / audio
        layers do not need decoder layer k: on a CUDA device the chain of decoder layers (a few dozen small launches each)
        runs on its own stream, one block behind the encoders' large kernels, and is joined before the final norm
        (``BIST_DECODER_STREAM=0`` keeps it on the current stream)."""
        eq = ft['encoded_query']
        in_ft = {'t2s': eq, 's2t': eq, 'audio': eq, 'cap': eq}
        ahead = (eq.is_cuda and len(self.layers) > 1 and (self.v_N > 0 or self.c_N > 0 or self.a_N > 0)
                 and _branch_streams_enabled() and os.environ.get("BIST_DECODER_STREAM", "1") != "0")
        if ahead:
            main = torch.cuda.current_stream(eq.device)
            s_dec = _decoder_stream(eq.device)
            refresh_weight_copies(eq.device)
        for count, layer in enumerate(self.layers):
            in_ft = self._block_memories(count, in_ft, ft, b)
            if ahead:
                s_dec.wait_stream(main)                       # block `count`'s memories are ready
                read = [t for t in ft.values() if torch.is_tensor(t)] + [x]
                with torch.cuda.stream(s_dec):
                    x = layer(b, ft, x)
                for t in read:                                # the next block may drop them while s_dec still reads
                    t.record_stream(s_dec)
            else:
                x = layer(b, ft, x)
        if ahead:
            main.wait_stream(s_dec)
            x.record_stream(main)
        ft['decoded_text'] = self.norm(x)
        return ft

    # ---- decoding with the target-independent part cached ----------------------------------
    def encode_memories(self, b, ft):
        """Run every block's video / caption / audio layers and gate ONCE and return one ft snapshot per block — what
        each decoder layer attends over.  The reference recomputes all of this for every hypothesis at every step of
        ``beam_search_decode`` (decode.py:63-66 calls the whole ``MultimodalDecoder8.forward``)."""
        eq = ft['encoded_query']
        in_ft = {'t2s': eq, 's2t': eq, 'audio': eq, 'cap': eq}
        snaps = []
        for count in range(len(self.layers)):
            in_ft = self._block_memories(count, in_ft, ft, b)
            snaps.append(dict(ft))
        return snaps

    def decode_with_memories(self, b, snaps, x):
        """``decoded_text`` for target embeddings x given the cached per-block memories (decoder.py:182-185)."""
        for layer, ft_k in zip(self.layers, snaps):
            x = layer(b, ft_k, x)
        return self.norm(x)


def make_decoder(args, precision="fp32"):
    """The ``mutlimodal_decoder`` that mtn.py:make_model builds (mtn.py:66-153), from the library-backed classes:
    same clone counts (``nb_attn`` of the decoder layer, mtn.py:136-150) and therefore the same ``state_dict`` keys."""
    import copy
    from .encoder import AudioEncoderLayer, CapEncoderLayer, VidEncoderLayer4, set_precision
    from .modules import MultiHeadedAttention, PositionwiseFeedForward
    _require_plain_st(args)
    d_model, h, dropout = args.d_model, args.att_h, args.dropout
    c = copy.deepcopy
    attn = MultiHeadedAttention(h, d_model)
    ff = PositionwiseFeedForward(d_model, 4 * d_model, dropout)
    both = bool(_opt(args, "t2s", 1)) and bool(_opt(args, "s2t", 1))
    v_layer = VidEncoderLayer4(d_model, c(attn), 6 if both else 3, c(ff), 2 if both else 1, dropout, args)
    c_layer = CapEncoderLayer(d_model, c(attn), 2, c(ff), dropout)
    a_layer = AudioEncoderLayer(d_model, c(attn), 2, c(ff), dropout)
    nb_attn = 3
    if args.nb_cenc_blocks > 0 and args.nb_venc_blocks > 0 and _opt(args, "enc_vc_combine", "none") != 'none':
        nb_attn += 1
    else:
        nb_attn += int(args.nb_cenc_blocks > 0) + int(_opt(args, "nb_aenc_blocks", 0) > 0)
        nb_attn += 2 if args.nb_venc_blocks > 0 else 0
    layer = MultimodalDecoderLayer12(d_model, c(attn), nb_attn, c(ff), dropout, args)
    dec = MultimodalDecoder8(v_layer, c_layer, a_layer, layer, args.nb_venc_blocks, args.nb_cenc_blocks,
                             _opt(args, "nb_aenc_blocks", 0), args.nb_blocks, args)
    for m in dec.modules():                        # nobody reads .attn of these attentions (only the generator's pointer
        if isinstance(m, MultiHeadedAttention):    # attention is read, generator.py:62-63): skip storing the probabilities
            m.store_attn = False
    set_precision(dec, precision)
    return dec
