"""Drop-in mirror of /root/reference/model/mtn.py over libbist_b200.so: the ``MTN`` module (mtn.py:12-61) and
``make_model`` (mtn.py:63-167) with the reference's attribute names — ``text_encoder``, ``vid_encoder``,
``mutlimodal_decoder`` (sic), ``query_embed`` / ``tgt_embed`` (one shared module), ``generator``, ``ae_generator`` —
so that a reference ``state_dict`` loads with ``strict=True``.  Every arithmetic step of the forward, the loss
(optimize.py) and their backward runs in the library.

Built: ``enc_st_combine='none'``, ``ptr_gen=1`` with ``ptr_ft`` a comma list of query / his / cap, tied embeddings.
"""
import copy

import torch
import torch.nn as nn

from .decoder import make_decoder
from .encoder import Encoder, VidEncoder8, set_precision
from .generator import Generator, MultiPointerGenerator, PointerGenerator
from .modules import EmbedSequence, Embeddings, MultiHeadedAttention, PositionalEncoding


class MTN(nn.Module):
    "mtn.py:12-61."

    def __init__(self, text_encoder, vid_encoder, text_decoder, mutlimodal_decoder, query_embed, his_embed, cap_embed,
                 tgt_embed, generator, ae_generator, ptr_gen, args):
        super(MTN, self).__init__()
        self.text_encoder = text_encoder
        self.vid_encoder = vid_encoder
        self.text_decoder = text_decoder
        self.mutlimodal_decoder = mutlimodal_decoder
        self.query_embed = query_embed
        self.tgt_embed = tgt_embed
        self.generator = generator
        self.ae_generator = ae_generator
        self.ptr_gen = ptr_gen
        self.args = args

    def forward(self, b):
        return self.decode(b, self.encode(b))

    def encode(self, b):
        """mtn.py:36-52.  The text side (three embedding lookups and norms: small kernels) is enqueued on a branch stream
        beside the video projection (decoder.py has the rest of the model's stream branches); joined before the decoder."""
        from .decoder import _branch_streams_enabled, _loss_streams
        from .modules import refresh_weight_copies
        if not (b.fts.is_cuda and _branch_streams_enabled()):
            return self.encode_vid(b, self.encode_text(b, {}))
        main = torch.cuda.current_stream(b.fts.device)
        refresh_weight_copies(b.fts.device)
        s = _loss_streams(b.fts.device, 1)[0]
        s.wait_stream(main)
        with torch.cuda.stream(s):
            ft = self.encode_text(b, {})
        made = list(ft.values())
        ft = self.encode_vid(b, ft)
        main.wait_stream(s)
        for t in made:
            t.record_stream(main)
        return ft

    def encode_text(self, b, ft):
        ft['encoded_query'], ft['encoded_cap'], ft['encoded_his'] = self.text_encoder(
            self.query_embed(b.query), self.query_embed(b.cap), self.query_embed(b.his))
        return ft

    def encode_vid(self, b, ft):
        return self.vid_encoder(b, ft)

    def decode(self, b, ft):
        return self.multimodal_decode_text(b, ft)

    def multimodal_decode_text(self, b, ft):
        ft['encoded_tgt'] = self.tgt_embed(b.trg)
        return self.mutlimodal_decoder(b, ft, ft['encoded_tgt'])


def set_model_precision(model, precision):
    """fp32 or bf16 activations for the whole model (parameters stay fp32 masters)."""
    set_precision(model, precision)
    for m in model.modules():
        if isinstance(m, EmbedSequence):
            m.precision = precision
    return model


def make_model(src_vocab, tgt_vocab, args, ft_sizes=None, embeddings=None, precision="fp32"):
    "mtn.py:63-167."
    if not args.ptr_gen:
        raise NotImplementedError("ptr_gen=0 raises NameError in the reference (mtn.py:95); only ptr_gen=1 is built")
    ft_sizes = list(ft_sizes or [])
    if len(ft_sizes) not in (1, 2):
        raise NotImplementedError("ft_sizes = [video feature size] or [video, audio feature size] (mtn.py:110-115)")
    if getattr(args, "nb_aenc_blocks", 0) > 0 and len(ft_sizes) != 2:
        raise ValueError("nb_aenc_blocks > 0 needs the audio feature size: ft_sizes=[2048, 128] (mtn.py:112-113)")
    d_model, dropout = args.d_model, args.dropout
    position = PositionalEncoding(d_model, dropout)
    query_embed = EmbedSequence(Embeddings(d_model, src_vocab), copy.deepcopy(position))
    tgt_embed = query_embed
    names = args.ptr_ft.split(',')
    if len(names) > 1:
        pointer_attn = nn.ModuleList(MultiHeadedAttention(1, d_model, dropout=0) for _ in names)
        generator = MultiPointerGenerator(d_model, tgt_embed[0].lut.weight, pointer_attn, len(names))
    else:
        generator = PointerGenerator(d_model, tgt_embed[0].lut.weight, MultiHeadedAttention(1, d_model, dropout=0))
    ae_generator = Generator(d_model, tgt_vocab, tgt_embed[0].lut.weight) if args.auto_encoder else None
    text_encoder = Encoder(d_model, nb_layers=3)
    a_W = nn.Linear(ft_sizes[1], d_model) if len(ft_sizes) > 1 else None          # mtn.py:112-115
    vid_encoder = VidEncoder8(nn.Linear(ft_sizes[0], d_model), a_W, None, args.nb_venc_blocks,
                              getattr(args, "nb_aenc_blocks", 0), d_model, args)
    model = MTN(text_encoder, vid_encoder, None, make_decoder(args, precision), query_embed, None, None, tgt_embed,
                generator, ae_generator, args.ptr_gen, args)
    for p in model.parameters():                       # mtn.py:163-165
        if p.dim() > 1:
            nn.init.xavier_uniform_(p)
    return set_model_precision(model, precision)
