"""The model core at the shapes of BASELINE config 3: what ``MTN.forward`` (mtn.py:31-61) runs between the text
encoder and the generator — ``vid_encoder`` (VidEncoder8) followed by ``mutlimodal_decoder`` (MultimodalDecoder8 with
its video, caption and decoder layers) — built from the library-backed classes with the reference's attribute names.
The embeddings / text encoder and the generator + loss are not built yet: their outputs
and the loss are stood in for by synthetic activations and an MSE in bench.py.
"""
from argparse import Namespace

import torch
import torch.nn as nn

from . import synth
from .decoder import make_decoder
from .encoder import VidEncoder8, set_precision

CORE_SHAPES = dict(B=64, T=40, S=16, D=2048, d=512, h=8, Lq=16, Lh=64, Lc=32, Lt=12, Ta=40)


def core_args(d_model=512, h=8, n_blocks=3, dropout=0.1, nb_cenc=3, nb_aenc=0, vc="dyn", cap="summary"):
    """The args fields the core reads, at exec.sh's defaults for BASELINE's d_model."""
    return Namespace(nb_blocks=n_blocks, nb_venc_blocks=n_blocks, nb_cenc_blocks=nb_cenc, nb_aenc_blocks=nb_aenc,
                     d_model=d_model, att_h=h, dropout=dropout, t2s=1, s2t=1, enc_st_combine="none",
                     enc_vc_combine=vc, dec_st_combine="seq", include_caption=cap)


class BiSTCore(nn.Module):
    def __init__(self, args=None, ft_size=2048, precision="fp32"):
        super(BiSTCore, self).__init__()
        self.args = args = args or core_args()
        self.vid_encoder = VidEncoder8(nn.Linear(ft_size, args.d_model), None, None, args.nb_venc_blocks, 0,
                                       args.d_model, args)
        self.mutlimodal_decoder = make_decoder(args, precision)               # (sic) the reference's spelling
        for p in self.parameters():                                           # mtn.py:163-165
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)
        set_precision(self, precision)

    def forward(self, b, ft, encoded_tgt):
        ft = self.vid_encoder(b, ft)                                          # mtn.py:49-51
        return self.mutlimodal_decoder(b, ft, encoded_tgt)                    # mtn.py:57-61


def core_inputs(seed, device, precision="fp32", shapes=None):
    """Seeded config-3 batch: video features and masks, the text encoder's outputs as leaf activations that take
    gradients, the target-side embedding and an MSE target.  Returns (batch, ft, encoded_tgt, target)."""
    c = dict(CORE_SHAPES, **(shapes or {}))
    dt = torch.float32 if precision == "fp32" else torch.bfloat16
    vb = synth.make_batch(seed, c["B"], c["T"], c["S"], c["D"], c["Lq"])
    b, ft, x = synth.make_decoder_inputs(seed, c["B"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["Ta"], c["T"], c["S"], 8)
    b.fts, b.temporal_mask, b.query_mask = vb.fts, vb.temporal_mask, vb.query_mask
    g = torch.Generator().manual_seed(seed)
    ft = {k: torch.randn(c["B"], L, c["d"], generator=g).to(device, dt).requires_grad_(True)
          for k, L in (("encoded_query", c["Lq"]), ("encoded_his", c["Lh"]), ("encoded_cap", c["Lc"]))}
    x = torch.randn(c["B"], c["Lt"], c["d"], generator=g).to(device, dt).requires_grad_(True)
    tgt = torch.randn(c["B"], c["Lt"], c["d"], generator=g).to(device)
    return b.to(device), ft, x, tgt
