"""BASELINE config 3 — the full BiST AVSD training step — as a ready-made model + batch: ``make_model`` (mtn.py) at
exec.sh's settings for BASELINE's d_model (3 decoder / 3 video / 3 caption blocks, softmax gate, pointer generator
over query and caption, auto-encoder heads, label smoothing 0.1) and the seeded synthetic batch of SURVEY.md §8(d).
Used by bench.py (``train_full``) and scripts/core_bench.py.
"""
from argparse import Namespace

from . import synth
from .generator import LabelSmoothing
from .mtn import make_model
from .optimize import SimpleLossCompute

FULL_SHAPES = dict(B=64, T=40, S=16, D=2048, d=512, h=8, Lq=16, Lh=64, Lc=32, Lt=12, V=4096,
                   Ta=40, Da=128)                            # audio stream: Ta = T clips of 128-d VGGish features (§8(c), (d))
PAD, SMOOTHING = 1, 0.1                                                                        # train.py:93


def full_args(d_model=512, h=8, n_blocks=3, dropout=0.1, audio=False):
    """exec.sh's settings; ``audio=True`` adds the audio encoder config 3 names (nb_aenc_blocks = n_blocks,
    ft_sizes=[2048, 128], args.noW_venc=0 as SURVEY.md §8(c) injects it; train_configs.py:30 defaults to 0 blocks)."""
    return Namespace(nb_blocks=n_blocks, nb_venc_blocks=n_blocks, nb_cenc_blocks=n_blocks,
                     nb_aenc_blocks=n_blocks if audio else 0, noW_venc=0,
                     d_model=d_model, att_h=h, dropout=dropout, ptr_gen=1, ptr_ft="query,cap", auto_encoder=1,
                     t2s=1, s2t=1, enc_st_combine="none", enc_vc_combine="dyn", dec_st_combine="seq",
                     include_caption="summary", mask_unk=1)


def full_model(precision="fp32", shapes=None, args=None, audio=False):
    """(model, loss_compute) for config 3; ``loss_compute.loss(model(batch), batch)`` is the training loss."""
    c = dict(FULL_SHAPES, **(shapes or {}))
    args = args or full_args(c["d"], c["h"], audio=audio)
    ft_sizes = [c["D"]] + ([c["Da"]] if getattr(args, "nb_aenc_blocks", 0) > 0 else [])
    model = make_model(c["V"], c["V"], args, ft_sizes=ft_sizes, precision=precision)
    lc = SimpleLossCompute(model.generator, model.ae_generator, LabelSmoothing(c["V"], PAD, SMOOTHING), args=args)
    return model, lc


def full_batch(seed, device, shapes=None, audio=False):
    c = dict(FULL_SHAPES, **(shapes or {}))
    return synth.make_full_batch(seed, c["B"], c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                                 Ta=c["Ta"] if audio else 0, Da=c["Da"] if audio else 0).to(device)
