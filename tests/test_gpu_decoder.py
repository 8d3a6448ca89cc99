"""GPU parity of the decoder side (run with -m gpu on the B200 box): bist_b200.decoder's
MultimodalDecoder8 / MultimodalDecoderLayer12 and encoder's CapEncoderLayer / AudioEncoderLayer

  * forward against tests/golden/decoder_*.npz — outputs of the UNMODIFIED reference on the same seeded inputs and
    weights (oracle/make_golden.py): fp32 <= 1e-4, bf16 <= 2e-2 (max|a-b| / max|b| per tensor);
  * the mirror's state_dict keys and shapes against the reference's (stored in the fixture), so a reference
    checkpoint loads strict=True;
  * gradients (eval mode, and train() mode with the kernels' dropout masks injected into the oracle) against
    torch.autograd over the CPU oracle: fp32 <= 1e-4 on every parameter and input gradient.
"""
import argparse
import glob
import json
import os

import pytest
import torch

from bist_b200 import _lib, make_decoder
from oracle import bist_oracle as O
from tests.test_oracle_decoder import load_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "decoder_*.npz")))
DT = {"fp32": torch.float32, "bf16": torch.bfloat16}


def _args(c):
    return argparse.Namespace(
        nb_blocks=c["n_blocks"], nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"], nb_aenc_blocks=c["nb_aenc"],
        d_model=c["d"], att_h=c["h"], dropout=0.1, t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none",
        enc_vc_combine=c["vc"], dec_st_combine=c["dec_st"], include_caption=c["cap"])


def _build(c, P, precision):
    dec = make_decoder(_args(c), precision)
    sd = dec.state_dict()
    assert sorted(sd) == sorted(P), set(sd) ^ set(P)                       # the reference's keys, exactly
    assert all(tuple(sd[k].shape) == tuple(P[k].shape) for k in P)
    dec.load_state_dict(P, strict=True)
    return dec.to(DEV)


def _inputs(ft, x, precision, grad=False):
    dt = DT[precision]
    ftd = {k: v.to(DEV, dt).requires_grad_(grad) for k, v in ft.items()}
    return ftd, x.to(DEV, dt).requires_grad_(grad)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[8:-4])
def test_decoder_forward_matches_reference_outputs(path, precision):
    c, cfg, P, b, ft, x, masks, gold = load_case(path)
    dec = _build(c, P, precision).eval()
    ftd, xd = _inputs(ft, x, precision)
    n0 = _lib.lib().bist_launch_count()
    with torch.no_grad():
        out = dec(b.to(DEV), ftd, xd)
    torch.cuda.synchronize()
    assert _lib.lib().bist_launch_count() > n0
    tol = 1e-4 if precision == "fp32" else 2e-2
    for k, g in gold.items():
        assert O.norm_err(out[k].float().cpu(), g) < tol, (k, O.norm_err(out[k].float().cpu(), g))


def _loss(out, seed):
    w = torch.randn(out["decoded_text"].shape, generator=torch.Generator().manual_seed(seed))
    return (out["decoded_text"].float() * w.to(out["decoded_text"].device)).sum()


GRAD_INPUTS = ("encoded_query", "encoded_his", "encoded_cap", "encoded_audio", "spatiotemporal_ft")


def _compare_grads(dec, ftd, xd, Pg, ftr, xr, c):
    scale = max(float(g.abs().max()) for g in (p.grad for p in Pg.values()) if g is not None)
    for name, p in dec.named_parameters():
        ref = Pg[name].grad
        if ref is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
            continue
        assert p.grad is not None, name
        if name.endswith("linears.1.bias"):                 # zero by softmax shift invariance: rounding noise only
            assert float(p.grad.abs().max()) < 1e-4 * scale, name
            continue
        assert O.norm_err(p.grad.cpu(), ref) < 1e-4, (name, O.norm_err(p.grad.cpu(), ref))
    assert O.norm_err(xd.grad.cpu(), xr.grad) < 1e-4
    for k in GRAD_INPUTS:
        if ftr[k].grad is not None:
            assert O.norm_err(ftd[k].grad.cpu(), ftr[k].grad) < 1e-4, k


@pytest.mark.parametrize("name", ["avsd_dyn", "audio_dyn", "nocap_par", "cenc_aenc"])
def test_decoder_gradients_match_oracle(name):
    path = [p for p in GOLD if p.endswith("decoder_%s.npz" % name)][0]
    c, cfg, P, b, ft, x, masks, gold = load_case(path)
    dec = _build(c, P, "fp32").eval()
    ftd, xd = _inputs(ft, x, "fp32", grad=True)
    _loss(dec(b.to(DEV), dict(ftd), xd), 5).backward()
    torch.cuda.synchronize()
    Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    ftr = {k: v.clone().requires_grad_(True) for k, v in ft.items()}
    xr = x.clone().requires_grad_(True)
    _loss(O.multimodal_decoder8(Pg, "", c["h"], cfg, ftr, xr, masks, c["n_blocks"]), 5).backward()
    _compare_grads(dec, ftd, xd, Pg, ftr, xr, c)


def test_decoder_train_mode_dropout_matches_oracle_with_the_same_masks():
    """train() mode on the caption / audio / decoder layers (the video blocks stay in eval: their dropout is covered by
    tests/test_gpu_backward.py): every module's mask is rebuilt from module.last_dropout and injected into the oracle."""
    path = [p for p in GOLD if p.endswith("decoder_audio_dyn.npz")][0]
    c, cfg, P, b, ft, x, masks, gold = load_case(path)
    dec = _build(c, P, "fp32").train()
    for v in dec.v_layers:
        v.eval()
    ftd, xd = _inputs(ft, x, "fp32", grad=True)
    torch.manual_seed(4321)
    out = dec(b.to(DEV), dict(ftd), xd)
    _loss(out, 6).backward()
    torch.cuda.synchronize()
    used = []

    def drop(site, t):
        if ".v_layers." in "." + site:
            return t
        mod_name = site.rsplit(".", 1)[0]                    # 'layers.0.attn.1.p' -> module 'layers.0.attn.1'
        cfg_d = dec.get_submodule(mod_name).last_dropout
        assert cfg_d is not None, site
        p, seed, s_id = cfg_d
        buf = torch.empty(t.numel(), dtype=torch.float32, device=DEV)
        _lib.check(_lib.lib().bist_dropout_mask(seed, s_id, p, t.numel(), _lib.ptr(buf), _lib.stream_ptr()), "bist_dropout_mask")
        used.append(site)
        return t * buf.cpu().view(t.shape)

    Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    ftr = {k: v.clone().requires_grad_(True) for k, v in ft.items()}
    xr = x.clone().requires_grad_(True)
    ref = O.multimodal_decoder8(Pg, "", c["h"], cfg, ftr, xr, masks, c["n_blocks"], drop=drop)
    _loss(ref, 6).backward()
    assert len(used) > 20
    assert O.norm_err(out["decoded_text"].detach().cpu(), ref["decoded_text"].detach()) < 1e-4
    _compare_grads(dec, ftd, xd, Pg, ftr, xr, c)
