"""GPU parity of the whole model (run with -m gpu on the B200 box): bist_b200.make_model
(MTN with embeddings, text-encoder norms, video encoder, multimodal decoder, pointer generators) and
SimpleLossCompute (label-smoothed KL on the generator and the auto-encoder heads, fused forward + gradient kernel)

  * against tests/golden/model_*.npz — ``MTN.forward``, ``generator`` log-probabilities ("logits") and the summed
    loss of the UNMODIFIED reference on the same seeded batch and weights: fp32 <= 1e-4, bf16 <= 2e-2;
  * the mirror's state_dict keys and shapes against the reference's (a reference checkpoint loads strict=True);
  * every parameter gradient of the total loss against torch.autograd over the CPU oracle (fp32 <= 1e-4);
  * greedy token ids (argmax of the log-probabilities at every position) identical to the reference's in fp32.
"""
import argparse
import glob
import os

import pytest
import torch

from bist_b200 import LabelSmoothing, SimpleLossCompute, _lib, make_model
from oracle import bist_oracle as O
from tests.test_oracle_model import PAD, SMOOTHING, load_model_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "model_*.npz")))


def _args(c):
    return argparse.Namespace(
        nb_blocks=c["n_blocks"], nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"],
        nb_aenc_blocks=c.get("nb_aenc", 0), noW_venc=0, d_model=c["d"], att_h=c["h"], dropout=0.1, ptr_gen=1, ptr_ft=c["ptr_ft"], auto_encoder=c["ae"],
        t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none", enc_vc_combine=c["vc"],
        dec_st_combine="seq", include_caption=c["cap"], mask_unk=1)


def _build(c, P, shapes, precision):
    m = make_model(c["V"], c["V"], _args(c), ft_sizes=[c["D"]] + ([c["Da"]] if c.get("nb_aenc", 0) else []), precision=precision)
    sd = m.state_dict()
    assert sorted(sd) == sorted(shapes), set(sd) ^ set(shapes)             # the reference's keys, buffers included
    assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in shapes)
    missing = m.load_state_dict(P, strict=False)
    assert all(k.endswith(".pe") for k in missing.missing_keys) and not missing.unexpected_keys
    return m.to(DEV)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-4])
def test_model_forward_logits_and_loss_match_reference(path, precision):
    c, args, P, shapes, b, gold = load_model_case(path)
    m = _build(c, P, shapes, precision).eval()
    bd = b.to(DEV)
    n0 = _lib.lib().bist_launch_count()
    with torch.no_grad():
        ft = m(bd)
        logp = m.generator(ft, bd, m.args)
        total, parts = SimpleLossCompute(m.generator, m.ae_generator, LabelSmoothing(c["V"], PAD, SMOOTHING), args=m.args).loss(ft, bd)
    torch.cuda.synchronize()
    assert _lib.lib().bist_launch_count() > n0
    tol = 1e-4 if precision == "fp32" else 2e-2
    for k in ("encoded_query", "encoded_his", "encoded_cap", "encoded_tgt", "decoded_text", "encoded_audio", "audio_ft"):
        if k not in gold:
            continue
        assert O.norm_err(ft[k].float().cpu(), gold[k]) < tol, (k, O.norm_err(ft[k].float().cpu(), gold[k]))
    assert O.norm_err(logp.cpu(), gold["logp"]) < tol, O.norm_err(logp.cpu(), gold["logp"])
    assert abs(float(total) - float(gold["loss_total"])) < tol * abs(float(gold["loss_total"]))
    assert abs(float(parts["out"] * bd.ntokens) - float(gold["loss_out_sum"])) < tol * abs(float(gold["loss_out_sum"]))
    if precision == "fp32":                              # "greedy-decoded token IDs identical in fp32"
        assert torch.equal(logp.argmax(-1).cpu(), gold["logp"].argmax(-1))


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-4])
def test_model_loss_gradients_match_oracle(path):
    c, args, P, shapes, b, gold = load_model_case(path)
    m = _build(c, P, shapes, "fp32").eval()
    bd = b.to(DEV)
    total, _ = SimpleLossCompute(m.generator, m.ae_generator, LabelSmoothing(c["V"], PAD, SMOOTHING), args=m.args).loss(m(bd), bd)
    total.backward()
    torch.cuda.synchronize()
    leaves = {}
    Pg = {}
    for k, v in P.items():                               # tied keys share one leaf
        if id(v) not in leaves:
            leaves[id(v)] = v.clone().requires_grad_(True)
        Pg[k] = leaves[id(v)]
    ft = O.mtn_forward(Pg, args, b, c["h"], c["n_blocks"])
    ref_total, _ = O.loss_compute(Pg, args, ft, b, PAD, SMOOTHING)
    assert abs(float(total.detach()) - float(ref_total.detach())) < 1e-4 * abs(float(ref_total.detach()))
    ref_total.backward()
    scale = max(float(g.grad.abs().max()) for g in leaves.values() if g.grad is not None)
    checked = 0
    for name, p in m.named_parameters():
        ref = Pg[name].grad
        if ref is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
            continue
        assert p.grad is not None, name
        if name.endswith("linears.1.bias"):              # zero by softmax shift invariance: rounding noise only
            assert float(p.grad.abs().max()) < 1e-4 * scale, name
            continue
        assert O.norm_err(p.grad.cpu(), ref) < 1e-4, (name, O.norm_err(p.grad.cpu(), ref))
        checked += 1
    assert checked > 50


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_branch_streams_equal_single_stream(monkeypatch, precision):
    """The caption / audio layers of a block run on branch streams beside the video layer (decoder.py
    ``_block_memories``): outputs, loss and every gradient equal the single-stream run (same kernels, same operands;
    only atomically accumulated gradients may differ in the last bits), eagerly and over repeated steps."""
    path = [p for p in GOLD if p.endswith("model_avsd_audio.npz")][0]
    c, args, P, shapes, b, gold = load_model_case(path)
    bd = b.to(DEV)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("BIST_BRANCH_STREAMS", mode)
        m = _build(c, P, shapes, precision).eval()
        lc = SimpleLossCompute(m.generator, m.ae_generator, LabelSmoothing(c["V"], PAD, SMOOTHING), args=m.args)
        for _ in range(3):                               # repeated steps: allocator reuse across the branch streams
            for p in m.parameters():
                p.grad = None
            ft = m(bd)
            total, _ = lc.loss(ft, bd)
            total.backward()
        torch.cuda.synchronize()
        res[mode] = (ft["decoded_text"].float().cpu(), float(total), {n: p.grad.cpu() for n, p in m.named_parameters() if p.grad is not None})
    assert torch.equal(res["0"][0], res["1"][0])
    assert res["0"][1] == res["1"][1]
    assert res["0"][2].keys() == res["1"][2].keys()
    tol = 1e-5 if precision == "fp32" else 1e-3
    for n, g0 in res["0"][2].items():
        if n.endswith("linears.1.bias"):                 # zero up to rounding noise (softmax shift invariance)
            continue
        assert O.norm_err(res["1"][2][n], g0) <= tol, (n, O.norm_err(res["1"][2][n], g0))
