"""CPU: the oracle's restatement of the whole forward + loss (SURVEY.md §8(f) rows 1-2 — embeddings, positional
encoding, text-encoder norms, video encoder, multimodal decoder, pointer generators, label-smoothed KL and the
auto-encoder heads) against tests/golden/model_*.npz: outputs of the UNMODIFIED reference (``MTN.forward``,
``generator`` and ``SimpleLossCompute``) on the same seeded batch and weights (oracle/make_golden.py)."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from bist_b200 import synth
from oracle import bist_oracle as O

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "model_*.npz")))
PAD, SMOOTHING = 1, 0.1
TIED = ("tgt_embed.0.lut.weight", "generator.vocab_gen", "ae_generator.proj")


def load_model_case(path):
    z = np.load(path)
    c = json.loads(str(z["meta"]))
    shapes = {k: tuple(v) for k, v in json.loads(str(z["param_shapes"]))}
    P = synth.fill_params({k: v for k, v in shapes.items() if not k.endswith(".pe")}, c["seed"])
    for k in TIED:
        if k in P:
            P[k] = P["query_embed.0.lut.weight"]
    b = synth.make_full_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                              Ta=c.get("Ta", 0), Da=c.get("Da", 0))
    args = dict(nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"], nb_aenc_blocks=c.get("nb_aenc", 0), enc_vc_combine=c["vc"],
                include_caption=c["cap"], dec_st_combine="seq", t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none",
                ptr_ft=c["ptr_ft"], auto_encoder=c["ae"], mask_unk=1)
    gold = {k: torch.from_numpy(z[k]) for k in z.files if k not in ("meta", "param_shapes")}
    return c, args, P, shapes, b, gold


def test_fixtures_present():
    assert len(GOLD) >= 5
    assert sum("audio" in os.path.basename(p) for p in GOLD) >= 2        # config 3's audio encoder is pinned too


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-4])
def test_oracle_model_matches_reference(path):
    c, args, P, shapes, b, gold = load_model_case(path)
    ft = O.mtn_forward(P, args, b, c["h"], c["n_blocks"])
    for k in ("encoded_query", "encoded_his", "encoded_cap", "encoded_tgt", "decoded_text", "encoded_audio", "audio_ft"):
        if k in gold:
            assert O.norm_err(ft[k], gold[k]) < 2e-6, (k, O.norm_err(ft[k], gold[k]))
    logp = O.generator_logp(P, "generator.", ft, b, c["ptr_ft"], 1)
    assert O.norm_err(logp, gold["logp"]) < 5e-6
    total, parts = O.loss_compute(P, args, ft, b, PAD, SMOOTHING)
    assert abs(float(total) - float(gold["loss_total"])) < 2e-5 * abs(float(gold["loss_total"]))
    assert abs(float(parts["out"] * b.ntokens) - float(gold["loss_out_sum"])) < 2e-5 * abs(float(gold["loss_out_sum"]))


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-4])
def test_mirror_state_dict_matches_reference_keys(path):
    """Host logic, no GPU: bist_b200.make_model builds the reference's state_dict keys and shapes for every golden
    configuration (with ft_sizes=[D, Da] and the audio encoder where the case has one), so a reference checkpoint
    loads strict=True; the synthetic audio stream follows data/dataset.py:83-87."""
    import argparse
    from bist_b200 import make_model
    c, args, P, shapes, b, gold = load_model_case(path)
    ns = argparse.Namespace(nb_blocks=c["n_blocks"], d_model=c["d"], att_h=c["h"], dropout=0.1, ptr_gen=1, **args)
    m = make_model(c["V"], c["V"], ns, ft_sizes=[c["D"]] + ([c["Da"]] if c.get("nb_aenc", 0) else []))
    sd = m.state_dict()
    assert sorted(sd) == sorted(shapes)
    assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in shapes)
    if c.get("nb_aenc", 0):
        assert "vid_encoder.a_W.weight" in sd and sd["vid_encoder.a_W.weight"].shape == (c["d"], c["Da"])
        assert b.audio_fts.shape == (c["B"], c["Ta"], c["Da"])
        assert torch.equal(b.audio_mask, (b.audio_fts.sum(-1) != 0).unsqueeze(-2))
        assert bool(b.audio_mask[0].all()) and not bool(b.audio_mask.all())          # ragged, first sample full
    else:
        assert not any(k.startswith("vid_encoder.a_") for k in sd)
