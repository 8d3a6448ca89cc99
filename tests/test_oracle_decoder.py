"""CPU: the oracle's restatement of the decoder side (SURVEY.md §8(f) row 1 — CapEncoderLayer / AudioEncoderLayer,
MultimodalDecoderLayer12, MultimodalDecoder8 with its softmax gate) against tests/golden/decoder_*.npz, which hold
the outputs of the UNMODIFIED reference on the same seeded inputs and weights (oracle/make_golden.py)."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from bist_b200 import synth
from oracle import bist_oracle as O

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "decoder_*.npz")))


def load_case(path):
    z = np.load(path)
    c = json.loads(str(z["meta"]))
    shapes = {k: tuple(v) for k, v in json.loads(str(z["param_shapes"]))}
    P = synth.fill_params(shapes, c["seed"])
    b, ft, x = synth.make_decoder_inputs(c["seed"], c["B"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["Ta"], c["T"], c["S"], c["d"])
    cfg = dict(nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"], nb_aenc_blocks=c["nb_aenc"],
               enc_vc_combine=c["vc"], include_caption=c["cap"], dec_st_combine=c["dec_st"], t2s=c["t2s"], s2t=c["s2t"],
               enc_st_combine="none")
    masks = dict(trg=b.trg_mask, his=b.his_mask, query=b.query_mask, cap=b.cap_mask, audio=b.audio_mask,
                 temporal=b.temporal_mask)
    gold = {k: torch.from_numpy(z[k]) for k in z.files if k not in ("meta", "param_shapes")}
    return c, cfg, P, b, ft, x, masks, gold


def test_fixtures_present():
    assert len(GOLD) >= 7


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[8:-4])
def test_oracle_decoder_matches_reference_outputs(path):
    c, cfg, P, b, ft, x, masks, gold = load_case(path)
    out = O.multimodal_decoder8(P, "", c["h"], cfg, ft, x, masks, c["n_blocks"])
    assert "decoded_text" in gold
    for k, g in gold.items():
        assert O.norm_err(out[k], g) < 2e-6, (k, O.norm_err(out[k], g))
