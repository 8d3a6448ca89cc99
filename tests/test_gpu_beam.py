"""GPU: bist_b200.decode.beam_search_decode / beam_search_decode_batch against tests/golden/beam_*.npz — the n-best
hypotheses (token ids, scores, best_state) of the UNMODIFIED reference's ``beam_search_decode`` (decode.py:53-104) run
one dialogue turn at a time on the same seeded batch and weights.  Token ids must be identical in fp32; scores within
1e-4 relative.  The batched variant (all turns and all hypotheses through the decoder at once, encoder-side blocks
cached) must give the same hypotheses as the turn-by-turn calls."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from bist_b200 import synth
from bist_b200.decode import beam_search_decode, beam_search_decode_batch
from tests.test_gpu_model import _build
from tests.test_oracle_model import load_model_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "beam_*.npz")))


def _check(hyps, best, z, i, g):
    toks = np.full((g["nbest"], g["max_len"]), -1, dtype=np.int64)
    for j, (seq, _s) in enumerate(hyps):
        toks[j, :len(seq)] = [int(t) for t in seq]
    assert np.array_equal(toks, z[f"tokens_{i}"]), (i, toks.tolist(), z[f"tokens_{i}"].tolist())
    ref = z[f"scores_{i}"]
    got = np.array([float(s) for _seq, s in hyps])
    assert np.allclose(got, ref, rtol=1e-4, atol=1e-4), (got, ref)
    assert abs(float(best) - float(z[f"best_{i}"])) < 1e-4 * max(1.0, abs(float(z[f"best_{i}"])))


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[5:-4])
def test_beam_search_matches_reference(path):
    z = np.load(path)
    g = json.loads(str(z["beam"]))
    model_path = path.replace("beam_", "model_")
    c, args, P, shapes, b, gold = load_model_case(model_path)
    m = _build(c, P, shapes, "fp32").eval()
    bd = b.to(DEV)
    kw = dict(beam=g["beam"], penalty=g["penalty"], nbest=g["nbest"], min_len=g["min_len"], train_args=m.args)
    for i in range(c["B"]):                                   # the reference's protocol: one turn per call
        hyps, best = beam_search_decode(m, synth.batch_rows(bd, i, i + 1), g["max_len"], g["start"], g["unk"], g["end"],
                                        g["pad"], **kw)
        _check(hyps, best, z, i, g)
    res = beam_search_decode_batch(m, bd, g["max_len"], g["start"], g["unk"], g["end"], g["pad"], **kw)
    assert len(res) == c["B"]
    for i, (hyps, best) in enumerate(res):                    # all turns and hypotheses in one decoder batch
        _check(hyps, best, z, i, g)
