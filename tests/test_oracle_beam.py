"""CPU: the oracle's restatement of ``beam_search_decode`` (decode.py:53-104) against
tests/golden/beam_*.npz — the n-best token ids, scores and best_state of the UNMODIFIED reference run one dialogue turn
at a time on the same seeded batch and weights (oracle/make_golden.py)."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from bist_b200 import synth
from oracle import bist_oracle as O
from tests.test_oracle_model import load_model_case

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "beam_*.npz")))


def test_fixtures_present():
    assert len(GOLD) >= 2


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[5:-4])
def test_oracle_beam_search_matches_reference(path):
    z = np.load(path)
    g = json.loads(str(z["beam"]))
    c, args, P, shapes, b, gold = load_model_case(path.replace("beam_", "model_"))
    with torch.no_grad():
        for i in range(c["B"]):
            hyps, best = O.beam_search(P, args, synth.batch_rows(b, i, i + 1), c["h"], c["n_blocks"], g["max_len"], g["start"],
                                       g["unk"], g["end"], beam=g["beam"], penalty=g["penalty"], nbest=g["nbest"],
                                       min_len=g["min_len"])
            toks = np.full((g["nbest"], g["max_len"]), -1, dtype=np.int64)
            for j, (seq, _s) in enumerate(hyps):
                toks[j, :len(seq)] = [int(t) for t in seq]
            assert np.array_equal(toks, z[f"tokens_{i}"]), (i, toks.tolist(), z[f"tokens_{i}"].tolist())
            got = np.array([float(s) for _seq, s in hyps])
            assert np.allclose(got, z[f"scores_{i}"], rtol=1e-5, atol=1e-5), (got, z[f"scores_{i}"])
            assert abs(float(best) - float(z[f"best_{i}"])) < 1e-5 * max(1.0, abs(float(z[f"best_{i}"])))
