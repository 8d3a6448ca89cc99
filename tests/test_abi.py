"""CPU tests of the boundary: the C-ABI library builds, loads and exports every symbol that
include/bist_b200.h declares; the Python mirrors keep the reference's parameter names and fail
loudly without a GPU (no CPU fallback); host-side helpers."""
import ctypes
import os
import re

import pytest
import torch

import __graft_entry__ as entry
from bist_b200 import VideoPath, _lib, synth
from bist_b200.modules import LayerNorm, MultiHeadedAttention, PositionwiseFeedForward, SublayerConnection

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libpath():
    return entry.build()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "bist_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bist_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(libpath):
    h = ctypes.CDLL(libpath)
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(h, n), "libbist_b200.so does not export %s" % n
    assert sorted(_lib.SIGNATURES) == names, "bist_b200/_lib.py SIGNATURES and the header disagree"


def test_version_and_size_queries_need_no_gpu(libpath):
    L = _lib.lib()
    assert L.bist_version() == 100
    assert L.bist_st_layer_packed_bytes(512, 2048, 2, _lib.F32) > 4 * 2 * (11 * 512 * 512 + 2 * 512 * 2048)
    assert L.bist_st_layer_packed_bytes(512, 2048, 2, _lib.BF16) < L.bist_st_layer_packed_bytes(512, 2048, 2, _lib.F32)
    assert L.bist_st_layer_packed_bytes(0, 2048, 2, _lib.F32) == 0
    w2 = L.bist_st_layer_workspace_bytes(64, 40, 16, 16, 512, 2048, 8, 2, _lib.F32)
    w1 = L.bist_st_layer_workspace_bytes(64, 40, 16, 16, 512, 2048, 8, 1, _lib.F32)
    assert 0 < w1 < w2 < 2 ** 31
    assert L.bist_vidproj_workspace_bytes(40960, 2048, 512, _lib.F32) >= 40960 * 512 * 4


def test_argument_errors_are_reported_not_thrown(libpath):
    L = _lib.lib()
    rc = L.bist_ln_fwd(None, None, None, None, 4, 8, 1e-6, 0, 0, None)
    assert rc == 1 and b"null" in L.bist_last_error()
    rc = L.bist_linear_fwd(None, None, None, None, None, 1, 1, 1, 0, 0, 0, 0, None)
    assert rc == 1
    with pytest.raises(RuntimeError, match="status 1"):
        _lib.check(rc, "bist_linear_fwd")


def test_modules_keep_reference_parameter_names():
    m = VideoPath(64, 4, 2, 80)
    keys = set(m.state_dict().keys())
    P = synth.make_params(1, 64, 80, 2)
    assert keys == set(P.keys())
    assert set(MultiHeadedAttention(4, 64).state_dict()) == {f"linears.{i}.{n}" for i in range(4) for n in ("weight", "bias")}
    assert set(PositionwiseFeedForward(64, 128).state_dict()) == {"w_1.weight", "w_1.bias", "w_2.weight", "w_2.bias"}
    assert set(SublayerConnection(64, 0.1).state_dict()) == {"norm.a_2", "norm.b_2"}
    assert set(LayerNorm(64).state_dict()) == {"a_2", "b_2"}
    one = VideoPath(64, 4, 2, 80, t2s=0, s2t=1)
    assert set(one.state_dict().keys()) == set(synth.make_params(1, 64, 80, 2, False, True).keys())


def test_no_cpu_fallback():
    m = VideoPath(64, 4, 1, 80)
    b = synth.make_batch(1, 2, 3, 4, 80, 5)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(b, synth.make_encoded_query(1, 2, 5, 64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        LayerNorm(8)(torch.zeros(2, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):       # train() mode runs the CUDA path too, never torch's
        SublayerConnection(8, 0.1).train()(torch.zeros(2, 8), lambda x: x)


def test_shard_rows_covers_batch_exactly():
    for n in (1, 7, 64, 65):
        for w in (1, 2, 3, 8):
            spans = [synth.shard_rows(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
