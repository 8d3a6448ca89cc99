"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the
drop-in modules and hence the C ABI, against (a) the committed golden fixtures produced by the
unmodified reference, (b) the CPU oracle on fresh seeded shapes, (c) size-independent
properties at BASELINE config-2 size.  Tolerances are the north_star's: normalised error
max|a-b|/max|b| <= 1e-4 in fp32 mode and <= 2e-2 in bf16 mode."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from bist_b200 import VideoPath, _lib, synth
from bist_b200.modules import LayerNorm, MultiHeadedAttention, PositionwiseFeedForward, SublayerConnection
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
TOL = {"fp32": 1e-4, "bf16": 2e-2}
DEV = "cuda:0"


def _case(path):
    z = np.load(path)
    return json.loads(str(z["meta"])), z


def _run_path(c, precision, P=None):
    P = P or synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c.get("t2s", 1)), bool(c.get("s2t", 1)))
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], t2s=c.get("t2s", 1), s2t=c.get("s2t", 1),
                  precision=precision).to(DEV)
    m.load_state_dict(P, strict=True)
    m.requires_grad_(False)          # inference path (the training path is covered by test_gpu_backward.py)
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(DEV)
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"]).to(DEV)
    return m, b, eq


def test_library_loaded_and_device_is_blackwell():
    _lib.check(_lib.lib().bist_device_check(), "bist_device_check")


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["both_small", "t2s_only", "s2t_only", "lq1_ragged", "exec_sh_d128",
                                  "long_video", "cfg1_d512"])
def test_video_path_matches_reference_golden(golden_dir, name, precision):
    c, z = _case(os.path.join(golden_dir, f"path_{name}.npz"))
    m, b, eq = _run_path(c, precision)
    tol = TOL[precision]
    n0 = _lib.lib().bist_launch_count()
    ft = {"encoded_query": eq}
    ft = m.vid_encoder(b, ft)
    st, sp = [int(v) for v in z["vft_stride"]]
    assert O.norm_err(ft["spatiotemporal_ft"][:, ::st, ::sp, :].float(), torch.from_numpy(z["vft_sample"])) < tol
    in_ft = {"t2s": eq, "s2t": eq}
    for blk, layer in enumerate(m.mutlimodal_decoder.v_layers):
        in_ft = layer(in_ft, ft, b)
        if c["t2s"]:
            assert O.norm_err(in_ft["t2s"], torch.from_numpy(z[f"t2s_{blk}"])) < tol, f"t2s block {blk}"
        if c["s2t"]:
            assert O.norm_err(in_ft["s2t"], torch.from_numpy(z[f"s2t_{blk}"])) < tol, f"s2t block {blk}"
    full = m(b, eq)
    if c["s2t"]:
        assert O.norm_err(full["temporal_ft"], torch.from_numpy(z["temporal_ft"])) < tol
    if c["t2s"]:
        assert O.norm_err(full["spatial_ft"], torch.from_numpy(z["spatial_ft"])) < tol
    assert _lib.lib().bist_launch_count() > n0, "no kernel of libbist_b200 was launched"


def test_generic_blocks_match_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "modules.npz"))
    meta = json.loads(str(z["meta"]))
    t = lambda k: torch.from_numpy(z[k]).to(DEV)
    x, kv = t("x"), t("kv")

    def load(mod, pre):
        mod.load_state_dict({k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)}, strict=True)
        return mod.to(DEV).eval()

    ln = load(LayerNorm(meta["d"]), "ln.")
    assert O.norm_err(ln(x), t("ln_out")) < 1e-5
    att = load(MultiHeadedAttention(meta["h"], meta["d"]), "att.")
    for tag, mask in (("k", t("mask_k")), ("qk", t("mask_qk")), ("nomask", None)):
        out = att(x, kv, kv, mask)
        assert O.norm_err(out, t(f"att_out_{tag}")) < 1e-5, tag
        if tag != "nomask":
            assert O.norm_err(att.attn, t(f"att_p_{tag}")) < 1e-5, tag
    att(x, kv, kv, t("mask_qk"))
    p = att.attn[1, :, 2, :]        # fully masked query row -> uniform (modules.py:60-61)
    assert torch.allclose(p, torch.full_like(p, 1.0 / meta["Lk"]), atol=1e-6)
    ff = load(PositionwiseFeedForward(meta["d"], meta["dff"]), "ff.")
    assert O.norm_err(ff(x), t("ff_out")) < 1e-5
    sc = load(SublayerConnection(meta["d"], 0.1), "sc.")
    assert O.norm_err(sc(x, ff), t("sc_out")) < 1e-5
    # bf16 storage of the same blocks
    xb = x.to(torch.bfloat16)
    assert O.norm_err(ln(xb).float(), t("ln_out")) < 2e-2
    assert O.norm_err(ff(xb).float(), t("ff_out")) < 2e-2
    assert O.norm_err(att(xb, kv.to(torch.bfloat16), kv.to(torch.bfloat16), t("mask_k")).float(), t("att_out_k")) < 2e-2


@pytest.mark.parametrize("c", [
    dict(seed=101, B=3, T=9, S=5, D=72, d=96, h=6, Lq=4, n_blocks=2, t2s=1, s2t=1),
    dict(seed=102, B=2, T=17, S=16, D=128, d=256, h=8, Lq=11, n_blocks=1, t2s=1, s2t=1),
    dict(seed=103, B=5, T=3, S=2, D=36, d=48, h=4, Lq=2, n_blocks=3, t2s=0, s2t=1),
    dict(seed=104, B=1, T=1, S=1, D=16, d=32, h=2, Lq=1, n_blocks=1, t2s=1, s2t=1),     # degenerate sizes
    dict(seed=105, B=2, T=130, S=3, D=32, d=64, h=8, Lq=3, n_blocks=1, t2s=1, s2t=0),    # > 128 clips
])
def test_video_path_matches_oracle_on_fresh_shapes(c):
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c["t2s"]), bool(c["s2t"]))
    bc = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eqc = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        ref = O.video_path(P, c["h"], bc.fts, eqc, bc.query_mask, bc.temporal_mask, c["n_blocks"],
                           t2s=bool(c["t2s"]), s2t=bool(c["s2t"]))
    for precision in ("fp32", "bf16"):
        m, b, eq = _run_path(c, precision, P)
        ft = m(b, eq)
        keys = ["spatiotemporal_ft"] + (["temporal_ft", "s2t"] if c["s2t"] else []) + (["spatial_ft", "t2s"] if c["t2s"] else [])
        for k in keys:
            if True:
                assert O.norm_err(ft[k].float(), ref[k]) < TOL[precision], (precision, k)


def test_config2_full_size_properties():
    """BASELINE config 2 (B=64,T=40,S=16,D=2048,d=512, 3 blocks): oracle on two rows, row
    independence, and invariance to the content of zero-padded clips."""
    c = dict(seed=1, B=64, T=40, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=3, t2s=1, s2t=1)
    P = synth.make_params(0, c["d"], c["D"], c["n_blocks"])
    m, b, eq = _run_path(c, "fp32", P)
    full = m(b, eq)
    # (a) oracle on rows 0 and 5 (row 5 is ragged)
    rows = [0, 5]
    bc = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eqc = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        ref = O.video_path(P, c["h"], bc.fts[rows], eqc[rows], bc.query_mask[rows], bc.temporal_mask[rows], 3)
    for k in ("spatiotemporal_ft", "temporal_ft", "spatial_ft"):
        assert O.norm_err(full[k][rows], ref[k]) < 1e-4, k
    # (b) dialogue turns are independent: a sub-batch reproduces its rows
    sub = synth.SynthBatch(fts=b.fts[8:24], query_mask=b.query_mask[8:24], temporal_mask=b.temporal_mask[8:24])
    part = m(sub, eq[8:24])
    for k in ("temporal_ft", "spatial_ft"):
        assert O.norm_err(part[k], full[k][8:24]) < 1e-5, k
    # (c) padded clips are excluded by temporal_mask: garbage there changes nothing
    ft = {"spatiotemporal_ft": full["spatiotemporal_ft"].clone()}
    pad = ~b.temporal_mask[:, 0, :]
    assert pad.any()
    ft["spatiotemporal_ft"][pad] = 77.0
    in_ft = {"t2s": eq, "s2t": eq}
    for layer in m.mutlimodal_decoder.v_layers:
        in_ft = layer(in_ft, ft, b)
    assert O.norm_err(in_ft["t2s"], full["t2s"]) < 1e-5
    assert O.norm_err(in_ft["s2t"], full["s2t"]) < 1e-5
    # bf16 mode at full size against the fp32 run
    mb, _, _ = _run_path(c, "bf16", P)
    fb = mb(b, eq)
    for k in ("temporal_ft", "spatial_ft"):
        assert O.norm_err(fb[k].float(), full[k]) < 2e-2, k


def test_errors_are_loud():
    c = dict(seed=7, B=2, T=3, S=4, D=40, d=32, h=2, Lq=3, n_blocks=1, t2s=1, s2t=1)
    m, b, eq = _run_path(c, "fp32")
    ft = m.vid_encoder(b, {})
    ft["spatiotemporal_ft"] = ft["spatiotemporal_ft"].to(torch.bfloat16)
    with pytest.raises(TypeError):
        m.mutlimodal_decoder.v_layers[0]({"t2s": eq, "s2t": eq}, ft, b)
    # a non-CUDA input is an error, never a CPU fallback
    with pytest.raises(RuntimeError):
        m(synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"]), eq.cpu())
