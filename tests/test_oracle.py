"""CPU tests: the oracle restatement against (a) the committed golden fixtures produced by
the unmodified reference (oracle/make_golden.py) and (b) the live reference when its tree
is present (build container only; /root/reference does not exist on the GPU box)."""
import glob
import json
import os
import sys

import numpy as np
import pytest
import torch

from bist_b200 import synth
from oracle import bist_oracle as O

REF = os.environ.get("BIST_REFERENCE", "/root/reference")
TOL = 2e-6  # fp32 re-run of the same formulation: only op-order noise


def _load(path):
    z = np.load(path)
    return json.loads(str(z["meta"])), z


def _cases(golden_dir):
    return sorted(glob.glob(os.path.join(golden_dir, "path_*.npz")))


def test_golden_fixtures_exist(golden_dir):
    assert len(_cases(golden_dir)) >= 6
    assert os.path.exists(os.path.join(golden_dir, "modules.npz"))


@pytest.mark.parametrize("name", ["both_small", "t2s_only", "s2t_only", "lq1_ragged",
                                  "exec_sh_d128", "long_video", "cfg1_d512"])
def test_oracle_matches_reference_golden(golden_dir, name):
    c, z = _load(os.path.join(golden_dir, f"path_{name}.npz"))
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c["t2s"]), bool(c["s2t"]))
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        vft, _ = O.vid_encoder8(P, "vid_encoder.", b.fts)
        st, sp = [int(v) for v in z["vft_stride"]]
        assert O.norm_err(vft[:, ::st, ::sp, :], torch.from_numpy(z["vft_sample"])) < TOL
        t2s = s2t = eq
        for blk in range(c["n_blocks"]):
            t2s, s2t = O.vid_encoder_layer4(P, f"mutlimodal_decoder.v_layers.{blk}.", c["h"], t2s, s2t,
                                            vft, b.query_mask, b.temporal_mask,
                                            t2s=bool(c["t2s"]), s2t=bool(c["s2t"]))
            if c["t2s"]:
                assert O.norm_err(t2s, torch.from_numpy(z[f"t2s_{blk}"])) < TOL
            if c["s2t"]:
                assert O.norm_err(s2t, torch.from_numpy(z[f"s2t_{blk}"])) < TOL
        full = O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask, c["n_blocks"],
                            t2s=bool(c["t2s"]), s2t=bool(c["s2t"]))
    if c["s2t"]:
        assert O.norm_err(full["temporal_ft"], torch.from_numpy(z["temporal_ft"])) < TOL
    if c["t2s"]:
        assert O.norm_err(full["spatial_ft"], torch.from_numpy(z["spatial_ft"])) < TOL


def test_oracle_generic_blocks_match_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "modules.npz"))
    m = json.loads(str(z["meta"]))
    t = lambda k: torch.from_numpy(z[k])
    x, kv = t("x"), t("kv")
    assert O.norm_err(O.layer_norm(x, t("ln.a_2"), t("ln.b_2")), t("ln_out")) < TOL
    P = {k[len("att."):]: t(k) for k in z.files if k.startswith("att.")}
    for tag, mask in (("k", t("mask_k")), ("qk", t("mask_qk")), ("nomask", None)):
        out, p = O.mha(P, "", m["h"], x, kv, kv, mask)
        assert O.norm_err(out, t(f"att_out_{tag}")) < TOL
        if tag != "nomask":
            assert O.norm_err(p, t(f"att_p_{tag}")) < TOL
    # the fully masked query row (mask_qk[1,2,:] == 0) must be uniform, as in modules.py:60-61
    _, p = O.mha(P, "", m["h"], x, kv, kv, t("mask_qk"))
    assert torch.allclose(p[1, :, 2, :], torch.full_like(p[1, :, 2, :], 1.0 / m["Lk"]))
    Pf = {k[len("ff."):]: t(k) for k in z.files if k.startswith("ff.")}
    assert O.norm_err(O.ffn(Pf, "", x), t("ff_out")) < TOL
    Ps = {k[len("sc."):]: t(k) for k in z.files if k.startswith("sc.")}
    assert O.norm_err(O.sublayer(Ps, "", x, lambda y: O.ffn(Pf, "", y)), t("sc_out")) < TOL


def test_masked_clips_do_not_influence_outputs():
    """Property from SURVEY.md §8(a) a4: zero-padded clips are excluded by temporal_mask."""
    c = dict(seed=5, B=2, T=6, S=3, D=24, d=32, h=4, Lq=4, n_blocks=1)
    P = synth.make_params(c["seed"], c["d"], c["D"], 1)
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        vft, _ = O.vid_encoder8(P, "vid_encoder.", b.fts)
        a = O.vid_encoder_layer4(P, "mutlimodal_decoder.v_layers.0.", 4, eq, eq, vft, b.query_mask, b.temporal_mask)
        vft2 = vft.clone()
        pad = ~b.temporal_mask[:, 0, :]
        vft2[pad] = 123.0  # garbage in the padded clips of the projected tensor
        bb = O.vid_encoder_layer4(P, "mutlimodal_decoder.v_layers.0.", 4, eq, eq, vft2, b.query_mask, b.temporal_mask)
    assert pad.any()
    assert O.norm_err(a[0], bb[0]) < 1e-6 and O.norm_err(a[1], bb[1]) < 1e-6


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_oracle_matches_live_reference():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle"))
    import make_golden as G
    mtn, _ = G.import_reference()
    c = dict(seed=77, B=2, T=7, S=5, D=56, d=96, h=6, Lq=4, n_blocks=2, t2s=1, s2t=1)
    ref = G.run_reference_case(mtn, c)
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        full = O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask, c["n_blocks"])
    assert O.norm_err(full["temporal_ft"], torch.from_numpy(ref["temporal_ft"])) < TOL
    assert O.norm_err(full["spatial_ft"], torch.from_numpy(ref["spatial_ft"])) < TOL
    assert O.norm_err(full["spatiotemporal_ft"], torch.from_numpy(ref["vft_sample"])) < TOL
