"""Fused query-side kernels of the BiST layer (csrc/qside.cu: LN0 + QKV + self-attention, LN1/LN2 + Q projections + key
folds, LN3 + first FFN layer — /root/reference/model/encoder.py:172-201 on the query rows) against (a) the same forward with
the separate launches (bist_debug_qside(0)) and (b) the CPU oracle, at shapes that exercise partial tiles (B % 8 != 0),
short queries (Lq < 16), masked query tokens, one direction only, and every supported head count."""
import pytest
import torch

from bist_b200 import VideoPath, _lib, synth
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

SHAPES = [  # B, T, S, D, d, h, Lq, n_blocks
    (3, 6, 16, 256, 128, 2, 7, 2),
    (8, 5, 9, 128, 256, 4, 16, 1),
    (13, 12, 16, 256, 384, 6, 11, 1),
    (17, 10, 16, 512, 512, 8, 16, 2),
    (64, 40, 16, 512, 512, 8, 16, 1),
]


def _run(model, b, eq, on):
    _lib.lib().bist_debug_qside(on)
    try:
        with torch.no_grad():
            ft = model(b, eq)
        torch.cuda.synchronize()
        return {k: ft[k].float().clone() for k in ("temporal_ft", "spatial_ft")}
    finally:
        _lib.lib().bist_debug_qside(-1)


@pytest.mark.parametrize("shape", SHAPES)
def test_fused_query_side_matches_separate_launches_and_oracle(shape):
    B, T, S, D, d, h, Lq, nb = shape
    P = synth.make_params(11, d, D, nb)
    hb = synth.make_batch(11, B, T, S, D, Lq)
    heq = synth.make_encoded_query(11, B, Lq, d)
    if Lq > 3:                                   # some padded query tokens (query_mask == 0 on the keys of the self-attention)
        hb.query_mask[1 % B, Lq - 2:] = 0
        hb.query_mask[B - 1, Lq - 1:] = 0
    model = VideoPath(d, h, nb, D, precision="bf16").to(DEV)
    model.load_state_dict(P, strict=True)
    b, eq = hb.to(DEV), heq.to(DEV)
    L = _lib.lib()
    _run(model, b, eq, -1)                       # the first call packs the weights
    n0 = L.bist_launch_count()
    fused = _run(model, b, eq, 1)
    n1 = L.bist_launch_count()
    plain = _run(model, b, eq, 0)
    n2 = L.bist_launch_count()
    assert n1 - n0 < n2 - n1, "the fused path must launch fewer kernels (%d vs %d)" % (n1 - n0, n2 - n1)
    for k in fused:
        assert torch.isfinite(fused[k]).all()
        err = O.norm_err(fused[k].cpu(), plain[k].cpu())
        assert err < 1e-2, "%s: fused vs separate launches %.3e" % (k, err)
    if B <= 17:
        with torch.no_grad():
            ref = O.video_path(P, h, hb.fts, heq, hb.query_mask, hb.temporal_mask, nb)
        for k in fused:
            err = O.norm_err(fused[k].cpu(), ref[k])
            assert err < 2e-2, "%s: fused vs oracle %.3e" % (k, err)
