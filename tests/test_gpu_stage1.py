"""Kernel-level parity of the fused stage-1 kernel (csrc/stage1_fused.cu, C-ABI bist_stage1_fused) against the torch restatement
of temporal2spatial / spatial2temporal stage 1 with folded keys (tests/stage1_ref.py; /root/reference/model/encoder.py:109-121,
143-148).  The layer-level parity against the oracle and the reference goldens is in test_gpu_parity.py."""
import pytest
import torch

from tests import stage1_ref as R

pytestmark = pytest.mark.gpu

# seed, B, T, S, Lq, d, h
CASES = [
    (1, 2, 8, 16, 16, 512, 8),      # config-1 shape
    (2, 3, 40, 16, 16, 512, 8),     # config-2 shape (3 regions x 40 clips per t2s tile, ragged clips)
    (3, 2, 33, 49, 7, 512, 8),      # odd sizes: keys padded to 8 / 16, Lq < 16
    (4, 2, 128, 49, 16, 512, 8),    # config-5 shape: one region per t2s tile, two clips per s2t tile
    (5, 5, 6, 16, 5, 128, 2),
    (6, 2, 64, 16, 1, 256, 4),      # config-4 clip count, Lq = 1
    (7, 2, 17, 20, 9, 384, 6),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "B%d_T%d_S%d_Lq%d_d%d" % c[1:6])
@pytest.mark.parametrize("kinds", [(0, 1), (1,), (0,)], ids=["both", "s2t", "t2s"])
def test_stage1_fused_matches_restatement(case, kinds):
    res = R.check_case(*case, kinds=kinds, want_dbg=True)
    for k in kinds:
        # bf16 output rounding is 2^-9 relative to the row scale; the restatement rounds values and probabilities like the kernel
        assert res[k] < 5e-3, "c1 of direction %d: normalised error %.3e" % (k, res[k])
        s_err, v_err = res[("raw", k)]
        assert s_err < 1e-5 and v_err < 1e-5, "raw accumulator tiles: scores %.3e values %.3e" % (s_err, v_err)


def test_stage1_fused_full_config2_finite_and_masked():
    "Config-2 size: every output finite; a masked clip does not influence t2s (padded-clip invariance, modules.py:60)."
    B, T, S, Lq, d, h = 64, 40, 16, 16, 512, 8
    vft, tmask, dirs = R.make_inputs(21, B, T, S, Lq, d, h, "cuda:0")
    outs, _, _ = R.run_kernel(vft, tmask, dirs, [0, 1], Lq, h)
    assert all(torch.isfinite(o.float()).all() for o in outs)
    v2 = vft.clone()
    v2[tmask == 0] = 7.0          # garbage in the padded clips
    outs2, _, _ = R.run_kernel(v2, tmask, dirs, [0, 1], Lq, h)
    assert torch.equal(outs[0], outs2[0])          # t2s: masked keys have exactly zero probability
    keep = tmask.bool()[:, None, :, None].expand_as(outs[1])
    assert torch.equal(outs[1][keep], outs2[1][keep])   # s2t: summaries of valid clips unchanged


def test_stage1_fused_unsupported_shapes_are_declined():
    from bist_b200 import _lib
    L = _lib.lib()
    assert L.bist_stage1_fused_supported(4, 40, 16, 16, 512, 8) == 1
    assert L.bist_stage1_fused_supported(4, 40, 16, 20, 512, 8) == 0     # Lq > 16
    assert L.bist_stage1_fused_supported(4, 130, 16, 16, 512, 8) == 0    # T > 128
    assert L.bist_stage1_fused_supported(4, 40, 16, 16, 64, 8) == 0      # d_k != 64
