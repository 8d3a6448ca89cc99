"""Guard-band tests (compute-sanitizer is closed on this pool, profiles/r02_sanitizer_unavailable.log): every workspace, tape and
output buffer of the video path's forward AND backward is carved out of a larger allocation whose borders hold a byte pattern;
after the run the borders must be intact — an out-of-bounds write of any kernel (TMA stores included) lands in them.  Covers the
fused stage-1 kernel, the fused query-side kernels, the projection kernels and the hand-written backward at shapes with partial
tiles."""
import pytest
import torch

import bist_b200.encoder as E
import bist_b200.modules as M
from bist_b200 import VideoPath, _lib, synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 4096
PATTERN = 0xA5


class _Guarded:
    def __init__(self):
        self.bufs = []

    def ws(self, nbytes, device):
        n = (int(nbytes) + 255) // 256 * 256
        raw = torch.full((n + 2 * GUARD,), PATTERN, dtype=torch.uint8, device=device)
        self.bufs.append((raw, n))
        return raw[GUARD:GUARD + int(nbytes)]

    def check(self):
        assert self.bufs
        for raw, n in self.bufs:
            assert bool((raw[:GUARD] == PATTERN).all()), "write below a workspace"
            assert bool((raw[GUARD + n:] == PATTERN).all()), "write above a workspace"


@pytest.mark.parametrize("shape", [(3, 6, 16, 256, 128, 2, 7, 2), (13, 12, 9, 256, 384, 6, 11, 1), (9, 40, 16, 512, 512, 8, 16, 1)])
@pytest.mark.parametrize("train", [False, True])
def test_workspaces_and_outputs_keep_their_borders(monkeypatch, shape, train):
    B, T, S, D, d, h, Lq, nb = shape
    g = _Guarded()
    monkeypatch.setattr(E, "_ws", g.ws)
    monkeypatch.setattr(M, "_ws", g.ws)
    model = VideoPath(d, h, nb, D, dropout=0.1 if train else 0.0, precision="bf16").to(DEV)
    model.load_state_dict(synth.make_params(5, d, D, nb), strict=True)
    model.train(train)
    b = synth.make_batch(5, B, T, S, D, Lq).to(DEV)
    eq = synth.make_encoded_query(5, B, Lq, d).to(DEV).requires_grad_(train)
    if train:
        o = model(b, eq)
        (o["temporal_ft"].float().sum() + o["spatial_ft"].float().square().sum()).backward()
    else:
        with torch.no_grad():
            o = model(b, eq)
    torch.cuda.synchronize()
    assert torch.isfinite(o["temporal_ft"].float()).all() and torch.isfinite(o["spatial_ft"].float()).all()
    g.check()


def test_stage1_and_projection_outputs_keep_their_borders():
    import ctypes
    from tests import stage1_ref as R
    L = _lib.lib()
    B, T, S, Lq, d, h = 5, 37, 9, 11, 256, 4
    if not L.bist_stage1_fused_supported(B, T, S, Lq, d, h):
        pytest.skip("fused stage 1 switched off (BIST_STAGE1_UNFUSED)")
    vft, tmask, dirs = R.make_inputs(3, B, T, S, Lq, d, h, DEV)
    raws, outs = [], []
    for k in (0, 1):
        G = S if k == 0 else T
        n = B * Lq * G * d
        raw = torch.full((n + 2 * GUARD,), float("nan"), dtype=torch.bfloat16, device=DEV)
        raw[:GUARD] = 1.5
        raw[GUARD + n:] = 1.5
        raws.append((raw, n))
        outs.append(raw[GUARD:GUARD + n].view(B, Lq, G, d))
    arr_i = (ctypes.c_int * 2)(0, 1)
    mk = lambda ts: (ctypes.c_void_p * 2)(*[t.data_ptr() for t in ts])
    rc = L.bist_stage1_fused(_lib.ptr(vft), _lib.ptr(tmask), 2, arr_i, mk([x[0] for x in dirs]), mk([x[1] for x in dirs]),
                             mk([x[2] for x in dirs]), mk(outs), B, T, S, Lq, d, h, None, _lib.stream_ptr())
    _lib.check(rc, "bist_stage1_fused")
    torch.cuda.synchronize()
    for (raw, n), o in zip(raws, outs):
        assert bool((raw[:GUARD] == 1.5).all()) and bool((raw[GUARD + n:] == 1.5).all()), "stage1_fused wrote outside c1"
        assert torch.isfinite(o.float()).all()
    # projection (fused kernel, last tile partial) and its no-allocation variant
    N, D2, d2 = 1000, 512, 256
    fts = torch.rand(N, D2, device=DEV)
    W = torch.randn(d2, D2, device=DEV) / D2 ** 0.5
    bias, a2, b2 = torch.zeros(d2, device=DEV), torch.ones(d2, device=DEV), torch.zeros(d2, device=DEV)
    for noalloc in (False, True):
        raw = torch.full((N * d2 + 2 * GUARD,), 1.5, dtype=torch.bfloat16, device=DEV)
        out = raw[GUARD:GUARD + N * d2]
        if noalloc:
            rc = L.bist_vidproj_fwd_noalloc(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D2, d2,
                                            1e-6, _lib.BF16, _lib.stream_ptr())
        else:
            ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D2, d2, _lib.BF16), dtype=torch.uint8, device=DEV)
            rc = L.bist_vidproj_fwd(_lib.ptr(fts), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(a2), _lib.ptr(b2), _lib.ptr(out), N, D2, d2, 1e-6,
                                    _lib.ptr(ws), ws.numel(), _lib.BF16, _lib.BF16, _lib.stream_ptr())
        _lib.check(rc, "vidproj")
        torch.cuda.synchronize()
        assert bool((raw[:GUARD] == 1.5).all()) and bool((raw[GUARD + N * d2:] == 1.5).all()), "projection wrote outside its output"
        assert torch.isfinite(out.float()).all()
