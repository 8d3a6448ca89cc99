"""GPU parity of the training path (run with -m gpu on the B200 box): forward with a tape + backward through the
C ABI (bist_vidproj_train_fwd/bwd, bist_st_layer_train_fwd/bwd, bist_ln_bwd) against torch.autograd over the CPU
oracle (the reference formulation) on the same seeded inputs and weights.

Metric: normalised error max|a-b| / max|b| per gradient tensor.
  fp32 mode: <= 1e-4 on every parameter gradient and on d(encoded_query) (the north_star's fp32 bound).
  bf16 mode (bf16 activations AND bf16 gradient operands, fp32 accumulation): relative L2 error
  ||a-b|| / ||b|| <= 8e-2 and max-normalised error <= 1e-1 on every gradient that does not sit directly behind a
  ReLU gate, relative L2 <= 2e-1 on those that do (the bound is reached only by the single-row degenerate case).  A ReLU unit whose pre-activation is within bf16 rounding of zero opens in one
  precision and closes in the other, which changes single rows of dW_1 / db_1 (and of the LayerNorm in front of
  them) by O(1) while leaving the tensor as a whole within a few percent — a property of bf16 training, not of
  the kernels (the fp32 mode runs the same kernels and meets 1e-4).
Key-projection biases have a mathematically zero gradient (softmax shift invariance): the reference's autograd
leaves rounding noise there, the folded stage-1/stage-2 attentions give exactly 0.
"""
import pytest
import torch

from bist_b200 import VideoPath, _lib, synth
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_FWD = {"fp32": 1e-4, "bf16": 2e-2}
TOL_BWD = {"fp32": 1e-4, "bf16": 1e-1}
TOL_BWD_L2 = 8e-2
TOL_BWD_L2_RELU = 2e-1
RELU_ADJACENT = ("w_1.", "sublayer.3.", "sublayer.7.", "vid_encoder.")


def _loss_weights(seed, shape):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g)


def _oracle_grads(c, P, b, eq):
    Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    eqg = eq.clone().requires_grad_(True)
    out = O.video_path(Pg, c["h"], b.fts, eqg, b.query_mask, b.temporal_mask, c["n_blocks"],
                       t2s=bool(c["t2s"]), s2t=bool(c["s2t"]))
    loss = 0.0
    if c["s2t"]:
        loss = loss + (out["temporal_ft"] * _loss_weights(11, out["temporal_ft"].shape)).sum()
    if c["t2s"]:
        loss = loss + (out["spatial_ft"] * _loss_weights(12, out["spatial_ft"].shape)).sum()
    loss.backward()
    return out, {k: v.grad for k, v in Pg.items()}, eqg.grad


def _ours(c, P, b, eq, precision):
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], t2s=c["t2s"], s2t=c["s2t"], dropout=0.0,
                  precision=precision).to(DEV)
    m.load_state_dict(P, strict=True)
    m.eval()                      # dropout identity; gradients flow through the training formulation
    eqd = eq.to(DEV).requires_grad_(True)
    out = m(b.to(DEV), eqd)
    loss = 0.0
    if c["s2t"]:
        loss = loss + (out["temporal_ft"].float() * _loss_weights(11, out["temporal_ft"].shape).to(DEV)).sum()
    if c["t2s"]:
        loss = loss + (out["spatial_ft"].float() * _loss_weights(12, out["spatial_ft"].shape).to(DEV)).sum()
    loss.backward()
    torch.cuda.synchronize()
    return m, out, eqd.grad


def _check(c, precision):
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c["t2s"]), bool(c["s2t"]))
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    ref_out, ref_g, ref_geq = _oracle_grads(c, P, b, eq)
    n0 = _lib.lib().bist_launch_count()
    m, out, geq = _ours(c, P, b, eq, precision)
    assert _lib.lib().bist_launch_count() > n0
    for k in (["temporal_ft"] if c["s2t"] else []) + (["spatial_ft"] if c["t2s"] else []) + ["spatiotemporal_ft"]:
        assert O.norm_err(out[k].float(), ref_out[k]) < TOL_FWD[precision], ("forward", precision, k)
    tol = TOL_BWD[precision]
    worst = ("", 0.0)
    scale = max(float(g.abs().max()) for g in ref_g.values() if g is not None)
    for name, p in m.named_parameters():
        ref = ref_g[name]
        if ref is None:                                               # parameter not on the path (a direction is off)
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
            continue
        assert p.grad is not None, name
        if float(ref.abs().max()) < 1e-6 * scale:                     # mathematically zero (key biases)
            assert float(p.grad.abs().max()) <= 0.1 * tol * scale, (name, float(p.grad.abs().max()))
            continue
        e = O.norm_err(p.grad.float(), ref)
        if precision == "bf16":
            l2 = float((p.grad.float().cpu() - ref).norm() / ref.norm())
            gated = any(t in name for t in RELU_ADJACENT)
            assert l2 < (TOL_BWD_L2_RELU if gated else TOL_BWD_L2), (precision, name, "relative L2", l2)
            if gated:
                continue
        worst = max(worst, (name, e), key=lambda t: t[1])
        assert e < tol, (precision, name, e)
    print("worst gradient error %s: %.3e (%s)" % (precision, worst[1], worst[0]))
    e = O.norm_err(geq, ref_geq)
    assert e < tol, (precision, "encoded_query", e)
    return worst


CASES = [
    dict(seed=201, B=2, T=4, S=3, D=40, d=32, h=4, Lq=3, n_blocks=1, t2s=1, s2t=1),
    dict(seed=202, B=3, T=9, S=5, D=72, d=96, h=6, Lq=4, n_blocks=2, t2s=1, s2t=1),
    dict(seed=203, B=2, T=6, S=16, D=128, d=128, h=8, Lq=7, n_blocks=2, t2s=1, s2t=1),
    dict(seed=204, B=4, T=5, S=2, D=36, d=64, h=4, Lq=2, n_blocks=2, t2s=0, s2t=1),
    dict(seed=205, B=2, T=7, S=4, D=64, d=64, h=8, Lq=5, n_blocks=1, t2s=1, s2t=0),
    dict(seed=206, B=1, T=1, S=1, D=16, d=32, h=2, Lq=1, n_blocks=1, t2s=1, s2t=1),
]


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("c", CASES, ids=lambda c: "seed%d" % c["seed"])
def test_training_path_gradients_match_oracle_autograd(c, precision):
    _check(c, precision)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_training_path_at_model_width(precision):
    """d_model = 512, h = 8, d_ff = 2048, D = 2048 (BASELINE widths; tensor-core GEMM / mma wgrad shapes), small batch."""
    c = dict(seed=207, B=4, T=10, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=2, t2s=1, s2t=1)
    _check(c, precision)


def test_layer_norm_module_backward():
    from bist_b200.modules import LayerNorm
    torch.manual_seed(0)
    x = torch.randn(37, 96)
    a, b2 = 1 + 0.1 * torch.randn(96), 0.1 * torch.randn(96)
    w = torch.randn(37, 96)
    xr, ar, br = x.clone().requires_grad_(True), a.clone().requires_grad_(True), b2.clone().requires_grad_(True)
    (O.layer_norm(xr, ar, br) * w).sum().backward()
    ln = LayerNorm(96).to(DEV)
    ln.load_state_dict({"a_2": a, "b_2": b2})
    xd = x.to(DEV).requires_grad_(True)
    (ln(xd) * w.to(DEV)).sum().backward()
    assert O.norm_err(xd.grad, xr.grad) < 1e-5
    assert O.norm_err(ln.a_2.grad, ar.grad) < 1e-5
    assert O.norm_err(ln.b_2.grad, br.grad) < 1e-5


def test_optimizer_step_refreshes_packed_weights():
    """The packed (inference and training) weight buffers are keyed on the parameters' version counters: after an
    optimizer step both paths must see the new weights."""
    c = dict(seed=208, B=2, T=4, S=3, D=40, d=32, h=4, Lq=3, n_blocks=1, t2s=1, s2t=1)
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(DEV)
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"]).to(DEV)
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.0, precision="fp32").to(DEV)
    m.load_state_dict(P, strict=True)
    opt = torch.optim.Adam(m.parameters(), lr=1e-2, fused=True)
    with torch.no_grad():
        before = m(b, eq)["temporal_ft"].clone()
    for _ in range(2):
        opt.zero_grad(set_to_none=True)
        out = m(b, eq)
        (out["temporal_ft"].square().mean() + out["spatial_ft"].square().mean()).backward()
        opt.step()
    fresh = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.0, precision="fp32").to(DEV)
    fresh.load_state_dict(m.state_dict(), strict=True)
    with torch.no_grad():
        want = fresh(b, eq)
        got_inf = m(b, eq)
    got_train = m(b, eq)
    assert O.norm_err(before, want["temporal_ft"]) > 1e-3            # the step did move the weights
    for k in ("temporal_ft", "spatial_ft"):
        assert O.norm_err(got_inf[k], want[k]) < 1e-6, k
        assert O.norm_err(got_train[k], want[k]) < 1e-5, k


# ----------------------------------------------------------------------------------------------- dropout (train mode)
def _site_map(layer_pre, u, B, Lq, h, d, d_ff):
    """reference dropout site (oracle name) -> (library site id, flat offset inside that site) for direction slot u."""
    Q = B * Lq
    return {
        f"{layer_pre}attn.{3 * u}.p": (0, u * B * h * Lq * Lq),
        f"{layer_pre}sublayer.{4 * u}.out": (1, u * Q * d),
        f"{layer_pre}attn.{3 * u + 1}.p": (2 + u, 0),
        f"{layer_pre}sublayer.{4 * u + 1}.out": (4 + u, 0),
        f"{layer_pre}attn.{3 * u + 2}.p": (6 + u, 0),
        f"{layer_pre}sublayer.{4 * u + 2}.out": (8, u * Q * d),
        f"{layer_pre}ff.{u}.hidden": (9, u * Q * d_ff),
        f"{layer_pre}sublayer.{4 * u + 3}.out": (10, u * Q * d),
    }


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("c", [
    dict(seed=211, B=3, T=5, S=4, D=48, d=64, h=4, Lq=5, n_blocks=2, t2s=1, s2t=1),
    dict(seed=212, B=2, T=6, S=16, D=128, d=128, h=8, Lq=7, n_blocks=1, t2s=1, s2t=1),
], ids=lambda c: "seed%d" % c["seed"])
def test_train_mode_dropout_matches_oracle_with_the_same_masks(c, precision):
    """model.train() with dropout 0.2 / attention dropout 0.1 (exec.sh:35, mtn.py:77): the kernels' counter-based masks
    are read back through bist_dropout_mask and injected into the oracle, so forward AND gradients must agree."""
    import ctypes
    p_sub, p_attn = 0.2, 0.1
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=p_sub, precision=precision).to(DEV)
    m.load_state_dict(P, strict=True)
    m.train()
    torch.manual_seed(1234)
    eqd = eq.to(DEV).requires_grad_(True)
    out = m(b.to(DEV), eqd)
    (out["temporal_ft"].float() * _loss_weights(11, out["temporal_ft"].shape).to(DEV)).sum().add(
        (out["spatial_ft"].float() * _loss_weights(12, out["spatial_ft"].shape).to(DEV)).sum()).backward()
    torch.cuda.synchronize()

    d_ff = 4 * c["d"]
    table, cache = {}, {}
    for blk, layer in enumerate(m.mutlimodal_decoder.v_layers):
        cfg = layer.last_dropout
        assert cfg is not None and abs(cfg.p_attn - p_attn) < 1e-7 and abs(cfg.p_sublayer - p_sub) < 1e-7
        for u in range(2):
            for name, (site, off) in _site_map(f"mutlimodal_decoder.v_layers.{blk}.", u, c["B"], c["Lq"], c["h"], c["d"], d_ff).items():
                p = p_attn if name.endswith(".p") else p_sub
                table[name] = (int(cfg.seed), site, off, p)

    def drop(name, x):
        seed, site, off, p = table[name]
        n = off + x.numel()
        key = (seed, site, n)
        if key not in cache:
            buf = torch.empty(n, dtype=torch.float32, device=DEV)
            _lib.check(_lib.lib().bist_dropout_mask(seed, site, p, n, _lib.ptr(buf), _lib.stream_ptr()), "bist_dropout_mask")
            cache[key] = buf.cpu()
        mul = cache[key][off:off + x.numel()].view(x.shape)
        frac = float((mul == 0).float().mean())
        assert abs(frac - p) < 0.08, (name, frac)             # the masks drop about p of the elements
        return x * mul

    Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    eqg = eq.clone().requires_grad_(True)
    ref = O.video_path(Pg, c["h"], b.fts, eqg, b.query_mask, b.temporal_mask, c["n_blocks"], drop=drop)
    ((ref["temporal_ft"] * _loss_weights(11, ref["temporal_ft"].shape)).sum()
     + (ref["spatial_ft"] * _loss_weights(12, ref["spatial_ft"].shape)).sum()).backward()
    for k in ("temporal_ft", "spatial_ft"):
        assert O.norm_err(out[k].float(), ref[k]) < TOL_FWD[precision], (precision, k)
    with torch.no_grad():                                    # and dropout really changed the result
        assert O.norm_err(ref["temporal_ft"], O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask,
                                                           c["n_blocks"])["temporal_ft"]) > 1e-2
    if precision == "fp32":
        for name, prm in m.named_parameters():
            r = Pg[name].grad
            scale = max(float(g.abs().max()) for g in (v.grad for v in Pg.values()) if g is not None)
            if float(r.abs().max()) < 1e-6 * scale:
                assert float(prm.grad.abs().max()) <= 1e-5 * scale, name
                continue
            assert O.norm_err(prm.grad, r) < 1e-4, (name, O.norm_err(prm.grad, r))
        assert O.norm_err(eqd.grad, eqg.grad) < 1e-4
    else:
        scale = max(float(v.grad.abs().max()) for v in Pg.values() if v.grad is not None)
        for name, prm in m.named_parameters():
            r = Pg[name].grad
            if float(r.abs().max()) < 1e-6 * scale:          # key biases: mathematically zero
                continue
            l2 = float((prm.grad.float().cpu() - r).norm() / r.norm())
            assert l2 < (TOL_BWD_L2_RELU if any(t in name for t in RELU_ADJACENT) else TOL_BWD_L2), (name, l2)


def test_train_mode_dropout_is_seeded_by_torch():
    c = dict(seed=213, B=2, T=4, S=3, D=40, d=32, h=4, Lq=3, n_blocks=1)
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.3, precision="fp32").to(DEV)
    m.load_state_dict(synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"]), strict=True)
    m.train().requires_grad_(False)
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(DEV)
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"]).to(DEV)
    torch.manual_seed(7); a1 = m(b, eq)["temporal_ft"].clone()
    torch.manual_seed(7); a2 = m(b, eq)["temporal_ft"].clone()
    a3 = m(b, eq)["temporal_ft"].clone()
    assert torch.equal(a1, a2) and not torch.equal(a1, a3)
    assert torch.equal(m.eval()(b, eq)["temporal_ft"], m(b, eq)["temporal_ft"])      # eval mode: no dropout
