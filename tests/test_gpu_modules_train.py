"""GPU parity of the stand-alone modules in training (run with -m gpu on the B200 box): MultiHeadedAttention,
PositionwiseFeedForward and SublayerConnection of bist_b200/modules.py (the blocks the reference's decoder, caption
and audio layers are built from, decoder.py:11-60 / encoder.py:203-235) against torch.autograd over the CPU oracle.

Forward, input gradients and every parameter gradient are compared with the normalised error max|a-b| / max|b|:
fp32 <= 1e-4; bf16 forward <= 2e-2, bf16 gradients relative L2 <= 8e-2 (2e-1 behind a ReLU gate, see
tests/test_gpu_backward.py).  In train() mode the kernels' counter-based dropout masks are read back with
bist_dropout_mask from module.last_dropout = (p, seed, site) and injected into the oracle.
"""
import pytest
import torch

from bist_b200 import _lib
from bist_b200.modules import (LayerNorm, MultiHeadedAttention, PositionwiseFeedForward, SublayerConnection)
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
DT = {"fp32": torch.float32, "bf16": torch.bfloat16}


def _rand(seed, *shape, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def _randomise(mod, seed):
    """Weights ~ N(0, 1.5/sqrt(fan_in)), LayerNorm gains around 1, every other vector around 0."""
    g = torch.Generator().manual_seed(seed)
    gains = {id(m.a_2) for m in mod.modules() if isinstance(m, LayerNorm)}
    with torch.no_grad():
        for p in mod.parameters():
            if p.dim() > 1:
                p.copy_(torch.randn(p.shape, generator=g) * (1.5 / p.shape[1] ** 0.5))
            elif id(p) in gains:
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
    return mod


def _mask_for(cfg, seed):
    """Multipliers of one dropout site as a CPU tensor (what the kernels applied)."""
    p, s, site = cfg

    def get(n):
        buf = torch.empty(n, dtype=torch.float32, device=DEV)
        _lib.check(_lib.lib().bist_dropout_mask(s, site, p, n, _lib.ptr(buf), _lib.stream_ptr()), "bist_dropout_mask")
        return buf.cpu()
    return get


def _cmp(name, ours, ref, precision, relu=False, fwd=False):
    ours, ref = ours.detach().float().cpu(), ref.detach().float()
    if precision == "fp32":
        assert O.norm_err(ours, ref) < 1e-4, (name, O.norm_err(ours, ref))
    elif fwd:
        assert O.norm_err(ours, ref) < 2e-2, (name, O.norm_err(ours, ref))
    else:
        l2 = float((ours - ref).norm() / ref.norm().clamp_min(1e-30))
        assert l2 < (2e-1 if relu else 8e-2), (name, l2)


MHA_CASES = [
    dict(id="self_keymask", Bp=3, Lq=7, Lk=7, d=64, d_in=-1, h=4, mask="key", self_attn=True),
    dict(id="cross_fullmask", Bp=2, Lq=5, Lk=9, d=64, d_in=-1, h=8, mask="full", self_attn=False),
    dict(id="cross_nomask_din", Bp=2, Lq=4, Lk=6, d=64, d_in=48, h=4, mask=None, self_attn=False),
    dict(id="decoder_width", Bp=2, Lq=12, Lk=64, d=512, d_in=-1, h=8, mask="key", self_attn=False),
    dict(id="causal_width", Bp=2, Lq=12, Lk=12, d=512, d_in=-1, h=8, mask="causal", self_attn=True),
]


def _mha_inputs(c, seed):
    d_in = c["d"] if c["d_in"] < 0 else c["d_in"]
    q = _rand(seed, c["Bp"], c["Lq"], d_in)
    kv = q if c["self_attn"] else _rand(seed + 1, c["Bp"], c["Lk"], d_in)
    mask = None
    if c["mask"] == "key":
        lens = torch.randint(1, c["Lk"] + 1, (c["Bp"],), generator=torch.Generator().manual_seed(seed + 2))
        lens[0] = c["Lk"]
        mask = (torch.arange(c["Lk"])[None, :] < lens[:, None]).unsqueeze(1)           # [Bp,1,Lk]
    elif c["mask"] == "full":
        mask = torch.rand(c["Bp"], c["Lq"], c["Lk"], generator=torch.Generator().manual_seed(seed + 2)) > 0.3
        mask[:, :, 0] = True
    elif c["mask"] == "causal":                                                          # data/dataset.py subsequent_mask & pad
        mask = torch.tril(torch.ones(c["Lq"], c["Lk"], dtype=torch.bool)).unsqueeze(0).expand(c["Bp"], -1, -1).clone()
    return q, kv, mask


def _run_mha(c, precision, train):
    seed = 300 + len(c["id"])
    mod = _randomise(MultiHeadedAttention(c["h"], c["d"], c["d_in"], dropout=0.1), seed).to(DEV)
    mod.train(train)
    q, kv, mask = _mha_inputs(c, seed)
    w = _rand(seed + 5, *q.shape)
    dt = DT[precision]
    q = q.to(dt).float()                      # both sides start from the same (bf16-representable) inputs
    kv = q if c["self_attn"] else kv.to(dt).float()
    qd = q.to(DEV, dt).requires_grad_(True)
    kvd = qd if c["self_attn"] else kv.to(DEV, dt).requires_grad_(True)
    torch.manual_seed(77)
    n0 = _lib.lib().bist_launch_count()
    out = mod(qd, kvd, kvd, mask.to(DEV) if mask is not None else None)
    (out.float() * w.to(DEV)).sum().backward()
    torch.cuda.synchronize()
    assert _lib.lib().bist_launch_count() > n0

    P = {"a." + k: v.detach().cpu().clone().requires_grad_(True) for k, v in mod.state_dict().items()}
    drop = None
    if train:
        get = _mask_for(mod.last_dropout, seed)

        def drop(site, x):
            mul = get(x.numel()).view(x.shape)
            assert abs(float((mul == 0).float().mean()) - 0.1) < 0.06
            return x * mul
    qr = q.clone().requires_grad_(True)
    kvr = qr if c["self_attn"] else kv.clone().requires_grad_(True)
    ref, p_ref = O.mha(P, "a.", c["h"], qr, kvr, kvr, mask, drop)
    (ref * w).sum().backward()
    _cmp("out", out, ref, precision, fwd=True)
    _cmp("attn", mod.attn, p_ref, precision, fwd=True)
    _cmp("d_query", qd.grad, qr.grad, precision)
    if not c["self_attn"]:
        _cmp("d_key_value", kvd.grad, kvr.grad, precision)
    for name, p in mod.named_parameters():
        g = P["a." + name].grad
        if name == "linears.1.bias":          # zero by softmax shift invariance; rounding noise on both sides
            assert float(p.grad.abs().max()) < 1e-3 * float(P["a.linears.1.weight"].grad.abs().max()) * (
                1 if precision == "fp32" else 30)
            continue
        _cmp(name, p.grad, g, precision)
    return mod


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("c", MHA_CASES, ids=lambda c: c["id"])
def test_mha_gradients_match_oracle(c, precision):
    _run_mha(c, precision, train=False)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("c", [MHA_CASES[0], MHA_CASES[3]], ids=lambda c: c["id"])
def test_mha_train_mode_dropout_matches_oracle_with_the_same_mask(c, precision):
    mod = _run_mha(c, precision, train=True)
    assert mod.last_dropout is not None and mod.last_dropout[0] == 0.1


FFN_CASES = [dict(id="small", rows=(3, 5), d=32, d_ff=96, d_out=-1), dict(id="dout", rows=(7,), d=64, d_ff=128, d_out=48),
             dict(id="width", rows=(4, 16), d=512, d_ff=2048, d_out=-1)]


def _run_ffn(c, precision, train):
    seed = 500 + len(c["id"])
    mod = _randomise(PositionwiseFeedForward(c["d"], c["d_ff"], dropout=0.2, d_out=c["d_out"]), seed).to(DEV)
    mod.train(train)
    x = _rand(seed, *c["rows"], c["d"]).to(DT[precision]).float()
    d_out = c["d"] if c["d_out"] < 0 else c["d_out"]
    w = _rand(seed + 1, *c["rows"], d_out)
    xd = x.to(DEV, DT[precision]).requires_grad_(True)
    out = mod(xd)
    (out.float() * w.to(DEV)).sum().backward()
    torch.cuda.synchronize()
    P = {"f." + k: v.detach().cpu().clone().requires_grad_(True) for k, v in mod.state_dict().items()}
    drop = None
    if train:
        get = _mask_for(mod.last_dropout, seed)
        drop = lambda site, t: t * get(t.numel()).view(t.shape)
    xr = x.clone().requires_grad_(True)
    ref = O.ffn(P, "f.", xr, drop)
    (ref * w).sum().backward()
    _cmp("out", out, ref, precision, fwd=True)
    _cmp("d_x", xd.grad, xr.grad, precision, relu=True)
    for name, p in mod.named_parameters():
        _cmp(name, p.grad, P["f." + name].grad, precision, relu=name.startswith("w_1"))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("train", [False, True], ids=["eval", "train"])
@pytest.mark.parametrize("c", FFN_CASES, ids=lambda c: c["id"])
def test_ffn_gradients_match_oracle(c, train, precision):
    _run_ffn(c, precision, train)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_decoder_style_block_trains(precision):
    """SublayerConnection around a cross attention and around an FFN, train() mode (the shape of one sub-block of
    MultimodalDecoderLayer12, decoder.py:30-60): x -> x + D(MHA(LN(x), mem, mem, mask)) -> + D(FFN(LN(.)))."""
    B, L, Lm, d, h = 2, 6, 10, 64, 4
    sub = [_randomise(SublayerConnection(d, 0.2), 700 + i).to(DEV) for i in range(2)]
    att = _randomise(MultiHeadedAttention(h, d, dropout=0.1), 710).to(DEV)
    ff = _randomise(PositionwiseFeedForward(d, 4 * d, dropout=0.2), 711).to(DEV)
    for m in sub + [att, ff]:
        m.train()
    x, mem, w = _rand(720, B, L, d), _rand(721, B, Lm, d), _rand(722, B, L, d)
    mask = torch.ones(B, 1, Lm, dtype=torch.bool)
    mask[1, 0, 7:] = False
    dt = DT[precision]
    x, mem = x.to(dt).float(), mem.to(dt).float()
    xd, memd = x.to(DEV, dt).requires_grad_(True), mem.to(DEV, dt).requires_grad_(True)
    y1 = sub[0](xd, lambda z: att(z, memd, memd, mask.to(DEV)))
    y2 = sub[1](y1, ff)
    (y2.float() * w.to(DEV)).sum().backward()
    torch.cuda.synchronize()

    P = {}
    for pre, m in (("s0.", sub[0]), ("s1.", sub[1]), ("a.", att), ("f.", ff)):
        P.update({pre + k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()})
    gets = {"s0.out": _mask_for(sub[0].last_dropout, 0), "s1.out": _mask_for(sub[1].last_dropout, 0),
            "a.p": _mask_for(att.last_dropout, 0), "f.hidden": _mask_for(ff.last_dropout, 0)}
    drop = lambda site, t: t * gets[site](t.numel()).view(t.shape)
    xr, memr = x.clone().requires_grad_(True), mem.clone().requires_grad_(True)
    r1 = O.sublayer(P, "s0.", xr, lambda z: O.mha(P, "a.", h, z, memr, memr, mask, drop)[0], drop)
    r2 = O.sublayer(P, "s1.", r1, lambda z: O.ffn(P, "f.", z, drop), drop)
    (r2 * w).sum().backward()
    _cmp("out", y2, r2, precision, fwd=True)
    _cmp("d_x", xd.grad, xr.grad, precision)
    _cmp("d_mem", memd.grad, memr.grad, precision)
    for pre, m in (("s0.", sub[0]), ("s1.", sub[1]), ("a.", att), ("f.", ff)):
        for name, p in m.named_parameters():
            if pre + name == "a.linears.1.bias":
                continue
            _cmp(pre + name, p.grad, P[pre + name].grad, precision, relu=(pre in ("f.", "s1.")))


def test_module_gradients_share_one_block_per_module():
    """The parameter gradients of one module come back as slices of one fp32 block (one all-reduce per module)."""
    att = MultiHeadedAttention(4, 64).to(DEV)
    x = torch.randn(2, 5, 64, device=DEV, requires_grad=True)
    att.eval()
    att(x, x, x).sum().backward()
    ptrs = {p.grad.untyped_storage().data_ptr() for p in att.parameters()}
    assert len(ptrs) == 1
