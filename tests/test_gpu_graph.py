"""GPU: a training step captured into one CUDA graph (bist_b200/graph.py) equals the eager step — gradients, weight
refresh after an optimizer step, and fresh dropout masks on every replay."""
import pytest
import torch

from bist_b200 import VideoPath, _lib, synth
from bist_b200.graph import GraphedTrainStep
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
C = dict(seed=401, B=3, T=5, S=4, D=48, d=64, h=4, Lq=5, n_blocks=2)


def _setup(dropout, train):
    c = C
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=dropout, precision="fp32").to(DEV)
    m.load_state_dict(synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"]), strict=True)
    m.train(train)
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(DEV)
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"]).to(DEV)
    w = torch.randn(c["B"], c["Lq"], c["d"], generator=torch.Generator().manual_seed(9)).to(DEV)

    def loss_fn():
        out = m(b, eq)
        return ((out["temporal_ft"] + out["spatial_ft"]) * w).sum()
    return m, loss_fn


def test_graphed_step_equals_eager_and_follows_weight_updates():
    m, loss_fn = _setup(0.0, train=False)
    params = list(m.parameters())
    loss_e = loss_fn()
    loss_e.backward()
    eager = [p.grad.clone() for p in params]
    loss_e = float(loss_e.detach())     # drop the eager autograd graph: its AccumulateGrad nodes live on the default stream
    step = GraphedTrainStep(loss_fn, params)
    try:
        loss_g = step()
        torch.cuda.synchronize()
        assert abs(float(loss_g.detach()) - float(loss_e)) <= 1e-5 * abs(float(loss_e))
        for p, g in zip(params, eager):
            assert O.norm_err(p.grad.cpu(), g.cpu()) < 1e-5
        opt = torch.optim.SGD(params, lr=5e-3)
        opt.step()                                       # eager optimizer on the graph's static gradients
        loss_g2 = float(step())
        with torch.no_grad():
            loss_e2 = float(loss_fn())                   # eager forward with the updated weights
        assert abs(loss_g2 - float(loss_e)) > 1e-3 * abs(float(loss_e))      # the replay saw the new weights
        assert abs(loss_g2 - loss_e2) <= 1e-5 * abs(loss_e2)
    finally:
        step.close()


def test_graphed_step_draws_fresh_dropout_masks_per_replay():
    m, loss_fn = _setup(0.3, train=True)
    params = list(m.parameters())
    torch.manual_seed(5)
    step = GraphedTrainStep(loss_fn, params)
    try:
        a = float(step())
        b = float(step())
        assert a != b                                    # new masks on every replay
        step.counter.fill_(0)                            # replaying from the same counter reproduces the step
        assert float(step()) == a
        assert float(step()) == b
    finally:
        step.close()


def test_pipelined_forward_equals_direct_forward_for_every_batch():
    """PipelinedVideoPath overlaps the next batch's H2D with the current replay: every batch's outputs must still be
    exactly those of a direct forward on that batch."""
    from bist_b200 import PipelinedVideoPath
    c = dict(seed=411, B=4, T=6, S=4, D=64, d=64, h=4, Lq=5, n_blocks=2)
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision="fp32").to(DEV)
    m.load_state_dict(synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"]), strict=True)
    m.eval()
    batches = []
    for i in range(5):
        hb = synth.make_batch(500 + i, c["B"], c["T"], c["S"], c["D"], c["Lq"], full_clips=True)
        hb.fts, hb.query_mask, hb.temporal_mask = hb.fts.pin_memory(), hb.query_mask.pin_memory(), hb.temporal_mask.pin_memory()
        batches.append((hb, synth.make_encoded_query(500 + i, c["B"], c["Lq"], c["d"]).pin_memory()))
    with torch.no_grad():
        direct = [m(hb.to(DEV), heq.to(DEV)) for hb, heq in batches]
        direct = [(o["temporal_ft"].cpu(), o["spatial_ft"].cpu()) for o in direct]
    pipe = PipelinedVideoPath(m, batches[0][0].to(DEV), batches[0][1].to(DEV), depth=2)
    outs = [(torch.empty(c["B"], c["Lq"], c["d"]).pin_memory(), torch.empty(c["B"], c["Lq"], c["d"]).pin_memory())
            for _ in batches]
    for (hb, heq), (ot, os_) in zip(batches, outs):
        pipe.submit(hb, heq, ot, os_)
    pipe.drain()
    for (ot, os_), (dt_, ds_) in zip(outs, direct):
        assert torch.equal(ot, dt_) and torch.equal(os_, ds_)
