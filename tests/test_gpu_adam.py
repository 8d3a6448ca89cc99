"""FusedAdam (bist_adam_step: one launch for all parameters) against torch.optim.Adam — the optimizer the reference's NoamOpt wraps
(/root/reference/model/optimize.py:9-34): same parameters after several steps, odd sizes and a changing learning rate included."""
import pytest
import torch

from bist_b200.optimize import FusedAdam, NoamOpt

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_fused_adam_matches_torch_adam():
    g = torch.Generator().manual_seed(0)
    shapes = [(512, 512), (7,), (1000, 33), (16385,), (3, 5, 7), (2048, 512)]
    p_ref = [torch.randn(s, generator=g).to(DEV).requires_grad_(True) for s in shapes]
    p_our = [p.detach().clone().requires_grad_(True) for p in p_ref]
    ref = NoamOpt(512, 1.0, 4, torch.optim.Adam(p_ref, lr=0.0, betas=(0.9, 0.98), eps=1e-9))
    our = NoamOpt(512, 1.0, 4, FusedAdam(p_our, lr=0.0, betas=(0.9, 0.98), eps=1e-9))
    for it in range(7):
        for a, b in zip(p_ref, p_our):
            gr = torch.randn(a.shape, generator=g).to(DEV) * (0.1 + it)
            a.grad = gr.clone()
            b.grad = gr.clone() if it % 2 else gr        # the gradient tensors' addresses change between steps
        ref.step()
        our.step()
    torch.cuda.synchronize()
    for a, b in zip(p_ref, p_our):
        err = float((a - b).abs().max() / a.abs().max())
        assert err < 2e-6, err
    sd = our.optimizer.state_dict()
    assert sd["step"] == 7 and sd["exp_avg"].numel() == sum(p.numel() for p in p_our)
