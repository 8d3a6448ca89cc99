"""Workload for compute-sanitizer (memcheck / racecheck): the smoke shapes (VideoPath forward fp32 + bf16, backward) and one small
full-model training step (config-3 architecture with the audio stream at reduced sizes, train() mode with dropout, bf16 and fp32),
plus one launch of the fused stage-1 kernel at a shape that takes it (d_k = 64).  GPU box only:
  compute-sanitizer --tool memcheck --log-file gpurun_out/memcheck.log python scripts/sanitize_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

import __graft_entry__ as g
from bist_b200 import synth
from bist_b200.core import full_args, full_model

g.smoke()
dev = torch.device("cuda:0")
shapes = dict(B=4, T=12, S=16, D=256, d=128, h=2, Lq=9, Lh=20, Lc=11, Lt=7, V=501, Ta=12, Da=128)
for precision in ("bf16", "fp32"):
    args = full_args(shapes["d"], shapes["h"], n_blocks=2, dropout=0.1, audio=True)
    torch.manual_seed(0)
    model, lc = full_model(precision, shapes, args=args, audio=True)
    model = model.to(dev).train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
    b = synth.make_full_batch(3, shapes["B"], shapes["T"], shapes["S"], shapes["D"], shapes["Lq"], shapes["Lh"], shapes["Lc"],
                              shapes["Lt"], shapes["V"], Ta=shapes["Ta"], Da=shapes["Da"]).to(dev)
    for _ in range(2):
        opt.zero_grad(set_to_none=True)
        loss = lc.loss(model(b), b)[0]
        loss.backward()
        opt.step()
    torch.cuda.synchronize()
    print("train step ok (%s): loss %.4f" % (precision, float(loss)), flush=True)
import stage1_ref
r = stage1_ref.check_case(5, 3, 12, 16, 9, 128, 2)
print("stage1_fused ok:", r, flush=True)
assert max(r.values()) < 2e-2
print("sanitize_small: done")
