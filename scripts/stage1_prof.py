"""One config-2 launch of the fused stage-1 kernel (both directions) for ncu.  Usage: python scripts/stage1_prof.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tests import stage1_ref as R

B, T, S, Lq, d, h = 64, 40, 16, 16, 512, 8
vft, tmask, dirs = R.make_inputs(11, B, T, S, Lq, d, h, "cuda:0")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    outs, _, _ = R.run_kernel(vft, tmask, dirs, [0, 1], Lq, h)
print("ok", float(outs[0].float().abs().mean()), float(outs[1].float().abs().mean()))
