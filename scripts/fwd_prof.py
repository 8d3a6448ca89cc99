"""A few config-2 encoder forwards (eager launches) for ncu: python scripts/fwd_prof.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bist_b200 import VideoPath, synth

dev = "cuda:0"
B, T, S, D, d, h, Lq, nb = 64, 40, 16, 2048, 512, 8, 16, 3
model = VideoPath(d, h, nb, D, precision="bf16").to(dev)
model.load_state_dict(synth.make_params(0, d, D, nb), strict=True)
b = synth.make_batch(1, B, T, S, D, Lq).to(dev)
eq = synth.make_encoded_query(1, B, Lq, d).to(dev)
with torch.no_grad():
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
        ft = model(b, eq)
torch.cuda.synchronize()
print("ok", float(ft["temporal_ft"].float().abs().mean()))
