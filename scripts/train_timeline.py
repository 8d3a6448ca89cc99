"""Where the headline training step's time goes: torch.profiler (CUPTI) kernel records of a few CUDA-graph replays of the config-3
(+audio) step -> GPU-busy time (union of kernel intervals), concurrency, and the kernels by total time.  GPU box only."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from bist_b200 import synth
from bist_b200.core import FULL_SHAPES, full_model
from bist_b200.graph import GraphedTrainStep

dev = torch.device("cuda:0")
fc = FULL_SHAPES
FORWARD = "--forward" in sys.argv      # the config-2 encoder forward (graph replay) instead of the training step
if FORWARD:
    from bist_b200 import VideoPath
    vm = VideoPath(512, 8, 3, 2048, precision="bf16").to(dev)
    vm.load_state_dict(synth.make_params(0, 512, 2048, 3), strict=True)
    vb = synth.make_batch(1, 64, 40, 16, 2048, 16).to(dev)
    veq = synth.make_encoded_query(1, 64, 16, 512).to(dev)
    with torch.no_grad():
        for _ in range(3):
            vm(vb, veq)
        gr = vm.capture(vb, veq)

    class _Opt:
        def step(self):
            pass
    step, opt = gr.replay, _Opt()
    for _ in range(5):
        step()
model, lc = (None, None) if (FORWARD or "--video" in sys.argv) else full_model("bf16", audio=True)
VIDEO = "--video" in sys.argv          # the video path alone in training (projection + 3 blocks + norms, MSE loss)
if VIDEO and not FORWARD:
    from bist_b200 import VideoPath
    from bist_b200.optimize import FusedAdam
    tm = VideoPath(512, 8, 3, 2048, dropout=0.1, precision="bf16").to(dev)
    tm.load_state_dict(synth.make_params(0, 512, 2048, 3), strict=True)
    tm.train()
    vb = synth.make_batch(1, 64, 40, 16, 2048, 16).to(dev)
    veq = synth.make_encoded_query(1, 64, 16, 512).to(dev)
    tgt = torch.randn(64, 16, 512, device=dev)

    def vloss():
        o = tm(vb, veq)
        return ((o["temporal_ft"] - tgt) ** 2).mean() + ((o["spatial_ft"] - tgt) ** 2).mean()
    params = list(tm.parameters())
    opt = FusedAdam(params, lr=1e-4)
    step = GraphedTrainStep(vloss, params, warmup=3)
    for _ in range(5):
        step(); opt.step()
if not FORWARD and not VIDEO:
    model = model.to(dev).train()
    b = synth.make_full_batch(1, 64, fc["T"], fc["S"], fc["D"], fc["Lq"], fc["Lh"], fc["Lc"], fc["Lt"], fc["V"], Ta=fc["Ta"], Da=fc["Da"]).to(dev)
    params = list(model.parameters())
    from bist_b200.optimize import FusedAdam
    opt = torch.optim.Adam(params, lr=1e-4, fused=True) if os.environ.get('BIST_TORCH_ADAM') else FusedAdam(params, lr=1e-4)
    step = GraphedTrainStep(lambda: lc.loss(model(b), b)[0], params, warmup=3)
    for _ in range(5):
        step(); opt.step()
torch.cuda.synchronize()
n = 3
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(n):
        step(); opt.step()
    torch.cuda.synchronize()
iv = []
named = []
import collections
by = collections.defaultdict(lambda: [0, 0.0])
for e in prof.events():
    if e.device_type != torch.autograd.DeviceType.CUDA:
        continue
    t0 = e.time_range.start
    t1 = e.time_range.end
    if t1 <= t0:
        continue
    iv.append((t0, t1))
    import re
    m = re.search(r"(\w+_kernel\w*|ln_rows\w*|ln2_rows\w*)(<[^(]*>)?\(", e.name)
    k = (m.group(1) + (m.group(2) or ""))[:60] if m else e.name.split("(")[0][-60:]
    by[k][0] += 1
    by[k][1] += (t1 - t0)
    named.append((t0, t1, k))
iv.sort()
busy, cur0, cur1 = 0.0, None, None
for a, c in iv:
    if cur1 is None or a > cur1:
        if cur1 is not None:
            busy += cur1 - cur0
        cur0, cur1 = a, c
    else:
        cur1 = max(cur1, c)
busy += (cur1 - cur0) if cur1 is not None else 0
span = iv[-1][1] - iv[0][0]
tot = sum(c - a for a, c in iv)
print("kernel records %d over %d steps; span %.2f ms/step, GPU busy (union) %.2f ms/step, sum of kernel times %.2f ms/step (avg concurrency %.2f), idle %.2f ms/step"
      % (len(iv), n, span / n / 1e3, busy / n / 1e3, tot / n / 1e3, tot / max(busy, 1), (span - busy) / n / 1e3))
for k, (c, t) in sorted(by.items(), key=lambda kv: -kv[1][1])[:45]:
    print("%-62s n/step %6.1f  %8.3f ms/step  avg %7.1f us" % (k, c / n, t / n / 1e3, t / c))

# exclusive time: how long each kernel is the ONLY one on the GPU (the step's critical path runs through these)
pts = []
for a, c, k in named:
    pts.append((a, 1, k))
    pts.append((c, -1, k))
pts.sort(key=lambda x: (x[0], x[1]))
active, last, excl = {}, None, collections.defaultdict(float)
for t, d_, k in pts:
    if last is not None and len(active) == 1:
        excl[next(iter(active))] += t - last
    if d_ == 1:
        active[k] = active.get(k, 0) + 1
    else:
        active[k] -= 1
        if active[k] == 0:
            del active[k]
    last = t
print("exclusive time (only kernel running), ms/step: total %.2f" % (sum(excl.values()) / n / 1e3))
for k, t in sorted(excl.items(), key=lambda kv: -kv[1])[:22]:
    print("   %-62s %7.3f" % (k, t / n / 1e3))
