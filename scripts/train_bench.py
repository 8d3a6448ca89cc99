"""Video-path training step at BASELINE config 2 sizes: forward (tape) + backward + Adam; per-kernel CUDA-event times."""
import argparse, ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bist_b200 import VideoPath, _lib, synth

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--no-adam", action="store_true")
ap.add_argument("--dropout", type=float, default=0.1)
a = ap.parse_args()
c = dict(B=a.B, T=40, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=3)
dev = torch.device("cuda:0")
L = _lib.lib()
m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=a.dropout, precision=a.precision).to(dev)
m.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
m.train()
if a.dropout == 0:
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
b = synth.make_batch(1, c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(dev)
eq = synth.make_encoded_query(1, c["B"], c["Lq"], c["d"]).to(dev)
tgt = torch.randn(c["B"], c["Lq"], c["d"], device=dev)
opt = torch.optim.Adam(m.parameters(), lr=1e-4, fused=True)

def step():
    opt.zero_grad(set_to_none=True)
    out = m(b, eq)
    loss = ((out["temporal_ft"] - tgt) ** 2).mean() + ((out["spatial_ft"] - tgt) ** 2).mean()
    loss.backward()
    if not a.no_adam:
        opt.step()
    return loss

for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
print("train step %.3f ms  -> %.0f turns/s (%s, B=%d, dropout %.2f)" % (ms, c["B"] / ms * 1e3, a.precision, c["B"], a.dropout))
# phases
def t(fn, n=5):
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): r = fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n, r
fw, out = t(lambda: m(b, eq))
loss = ((out["temporal_ft"] - tgt) ** 2).mean() + ((out["spatial_ft"] - tgt) ** 2).mean()
torch.cuda.synchronize(); e0.record(); loss.backward(); e1.record(); torch.cuda.synchronize()
print("forward(train) %.3f ms, backward %.3f ms" % (fw, e0.elapsed_time(e1)))
with torch.no_grad():
    m.eval(); fi, _ = t(lambda: m(b, eq)); m.train()
print("forward(inference path, eager) %.3f ms" % fi)
L.bist_prof_enable(1)
n = 3
for _ in range(n):
    torch.cuda._sleep(20_000_000)
    step()
buf = ctypes.create_string_buffer(1 << 16)
_lib.check(L.bist_prof_collect(buf, len(buf)), "prof")
L.bist_prof_enable(0)
rows = []
for ln in buf.value.decode().strip().splitlines():
    tag, k, tot = ln.split(",")
    rows.append((float(tot) / n, tag, int(k) / n))
tot = sum(r[0] for r in rows)
print("kernel total %.3f ms/step" % tot)
for ms_, tag, k in sorted(rows, reverse=True):
    print("  %-20s n=%6.1f %8.3f ms/step %8.1f us/launch  %4.1f%%" % (tag, k, ms_, 1e3 * ms_ / k, 100 * ms_ / tot))
