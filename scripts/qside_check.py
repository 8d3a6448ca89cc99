"""Timeline of the fused query-side kernels (csrc/qside.cu) at config 2: BIST_QSIDE_TRACE=1 python scripts/qside_check.py
Prints, for CTA 0, 1, last-1, last of each kernel, the globaltimer stamps (ns, relative to the first) of: start, set-up done,
previous kernel done (griddepcontrol.wait), rows normalised, main product done, epilogue done, stores drained, exit."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bist_b200 import VideoPath, _lib, synth

dev = "cuda:0"
B, T, S, D, d, h, Lq, nb = 64, 40, 16, 2048, 512, 8, 16, 3
model = VideoPath(d, h, nb, D, precision="bf16").to(dev)
model.load_state_dict(synth.make_params(0, d, D, nb), strict=True)
b = synth.make_batch(1, B, T, S, D, Lq).to(dev)
eq = synth.make_encoded_query(1, B, Lq, d).to(dev)
L = _lib.lib()
with torch.no_grad():
    for _ in range(200):          # long enough for the SM clock to ramp: the stamps of the LAST pass are read
        model(b, eq)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (3 * 4 * 16))()
assert L.bist_debug_qside_trace(buf) == 0
names = ["attn", "fold", "ffn1"]
ev = ["start", "setup", "prev_done", "ln_done", "acc_full", "epi_done", "stores", "exit", "a_land", "gains", "ln_loop", "attn_staged", "attn_bar", "attn_scores", "attn_pv", "-"]
for m in range(3):
    if not any(buf[(m * 4 + s) * 16] for s in range(4)):
        print("(no timeline: run with BIST_QSIDE_TRACE=1)")
        break
    t0 = min(buf[(m * 4 + s) * 16] for s in range(4) if buf[(m * 4 + s) * 16])
    for s in range(4):
        row = [buf[(m * 4 + s) * 16 + e] for e in range(16)]
        print("%-5s cta-slot %d: " % (names[m], s) + "  ".join("%s %6d" % (ev[e], row[e] - t0) for e in sorted(range(15), key=lambda e: row[e]) if row[e]))
L.bist_prof_enable(1)
with torch.no_grad():
    for _ in range(5):
        torch.cuda._sleep(5_000_000)
        model(b, eq)
torch.cuda.synchronize()
out = ctypes.create_string_buffer(1 << 16)
L.bist_prof_collect(out, len(out))
L.bist_prof_enable(0)
for ln in out.value.decode().strip().splitlines():
    tag, n, tot = ln.split(",")
    print("%-18s %4d launches  %7.1f us each" % (tag, int(n) // 5, 1e3 * float(tot) / int(n)))
