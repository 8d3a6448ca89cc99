"""Full BiST training step at config-3 shapes: embeddings, text-encoder norms, VidEncoder8,
MultimodalDecoder8 (3 video / 3 caption / 3 decoder blocks, softmax gate), pointer generator over query + caption,
auto-encoder heads, label-smoothed loss, backward, fused Adam — per-step time, host enqueue time, per-tag kernel times.

    python scripts/core_bench.py [--precision bf16|fp32] [--steps 10] [--tags]
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bist_b200 import _lib, synth  # noqa: E402
from bist_b200.core import full_batch, full_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--eval", action="store_true")
    ap.add_argument("--tags", action="store_true")
    ap.add_argument("--ncu", action="store_true", help="bracket the timed steps with cudaProfilerStart/Stop (ncu --profile-from-start off)")
    ap.add_argument("--graph", action="store_true", help="forward+loss+backward as one CUDA graph (bist_b200/graph.py)")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    m, lc = full_model(a.precision)
    m = m.to(dev)
    m.train(not a.eval)
    b = full_batch(1, dev)
    opt = torch.optim.Adam(m.parameters(), lr=1e-4, fused=True)
    L = _lib.lib()

    def loss_fn():
        return lc.loss(m(b), b)[0]

    if a.graph:
        from bist_b200.graph import GraphedTrainStep
        graphed = GraphedTrainStep(loss_fn, list(m.parameters()))

        def step():
            graphed()
            opt.step()
    else:
        def step():
            opt.zero_grad(set_to_none=True)
            loss_fn().backward()
            opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    n0 = L.bist_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    if a.ncu:
        torch.cuda.profiler.start()
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    if a.ncu:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    print("full train step: %.3f ms (%d turns/s), host enqueue %.3f ms, %d library launches/step" % (
        ms, 64 / ms * 1e3, (t1 - t0) / a.steps * 1e3, (L.bist_launch_count() - n0) // a.steps))
    if a.tags:
        import ctypes
        L.bist_prof_enable(1)
        for _ in range(3):
            step()
        buf = ctypes.create_string_buffer(1 << 16)
        _lib.check(L.bist_prof_collect(buf, len(buf)), "prof")
        L.bist_prof_enable(0)
        rows = [ln.split(",") for ln in buf.value.decode().strip().splitlines()]
        for tag, n, tot in sorted(rows, key=lambda r: -float(r[2])):
            print("%-22s n=%6.1f  ms/step=%.4f  us/launch=%.1f" % (tag, int(n) / 3, float(tot) / 3, 1e3 * float(tot) / int(n)))


if __name__ == "__main__":
    main()
