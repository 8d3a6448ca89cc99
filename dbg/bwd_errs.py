"""Print per-parameter gradient errors (max-normalised and relative L2) of the bf16 training path vs oracle autograd."""
import sys, torch
sys.path.insert(0, ".")
from tests.test_gpu_backward import _oracle_grads, _ours, CASES
from bist_b200 import synth
from oracle import bist_oracle as O
cases = [CASES[2], dict(seed=207, B=4, T=10, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=2, t2s=1, s2t=1)]
for c in cases:
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c["t2s"]), bool(c["s2t"]))
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    ref_out, ref_g, ref_geq = _oracle_grads(c, P, b, eq)
    for prec in ("bf16",):
        m, out, geq = _ours(c, P, b, eq, prec)
        print("case", c["seed"], prec)
        for name, p in m.named_parameters():
            r = ref_g[name]
            if r is None or p.grad is None: continue
            g = p.grad.float().cpu()
            l2 = float((g - r).norm() / r.norm().clamp_min(1e-30))
            print("  %-60s max %.3e  l2 %.3e  |ref|max %.2e" % (name.replace("mutlimodal_decoder.", ""), O.norm_err(g, r), l2, float(r.abs().max())))
