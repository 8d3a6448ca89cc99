"""Bring-up / regression check of the fused stage-1 kernel on a B200: several shapes, raw accumulator tiles and final c1
against the torch restatement (tests/stage1_ref.py), then a timing of config 2.  Usage: python scripts/stage1_check.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tests import stage1_ref as R

CASES = [
    # seed, B, T, S, Lq, d, h
    (1, 2, 8, 16, 16, 512, 8),
    (2, 3, 40, 16, 16, 512, 8),
    (3, 2, 33, 49, 7, 512, 8),
    (4, 2, 128, 49, 16, 512, 8),
    (5, 5, 6, 16, 5, 128, 2),
    (6, 2, 64, 16, 1, 256, 4),
    (7, 2, 17, 20, 9, 384, 6),
    (8, 3, 9, 12, 3, 128, 2),
]


def main():
    only = os.environ.get("S1_KINDS")
    for c in CASES:
        print("case seed=%d B=%d T=%d S=%d Lq=%d d=%d h=%d" % c, flush=True)
        for kinds in ([(1,), (0,), (0, 1)] if not only else [tuple(int(x) for x in only.split(","))]):
            print(" kinds", kinds, flush=True)
            r = R.check_case(*c, kinds=kinds, verbose=True, want_dbg=True)
            bad = [k for k, v in r.items() if not isinstance(k, tuple) and not (v < 5e-3)]
            print("  ->", "FAIL" if bad else "ok", flush=True)
    # timing at config 2
    B, T, S, Lq, d, h = 64, 40, 16, 16, 512, 8
    vft, tmask, dirs = R.make_inputs(11, B, T, S, Lq, d, h, "cuda:0")
    for kinds in ((0, 1), (1,), (0,)):
        use = [dirs[k] for k in kinds]
        R.run_kernel(vft, tmask, use, list(kinds), Lq, h)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 20
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
        import ctypes
        from bist_b200 import _lib
        L = _lib.lib()
        nd = len(kinds)
        outs = [torch.empty(B, Lq, (S if k == 0 else T), d, dtype=torch.bfloat16, device="cuda:0") for k in kinds]
        arr_i = (ctypes.c_int * nd)(*kinds)
        mk = lambda tl: (ctypes.c_void_p * nd)(*[t.data_ptr() for t in tl])
        a_qf, a_wv, a_bv, a_out = mk([x[0] for x in use]), mk([x[1] for x in use]), mk([x[2] for x in use]), mk(outs)
        ts = []
        for i in range(n + 3):
            flush.zero_()
            ev0.record()
            L.bist_stage1_fused(_lib.ptr(vft), _lib.ptr(tmask), nd, arr_i, a_qf, a_wv, a_bv, a_out, B, T, S, Lq, d, h, None,
                                _lib.stream_ptr())
            ev1.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(ev0.elapsed_time(ev1) * 1e3)
        ts.sort()
        print("config 2 kinds %s: median %.1f us, min %.1f us (L2 flushed between launches)" % (kinds, ts[len(ts) // 2], ts[0]), flush=True)
        if os.environ.get("BIST_S1_TRACE"):
            buf = (ctypes.c_longlong * (6 * 256))()
            L.bist_debug_s1_trace.argtypes = [ctypes.c_void_p]
            L.bist_debug_s1_trace(buf)
            tr = [[buf[r * 256 + i] for i in range(256)] for r in range(6)]
            t0 = min(x for r in tr for x in r if x > 0)
            names = ["x producer: slab requested", "mma: operands landed", "b producer: slot free", "epi0 (wait,acc,staged,softmax,pv,store)", "epi1", "epi0 fine (prebar,bar,ld sc,staged,smx,bar,Pst,fence,bar,-)"]
            for r in range(6):
                vals = [x - t0 for x in tr[r][:(60 if r == 5 else 96)] if x > 0]
                print("   trace %-32s %s" % (names[r], " ".join(str(v) for v in vals)), flush=True)


if __name__ == "__main__":
    t0 = time.time()
    main()
    print("done in %.1f s" % (time.time() - t0))
