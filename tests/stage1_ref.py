"""Kernel-level check of bist_stage1_fused (csrc/stage1_fused.cu) against a plain torch restatement of stage 1 of
temporal2spatial / spatial2temporal with folded keys (/root/reference/model/encoder.py:109-121,143-148; the fold is the
one oracle/bist_oracle.py pins against the reference formulation).  Shared by tests/test_gpu_stage1.py and
scripts/stage1_check.py; runs on the GPU in fp32 from the same bf16-rounded operands the kernel sees."""
import ctypes

import torch

from bist_b200 import _lib


def tile_geometry(Kn):
    "Mirror of the host-side tiling in stage1_fused(): keys padded to 8 / 16, groups per 128-lane tile."
    kp8, kp16 = (Kn + 7) // 8 * 8, (Kn + 15) // 16 * 16
    R = min(8, 128 // kp8)
    while R > 1 and R * kp16 > 160:
        R -= 1
    return kp8, kp16, R


def make_inputs(seed, B, T, S, Lq, d, h, device, ragged=True):
    g = torch.Generator().manual_seed(seed)
    vft = torch.randn(B, T, S, d, generator=g).to(torch.bfloat16)
    tmask = torch.ones(B, T, dtype=torch.uint8)
    if ragged:
        for b in range(1, B):
            tb = int(torch.randint((T + 1) // 2, T + 1, (1,), generator=g))
            tmask[b, tb:] = 0
    dirs = []
    for _ in range(2):
        qf = (torch.randn(B * Lq, h, d, generator=g) * (2.0 / d ** 0.5)).to(torch.bfloat16)
        wv = (torch.randn(d, d, generator=g) / d ** 0.5).to(torch.bfloat16)
        bv = torch.randn(d, generator=g) * 0.1
        dirs.append((qf, wv, bv))
    dev = torch.device(device)
    return vft.to(dev), tmask.to(dev), [tuple(t.to(dev) for t in x) for x in dirs]


def reference(vft, tmask, qf, wv, bv, kind, Lq, h, round_bf16=True):
    """c1 [B, Lq, G, d] fp32.  kind 0 = t2s (groups = regions, keys = clips, key mask), 1 = s2t."""
    B, T, S, d = vft.shape
    dk = d // h
    x = vft.float()
    xv = x @ wv.float().t()                                   # [B,T,S,d]
    if round_bf16:
        xv = xv.to(torch.bfloat16).float()
    q = qf.float().view(B, Lq, h, d)
    sc = torch.einsum("blhd,btsd->blhts", q, x)               # [B,Lq,h,T,S]
    if kind == 0:
        m = tmask.view(B, 1, 1, T, 1) == 0
        sc = sc.masked_fill(m, -1e9)
        p = torch.softmax(sc, dim=3)
    else:
        p = torch.softmax(sc, dim=4)
    if round_bf16:
        p = p.to(torch.bfloat16).float()
    xvh = xv.view(B, T, S, h, dk)
    if kind == 0:
        c = torch.einsum("blhts,btshj->blshj", p, xvh)        # [B,Lq,S,h,dk]
    else:
        c = torch.einsum("blhts,btshj->blthj", p, xvh)        # [B,Lq,T,h,dk]
    G = S if kind == 0 else T
    return c.reshape(B, Lq, G, d) + bv.float()


def run_kernel(vft, tmask, dirs, kinds, Lq, h, want_dbg=False):
    B, T, S, d = vft.shape
    L = _lib.lib()
    nd = len(kinds)
    outs, units = [], []
    for k in kinds:
        G, Kn = (S, T) if k == 0 else (T, S)
        outs.append(torch.full((B, Lq, G, d), float("nan"), dtype=torch.bfloat16, device=vft.device))
        _, _, R = tile_geometry(Kn)
        units.append(B * ((G + R - 1) // R))                  # a unit = one tile of up to R groups
    dbg = None
    if want_dbg:
        dbg = torch.zeros(sum(units), h, 128, 80, dtype=torch.float32, device=vft.device)
    arr_i = (ctypes.c_int * nd)(*kinds)
    mk = lambda ts: (ctypes.c_void_p * nd)(*[t.data_ptr() for t in ts])
    keep = [dirs[i] for i in range(nd)]
    rc = L.bist_stage1_fused(_lib.ptr(vft), _lib.ptr(tmask), nd, arr_i, mk([x[0] for x in keep]), mk([x[1] for x in keep]),
                             mk([x[2] for x in keep]), mk(outs), B, T, S, Lq, d, h, _lib.ptr(dbg), _lib.stream_ptr())
    _lib.check(rc, "bist_stage1_fused")
    torch.cuda.synchronize()
    return outs, dbg, units


def nerr(a, b):
    a, b = a.float(), b.float()
    if not torch.isfinite(a).all():
        return float("inf")
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def check_case(seed, B, T, S, Lq, d, h, kinds=(0, 1), device="cuda:0", verbose=False, want_dbg=False):
    """Returns {kind: normalised error of c1 vs the bf16-emulating restatement}; with want_dbg also checks raw tiles."""
    vft, tmask, dirs = make_inputs(seed, B, T, S, Lq, d, h, device)
    use = [dirs[k] for k in kinds]
    outs, dbg, units = run_kernel(vft, tmask, use, list(kinds), Lq, h, want_dbg)
    res = {}
    off = 0
    for i, k in enumerate(kinds):
        qf, wv, bv = use[i]
        ref = reference(vft, tmask, qf, wv, bv, k, Lq, h)
        res[k] = nerr(outs[i], ref)
        if verbose:
            exact = reference(vft, tmask, qf, wv, bv, k, Lq, h, round_bf16=False)
            print("  kind %d: c1 err vs bf16-emulating ref %.3e, vs fp32 ref %.3e" % (k, res[k], nerr(outs[i], exact)))
        if want_dbg:
            G, Kn = (S, T) if k == 0 else (T, S)
            kp8, _, R = tile_geometry(Kn)
            tpb = (G + R - 1) // R
            x = vft.float()
            q = qf.float().view(B, Lq, h, d)
            worst_s = worst_v = 0.0
            per_b = sum(units) // B                     # tiles of one sample, both directions (sample-major unit order)
            base = sum(units[:i]) // B                  # tiles of the directions launched before this one
            for u in (0, units[i] // 2, units[i] - 1):
                b, j = u // tpb, u % tpb
                g0 = j * R
                gu = b * per_b + base + j               # global unit index of the kernel
                for r in range(min(R, G - g0)):
                    g = g0 + r
                    rows = x[b, :, g, :] if k == 0 else x[b, g, :, :]          # [Kn, d]
                    for hh in (0, h - 1):
                        tile = dbg[gu, hh, r * kp8:r * kp8 + Kn]                 # group-major tile rows
                        es = rows @ q[b, :, hh, :].t()                           # [Kn, Lq]
                        ev = rows @ wv.float()[hh * 64:(hh + 1) * 64].t()        # [Kn, 64]
                        worst_s = max(worst_s, nerr(tile[:, :Lq], es))
                        worst_v = max(worst_v, nerr(tile[:, 16:80], ev))
            res[("raw", k)] = (worst_s, worst_v)
            if verbose:
                print("  kind %d: raw tile errors: scores %.3e values %.3e" % (k, worst_s, worst_v))
        off += units[i]
    return res
