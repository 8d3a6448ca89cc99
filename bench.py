#!/usr/bin/env python
"""bench.py — BiST video-encoder forward throughput on B200 (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W [--precision bf16|fp32] [--impl ours|reference]

A step is one forward of the hot path (VidEncoder8 projection + 3 x VidEncoderLayer4 + output
norms) over one synthetic batch of B=64 dialogue turns, T=40 clips, S=16 regions, D=2048,
d_model=512, Lq=16.  One process per GPU; for N>1 every rank runs its own batch of 64 turns
(dialogue turns are independent, no data-path collective; weak scaling) and the time is the max
over ranks.  Prints ONE JSON line on rank 0 (contract in the task statement):
  value     turns/s with inputs resident in HBM
  e2e       turns/s through the public module API with pinned HOST inputs, H2D and D2H inside
  roofline  dominant kernel, per-launch CUDA-event time from the library's profiling hooks
  cpu_baseline  the CPU oracle (port of the reference formulation) on this box's host cores
`--impl reference` times that CPU oracle alone on the same config (the reference is pure
Python/PyTorch and /root/reference does not travel to the GPU box; oracle/bist_oracle.py keeps
the reference's formulation op for op).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

CFG = dict(B=64, T=40, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=3)
WORKLOAD = "BiST video encoder forward (VidEncoder8 + 3x VidEncoderLayer4 s2t+t2s + out norms), B=64 T=40 S=16 D=2048 d_model=512 Lq=16"
METRIC, UNIT = "dialogue_turns_per_sec", "turns/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm=j["hbm_gbs"], tc_burst=j["bf16_tflops"], tc_sust=j["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sust=1400.0, src="fallback")


def algorithmic_bytes(c, e_act):
    """fts read (fp32) + vft write + one vft read per block + weights + query streams."""
    N, Q, d, D, L = c["B"] * c["T"] * c["S"], c["B"] * c["Lq"], c["d"], c["D"], c["n_blocks"]
    return N * D * 4 + N * d * e_act + L * N * d * e_act + (D * d + 40 * L * d * d) * e_act + (1 + 2 * L) * 2 * Q * d * 4


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 100 ms during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_oracle_turns_per_sec(c, rows, reps, threads):
    """Time the CPU oracle (reference formulation) on `rows` dialogue turns of the workload."""
    from bist_b200 import synth
    from oracle import bist_oracle as O
    torch.set_num_threads(threads)
    P = synth.make_params(0, c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(1, rows, c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(1, rows, c["Lq"], c["d"])
    best = float("inf")
    with torch.no_grad():
        for i in range(reps + 1):          # first pass is the warm-up
            t0 = time.perf_counter()
            O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask, c["n_blocks"])
            dt = time.perf_counter() - t0
            if i > 0:
                best = min(best, dt)
    return rows / best, best


def cpu_oracle_train_turns_per_sec(model, rows, reps, threads):
    """Time the CPU oracle's full training step (forward + label-smoothed losses + autograd backward, fp32; no
    optimizer) on `rows` dialogue turns of config 3, with the model's own weights."""
    from bist_b200 import synth
    from bist_b200.core import FULL_SHAPES, PAD, SMOOTHING
    from oracle import bist_oracle as O
    torch.set_num_threads(threads)
    c = FULL_SHAPES
    leaves, P = {}, {}
    for k, v in model.state_dict().items():
        if k.endswith(".pe"):
            continue
        key = v.data_ptr()
        if key not in leaves:
            leaves[key] = v.detach().float().cpu().clone().requires_grad_(True)
        P[k] = leaves[key]
    a = model.args
    audio = getattr(a, "nb_aenc_blocks", 0) > 0
    args = dict(nb_venc_blocks=a.nb_venc_blocks, nb_cenc_blocks=a.nb_cenc_blocks, nb_aenc_blocks=a.nb_aenc_blocks if audio else 0,
                enc_vc_combine=a.enc_vc_combine, include_caption=a.include_caption, dec_st_combine=a.dec_st_combine,
                t2s=a.t2s, s2t=a.s2t, enc_st_combine="none", ptr_ft=a.ptr_ft, auto_encoder=a.auto_encoder, mask_unk=1)
    b = synth.make_full_batch(1, rows, c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                              Ta=c["Ta"] if audio else 0, Da=c["Da"] if audio else 0)
    best = float("inf")
    for i in range(reps + 1):              # first pass is the warm-up
        for v in leaves.values():
            v.grad = None
        t0 = time.perf_counter()
        ft = O.mtn_forward(P, args, b, a.att_h, a.nb_blocks)
        O.loss_compute(P, args, ft, b, PAD, SMOOTHING)[0].backward()
        dt = time.perf_counter() - t0
        if i > 0:
            best = min(best, dt)
    return rows / best, best


def cpu_oracle_decode_turns_per_sec(model, shapes, max_len_sample, max_len, beam, threads):
    """Time the CPU oracle's beam search (the reference's procedure: every hypothesis of every step re-runs the whole
    decoder, video blocks included) on ONE dialogue turn of config 5 for `max_len_sample` of the `max_len` steps, and
    scale by the count of hypothesis-decodes (1 at step 0, `beam` afterwards)."""
    from bist_b200 import synth
    from bist_b200.core import FULL_SHAPES
    from oracle import bist_oracle as O
    torch.set_num_threads(threads)
    c = dict(FULL_SHAPES, **shapes)
    P = {k: v.detach().float().cpu() for k, v in model.state_dict().items() if not k.endswith(".pe")}
    a = model.args
    args = dict(nb_venc_blocks=a.nb_venc_blocks, nb_cenc_blocks=a.nb_cenc_blocks, nb_aenc_blocks=0,
                enc_vc_combine=a.enc_vc_combine, include_caption=a.include_caption, dec_st_combine=a.dec_st_combine,
                t2s=a.t2s, s2t=a.s2t, enc_st_combine="none", ptr_ft=a.ptr_ft, auto_encoder=a.auto_encoder, mask_unk=1)
    b = synth.make_full_batch(5, 1, c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"])
    with torch.no_grad():
        t0 = time.perf_counter()
        O.beam_search(P, args, b, a.att_h, a.nb_blocks, max_len_sample, 2, 0, 3, beam=beam, penalty=1.0, nbest=5, min_len=1)
        dt = time.perf_counter() - t0
    done, full = 1 + beam * (max_len_sample - 1), 1 + beam * (max_len - 1)
    return 1.0 / (dt * full / done), dt, done, full


def kernel_model(c, e_act, nd=2):
    """Algorithmic (flops, bytes) per step summed over the launches of each kernel tag."""
    B, T, S, D, d, h, Lq, L = c["B"], c["T"], c["S"], c["D"], c["d"], c["h"], c["Lq"], c["n_blocks"]
    N, TS, Q = B * T * S, T * S, B * Lq
    dff = 4 * d
    c1 = Q * (S + T) * d                      # stage-1 context elements, both directions
    m = {
        "vidproj_fused": (2.0 * N * D * d, N * D * 4 + d * D * 2 + N * d * e_act),
        "vidproj_gemm": (2.0 * N * D * d, N * D * 4 + d * D * 4 + N * d * 4),
        "xv_gemm": (L * 2.0 * N * d * nd * d, L * (N * d * e_act + nd * d * d * e_act + N * nd * d * e_act)),
        "scores1_gemm": (L * nd * 2.0 * Q * h * TS * d, L * nd * (N * d * e_act + Q * h * d * e_act + Q * h * TS * 4)),
        "softmax_pv_mma": (L * nd * 2.0 * Q * TS * d, L * (nd * Q * h * TS * 4 + nd * N * d * e_act + c1 * e_act)),
        "stage2_mma": (L * 4.0 * c1 * h, L * (c1 * e_act + 2 * nd * Q * h * d * e_act)),
        "fold_gemm": (L * 2 * nd * 2.0 * Q * d * d, L * 2 * nd * (Q * d * e_act + d * d * e_act + Q * h * d * e_act)),
        "vfold2_gemm": (L * nd * 2.0 * Q * d * d, L * nd * (Q * h * d * e_act + d * d * e_act + 2 * Q * d * e_act)),
        # per block: QKV (3), O0, {Q1,Q2,xV2} (3), O2 (d x d each), FFN (2 x d x dff), all on nd*Q rows
        "query_side_gemm": (L * nd * 2.0 * Q * d * (8 * d + 2 * dff),
                            L * nd * ((8 * d * d + 2 * d * dff) * e_act + Q * (16 * d + 2 * dff) * e_act)),
        "layer_norm": (0.0, L * nd * Q * d * (4 * 4 + 5 * e_act) + 2 * Q * d * 8),
        "self_attn_mma": (L * nd * 4.0 * Q * Lq * d, L * nd * Q * 4 * d * e_act),
    }
    m["softmax_pv"], m["stage2_v2"] = m["softmax_pv_mma"], m["stage2_mma"]
    return m


def roofline_for(tag, kern, km, peaks, traffic):
    """Roofline entry of one kernel tag: per-launch algorithmic work / per-launch CUDA-event time."""
    fl, by = km[tag]
    n = kern[tag]["launches_per_step"]
    sec = kern[tag]["us_per_launch"] * 1e-6
    fl, by = fl / n, by / n
    ridge = peaks["tc_sust"] * 1e12 / (peaks["hbm"] * 1e9)
    if by > 0 and fl / by > ridge:
        a, p = fl / sec / 1e12, peaks["tc_sust"]
        r = {"bound": "tensor", "achieved": a, "peak": p, "unit": "TFLOP/s", "frac": a / p}
    else:
        a, p = by / sec / 1e9, peaks["hbm"]
        r = {"bound": "hbm", "achieved": a, "peak": p, "unit": "GB/s", "frac": a / p}
    r.update({"traffic": traffic, "kernel": tag, "us_per_launch": kern[tag]["us_per_launch"],
              "algorithmic_bytes_per_launch": by, "algorithmic_flops_per_launch": fl})
    return r


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    rows = 16
    from bist_b200 import synth
    from oracle import bist_oracle as O
    torch.set_num_threads(threads)
    c = CFG
    P = synth.make_params(0, c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(1, rows, c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(1, rows, c["Lq"], c["d"])
    times = []
    with torch.no_grad():
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask, c["n_blocks"])
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
    total = sum(times)
    v = rows * len(times) / total
    sample = "%d of the %d dialogue turns of the workload per step, fp32, torch CPU, %d threads" % (rows, c["B"], threads)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_step": rows},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from bist_b200 import PipelinedVideoPath, VideoPath, _lib, synth
    L = _lib.lib()
    _lib.check(L.bist_device_check(), "bist_device_check")
    c = CFG
    e_act = 2 if args.precision == "bf16" else 4
    model = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision=args.precision).to(dev)
    model.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
    hb = synth.make_batch(1 + rank, c["B"], c["T"], c["S"], c["D"], c["Lq"])
    heq = synth.make_encoded_query(1 + rank, c["B"], c["Lq"], c["d"]).pin_memory()
    hb.fts, hb.query_mask, hb.temporal_mask = hb.fts.pin_memory(), hb.query_mask.pin_memory(), hb.temporal_mask.pin_memory()
    db, deq = hb.to(dev), heq.to(dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident throughput -----------------------------------------------------------
    with torch.no_grad():
        n0 = L.bist_launch_count()
        model(db, deq)
        torch.cuda.synchronize()
        launches_per_step = L.bist_launch_count() - n0      # kernels of libbist_b200 in one forward
        for _ in range(max(args.warmup, 3)):
            model(db, deq)
        if args.no_graph:
            step = lambda: model(db, deq)
        else:
            graphed = model.capture(db, deq)                # CUDA graph of the same launches
            step = graphed.replay
            for _ in range(max(args.warmup, 3)):
                step()
        sampler = ClockSampler(local)
        ms_total = timed(step, args.steps)
        launches = launches_per_step * args.steps
        clocks = sampler.stop()
        ms_step = ms_total / args.steps
        value = world * c["B"] / (ms_step * 1e-3)

        # ---- end to end: pinned host inputs -> H2D -> forward -> D2H of the encoder outputs ------
        # two pinned output buffers and two captured copies of the forward: batch i+1 crosses PCIe while batch i computes
        outs = [(torch.empty((c["B"], c["Lq"], c["d"]), dtype=torch.float32).pin_memory(),
                 torch.empty((c["B"], c["Lq"], c["d"]), dtype=torch.float32).pin_memory()) for _ in range(2)]
        out_t = outs[0][0]
        pipe = None if args.no_graph else PipelinedVideoPath(model, db, deq, depth=2)
        e2e_n = [0]

        def e2e_step():
            o_t, o_s = outs[e2e_n[0] % 2]
            e2e_n[0] += 1
            if args.no_graph:
                ft = model(hb.to(dev, non_blocking=True), heq.to(dev, non_blocking=True))
                o_t.copy_(ft["temporal_ft"], non_blocking=True)
                o_s.copy_(ft["spatial_ft"], non_blocking=True)
            else:
                pipe.submit(hb, heq, o_t, o_s)              # H2D of this step's inputs, replay, D2H of its outputs

        for _ in range(2):
            e2e_step()
        e2e_ms = timed(e2e_step, args.steps) / args.steps

        # ---- the same end-to-end step fed by the ingest cache: features held in pinned host memory
        # in bf16, ragged videos copied straight into the zero-padded static batch, 2-byte projection -------------------
        ingest = None
        if pipe is not None and args.precision == "bf16":
            import numpy as np
            from bist_b200.ingest import FeatureCache
            rs = np.random.RandomState(100 + rank)
            n_vid = 96
            lens = rs.randint(c["T"] // 2, c["T"] + 1, size=n_vid)
            cache = FeatureCache(c["D"], c["S"], int(lens.sum()), dtype=torch.bfloat16)
            picks = [[int(v) for v in rs.permutation(n_vid)[:c["B"]]] for _ in range(8)]
            keep = {}
            for v in range(n_vid):
                arr = rs.uniform(0, 1, size=(int(lens[v]), c["S"], c["D"])).astype(np.float32)
                cache.add(v, arr)
                if v in picks[0]:
                    keep[v] = arr
            db16 = db.to(dev)
            db16.fts = db.fts.to(torch.bfloat16)
            pipe16 = PipelinedVideoPath(model, db16, deq, depth=2)
            ing_n, ing_bytes = [0], [0]

            def ingest_step():
                vids = picks[ing_n[0] % len(picks)]
                o_t, o_s = outs[ing_n[0] % 2]
                ing_n[0] += 1
                ing_bytes[0] = cache.host_bytes(vids)

                def fill(g):
                    cache.collate(vids, dev, out=(g.static_b.fts, g.static_b.temporal_mask), T=c["T"])
                    g.static_b.query_mask.copy_(hb.query_mask, non_blocking=True)
                    g.static_eq.copy_(heq, non_blocking=True)
                pipe16.submit(None, None, o_t, o_s, fill=fill)

            for _ in range(2):
                ingest_step()
            ing_ms = timed(ingest_step, args.steps) / args.steps
            ingest = {"metric": "e2e_from_bf16_feature_cache", "value": world * c["B"] / (ing_ms * 1e-3), "unit": UNIT,
                      "ms_per_step": ing_ms, "h2d_bytes_per_step": int(ing_bytes[0]) + hb.query_mask.numel() + heq.numel() * 4,
                      "d2h_bytes_per_step": 2 * out_t.numel() * 4,
                      "what": "same step as e2e, but the batch is collated from bist_b200.ingest.FeatureCache: 96 ragged videos "
                              "(T_i in [T/2, T]) held in pinned host memory in bf16, 64 picked per step, each copied straight "
                              "into the zero-padded static batch (no padded host copy), temporal mask from the clip counts, "
                              "2-byte projection (bist_vidproj_fwd_bf16in); results equal bf16 mode on the fp32 array"}
            if rank == 0 and world == 1 and not args.no_cpu_baseline:
                # what data/dataset.py:164-174,79 + train.py's .cuda() do for the same 64 videos, arrays already in memory:
                # zero-padded fp32 host batch, mask from the sums, pageable H2D
                from bist_b200.ingest import reference_collate
                arrs = [keep[v] for v in picks[0]]
                best = float("inf")
                for _ in range(3):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    f_ref, m_ref = reference_collate(arrs)
                    f_dev, m_dev = f_ref.to(dev), m_ref.to(dev)
                    torch.cuda.synchronize()
                    best = min(best, time.perf_counter() - t0)
                ingest["reference_collate"] = {"ms_per_batch": best * 1e3, "cores": os.cpu_count() or 1,
                                               "what": "the reference's collate for the same 64 videos (numpy zero-pad to the "
                                                       "longest video, torch.from_numpy, mask = sum != 0, fp32 pageable "
                                                       "H2D), best of 3, arrays already in memory (no .npy reads)"}
                del f_ref, m_ref, f_dev, m_dev, arrs
            del pipe16, cache, keep
        h2d = hb.fts.numel() * 4 + hb.query_mask.numel() + hb.temporal_mask.numel() + heq.numel() * 4
        if args.no_graph:
            h2d += hb.query.numel() * 8 + hb.spatial_mask.numel() + hb.t_len.numel() * 8 + hb.q_len.numel() * 8
        d2h = 2 * out_t.numel() * 4

        # ---- per-kernel CUDA-event times (library hooks), separate pass --------------------------
        torch.cuda.synchronize()
        L.bist_prof_enable(1)
        prof_steps = min(args.steps, 5)
        for _ in range(prof_steps):
            torch.cuda._sleep(6_000_000)      # ~3 ms of GPU spin: the host enqueues the whole step before it runs,
            model(db, deq)                    # so the per-launch event times are not inflated by launch gaps
        import ctypes
        buf = ctypes.create_string_buffer(1 << 16)
        _lib.check(L.bist_prof_collect(buf, len(buf)), "bist_prof_collect")
        L.bist_prof_enable(0)
    kern = {}
    for ln in buf.value.decode().strip().splitlines():
        tag, n, tot = ln.split(",")
        kern[tag] = {"launches_per_step": int(n) / prof_steps, "ms_per_step": float(tot) / prof_steps,
                     "us_per_launch": 1e3 * float(tot) / int(n)}
    peaks = load_peaks()
    km = kernel_model(c, e_act)
    total_k = sum(v["ms_per_step"] for v in kern.values()) or 1.0
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    traffic_db = json.load(open(tp)).get(args.precision, {}) if os.path.exists(tp) else {}
    per_kernel = {}
    for tag in kern:
        if tag in km:
            per_kernel[tag] = roofline_for(tag, kern, km, peaks, traffic_db.get(tag))
            per_kernel[tag]["share_of_step"] = kern[tag]["ms_per_step"] / total_k
    # dominant kernel = the single kernel (not an aggregate of many small launches) with the largest share of the step
    single = [t for t in per_kernel if kern[t]["launches_per_step"] <= 2 * c["n_blocks"]]
    top = max(single, key=lambda t: kern[t]["ms_per_step"]) if single else None
    roof = None
    if top:
        roof = dict(per_kernel[top])
        roof["peak_source"] = peaks["src"] + " (sustained bf16 / copy bandwidth, MEASURED_PEAKS.json)"
    whole = algorithmic_bytes(c, e_act) / (ms_step * 1e-3) / 1e9

    # ---- training steps (forward with tape + backward + grad all-reduce + Adam) ------------------------------
    # forward + loss + backward replay as ONE CUDA graph (bist_b200/graph.py: weights re-packed and dropout masks
    # re-drawn inside the graph); the gradient all-reduce (NCCL, N>1) and the fused Adam step run eagerly after it.
    train = train_full = train_full_audio = None
    if not args.no_train:
        from bist_b200.dist import allreduce_grads, overlap_grad_allreduce
        from bist_b200.graph import GraphedTrainStep

        def measure_train(mod, loss_fn, extra_leaves, what):
            plist = list(mod.parameters())
            opt = torch.optim.Adam(plist, lr=1e-4, fused=True)
            ncoll = [0]
            n0 = L.bist_launch_count()
            if args.no_graph:
                def train_step():
                    opt.zero_grad(set_to_none=True)
                    with overlap_grad_allreduce():           # each block's all-reduce starts as its backward is enqueued
                        loss_fn().backward()
                    ncoll[0] = allreduce_grads(plist)
                    opt.step()
                for _ in range(3):
                    train_step()
                per_step = (L.bist_launch_count() - n0) // 3
            else:
                in_graph = world > 1 and os.environ.get("BIST_GRAPH_ALLREDUCE", "0") == "1"
                gstep = GraphedTrainStep(loss_fn, plist + extra_leaves, warmup=3, allreduce=in_graph)
                per_step = (L.bist_launch_count() - n0) // 4   # 3 warm-up steps + the captured one

                def train_step():
                    gstep()
                    ncoll[0] = gstep.collectives if in_graph else allreduce_grads(plist)
                    opt.step()
                for _ in range(3):
                    train_step()
            tsteps = max(3, min(args.steps, 10))
            tms = timed(train_step, tsteps) / tsteps
            if not args.no_graph:
                gstep.close()
            return {"value": world * c["B"] / (tms * 1e-3), "unit": UNIT, "ms_per_step": tms,
                    "gpu_launches_per_step": int(per_step), "collectives_per_step": ncoll[0],
                    "launch": "eager" if args.no_graph else "cuda_graph(fwd+loss+bwd) + eager all-reduce + fused Adam",
                    "what": what}

        tm = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.1, precision=args.precision).to(dev)
        tm.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
        tm.train()                                       # dropout on: 0.1 sublayer/FFN (train_configs.py:34), 0.1 attention (mtn.py:77)
        tgt = torch.randn(c["B"], c["Lq"], c["d"], device=dev)

        def video_loss():
            out = tm(db, deq)
            return ((out["temporal_ft"] - tgt) ** 2).mean() + ((out["spatial_ft"] - tgt) ** 2).mean()

        train = measure_train(tm, video_loss, [],
                              "video path only (projection + 3 blocks + out norms): forward with tape, MSE loss, backward "
                              "through libbist_b200, gradient all-reduce (NCCL, N>1), fused Adam; train() mode, dropout 0.1")
        train["metric"] = "video_path_train_turns_per_sec"
        del tm

        from bist_b200.core import full_batch, full_model
        fm, flc = full_model(args.precision)
        fm = fm.to(dev).train()
        fb = full_batch(1 + rank, dev)

        def full_loss():
            return flc.loss(fm(fb), fb)[0]

        train_full = measure_train(fm, full_loss, [],
                                   "BASELINE config 3, the full BiST AVSD training step, per GPU B=64 dialogue turns (T=40, S=16, "
                                   "D=2048, Lq=16, L_his=64, L_cap=32, L_tgt=12, V=4096, d=512, 3 decoder / 3 video / 3 caption "
                                   "blocks): embeddings + positional encoding, text-encoder norms, VidEncoder8, MultimodalDecoder8 "
                                   "with softmax gate, pointer generator over query and caption, auto-encoder heads on cap/temporal/"
                                   "spatial, label smoothing 0.1, backward, gradient all-reduce (NCCL, N>1), fused Adam; train() "
                                   "mode, dropout 0.1; all arithmetic in libbist_b200 (exec.sh's configuration: nb_aenc_blocks=0, "
                                   "train_configs.py:30; the audio encoder is the train_full_audio leg)")
        train_full["metric"] = "bist_train_turns_per_sec"
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, best = cpu_oracle_train_turns_per_sec(fm, 8, 2, threads)
            train_full["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                          "sample": "8 dialogue turns of config 3, forward + losses + autograd backward "
                                                    "(no optimizer), best of 2 after 1 warm-up (%.2f s each), fp32 torch CPU" % best}
        del fm, fb

        # config 3 with the audio encoder it names: ft_sizes=[2048, 128], three AudioEncoderLayer blocks over Ta=40 VGGish-like
        # frames, a fourth gate column and the audio auto-encoder head (encoder.py:83-91,220-235; decoder.py:154,167)
        am, alc = full_model(args.precision, audio=True)
        am = am.to(dev).train()
        ab = full_batch(1 + rank, dev, audio=True)

        def audio_loss():
            return alc.loss(am(ab), ab)[0]

        train_full_audio = measure_train(am, audio_loss, [],
                                         "train_full plus the audio stream of BASELINE config 3: audio features [B, Ta=40, 128] "
                                         "through VidEncoder8.a_W + a_in_norm, 3 AudioEncoderLayer blocks, audio column of the "
                                         "softmax gate, audio auto-encoder head; same step otherwise")
        train_full_audio["metric"] = "bist_train_turns_per_sec"
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            v, best = cpu_oracle_train_turns_per_sec(am, 8, 1, threads)
            train_full_audio["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                                "sample": "8 dialogue turns, forward + losses + autograd backward (no optimizer), "
                                                          "1 pass after 1 warm-up (%.2f s), fp32 torch CPU" % best}
        del am, ab

    # ---- beam-search decode at BASELINE config 5 (long video: S=49, T=128; beam 5, max_len 12), B=8 turns per GPU ---
    decode = None
    if not args.no_train:
        from bist_b200.core import full_batch, full_model
        from bist_b200.decode import beam_search_decode_batch
        dshape = dict(B=8, T=128, S=49)
        dm, _ = full_model(args.precision, dshape)
        dm = dm.to(dev).eval()
        dbatch = full_batch(5 + rank, dev, dshape)

        def decode_step():
            beam_search_decode_batch(dm, dbatch, 12, 2, 0, 3, 1, beam=5, penalty=1.0, nbest=5, min_len=1)

        for _ in range(2):
            decode_step()
        dsteps = 3
        dms = timed(decode_step, dsteps) / dsteps
        decode = {"metric": "beam_search_turns_per_sec", "value": world * dshape["B"] / (dms * 1e-3), "unit": UNIT,
                  "ms_per_batch": dms,
                  "what": "BASELINE config 5: generate.py-style beam search (beam 5, max_len 12, n-best 5) for a batch of 8 "
                          "dialogue turns per GPU over long videos (T=128 clips x S=49 regions x 2048), whole model in "
                          "libbist_b200; the video / caption blocks run once per turn and all hypotheses of all turns go "
                          "through the decoder layers as one batch (the reference re-runs everything per hypothesis per step)"}
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, dt, done, full = cpu_oracle_decode_turns_per_sec(dm, dshape, 12, 12, 5, threads)
            decode["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                      "sample": "1 dialogue turn of config 5, the whole search (beam 5, max_len 12: %d of %d "
                                                "hypothesis decodes) in %.1f s, fp32 torch CPU, the reference's procedure "
                                                "(whole decoder incl. video blocks per hypothesis per step)" % (done, full, dt)}
        del dm

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, best = cpu_oracle_turns_per_sec(c, 16, 3, threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "16 of the 64 dialogue turns, best of 3 after 1 warm-up (%.2f s each), fp32 torch CPU" % best}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "turns_per_gpu": c["B"], "precision": args.precision,
                           "l2": "inputs larger than L2 (335 MB fts read per step vs 126 MB L2), no explicit flush",
                           "launch": "eager" if args.no_graph else "cuda_graph",
                           "parallelism": "dialogue turns sharded over %d GPU(s), no collective" % world},
                "clocks": clocks,
                "e2e": {"value": world * c["B"] / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "ingest": ingest,
                "gpu_launches": int(launches),
                "roofline": roof,
                "encoder_algorithmic_gbs": whole, "encoder_frac_of_hbm_peak": whole / peaks["hbm"],
                "kernels": kern, "kernel_rooflines": per_kernel,
                "train": train, "train_full": train_full, "train_full_audio": train_full_audio, "decode": decode,
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("BIST_PRECISION", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step measurement of the video path")
    args = ap.parse_args()
    # stdout carries exactly ONE line (the JSON): libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
