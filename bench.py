#!/usr/bin/env python
"""bench.py — BiST training throughput (BASELINE.json metric: "BiST train dialogue-turns/sec at 1/2/4/8 B200;
s2t/t2s HBM GB/s vs peak") and the encoder-forward / attention-kernel numbers that explain it.

    python bench.py --gpus N --steps K --warmup W [--precision bf16|fp32] [--impl ours|reference] [--verbose]

Headline (`value`, the same at every N so that the driver's scaling efficiency is computed on the step that HAS a
collective): one full BiST AVSD training step — BASELINE config 3 with the audio encoder it names — per GPU B=64
dialogue turns (T=40, S=16, D=2048, d=512, Lq=16, L_his=64, L_cap=32, L_tgt=12, V=4096, 3/3/3/3 blocks): forward with
tape, generator + label-smoothed loss + auto-encoder heads, backward, gradient all-reduce (NCCL, N>1), fused Adam;
dropout on.  One process per GPU, every rank its own 64 turns (weak scaling), time = max over ranks.
  value      turns/s, batch resident in HBM
  e2e        the same step fed from pinned HOST buffers: H2D of the whole batch and D2H of the loss inside the timed region
  roofline   dominant kernel of the step (per-launch CUDA-event times from the library's hooks over the same step)
  encoder_forward   BASELINE config 2 (video encoder forward, B=64): turns/s, e2e, fp32 mode, the fused s2t/t2s stage-1
             kernel's roofline (`attention_roofline`), the reference formulation in PyTorch eager on the same GPU
  cpu_baseline      the CPU oracle (port of the reference formulation) on this box's host cores, same step, bounded sample
`--impl reference` times that CPU oracle's training step alone (the reference is pure Python/PyTorch and /root/reference
does not travel to the GPU box; oracle/bist_oracle.py keeps the reference's formulation op for op).
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
WORKLOAD = ("full BiST AVSD training step (BASELINE config 3 with the audio encoder): per GPU B=64 turns, T=40 S=16 D=2048 d=512 Lq=16 "
            "L_his=64 L_cap=32 L_tgt=12 V=4096, 3 decoder/video/caption/audio blocks; fwd + label-smoothed losses + bwd + all-reduce + Adam")
FWD_WORKLOAD = "BiST video encoder forward (VidEncoder8 + 3x VidEncoderLayer4 s2t+t2s + out norms), B=64 T=40 S=16 D=2048 d_model=512 Lq=16"
METRIC, UNIT = "bist_train_turns_per_sec", "turns/s"
CPU_TRAIN_ROWS, CPU_FWD_ROWS = 8, 16


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


def _oracle_train_setup(model, rows, audio, device="cpu"):
    """Oracle parameters (leaves sharing the model's tying), args and a `rows`-turn batch of config 3."""
    from bist_b200 import synth
    from bist_b200.core import FULL_SHAPES
    c = FULL_SHAPES
    leaves, P = {}, {}
    for k, v in model.state_dict().items():
        if k.endswith(".pe"):
            continue
        key = v.data_ptr()
        if key not in leaves:
            leaves[key] = v.detach().float().to(device).clone().requires_grad_(True)
        P[k] = leaves[key]
    a = model.args
    args = dict(nb_venc_blocks=a.nb_venc_blocks, nb_cenc_blocks=a.nb_cenc_blocks, nb_aenc_blocks=a.nb_aenc_blocks if audio else 0,
                enc_vc_combine=a.enc_vc_combine, include_caption=a.include_caption, dec_st_combine=a.dec_st_combine,
                t2s=a.t2s, s2t=a.s2t, enc_st_combine="none", ptr_ft=a.ptr_ft, auto_encoder=a.auto_encoder, mask_unk=1)
    b = synth.make_full_batch(1, rows, c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                              Ta=c["Ta"] if audio else 0, Da=c["Da"] if audio else 0)
    return leaves, P, args, (b.to(device) if device != "cpu" else b), a


def oracle_train_steps(model, rows, warmup, steps, audio, threads=None, device="cpu"):
    """The oracle's full training step (reference formulation: forward, label-smoothed losses, autograd backward, Adam) on
    `rows` dialogue turns of config 3, fp32.  Returns the list of per-step seconds (after `warmup` untimed steps)."""
    from bist_b200.core import PAD, SMOOTHING
    from oracle import bist_oracle as O
    if threads:
        torch.set_num_threads(threads)
    leaves, P, args, b, a = _oracle_train_setup(model, rows, audio, device)
    opt = torch.optim.Adam(list(leaves.values()), lr=1e-4)
    times = []
    for i in range(warmup + steps):
        if device != "cpu":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        ft = O.mtn_forward(P, args, b, a.att_h, a.nb_blocks)
        O.loss_compute(P, args, ft, b, PAD, SMOOTHING)[0].backward()
        opt.step()
        if device != "cpu":
            torch.cuda.synchronize()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def oracle_forward_steps(c, rows, warmup, steps, threads=None, device="cpu", dtype=torch.float32):
    """The oracle's video-encoder forward (config 2 shape, `rows` turns).  Per-step seconds."""
    from bist_b200 import synth
    from oracle import bist_oracle as O
    if threads:
        torch.set_num_threads(threads)
    P = {k: v.to(device=device, dtype=dtype) for k, v in synth.make_params(0, c["d"], c["D"], c["n_blocks"]).items()}
    b = synth.make_batch(1, rows, c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(1, rows, c["Lq"], c["d"]).to(device=device, dtype=dtype)
    fts, qm, tm = b.fts.to(device=device, dtype=dtype), b.query_mask.to(device), b.temporal_mask.to(device)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            if device != "cpu":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            O.video_path(P, c["h"], fts, eq, qm, tm, c["n_blocks"])
            if device != "cpu":
                torch.cuda.synchronize()
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    return times


def cpu_oracle_decode_turns_per_sec(model, shapes, max_len_sample, max_len, beam, threads):
    """The CPU oracle's beam search (the reference's procedure: every hypothesis of every step re-runs the whole decoder,
    video blocks included) on ONE dialogue turn of config 5."""
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


def baseline_entry(times, rows, cores, sample):
    "Same statistic everywhere: mean over the timed steps."
    mean = sum(times) / len(times)
    return {"value": rows / mean, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_step": 1e3 * mean,
            "best_ms": 1e3 * min(times), "sample": sample}


def kernel_model(c, e_act, nd=2):
    """ALGORITHMIC (flops, bytes) per step, summed over the launches of a kernel tag.
    Bytes count only what the formulation must move: operands read once, results written once — no intermediates."""
    B, T, S, D, d, h, Lq, L = c["B"], c["T"], c["S"], c["D"], c["d"], c["h"], c["Lq"], c["n_blocks"]
    N, Q = B * T * S, B * Lq
    c1 = Q * (S + T) * d                      # stage-1 summaries, both directions
    return {
        # stage 1 of both directions as ONE kernel: vft read once (N*d*e per block), folded queries + value
        # weights, summaries written once; flops = [qf ; W_v] product on every row of both directions + P.V
        "stage1_fused": (L * (nd * 2.0 * N * d * (d + h * Lq) + nd * 2.0 * N * Lq * d),
                         L * (N * d * e_act + nd * (Q * h * d + d * d) * e_act + c1 * e_act)),
        "vidproj_fused": (2.0 * N * D * d, N * D * 4 + d * D * 2 + N * d * e_act),
        "vidproj_gemm": (2.0 * N * D * d, N * D * e_act + d * D * e_act + N * d * 4),
        "wgrad_vidproj": (2.0 * N * D * d, N * D * e_act + N * d * e_act + d * D * 4),
        "xv_gemm": (L * 2.0 * N * d * nd * d, L * (N * d * e_act + nd * d * d * e_act + N * nd * d * e_act)),
        "wgrad_xv": (L * 2.0 * N * d * nd * d, L * (N * d * e_act + N * nd * d * e_act + nd * d * d * 4)),
        # one K-concatenated launch per block: both directions' value gradients and weights read once, d_vft (fp32) read and written once
        "bwd_xv_gemm": (L * nd * 2.0 * N * d * d, L * (nd * (N * d * e_act + d * d * e_act) + 2 * N * d * 4)),
        "stage2_mma": (L * 4.0 * c1 * h, L * (c1 * e_act + 2 * nd * Q * h * d * e_act)),
    }


def roofline_for(tag, kern, km, peaks, traffic):
    """Roofline entry of one kernel tag: per-launch ALGORITHMIC work / per-launch CUDA-event time; `traffic` = measured
    dram__bytes per launch from the committed ncu capture (profiles/roofline_traffic.json), reported separately."""
    fl, by = km[tag]
    n = kern[tag]["launches_per_step"]
    sec = kern[tag]["us_per_launch"] * 1e-6
    fl, by = fl / n, by / n
    ridge = peaks["tc_sust"] * 1e12 / (peaks["hbm"] * 1e9)
    hbm_gbs, tflops = by / sec / 1e9, fl / sec / 1e12
    if by > 0 and fl / by > ridge:
        r = {"bound": "tensor", "achieved": tflops, "peak": peaks["tc_sust"], "unit": "TFLOP/s", "frac": tflops / peaks["tc_sust"]}
    else:
        r = {"bound": "hbm", "achieved": hbm_gbs, "peak": peaks["hbm"], "unit": "GB/s", "frac": hbm_gbs / peaks["hbm"]}
    r.update({"traffic": traffic, "kernel": tag, "us_per_launch": kern[tag]["us_per_launch"],
              "algorithmic_bytes_per_launch": by, "algorithmic_flops_per_launch": fl,
              "hbm_gbs": hbm_gbs, "hbm_frac": hbm_gbs / peaks["hbm"], "tflops": tflops, "tensor_frac": tflops / peaks["tc_sust"],
              "peak_source": peaks["src"] + " (MEASURED_PEAKS.json: copy bandwidth / sustained bf16)"})
    return r


def collect_kernels(L, steps):
    import ctypes
    from bist_b200 import _lib
    buf = ctypes.create_string_buffer(1 << 16)
    _lib.check(L.bist_prof_collect(buf, len(buf)), "bist_prof_collect")
    kern = {}
    for ln in buf.value.decode().strip().splitlines():
        tag, n, tot = ln.split(",")
        kern[tag] = {"launches_per_step": int(n) / steps, "ms_per_step": float(tot) / steps, "us_per_launch": 1e3 * float(tot) / int(n)}
    return kern


def run_reference(args):
    """The reference's own CPU implementation of the step (oracle port: reference formulation op for op) on this box's host
    cores: the same metric on a bounded sample (CPU_TRAIN_ROWS dialogue turns of config 3 with audio per step)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    from bist_b200.core import full_model
    torch.manual_seed(0)
    model, _ = full_model("fp32", audio=True)
    rows = CPU_TRAIN_ROWS
    times = oracle_train_steps(model, rows, args.warmup, args.steps, True, threads)
    ent = baseline_entry(times, rows, threads, "%d dialogue turns of config 3 (with audio) per step: forward + label-smoothed losses + "
                         "autograd backward + Adam, fp32 torch CPU, %d threads, mean of %d steps" % (rows, threads, len(times)))
    line = {"impl": "reference", "metric": METRIC, "value": ent["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ent["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "rows_per_step": rows},
            "cpu_baseline": ent,
            "e2e": {"value": ent["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def measure_encoder_forward(args, env):
    """BASELINE config 2: video encoder forward (a1 + 3 x a2 + a10), device-resident and end to end, plus the fused
    s2t/t2s stage-1 kernel's roofline, fp32 mode and the reference formulation in PyTorch eager on the same GPU."""
    from bist_b200 import PipelinedVideoPath, VideoPath, _lib, synth
    L, dev, rank, world, timed = env["L"], env["dev"], env["rank"], env["world"], env["timed"]
    c = CFG
    e_act = 2 if args.precision == "bf16" else 4
    model = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision=args.precision).to(dev)
    model.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
    hb = synth.make_batch(1 + rank, c["B"], c["T"], c["S"], c["D"], c["Lq"])
    heq = synth.make_encoded_query(1 + rank, c["B"], c["Lq"], c["d"]).pin_memory()
    hb.fts, hb.query_mask, hb.temporal_mask = hb.fts.pin_memory(), hb.query_mask.pin_memory(), hb.temporal_mask.pin_memory()
    db, deq = hb.to(dev), heq.to(dev)
    out = {"workload": FWD_WORKLOAD, "precision": args.precision}
    with torch.no_grad():
        n0 = L.bist_launch_count()
        model(db, deq)
        torch.cuda.synchronize()
        out["gpu_launches_per_step"] = int(L.bist_launch_count() - n0)
        for _ in range(max(args.warmup, 3)):
            model(db, deq)
        if args.no_graph:
            step = lambda: model(db, deq)
        else:
            graphed = model.capture(db, deq)
            step = graphed.replay
            for _ in range(max(args.warmup, 3)):
                step()
        ms_step = timed(step, args.steps) / args.steps
        out.update({"value": world * c["B"] / (ms_step * 1e-3), "unit": UNIT, "ms_per_step": ms_step,
                    "launch": "eager" if args.no_graph else "cuda_graph"})
        # ---- end to end: pinned host inputs -> H2D -> forward -> D2H of the encoder outputs (double-buffered graphs) ----
        outs = [(torch.empty((c["B"], c["Lq"], c["d"]), dtype=torch.float32).pin_memory(),
                 torch.empty((c["B"], c["Lq"], c["d"]), dtype=torch.float32).pin_memory()) for _ in range(2)]
        pipe = None if args.no_graph else PipelinedVideoPath(model, db, deq, depth=2)
        n_e2e = [0]

        def e2e_step():
            o_t, o_s = outs[n_e2e[0] % 2]
            n_e2e[0] += 1
            if pipe is None:
                ft = model(hb.to(dev, non_blocking=True), heq.to(dev, non_blocking=True))
                o_t.copy_(ft["temporal_ft"], non_blocking=True)
                o_s.copy_(ft["spatial_ft"], non_blocking=True)
            else:
                pipe.submit(hb, heq, o_t, o_s)

        for _ in range(2):
            e2e_step()
        e2e_ms = timed(e2e_step, args.steps) / args.steps
        out["e2e"] = {"value": world * c["B"] / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                      "h2d_bytes_per_step": hb.fts.numel() * 4 + hb.query_mask.numel() + hb.temporal_mask.numel() + heq.numel() * 4,
                      "d2h_bytes_per_step": 2 * outs[0][0].numel() * 4}
        # ---- the same step fed by the bf16 ingest cache -------------------------------------------
        if pipe is not None and args.precision == "bf16":
            out["ingest"] = measure_ingest(args, env, model, db, deq, hb, heq, outs)
        del pipe
        # ---- per-kernel CUDA-event times (library hooks), separate eager pass ----------------------------------------------
        torch.cuda.synchronize()
        L.bist_prof_enable(1)
        prof_steps = min(args.steps, 5)
        for _ in range(prof_steps):
            torch.cuda._sleep(6_000_000)      # the host enqueues the whole step before it runs: no launch gaps in the event times
            model(db, deq)
        kern = collect_kernels(L, prof_steps)
        L.bist_prof_enable(0)
    peaks = env["peaks"]
    km = kernel_model(c, e_act)
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    traffic_db = json.load(open(tp)).get(args.precision, {}) if os.path.exists(tp) else {}
    total_k = sum(v["ms_per_step"] for v in kern.values()) or 1.0
    roofs = {}
    for tag in kern:
        if tag in km:
            roofs[tag] = roofline_for(tag, kern, km, peaks, traffic_db.get(tag))
            roofs[tag]["share_of_step"] = kern[tag]["ms_per_step"] / total_k
    out["attention_roofline"] = roofs.get("stage1_fused")
    out["projection_roofline"] = roofs.get("vidproj_fused") or roofs.get("vidproj_gemm")
    whole = algorithmic_bytes(c, e_act) / (ms_step * 1e-3) / 1e9
    out["encoder_algorithmic_gbs"], out["encoder_frac_of_hbm_peak"] = whole, whole / peaks["hbm"]
    out["kernels_by_tag_ms"] = {k: round(v["ms_per_step"], 4) for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["ms_per_step"])}
    if args.verbose:
        out["kernels"], out["kernel_rooflines"] = kern, roofs
    # ---- the other precision, briefly (bf16 line carries an fp32 object and vice versa) --------------------------------
    other = "fp32" if args.precision == "bf16" else "bf16"
    try:
        with torch.no_grad():
            m2 = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision=other).to(dev)
            m2.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
            for _ in range(3):
                m2(db, deq)
            g2 = m2.capture(db, deq)
            for _ in range(3):
                g2.replay()
            n2 = max(3, min(args.steps, 10))
            ms2 = timed(g2.replay, n2) / n2
        out[other] = {"value": world * c["B"] / (ms2 * 1e-3), "unit": UNIT, "ms_per_step": ms2}
        del m2, g2
    except Exception as ex:   # noqa: BLE001
        out[other] = {"error": str(ex)[:200]}
    # ---- GPU-side bar: the reference formulation (oracle port) in PyTorch eager on this GPU
    if rank == 0 and not args.no_gpu_eager:
        eager = {}
        for name, dt in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
            try:
                ts = oracle_forward_steps(c, c["B"], 10, 20, device=dev, dtype=dt)
                eager[name] = {"value": c["B"] / (sum(ts) / len(ts)), "unit": UNIT, "ms_per_step": 1e3 * sum(ts) / len(ts),
                               "best_ms": 1e3 * min(ts)}
            except Exception as ex:   # noqa: BLE001
                eager[name] = {"error": str(ex)[:200]}
            torch.cuda.empty_cache()
        eager["what"] = "oracle/bist_oracle.py video_path (reference formulation: expanded queries, K/V projections of every row, " \
                        "materialised scores) on cuda, all 64 turns, 10 warm-up + 20 timed steps (mean and best), one GPU"
        out["gpu_eager_baseline"] = eager
    del model
    return out


def measure_ingest(args, env, model, db, deq, hb, heq, outs):
    import numpy as np
    from bist_b200 import PipelinedVideoPath
    from bist_b200.ingest import FeatureCache
    dev, rank, world, timed = env["dev"], env["rank"], env["world"], env["timed"]
    c = CFG
    rs = np.random.RandomState(100 + rank)
    n_vid = 96
    lens = rs.randint(c["T"] // 2, c["T"] + 1, size=n_vid)
    cache = FeatureCache(c["D"], c["S"], int(lens.sum()), dtype=torch.bfloat16)
    picks = [[int(v) for v in rs.permutation(n_vid)[:c["B"]]] for _ in range(8)]
    for v in range(n_vid):
        cache.add(v, rs.uniform(0, 1, size=(int(lens[v]), c["S"], c["D"])).astype(np.float32))
    db16 = db.to(dev)
    db16.fts = db.fts.to(torch.bfloat16)
    pipe16 = PipelinedVideoPath(model, db16, deq, depth=2)
    n, nbytes = [0], [0]

    def ingest_step():
        vids = picks[n[0] % len(picks)]
        o_t, o_s = outs[n[0] % 2]
        n[0] += 1
        nbytes[0] = cache.host_bytes(vids)

        def fill(g):
            cache.collate(vids, dev, out=(g.static_b.fts, g.static_b.temporal_mask), T=c["T"])
            g.static_b.query_mask.copy_(hb.query_mask, non_blocking=True)
            g.static_eq.copy_(heq, non_blocking=True)
        pipe16.submit(None, None, o_t, o_s, fill=fill)

    for _ in range(2):
        ingest_step()
    ms = timed(ingest_step, args.steps) / args.steps
    r = {"metric": "e2e_from_bf16_feature_cache", "value": world * c["B"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
         "h2d_bytes_per_step": int(nbytes[0]) + hb.query_mask.numel() + heq.numel() * 4, "d2h_bytes_per_step": 2 * outs[0][0].numel() * 4}
    del pipe16, cache
    return r


def measure_train(args, env, mod, loss_fn, what, static_batch=None, host_batch=None, profile=False):
    """One training step = forward with tape + loss + backward (ONE CUDA graph: bist_b200/graph.py) + gradient all-reduce
    (NCCL, N > 1) + fused Adam.  With `host_batch`, also the end-to-end step: the whole batch crosses PCIe from pinned host
    memory into the static device batch and the loss comes back, inside the timed region."""
    from bist_b200.dist import allreduce_grads, overlap_grad_allreduce, set_grad_comm_dtype, sync_token_counts
    from bist_b200.graph import GraphedTrainStep
    L, dev, world, timed = env["L"], env["dev"], env["world"], env["timed"]
    # bf16 mode exchanges bf16 gradients (fp32 accumulation inside every rank; BIST_GRAD_COMM=fp32 keeps the wire in fp32)
    wire_bf16 = args.precision == "bf16" and os.environ.get("BIST_GRAD_COMM", "bf16") == "bf16"
    set_grad_comm_dtype(torch.bfloat16 if wire_bf16 else torch.float32)
    B = env["rows_per_gpu"]
    plist = list(mod.parameters())
    # Adam (what the reference's NoamOpt wraps) as ONE launch over all parameters (bist_adam_step); BIST_TORCH_ADAM=1: torch's fused Adam
    if os.environ.get("BIST_TORCH_ADAM"):
        opt = torch.optim.Adam(plist, lr=1e-4, fused=True)
    else:
        from bist_b200.optimize import FusedAdam
        opt = FusedAdam(plist, lr=1e-4)
    ncoll = [0]
    has_counts = static_batch is not None and hasattr(static_batch, "ntokens")
    if has_counts:
        sync_token_counts(static_batch)          # global token counts / world: the N-GPU step is the 1-GPU step on the concatenated batch
    n0 = L.bist_launch_count()
    if args.no_graph:
        gstep = None

        def train_step():
            opt.zero_grad(set_to_none=True)
            if has_counts and world > 1:
                sync_token_counts(static_batch)
            with overlap_grad_allreduce():
                loss = loss_fn()
                loss.backward()
            ncoll[0] = allreduce_grads(plist)
            opt.step()
            return loss
        for _ in range(3):
            train_step()
        per_step = (L.bist_launch_count() - n0) // 3
    else:
        in_graph = world > 1 and os.environ.get("BIST_GRAPH_ALLREDUCE", "1") == "1"
        gstep = GraphedTrainStep(loss_fn, plist, warmup=3, allreduce=in_graph)
        per_step = (L.bist_launch_count() - n0) // 4   # 3 warm-up steps + the captured one

        def train_step():
            if has_counts and world > 1:
                sync_token_counts(static_batch)
            loss = gstep()
            ncoll[0] = gstep.collectives if in_graph else allreduce_grads(plist)
            opt.step()
            return loss
        for _ in range(3):
            train_step()
    tsteps = max(3, min(args.steps, 10))
    sampler = ClockSampler(env["local"]) if profile else None
    tms = timed(train_step, tsteps) / tsteps
    clocks = sampler.stop() if sampler else None
    r = {"value": world * B / (tms * 1e-3), "unit": UNIT, "ms_per_step": tms, "steps": tsteps,
         "gpu_launches_per_step": int(per_step), "collectives_per_step": ncoll[0],
         "launch": "eager" if args.no_graph else ("cuda_graph(fwd+loss+bwd%s) + one-launch Adam" % (" + all-reduce" if world > 1 and gstep and gstep.collectives else "")),
         "what": what}
    if world > 1:
        r["grad_comm_dtype"] = "bf16" if wire_bf16 else "fp32"
    if clocks:
        r["clocks"] = clocks
    if host_batch is not None:
        # End to end: every step's batch crosses PCIe from pinned host memory (copy stream) into a device staging batch while the
        # previous step computes; the step then starts with a device-to-device copy staging -> static batch (the captured graph
        # reads the static one).  The loss comes back to pinned host memory.  Like the reference's loader + batch.move_to_cuda()
        # (train.py:30), but prefetched one step ahead.
        names = [k for k, v in host_batch.__dict__.items() if torch.is_tensor(v)]
        stage = {k: torch.empty_like(getattr(static_batch, k)) for k in names}
        loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        h2d = sum(getattr(host_batch, k).numel() * getattr(host_batch, k).element_size() for k in names)
        cs = torch.cuda.Stream(device=dev)
        landed, consumed = torch.cuda.Event(), torch.cuda.Event()

        def prefetch():
            with torch.cuda.stream(cs):
                cs.wait_event(consumed)               # the staging batch has been copied out
                for k in names:
                    stage[k].copy_(getattr(host_batch, k), non_blocking=True)
                landed.record(cs)

        consumed.record(torch.cuda.current_stream(dev))
        prefetch()

        def e2e_step():
            main = torch.cuda.current_stream(dev)
            main.wait_event(landed)
            for k in names:
                getattr(static_batch, k).copy_(stage[k], non_blocking=True)
            consumed.record(main)
            prefetch()                                # next step's batch crosses PCIe while this step computes
            loss = train_step()
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        for _ in range(2):
            e2e_step()
        ems = timed(e2e_step, tsteps) / tsteps
        torch.cuda.synchronize()
        r["e2e"] = {"value": world * B / (ems * 1e-3), "unit": UNIT, "ms_per_step": ems, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": 4, "feed": "pinned host batch -> device staging batch on a copy stream, one step ahead"}
        del stage
    if profile:
        # per-kernel CUDA-event times of the same step, eager (the hooks time launches, not graph nodes)
        L.bist_prof_enable(1)
        psteps = 2
        for _ in range(psteps):
            torch.cuda._sleep(20_000_000)
            for p in plist:
                p.grad = None
            loss_fn().backward()
        torch.cuda.synchronize()
        r["_kern"] = collect_kernels(L, psteps)
        L.bist_prof_enable(0)
        for p in plist:
            p.grad = None
    if gstep is not None:
        gstep.close()
        del gstep
    return r


def pin_batch(b):
    for k, v in b.__dict__.items():
        if torch.is_tensor(v):
            setattr(b, k, v.contiguous().pin_memory() if v.dim() > 0 else v.clone().pin_memory())
    return b


def measure_training(args, env):
    from bist_b200 import VideoPath, synth
    from bist_b200.core import FULL_SHAPES, full_model
    dev, rank, world = env["dev"], env["rank"], env["world"]
    c = CFG
    out = {}
    fc = FULL_SHAPES

    def host_full_batch(seed, B, audio):
        return pin_batch(synth.make_full_batch(seed, B, fc["T"], fc["S"], fc["D"], fc["Lq"], fc["Lh"], fc["Lc"], fc["Lt"], fc["V"],
                                               Ta=fc["Ta"] if audio else 0, Da=fc["Da"] if audio else 0))

    # ---- headline: config 3 with the audio encoder -------------------------------------------------------------------------
    am, alc = full_model(args.precision, audio=True)
    am = am.to(dev).train()
    hab = host_full_batch(1 + rank, c["B"], True)
    ab = hab.to(dev)
    env["rows_per_gpu"] = c["B"]
    head = measure_train(args, env, am, lambda: alc.loss(am(ab), ab)[0],
                         "BASELINE config 3 with its audio stream: embeddings + positional encoding, text-encoder norms, VidEncoder8 (video "
                         "and audio projections), MultimodalDecoder8 (3 video / caption / audio / decoder blocks, softmax gate), pointer "
                         "generator, 4 auto-encoder heads, label smoothing 0.1, backward, all-reduce, fused Adam; train() mode, dropout 0.1",
                         static_batch=ab, host_batch=hab, profile=True)
    out["headline"] = head
    if rank == 0 and not args.no_gpu_eager:
        try:   # GPU-side bar for the step: the reference formulation (oracle) in PyTorch eager + autograd + Adam on this GPU, all 64 turns
            ts = oracle_train_steps(am, c["B"], 3, 5, True, device=dev)
            head["gpu_eager_baseline"] = {"value": c["B"] / (sum(ts) / len(ts)), "unit": UNIT, "ms_per_step": 1e3 * sum(ts) / len(ts),
                                          "best_ms": 1e3 * min(ts), "what": "oracle mtn_forward + loss_compute + autograd backward + Adam "
                                          "on cuda, fp32, 64 turns, 3 warm-up + 5 timed steps, one GPU"}
        except Exception as ex:   # noqa: BLE001
            head["gpu_eager_baseline"] = {"error": str(ex)[:200]}
        torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        ts = oracle_train_steps(am, CPU_TRAIN_ROWS, 1, 3, True, threads)
        out["cpu_baseline"] = baseline_entry(ts, CPU_TRAIN_ROWS, threads, "%d dialogue turns of config 3 (with audio) per step: forward + "
                                             "label-smoothed losses + autograd backward + Adam, fp32 torch CPU, mean of %d steps after 1 warm-up"
                                             % (CPU_TRAIN_ROWS, len(ts)))
    # ---- strong scaling: global B = 512 dialogue turns over the N GPUs -----------------------------
    if not args.no_strong:
        Bs = 512 // world
        hsb = host_full_batch(100 + rank, Bs, True)
        sb = hsb.to(dev)
        env["rows_per_gpu"] = Bs
        st = measure_train(args, env, am, lambda: alc.loss(am(sb), sb)[0], "same step, global batch 512 dialogue turns split over the GPUs", static_batch=sb)
        st["global_batch"] = 512
        st.pop("what", None)
        out["train_strong"] = st
        del sb, hsb
        env["rows_per_gpu"] = c["B"]
    del am, ab, hab
    torch.cuda.empty_cache()
    if not args.quick:
        # ---- config 3 without the audio encoder (exec.sh's default nb_aenc_blocks = 0) ---------------------------------------
        fm, flc = full_model(args.precision)
        fm = fm.to(dev).train()
        fb = host_full_batch(1 + rank, c["B"], False).to(dev)
        out["train_full"] = measure_train(args, env, fm, lambda: flc.loss(fm(fb), fb)[0], "config 3 without the audio encoder", static_batch=fb)
        del fm, fb
        # ---- the video path alone (config-2 sizes) --------------------------------------------------------------------------
        tm = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.1, precision=args.precision).to(dev)
        tm.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
        tm.train()
        vb = synth.make_batch(1 + rank, c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(dev)
        veq = synth.make_encoded_query(1 + rank, c["B"], c["Lq"], c["d"]).to(dev)
        tgt = torch.randn(c["B"], c["Lq"], c["d"], device=dev)

        def video_loss():
            o = tm(vb, veq)
            return ((o["temporal_ft"] - tgt) ** 2).mean() + ((o["spatial_ft"] - tgt) ** 2).mean()
        out["train_video_path"] = measure_train(args, env, tm, video_loss, "video path only (projection + 3 blocks + out norms), MSE loss")
        del tm, vb, veq
    torch.cuda.empty_cache()
    return out


def measure_decode(args, env):
    "BASELINE config 5: generate.py-style beam search (beam 5, max_len 12) for 8 turns per GPU over long videos (T=128, S=49)."
    from bist_b200.core import full_batch, full_model
    from bist_b200.decode import beam_search_decode_batch
    dev, rank, world, timed = env["dev"], env["rank"], env["world"], env["timed"]
    dshape = dict(B=8, T=128, S=49)
    dm, _ = full_model(args.precision, dshape)
    dm = dm.to(dev).eval()
    dbatch = full_batch(5 + rank, dev, dshape)

    def decode_step():
        beam_search_decode_batch(dm, dbatch, 12, 2, 0, 3, 1, beam=5, penalty=1.0, nbest=5, min_len=1)
    for _ in range(2):
        decode_step()
    dms = timed(decode_step, 3) / 3
    r = {"metric": "beam_search_turns_per_sec", "value": world * dshape["B"] / (dms * 1e-3), "unit": UNIT, "ms_per_batch": dms,
         "workload": "config 5: B=8 turns/GPU, T=128, S=49, beam 5, max_len 12; video/caption blocks once per turn, all hypotheses batched"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, dt, done, full = cpu_oracle_decode_turns_per_sec(dm, dshape, 12, 12, 5, threads)
        r["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "1 dialogue turn, whole search (%d hypothesis decodes) in %.1f s, fp32 torch CPU, reference procedure" % (done, dt)}
    del dm
    return r


def run_ours(args):
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from bist_b200 import _lib
    L = _lib.lib()
    _lib.check(L.bist_device_check(), "bist_device_check")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        "K steps between barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks (ms)."
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

    peaks = load_peaks()
    env = {"L": L, "dev": dev, "rank": rank, "world": world, "local": local, "timed": timed, "peaks": peaks, "rows_per_gpu": CFG["B"]}
    e_act = 2 if args.precision == "bf16" else 4
    n_launch0 = L.bist_launch_count()
    tr = measure_training(args, env)
    head = tr["headline"]
    kern = head.pop("_kern", {})
    # dominant kernel of the step: the single kernel (not an aggregate of many small launches) with the largest time per step
    km = kernel_model(CFG, e_act)
    total_k = sum(v["ms_per_step"] for v in kern.values()) or 1.0
    cand = [t for t in kern if t in km and kern[t]["launches_per_step"] <= 2 * CFG["n_blocks"]]
    roof = None
    if cand:
        top = max(cand, key=lambda t: kern[t]["ms_per_step"])
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        traffic_db = json.load(open(tp)).get(args.precision, {}) if os.path.exists(tp) else {}
        roof = roofline_for(top, kern, km, peaks, traffic_db.get(top))
        roof["share_of_step"] = kern[top]["ms_per_step"] / total_k
    enc = measure_encoder_forward(args, env) if not args.no_forward else None
    decode = measure_decode(args, env) if not (args.quick or args.no_forward) else None
    tgif = None
    if not (args.quick or args.no_forward):
        try:
            from bist_b200.tgifqa import bench_leg
            tgif = bench_leg(args, env)
        except ImportError:
            tgif = None
    if rank == 0:
        line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": head["steps"],
                "warmup": 3, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "turns_per_gpu": CFG["B"], "precision": args.precision,
                           "l2": "inputs larger than L2 (335 MB of features + 0.6 GB of tape per step vs 126 MB L2), no explicit flush",
                           "launch": head["launch"], "parallelism": "dp%d: dialogue turns sharded over the GPUs, one gradient all-reduce per step" % world},
                "clocks": head.get("clocks"),
                "e2e": head.get("e2e"),
                "gpu_launches": int(head["gpu_launches_per_step"] * head["steps"]),
                "collectives_per_step": head["collectives_per_step"],
                "roofline": roof,
                "attention_roofline": enc.get("attention_roofline") if enc else None,
                "cpu_baseline": tr.get("cpu_baseline"),
                "gpu_eager_baseline": head.get("gpu_eager_baseline"),
                "train_strong": tr.get("train_strong"),
                "train_full": tr.get("train_full"), "train_video_path": tr.get("train_video_path"),
                "encoder_forward": enc, "decode": decode, "tgifqa": tgif,
                "train_kernels_by_tag_ms": {k: round(v["ms_per_step"], 4) for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["ms_per_step"])[:12]},
                "library_launches_total": int(L.bist_launch_count() - n_launch0)}
        if args.verbose:
            line["train_kernels"] = kern
            line["what"] = head["what"]
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.cuda.synchronize()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("BIST_PRECISION", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-eager", action="store_true", help="skip the PyTorch-eager legs of the reference formulation on the GPU")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-forward", action="store_true", help="skip the encoder-forward / decode legs")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (global batch 512)")
    ap.add_argument("--quick", action="store_true", help="headline + encoder forward only")
    ap.add_argument("--verbose", action="store_true", help="add the per-kernel tables to the JSON line")
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
