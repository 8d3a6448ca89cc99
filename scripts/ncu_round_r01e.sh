# Round-1 closing capture: tests, smoke, bench (bf16, fp32), ncu launch lists (forward bench, training step) and
# --set full of the new backward kernels.  Outputs in gpurun_out/ (copied to profiles/ by scripts/export_profile.py).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/pytest_e.log; cat gpurun_out/pytest_e.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_e.log 2>&1; tail -1 gpurun_out/smoke_e.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; tail -2 gpurun_out/bench_bf16.err
timeout 600 python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python scripts/train_bench.py > gpurun_out/train_bf16_do.log 2>&1; head -3 gpurun_out/train_bf16_do.log
timeout 300 python scripts/train_bench.py --dropout 0 > gpurun_out/train_bf16.log 2>&1; head -3 gpurun_out/train_bf16.log
# launch lists (cold-cache, serialised: compare shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 150 --csv --log-file gpurun_out/launches_r01e_fwd.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-train > gpurun_out/ncu_launches_fwd.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2100 -c 700 --csv --log-file gpurun_out/launches_r01e_train.csv \
  python scripts/train_bench.py --steps 2 > gpurun_out/ncu_launches_train.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wgrad_tc_kernel|stage1_bwd_mma_kernel|stage2_bwd_mma_kernel|vidproj_kernel" \
  -s 8 -c 8 -o gpurun_out/prof_r01e python scripts/train_bench.py --steps 2 > gpurun_out/ncu_full_r01e.log 2>&1
tail -2 gpurun_out/ncu_launches_train.log gpurun_out/ncu_full_r01e.log
python - <<'PY'
import json
for f in ("gpurun_out/bench_bf16.json","gpurun_out/bench_fp32.json"):
    try: j=json.load(open(f))
    except Exception as e: print(f,"ERR",e); continue
    print(f, "turns/s", round(j["value"]), "ms", round(j["ms_per_step"],3), "e2e", round(j["e2e"]["value"]), "train", j.get("train") and round(j["train"]["value"]), j["roofline"] and (j["roofline"]["kernel"], round(j["roofline"]["frac"],3)))
PY
