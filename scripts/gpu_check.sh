# Round-trip check on the GPU box: unit tests, parity tests, smoke, bench (bf16 + fp32). Outputs in gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_gemm.log; tail -5 gpurun_out/pytest_gemm.log
timeout 300 python scripts/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; cat gpurun_out/gemm_bench.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/pytest.log; tail -12 gpurun_out/pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --precision bf16 > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; tail -3 gpurun_out/bench_bf16.err
timeout 600 python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err
python - <<'PY'
import json
for f in ("gpurun_out/bench_bf16.json","gpurun_out/bench_fp32.json"):
    try: j=json.load(open(f))
    except Exception as e: print(f,"ERR",e); continue
    print(f, "turns/s", round(j["value"]), "ms", round(j["ms_per_step"],3), "e2e", round(j["e2e"]["value"]), j["roofline"])
    for k,v in sorted(j.get("kernels",{}).items(), key=lambda kv:-kv[1]["ms_per_step"]):
        print("   %-18s n=%5.1f  %8.3f ms/step  %9.1f us/launch"%(k,v["launches_per_step"],v["ms_per_step"],v["us_per_launch"]))
PY
