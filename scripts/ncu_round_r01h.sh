# r01h closing capture of round 1: full GPU test suite, smoke, bench (bf16 / fp32 / reference arm), launch list of the
# eager forward bench (overlap on), per-tag times of the full training step.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -2 > gpurun_out/pytest_h.log; cat gpurun_out/pytest_h.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_h.log 2>&1; tail -1 gpurun_out/smoke_h.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; tail -2 gpurun_out/bench_bf16.err
timeout 600 python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python scripts/core_bench.py --tags > gpurun_out/full_step_kernels_h.txt 2>&1
timeout 300 python scripts/core_bench.py --graph >> gpurun_out/full_step_kernels_h.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 150 --csv --log-file gpurun_out/launches_r01h_fwd.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-train > gpurun_out/ncu_launches_fwd_h.log 2>&1
tail -1 gpurun_out/ncu_launches_fwd_h.log
