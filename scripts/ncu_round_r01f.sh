# r01f capture: the full BiST training step (config 3) — bench JSON, per-tag kernel times, ncu launch list of ONE eager
# step (cudaProfilerStart/Stop around it) and --set full of the new generator / pointer / gate / embedding kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/pytest_f.log; cat gpurun_out/pytest_f.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; tail -2 gpurun_out/bench_bf16.err
timeout 300 python scripts/core_bench.py --tags > gpurun_out/full_step_kernels_bf16.txt 2>&1; head -2 gpurun_out/full_step_kernels_bf16.txt
timeout 300 python scripts/core_bench.py --graph > gpurun_out/full_step_graph_bf16.txt 2>&1; cat gpurun_out/full_step_graph_bf16.txt
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/launches_r01f_full_step.csv python scripts/core_bench.py --steps 1 --ncu > gpurun_out/ncu_launches_full.log 2>&1
tail -2 gpurun_out/ncu_launches_full.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:"genloss_kernel|ptr_bwd_kernel|ptr_fwd_kernel|gate_bwd_kernel|embed_bwd_kernel" -c 12 \
  -o gpurun_out/prof_r01f python scripts/core_bench.py --steps 1 --ncu > gpurun_out/ncu_full_r01f.log 2>&1
tail -2 gpurun_out/ncu_full_r01f.log
