# r01j capture: state after the CTA-pair projection and the audio encoder.  fp32 bench line, launch list of the eager
# forward bench, --set full of the projection kernel (CTA-pair form) and the tcgen05 GEMM, per-tag times of the full step
# with the audio stream.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > gpurun_out/r01j_bench_fp32.json 2> gpurun_out/r01j_bench_fp32.err; echo "fp32 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 150 --csv --log-file gpurun_out/r01j_launches_forward_bf16.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-train > gpurun_out/r01j_ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vidproj_kernel -s 3 -c 1 -f -o gpurun_out/r01j_vidproj \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-train > gpurun_out/r01j_ncu_vidproj.log 2>&1; echo "vidproj rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 60 -c 12 -f -o gpurun_out/r01j_gemm_tc \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-train > gpurun_out/r01j_ncu_gemm.log 2>&1; echo "gemm rc=$?"
