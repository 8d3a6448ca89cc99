# r02c capture (one gpurun call, one GPU): launch list of the bench command in eager mode (the graphed step launches the same
# kernels) and --set full of the large plain GEMM (CTA-pair form, TMA-store epilogue).  Each ncu run follows a plain run of the SAME
# command that exited 0 (&&, no pipe).
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --quick --no-cpu-baseline --no-gpu-eager --no-strong --no-graph"
timeout 300 $CMD > gpurun_out/r02c_plain_bench.json 2> gpurun_out/r02c_plain_bench.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02c_launches_bench_bf16_eager.csv \
  $CMD > gpurun_out/r02c_ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 200 python scripts/gemm_prof.py > gpurun_out/r02c_plain_gemm.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 1 -f -o gpurun_out/r02c_gemm_tc \
  python scripts/gemm_prof.py > gpurun_out/r02c_ncu_gemm.log 2>&1; echo "gemm rc=$?"
