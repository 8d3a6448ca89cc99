# r02 capture (one gpurun call, one GPU): launch list of the bench command in eager mode (the graphed step launches the same
# kernels) and --set full of the fused stage-1 kernel inside the encoder forward.  Each ncu run follows a plain run of the SAME
# command that exited 0 (&&, no pipe).
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --quick --no-cpu-baseline --no-gpu-eager --no-strong --no-graph"
timeout 300 $CMD > gpurun_out/r02_plain_bench.json 2> gpurun_out/r02_plain_bench.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02_launches_bench_bf16_eager.csv \
  $CMD > gpurun_out/r02_ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 200 python scripts/stage1_prof.py > gpurun_out/r02_plain_stage1.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stage1_fused -s 1 -c 1 -f -o gpurun_out/r02_stage1_fused \
  python scripts/stage1_prof.py > gpurun_out/r02_ncu_stage1.log 2>&1; echo "stage1 rc=$?"
