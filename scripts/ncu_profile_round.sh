# One-stop profile capture for profiles/: launch list of a full eager bench run + --set full of the heavy kernels.
# Usage: bash scripts/ncu_profile_round.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 150 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_launches_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vidproj_kernel|softmax_pv_mma|stage2_mma|gemm_tc_kernel<256" -s 12 -c 12 -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/ncu_launches_$TAG.log gpurun_out/ncu_full_$TAG.log
