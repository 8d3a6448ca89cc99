# ncu --set full capture of kernels matching a regex during one eager bench run. Usage: bash scripts/ncu_full.sh <regex> <outname> [skip] [count]
RX=$1; OUT=$2; SKIP=${3:-6}; COUNT=${4:-2}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain_$OUT.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c $COUNT -o gpurun_out/$OUT python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_$OUT.log 2>&1
tail -2 gpurun_out/ncu_$OUT.log
