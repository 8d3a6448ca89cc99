# ncu launch list of one bench run (eager launches so that every kernel is visible to ncu). Usage: bash scripts/ncu_launches.sh <tag> [skip] [count]
TAG=${1:-r01}; SKIP=${2:-420}; COUNT=${3:-160}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c $COUNT --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_launches_$TAG.log 2>&1
tail -2 gpurun_out/ncu_launches_$TAG.log
