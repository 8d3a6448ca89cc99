for v in 0 8192 2048; do
  BIST_GEMM_2CTA=$v timeout 600 python bench.py --no-train --no-cpu-baseline --steps 30 > gpurun_out/bench_g$v.json 2> gpurun_out/bench_g$v.err
  python - <<PY
import json
j=json.loads(open('gpurun_out/bench_g$v.json').read())
k=j['kernels']
print($v, round(j['ms_per_step'],4), {n:round(k[n]['ms_per_step'],4) for n in ('xv_gemm','fold_gemm','query_side_gemm','scores1_gemm','vfold2_gemm')})
PY
done
BIST_GEMM_2CTA=2048 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
