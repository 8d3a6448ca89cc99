set -x
timeout 300 python -m pytest tests/test_gpu_gemm.py -x -q 2>&1 | tail -3
timeout 120 python dbg/vidproj_bench.py 2>&1 | tail -5
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err; tail -c 300 gpurun_out/bench_i.err
python - <<'PY'
import json
j=json.loads(open('gpurun_out/bench_i.json').read())
print({k:j.get(k) for k in ('value','ms_per_step')}, j['e2e']['value'], j['ingest'], j['decode'], [k for k in j['kernels'] if 'vidproj' in str(k)][:3])
print(j['kernels'])
PY
