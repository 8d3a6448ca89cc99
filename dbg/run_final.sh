set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_bf16_i.json 2> gpurun_out/bench_bf16_i.err; tail -c 200 gpurun_out/bench_bf16_i.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_i.json 2> gpurun_out/bench_ref_i.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01i_fwd.csv python bench.py --no-graph --no-train --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/ncu_i.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:vidproj_kernel -c 1 -s 2 -o gpurun_out/vidproj_pair_full -f python dbg/vidproj_bench.py > gpurun_out/ncu_vp.log 2>&1; tail -2 gpurun_out/ncu_vp.log
