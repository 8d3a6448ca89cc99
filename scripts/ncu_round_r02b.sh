# r02b capture (one gpurun call, one GPU) after the query-side fusion: launch list of three eager config-2 forwards, and --set full of
# one launch each of qside_attn / qside_fold / qside_ffn1 and the projection kernel.  Each ncu run follows a plain run of the SAME
# command that exited 0 (&&, no pipe).
mkdir -p gpurun_out
timeout 200 python scripts/fwd_prof.py 3 > gpurun_out/r02b_plain_fwd.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r02b_launches_forward_bf16.csv \
  python scripts/fwd_prof.py 3 > gpurun_out/r02b_ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 200 python scripts/fwd_prof.py 2 > gpurun_out/r02b_plain_fwd2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"qside_kernel|vidproj_kernel" -s 4 -c 4 -f -o gpurun_out/r02b_qside \
  python scripts/fwd_prof.py 2 > gpurun_out/r02b_ncu_qside.log 2>&1; echo "qside rc=$?"
