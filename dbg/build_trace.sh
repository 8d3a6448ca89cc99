set -e
cd /root/repo
mkdir -p dbg/obj
for f in bist_b200/csrc/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DBIST_GEMM_TRACE -c $f -o dbg/obj/$(basename $f .cu).o &
done
wait
nvcc -shared -o dbg/libbist_trace.so dbg/obj/*.o -lcudart
