// tcgen05 / TMEM / TMA GEMM for sm_100a (bf16 operands, fp32 accumulate).
// Placeholder dispatcher: until the tensor-core kernel lands every GEMM takes the CUDA-core path.
#include "common.cuh"

namespace bist {
int gemm_tc(const GemmArgs& g, cudaStream_t st, bool* taken) {
  (void)g; (void)st;
  *taken = false;
  return 0;
}
}  // namespace bist
