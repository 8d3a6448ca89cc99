// VidEncoder8.forward, video branch — /root/reference/model/encoder.py:72-80:
//   out = LayerNorm(relu(fts . W^T + bias)).
#include "common.cuh"

namespace bist {
int vidproj_fused(const float* fts, const float* W, const float* bias, const float* a2, const float* b2, void* out, int64_t N,
                  int64_t D, int64_t d, float eps, void* wb, int out_dtype, cudaStream_t st, bool* taken);  // vidproj_tc.cu
}
using namespace bist;

extern "C" {

size_t bist_vidproj_workspace_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype) {
  size_t n = align_up((size_t)N * d * 4, 256);                        // relu(fts W^T + b), fp32
  if (compute_dtype == BIST_BF16) n += align_up((size_t)d * D * 2, 256);  // bf16 copy of W
  return n + 512;
}

int bist_vidproj_fwd(const float* fts, const float* W, const float* bias, const float* a_2, const float* b_2,
                     void* out, int64_t N, int64_t D, int64_t d, float eps, void* ws, size_t ws_bytes,
                     int compute_dtype, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(fts && W && bias && a_2 && b_2 && out && ws, "vidproj: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31), "vidproj: bad shape N=%lld D=%lld d=%lld",
                 (long long)N, (long long)D, (long long)d);
  BIST_CHECK_ARG(dtype_ok(compute_dtype) && dtype_ok(out_dtype), "vidproj: bad dtype");
  Arena A(ws, ws_bytes);
  float* act = (float*)A.take((size_t)N * d * 4);
  void* wb = compute_dtype == BIST_BF16 ? A.take((size_t)d * D * 2) : nullptr;
  if (!A.ok) { set_error("vidproj: workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }
  if (compute_dtype == BIST_BF16) {
    // bf16 mode: one fused kernel (TMA fp32 -> in-smem bf16 -> tcgen05 -> LayerNorm epilogue), see vidproj_tc.cu
    bool taken = false;
    int rc = vidproj_fused(fts, W, bias, a_2, b_2, out, N, D, d, eps, wb, out_dtype, st, &taken);
    if (taken) return rc;
  }
  // other shapes in bf16 mode: the fp32 features are multiplied as tf32 on the tensor cores (kind::tf32 reads
  // the fp32 bits directly), followed by the LayerNorm kernel.
  const void* w_used = W;
  const int w_dtype = BIST_F32;
  GemmArgs g;  // encoder.py:75 — relu(W fts + b)
  g.A = fts; g.W = w_used; g.C = act; g.bias = bias;
  g.M = (int)N; g.N = (int)d; g.K = (int)D;
  g.a = rows_dense(D); g.w = rows_dense(D); g.c = rows_dense(d);
  g.relu = 1; g.w_dtype = w_dtype; g.tag = "vidproj_gemm";
  g.allow_tf32 = compute_dtype == BIST_BF16;
  BIST_TRY(gemm(g, st));
  return layer_norm(act, a_2, b_2, out, N, d, eps, BIST_F32, out_dtype, st);                                  // :80
}

}  // extern "C"
