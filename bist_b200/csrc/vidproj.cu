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


/* Training forward: same result as bist_vidproj_fwd, and the post-ReLU activations h = relu(fts W^T + b) are kept as
 * fp32 [N, d] in `act` (caller-owned) for the backward (the LayerNorm cannot be inverted through a bf16 output). */
int bist_vidproj_train_fwd(const float* fts, const float* W, const float* bias, const float* a_2, const float* b_2,
                           void* out, float* act, int64_t N, int64_t D, int64_t d, float eps, int compute_dtype,
                           int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(fts && W && bias && a_2 && b_2 && out && act, "vidproj_train: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31), "vidproj_train: bad shape");
  BIST_CHECK_ARG(dtype_ok(compute_dtype) && dtype_ok(out_dtype), "vidproj_train: bad dtype");
  GemmArgs g;  // encoder.py:75 — relu(W fts + b); bf16 mode multiplies the fp32 operands as tf32 on the tensor cores
  g.A = fts; g.W = W; g.C = act; g.bias = bias;
  g.M = (int)N; g.N = (int)d; g.K = (int)D;
  g.a = rows_dense(D); g.w = rows_dense(D); g.c = rows_dense(d);
  g.relu = 1; g.tag = "vidproj_gemm";
  g.allow_tf32 = compute_dtype == BIST_BF16;
  BIST_TRY(gemm(g, st));
  return layer_norm(act, a_2, b_2, out, N, d, eps, BIST_F32, out_dtype, st);                                  // :80
}

size_t bist_vidproj_bwd_workspace_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype) {
  (void)D;
  return align_up((size_t)N * d * dtype_size(compute_dtype), 256) + 512;
}

/* Backward of LayerNorm(relu(fts W^T + b)) w.r.t. W, bias, a_2, b_2 (fts is data: no gradient).
 *   d_out fp32 [N, d] (gradient of the projected video tensor, accumulated over blocks and directions);
 *   act = the tape of bist_vidproj_train_fwd; dW [d, D], dbias, da_2, db_2 fp32, accumulated (+=). */
int bist_vidproj_bwd(const float* d_out, const float* fts, const float* act, const float* a_2, float* dW, float* dbias,
                     float* da_2, float* db_2, int64_t N, int64_t D, int64_t d, float eps, void* ws, size_t ws_bytes,
                     int compute_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(d_out && fts && act && a_2 && dW && dbias && da_2 && db_2 && ws, "vidproj_bwd: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31) && dtype_ok(compute_dtype), "vidproj_bwd: bad argument");
  Arena A(ws, ws_bytes);
  void* dh = A.take((size_t)N * d * dtype_size(compute_dtype));
  if (!A.ok) { set_error("vidproj_bwd: workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }
  LnBwdArgs l;  // dh = relu'(h) * dLN(d_out; h)
  l.dy = d_out; l.x = act; l.a2 = a_2; l.dx = dh; l.da = da_2; l.db = db_2;
  l.rows = N; l.d = d; l.eps = eps;
  l.dy_dtype = BIST_F32; l.x_dtype = BIST_F32; l.dx_dtype = compute_dtype; l.relu_gate = 1;
  BIST_TRY(layer_norm_bwd(l, st));
  WgradArgs w;  // dW += dh^T fts ; dbias += colsum(dh)
  w.A = dh; w.B = fts; w.C = dW; w.colsum = dbias;
  w.M = (int)N; w.N = (int)d; w.K = (int)D; w.lda = d; w.ldb = D; w.ldc = D;
  w.a_dtype = compute_dtype; w.b_dtype = BIST_F32; w.tag = "wgrad_vidproj";
  return wgrad(w, st);
}

}  // extern "C"
