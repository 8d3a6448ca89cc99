// VidEncoder8.forward, video branch — /root/reference/model/encoder.py:72-80:
//   out = LayerNorm(relu(fts . W^T + bias)).
#include <mutex>

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
  else n += 2 * align_up((size_t)N * D * 4, 256) + 2 * align_up((size_t)d * D * 4, 256);   // hi / lo splits of fts and W (gemm_x3.cu)
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
  SplitRef sa, sw;
  if (compute_dtype == BIST_F32) {
    sa.src = fts; sa.n = N * D; sa.hi = (float*)A.take((size_t)N * D * 4); sa.lo = (float*)A.take((size_t)N * D * 4);
    sw.src = W; sw.n = d * D; sw.hi = (float*)A.take((size_t)d * D * 4); sw.lo = (float*)A.take((size_t)d * D * 4);
  }
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
  // fp32 mode: three tf32 products over hi / lo splits of the features and the weight (fp32 accuracy on the tensor cores)
  if (compute_dtype == BIST_F32) BIST_TRY(gemm_fp32x3(g, sa, sw, st));
  else BIST_TRY(gemm(g, st));
  return layer_norm(act, a_2, b_2, out, N, d, eps, BIST_F32, out_dtype, st);                                  // :80
}


/* No-allocation variant of the bf16-mode projection.  It has no counterpart in the reference; it was added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  The one piece of scratch the fused kernel needs -- the bf16 copy of W [d, D] -- lives in a
 * file-scope device array of BIST_VIDPROJ_NOALLOC_MAX_BYTES (2 MiB: d = 512, D = 2048) that every GPU loading the library
 * holds until exit.  In host code `g_vp_noalloc_scratch` is only a stand-in for the symbol: the device address is taken once
 * per device with cudaGetSymbolAddress (that device current) and passed to the kernels.  Every call rewrites the whole
 * [d, D] region it reads, so nothing has to start at zero; all calls share the array, so calls that overlap on the GPU
 * (different streams, or graphs replayed concurrently) race. */
static __device__ __align__(256) unsigned char g_vp_noalloc_scratch[BIST_VIDPROJ_NOALLOC_MAX_BYTES];

namespace {
constexpr int VP_NOALLOC_MAX_DEVICES = 64;
std::mutex g_vp_noalloc_mu;
void* g_vp_noalloc_ptr[VP_NOALLOC_MAX_DEVICES] = {};

int vp_noalloc_scratch(void** out) {
  int dev = -1;
  BIST_CUDA(cudaGetDevice(&dev));
  BIST_CHECK_ARG(dev >= 0 && dev < VP_NOALLOC_MAX_DEVICES, "vidproj_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_vp_noalloc_mu);
  if (!g_vp_noalloc_ptr[dev]) {
    void* ptr = nullptr;
    BIST_CUDA(cudaGetSymbolAddress((void**)&ptr, g_vp_noalloc_scratch));   // this device's instance of the array
    g_vp_noalloc_ptr[dev] = ptr;
  }
  *out = g_vp_noalloc_ptr[dev];
  return 0;
}
}  // namespace

int bist_vidproj_fwd_noalloc(const float* fts, const float* W, const float* bias, const float* a_2, const float* b_2,
                             void* out, int64_t N, int64_t D, int64_t d, float eps, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(fts && W && bias && a_2 && b_2 && out, "vidproj_noalloc: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31), "vidproj_noalloc: bad shape N=%lld D=%lld d=%lld",
                 (long long)N, (long long)D, (long long)d);
  BIST_CHECK_ARG(dtype_ok(out_dtype), "vidproj_noalloc: bad dtype");
  if ((unsigned long long)d * (unsigned long long)D * 2ull > (unsigned long long)BIST_VIDPROJ_NOALLOC_MAX_BYTES) {
    set_error("vidproj_noalloc: d*D*2 = %llu bytes of weight scratch exceed the built-in %d-byte array (limit d*D <= %d); "
              "use bist_vidproj_fwd with a caller-owned workspace", (unsigned long long)d * D * 2ull,
              (int)BIST_VIDPROJ_NOALLOC_MAX_BYTES, (int)(BIST_VIDPROJ_NOALLOC_MAX_BYTES / 2));
    return BIST_ERR_WORKSPACE;
  }
  void* wb = nullptr;
  BIST_TRY(vp_noalloc_scratch(&wb));
  bool taken = false;
  const int rc = vidproj_fused(fts, W, bias, a_2, b_2, out, N, D, d, eps, wb, out_dtype, st, &taken);
  if (!taken) {   // the GEMM + LayerNorm pair needs an [N, d] fp32 buffer: not available without a workspace
    set_error("vidproj_noalloc: shape d=%lld D=%lld is not taken by the fused kernel (needs d %% 64 == 0, d <= 512, D %% 32 == 0, "
              "16-byte aligned pointers); use bist_vidproj_fwd with a caller-owned workspace", (long long)d, (long long)D);
    return BIST_ERR_UNSUPPORTED;
  }
  return rc;
}


/* The same projection for features that are ALREADY bf16 (the ingest cache of bist_b200/ingest.py keeps them in 2 bytes
 * on the host and on the device: half the PCIe and HBM traffic of the fp32 array the reference loads).  Rounding the fp32
 * features to bf16 is exactly what the fused kernel above does in shared memory, so the result is the same; with a 2-byte
 * A operand the plain tcgen05 GEMM (bias + ReLU epilogue, fp32 activations) followed by the LayerNorm kernel is faster than
 * the fused fp32-input kernel (the contraction reads 168 MB instead of 335 MB at config 2).
 * W16: bf16 [d, D]; ws >= N*d*4 bytes. */
int bist_vidproj_fwd_bf16in(const void* fts16, const void* W16, const float* bias, const float* a_2, const float* b_2,
                            void* out, int64_t N, int64_t D, int64_t d, float eps, void* ws, size_t ws_bytes,
                            int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(fts16 && W16 && bias && a_2 && b_2 && out && ws, "vidproj(bf16 in): null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31), "vidproj(bf16 in): bad shape");
  BIST_CHECK_ARG(dtype_ok(out_dtype), "vidproj(bf16 in): bad dtype");
  if (ws_bytes < (size_t)N * d * 4) { set_error("vidproj(bf16 in): workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }
  float* act = (float*)ws;
  GemmArgs g;  // encoder.py:75 — relu(W fts + b)
  g.A = fts16; g.W = W16; g.C = act; g.bias = bias;
  g.M = (int)N; g.N = (int)d; g.K = (int)D;
  g.a = rows_dense(D); g.w = rows_dense(D); g.c = rows_dense(d);
  g.relu = 1; g.a_dtype = BIST_BF16; g.w_dtype = BIST_BF16; g.tag = "vidproj_gemm";
  BIST_TRY(gemm(g, st));
  return layer_norm(act, a_2, b_2, out, N, d, eps, BIST_F32, out_dtype, st);                                  // :80
}


/* Tape of the training forward: h = relu(fts W^T + b) as fp32 [N, d] (the LayerNorm cannot be inverted through a bf16
 * output) and, in bf16 mode, bf16 copies of fts [N, D] and W [d, D]: the forward contraction and the weight gradient
 * dW = dh^T fts both read the 2-byte copy (168 MB instead of 335 MB at config 2) on the tcgen05 kernels. */
namespace {
struct VpTape {
  float* act;
  void* fts16;
  void* w16;
  size_t bytes;
  bool ok;
};
VpTape vp_tape(void* base, size_t cap, int64_t N, int64_t D, int64_t d, int compute_dtype) {
  Arena A(base, cap);
  VpTape t;
  t.act = (float*)A.take((size_t)N * d * 4);
  t.fts16 = compute_dtype == BIST_BF16 ? A.take((size_t)N * D * 2) : nullptr;
  t.w16 = compute_dtype == BIST_BF16 ? A.take((size_t)d * D * 2) : nullptr;
  t.bytes = A.used + 256;
  t.ok = A.ok;
  return t;
}
}  // namespace

size_t bist_vidproj_tape_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype) {
  if (N <= 0 || D <= 0 || d <= 0 || !dtype_ok(compute_dtype)) return 0;
  return vp_tape(nullptr, (size_t)-1, N, D, d, compute_dtype).bytes;
}

int bist_vidproj_train_fwd(const float* fts, const float* W, const float* bias, const float* a_2, const float* b_2,
                           void* out, void* tape, size_t tape_bytes, int64_t N, int64_t D, int64_t d, float eps,
                           int compute_dtype, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(fts && W && bias && a_2 && b_2 && out && tape, "vidproj_train: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31), "vidproj_train: bad shape");
  BIST_CHECK_ARG(dtype_ok(compute_dtype) && dtype_ok(out_dtype), "vidproj_train: bad dtype");
  const VpTape T = vp_tape(tape, tape_bytes, N, D, d, compute_dtype);
  if (!T.ok) { set_error("vidproj_train: tape too small (%zu bytes)", tape_bytes); return BIST_ERR_WORKSPACE; }
  GemmArgs g;  // encoder.py:75 — relu(W fts + b)
  g.A = fts; g.W = W; g.C = T.act; g.bias = bias;
  g.M = (int)N; g.N = (int)d; g.K = (int)D;
  g.a = rows_dense(D); g.w = rows_dense(D); g.c = rows_dense(d);
  g.relu = 1; g.tag = "vidproj_gemm";
  if (compute_dtype == BIST_BF16) {
    if (D % 4 == 0) BIST_TRY(dropout_scale(fts, T.fts16, N, D, DropCfg(), 0, 0, BIST_F32, BIST_BF16, st));  // vectorised cast
    else BIST_TRY(cast_copy(fts, T.fts16, N * D, BIST_F32, BIST_BF16, st));
    BIST_TRY(pack_matrix(W, T.w16, (int)d, (int)D, 0, 1.f, BIST_BF16, st));
    g.A = T.fts16; g.W = T.w16; g.a_dtype = BIST_BF16; g.w_dtype = BIST_BF16;
  }
  BIST_TRY(gemm(g, st));
  return layer_norm(T.act, a_2, b_2, out, N, d, eps, BIST_F32, out_dtype, st);                                // :80
}

size_t bist_vidproj_bwd_workspace_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype) {
  (void)D;
  return align_up((size_t)N * d * dtype_size(compute_dtype), 256) + 512;
}

/* Backward of LayerNorm(relu(fts W^T + b)) w.r.t. W, bias, a_2, b_2 (fts is data: no gradient).
 *   d_out fp32 [N, d] (gradient of the projected video tensor, accumulated over blocks and directions);
 *   tape = what bist_vidproj_train_fwd wrote; dW [d, D], dbias, da_2, db_2 fp32, accumulated (+=). */
int bist_vidproj_bwd(const float* d_out, const float* fts, const void* tape, size_t tape_bytes, const float* a_2, float* dW,
                     float* dbias, float* da_2, float* db_2, int64_t N, int64_t D, int64_t d, float eps, void* ws,
                     size_t ws_bytes, int compute_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(d_out && fts && tape && a_2 && dW && dbias && da_2 && db_2 && ws, "vidproj_bwd: null pointer");
  BIST_CHECK_ARG(N > 0 && D > 0 && d > 1 && N < (1LL << 31) && dtype_ok(compute_dtype), "vidproj_bwd: bad argument");
  const VpTape T = vp_tape(const_cast<void*>(tape), tape_bytes, N, D, d, compute_dtype);
  if (!T.ok) { set_error("vidproj_bwd: tape too small (%zu bytes)", tape_bytes); return BIST_ERR_WORKSPACE; }
  Arena A(ws, ws_bytes);
  void* dh = A.take((size_t)N * d * dtype_size(compute_dtype));
  if (!A.ok) { set_error("vidproj_bwd: workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }
  LnBwdArgs l;  // dh = relu'(h) * dLN(d_out; h)
  l.dy = d_out; l.x = T.act; l.a2 = a_2; l.dx = dh; l.da = da_2; l.db = db_2;
  l.rows = N; l.d = d; l.eps = eps;
  l.dy_dtype = BIST_F32; l.x_dtype = BIST_F32; l.dx_dtype = compute_dtype; l.relu_gate = 1;
  BIST_TRY(layer_norm_bwd(l, st));
  WgradArgs w;  // dW += dh^T fts ; dbias += colsum(dh)
  w.A = dh; w.B = compute_dtype == BIST_BF16 ? T.fts16 : (const void*)fts; w.C = dW; w.colsum = dbias;
  w.M = (int)N; w.N = (int)d; w.K = (int)D; w.lda = d; w.ldb = D; w.ldc = D;
  w.a_dtype = compute_dtype; w.b_dtype = compute_dtype; w.tag = "wgrad_vidproj";
  return wgrad(w, st);
}

}  // extern "C"
