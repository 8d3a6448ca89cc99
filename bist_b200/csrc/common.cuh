// Internal helpers shared by the kernels of libbist_b200.so (not part of the ABI).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/bist_b200.h"

namespace bist {

// ---- error plumbing (api.cu owns the storage) -------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
// Optional per-launch CUDA-event timing (bist_prof_enable): declare one right before a launch.
struct ProfScope {
  void* rec;
  cudaStream_t st;
  ProfScope(const char* tag, cudaStream_t st);
  ~ProfScope();
};

#define BIST_CHECK_ARG(cond, ...)                 \
  do {                                            \
    if (!(cond)) {                                \
      ::bist::set_error(__VA_ARGS__);             \
      return BIST_ERR_ARG;                        \
    }                                             \
  } while (0)

#define BIST_CUDA(call)                                                                   \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      ::bist::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return BIST_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

#define BIST_TRY(call)        \
  do {                        \
    int rc__ = (call);        \
    if (rc__ != 0) return rc__; \
  } while (0)

// Check the launch that just happened (no sync) and count it.
#define BIST_LAUNCHED()                     \
  do {                                      \
    ::bist::count_launch();                 \
    BIST_CUDA(cudaGetLastError());          \
  } while (0)

// ---- launches: programmatic dependent launch (PDL) ------------------------------------------------------------
// Every kernel of the library is launched with programmatic stream serialization allowed: it calls
// pdl_trigger() first (the next kernel in the stream may start launching once all CTAs of this one have
// started) and pdl_wait() before its first access to global memory (blocks until the previous kernel has
// completed and flushed).  Prologues (barrier init, TMEM allocation, smem carve-up) therefore overlap the tail
// of the previous kernel, inside CUDA graphs too.  BIST_NO_PDL=1 turns the attribute off.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename... P, typename... A>
inline cudaError_t launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, P(args)...);
}
#endif
#define BIST_LAUNCH(kern, grid, block, smem, st, ...) \
  BIST_CUDA(::bist::launch_pdl(kern, dim3(grid), dim3(block), (size_t)(smem), st, __VA_ARGS__))

static inline size_t dtype_size(int dt) { return dt == BIST_BF16 ? 2 : 4; }
static inline bool dtype_ok(int dt) { return dt == BIST_F32 || dt == BIST_BF16; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// A bump allocator over the caller's workspace.
struct Arena {
  char* base;
  size_t cap, used = 0;
  bool ok = true;
  Arena(void* p, size_t n) : base(reinterpret_cast<char*>(p)), cap(n) {}
  void* take(size_t bytes) {
    size_t off = align_up(used, 256);
    if (off + bytes > cap) { ok = false; return nullptr; }
    used = off + bytes;
    return base + off;
  }
};

// ---- element access ---------------------------------------------------------------------
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// Load 4 consecutive elements as floats (address must be 16 B aligned for float, 8 B for bf16).
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&u.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&u.y);
  float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- dropout (training mode, modules.py:44,62-63,113) --------------------------------------------
// Counter-based: element `idx` of dropout site `site` is kept iff hash(seed, site, idx) >= thresh; kept values are
// scaled by 1/(1-p).  Forward and backward regenerate the same mask; nothing is stored.  `idx` is the flat index of
// the element in the REFERENCE's tensor at that site (so that tests can inject identical masks into the oracle
// through bist_dropout_mask).  p == 0 (thresh == 0) disables a site.
struct DropCfg {
  unsigned long long seed = 0;
  const unsigned long long* step = nullptr;  // optional device counter added to the seed at run time (bist_set_dropout_step)
  uint32_t thresh = 0;   // floor(p * 2^32)
  uint32_t site = 0;
  float scale = 1.f;     // 1 / (1 - p)
};
const unsigned long long* dropout_step_ptr();   // api.cu
static inline DropCfg make_drop(float p, unsigned long long seed, uint32_t site) {
  DropCfg c;
  if (p > 0.f) {
    double t = (double)p * 4294967296.0;
    c.thresh = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
    c.scale = 1.f / (1.f - p);
    c.seed = seed; c.site = site;
    c.step = dropout_step_ptr();
  }
  return c;
}
#ifdef __CUDACC__
// 32-bit avalanche mixer ("lowbias32"): a handful of integer instructions per element instead of 64-bit multiplies.
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t drop_hash(unsigned long long seed, uint32_t site, unsigned long long idx) {
  const uint32_t key = mix32((uint32_t)(seed >> 32) + (uint32_t)(idx >> 32) * 0x9E3779B9u + site * 0x85EBCA6Bu) ^ (uint32_t)seed;
  return mix32((uint32_t)idx ^ key);
}
// multiplier of element idx: 0 (dropped) or 1/(1-p)
__device__ __forceinline__ float drop_mul(const DropCfg& c, unsigned long long idx) {
  if (c.thresh == 0u) return c.scale;
  const unsigned long long seed = c.step ? c.seed + __ldg(c.step) : c.seed;   // a CUDA-graph replay advances *step
  return drop_hash(seed, c.site, idx) >= c.thresh ? c.scale : 0.f;
}
#endif

// ---- two-level row addressing -----------------------------------------------------------
// Element offset of logical row r: (r / inner) * outer_stride + (r % inner) * inner_stride.
// Lets one GEMM / attention kernel walk [B,T,S,d] as rows (b,t,s) or (b,s,t), broadcast a
// [B,Lq,d] residual over a group dimension, or pick a per-head column slice, without a copy.
// An optional modulus on the inner index (`mod`) broadcasts rows: (b, g, l) -> (b, l).
struct RowMap {
  int inner;
  long long so, si;
  int mod;
  __host__ __device__ __forceinline__ long long off(int r) const {
    return (long long)(r / inner) * so + (long long)((r % inner) % mod) * si;
  }
};
constexpr int kNoSplit = 1 << 30;
static inline RowMap rows_dense(long long ld) { return RowMap{kNoSplit, 0, ld, kNoSplit}; }
static inline RowMap rows_2level(int inner, long long so, long long si) { return RowMap{inner, so, si, kNoSplit}; }
static inline RowMap rows_bcast(int inner, long long so, int mod, long long si) { return RowMap{inner, so, si, mod}; }

// Batch index z -> inner = z % zin (stride s1), outer = z / zin split once more into outer / zmid (stride s0)
// and outer % zmid (stride s2).  zmid = 1 (default) gives the plain two-level map.
struct BatchMap {
  long long s0, s1, s2 = 0;
  __host__ __device__ __forceinline__ long long off(int z, int zin, int zmid) const {
    const int outer = z / zin;
    return (long long)(outer / zmid) * s0 + (long long)(outer % zmid) * s2 + (long long)(z % zin) * s1;
  }
};

// ---- generic GEMM:  C = act(alpha * A . W^T + bias) + R ------------------------------------
struct GemmArgs {
  const void* A = nullptr;  // [M,K] rows via `a`, K contiguous
  const void* W = nullptr;  // [N,K] rows via `w`, K contiguous
  void* C = nullptr;        // [M,N] rows via `c`, N contiguous
  const void* A2 = nullptr; // split-tf32 products (gemm_x3.cu; tensor-core path only): low halves with the layout of A and W;
  const void* W2 = nullptr; //   C = act(alpha (A W^T + A2 W^T + A W2^T) + bias) + R in one launch, summed in the accumulator
  int kcat = 1;             // K concatenation: C = act(alpha sum_p A_p W_p^T + bias) + R with A_p = A + p * a_cat (columns of the same
  long long a_cat = 0;      //   rows), W_p = W + p * w_cat (elements; a multiple of w's row stride), each part [., K], in every batch.  One
  long long w_cat = 0;      //   launch on the tensor-core path; gemm() falls back to kcat accumulating launches
  void* C2 = nullptr;       // optional bf16 copy of the output at the same element offsets as C (tensor-core path only)
  const float* bias = nullptr;  // [N] fp32 or null
  const void* R = nullptr;      // residual rows via `r`, N contiguous, or null
  int M = 0, N = 0, K = 0;
  RowMap a{kNoSplit, 0, 0, kNoSplit}, w{kNoSplit, 0, 0, kNoSplit}, c{kNoSplit, 0, 0, kNoSplit}, r{kNoSplit, 0, 0, kNoSplit};
  int Z = 1, zin = 1, zmid = 1;  // batch count, inner batch extent, optional split of the outer index
  BatchMap az{0, 0}, wz{0, 0}, cz{0, 0}, rz{0, 0}, bz{0, 0};
  float alpha = 1.f;
  int relu = 0;
  const char* tag = "gemm";  // label for bist_prof_collect
  int allow_tf32 = 0;        // fp32 operands may be multiplied as tf32 on the tensor cores
  int accumulate = 0;        // tensor-core path: the fp32 values already in C are added before the activation (split products)
  int max_ctas = 0;          // persistent tensor-core GEMM: cap on the grid (0 = one CTA per SM); a GEMM forked onto the side
                             // stream leaves some SMs to the serial chain it overlaps with
  // Region-major W (tensor-core path only): W is a [*, wr_T, wr_S, K] tensor (row (bt, s) at (bt*wr_S + s)*w.si) and
  // output column n = s*wr_T + t multiplies row (t, s) of the batch's wr_T x wr_S block; N = wr_T*wr_S.
  int wr_T = 0, wr_S = 0;
  int a_dtype = BIST_F32, w_dtype = BIST_F32, c_dtype = BIST_F32, r_dtype = BIST_F32;
};
// One internal non-blocking stream (+ fork / join events) per host thread and device, for fork-join overlap INSIDE one
// entry point: work that does not depend on the call's serial chain is launched there between an event recorded on the
// caller's stream and an event the caller's stream waits for before the call returns.  Capturable (the side stream joins
// the caller's capture through the fork event).  BIST_NO_OVERLAP=1 disables it.
struct SideStream { cudaStream_t s; cudaEvent_t fork, join, fork2, join2; };
bool dir_overlap_enabled();  // second direction's stage-1/stage-2 kernels on the side stream (BIST_DIR_OVERLAP)
int side_stream(SideStream** out);
bool overlap_enabled();

// fp32 accuracy on the tensor cores: three tf32 products over hi / lo splits (gemm_x3.cu)
struct SplitRef {
  const float* src = nullptr;   // first element of the memory range that holds the operand (all batches)
  long long n = 0;              // elements in the range (multiple of 4)
  float* hi = nullptr;          // scratch mirrors of the range
  float* lo = nullptr;
  int ready = 0;                // the mirrors have been written already
};
int split_tf32(const float* x, float* hi, float* lo, long long n, cudaStream_t st);
bool fp32x3_enabled();
int gemm_fp32x3(const GemmArgs& g, const SplitRef& sa, const SplitRef& sw, cudaStream_t st);

int gemm(const GemmArgs& g, cudaStream_t st);       // dispatches to the kernels below
int gemm_simt(const GemmArgs& g, cudaStream_t st);  // CUDA-core fp32-accumulate GEMM, any shape
int gemm_tc(const GemmArgs& g, cudaStream_t st, bool* taken);  // tcgen05 GEMM; *taken = false when it declines the shape

// ---- LayerNorm ----------------------------------------------------------------------------
// Rows [k*group_rows, (k+1)*group_rows) use a2 + k*pstride, b2 + k*pstride (batched directions).
// With x_strided, group k of x starts at x + k * x_gstride elements (the groups need not be adjacent; the
// stride may be zero or negative).
int layer_norm(const void* x, const float* a2, const float* b2, void* y, int64_t rows, int64_t d,
               float eps, int x_dtype, int y_dtype, cudaStream_t st, int64_t group_rows = 0,
               int64_t pstride = 0, int64_t x_gstride = 0, bool x_strided = false);

int layer_norm2(const float* x, const float* a2a, const float* b2a, const float* a2b, const float* b2b, void* ya, void* yb,
                void* xc, int64_t rows, int64_t d, float eps, int y_dtype, cudaStream_t st, int64_t group_rows = 0,
                int64_t pstride = 0);

// ---- softmax + weighted sum over projected values ---------------------------------------------
// For every group g in [0,G): scores S[g] [Lq*h rows (l-major, h-minor), nk] fp32 (row stride
// ld_s), optional mask, softmax over nk, ctx[g][l, hh*dk + j] = sum_k p[l,hh,k] * V[g][k, hh*dk+j].
struct SoftmaxPvArgs {
  const float* S = nullptr;
  RowMap sg{kNoSplit, 0, 0, kNoSplit};  // group g -> base offset of S
  long long s_l = 0, s_h = 0;  // strides of the (l, h) row of S
  long long s_k = 1;           // stride between consecutive keys of a group
  const uint8_t* mask = nullptr;  // null or uint8, 0 = masked
  int mask_gdiv = 1, mask_gmod = kNoSplit;  // mask batch index = (g / mask_gdiv) % mask_gmod
  long long m_b = 0, m_l = 0;  // mask strides (batch, query row; m_l = 0 for key masks)
  const void* V = nullptr;
  RowMap v{kNoSplit, 0, 0, kNoSplit};   // key row k -> offset within the group's V
  RowMap vg{kNoSplit, 0, 0, kNoSplit};  // group g  -> base offset of V
  void* ctx = nullptr;      // element (g, l, c) at og.off(g) + l * o_l + c
  RowMap og{kNoSplit, 0, 0, kNoSplit};
  long long o_l = 0;
  float* p_out = nullptr;   // optional fp32 [G, h, Lq, nk]
  int G = 0, Lq = 0, h = 0, dk = 0, nk = 0;
  int v_dtype = BIST_F32, ctx_dtype = BIST_F32;
  DropCfg drop;  // dropout on the probabilities (modules.py:62-63); element index ((g*h + hh)*Lq + l)*nk + k
};
int softmax_pv(const SoftmaxPvArgs& a, cudaStream_t st);  // tensor-core kernel when the shape allows, else CUDA cores
int softmax_pv_mma(const SoftmaxPvArgs& a, cudaStream_t st, bool* taken);  // attn_mma.cu

// ---- stage 2 of t2s / s2t: one query row attends over its G stage-1 summaries ---------------------
// For every (b,l): scores[hh,g] = qf[b,l,hh,:] . X1[b,l,g,:] (qf pre-scaled), optional key mask
// [B,G], softmax over g, ctxx[b,l,hh,:] = sum_g p[hh,g] X1[b,l,g,:]   (fp32 [B*Lq, h, d]).
// x1 is laid out [B, Lq, G, d] (the G summaries of one (b,l) are contiguous).
// Training-mode extras of stage 2: dropout on the probabilities (element index (q*h + hh)*G + g) and, because the
// dropped probabilities no longer sum to one, rsum[q,hh] = sum_g p'[hh,g] (fp32 [B*Lq, h]) for the value bias.
struct Stage2Extra {
  DropCfg drop;
  float* rsum = nullptr;
};
int stage2_attn(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G,
                int Lq, int h, int d, int x1_dtype, int ctxx_dtype, cudaStream_t st, const Stage2Extra* ex = nullptr);
int stage2_mma(const void* q, const void* x1, const uint8_t* mask, void* out, int B, int G, int Lq, int h, int d,
               cudaStream_t st, bool* taken, const Stage2Extra* ex = nullptr);  // bf16 q / x1 / out; attn_mma.cu
bool stage2_mma_supported(int G, int h, int d);
// persistent, TMA-pipelined form of stage2_mma (stage2_pipe.cu); *taken = false when the shape is left to stage2_mma's own kernel
int stage2_pipe(const void* q, const void* x1, const uint8_t* mask, void* out, int B, int G, int Lq, int h, int d, cudaStream_t st,
                bool* taken, const Stage2Extra* ex);
// Fused self-attention over the query tokens (bf16): qkv [G*Lq, 3d] -> ctx [G*Lq, d]; key mask uint8 [(g % mask_mod), Lq].
int self_attn_mma(const void* qkv, const uint8_t* mask, int mask_mod, void* ctx, int G, int Lq, int h, int dk, cudaStream_t st,
                  bool* taken);
int stage2_v2(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int h, int d,
              int x1_dtype, int ctxx_dtype, cudaStream_t st, bool* taken);  // attn_mma.cu

// ---- stage 1 of both directions as one kernel (stage1_fused.cu; bf16, d_k = 64, Lq <= 16, T, S <= 128) ----------------------
// kinds[u]: 0 = t2s (groups = regions, keys = clips, temporal key mask), 1 = s2t.  qf[u] bf16 [B*Lq, h, d] folded queries,
// wv[u] bf16 [d, d], bv[u] fp32 [d], c1[u] bf16 [B, Lq, G, d].  Host arrays of device pointers.
bool stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h);
int stage1_fused(const void* vft, const uint8_t* tmask, int nd, const int* kinds, const void* const* qf, const void* const* wv,
                 const float* const* bv, void* const* c1, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h,
                 float* dbg, cudaStream_t st);

// ---- query side of the BiST layer in three kernels (qside.cu; bf16, d_k = 64, h even, Lq <= 16, d_ff % 256 == 0) ---------------
// Host arrays of per-direction device pointers; see qside.cu for layouts.
bool qside_supported(int64_t B, int64_t Lq, int64_t d, int64_t dff, int64_t h);
int qside_attn(const void* xn, const void* const* wqkv, const float* const* bqkv, const uint8_t* qmask, void* ctx, int nd, int64_t B,
               int64_t Lq, int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st);
int qside_fold(const void* xb, const float* const* ln1_a, const float* const* ln1_b, const float* const* ln2_a, const float* const* ln2_b,
               const void* const* wq, const float* const* bq, const void* const* wkt, void* qf, void* xv2, int nd, int64_t B, int64_t Lq,
               int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st);
int qside_ffn1(const void* tsb, const float* const* ln_a, const float* const* ln_b, const void* const* w1, const float* const* b1, void* ffh,
               int nd, int64_t B, int64_t Lq, int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st);

// ---- weight packing helpers ------------------------------------------------------------------
// dst[r, c] = scale * src[r, c] (transpose = 0) or scale * src[c, r] (transpose = 1); dst is
// [rows, cols] dense of dst_dtype, src fp32 dense.
int pack_matrix(const float* src, void* dst, int rows, int cols, int transpose, float scale,
                int dst_dtype, cudaStream_t st);
// Many pack_matrix calls as ONE launch (weights are re-packed after every optimizer step): item i writes
// dst_i[r, c] = scale_i * (transpose_i ? src_i[c, r] : src_i[r, c]) with dst_i = (to_f32_i ? f32_base : base of dtype) + dst_off_i.
struct PackItem {
  const float* src;
  long long dst_off;   // element offset from the destination base
  int rows, cols;      // of the destination
  int transpose;
  int to_f32;          // destination is the fp32 vector section
  float scale;
  int tile_begin;      // filled in by pack_many
};
constexpr int kMaxPackItems = 72;
int pack_many(const PackItem* items, int n, void* base, float* f32_base, int dtype, cudaStream_t st);

}  // namespace bist

// =============================================================================================
// Backward-pass building blocks (bwd.cu)
// =============================================================================================
namespace bist {

// Weight-gradient product (contraction over the ROWS of two row-major operands):
//   C[z][n,k] += alpha * sum_m A[z][m,n] * B[z][m,k]          C fp32, accumulated with red.global.add
//   colsum[z][n] += sum_m A[z][m,n]                            (optional: bias gradients)
// The row range is split across CTAs; C (and colsum) must hold the running sum (zero for a fresh gradient).
struct WgradArgs {
  const void* A = nullptr;
  const void* B = nullptr;
  float* C = nullptr;
  float* colsum = nullptr;
  int M = 0, N = 0, K = 0;
  long long lda = 0, ldb = 0, ldc = 0;
  int Z = 1, zin = 1;
  BatchMap az{0, 0}, bz{0, 0}, cz{0, 0}, sz{0, 0};
  float alpha = 1.f;
  int a_dtype = BIST_F32, b_dtype = BIST_F32;
  const char* tag = "wgrad";
};
int wgrad(const WgradArgs& w, cudaStream_t st);
int wgrad_tc(const WgradArgs& w, cudaStream_t st, bool* taken);  // tcgen05 path (wgrad_tc.cu)

// LayerNorm backward (modules.py:28-31: unbiased std, eps on the std).
//   dx = [accumulate ? dx : 0] + dLN(dy; x)         (optionally gated by x > 0: the ReLU in front of in_norm)
//   da += sum_rows dy * xhat, db += sum_rows dy     (atomics; rows [k*group_rows, ..) use a2 + k*pstride, da/db + k*gstride)
struct LnBwdArgs {
  const void* dy = nullptr;
  const void* x = nullptr;
  const float* a2 = nullptr;
  void* dx = nullptr;
  float* da = nullptr;
  float* db = nullptr;
  int64_t rows = 0, d = 0, group_rows = 0, pstride = 0, gstride = 0;
  float eps = 1e-6f;
  int dy_dtype = BIST_F32, x_dtype = BIST_F32, dx_dtype = BIST_F32;
  int accumulate = 0, relu_gate = 0;
};
int layer_norm_bwd(const LnBwdArgs& a, cudaStream_t st);

// dy[i] = y[i] > 0 ? scale * dy[i] : 0   (in place; y and dy of dtype `dtype`).  With dropout after the ReLU, y is the
// dropped activation (positive only where the unit was active AND kept) and scale = 1/(1-p).
int relu_gate(void* dy, const void* y, int64_t n, int dtype, cudaStream_t st, float scale = 1.f);
// Sublayer / hidden dropout, forward (modules.py:44,113):  out[r,c] = (res ? res[r / res_div, c] : 0) + out[r,c] * m(idx),
// idx = ref_row(r) * d + c.  ref_row(r) = r, or with perm_L > 0 the reference's row of our (b, l, g) row:
// r = (b*perm_L + l)*perm_G + g  ->  (b*perm_G + g)*perm_L + l   (stage-1 outputs are kept [B,Lq,G,d], the reference's are [B*G,Lq,d]).
int dropout_add(void* out, const float* res, int64_t rows, int64_t d, int64_t res_div, const DropCfg& drop, int perm_L, int perm_G,
                int out_dtype, cudaStream_t st, int64_t row0 = 0);   // row0: offset added to ref_row (per-direction calls of a batched site)
// Backward of the same: dst[r,c] = src[r,c] * m(idx)  (also the fp32 -> compute-dtype cast when drop is off; src may equal dst).
int dropout_scale(const void* src, void* dst, int64_t rows, int64_t d, const DropCfg& drop, int perm_L, int perm_G, int src_dtype,
                  int dst_dtype, cudaStream_t st, int64_t row0 = 0);
// Stage-2 value bias under dropout: c2[q, hh*dk + j] += bv[hh*dk + j] * (rsum[q,hh] - 1)   (the GEMM already added bv once)
int bias_rsum_fwd(void* c2, const float* bv, const float* rsum, int64_t Q, int h, int dk, int dtype, cudaStream_t st);
// and its backward: drsum[q,hh] = sum_j dc2[q, hh*dk+j] bv[hh*dk+j];  dbv[hh*dk+j] += sum_q dc2[q, hh*dk+j] rsum[q,hh]
int bias_rsum_bwd(const void* dc2, const float* bv, const float* rsum, float* drsum, float* dbv, int64_t Q, int h, int dk, int dtype,
                  cudaStream_t st);
// out[i] = keep ? 1/(1-p) : 0 for i in [0, n): the multipliers of one dropout site (tests / bist_dropout_mask)
int dropout_mask(float* out, int64_t n, const DropCfg& drop, cudaStream_t st);
// out[q, c] (fp32) += sum_g in[q, g, c]
int group_sum(const void* in, float* out, int64_t Qrows, int64_t G, int64_t d, int in_dtype, cudaStream_t st);
// dst (dst_dtype) = src (src_dtype), n elements
int cast_copy(const void* src, void* dst, int64_t n, int src_dtype, int dst_dtype, cudaStream_t st);

// Backward of scaled-dot-product attention with projected q, k, v (attention(), modules.py:54-64); probabilities are
// recomputed.  Element (b, row, hh*dk + j) of q / dq / dctx at  b*q_b + row*q_l + hh*dk + j  (k, v, dk, dv: k_b, k_l).
struct AttnBwdArgs {
  const void *q = nullptr, *k = nullptr, *v = nullptr, *dctx = nullptr;
  void *dq = nullptr, *dk_ = nullptr, *dv = nullptr;
  long long q_b = 0, q_l = 0, k_b = 0, k_l = 0, c_b = 0, c_l = 0;  // strides of q/dq, k/v/dk/dv, dctx
  const uint8_t* mask = nullptr;
  int mask_mod = kNoSplit;
  long long m_b = 0, m_l = 0;  // mask[(b % mask_mod)*m_b + l*m_l + k]
  int Bp = 0, Lq = 0, Lk = 0, h = 0, dk = 0;
  int dtype = BIST_F32;  // of q, k, v, dctx, dq, dk, dv
  DropCfg drop;          // dropout that the forward applied to the probabilities: element ((b*h + hh)*Lq + l)*Lk + k
};
int attn_core_bwd(const AttnBwdArgs& a, cudaStream_t st);

// Backward of stage 1 (softmax over precomputed folded scores + weighted sum of projected values).
// `f` is the forward's argument block (S = raw scores, mask, V, group/row maps); dctx is addressed like f.ctx.
//   dV  (same addressing as f.V, dtype f.v_dtype)   = sum_l p[l,hh,k] dctx[l, hh-slice]
//   ds  [b][(l*h+hh)][ts]  and  dsT [b][ts][(l*h+hh)]   (dtype ds_dtype), ts = (g % gdiv)*ts_g + k*ts_k, b = g / gdiv
struct Stage1BwdArgs {
  SoftmaxPvArgs f;
  const void* dctx = nullptr;
  void* dV = nullptr;
  void* ds = nullptr;
  void* dsT = nullptr;
  int gdiv = 1, TS = 0;
  long long ts_g = 0, ts_k = 0;
  int ds_dtype = BIST_F32;
  int dsT_ld = 0;   // elements between rows of dsT (0 = Lq * h: dense); wider when several directions' dsT sit side by side
};
int stage1_bwd(const Stage1BwdArgs& a, cudaStream_t st);      // a.ds may be null (only dsT is written); tensor cores when possible
int stage1_bwd_mma(const Stage1BwdArgs& a, cudaStream_t st, bool* taken);  // bwd_mma.cu

// Backward of stage 2: per (b,l) one folded query row per head against its G summaries x1[b,l,g,:].
//   scores s[hh,g] = qf[q,hh,:].x1[q,g,:] (recomputed), p = softmax_g(mask), u[hh,:] = sum_g p x1[g,:]
//   in : du [Q,h,d];  out: dx1 [Q,G,d] = sum_hh p du + ds qf ;  dqf [Q,h,d] = sum_g ds x1
struct Stage2BwdArgs {
  const void *qf = nullptr, *x1 = nullptr, *du = nullptr;
  const uint8_t* mask = nullptr;  // [B, G] or null
  void *dx1 = nullptr, *dqf = nullptr;
  int B = 0, G = 0, Lq = 0, h = 0, d = 0;
  int dtype = BIST_F32;  // of all six tensors
  DropCfg drop;                  // dropout of the forward's probabilities (Stage2Extra)
  const float* drsum = nullptr;  // fp32 [Q,h] gradient of rsum (added to every dP'[hh,g]) or null
};
int stage2_bwd(const Stage2BwdArgs& a, cudaStream_t st);
int stage2_bwd_mma(const Stage2BwdArgs& a, cudaStream_t st, bool* taken);  // bwd_mma.cu
// dst[z][c][r] = src[z][r][c]  (Z matrices of [rows, cols], dense)
int transpose_batched(const void* src, void* dst, int64_t Z, int64_t rows, int64_t cols, int dtype, cudaStream_t st);

}  // namespace bist
