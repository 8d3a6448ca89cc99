// C-ABI entry points of libbist_b200.so (include/bist_b200.h): error plumbing and the generic
// blocks of /root/reference/model/modules.py (LayerNorm, Linear, MultiHeadedAttention,
// PositionwiseFeedForward) plus the VidEncoder8 projection.  The VidEncoderLayer4 entry points
// live in st_layer.cu.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace bist {

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool pdl_enabled() {
  static const bool on = getenv("BIST_NO_PDL") == nullptr;
  return on;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static std::atomic<const unsigned long long*> g_drop_step{nullptr};
const unsigned long long* dropout_step_ptr() { return g_drop_step.load(std::memory_order_relaxed); }

// ---- per-launch event timing (off by default) ---------------------------------------------
struct ProfRec { const char* tag; cudaEvent_t e0, e1; };
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
static std::vector<ProfRec*> g_prof_recs;

ProfScope::ProfScope(const char* tag, cudaStream_t s) : rec(nullptr), st(s) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  ProfRec* r = new ProfRec{tag, nullptr, nullptr};
  if (cudaEventCreate(&r->e0) != cudaSuccess || cudaEventCreate(&r->e1) != cudaSuccess) { delete r; return; }
  cudaEventRecord(r->e0, st);
  rec = r;
}
ProfScope::~ProfScope() {
  if (!rec) return;
  ProfRec* r = (ProfRec*)rec;
  cudaEventRecord(r->e1, st);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_recs.push_back(r);
}

bool overlap_enabled() {   // off while per-launch timing is on: a kernel timed beside another is not its own duration
  static const bool on = getenv("BIST_NO_OVERLAP") == nullptr;
  return on && !g_prof_on.load(std::memory_order_relaxed);
}
bool dir_overlap_enabled() {
  static const bool on = [] { const char* v = getenv("BIST_DIR_OVERLAP"); return v ? atoi(v) != 0 : true; }();
  return on && overlap_enabled();
}
int side_stream(SideStream** out) {
  static thread_local std::map<int, SideStream> tab;
  int dev = 0;
  BIST_CUDA(cudaGetDevice(&dev));
  auto it = tab.find(dev);
  if (it == tab.end()) {
    SideStream ss{};
    BIST_CUDA(cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking));
    BIST_CUDA(cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming));
    BIST_CUDA(cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming));
    BIST_CUDA(cudaEventCreateWithFlags(&ss.fork2, cudaEventDisableTiming));
    BIST_CUDA(cudaEventCreateWithFlags(&ss.join2, cudaEventDisableTiming));
    it = tab.emplace(dev, ss).first;
  }
  *out = &it->second;
  return 0;
}

int gemm_tc(const GemmArgs& g, cudaStream_t st, bool* taken);  // gemm_tc.cu

int gemm(const GemmArgs& g, cudaStream_t st) {
  bool taken = false;
  int rc = gemm_tc(g, st, &taken);
  if (taken) return rc;
  if (g.kcat > 1) {   // K-concatenated product outside the tensor-core path: one launch per part, parts after the first add onto C
    BIST_CHECK_ARG(!g.relu && !g.C2, "gemm: a K-concatenated product takes no ReLU or second output");
    for (int p = 0; p < g.kcat; ++p) {
      GemmArgs q = g;
      q.kcat = 1;
      const int asz = g.a_dtype == BIST_BF16 ? 2 : 4, wsz = g.w_dtype == BIST_BF16 ? 2 : 4;
      q.A = reinterpret_cast<const char*>(g.A) + (size_t)p * g.a_cat * asz;
      q.W = reinterpret_cast<const char*>(g.W) + (size_t)p * g.w_cat * wsz;
      if (p > 0) { q.bias = nullptr; q.R = g.C; q.r = g.c; q.rz = g.cz; q.r_dtype = g.c_dtype; }
      BIST_TRY(gemm(q, st));
    }
    return 0;
  }
  return gemm_simt(g, st);
}

// ---- skinny Linear (the gate / switch heads: a handful of outputs per row, decoder.py:147-167, generator.py:71,120) ----
// N <= 8 outputs: one warp per row keeps the row of A in flight once and reduces N dot products with shuffles.
template <typename T>
__global__ void __launch_bounds__(256) skinny_fwd_kernel(const T* __restrict__ A, const T* __restrict__ W, const float* __restrict__ bias,
                                                         T* __restrict__ C, int64_t M, int N, int K) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t m = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (m >= M) return;
  float acc[8];
#pragma unroll
  for (int n = 0; n < 8; ++n) acc[n] = 0.f;
  const T* a = A + m * K;
  if (sizeof(T) == 2 && K % 8 == 0 && ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(W)) & 15) == 0) {
    // 16 bytes (8 bf16) per lane and operand per iteration: the row is a handful of independent wide loads instead of K / 32
    // dependent two-byte ones (47 us -> a few us for 768 x 2048 on the step's critical path)
    for (int k = lane * 8; k < K; k += 256) {
      const uint4 xa = *reinterpret_cast<const uint4*>(a + k);
      const __nv_bfloat162* xp = reinterpret_cast<const __nv_bfloat162*>(&xa);
      float xf[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float2 f = __bfloat1622float2(xp[j]); xf[2 * j] = f.x; xf[2 * j + 1] = f.y; }
#pragma unroll
      for (int n = 0; n < 8; ++n) {
        if (n < N) {
          const uint4 wa = *reinterpret_cast<const uint4*>(W + (int64_t)n * K + k);
          const __nv_bfloat162* wp = reinterpret_cast<const __nv_bfloat162*>(&wa);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = __bfloat1622float2(wp[j]);
            acc[n] = fmaf(xf[2 * j], f.x, acc[n]);
            acc[n] = fmaf(xf[2 * j + 1], f.y, acc[n]);
          }
        }
      }
    }
  } else {
    for (int k = lane; k < K; k += 32) {
      const float x = to_f32(a[k]);
#pragma unroll
      for (int n = 0; n < 8; ++n)
        if (n < N) acc[n] = fmaf(x, to_f32(W[(int64_t)n * K + k]), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    const float v = warp_sum(acc[n]);
    if (n < N && lane == 0) C[m * N + n] = from_f32<T>(v + (bias ? bias[n] : 0.f));
  }
}
// dx[m,k] = sum_n dy[m,n] wT[k,n] with N <= 8: one thread per output element.
template <typename T>
__global__ void __launch_bounds__(256) skinny_dx_kernel(const T* __restrict__ dy, const T* __restrict__ wT, T* __restrict__ dx, int64_t M,
                                                        int N, int K) {
  pdl_trigger();
  pdl_wait();
  const int64_t total = M * K;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / K;
    const int k = (int)(i - m * K);
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc = fmaf(to_f32(dy[m * N + n]), to_f32(wT[(int64_t)k * N + n]), acc);
    dx[i] = from_f32<T>(acc);
  }
}

// out = x + y * m(idx): the residual add of SublayerConnection fused with its dropout (modules.py:44)
template <typename T>
__global__ void __launch_bounds__(256) dropout_residual_kernel(const T* __restrict__ y, const T* __restrict__ x, T* __restrict__ out,
                                                               int64_t n, DropCfg drop) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = from_f32<T>(fmaf(to_f32(y[i]), drop_mul(drop, (unsigned long long)i), to_f32(x[i])));
}

// eight elements per thread and iteration, hash key computed once (n % 8 == 0, n < 2^32, 16-byte aligned pointers)
template <typename T>
__global__ void __launch_bounds__(256) dropout_residual8_kernel(const T* __restrict__ y, const T* __restrict__ x, T* __restrict__ out,
                                                                uint32_t n8, DropCfg drop) {
  pdl_trigger();
  pdl_wait();
  uint32_t key = 0u;
  const bool on = drop.thresh != 0u;
  if (on) {   // drop_hash() with the high half of the element index equal to zero
    const unsigned long long seed = drop.step ? drop.seed + __ldg(drop.step) : drop.seed;
    key = mix32((uint32_t)(seed >> 32) + drop.site * 0x85EBCA6Bu) ^ (uint32_t)seed;
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += gridDim.x * blockDim.x) {
    float yv[8], xv[8];
    if (sizeof(T) == 2) {
      const uint4 uy = *reinterpret_cast<const uint4*>(y + (size_t)i * 8), ux = *reinterpret_cast<const uint4*>(x + (size_t)i * 8);
      const uint32_t wy[4] = {uy.x, uy.y, uy.z, uy.w}, wx[4] = {ux.x, ux.y, ux.z, ux.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        yv[2 * j] = __uint_as_float(wy[j] << 16); yv[2 * j + 1] = __uint_as_float(wy[j] & 0xffff0000u);
        xv[2 * j] = __uint_as_float(wx[j] << 16); xv[2 * j + 1] = __uint_as_float(wx[j] & 0xffff0000u);
      }
    } else {
      const float4 a = reinterpret_cast<const float4*>(y)[(size_t)i * 2], b = reinterpret_cast<const float4*>(y)[(size_t)i * 2 + 1];
      const float4 c = reinterpret_cast<const float4*>(x)[(size_t)i * 2], e = reinterpret_cast<const float4*>(x)[(size_t)i * 2 + 1];
      yv[0] = a.x; yv[1] = a.y; yv[2] = a.z; yv[3] = a.w; yv[4] = b.x; yv[5] = b.y; yv[6] = b.z; yv[7] = b.w;
      xv[0] = c.x; xv[1] = c.y; xv[2] = c.z; xv[3] = c.w; xv[4] = e.x; xv[5] = e.y; xv[6] = e.z; xv[7] = e.w;
    }
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float m = on ? (mix32((i * 8u + (uint32_t)j) ^ key) >= drop.thresh ? drop.scale : 0.f) : drop.scale;
      o[j] = fmaf(yv[j], m, xv[j]);
    }
    if (sizeof(T) == 2) {
      uint4 u;
      __nv_bfloat162 h0 = __floats2bfloat162_rn(o[0], o[1]), h1 = __floats2bfloat162_rn(o[2], o[3]);
      __nv_bfloat162 h2 = __floats2bfloat162_rn(o[4], o[5]), h3 = __floats2bfloat162_rn(o[6], o[7]);
      u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
      u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
      *reinterpret_cast<uint4*>(out + (size_t)i * 8) = u;
    } else {
      reinterpret_cast<float4*>(out)[(size_t)i * 2] = make_float4(o[0], o[1], o[2], o[3]);
      reinterpret_cast<float4*>(out)[(size_t)i * 2 + 1] = make_float4(o[4], o[5], o[6], o[7]);
    }
  }
}

// Ingest gather: the zero-padded device batch is written by ONE kernel whose threads read the page-locked host rows of
// each video directly over PCIe (unified addressing maps pinned host memory into the device's address space) — every
// video's transfer is in flight at once, instead of one DMA descriptor after the other with a set-up gap between them.
constexpr int kIngestMax = 64;
struct IngestTable {
  const uint4* src[kIngestMax];
  long long chunks[kIngestMax];     // 16-byte chunks of real data per video
  int n;
  long long slot_chunks;            // 16-byte chunks per batch row (T * S * D * elem / 16)
};
__global__ void __launch_bounds__(256) ingest_gather_kernel(const __grid_constant__ IngestTable tab, uint4* __restrict__ dst) {
  const long long total = (long long)tab.n * tab.slot_chunks;
  const long long stride = (long long)gridDim.x * blockDim.x;
  // four independent 16-byte host reads in flight per thread: a few CTAs keep PCIe busy and the SMs stay free for the
  // forward that runs beside this kernel
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += 4 * stride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * stride;
      v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (i < total) {
        const int b = (int)(i / tab.slot_chunks);
        const long long c = i - (long long)b * tab.slot_chunks;
        if (c < tab.chunks[b]) v[u] = __ldcs(tab.src[b] + c);      // streaming read from host memory
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * stride;
      if (i < total) dst[i] = v[u];
    }
  }
}

}  // namespace bist

using namespace bist;

extern "C" {

int bist_version(void) { return BIST_B200_VERSION; }
const char* bist_last_error(void) { return g_err; }
int64_t bist_launch_count(void) { return (int64_t)g_launches.load(); }

int bist_set_dropout_step(const uint64_t* device_counter) {
  g_drop_step.store(reinterpret_cast<const unsigned long long*>(device_counter));
  return 0;
}

int bist_prof_enable(int on) {
  g_prof_on.store(on ? 1 : 0);
  return 0;
}

// Synchronises the device, then writes "tag,launches,total_ms\n" lines (one per tag) into buf.
int bist_prof_collect(char* buf, size_t cap) {
  BIST_CHECK_ARG(buf && cap > 0, "prof_collect: bad buffer");
  BIST_CUDA(cudaDeviceSynchronize());
  std::map<std::string, std::pair<long long, double>> agg;
  {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (ProfRec* r : g_prof_recs) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, r->e0, r->e1) == cudaSuccess) {
        auto& a = agg[r->tag];
        a.first += 1;
        a.second += ms;
      }
      cudaEventDestroy(r->e0);
      cudaEventDestroy(r->e1);
      delete r;
    }
    g_prof_recs.clear();
  }
  size_t used = 0;
  buf[0] = 0;
  for (auto& kv : agg) {
    int n = snprintf(buf + used, cap - used, "%s,%lld,%.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    if (n < 0 || (size_t)n >= cap - used) break;
    used += n;
  }
  return 0;
}

int bist_device_check(void) {
  int dev = 0;
  BIST_CUDA(cudaGetDevice(&dev));
  int major = 0;
  BIST_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    set_error("libbist_b200 is built for sm_100a only; device %d has compute capability %d.x", dev, major);
    return BIST_ERR_DEVICE;
  }
  return 0;
}

int bist_ln_fwd(const void* x, const float* a_2, const float* b_2, void* y, int64_t rows, int64_t d, float eps,
                int x_dtype, int y_dtype, void* stream) {
  return layer_norm(x, a_2, b_2, y, rows, d, eps, x_dtype, y_dtype, (cudaStream_t)stream);
}

int bist_linear_fwd(const void* x, const void* w, const float* bias, const void* residual, void* y, int64_t M,
                    int64_t N, int64_t K, int relu, int x_dtype, int w_dtype, int y_dtype, void* stream) {
  BIST_CHECK_ARG(x && w && y, "linear: null pointer");
  BIST_CHECK_ARG(M > 0 && N > 0 && K > 0 && M < (1LL << 31) && N < (1LL << 31) && K < (1LL << 31),
                 "linear: bad shape");
  BIST_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(w_dtype) && dtype_ok(y_dtype), "linear: bad dtype");
  if (N <= 8 && !relu && !residual && x_dtype == w_dtype && x_dtype == y_dtype) {
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope ps("skinny_linear", st);
    const unsigned grid = (unsigned)((M + 7) / 8);
    if (x_dtype == BIST_BF16)
      BIST_LAUNCH(skinny_fwd_kernel<__nv_bfloat16>, grid, 256, 0, st, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, bias, (__nv_bfloat16*)y, M, (int)N, (int)K);
    else
      BIST_LAUNCH(skinny_fwd_kernel<float>, grid, 256, 0, st, (const float*)x, (const float*)w, bias, (float*)y, M, (int)N, (int)K);
    BIST_LAUNCHED();
    return 0;
  }
  GemmArgs g;
  g.A = x; g.W = w; g.C = y; g.bias = bias; g.R = residual;
  g.M = (int)M; g.N = (int)N; g.K = (int)K;
  g.a = rows_dense(K); g.w = rows_dense(K); g.c = rows_dense(N); g.r = rows_dense(N);
  g.relu = relu;
  g.a_dtype = x_dtype; g.w_dtype = w_dtype; g.c_dtype = y_dtype; g.r_dtype = y_dtype;
  return gemm(g, (cudaStream_t)stream);
}

// ----------------------------------------------------------------------------- MultiHeadedAttention
size_t bist_mha_workspace_bytes(int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int64_t d_in, int64_t h) {
  (void)d_in;
  size_t e = 4;  // sized for fp32; bf16 needs less
  size_t n = 0;
  n += align_up((size_t)Bp * Lq * d * e, 256);      // q
  n += 2 * align_up((size_t)Bp * Lk * d * e, 256);  // k, v
  n += align_up((size_t)Bp * Lq * d * e, 256);      // ctx
  n += align_up((size_t)Bp * Lq * h * Lk * 4, 256); // scores
  return n + 1024;
}

}  // extern "C"

namespace {

// Carving of the MultiHeadedAttention workspace; the training forward leaves q, k, v, ctx there for the backward.
struct MhaWs {
  void *q, *k, *v, *ctx;
  float* sc;
  bool ok;
  MhaWs(void* ws, size_t ws_bytes, int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int64_t h, size_t e) {
    Arena A(ws, ws_bytes);
    q = A.take((size_t)Bp * Lq * d * e);
    k = A.take((size_t)Bp * Lk * d * e);
    v = A.take((size_t)Bp * Lk * d * e);
    ctx = A.take((size_t)Bp * Lq * d * e);
    sc = (float*)A.take((size_t)Bp * Lq * h * Lk * 4);
    ok = A.ok;
  }
};

int mha_forward(const void* query, const void* key, const void* value, const uint8_t* mask, int mask_mode,
                const bist_mha_params_t* P, void* out, float* attn_out, int64_t Bp, int64_t Lq, int64_t Lk,
                int64_t d, int64_t d_in, int64_t h, void* ws, size_t ws_bytes, int dtype, const DropCfg& drop,
                void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(query && key && value && P && out && ws, "mha: null pointer");
  BIST_CHECK_ARG(Bp > 0 && Lq > 0 && Lk > 0 && d > 0 && d_in > 0 && h > 0 && d % h == 0, "mha: bad shape");
  BIST_CHECK_ARG(dtype_ok(dtype), "mha: bad dtype");
  BIST_CHECK_ARG((mask_mode == BIST_MASK_NONE) == (mask == nullptr) && mask_mode >= 0 && mask_mode <= 2,
                 "mha: mask and mask_mode disagree");
  for (int i = 0; i < 4; ++i) BIST_CHECK_ARG(P->w[i] && P->b[i], "mha: null parameter %d", i);
  BIST_CHECK_ARG(Bp * h <= 65535 * 1LL, "mha: Bp*h too large");
  const int dk = (int)(d / h);
  const size_t e = dtype_size(dtype);
  MhaWs W(ws, ws_bytes, Bp, Lq, Lk, d, h, e);
  if (!W.ok) { set_error("mha: workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }
  void *q = W.q, *k = W.k, *v = W.v, *ctx = W.ctx;
  float* sc = W.sc;

  // modules.py:89-91 — the three input projections
  const void* in[3] = {query, key, value};
  void* proj[3] = {q, k, v};
  const int64_t rows[3] = {Bp * Lq, Bp * Lk, Bp * Lk};
  for (int i = 0; i < 3; ++i)
    BIST_TRY(bist_linear_fwd(in[i], P->w[i], P->b[i], nullptr, proj[i], rows[i], d, d_in, 0, dtype, dtype, dtype, st));

  // modules.py:57-58 — scores[b,l,hh,k] = q[b,l,hh,:].k[b,k,hh,:] / sqrt(dk), batched over (b,hh)
  GemmArgs g;
  g.A = q; g.W = k; g.C = sc;
  g.M = (int)Lq; g.N = (int)Lk; g.K = dk;
  g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(h * Lk);
  g.Z = (int)(Bp * h); g.zin = (int)h;
  g.az = {Lq * d, dk}; g.wz = {Lk * d, dk}; g.cz = {Lq * h * Lk, Lk};
  g.alpha = 1.f / sqrtf((float)dk);
  g.a_dtype = dtype; g.w_dtype = dtype; g.c_dtype = BIST_F32;
  BIST_TRY(gemm(g, st));

  // modules.py:59-64 — mask, softmax, weighted sum
  SoftmaxPvArgs a;
  a.S = sc; a.sg = rows_dense(Lq * h * Lk); a.s_l = h * Lk; a.s_h = Lk;
  a.mask = mask; a.m_b = (mask_mode == BIST_MASK_FULL) ? Lq * Lk : Lk; a.m_l = (mask_mode == BIST_MASK_FULL) ? Lk : 0;
  a.V = v; a.v = rows_dense(d); a.vg = rows_dense(Lk * d);
  a.ctx = ctx; a.og = rows_dense(Lq * d); a.o_l = d; a.p_out = attn_out;
  a.G = (int)Bp; a.Lq = (int)Lq; a.h = (int)h; a.dk = dk; a.nk = (int)Lk;
  a.v_dtype = dtype; a.ctx_dtype = dtype;
  a.drop = drop;                      // modules.py:62-63 (element ((b*h + hh)*Lq + l)*Lk + k of p_attn)
  BIST_TRY(softmax_pv(a, st));

  // modules.py:98-100 — output projection
  return bist_linear_fwd(ctx, P->w[3], P->b[3], nullptr, out, Bp * Lq, d_in, d, 0, dtype, dtype, dtype, st);
}

}  // namespace

extern "C" {

int bist_mha_fwd(const void* query, const void* key, const void* value, const uint8_t* mask, int mask_mode,
                 const bist_mha_params_t* P, void* out, float* attn_out, int64_t Bp, int64_t Lq, int64_t Lk,
                 int64_t d, int64_t d_in, int64_t h, void* ws, size_t ws_bytes, int dtype, void* stream) {
  return mha_forward(query, key, value, mask, mask_mode, P, out, attn_out, Bp, Lq, Lk, d, d_in, h, ws, ws_bytes, dtype,
                     DropCfg(), stream);
}

// ----------------------------------------------------------------------------- PositionwiseFeedForward
int bist_ffn_fwd(const void* x, const void* w1, const float* b1, const void* w2, const float* b2, void* y,
                 int64_t rows, int64_t d, int64_t d_ff, int64_t d_out, void* ws, size_t ws_bytes, int dtype,
                 void* stream) {
  BIST_CHECK_ARG(x && w1 && w2 && y && ws, "ffn: null pointer");
  BIST_CHECK_ARG(dtype_ok(dtype), "ffn: bad dtype");
  if (ws_bytes < (size_t)rows * d_ff * dtype_size(dtype)) { set_error("ffn: workspace too small"); return BIST_ERR_WORKSPACE; }
  BIST_TRY(bist_linear_fwd(x, w1, b1, nullptr, ws, rows, d_ff, d, 1, dtype, dtype, dtype, stream));
  return bist_linear_fwd(ws, w2, b2, nullptr, y, rows, d_out, d_ff, 0, dtype, dtype, dtype, stream);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------- backward building blocks
extern "C" {

int bist_ln_bwd(const void* dy, const void* x, const float* a_2, void* dx, float* da_2, float* db_2, int64_t rows, int64_t d,
                float eps, int dy_dtype, int x_dtype, int dx_dtype, void* stream) {
  BIST_CHECK_ARG(dtype_ok(dy_dtype) && dtype_ok(x_dtype) && dtype_ok(dx_dtype), "ln_bwd: bad dtype");
  LnBwdArgs a;
  a.dy = dy; a.x = x; a.a2 = a_2; a.dx = dx; a.da = da_2; a.db = db_2;
  a.rows = rows; a.d = d; a.eps = eps;
  a.dy_dtype = dy_dtype; a.x_dtype = x_dtype; a.dx_dtype = dx_dtype;
  return layer_norm_bwd(a, (cudaStream_t)stream);
}

int bist_linear_bwd(const void* dy, const void* x, const void* wT, void* dx, float* dw, float* db, int64_t M, int64_t N,
                    int64_t K, int dy_dtype, int x_dtype, int w_dtype, int dx_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(dy && M > 0 && N > 0 && K > 0 && M < (1LL << 31), "linear_bwd: bad argument");
  BIST_CHECK_ARG(dtype_ok(dy_dtype) && dtype_ok(x_dtype) && dtype_ok(w_dtype) && dtype_ok(dx_dtype), "linear_bwd: bad dtype");
  if (dx && N <= 8 && dy_dtype == w_dtype && dy_dtype == dx_dtype) {
    BIST_CHECK_ARG(wT, "linear_bwd: dx needs the transposed weight");
    ProfScope ps("skinny_linear_dx", st);
    const unsigned grid = (unsigned)std::min<int64_t>((M * K + 255) / 256, 148 * 8);
    if (dy_dtype == BIST_BF16)
      BIST_LAUNCH(skinny_dx_kernel<__nv_bfloat16>, grid, 256, 0, st, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)wT, (__nv_bfloat16*)dx, M, (int)N, (int)K);
    else
      BIST_LAUNCH(skinny_dx_kernel<float>, grid, 256, 0, st, (const float*)dy, (const float*)wT, (float*)dx, M, (int)N, (int)K);
    BIST_LAUNCHED();
  } else if (dx) {
    BIST_CHECK_ARG(wT, "linear_bwd: dx needs the transposed weight");
    GemmArgs g;
    g.A = dy; g.W = wT; g.C = dx;
    g.M = (int)M; g.N = (int)K; g.K = (int)N;
    g.a = rows_dense(N); g.w = rows_dense(N); g.c = rows_dense(K);
    g.a_dtype = dy_dtype; g.w_dtype = w_dtype; g.c_dtype = dx_dtype; g.tag = "bwd_dx_gemm";
    BIST_TRY(gemm(g, st));
  }
  if (dw) {
    BIST_CHECK_ARG(x, "linear_bwd: dw needs x");
    WgradArgs w;
    w.A = dy; w.B = x; w.C = dw; w.colsum = db;
    w.M = (int)M; w.N = (int)N; w.K = (int)K; w.lda = N; w.ldb = K; w.ldc = K;
    w.a_dtype = dy_dtype; w.b_dtype = x_dtype; w.tag = "wgrad";
    BIST_TRY(wgrad(w, st));
  }
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------- stand-alone modules, training
// MultiHeadedAttention / PositionwiseFeedForward / nn.Dropout of /root/reference/model/modules.py used on their own
// (the decoder, caption and audio layers are built from them: decoder.py:11-60, encoder.py:203-235).
extern "C" {

int bist_dropout_fwd(const void* x, void* y, int64_t rows, int64_t d, float p, uint64_t seed, uint32_t site, int dtype,
                     void* stream) {
  BIST_CHECK_ARG(x && y && rows > 0 && d > 0, "dropout: bad argument");
  BIST_CHECK_ARG(dtype_ok(dtype) && p >= 0.f && p < 1.f, "dropout: bad dtype or probability");
  return dropout_scale(x, y, rows, d, make_drop(p, seed, site), 0, 0, dtype, dtype, (cudaStream_t)stream);
}

// Compute-dtype (optionally transposed) copies of many fp32 parameters into ONE destination buffer, in as few launches
// as the 72-item table of pack_many allows (the stand-alone modules refresh all their weight copies together after an
// optimizer step, bist_b200/modules.py: _CopyArena).
int bist_pack_params(const bist_pack_item_t* items, int64_t n, void* base, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(items && n > 0 && base && dtype_ok(dtype), "pack_params: bad argument");
  PackItem tab[kMaxPackItems];
  static float dummy_f32;   // pack_many wants a non-null fp32 base; no item here targets it
  for (int64_t i0 = 0; i0 < n; i0 += kMaxPackItems) {
    const int m = (int)std::min<int64_t>(kMaxPackItems, n - i0);
    for (int i = 0; i < m; ++i) {
      const bist_pack_item_t& it = items[i0 + i];
      BIST_CHECK_ARG(it.src && it.rows > 0 && it.cols > 0 && it.dst_off >= 0, "pack_params: bad item %lld", (long long)(i0 + i));
      tab[i] = PackItem{it.src, (long long)it.dst_off, (int)it.rows, (int)it.cols, it.transpose ? 1 : 0, 0, 1.f, 0};
    }
    BIST_TRY(pack_many(tab, m, base, &dummy_f32, dtype, st));
  }
  return 0;
}

// Ingest: n ragged host->device row copies and the zero fill of each video's padding, issued back to back on one stream
// (bist_b200/ingest.py: FeatureCache.collate).  src[i]: page-locked host rows; offsets in bytes from dst_base.
int bist_ingest_gather(const void* const* src, void* dst_base, const int64_t* dst_off, const int64_t* nbytes,
                       const int64_t* pad_off, const int64_t* pad_nbytes, int64_t n, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(src && dst_base && dst_off && nbytes && pad_off && pad_nbytes && n > 0, "ingest_gather: bad argument");
  char* base = (char*)dst_base;
  // Fast path: dense batch rows of equal size, 16-byte granularity -> one gather kernel per 64 videos (see above).
  {
    static const bool kernel_on = getenv("BIST_INGEST_DMA") == nullptr;
    bool dense = kernel_on && ((uintptr_t)dst_base % 16 == 0);
    const int64_t slot = n > 0 ? nbytes[0] + pad_nbytes[0] : 0;
    for (int64_t i = 0; i < n && dense; ++i)
      dense = src[i] && ((uintptr_t)src[i] % 16 == 0) && nbytes[i] % 16 == 0 && nbytes[i] + pad_nbytes[i] == slot &&
              dst_off[i] == i * slot && (pad_nbytes[i] == 0 || pad_off[i] == i * slot + nbytes[i]);
    if (dense && slot > 0 && slot % 16 == 0) {
      for (int64_t i0 = 0; i0 < n; i0 += kIngestMax) {
        IngestTable tab{};
        tab.n = (int)std::min<int64_t>(kIngestMax, n - i0);
        tab.slot_chunks = slot / 16;
        for (int i = 0; i < tab.n; ++i) { tab.src[i] = (const uint4*)src[i0 + i]; tab.chunks[i] = nbytes[i0 + i] / 16; }
        const long long total = (long long)tab.n * tab.slot_chunks;
        static const int max_ctas = [] { const char* e = getenv("BIST_INGEST_CTAS"); return e ? atoi(e) : 64; }();
        const unsigned grid = (unsigned)std::min<long long>((total + 1023) / 1024, (long long)max_ctas);
        ProfScope ps("ingest_gather", st);
        BIST_LAUNCH(ingest_gather_kernel, grid, 256, 0, st, tab, (uint4*)(base + i0 * slot));
        BIST_LAUNCHED();
      }
      return 0;
    }
  }
  // General path (BIST_INGEST_DMA=1 forces it): copy-engine transfers, all fills first, then the copies back to back.
  for (int64_t i = 0; i < n; ++i) {
    BIST_CHECK_ARG(src[i] && nbytes[i] >= 0 && pad_nbytes[i] >= 0, "ingest_gather: bad item %lld", (long long)i);
    if (pad_nbytes[i] > 0) BIST_CUDA(cudaMemsetAsync(base + pad_off[i], 0, (size_t)pad_nbytes[i], st));
  }
  for (int64_t i = 0; i < n; ++i)
    if (nbytes[i] > 0) BIST_CUDA(cudaMemcpyAsync(base + dst_off[i], src[i], (size_t)nbytes[i], cudaMemcpyHostToDevice, st));
  return 0;
}

int bist_dropout_add_fwd(const void* y, const void* x, void* out, int64_t rows, int64_t d, float p, uint64_t seed,
                         uint32_t site, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(y && x && out && rows > 0 && d > 0, "dropout_add: bad argument");
  BIST_CHECK_ARG(dtype_ok(dtype) && p >= 0.f && p < 1.f, "dropout_add: bad dtype or probability");
  const int64_t n = rows * d;
  const DropCfg drop = make_drop(p, seed, site);
  ProfScope ps("dropout", st);
  if (n % 8 == 0 && n < (1LL << 32) && (((uintptr_t)y | (uintptr_t)x | (uintptr_t)out) & 15) == 0) {
    const unsigned g8 = (unsigned)std::min<int64_t>((n / 8 + 255) / 256, 148 * 8);
    if (dtype == BIST_BF16)
      BIST_LAUNCH(dropout_residual8_kernel<__nv_bfloat16>, g8, 256, 0, st, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x, (__nv_bfloat16*)out, (uint32_t)(n / 8), drop);
    else
      BIST_LAUNCH(dropout_residual8_kernel<float>, g8, 256, 0, st, (const float*)y, (const float*)x, (float*)out, (uint32_t)(n / 8), drop);
    BIST_LAUNCHED();
    return 0;
  }
  const unsigned grid = (unsigned)std::min<int64_t>((n + 255) / 256, 148 * 8);
  if (dtype == BIST_BF16)
    BIST_LAUNCH(dropout_residual_kernel<__nv_bfloat16>, grid, 256, 0, st, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x, (__nv_bfloat16*)out, n, drop);
  else
    BIST_LAUNCH(dropout_residual_kernel<float>, grid, 256, 0, st, (const float*)y, (const float*)x, (float*)out, n, drop);
  BIST_LAUNCHED();
  return 0;
}

int bist_mha_train_fwd(const void* query, const void* key, const void* value, const uint8_t* mask, int mask_mode,
                       const bist_mha_params_t* P, void* out, float* attn_out, int64_t Bp, int64_t Lq, int64_t Lk,
                       int64_t d, int64_t d_in, int64_t h, void* tape, size_t tape_bytes, int dtype, float p_drop,
                       uint64_t seed, uint32_t site, void* stream) {
  BIST_CHECK_ARG(p_drop >= 0.f && p_drop < 1.f, "mha(train): dropout probability must be in [0, 1)");
  return mha_forward(query, key, value, mask, mask_mode, P, out, attn_out, Bp, Lq, Lk, d, d_in, h, tape, tape_bytes, dtype,
                     make_drop(p_drop, seed, site), stream);
}

size_t bist_mha_bwd_workspace_bytes(int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int dtype) {
  const size_t e = dtype_size(dtype);
  return 2 * align_up((size_t)Bp * Lq * d * e, 256) + 2 * align_up((size_t)Bp * Lk * d * e, 256) + 1024;
}

// PT->w[i] = linears[i].weight transposed ([in, out] row-major, dtype `dtype`); PT->b is not read.
// dW[i] / db[i] (fp32, zero-initialised or holding a running sum) are accumulated into; any of d_query / d_key /
// d_value may be null (that input needs no gradient).
int bist_mha_bwd(const void* d_out, const void* query, const void* key, const void* value, const uint8_t* mask,
                 int mask_mode, const bist_mha_params_t* PT, const void* tape, size_t tape_bytes, void* d_query,
                 void* d_key, void* d_value, float* const* dW, float* const* db, int64_t Bp, int64_t Lq, int64_t Lk,
                 int64_t d, int64_t d_in, int64_t h, void* ws, size_t ws_bytes, int dtype, float p_drop, uint64_t seed,
                 uint32_t site, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(d_out && query && key && value && PT && tape && dW && db && ws, "mha_bwd: null pointer");
  BIST_CHECK_ARG(Bp > 0 && Lq > 0 && Lk > 0 && d > 0 && d_in > 0 && h > 0 && d % h == 0, "mha_bwd: bad shape");
  BIST_CHECK_ARG(dtype_ok(dtype), "mha_bwd: bad dtype");
  BIST_CHECK_ARG((mask_mode == BIST_MASK_NONE) == (mask == nullptr) && mask_mode >= 0 && mask_mode <= 2,
                 "mha_bwd: mask and mask_mode disagree");
  for (int i = 0; i < 4; ++i) BIST_CHECK_ARG(PT->w[i] && dW[i] && db[i], "mha_bwd: null parameter %d", i);
  const size_t e = dtype_size(dtype);
  MhaWs T(const_cast<void*>(tape), tape_bytes, Bp, Lq, Lk, d, h, e);
  if (!T.ok) { set_error("mha_bwd: tape too small (%zu bytes)", tape_bytes); return BIST_ERR_WORKSPACE; }
  Arena A(ws, ws_bytes);
  void* dctx = A.take((size_t)Bp * Lq * d * e);
  void* dq = A.take((size_t)Bp * Lq * d * e);
  void* dk = A.take((size_t)Bp * Lk * d * e);
  void* dv = A.take((size_t)Bp * Lk * d * e);
  if (!A.ok) { set_error("mha_bwd: workspace too small (%zu bytes)", ws_bytes); return BIST_ERR_WORKSPACE; }

  // modules.py:98-100 backwards: dctx = d_out W_o ; dW_o += d_out^T ctx ; db_o += colsum(d_out)
  BIST_TRY(bist_linear_bwd(d_out, T.ctx, PT->w[3], dctx, dW[3], db[3], Bp * Lq, d_in, d, dtype, dtype, dtype, dtype, st));

  // modules.py:54-64 backwards (probabilities recomputed from the saved q, k)
  AttnBwdArgs a;
  a.q = T.q; a.k = T.k; a.v = T.v; a.dctx = dctx; a.dq = dq; a.dk_ = dk; a.dv = dv;
  a.q_b = Lq * d; a.q_l = d; a.k_b = Lk * d; a.k_l = d; a.c_b = Lq * d; a.c_l = d;
  a.mask = mask;
  a.m_b = (mask_mode == BIST_MASK_FULL) ? Lq * Lk : Lk;
  a.m_l = (mask_mode == BIST_MASK_FULL) ? Lk : 0;
  a.Bp = (int)Bp; a.Lq = (int)Lq; a.Lk = (int)Lk; a.h = (int)h; a.dk = (int)(d / h);
  a.dtype = dtype;
  a.drop = make_drop(p_drop, seed, site);
  BIST_TRY(attn_core_bwd(a, st));

  // modules.py:89-91 backwards: the three input projections
  const void* in[3] = {query, key, value};
  const void* dproj[3] = {dq, dk, dv};
  void* din[3] = {d_query, d_key, d_value};
  const int64_t rows[3] = {Bp * Lq, Bp * Lk, Bp * Lk};
  for (int i = 0; i < 3; ++i)
    BIST_TRY(bist_linear_bwd(dproj[i], in[i], PT->w[i], din[i], dW[i], db[i], rows[i], d, d_in, dtype, dtype, dtype, dtype, st));
  return 0;
}

// hidden [rows, d_ff] (dtype) is the tape: relu(x W1^T + b1) after dropout (modules.py:113).
int bist_ffn_train_fwd(const void* x, const void* w1, const float* b1, const void* w2, const float* b2, void* y,
                       void* hidden, int64_t rows, int64_t d, int64_t d_ff, int64_t d_out, int dtype, float p_drop,
                       uint64_t seed, uint32_t site, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(x && w1 && w2 && y && hidden, "ffn(train): null pointer");
  BIST_CHECK_ARG(dtype_ok(dtype) && p_drop >= 0.f && p_drop < 1.f, "ffn(train): bad dtype or dropout probability");
  BIST_TRY(bist_linear_fwd(x, w1, b1, nullptr, hidden, rows, d_ff, d, 1, dtype, dtype, dtype, st));
  if (p_drop > 0.f)
    BIST_TRY(dropout_add(hidden, nullptr, rows, d_ff, 1, make_drop(p_drop, seed, site), 0, 0, dtype, st));
  return bist_linear_fwd(hidden, w2, b2, nullptr, y, rows, d_out, d_ff, 0, dtype, dtype, dtype, st);
}

// w1T [d, d_ff], w2T [d_ff, d_out] (dtype); ws >= rows*d_ff*sizeof(dtype); dW / db fp32, accumulated into; dx may be null.
int bist_ffn_bwd(const void* dy, const void* x, const void* w1T, const void* w2T, const void* hidden, void* dx, float* dW1,
                 float* db1, float* dW2, float* db2, int64_t rows, int64_t d, int64_t d_ff, int64_t d_out, void* ws,
                 size_t ws_bytes, int dtype, float p_drop, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(dy && x && w1T && w2T && hidden && dW1 && db1 && dW2 && db2 && ws, "ffn_bwd: null pointer");
  BIST_CHECK_ARG(dtype_ok(dtype) && p_drop >= 0.f && p_drop < 1.f, "ffn_bwd: bad dtype or dropout probability");
  if (ws_bytes < (size_t)rows * d_ff * dtype_size(dtype)) { set_error("ffn_bwd: workspace too small"); return BIST_ERR_WORKSPACE; }
  // d(hidden') = dy W2 ; dW2 += dy^T hidden' ; then through the dropout multiplier and the ReLU gate in one pass
  BIST_TRY(bist_linear_bwd(dy, hidden, w2T, ws, dW2, db2, rows, d_out, d_ff, dtype, dtype, dtype, dtype, st));
  BIST_TRY(relu_gate(ws, hidden, rows * d_ff, dtype, st, p_drop > 0.f ? 1.f / (1.f - p_drop) : 1.f));
  return bist_linear_bwd(ws, x, w1T, dx, dW1, db1, rows, d_ff, d, dtype, dtype, dtype, dtype, st);
}

}  // extern "C"
