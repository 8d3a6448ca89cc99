// Backward-pass kernels of the BiST video hot path (training step).
//   wgrad           weight / bias gradients: C[n,k] += sum_m A[m,n] B[m,k]   (nn.Linear backward w.r.t. weight)
//   layer_norm_bwd  LayerNorm backward (modules.py:28-31: unbiased std, eps added to the std)
//   attn_core_bwd   attention() backward with projected q,k,v (modules.py:54-64)       -> query self-attention
//   stage1_bwd      softmax / weighted-sum backward of stage 1 over the folded scores   (encoder.py:111-121 / 143-148)
//   stage2_bwd      stage-2 backward, projections folded to the query side            (encoder.py:125-132 / 152-161)
//   relu_gate, group_sum, cast_copy   small element-wise helpers
// All probabilities are recomputed from saved scores / operands; nothing here allocates.
#include <math.h>

#include "common.cuh"

namespace bist {
namespace {

constexpr float kMaskFill = -1e9f;  // modules.py:60

template <typename T>
__device__ __forceinline__ float ldf(const T* p) { return to_f32(*p); }

// ------------------------------------------------------------------------------------------------ wgrad (CUDA cores)
// grid (tiles_n * tiles_k, splits, Z); block 256 = 16 x 16 threads, 4 x 4 outputs each; tile 64 x 64, 16 rows per step.
constexpr int WG_T = 64, WG_R = 16;

template <typename TA, typename TB>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(WgradArgs w, int tiles_k, int rows_per_split) {
  __shared__ __align__(16) float As[2][WG_R][WG_T];
  __shared__ __align__(16) float Bs[2][WG_R][WG_T];
  pdl_trigger();
  pdl_wait();
  const int z = blockIdx.z;
  const int tn = blockIdx.x / tiles_k, tk = blockIdx.x % tiles_k;
  const int n0 = tn * WG_T, k0 = tk * WG_T;
  const int m_begin = blockIdx.y * rows_per_split;
  const int m_end = min(w.M, m_begin + rows_per_split);
  if (m_begin >= m_end) return;
  const TA* A = reinterpret_cast<const TA*>(w.A) + w.az.off(z, w.zin, 1);
  const TB* B = reinterpret_cast<const TB*>(w.B) + w.bz.off(z, w.zin, 1);
  const int t = threadIdx.x;
  const int lc = t & 63, lr = t >> 6;  // loader: column lc, rows lr + 4*i
  const int ty = t >> 4, tx = t & 15;  // compute: n = n0 + ty*4 + i, k = k0 + tx*4 + j
  const bool do_cs = w.colsum != nullptr && tk == 0;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float cs = 0.f;
  float ra[4], rb[4];
  auto gload = [&](int m0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + lr + 4 * i;
      ra[i] = (m < m_end && n0 + lc < w.N) ? ldf(A + (long long)m * w.lda + n0 + lc) : 0.f;
      rb[i] = (m < m_end && k0 + lc < w.K) ? ldf(B + (long long)m * w.ldb + k0 + lc) : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[buf][lr + 4 * i][lc] = ra[i];
      Bs[buf][lr + 4 * i][lc] = rb[i];
    }
  };
  gload(m_begin);
  sstore(0);
  __syncthreads();
  const int steps = (m_end - m_begin + WG_R - 1) / WG_R;
  for (int s = 0; s < steps; ++s) {
    const int buf = s & 1;
    if (s + 1 < steps) gload(m_begin + (s + 1) * WG_R);
#pragma unroll
    for (int mm = 0; mm < WG_R; ++mm) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[buf][mm][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[buf][mm][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (do_cs && t < WG_T) {
#pragma unroll
      for (int mm = 0; mm < WG_R; ++mm) cs += As[buf][mm][t];
    }
    if (s + 1 < steps) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }
  float* C = w.C + w.cz.off(z, w.zin, 1);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int n = n0 + ty * 4 + i;
    if (n >= w.N) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + tx * 4 + j;
      if (k < w.K) atomicAdd(C + (long long)n * w.ldc + k, w.alpha * acc[i][j]);
    }
  }
  if (do_cs && t < WG_T && n0 + t < w.N) atomicAdd(w.colsum + w.sz.off(z, w.zin, 1) + n0 + t, cs);
}

// ------------------------------------------------------------------------------------------------ wgrad (tensor cores)
// bf16 x bf16 (B may also be fp32, converted while staging) -> fp32, mma.sync.m16n8k16.  The contraction runs over
// the rows of both operands, so both fragments come out of row-major smem tiles through ldmatrix.trans.
// CTA tile 128 (n) x 128 (k), 32 rows per step, 8 warps as 4 (n) x 2 (k): warp tile 32 x 64.
constexpr int WM_TN = 128, WM_TK = 128, WM_R = 32, WM_LD = 136;  // smem row stride in bf16 (272 B: conflict-free ldmatrix)

__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint4 load8_bf16(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ uint4 load8_bf16(const float* p) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  __nv_bfloat162 v0 = __floats2bfloat162_rn(a.x, a.y), v1 = __floats2bfloat162_rn(a.z, a.w);
  __nv_bfloat162 v2 = __floats2bfloat162_rn(b.x, b.y), v3 = __floats2bfloat162_rn(b.z, b.w);
  uint4 u;
  u.x = *reinterpret_cast<uint32_t*>(&v0); u.y = *reinterpret_cast<uint32_t*>(&v1);
  u.z = *reinterpret_cast<uint32_t*>(&v2); u.w = *reinterpret_cast<uint32_t*>(&v3);
  return u;
}

template <typename TB>
__global__ void __launch_bounds__(256) wgrad_mma_kernel(WgradArgs w, int tiles_k, int rows_per_split) {
  __shared__ __align__(16) __nv_bfloat16 As[2][WM_R][WM_LD];
  __shared__ __align__(16) __nv_bfloat16 Bs[2][WM_R][WM_LD];
  pdl_trigger();
  pdl_wait();
  const int z = blockIdx.z;
  const int tn = blockIdx.x / tiles_k, tk = blockIdx.x % tiles_k;
  const int n0 = tn * WM_TN, k0 = tk * WM_TK;
  const int m_begin = blockIdx.y * rows_per_split;
  const int m_end = min(w.M, m_begin + rows_per_split);
  if (m_begin >= m_end) return;
  const __nv_bfloat16* A = reinterpret_cast<const __nv_bfloat16*>(w.A) + w.az.off(z, w.zin, 1);
  const TB* B = reinterpret_cast<const TB*>(w.B) + w.bz.off(z, w.zin, 1);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int wn = warp >> 1, wk = warp & 1;  // warp tile: n in [wn*32, +32), k in [wk*64, +64)
  // loader: 32 rows x 128 cols = 512 chunks of 8 elements; thread t takes chunks t and t + 256
  const int lrow = t >> 4, lcol = (t & 15) * 8;
  const bool do_cs = w.colsum != nullptr && tk == 0;
  uint4 ra[2], rb[2];
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  auto gload = [&](int m0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int m = m0 + lrow + 16 * i;
      // N and K are multiples of 8 (checked by the host), so a chunk is either fully inside or fully outside
      ra[i] = (m < m_end && n0 + lcol < w.N) ? load8_bf16(A + (long long)m * w.lda + n0 + lcol) : zero4;
      rb[i] = (m < m_end && k0 + lcol < w.K) ? load8_bf16(B + (long long)m * w.ldb + k0 + lcol) : zero4;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      *reinterpret_cast<uint4*>(&As[buf][lrow + 16 * i][lcol]) = ra[i];
      *reinterpret_cast<uint4*>(&Bs[buf][lrow + 16 * i][lcol]) = rb[i];
    }
  };
  float acc[2][8][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;
  float cs = 0.f;
  gload(m_begin);
  sstore(0);
  __syncthreads();
  const int steps = (m_end - m_begin + WM_R - 1) / WM_R;
  const int lm = lane & 7, lq = lane >> 3;  // ldmatrix: lane supplies row lm of 8x8 matrix number lq
  for (int s = 0; s < steps; ++s) {
    const int buf = s & 1;
    if (s + 1 < steps) gload(m_begin + (s + 1) * WM_R);
#pragma unroll
    for (int ks = 0; ks < WM_R / 16; ++ks) {
      // A fragments (mma "row" operand [16 n x 16 m]) from As[m][n] via .trans:
      //   matrix 0: n 0-7,  m 0-7 ; 1: n 8-15, m 0-7 ; 2: n 0-7, m 8-15 ; 3: n 8-15, m 8-15
      uint32_t af[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int mrow = ks * 16 + (lq >> 1) * 8 + lm;
        const int ncol = wn * 32 + i * 16 + (lq & 1) * 8;
        ldsm_x4_t((uint32_t)__cvta_generic_to_shared(&As[buf][mrow][ncol]), af[i][0], af[i][1], af[i][2], af[i][3]);
      }
      // B fragments (mma "col" operand [16 m x 8 k]) from Bs[m][k] via .trans: matrices (m 0-7 | m 8-15) x (k 0-7 | k 8-15)
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        const int mrow = ks * 16 + (lq & 1) * 8 + lm;
        const int kcol = wk * 64 + jp * 16 + (lq >> 1) * 8;
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t((uint32_t)__cvta_generic_to_shared(&Bs[buf][mrow][kcol]), b0, b1, b2, b3);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          mma_bf16(acc[i][jp * 2], af[i], b0, b1);
          mma_bf16(acc[i][jp * 2 + 1], af[i], b2, b3);
        }
      }
    }
    if (do_cs && t < WM_TN) {
#pragma unroll 8
      for (int mm = 0; mm < WM_R; ++mm) cs += __bfloat162float(As[buf][mm][t]);
    }
    if (s + 1 < steps) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }
  float* C = w.C + w.cz.off(z, w.zin, 1);
  const int gr = lane >> 2, gc = (lane & 3) * 2;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int n = n0 + wn * 32 + i * 16 + gr + (q >> 1) * 8;
        const int k = k0 + wk * 64 + j * 8 + gc + (q & 1);
        if (n < w.N && k < w.K) atomicAdd(C + (long long)n * w.ldc + k, w.alpha * acc[i][j][q]);
      }
  if (do_cs && t < WM_TN && n0 + t < w.N) atomicAdd(w.colsum + w.sz.off(z, w.zin, 1) + n0 + t, cs);
}

int num_sms_cached() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <typename TA, typename TB>
int launch_wgrad_simt(const WgradArgs& w, cudaStream_t st) {
  const int tiles_n = (w.N + WG_T - 1) / WG_T, tiles_k = (w.K + WG_T - 1) / WG_T;
  const long long base = (long long)tiles_n * tiles_k * w.Z;
  long long splits = (2LL * num_sms_cached() + base - 1) / base;
  const long long max_splits = (w.M + 4 * WG_R - 1) / (4 * WG_R);
  splits = splits < 1 ? 1 : (splits > max_splits ? max_splits : splits);
  if (splits > 65535) splits = 65535;
  int rows = (int)((w.M + splits - 1) / splits);
  rows = (rows + WG_R - 1) / WG_R * WG_R;
  splits = (w.M + rows - 1) / rows;
  dim3 grid(tiles_n * tiles_k, (unsigned)splits, w.Z);
  ProfScope ps(w.tag, st);
  BIST_LAUNCH((wgrad_simt_kernel<TA, TB>), grid, 256, 0, st, w, tiles_k, rows);
  BIST_LAUNCHED();
  return 0;
}

template <typename TB>
int launch_wgrad_mma(const WgradArgs& w, cudaStream_t st) {
  const int tiles_n = (w.N + WM_TN - 1) / WM_TN, tiles_k = (w.K + WM_TK - 1) / WM_TK;
  const long long base = (long long)tiles_n * tiles_k * w.Z;
  long long splits = (1LL * num_sms_cached() + base - 1) / base;   // every split costs a full tile of reductions
  const long long max_splits = (w.M + 2 * WM_R - 1) / (2 * WM_R);
  splits = splits < 1 ? 1 : (splits > max_splits ? max_splits : splits);
  if (splits > 65535) splits = 65535;
  int rows = (int)((w.M + splits - 1) / splits);
  rows = (rows + WM_R - 1) / WM_R * WM_R;
  splits = (w.M + rows - 1) / rows;
  dim3 grid(tiles_n * tiles_k, (unsigned)splits, w.Z);
  ProfScope ps(w.tag, st);
  BIST_LAUNCH((wgrad_mma_kernel<TB>), grid, 256, 0, st, w, tiles_k, rows);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace

int wgrad(const WgradArgs& w, cudaStream_t st) {
  BIST_CHECK_ARG(w.A && w.B && w.C && w.M > 0 && w.N > 0 && w.K > 0 && w.Z > 0 && w.Z <= 65535, "wgrad: bad argument");
  {
    bool taken = false;
    int rc = wgrad_tc(w, st, &taken);
    if (taken) return rc;
  }
  static const bool no_mma = getenv("BIST_DISABLE_TC") != nullptr;
  // tensor-core path: bf16 A, bf16 or fp32 B, 16-byte aligned 8-element chunks everywhere
  auto al = [](long long v, int esz) { return (v * esz) % 16 == 0; };
  const int bsz = w.b_dtype == BIST_BF16 ? 2 : 4;
  bool mma_ok = !no_mma && w.a_dtype == BIST_BF16 && w.N % 8 == 0 && w.K % 8 == 0 && w.M >= 64 &&
                ((uintptr_t)w.A % 16 == 0) && ((uintptr_t)w.B % 16 == 0) && al(w.lda, 2) && al(w.ldb, bsz) &&
                al(w.az.s0, 2) && al(w.az.s1, 2) && al(w.bz.s0, bsz) && al(w.bz.s1, bsz);
  if (mma_ok) return w.b_dtype == BIST_BF16 ? launch_wgrad_mma<__nv_bfloat16>(w, st) : launch_wgrad_mma<float>(w, st);
  if (w.a_dtype == BIST_F32 && w.b_dtype == BIST_F32) return launch_wgrad_simt<float, float>(w, st);
  if (w.a_dtype == BIST_BF16 && w.b_dtype == BIST_BF16) return launch_wgrad_simt<__nv_bfloat16, __nv_bfloat16>(w, st);
  if (w.a_dtype == BIST_BF16 && w.b_dtype == BIST_F32) return launch_wgrad_simt<__nv_bfloat16, float>(w, st);
  return launch_wgrad_simt<float, __nv_bfloat16>(w, st);
}

// ------------------------------------------------------------------------------------------------ LayerNorm backward
namespace {

// grid: ceil(rows / rows_per_cta) CTAs inside one parameter group each; block 256 (8 warps, one row per warp at a time).
template <typename TDY, typename TX, typename TDX>
__global__ void __launch_bounds__(256) ln_bwd_kernel(LnBwdArgs a, int rows_per_cta, int ctas_per_group) {
  extern __shared__ float sacc[];  // [2][d]: da, db partial sums of this CTA
  pdl_trigger();
  pdl_wait();
  const int d = (int)a.d;
  const int grp = blockIdx.x / ctas_per_group, cg = blockIdx.x % ctas_per_group;
  const int64_t grows = a.group_rows > 0 ? a.group_rows : a.rows;
  const int64_t r_begin = grp * grows + (int64_t)cg * rows_per_cta;
  const int64_t r_end = min(grp * grows + grows, r_begin + rows_per_cta);
  const bool want_p = a.da != nullptr;
  if (want_p)
    for (int i = threadIdx.x; i < 2 * d; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* a2 = a.a2 + grp * a.pstride;
  for (int64_t r = r_begin + warp; r < r_end; r += 8) {
    const TX* x = reinterpret_cast<const TX*>(a.x) + r * d;
    const TDY* dy = reinterpret_cast<const TDY*>(a.dy) + r * d;
    float s = 0.f;
    for (int c = lane; c < d; c += 32) s += to_f32(x[c]);
    const float mu = warp_sum(s) / (float)d;
    float ss = 0.f, sg = 0.f, sgc = 0.f;
    for (int c = lane; c < d; c += 32) {
      const float cc = to_f32(x[c]) - mu, g = to_f32(dy[c]) * a2[c];
      ss = fmaf(cc, cc, ss);
      sg += g;
      sgc = fmaf(g, cc, sgc);
    }
    ss = warp_sum(ss); sg = warp_sum(sg); sgc = warp_sum(sgc);
    const float sd = sqrtf(ss / (float)(d - 1));
    const float rinv = 1.f / (sd + a.eps);
    const float mg = sg / (float)d;
    const float k2 = sd > 0.f ? rinv * rinv * sgc / ((float)(d - 1) * sd) : 0.f;
    TDX* dx = reinterpret_cast<TDX*>(a.dx) + r * d;
    for (int c = lane; c < d; c += 32) {
      const float xv = to_f32(x[c]), cc = xv - mu, dyv = to_f32(dy[c]);
      float v = rinv * (dyv * a2[c] - mg) - k2 * cc;
      if (a.relu_gate && !(xv > 0.f)) v = 0.f;
      if (a.accumulate) v += to_f32(dx[c]);
      dx[c] = from_f32<TDX>(v);
      if (want_p) {
        atomicAdd(&sacc[c], dyv * cc * rinv);
        atomicAdd(&sacc[d + c], dyv);
      }
    }
  }
  if (want_p) {
    __syncthreads();
    float* da = a.da + grp * a.gstride;
    float* db = a.db + grp * a.gstride;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
      atomicAdd(da + i, sacc[i]);
      atomicAdd(db + i, sacc[d + i]);
    }
  }
}

// d % 4 == 0, d <= 128 * MAXV: a row lives in registers (one vector load of x and dy per lane and column group), and the
// lane keeps the gain/offset gradients of its own columns in registers over all rows of the warp.
template <typename TDY, typename TX, typename TDX, int MAXV>
__global__ void __launch_bounds__(256) ln_bwd_vec_kernel(LnBwdArgs a, int rows_per_cta, int ctas_per_group) {
  extern __shared__ float sacc[];  // [8 warps][2][d]: per-warp partial sums, added up without atomics
  pdl_trigger();
  pdl_wait();
  const int d = (int)a.d, nv = d >> 2;
  const int grp = blockIdx.x / ctas_per_group, cg = blockIdx.x % ctas_per_group;
  const int64_t grows = a.group_rows > 0 ? a.group_rows : a.rows;
  const int64_t r_begin = grp * grows + (int64_t)cg * rows_per_cta;
  const int64_t r_end = min(grp * grows + grows, r_begin + rows_per_cta);
  const bool want_p = a.da != nullptr;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* a2 = a.a2 + grp * a.pstride;
  float4 ga[MAXV], pa[MAXV], pb[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + i * 32;
    ga[i] = c < nv ? *reinterpret_cast<const float4*>(a2 + c * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    pa[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int64_t r = r_begin + warp; r < r_end; r += 8) {
    const TX* x = reinterpret_cast<const TX*>(a.x) + r * d;
    const TDY* dy = reinterpret_cast<const TDY*>(a.dy) + r * d;
    float4 xv[MAXV], gv[MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        xv[i] = ld4(x + c * 4);
        gv[i] = ld4(dy + c * 4);
        s += (xv[i].x + xv[i].y) + (xv[i].z + xv[i].w);
      }
    }
    const float mu = warp_sum(s) / (float)d;
    float ss = 0.f, sg = 0.f, sgc = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        const float c0 = xv[i].x - mu, c1 = xv[i].y - mu, c2 = xv[i].z - mu, c3 = xv[i].w - mu;
        const float g0 = gv[i].x * ga[i].x, g1 = gv[i].y * ga[i].y, g2 = gv[i].z * ga[i].z, g3 = gv[i].w * ga[i].w;
        ss += (c0 * c0 + c1 * c1) + (c2 * c2 + c3 * c3);
        sg += (g0 + g1) + (g2 + g3);
        sgc += (g0 * c0 + g1 * c1) + (g2 * c2 + g3 * c3);
      }
    }
    ss = warp_sum(ss); sg = warp_sum(sg); sgc = warp_sum(sgc);
    const float sd = sqrtf(ss / (float)(d - 1));
    const float rinv = 1.f / (sd + a.eps);
    const float mg = sg / (float)d;
    const float k2 = sd > 0.f ? rinv * rinv * sgc / ((float)(d - 1) * sd) : 0.f;
    TDX* dx = reinterpret_cast<TDX*>(a.dx) + r * d;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        const float c0 = xv[i].x - mu, c1 = xv[i].y - mu, c2 = xv[i].z - mu, c3 = xv[i].w - mu;
        float4 v = make_float4(rinv * (gv[i].x * ga[i].x - mg) - k2 * c0, rinv * (gv[i].y * ga[i].y - mg) - k2 * c1,
                               rinv * (gv[i].z * ga[i].z - mg) - k2 * c2, rinv * (gv[i].w * ga[i].w - mg) - k2 * c3);
        if (a.relu_gate) {
          if (!(xv[i].x > 0.f)) v.x = 0.f;
          if (!(xv[i].y > 0.f)) v.y = 0.f;
          if (!(xv[i].z > 0.f)) v.z = 0.f;
          if (!(xv[i].w > 0.f)) v.w = 0.f;
        }
        if (a.accumulate) {
          const float4 o = ld4(dx + c * 4);
          v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
        }
        st4(dx + c * 4, v);
        pa[i].x += gv[i].x * c0 * rinv; pa[i].y += gv[i].y * c1 * rinv; pa[i].z += gv[i].z * c2 * rinv; pa[i].w += gv[i].w * c3 * rinv;
        pb[i].x += gv[i].x; pb[i].y += gv[i].y; pb[i].z += gv[i].z; pb[i].w += gv[i].w;
      }
    }
  }
  if (want_p) {
    // (shared-memory atomics from eight warps onto the same 2*d floats were most of this kernel's time)
    float4* mine = reinterpret_cast<float4*>(sacc + (size_t)warp * 2 * d);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + i * 32;
      if (c < nv) { mine[c] = pa[i]; mine[nv + c] = pb[i]; }
    }
    __syncthreads();
    float* da = a.da + grp * a.gstride;
    float* db = a.db + grp * a.gstride;
    const bool vec = (((uintptr_t)da | (uintptr_t)db) & 15) == 0;
    for (int i = threadIdx.x; i < 2 * nv; i += blockDim.x) {   // float4 i of [da | db]
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const float4 v = reinterpret_cast<const float4*>(sacc + (size_t)w * 2 * d)[i];
        t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
      }
      float* dst = i < nv ? da + 4 * i : db + 4 * (i - nv);
      if (vec) atomicAdd(reinterpret_cast<float4*>(dst), t);   // 16-byte vector reductions: a quarter of the L2 atomic operations
      else { atomicAdd(dst, t.x); atomicAdd(dst + 1, t.y); atomicAdd(dst + 2, t.z); atomicAdd(dst + 3, t.w); }
    }
  }
}

template <typename TDY, typename TX, typename TDX>
int launch_ln_bwd(const LnBwdArgs& a, cudaStream_t st) {
  const int64_t grows = a.group_rows > 0 ? a.group_rows : a.rows;
  const int64_t groups = a.rows / grows;
  // enough CTAs to fill the machine, at least 8 rows (one per warp) each
  int64_t per = (a.rows + 4LL * num_sms_cached() - 1) / (4LL * num_sms_cached());
  per = per < 8 ? 8 : (per + 7) / 8 * 8;
  const bool vec = a.d % 4 == 0 && a.d <= 512 && ((uintptr_t)a.x % 16 == 0) && ((uintptr_t)a.dy % 16 == 0) && ((uintptr_t)a.dx % 16 == 0) &&
                   ((uintptr_t)a.a2 % 16 == 0) && a.pstride % 4 == 0;
  if (vec && per < 32) per = 32;   // four rows per warp: the per-lane gradient sums amortise their atomics
  const int64_t cpg = (grows + per - 1) / per;
  ProfScope ps("ln_bwd", st);
  if (vec)
    BIST_LAUNCH((ln_bwd_vec_kernel<TDY, TX, TDX, 4>), (unsigned)(groups * cpg), 256, 8 * 2 * a.d * sizeof(float), st, a, (int)per, (int)cpg);
  else
    BIST_LAUNCH((ln_bwd_kernel<TDY, TX, TDX>), (unsigned)(groups * cpg), 256, 2 * a.d * sizeof(float), st, a, (int)per, (int)cpg);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace

int layer_norm_bwd(const LnBwdArgs& a, cudaStream_t st) {
  BIST_CHECK_ARG(a.dy && a.x && a.a2 && a.dx && a.rows > 0 && a.d > 1 && a.d <= 8192, "ln_bwd: bad argument");
  BIST_CHECK_ARG((a.da == nullptr) == (a.db == nullptr), "ln_bwd: da and db go together");
  const int64_t grows = a.group_rows > 0 ? a.group_rows : a.rows;
  BIST_CHECK_ARG(a.rows % grows == 0, "ln_bwd: rows must be a multiple of group_rows");
  const int key = a.dy_dtype * 4 + a.x_dtype * 2 + a.dx_dtype;
  switch (key) {
    case 0: return launch_ln_bwd<float, float, float>(a, st);
    case 1: return launch_ln_bwd<float, float, __nv_bfloat16>(a, st);
    case 2: return launch_ln_bwd<float, __nv_bfloat16, float>(a, st);
    case 3: return launch_ln_bwd<float, __nv_bfloat16, __nv_bfloat16>(a, st);
    case 4: return launch_ln_bwd<__nv_bfloat16, float, float>(a, st);
    case 5: return launch_ln_bwd<__nv_bfloat16, float, __nv_bfloat16>(a, st);
    case 6: return launch_ln_bwd<__nv_bfloat16, __nv_bfloat16, float>(a, st);
    default: return launch_ln_bwd<__nv_bfloat16, __nv_bfloat16, __nv_bfloat16>(a, st);
  }
}

// ------------------------------------------------------------------------------------------------ element-wise helpers
namespace {

template <typename T>
__global__ void relu_gate_kernel(T* __restrict__ dy, const T* __restrict__ y, int64_t n, float scale) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (!(to_f32(y[i]) > 0.f)) dy[i] = from_f32<T>(0.f);
    else if (scale != 1.f) dy[i] = from_f32<T>(to_f32(dy[i]) * scale);
  }
}

// Index arithmetic in 32 bits whenever the tensor has fewer than 2^31 elements (64-bit division costs ~100 instructions).
template <typename I>
__device__ __forceinline__ I ref_row(I r, int perm_L, int perm_G) {
  if (perm_L <= 0) return r;
  const I g = r % (I)perm_G, bl = r / (I)perm_G;
  const I l = bl % (I)perm_L, b = bl / (I)perm_L;
  return (b * (I)perm_G + g) * (I)perm_L + l;
}

// Four consecutive elements per thread (d % 4 == 0): one row/column split and one vector access per four hashes.
template <typename T, typename I>
__global__ void dropout_add_kernel(T* __restrict__ out, const float* __restrict__ res, int64_t rows, int d, int64_t res_div,
                                   DropCfg drop, int perm_L, int perm_G, int64_t row0) {
  pdl_trigger();
  pdl_wait();
  const I d4 = (I)(d >> 2);
  const I total = (I)rows * d4;
  for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
    const I r = i / d4;
    const int c = (int)(i - r * d4) * 4;
    const unsigned long long ib = (unsigned long long)((int64_t)(ref_row<I>(r, perm_L, perm_G)) + row0) * d + c;
    float4 v = ld4(out + (int64_t)r * d + c);
    v.x *= drop_mul(drop, ib); v.y *= drop_mul(drop, ib + 1); v.z *= drop_mul(drop, ib + 2); v.w *= drop_mul(drop, ib + 3);
    if (res) {
      const float4 x = ld4(res + (int64_t)(r / (I)res_div) * d + c);
      v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
    }
    st4(out + (int64_t)r * d + c, v);
  }
}

template <typename TS, typename TD, typename I>
__global__ void dropout_scale_kernel(const TS* __restrict__ src, TD* __restrict__ dst, int64_t rows, int d, DropCfg drop, int perm_L,
                                     int perm_G, int64_t row0) {
  pdl_trigger();
  pdl_wait();
  const I d4 = (I)(d >> 2);
  const I total = (I)rows * d4;
  for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
    const I r = i / d4;
    const int c = (int)(i - r * d4) * 4;
    float4 v = ld4(src + (int64_t)r * d + c);
    if (drop.thresh != 0u) {
      const unsigned long long ib = (unsigned long long)((int64_t)(ref_row<I>(r, perm_L, perm_G)) + row0) * d + c;
      v.x *= drop_mul(drop, ib); v.y *= drop_mul(drop, ib + 1); v.z *= drop_mul(drop, ib + 2); v.w *= drop_mul(drop, ib + 3);
    }
    st4(dst + (int64_t)r * d + c, v);
  }
}

inline unsigned ew_grid_fwd(int64_t n) {   // blocks of 256 threads, capped at eight per SM (grid-stride loops)
  int64_t g = (n + 255) / 256;
  const int64_t cap = 8LL * num_sms_cached();
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

// Row-structured form of the two kernels above for tensors below 2^32 elements whose rows are a power-of-two number (<= 256) of
// 8-element chunks: a thread owns one chunk column and walks down the rows, so the row split, the reference-row permutation and the
// residual's row are computed once per EIGHT elements, the hash key (seed, step counter, site) once per thread, and the accesses are
// 16 bytes wide.  The four-elements-per-thread kernels spend ~250 instructions per 8 bytes (five 32-bit divisions, four reads of the
// step counter, 64-bit hash prologues) and ran at 2.2 TB/s of traffic on the [B Lq G, d] stage-1 outputs.
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) { v[2 * j] = __uint_as_float(w[j] << 16); v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u); }
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
  __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
  u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
  u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
  *reinterpret_cast<uint4*>(p) = u;
}
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) dropout_rows_kernel(const TS* __restrict__ src, TD* __restrict__ dst, const float* __restrict__ res,
                                                           uint32_t rows, uint32_t d, uint32_t res_div, DropCfg drop, int perm_L, int perm_G,
                                                           uint32_t row0, int cpr_shift) {
  pdl_trigger();
  pdl_wait();
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t c = (tid & ((1u << cpr_shift) - 1u)) * 8u, rstep = (gridDim.x * blockDim.x) >> cpr_shift;
  const bool on = drop.thresh != 0u;
  uint32_t key = 0u;
  if (on) {   // drop_hash() with the high half of the element index equal to zero
    const unsigned long long seed = drop.step ? drop.seed + __ldg(drop.step) : drop.seed;
    key = mix32((uint32_t)(seed >> 32) + drop.site * 0x85EBCA6Bu) ^ (uint32_t)seed;
  }
  for (uint32_t r = tid >> cpr_shift; r < rows; r += rstep) {
    float v[8];
    ld8(src + (size_t)r * d + c, v);
    if (on) {
      const uint32_t ib = (ref_row<uint32_t>(r, perm_L, perm_G) + row0) * d + c;
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] *= mix32((ib + (uint32_t)e) ^ key) >= drop.thresh ? drop.scale : 0.f;
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] *= drop.scale;   // (thresh == 0: drop_mul returns the scale, 1 for a disabled site)
    }
    if (res) {
      float x[8];
      ld8(res + (size_t)(r / res_div) * d + c, x);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] += x[e];
    }
    st8(dst + (size_t)r * d + c, v);
  }
}
// shape test + launch; *taken = false: left to the element-wise kernels
template <typename TS, typename TD>
int launch_dropout_rows(const void* src, void* dst, const float* res, int64_t rows, int64_t d, int64_t res_div, const DropCfg& drop,
                        int perm_L, int perm_G, int64_t row0, cudaStream_t st, bool* taken) {
  *taken = false;
  if (d % 8 != 0 || d > 2048 || ((d / 8) & (d / 8 - 1)) != 0) return 0;
  if ((rows + row0) * d >= (1LL << 32) || rows * d >= (1LL << 31) || rows >= (1LL << 31) || res_div >= (1LL << 31)) return 0;
  if (((uintptr_t)src | (uintptr_t)dst | (uintptr_t)res) % 16 != 0) return 0;
  *taken = true;
  int shift = 0;
  while ((1 << shift) < d / 8) ++shift;
  const unsigned grid = ew_grid_fwd(rows * d / 8);
  BIST_LAUNCH((dropout_rows_kernel<TS, TD>), grid, 256, 0, st, (const TS*)src, (TD*)dst, res, (uint32_t)rows, (uint32_t)d, (uint32_t)res_div,
              drop, perm_L, perm_G, (uint32_t)row0, shift);
  return 0;
}

template <typename T>
__global__ void bias_rsum_fwd_kernel(T* __restrict__ c2, const float* __restrict__ bv, const float* __restrict__ rsum, int64_t Q, int h,
                                     int dk) {
  pdl_trigger();
  pdl_wait();
  const int d = h * dk;
  const int64_t total = Q * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t q = i / d;
    const int c = (int)(i % d);
    c2[i] = from_f32<T>(to_f32(c2[i]) + bv[c] * (rsum[q * h + c / dk] - 1.f));
  }
}

// grid: ceil(Q / 32) CTAs of 256 threads (8 warps x 4 rows each); per-CTA dbv partial sums in smem
template <typename T>
__global__ void __launch_bounds__(256) bias_rsum_bwd_kernel(const T* __restrict__ dc2, const float* __restrict__ bv,
                                                            const float* __restrict__ rsum, float* __restrict__ drsum,
                                                            float* __restrict__ dbv, int64_t Q, int h, int dk) {
  extern __shared__ float sdb[];  // [d]
  pdl_trigger();
  pdl_wait();
  const int d = h * dk;
  for (int i = threadIdx.x; i < d; i += blockDim.x) sdb[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int rr = 0; rr < 4; ++rr) {
    const int64_t q = (int64_t)blockIdx.x * 32 + warp * 4 + rr;
    if (q >= Q) break;
    for (int hh = 0; hh < h; ++hh) {
      const float rs = rsum[q * h + hh];
      float part = 0.f;
      for (int j = lane; j < dk; j += 32) {
        const float g = to_f32(dc2[q * d + hh * dk + j]);
        part = fmaf(g, bv[hh * dk + j], part);
        atomicAdd(&sdb[hh * dk + j], g * rs);
      }
      part = warp_sum(part);
      if (lane == 0) drsum[q * h + hh] = part;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) atomicAdd(dbv + i, sdb[i]);
}

// Vector form of bias_rsum_bwd_kernel (d % 8 == 0, d <= 1024, dk / 8 a power of two <= 32): a lane owns up to four 8-column chunks
// for all the rows of its warp and keeps their db_v partial sums in registers; the per-head dot product with b_v is reduced over
// the dk / 8 lanes of the head with shuffles.  (One shared-memory atomic per element and 64 CTAs made the scalar form 49 us on
// 2048 x 512.)
template <typename T>
__global__ void __launch_bounds__(256) bias_rsum_bwd_vec_kernel(const T* __restrict__ dc2, const float* __restrict__ bv,
                                                                const float* __restrict__ rsum, float* __restrict__ drsum,
                                                                float* __restrict__ dbv, int64_t Q, int h, int dk, int rows_per_cta) {
  extern __shared__ float sdb[];  // [d]
  pdl_trigger();
  pdl_wait();
  const int d = h * dk, nch = d >> 3, lph = dk >> 3;   // chunks per row, lanes per head
  for (int i = threadIdx.x; i < d; i += blockDim.x) sdb[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float acc[4][8], bvr[4][8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ch = lane + 32 * k;
#pragma unroll
    for (int e = 0; e < 8; ++e) { acc[k][e] = 0.f; bvr[k][e] = ch < nch ? bv[ch * 8 + e] : 0.f; }
  }
  const int64_t q_end = min(Q, ((int64_t)blockIdx.x + 1) * rows_per_cta);
  for (int64_t q = (int64_t)blockIdx.x * rows_per_cta + warp; q < q_end; q += 8) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int ch = lane + 32 * k;
      if (ch - lane < nch) {                       // warp-uniform: the shuffles below need all lanes
        float g[8], part = 0.f;
        int hh = 0;
        if (ch < nch) {
          ld8(dc2 + q * d + ch * 8, g);
          hh = (ch * 8) / dk;
          const float rs = rsum[q * h + hh];
#pragma unroll
          for (int e = 0; e < 8; ++e) { part = fmaf(g[e], bvr[k][e], part); acc[k][e] = fmaf(g[e], rs, acc[k][e]); }
        }
        for (int o = 1; o < lph; o <<= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        if (ch < nch && (lane & (lph - 1)) == 0) drsum[q * h + hh] = part;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ch = lane + 32 * k;
    if (ch < nch) {
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(&sdb[ch * 8 + e], acc[k][e]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) atomicAdd(dbv + i, sdb[i]);
}

__global__ void dropout_mask_kernel(float* __restrict__ out, int64_t n, DropCfg drop) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = drop_mul(drop, (unsigned long long)i);
}

template <typename T>
__global__ void group_sum_kernel(const T* __restrict__ in, float* __restrict__ out, int64_t Qrows, int G, int d) {
  pdl_trigger();
  pdl_wait();
  const int64_t total = Qrows * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t q = i / d;
    const int c = (int)(i % d);
    const T* p = in + q * G * d + c;
    float s = 0.f;
    for (int g = 0; g < G; ++g) s += to_f32(p[(int64_t)g * d]);
    out[i] += s;
  }
}

template <typename TS, typename TD>
__global__ void cast_kernel(const TS* __restrict__ src, TD* __restrict__ dst, int64_t n) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = from_f32<TD>(to_f32(src[i]));
}

// eight elements per thread and iteration (n % 8 == 0, 16-byte aligned pointers): 16- and 32-byte accesses instead of 2- and 4-byte ones
template <typename TS, typename TD>
__global__ void cast8_kernel(const TS* __restrict__ src, TD* __restrict__ dst, int64_t n8) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    float v[8];
    ld8(src + i * 8, v);
    st8(dst + i * 8, v);
  }
}

// 32 x 32 tiles through shared memory; grid (ceil(cols/32), ceil(rows/32), Z), block (32, 8)
template <typename T>
__global__ void transpose_kernel(const T* __restrict__ src, T* __restrict__ dst, int rows, int cols) {
  __shared__ T tile[32][33];
  pdl_trigger();
  pdl_wait();
  const long long zoff = (long long)blockIdx.z * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8)
    if (r0 + i < rows && c0 + threadIdx.x < cols) tile[i][threadIdx.x] = src[zoff + (long long)(r0 + i) * cols + c0 + threadIdx.x];
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8)
    if (c0 + i < cols && r0 + threadIdx.x < rows) dst[zoff + (long long)(c0 + i) * rows + r0 + threadIdx.x] = tile[threadIdx.x][i];
}

inline unsigned ew_grid(int64_t n) {
  int64_t g = (n + 255) / 256;
  const int64_t cap = 8LL * num_sms_cached();
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

namespace {
// eight elements per thread and iteration (n % 8 == 0, 16-byte aligned pointers)
template <typename T>
__global__ void relu_gate8_kernel(T* __restrict__ dy, const T* __restrict__ y, int64_t n8, float scale) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    float g[8], v[8];
    ld8(dy + i * 8, g);
    ld8(y + i * 8, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = v[e] > 0.f ? g[e] * scale : 0.f;
    st8(dy + i * 8, g);
  }
}
// one thread sums eight columns of a (q, :) row group over its G rows (d % 8 == 0, 16-byte aligned input)
template <typename T>
__global__ void group_sum8_kernel(const T* __restrict__ in, float* __restrict__ out, int64_t Qrows, int G, int d) {
  pdl_trigger();
  pdl_wait();
  const int d8 = d >> 3;
  const int64_t total = Qrows * d8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t q = i / d8;
    const int c = (int)(i - q * d8) * 8;
    const T* p = in + q * G * d + c;
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int g = 0; g < G; ++g) {
      float v[8];
      ld8(p + (int64_t)g * d, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) s[e] += v[e];
    }
    float* o = out + q * d + c;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] += s[e];
  }
}
}  // namespace

int relu_gate(void* dy, const void* y, int64_t n, int dtype, cudaStream_t st, float scale) {
  ProfScope ps("relu_gate", st);
  if (n % 8 == 0 && (((uintptr_t)dy | (uintptr_t)y) & 15) == 0) {
    if (dtype == BIST_BF16) BIST_LAUNCH(relu_gate8_kernel<__nv_bfloat16>, ew_grid(n / 8), 256, 0, st, (__nv_bfloat16*)dy, (const __nv_bfloat16*)y, n / 8, scale);
    else BIST_LAUNCH(relu_gate8_kernel<float>, ew_grid(n / 8), 256, 0, st, (float*)dy, (const float*)y, n / 8, scale);
    BIST_LAUNCHED();
    return 0;
  }
  if (dtype == BIST_BF16) BIST_LAUNCH(relu_gate_kernel<__nv_bfloat16>, ew_grid(n), 256, 0, st, (__nv_bfloat16*)dy, (const __nv_bfloat16*)y, n, scale);
  else BIST_LAUNCH(relu_gate_kernel<float>, ew_grid(n), 256, 0, st, (float*)dy, (const float*)y, n, scale);
  BIST_LAUNCHED();
  return 0;
}

int dropout_add(void* out, const float* res, int64_t rows, int64_t d, int64_t res_div, const DropCfg& drop, int perm_L, int perm_G,
                int out_dtype, cudaStream_t st, int64_t row0) {
  BIST_CHECK_ARG(out && rows > 0 && d > 0 && d % 4 == 0 && res_div > 0, "dropout_add: bad argument (d %% 4 == 0)");
  BIST_CHECK_ARG(((uintptr_t)out % 16 == 0 || (out_dtype == BIST_BF16 && (uintptr_t)out % 8 == 0)) && (!res || (uintptr_t)res % 16 == 0),
                 "dropout_add: misaligned pointer");
  ProfScope ps("dropout", st);
  {
    static const bool rows_off = getenv("BIST_DROPOUT_ELEMENTWISE") != nullptr;   // debugging aid: the four-elements-per-thread kernels
    bool taken = false;
    if (!rows_off) {
      if (out_dtype == BIST_BF16) BIST_TRY((launch_dropout_rows<__nv_bfloat16, __nv_bfloat16>(out, out, res, rows, d, res_div, drop, perm_L, perm_G, row0, st, &taken)));
      else BIST_TRY((launch_dropout_rows<float, float>(out, out, res, rows, d, res_div, drop, perm_L, perm_G, row0, st, &taken)));
    }
    if (taken) { BIST_LAUNCHED(); return 0; }
  }
  const bool small = rows * d < (1LL << 31) - (1LL << 24);
  const unsigned grid = ew_grid(rows * d / 4);
  if (out_dtype == BIST_BF16) {
    if (small) BIST_LAUNCH((dropout_add_kernel<__nv_bfloat16, uint32_t>), grid, 256, 0, st, (__nv_bfloat16*)out, res, rows, (int)d, res_div, drop, perm_L, perm_G, row0);
    else BIST_LAUNCH((dropout_add_kernel<__nv_bfloat16, int64_t>), grid, 256, 0, st, (__nv_bfloat16*)out, res, rows, (int)d, res_div, drop, perm_L, perm_G, row0);
  } else {
    if (small) BIST_LAUNCH((dropout_add_kernel<float, uint32_t>), grid, 256, 0, st, (float*)out, res, rows, (int)d, res_div, drop, perm_L, perm_G, row0);
    else BIST_LAUNCH((dropout_add_kernel<float, int64_t>), grid, 256, 0, st, (float*)out, res, rows, (int)d, res_div, drop, perm_L, perm_G, row0);
  }
  BIST_LAUNCHED();
  return 0;
}

int dropout_scale(const void* src, void* dst, int64_t rows, int64_t d, const DropCfg& drop, int perm_L, int perm_G, int src_dtype,
                  int dst_dtype, cudaStream_t st, int64_t row0) {
  BIST_CHECK_ARG(src && dst && rows > 0 && d > 0 && d % 4 == 0, "dropout_scale: bad argument (d %% 4 == 0)");
  BIST_CHECK_ARG((uintptr_t)src % 8 == 0 && (uintptr_t)dst % 8 == 0 && (src_dtype == BIST_BF16 || (uintptr_t)src % 16 == 0) &&
                     (dst_dtype == BIST_BF16 || (uintptr_t)dst % 16 == 0), "dropout_scale: misaligned pointer");
  ProfScope ps("dropout", st);
  {
    static const bool rows_off = getenv("BIST_DROPOUT_ELEMENTWISE") != nullptr;
    bool taken = false;
    if (!rows_off) {
      if (src_dtype == BIST_F32 && dst_dtype == BIST_BF16) BIST_TRY((launch_dropout_rows<float, __nv_bfloat16>(src, dst, nullptr, rows, d, 1, drop, perm_L, perm_G, row0, st, &taken)));
      else if (src_dtype == BIST_F32) BIST_TRY((launch_dropout_rows<float, float>(src, dst, nullptr, rows, d, 1, drop, perm_L, perm_G, row0, st, &taken)));
      else if (dst_dtype == BIST_BF16) BIST_TRY((launch_dropout_rows<__nv_bfloat16, __nv_bfloat16>(src, dst, nullptr, rows, d, 1, drop, perm_L, perm_G, row0, st, &taken)));
      else BIST_TRY((launch_dropout_rows<__nv_bfloat16, float>(src, dst, nullptr, rows, d, 1, drop, perm_L, perm_G, row0, st, &taken)));
    }
    if (taken) { BIST_LAUNCHED(); return 0; }
  }
  const unsigned g = ew_grid(rows * d / 4);
  const int di = (int)d;
  const bool small = rows * d < (1LL << 31) - (1LL << 24);
#define BIST_DS(TS_, TD_)                                                                                                               \
  do {                                                                                                                                  \
    if (small) BIST_LAUNCH((dropout_scale_kernel<TS_, TD_, uint32_t>), g, 256, 0, st, (const TS_*)src, (TD_*)dst, rows, di, drop, perm_L, perm_G, row0); \
    else BIST_LAUNCH((dropout_scale_kernel<TS_, TD_, int64_t>), g, 256, 0, st, (const TS_*)src, (TD_*)dst, rows, di, drop, perm_L, perm_G, row0);        \
  } while (0)
  if (src_dtype == BIST_F32 && dst_dtype == BIST_BF16) BIST_DS(float, __nv_bfloat16);
  else if (src_dtype == BIST_F32) BIST_DS(float, float);
  else if (dst_dtype == BIST_BF16) BIST_DS(__nv_bfloat16, __nv_bfloat16);
  else BIST_DS(__nv_bfloat16, float);
#undef BIST_DS
  BIST_LAUNCHED();
  return 0;
}

int bias_rsum_fwd(void* c2, const float* bv, const float* rsum, int64_t Q, int h, int dk, int dtype, cudaStream_t st) {
  ProfScope ps("dropout", st);
  if (dtype == BIST_BF16) BIST_LAUNCH(bias_rsum_fwd_kernel<__nv_bfloat16>, ew_grid(Q * h * dk), 256, 0, st, (__nv_bfloat16*)c2, bv, rsum, Q, h, dk);
  else BIST_LAUNCH(bias_rsum_fwd_kernel<float>, ew_grid(Q * h * dk), 256, 0, st, (float*)c2, bv, rsum, Q, h, dk);
  BIST_LAUNCHED();
  return 0;
}

int bias_rsum_bwd(const void* dc2, const float* bv, const float* rsum, float* drsum, float* dbv, int64_t Q, int h, int dk, int dtype,
                  cudaStream_t st) {
  ProfScope ps("dropout", st);
  const size_t smem = (size_t)h * dk * 4;
  const int lph = dk / 8;
  if (dk % 8 == 0 && h * dk <= 1024 && lph >= 1 && lph <= 32 && (lph & (lph - 1)) == 0 && ((uintptr_t)dc2 % 16) == 0) {
    const int rpc = 16;   // two rows per warp: 128 CTAs at Q = 2048
    const unsigned grid = (unsigned)((Q + rpc - 1) / rpc);
    if (dtype == BIST_BF16) BIST_LAUNCH(bias_rsum_bwd_vec_kernel<__nv_bfloat16>, grid, 256, smem, st, (const __nv_bfloat16*)dc2, bv, rsum, drsum, dbv, Q, h, dk, rpc);
    else BIST_LAUNCH(bias_rsum_bwd_vec_kernel<float>, grid, 256, smem, st, (const float*)dc2, bv, rsum, drsum, dbv, Q, h, dk, rpc);
    BIST_LAUNCHED();
    return 0;
  }
  const unsigned grid = (unsigned)((Q + 31) / 32);
  if (dtype == BIST_BF16) BIST_LAUNCH(bias_rsum_bwd_kernel<__nv_bfloat16>, grid, 256, smem, st, (const __nv_bfloat16*)dc2, bv, rsum, drsum, dbv, Q, h, dk);
  else BIST_LAUNCH(bias_rsum_bwd_kernel<float>, grid, 256, smem, st, (const float*)dc2, bv, rsum, drsum, dbv, Q, h, dk);
  BIST_LAUNCHED();
  return 0;
}

int dropout_mask(float* out, int64_t n, const DropCfg& drop, cudaStream_t st) {
  BIST_CHECK_ARG(out && n > 0, "dropout_mask: bad argument");
  BIST_LAUNCH(dropout_mask_kernel, ew_grid(n), 256, 0, st, out, n, drop);
  BIST_LAUNCHED();
  return 0;
}

int group_sum(const void* in, float* out, int64_t Qrows, int64_t G, int64_t d, int in_dtype, cudaStream_t st) {
  ProfScope ps("group_sum", st);
  if (d % 8 == 0 && ((uintptr_t)in & 15) == 0) {
    if (in_dtype == BIST_BF16) BIST_LAUNCH(group_sum8_kernel<__nv_bfloat16>, ew_grid(Qrows * d / 8), 256, 0, st, (const __nv_bfloat16*)in, out, Qrows, (int)G, (int)d);
    else BIST_LAUNCH(group_sum8_kernel<float>, ew_grid(Qrows * d / 8), 256, 0, st, (const float*)in, out, Qrows, (int)G, (int)d);
    BIST_LAUNCHED();
    return 0;
  }
  if (in_dtype == BIST_BF16) BIST_LAUNCH(group_sum_kernel<__nv_bfloat16>, ew_grid(Qrows * d), 256, 0, st, (const __nv_bfloat16*)in, out, Qrows, (int)G, (int)d);
  else BIST_LAUNCH(group_sum_kernel<float>, ew_grid(Qrows * d), 256, 0, st, (const float*)in, out, Qrows, (int)G, (int)d);
  BIST_LAUNCHED();
  return 0;
}

int transpose_batched(const void* src, void* dst, int64_t Z, int64_t rows, int64_t cols, int dtype, cudaStream_t st) {
  BIST_CHECK_ARG(src && dst && Z > 0 && Z <= 65535 && rows > 0 && cols > 0, "transpose: bad argument");
  dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32), (unsigned)Z), block(32, 8);
  ProfScope ps("transpose", st);
  if (dtype == BIST_BF16) BIST_LAUNCH(transpose_kernel<__nv_bfloat16>, grid, block, 0, st, (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, (int)rows, (int)cols);
  else BIST_LAUNCH(transpose_kernel<float>, grid, block, 0, st, (const float*)src, (float*)dst, (int)rows, (int)cols);
  BIST_LAUNCHED();
  return 0;
}

int cast_copy(const void* src, void* dst, int64_t n, int src_dtype, int dst_dtype, cudaStream_t st) {
  ProfScope ps("cast", st);
  if (n % 8 == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0 && src_dtype != dst_dtype) {
    const unsigned g8 = ew_grid(n / 8);
    if (src_dtype == BIST_F32) BIST_LAUNCH((cast8_kernel<float, __nv_bfloat16>), g8, 256, 0, st, (const float*)src, (__nv_bfloat16*)dst, n / 8);
    else BIST_LAUNCH((cast8_kernel<__nv_bfloat16, float>), g8, 256, 0, st, (const __nv_bfloat16*)src, (float*)dst, n / 8);
    BIST_LAUNCHED();
    return 0;
  }
  const unsigned g = ew_grid(n);
  if (src_dtype == BIST_F32 && dst_dtype == BIST_BF16) BIST_LAUNCH((cast_kernel<float, __nv_bfloat16>), g, 256, 0, st, (const float*)src, (__nv_bfloat16*)dst, n);
  else if (src_dtype == BIST_BF16 && dst_dtype == BIST_F32) BIST_LAUNCH((cast_kernel<__nv_bfloat16, float>), g, 256, 0, st, (const __nv_bfloat16*)src, (float*)dst, n);
  else if (src_dtype == BIST_F32) BIST_LAUNCH((cast_kernel<float, float>), g, 256, 0, st, (const float*)src, (float*)dst, n);
  else BIST_LAUNCH((cast_kernel<__nv_bfloat16, __nv_bfloat16>), g, 256, 0, st, (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, n);
  BIST_LAUNCHED();
  return 0;
}

// ------------------------------------------------------------------------------------------------ attention() backward
namespace {

// grid (Bp * h); block 256.  smem (floats): Q[Lq][dkp] K[Lk][dkp] V[Lk][dkp] dC[Lq][dkp] P[Lq][Lkp] dS[Lq][Lkp]
template <typename T>
__global__ void __launch_bounds__(256) attn_bwd_kernel(AttnBwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x / a.h, hh = blockIdx.x % a.h;
  const int Lq = a.Lq, Lk = a.Lk, dk = a.dk, dkp = dk + 1, Lkp = Lk + 1;
  float* Qs = sm;
  float* Ks = Qs + Lq * dkp;
  float* Vs = Ks + Lk * dkp;
  float* Cs = Vs + Lk * dkp;
  float* Ps = Cs + Lq * dkp;
  float* Ds = Ps + Lq * Lkp;
  const T* q = reinterpret_cast<const T*>(a.q) + b * a.q_b + hh * dk;
  const T* k = reinterpret_cast<const T*>(a.k) + b * a.k_b + hh * dk;
  const T* v = reinterpret_cast<const T*>(a.v) + b * a.k_b + hh * dk;
  const T* dc = reinterpret_cast<const T*>(a.dctx) + b * a.c_b + hh * dk;
  const int t = threadIdx.x, nt = blockDim.x;
  for (int i = t; i < Lq * dk; i += nt) {
    const int l = i / dk, j = i % dk;
    Qs[l * dkp + j] = to_f32(q[l * a.q_l + j]);
    Cs[l * dkp + j] = to_f32(dc[l * a.c_l + j]);
  }
  for (int i = t; i < Lk * dk; i += nt) {
    const int kk = i / dk, j = i % dk;
    Ks[kk * dkp + j] = to_f32(k[kk * a.k_l + j]);
    Vs[kk * dkp + j] = to_f32(v[kk * a.k_l + j]);
  }
  __syncthreads();
  const float scale = rsqrtf((float)dk);
  const uint8_t* mb = a.mask ? a.mask + (long long)(b % a.mask_mod) * a.m_b : nullptr;
  for (int i = t; i < Lq * Lk; i += nt) {
    const int l = i / Lk, kk = i % Lk;
    float s = 0.f, dp = 0.f;
    for (int j = 0; j < dk; ++j) {
      s = fmaf(Qs[l * dkp + j], Ks[kk * dkp + j], s);
      dp = fmaf(Cs[l * dkp + j], Vs[kk * dkp + j], dp);
    }
    s *= scale;
    if (mb && mb[l * a.m_l + kk] == 0) s = kMaskFill;
    Ps[l * Lkp + kk] = s;
    Ds[l * Lkp + kk] = dp;
  }
  __syncthreads();
  const int lane = t & 31, warp = t >> 5, nw = nt >> 5;
  for (int l = warp; l < Lq; l += nw) {
    float mx = -INFINITY;
    for (int kk = lane; kk < Lk; kk += 32) mx = fmaxf(mx, Ps[l * Lkp + kk]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int kk = lane; kk < Lk; kk += 32) {
      const float e = expf(Ps[l * Lkp + kk] - mx);
      Ps[l * Lkp + kk] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float dot = 0.f;
    const unsigned long long ibase = (((unsigned long long)b * a.h + hh) * Lq + l) * Lk;
    for (int kk = lane; kk < Lk; kk += 32) {
      const float p = Ps[l * Lkp + kk] * inv;
      const float dpv = Ds[l * Lkp + kk] * drop_mul(a.drop, ibase + kk);  // dP = m o dP'
      Ps[l * Lkp + kk] = p;
      Ds[l * Lkp + kk] = dpv;
      dot = fmaf(p, dpv, dot);
    }
    dot = warp_sum(dot);
    for (int kk = lane; kk < Lk; kk += 32) {
      const float p = Ps[l * Lkp + kk];
      Ds[l * Lkp + kk] = p * (Ds[l * Lkp + kk] - dot) * scale;
      Ps[l * Lkp + kk] = p * drop_mul(a.drop, ibase + kk);                 // p' for dV below
    }
  }
  __syncthreads();
  T* dq = reinterpret_cast<T*>(a.dq) + b * a.q_b + hh * dk;
  T* dkk = reinterpret_cast<T*>(a.dk_) + b * a.k_b + hh * dk;
  T* dv = reinterpret_cast<T*>(a.dv) + b * a.k_b + hh * dk;
  for (int i = t; i < Lq * dk; i += nt) {
    const int l = i / dk, j = i % dk;
    float s = 0.f;
    for (int kk = 0; kk < Lk; ++kk) s = fmaf(Ds[l * Lkp + kk], Ks[kk * dkp + j], s);
    dq[l * a.q_l + j] = from_f32<T>(s);
  }
  for (int i = t; i < Lk * dk; i += nt) {
    const int kk = i / dk, j = i % dk;
    float sk = 0.f, sv = 0.f;
    for (int l = 0; l < Lq; ++l) {
      sk = fmaf(Ds[l * Lkp + kk], Qs[l * dkp + j], sk);
      sv = fmaf(Ps[l * Lkp + kk], Cs[l * dkp + j], sv);
    }
    dkk[kk * a.k_l + j] = from_f32<T>(sk);
    dv[kk * a.k_l + j] = from_f32<T>(sv);
  }
}

}  // namespace

int attn_core_bwd(const AttnBwdArgs& a, cudaStream_t st) {
  BIST_CHECK_ARG(a.q && a.k && a.v && a.dctx && a.dq && a.dk_ && a.dv, "attn_bwd: null pointer");
  BIST_CHECK_ARG(a.Bp > 0 && a.Lq > 0 && a.Lk > 0 && a.h > 0 && a.dk > 0, "attn_bwd: bad shape");
  const size_t smem = ((size_t)(2 * a.Lq + 2 * a.Lk) * (a.dk + 1) + 2 * (size_t)a.Lq * (a.Lk + 1)) * 4;
  if (smem > 200 * 1024) {
    set_error("attn_bwd: Lq=%d Lk=%d dk=%d does not fit one CTA", a.Lq, a.Lk, a.dk);
    return BIST_ERR_UNSUPPORTED;
  }
  ProfScope ps("attn_bwd", st);
  if (a.dtype == BIST_BF16) {
    auto kern = attn_bwd_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)(a.Bp * a.h), 256, smem, st, a);
  } else {
    auto kern = attn_bwd_kernel<float>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)(a.Bp * a.h), 256, smem, st, a);
  }
  BIST_LAUNCHED();
  return 0;
}

// ------------------------------------------------------------------------------------------------ stage-1 backward
namespace {

// grid G groups; block 256.  smem floats: P[Lq*h][nkp] (probabilities, then ds), dC[Lq][dkp], Vh[nk][dkp]
template <typename TV, typename TS>
__global__ void __launch_bounds__(256) stage1_bwd_kernel(Stage1BwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  pdl_trigger();
  pdl_wait();
  const SoftmaxPvArgs& f = a.f;
  const int g = blockIdx.x;
  const int h = f.h, nk = f.nk, Lq = f.Lq, dk = f.dk, dkp = dk + 1, nkp = nk + 1;
  float* Ps = sm;                      // [Lq*h][nkp], row = l*h + hh
  float* Cs = Ps + Lq * h * nkp;       // [Lq][dkp]
  float* Vs = Cs + Lq * dkp;           // [nk][dkp]
  float* Pd = Vs + nk * dkp;           // [Lq][nkp] dropped probabilities of the current head
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31, warp = t >> 5, nw = nt >> 5;

  // phase 1: probabilities of every (l, hh) row of this group, recomputed from the saved scores
  const float* Sg = f.S + f.sg.off(g);
  const uint8_t* Mg = f.mask ? f.mask + (long long)((g / f.mask_gdiv) % f.mask_gmod) * f.m_b : nullptr;
  for (int r = warp; r < Lq * h; r += nw) {
    const int l = r / h, hh = r % h;
    const float* srow = Sg + (long long)l * f.s_l + (long long)hh * f.s_h;
    const uint8_t* mrow = Mg ? Mg + (long long)l * f.m_l : nullptr;
    float mx = -INFINITY;
    for (int k = lane; k < nk; k += 32) {
      float s = srow[(long long)k * f.s_k];
      if (mrow && mrow[k] == 0) s = kMaskFill;
      Ps[r * nkp + k] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < nk; k += 32) {
      const float e = expf(Ps[r * nkp + k] - mx);
      Ps[r * nkp + k] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int k = lane; k < nk; k += 32) Ps[r * nkp + k] *= inv;
  }
  __syncthreads();

  const TV* Vg = reinterpret_cast<const TV*>(f.V) + f.vg.off(g);
  TV* dVg = reinterpret_cast<TV*>(a.dV) + f.vg.off(g);
  const TV* dCg = reinterpret_cast<const TV*>(a.dctx) + f.og.off(g);
  for (int hh = 0; hh < h; ++hh) {
    for (int i = t; i < Lq * dk; i += nt) {
      const int l = i / dk, j = i % dk;
      Cs[l * dkp + j] = to_f32(dCg[(long long)l * f.o_l + hh * dk + j]);
    }
    for (int i = t; i < nk * dk; i += nt) {
      const int k = i / dk, j = i % dk;
      Vs[k * dkp + j] = to_f32(Vg[f.v.off(k) + hh * dk + j]);
    }
    for (int i = t; i < Lq * nk; i += nt) {
      const int l = i / nk, k = i % nk;
      Pd[l * nkp + k] = Ps[(l * h + hh) * nkp + k] * drop_mul(f.drop, (((unsigned long long)g * h + hh) * Lq + l) * nk + k);
    }
    __syncthreads();
    // dV[k, j] = sum_l p'[l,hh,k] dC[l, j]
    for (int i = t; i < nk * dk; i += nt) {
      const int k = i / dk, j = i % dk;
      float s = 0.f;
      for (int l = 0; l < Lq; ++l) s = fmaf(Pd[l * nkp + k], Cs[l * dkp + j], s);
      dVg[f.v.off(k) + hh * dk + j] = from_f32<TV>(s);
    }
    // dP[l,k] = dC[l,:] . V[k,:] ;  ds = p * (dP - sum_k p dP): one warp per row l
    for (int l = warp; l < Lq; l += nw) {
      float* prow = Ps + (l * h + hh) * nkp;
      float dpv[8];  // nk <= 256
      float dot = 0.f;
      int c = 0;
      for (int k = lane; k < nk; k += 32, ++c) {
        float s = 0.f;
        for (int j = 0; j < dk; ++j) s = fmaf(Cs[l * dkp + j], Vs[k * dkp + j], s);
        s *= drop_mul(f.drop, (((unsigned long long)g * h + hh) * Lq + l) * nk + k);   // dP = m o dP'
        dpv[c] = s;
        dot = fmaf(prow[k], s, dot);
      }
      dot = warp_sum(dot);
      c = 0;
      for (int k = lane; k < nk; k += 32, ++c) prow[k] = prow[k] * (dpv[c] - dot);
    }
    __syncthreads();
  }

  // phase 3: write ds in both layouts
  const int b = g / a.gdiv, gi = g % a.gdiv;
  const int R = Lq * h;
  TS* ds = reinterpret_cast<TS*>(a.ds) + (long long)b * R * a.TS;
  const long long ldt = a.dsT_ld > 0 ? a.dsT_ld : R;
  TS* dsT = reinterpret_cast<TS*>(a.dsT) + (long long)b * a.TS * ldt;
  if (a.ds)
    for (int i = t; i < R * nk; i += nt) {  // k fastest: ds[b][r][ts]
      const int r = i / nk, k = i % nk;
      ds[(long long)r * a.TS + gi * a.ts_g + k * a.ts_k] = from_f32<TS>(Ps[r * nkp + k]);
    }
  for (int i = t; i < R * nk; i += nt) {  // r fastest: dsT[b][ts][r]
    const int k = i / R, r = i % R;
    dsT[((long long)gi * a.ts_g + k * a.ts_k) * ldt + r] = from_f32<TS>(Ps[r * nkp + k]);
  }
}

}  // namespace

int stage1_bwd(const Stage1BwdArgs& a, cudaStream_t st) {
  const SoftmaxPvArgs& f = a.f;
  BIST_CHECK_ARG(f.S && f.V && a.dctx && a.dV && a.dsT, "stage1_bwd: null pointer");
  {
    bool taken = false;
    int rc = stage1_bwd_mma(a, st, &taken);
    if (taken) return rc;
  }
  BIST_CHECK_ARG(f.G > 0 && f.Lq > 0 && f.h > 0 && f.dk > 0 && f.nk > 0 && f.nk <= 256, "stage1_bwd: bad shape (nk <= 256)");
  BIST_CHECK_ARG(a.ds_dtype == f.v_dtype, "stage1_bwd: ds dtype must equal the value dtype");
  const size_t smem = ((size_t)f.Lq * f.h * (f.nk + 1) + (size_t)(f.Lq + f.nk) * (f.dk + 1) + (size_t)f.Lq * (f.nk + 1)) * 4;
  if (smem > 200 * 1024) {
    set_error("stage1_bwd: Lq*h*nk = %d*%d*%d too large for one CTA", f.Lq, f.h, f.nk);
    return BIST_ERR_UNSUPPORTED;
  }
  ProfScope ps("stage1_bwd", st);
  if (f.v_dtype == BIST_BF16) {
    auto kern = stage1_bwd_kernel<__nv_bfloat16, __nv_bfloat16>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)f.G, 256, smem, st, a);
  } else {
    auto kern = stage1_bwd_kernel<float, float>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)f.G, 256, smem, st, a);
  }
  BIST_LAUNCHED();
  return 0;
}

// ------------------------------------------------------------------------------------------------ stage-2 backward
namespace {

// grid B*Lq; block 256.  smem floats: QF[h][d], DU[h][d], P[h][Gp], DS[h][Gp]
template <typename T>
__global__ void __launch_bounds__(256) stage2_bwd_kernel(Stage2BwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  pdl_trigger();
  pdl_wait();
  const int q = blockIdx.x, b = q / a.Lq;
  const int h = a.h, d = a.d, G = a.G, Gp = G + 1;
  float* QF = sm;
  float* DU = QF + h * d;
  float* Ps = DU + h * d;
  float* Ds = Ps + h * Gp;
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31, warp = t >> 5, nw = nt >> 5;
  const T* qf = reinterpret_cast<const T*>(a.qf) + (long long)q * h * d;
  const T* du = reinterpret_cast<const T*>(a.du) + (long long)q * h * d;
  const T* x1 = reinterpret_cast<const T*>(a.x1) + (long long)q * G * d;
  for (int i = t; i < h * d; i += nt) {
    QF[i] = to_f32(qf[i]);
    DU[i] = to_f32(du[i]);
  }
  __syncthreads();
  // scores and dP: one warp per key row g
  for (int g = warp; g < G; g += nw) {
    const T* xr = x1 + (long long)g * d;
    for (int hh = 0; hh < h; ++hh) {
      float s = 0.f, dp = 0.f;
      for (int c = lane; c < d; c += 32) {
        const float xv = to_f32(xr[c]);
        s = fmaf(QF[hh * d + c], xv, s);
        dp = fmaf(DU[hh * d + c], xv, dp);
      }
      s = warp_sum(s);
      dp = warp_sum(dp);
      if (lane == 0) {
        if (a.mask && a.mask[(long long)b * G + g] == 0) s = kMaskFill;
        if (a.drsum) dp += a.drsum[(long long)q * h + hh];
        Ps[hh * Gp + g] = s;
        Ds[hh * Gp + g] = dp * drop_mul(a.drop, ((unsigned long long)q * h + hh) * G + g);   // dP = m o dP'
      }
    }
  }
  __syncthreads();
  for (int hh = warp; hh < h; hh += nw) {
    float mx = -INFINITY;
    for (int g = lane; g < G; g += 32) mx = fmaxf(mx, Ps[hh * Gp + g]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int g = lane; g < G; g += 32) {
      const float e = expf(Ps[hh * Gp + g] - mx);
      Ps[hh * Gp + g] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float dot = 0.f;
    for (int g = lane; g < G; g += 32) {
      const float p = Ps[hh * Gp + g] * inv;
      Ps[hh * Gp + g] = p;
      dot = fmaf(p, Ds[hh * Gp + g], dot);
    }
    dot = warp_sum(dot);
    for (int g = lane; g < G; g += 32) {
      const float p = Ps[hh * Gp + g];
      Ds[hh * Gp + g] = p * (Ds[hh * Gp + g] - dot);
      Ps[hh * Gp + g] = p * drop_mul(a.drop, ((unsigned long long)q * h + hh) * G + g);       // p' for dx1 below
    }
  }
  __syncthreads();
  // dx1[g, c] = sum_hh p'[hh,g] du[hh,c] + ds[hh,g] qf[hh,c]
  T* dx1 = reinterpret_cast<T*>(a.dx1) + (long long)q * G * d;
  for (int i = t; i < G * d; i += nt) {
    const int g = i / d, c = i % d;
    float s = 0.f;
    for (int hh = 0; hh < h; ++hh) s = fmaf(Ps[hh * Gp + g], DU[hh * d + c], fmaf(Ds[hh * Gp + g], QF[hh * d + c], s));
    dx1[i] = from_f32<T>(s);
  }
  // dqf[hh, c] = sum_g ds[hh,g] x1[g,c]
  T* dqf = reinterpret_cast<T*>(a.dqf) + (long long)q * h * d;
  for (int i = t; i < h * d; i += nt) {
    const int hh = i / d, c = i % d;
    float s = 0.f;
    for (int g = 0; g < G; ++g) s = fmaf(Ds[hh * Gp + g], to_f32(x1[(long long)g * d + c]), s);
    dqf[i] = from_f32<T>(s);
  }
}

}  // namespace

int stage2_bwd(const Stage2BwdArgs& a, cudaStream_t st) {
  BIST_CHECK_ARG(a.qf && a.x1 && a.du && a.dx1 && a.dqf, "stage2_bwd: null pointer");
  BIST_CHECK_ARG(a.B > 0 && a.G > 0 && a.Lq > 0 && a.h > 0 && a.d > 0, "stage2_bwd: bad shape");
  {
    bool taken = false;
    int rc = stage2_bwd_mma(a, st, &taken);
    if (taken) return rc;
  }
  const size_t smem = ((size_t)2 * a.h * a.d + 2 * (size_t)a.h * (a.G + 1)) * 4;
  if (smem > 200 * 1024) {
    set_error("stage2_bwd: h*d = %d*%d too large for one CTA", a.h, a.d);
    return BIST_ERR_UNSUPPORTED;
  }
  ProfScope ps("stage2_bwd", st);
  if (a.dtype == BIST_BF16) {
    auto kern = stage2_bwd_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)(a.B * a.Lq), 256, smem, st, a);
  } else {
    auto kern = stage2_bwd_kernel<float>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)(a.B * a.Lq), 256, smem, st, a);
  }
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist
