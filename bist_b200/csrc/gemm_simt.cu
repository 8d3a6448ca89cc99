// Generic CUDA-core GEMM  C = act(alpha * A . W^T + bias) + R  with two-level row maps and
// two-level batching (common.cuh GemmArgs).  fp32 accumulate; A/W/C/R may be fp32 or bf16.
// This is the exact-fp32 path and the path for shapes the tcgen05 kernel does not take.
#include "common.cuh"

namespace bist {
namespace {

constexpr int BK = 16;

template <int BM, int BN, int TM, int TN, typename TA, typename TW, bool VEC>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_simt_kernel(GemmArgs g) {
  constexpr int NT = (BM / TM) * (BN / TN);
  constexpr int PAD = 4;
  __shared__ __align__(16) float As[2][BK][BM + PAD];
  __shared__ __align__(16) float Ws[2][BK][BN + PAD];
  constexpr int A_LD = BM * BK / 4 / NT;  // float4 loads per thread
  constexpr int W_LD = BN * BK / 4 / NT;
  static_assert(A_LD >= 1 && W_LD >= 1, "tile too small for the thread count");

  pdl_trigger();
  pdl_wait();
  const int tid = threadIdx.x;
  const int z = blockIdx.z;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const TA* A = reinterpret_cast<const TA*>(g.A) + g.az.off(z, g.zin, g.zmid);
  const TW* W = reinterpret_cast<const TW*>(g.W) + g.wz.off(z, g.zin, g.zmid);

  // rows this thread loads
  long long a_off[A_LD], w_off[W_LD];
  bool a_ok[A_LD], w_ok[W_LD];
  int a_row[A_LD], w_row[W_LD], a_kq[A_LD], w_kq[W_LD];
#pragma unroll
  for (int i = 0; i < A_LD; ++i) {
    int idx = tid + i * NT;
    a_row[i] = idx >> 2;
    a_kq[i] = (idx & 3) * 4;
    int r = m0 + a_row[i];
    a_ok[i] = r < g.M;
    a_off[i] = a_ok[i] ? g.a.off(r) : 0;
  }
#pragma unroll
  for (int i = 0; i < W_LD; ++i) {
    int idx = tid + i * NT;
    w_row[i] = idx >> 2;
    w_kq[i] = (idx & 3) * 4;
    int r = n0 + w_row[i];
    w_ok[i] = r < g.N;
    w_off[i] = w_ok[i] ? g.w.off(r) : 0;
  }

  auto load4 = [&](auto* base, long long off, bool ok, int k) -> float4 {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!ok) return v;
    if (VEC) {
      if (k < g.K) v = ld4(base + off + k);  // K % 4 == 0 guaranteed by the host for VEC
    } else {
      if (k + 0 < g.K) v.x = to_f32(base[off + k + 0]);
      if (k + 1 < g.K) v.y = to_f32(base[off + k + 1]);
      if (k + 2 < g.K) v.z = to_f32(base[off + k + 2]);
      if (k + 3 < g.K) v.w = to_f32(base[off + k + 3]);
    }
    return v;
  };

  float4 ra[A_LD], rw[W_LD];
  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < A_LD; ++i) ra[i] = load4(A, a_off[i], a_ok[i], k0 + a_kq[i]);
#pragma unroll
    for (int i = 0; i < W_LD; ++i) rw[i] = load4(W, w_off[i], w_ok[i], k0 + w_kq[i]);
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_LD; ++i) {
      As[buf][a_kq[i] + 0][a_row[i]] = ra[i].x;
      As[buf][a_kq[i] + 1][a_row[i]] = ra[i].y;
      As[buf][a_kq[i] + 2][a_row[i]] = ra[i].z;
      As[buf][a_kq[i] + 3][a_row[i]] = ra[i].w;
    }
#pragma unroll
    for (int i = 0; i < W_LD; ++i) {
      Ws[buf][w_kq[i] + 0][w_row[i]] = rw[i].x;
      Ws[buf][w_kq[i] + 1][w_row[i]] = rw[i].y;
      Ws[buf][w_kq[i] + 2][w_row[i]] = rw[i].z;
      Ws[buf][w_kq[i] + 3][w_row[i]] = rw[i].w;
    }
  };

  // thread tile: TM rows in chunks of 4 (chunk c at ty*4 + c*(BM/ (TM/4))), same for columns
  constexpr int MC = TM / 4, NC = TN / 4;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int nk = (g.K + BK - 1) / BK;
  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], w[TN];
#pragma unroll
      for (int c = 0; c < MC; ++c) {
        float4 v = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4 + c * (BM / MC)]);
        a[c * 4 + 0] = v.x; a[c * 4 + 1] = v.y; a[c * 4 + 2] = v.z; a[c * 4 + 3] = v.w;
      }
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        float4 v = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4 + c * (BN / NC)]);
        w[c * 4 + 0] = v.x; w[c * 4 + 1] = v.y; w[c * 4 + 2] = v.z; w[c * 4 + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }

  // epilogue
  const float* bias = g.bias ? g.bias + g.bz.off(z, g.zin, g.zmid) : nullptr;
  const long long c_base = g.cz.off(z, g.zin, g.zmid), r_base = g.rz.off(z, g.zin, g.zmid);
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int row = m0 + ty * 4 + (i / 4) * (BM / MC) + (i % 4);
    if (row >= g.M) continue;
    const long long co = c_base + g.c.off(row);
    const long long ro = g.R ? r_base + g.r.off(row) : 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int col = n0 + tx * 4 + c * (BN / NC);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int cc = col + j;
        if (cc >= g.N) continue;
        float v = g.alpha * acc[i][c * 4 + j];
        if (bias) v += bias[cc];
        if (g.relu) v = fmaxf(v, 0.f);
        if (g.R) {
          v += g.r_dtype == BIST_BF16
                   ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g.R)[ro + cc])
                   : reinterpret_cast<const float*>(g.R)[ro + cc];
        }
        if (g.c_dtype == BIST_BF16)
          reinterpret_cast<__nv_bfloat16*>(g.C)[co + cc] = __float2bfloat16_rn(v);
        else
          reinterpret_cast<float*>(g.C)[co + cc] = v;
      }
    }
  }
}

template <int BM, int BN, int TM, int TN, typename TA, typename TW>
int launch(const GemmArgs& g, bool vec, cudaStream_t st) {
  dim3 grid((g.N + BN - 1) / BN, (g.M + BM - 1) / BM, g.Z);
  constexpr int NT = (BM / TM) * (BN / TN);
  ProfScope ps(g.tag, st);
  if (vec)
    BIST_LAUNCH((gemm_simt_kernel<BM, BN, TM, TN, TA, TW, true>), grid, NT, 0, st, g);
  else
    BIST_LAUNCH((gemm_simt_kernel<BM, BN, TM, TN, TA, TW, false>), grid, NT, 0, st, g);
  BIST_LAUNCHED();
  return 0;
}

template <typename TA, typename TW>
int pick_tile(const GemmArgs& g, bool vec, cudaStream_t st) {
  // large problems: 128x128 tiles (8x8 per thread); otherwise 64x64 (4x4 per thread)
  long long tiles128 = (long long)((g.M + 127) / 128) * ((g.N + 127) / 128) * g.Z;
  if (g.M >= 128 && g.N >= 128 && tiles128 >= 148) return launch<128, 128, 8, 8, TA, TW>(g, vec, st);
  return launch<64, 64, 4, 4, TA, TW>(g, vec, st);
}

bool mult4(long long v) { return (v & 3) == 0; }

}  // namespace

int gemm_simt(const GemmArgs& g, cudaStream_t st) {
  BIST_CHECK_ARG(g.A && g.W && g.C, "gemm: null operand");
  BIST_CHECK_ARG(!g.C2, "gemm: the bf16 output copy is only written by the tensor-core kernel (shape declined)");
  BIST_CHECK_ARG(g.M > 0 && g.N > 0 && g.K > 0 && g.Z > 0 && g.zin > 0, "gemm: bad shape M=%d N=%d K=%d Z=%d",
                 g.M, g.N, g.K, g.Z);
  BIST_CHECK_ARG(g.Z <= 65535 && (g.M + 63) / 64 <= 65535, "gemm: grid too large");
  bool vec = mult4(g.K) && mult4(g.a.so) && mult4(g.a.si) && mult4(g.w.so) && mult4(g.w.si) &&
             mult4(g.az.s0) && mult4(g.az.s1) && mult4(g.az.s2) && mult4(g.wz.s0) && mult4(g.wz.s1) && mult4(g.wz.s2) &&
             ((uintptr_t)g.A % 16 == 0) && ((uintptr_t)g.W % 16 == 0);
  if (g.a_dtype == BIST_F32 && g.w_dtype == BIST_F32) return pick_tile<float, float>(g, vec, st);
  if (g.a_dtype == BIST_BF16 && g.w_dtype == BIST_BF16)
    return pick_tile<__nv_bfloat16, __nv_bfloat16>(g, vec, st);
  if (g.a_dtype == BIST_F32 && g.w_dtype == BIST_BF16) return pick_tile<float, __nv_bfloat16>(g, vec, st);
  if (g.a_dtype == BIST_BF16 && g.w_dtype == BIST_F32) return pick_tile<__nv_bfloat16, float>(g, vec, st);
  set_error("gemm: bad dtype");
  return BIST_ERR_ARG;
}

}  // namespace bist
