// Weight-gradient product on the 5th-generation tensor cores (tcgen05 / TMEM / TMA):
//   C[z][n,k] += alpha * sum_m A[z][m,n] * B[z][m,k]        A, B bf16 row-major, C fp32 (red.global.add)
// The contraction runs over the ROWS of both operands, so in UMMA terms both operands are MN-major: a TMA box of
// [64 rows (m) x 64 columns] with SWIZZLE_128B lands in shared memory exactly as the canonical MN-major SW128 atom
// (8 m-rows x 128 bytes of n, Swizzle<3,4,3>), no transposition anywhere.  One CTA = one (batch, 128-wide n tile,
// TN-wide k tile, m range); the m range is split over CTAs to fill the machine and the partial products meet in C
// through fp32 reductions.  192 threads, warp-specialised:
//   warp 0     TMA producer   2 + TN/64 boxes per 64-row slab into a ring of STAGES slots
//   warp 1     MMA issuer     tcgen05.mma kind::f16, M = 128, N = TN, K = 16 rows per instruction, fp32 accumulator in TMEM
//   warps 2-5  epilogue       tcgen05.ld -> per-warp smem transpose -> coalesced red.global.add.f32
#include <stdlib.h>

#include "tc_ptx.cuh"

namespace bist {
namespace {
using namespace tc;

constexpr int WT_TM = 128;   // n tile (UMMA M)
constexpr int WT_R = 64;     // m rows per ring slot
constexpr int WT_BOX = WT_R * 128;  // bytes of one [64 x 64] bf16 box
constexpr int WT_STG_LD = 33;

struct WtArgs {
  WgradArgs w;
  long long a_ld, b_ld;
  int tiles_n, tiles_k, splits, slabs_per_split, total_slabs;
  uint32_t lbo, sbo;  // descriptor fields (bytes >> 4)
};

template <int TN>
struct WtSmem {
  static constexpr int A_BYTES = 2 * WT_BOX, B_BYTES = (TN / 64) * WT_BOX, STAGE = A_BYTES + B_BYTES;
  static constexpr int STAGES = TN == 256 ? 4 : 6;
  static constexpr int STG_OFF = STAGES * STAGE;                 // 4 epilogue warps x [32][33] floats
  static constexpr int BAR_OFF = STG_OFF + 4 * 32 * WT_STG_LD * 4;
  static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 1) * 8 + 16 + 1024;
};

// MN-major SWIZZLE_128B descriptor: start address, LBO (stride between 64-element MN atoms), SBO (stride between 8-row K groups)
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)lbo << 16) | ((uint64_t)sbo << 32) | (1ull << 46) | (2ull << 61);
}

template <int TN>
__global__ void __launch_bounds__(192, 1) wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                          const __grid_constant__ CUtensorMap map_b, const WtArgs t) {
  using S = WtSmem<TN>;
  constexpr int STAGES = S::STAGES;
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + S::BAR_OFF);
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;
  uint32_t* tmem_slot = (uint32_t*)(acc_full + 1);
  const WgradArgs& w = t.w;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  // work item
  int item = blockIdx.x;
  const int sp = item % t.splits; item /= t.splits;
  const int tk = item % t.tiles_k; item /= t.tiles_k;
  const int tn = item % t.tiles_n;
  const int z = item / t.tiles_n;
  const int n0 = tn * WT_TM, k0 = tk * TN;
  const int slab0 = sp * t.slabs_per_split;
  const int nslab = min(t.slabs_per_split, t.total_slabs - slab0);  // >= 1 by construction

  if (warp == 0) {
    {   // converged warp, one elected lane issues (see gemm_tc.cu)
      const long long ao = w.az.off(z, w.zin, 1), bo = w.bz.off(z, w.zin, 1);
      const int a_row = (int)(ao / t.a_ld), a_col = (int)(ao % t.a_ld) + n0;
      const int b_row = (int)(bo / t.b_ld), b_col = (int)(bo % t.b_ld) + k0;
      for (int i = 0; i < nslab; ++i) {
        const int s = i % STAGES;
        mbar_wait(&empty[s], ((i / STAGES) & 1) ^ 1);
        if (elect_one_lane()) {
          mbar_expect_tx(&full[s], S::STAGE);
          uint8_t* st = smem + s * S::STAGE;
          const int m = (slab0 + i) * WT_R;
#pragma unroll
          for (int j = 0; j < 2; ++j) tma_load_2d(st + j * WT_BOX, &map_a, &full[s], a_col + j * 64, a_row + m);
#pragma unroll
          for (int j = 0; j < TN / 64; ++j) tma_load_2d(st + S::A_BYTES + j * WT_BOX, &map_b, &full[s], b_col + j * 64, b_row + m);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    {
      // D = f32, A = B = bf16, both MN-major (bits 15, 16), N = TN, M = 128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(TN >> 3) << 17) |
                             ((uint32_t)(WT_TM >> 4) << 24);
      for (int i = 0; i < nslab; ++i) {
        const int s = i % STAGES;
        mbar_wait(&full[s], (i / STAGES) & 1);
        tc_fence_after();
        if (elect_one_lane()) {
          const uint32_t sa = smem_u32(smem + s * S::STAGE);
          const uint64_t ad = make_desc_mn(sa, t.lbo, t.sbo), bd = make_desc_mn(sa + S::A_BYTES, t.lbo, t.sbo);
#pragma unroll
          for (int k = 0; k < WT_R / 16; ++k)  // 16 m-rows per instruction = two 8-row groups = 2048 bytes
            tc_mma<false>(tmem_base, ad + (uint64_t)(k * 128), bd + (uint64_t)(k * 128), idesc, (i | k) ? 1u : 0u);
          tc_commit(&empty[s]);
          if (i == nslab - 1) tc_commit(acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    const int lg = warp & 3;  // TMEM lane quarter this warp may read
    float* stg = reinterpret_cast<float*>(smem + S::STG_OFF) + (warp - 2) * 32 * WT_STG_LD;
    float* C = w.C + w.cz.off(z, w.zin, 1);
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int row0 = n0 + lg * 32;
    const int nrows = min(32, w.N - row0);
#pragma unroll 1
    for (int c0 = 0; c0 < TN; c0 += 32) {
      if (k0 + c0 >= w.K) break;  // warp-uniform
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, r);
#pragma unroll
      for (int j = 0; j < 32; ++j) stg[lane * WT_STG_LD + j] = __uint_as_float(r[j]);
      __syncwarp();
      const int col = k0 + c0 + lane;
      if (col < w.K)
        for (int rr = 0; rr < nrows; ++rr) atomicAdd(C + (long long)(row0 + rr) * w.ldc + col, w.alpha * stg[rr * WT_STG_LD + lane]);
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tmem_dealloc(tmem_base, TN);
  }
}

// colsum[z][n] += sum_m A[z][m,n]   (bias gradients); bf16 A, N % 8 == 0, N / 8 <= 256.  grid (row chunks, Z).
__global__ void __launch_bounds__(256) colsum_kernel(WgradArgs w, int rows_per_cta) {
  __shared__ float red[256][9];
  pdl_trigger();
  pdl_wait();
  const int z = blockIdx.y;
  const int ncg = w.N / 8, rl = 256 / ncg;  // column groups, row lanes
  const int cg = threadIdx.x % ncg, lane_r = threadIdx.x / ncg;
  const __nv_bfloat16* A = reinterpret_cast<const __nv_bfloat16*>(w.A) + w.az.off(z, w.zin, 1) + cg * 8;
  const int m0 = blockIdx.x * rows_per_cta, m1 = min(w.M, m0 + rows_per_cta);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (lane_r < rl) {
    int m = m0 + lane_r;
    for (; m + 3 * rl < m1; m += 4 * rl) {  // four independent 16-byte loads in flight
      uint4 u[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) u[q] = *reinterpret_cast<const uint4*>(A + (long long)(m + q * rl) * w.lda);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u[q]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = __bfloat1622float2(p[j]);
          acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
        }
      }
    }
    for (; m < m1; m += rl) {
      const uint4 u = *reinterpret_cast<const uint4*>(A + (long long)m * w.lda);
      const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(p[j]);
        acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.x][j] = acc[j];
  __syncthreads();
  if (lane_r == 0) {
    float* out = w.colsum + w.sz.off(z, w.zin, 1) + cg * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
      for (int q = 0; q < rl; ++q) s += red[q * ncg + cg][j];
      atomicAdd(out + j, s);
    }
  }
}

long long batch_span_rows(const BatchMap& bm, int Z, int zin, long long ld, long long M, bool* ok) {
  long long rows = 0;
  for (int z = 0; z < Z; ++z) {
    const long long off = bm.off(z, zin, 1);
    if (off < 0) { *ok = false; return 0; }
    const long long r = off / ld + M;
    rows = r > rows ? r : rows;
  }
  return rows;
}

template <int TN>
int launch_wt(const WtArgs& t, const CUtensorMap& ma, const CUtensorMap& mb, int ctas, cudaStream_t st) {
  using S = WtSmem<TN>;
  auto kern = wgrad_tc_kernel<TN>;
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    configured = true;
  }
  ProfScope ps(t.w.tag, st);
  BIST_LAUNCH(kern, (unsigned)ctas, 192, S::TOTAL, st, ma, mb, t);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace

// *taken = false when the shape is left to the mma.sync / CUDA-core kernels of bwd.cu.
int wgrad_tc(const WgradArgs& w, cudaStream_t st, bool* taken) {
  *taken = false;
  static const bool disabled = getenv("BIST_DISABLE_TC") != nullptr || getenv("BIST_DISABLE_WGRAD_TC") != nullptr;
  if (disabled) return 0;
  if (w.a_dtype != BIST_BF16 || w.b_dtype != BIST_BF16) return 0;
  if (w.M % WT_R != 0 || w.M < 8 * WT_R || w.N < 16 || w.K < 64 || w.N % 8 != 0 || w.K % 8 != 0) return 0;
  // small products are launch/latency-bound: the lighter mma.sync kernel (one launch, fused bias sums) wins there
  static const long long min_macs = getenv("BIST_WGRAD_TC_MIN") ? atoll(getenv("BIST_WGRAD_TC_MIN")) : (1LL << 31);
  if ((long long)w.M * w.N * w.K * w.Z < min_macs) return 0;
  if ((w.lda * 2) % 16 || (w.ldb * 2) % 16 || ((uintptr_t)w.A % 16) || ((uintptr_t)w.B % 16)) return 0;
  if (w.colsum && (w.N / 8 > 256)) return 0;
  bool ok = true;
  const long long a_rows = batch_span_rows(w.az, w.Z, w.zin, w.lda, w.M, &ok);
  const long long b_rows = batch_span_rows(w.bz, w.Z, w.zin, w.ldb, w.M, &ok);
  if (!ok) return 0;
  for (int z = 0; z < w.Z; ++z) {  // 16-byte aligned column offsets
    if (((w.az.off(z, w.zin, 1) % w.lda) * 2) % 16 || ((w.bz.off(z, w.zin, 1) % w.ldb) * 2) % 16) return 0;
  }
  *taken = true;

  WtArgs t;
  t.w = w;
  t.a_ld = w.lda; t.b_ld = w.ldb;
  const int TN = w.K > 128 ? 256 : 128;
  t.tiles_n = (w.N + WT_TM - 1) / WT_TM;
  t.tiles_k = (w.K + TN - 1) / TN;
  t.total_slabs = w.M / WT_R;
  const long long items = (long long)w.Z * t.tiles_n * t.tiles_k;
  long long splits = (num_sms() + items - 1) / items;
  const long long max_splits = t.total_slabs / 2 > 0 ? t.total_slabs / 2 : 1;
  splits = splits < 1 ? 1 : (splits > max_splits ? max_splits : splits);
  t.slabs_per_split = (int)((t.total_slabs + splits - 1) / splits);
  t.splits = (t.total_slabs + t.slabs_per_split - 1) / t.slabs_per_split;
  // descriptor strides: MN atoms (64 columns) are WT_BOX bytes apart, 8-row K groups 1024 bytes apart
  static const bool swap = getenv("BIST_WGRAD_TC_SWAP") != nullptr;  // bring-up aid
  t.lbo = (swap ? 1024u : (uint32_t)WT_BOX) >> 4;
  t.sbo = (swap ? (uint32_t)WT_BOX : 1024u) >> 4;
  CUtensorMap ma, mb;
  BIST_TRY(make_map_2d(&ma, w.A, false, w.lda, a_rows, 64, WT_R, 128));
  BIST_TRY(make_map_2d(&mb, w.B, false, w.ldb, b_rows, 64, WT_R, 128));
  const long long ctas = items * t.splits;
  if (ctas > 0x7fffffffLL) { *taken = false; return 0; }
  if (w.colsum) {
    const int rows_per_cta = 128;
    ProfScope ps("colsum", st);
    BIST_LAUNCH(colsum_kernel, dim3((unsigned)((w.M + rows_per_cta - 1) / rows_per_cta), (unsigned)w.Z), 256, 0, st, w, rows_per_cta);
    BIST_LAUNCHED();
  }
  return TN == 256 ? launch_wt<256>(t, ma, mb, (int)ctas, st) : launch_wt<128>(t, ma, mb, (int)ctas, st);
}

}  // namespace bist
