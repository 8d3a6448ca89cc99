// tcgen05 / TMEM / TMA GEMM for sm_100a:  C = act(alpha * A . W^T + bias) + R
//   A [M,K] and W [N,K] both K-major (dense rows), bf16 (kind::f16) or fp32 read as tf32
//   (kind::tf32); fp32 accumulation in tensor memory.
// One 128 x BN output tile per CTA, 192 threads, warp-specialised:
//   warp 0    TMA producer   cp.async.bulk.tensor (128B-swizzled, 128-byte-wide K slabs) -> smem ring
//   warp 1    MMA issuer     one elected lane issues tcgen05.mma (UMMA 128 x BN x 16|8) into TMEM;
//                            tcgen05.commit releases ring slots and publishes the accumulator;
//                            the warp also owns tcgen05.alloc / dealloc
//   warps 2-5 epilogue       tcgen05.ld (32 lanes x 32 columns) -> registers -> alpha / bias / ReLU /
//                            residual -> global; each thread owns one output row
// The ring is synchronised with mbarriers (full: TMA complete_tx; empty: tcgen05.commit).
#include "tc_ptx.cuh"

#include <stdlib.h>

#include <mutex>

namespace bist {
namespace {
using namespace tc;

struct TcArgs {
  GemmArgs g;
  long long a_ld, w_ld;  // leading dimensions (elements) of the 2-D tensors the TMA maps describe
  int nkb;               // K / (elements per 128-byte slab)
  int vec_c;             // 1 = 16-byte aligned output rows (vector stores)
};

template <int BN, int STAGES>
struct Smem {
  static constexpr int A_BYTES = BM * SLAB_BYTES, W_BYTES = BN * SLAB_BYTES;
  static constexpr int STAGE_BYTES = A_BYTES + W_BYTES;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 1) * 8 + 16 + 1024;  // + alignment slack
};

template <int BN, int STAGES, bool TF32>
__global__ void __launch_bounds__(192) gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                      const __grid_constant__ CUtensorMap map_w, const TcArgs t) {
  using S = Smem<BN, STAGES>;
  constexpr int ESZ = TF32 ? 4 : 2;
  constexpr int BKE = SLAB_BYTES / ESZ;        // elements of K per slab: 64 bf16 / 32 tf32
  constexpr int UMMA_K = 32 / ESZ;             // 16 bf16 / 8 tf32 (32 bytes of K per instruction)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + S::BAR_OFF);
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;
  uint32_t* tmem_slot = (uint32_t*)(acc_full + 1);

  const GemmArgs& g = t.g;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.z, m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // batch offsets become (row, column) coordinates inside the 2-D tensors
      const long long ao = g.az.off(z, g.zin), wo = g.wz.off(z, g.zin);
      const int a_row = (int)(ao / t.a_ld) + m0, a_col = (int)(ao % t.a_ld);
      const int w_row = (int)(wo / t.w_ld) + n0, w_col = (int)(wo % t.w_ld);
      for (int kb = 0; kb < t.nkb; ++kb) {
        const int s = kb % STAGES;
        if (kb >= STAGES) mbar_wait(&empty[s], ((kb / STAGES) - 1) & 1);
        mbar_expect_tx(&full[s], S::STAGE_BYTES);
        uint8_t* st = smem + s * S::STAGE_BYTES;
        tma_load_2d(st, &map_a, &full[s], a_col + kb * BKE, a_row);
        tma_load_2d(st + S::A_BYTES, &map_w, &full[s], w_col + kb * BKE, w_row);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): D = f32, A/B = bf16 | tf32, K-major both
      const uint32_t fmt = TF32 ? 2u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      for (int kb = 0; kb < t.nkb; ++kb) {
        const int s = kb % STAGES;
        mbar_wait(&full[s], (kb / STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
        const uint64_t ad = make_desc(sa), wd = make_desc(sa + S::A_BYTES);
#pragma unroll
        for (int k = 0; k < BKE / UMMA_K; ++k)  // advance 32 bytes of K inside the swizzle atom: +2 in the address field
          tc_mma<TF32>(tmem_base, ad + 2 * k, wd + 2 * k, idesc, (kb | k) ? 1u : 0u);
        tc_commit(&empty[s]);
      }
      tc_commit(acc_full);
    }
  } else {
    // ---- epilogue: warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32) ------------------------
    const int lg = warp & 3;
    const int row = m0 + lg * 32 + lane;
    const bool row_ok = row < g.M;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const float* bias = g.bias ? g.bias + g.bz.off(z, g.zin) : nullptr;
    const long long co = row_ok ? g.cz.off(z, g.zin) + g.c.off(row) : 0;
    const long long ro = (row_ok && g.R) ? g.rz.off(z, g.zin) + g.r.off(row) : 0;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      if (n0 + c0 >= g.N) break;  // warp-uniform
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + c0, r);
      if (!row_ok) continue;
      const int col = n0 + c0;
      if (t.vec_c && col + 32 <= g.N) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          float4 v = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                 __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
          v.x *= g.alpha; v.y *= g.alpha; v.z *= g.alpha; v.w *= g.alpha;
          if (bias) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col + 4 * q));
            v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
          }
          if (g.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
          if (g.R) {
            const float4 x = g.r_dtype == BIST_BF16 ? ld4(reinterpret_cast<const __nv_bfloat16*>(g.R) + ro + col + 4 * q)
                                                    : ld4(reinterpret_cast<const float*>(g.R) + ro + col + 4 * q);
            v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
          }
          if (g.c_dtype == BIST_BF16) st4(reinterpret_cast<__nv_bfloat16*>(g.C) + co + col + 4 * q, v);
          else st4(reinterpret_cast<float*>(g.C) + co + col + 4 * q, v);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int cc = col + j;
          if (cc < g.N) {
            float v = g.alpha * __uint_as_float(r[j]);
            if (bias) v += bias[cc];
            if (g.relu) v = fmaxf(v, 0.f);
            if (g.R) v += g.r_dtype == BIST_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g.R)[ro + cc])
                                                 : reinterpret_cast<const float*>(g.R)[ro + cc];
            if (g.c_dtype == BIST_BF16) reinterpret_cast<__nv_bfloat16*>(g.C)[co + cc] = __float2bfloat16_rn(v);
            else reinterpret_cast<float*>(g.C)[co + cc] = v;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tmem_dealloc(tmem_base, BN);
  }
}

// ---------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

// 2-D tensor [rows, ld] (ld contiguous) with a [box_rows, 128 bytes] box, 128B swizzle, zero OOB fill.
int make_map(CUtensorMap* m, const void* base, bool f32, long long ld, long long rows, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return BIST_ERR_CUDA; }
  const int esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(SLAB_BYTES / esz), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base),
                  dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d): ld=%lld rows=%lld", (int)r, ld, rows); return BIST_ERR_CUDA; }
  return 0;
}

template <int BN, int STAGES, bool TF32>
int launch(const TcArgs& t, const CUtensorMap& ma, const CUtensorMap& mw, cudaStream_t st) {
  auto kern = gemm_tc_kernel<BN, STAGES, TF32>;
  constexpr int smem = Smem<BN, STAGES>::TOTAL;
  static bool configured = false;  // per instantiation
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid((t.g.N + BN - 1) / BN, (t.g.M + BM - 1) / BM, t.g.Z);
  ProfScope ps(t.g.tag, st);
  kern<<<grid, 192, smem, st>>>(ma, mw, t);
  BIST_LAUNCHED();
  return 0;
}

// Can the (row offset, column offset) decomposition of every batch offset be used?  Returns the
// number of tensor rows needed, or -1.
long long batch_rows(const BatchMap& bm, int Z, int zin, long long ld, int K, int rows_per_batch) {
  long long max_row = 0;
  for (int z = 0; z < Z; ++z) {
    const long long off = bm.off(z, zin);
    if (off < 0 || off % ld + K > ld) return -1;
    max_row = off / ld > max_row ? off / ld : max_row;
  }
  return max_row + rows_per_batch;
}

}  // namespace

int gemm_tc(const GemmArgs& g, cudaStream_t st, bool* taken) {
  *taken = false;
  static const bool disabled = getenv("BIST_DISABLE_TC") != nullptr;  // debugging aid: force the CUDA-core GEMM
  if (disabled) return 0;
  const bool bf = g.a_dtype == BIST_BF16 && g.w_dtype == BIST_BF16;
  const bool tf = g.a_dtype == BIST_F32 && g.w_dtype == BIST_F32 && g.allow_tf32;
  if (!bf && !tf) return 0;
  const int esz = bf ? 2 : 4, bke = SLAB_BYTES / esz;
  if (g.K % bke != 0 || g.M < 1 || g.N < 8 || g.Z > 65535) return 0;
  if (g.a.inner != kNoSplit || g.w.inner != kNoSplit) return 0;  // operands must be dense row-major
  const long long a_ld = g.a.si, w_ld = g.w.si;
  if (a_ld < g.K || w_ld < g.K || (a_ld * esz) % 16 || (w_ld * esz) % 16) return 0;
  if (((uintptr_t)g.A % 16) || ((uintptr_t)g.W % 16)) return 0;
  if ((long long)g.M * g.N * g.Z < 64 * 64) return 0;  // tiny products stay on the CUDA cores
  const long long a_rows = batch_rows(g.az, g.Z, g.zin, a_ld, g.K, g.M);
  const long long w_rows = batch_rows(g.wz, g.Z, g.zin, w_ld, g.K, g.N);
  if (a_rows < 0 || w_rows < 0) return 0;
  *taken = true;

  TcArgs t;
  t.g = g;
  t.a_ld = a_ld; t.w_ld = w_ld; t.nkb = g.K / bke;
  const int csz = g.c_dtype == BIST_BF16 ? 2 : 4;
  auto al = [&](long long v, int sz) { return (v * sz) % 16 == 0; };
  t.vec_c = ((uintptr_t)g.C % 16 == 0) && al(g.c.si, csz) && al(g.c.so, csz) && al(g.cz.s0, csz) && al(g.cz.s1, csz) &&
            (!g.bias || (((uintptr_t)g.bias % 16 == 0) && g.bz.s0 % 4 == 0 && g.bz.s1 % 4 == 0));
  if (g.R) {
    const int rsz = g.r_dtype == BIST_BF16 ? 2 : 4;
    t.vec_c = t.vec_c && ((uintptr_t)g.R % 16 == 0) && al(g.r.si, rsz) && al(g.r.so, rsz) && al(g.rz.s0, rsz) && al(g.rz.s1, rsz);
  }
  const long long tiles256 = (long long)((g.M + BM - 1) / BM) * (g.N / 256) * g.Z;
  const int bn = (g.N % 256 == 0 && tiles256 >= 120) ? 256 : (g.N > 64 ? 128 : 64);
  CUtensorMap ma, mw;
  BIST_TRY(make_map(&ma, g.A, !bf, a_ld, a_rows, BM));
  BIST_TRY(make_map(&mw, g.W, !bf, w_ld, w_rows, bn));
  if (bf) {
    if (bn == 256) return launch<256, 4, false>(t, ma, mw, st);
    if (bn == 128) return launch<128, 3, false>(t, ma, mw, st);
    return launch<64, 4, false>(t, ma, mw, st);
  }
  if (bn == 256) return launch<256, 4, true>(t, ma, mw, st);
  if (bn == 128) return launch<128, 3, true>(t, ma, mw, st);
  return launch<64, 4, true>(t, ma, mw, st);
}

}  // namespace bist
