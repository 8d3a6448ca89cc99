// Query side of VidEncoderLayer4 (/root/reference/model/encoder.py:172-201) in bf16 mode: the chain of small dependent products
// on the B*Lq query rows of both directions, collapsed into three kernels that each start from fp32 rows, normalise them in the
// kernel (LayerNorm of modules.py:20-31: unbiased std, eps on the std) straight into the tensor core's A operand, and finish
// with an epilogue that used to be one or two more launches:
//   ATTN  x_in -> LN0 -> [q_h | k_h | v_h] (one UMMA, N = 192) -> masked self-attention of the head's 16-token samples
//         (encoder.py:176/:184, modules.py:54-64) -> ctx[:, head]                          (was: LN, QKV GEMM, attention)
//   FOLD  x -> LN1 / LN2 -> Q1 / Q2 head pair (N = 128) -> bf16 A operand in shared memory -> . (W_k / sqrt(dk))^T per head
//         (the key-projection fold of DESIGN.md §3) -> qf[which][q][head][d]; short units: x . W_v2^T (the raw-x value term)
//                                                                                           (was: LN2, Q12X GEMM, fold GEMM)
//   FFN1  ts -> LN3 -> relu(. W_1^T + b_1) (N = 256 per CTA) -> ffh                        (was: LN, GEMM)  modules.py:102-113
// A tile is 128 rows = 8 samples x 16 token slots of one direction (token slots >= Lq and samples >= B are zero rows whose
// results the TMA store clips), so the self-attention of a sample is local to half a warp of the epilogue.
// One CTA per unit (no persistence: at config 2 each kernel is 128-192 CTAs on 148 SMs), 320 threads:
//   warp 0  B-operand producer (TMA, ring of uniform slabs)      warp 1  TMEM allocation + UMMA issuer
//   warps 2-9  LayerNorm A-producer (one sample per warp), then the epilogue (two warps per TMEM lane quarter)
#include <stdlib.h>

#include "tc_ptx.cuh"

namespace bist {
__device__ long long g_qtrace[3][4][16];   // optional timeline (bist_debug_qside_trace): [mode][CTA 0, 1, last-1, last][event] clock64 / globaltimer
namespace {
using namespace tc;

enum { QS_ATTN = 0, QS_FOLD = 1, QS_FFN1 = 2 };
constexpr int QS_MAXKS = 8;                          // d <= 512
constexpr int QS_ASLAB = 128 * 128;                  // 128 rows x 64 k bf16, SWIZZLE_128B K-major
constexpr int QS_OFF_A = 0;                          // A operand; dead after the main product: epilogue staging
constexpr int QS_OFF_RING = QS_MAXKS * QS_ASLAB;     // 131072: B ring, 96 KB
constexpr int QS_RING_BYTES = 96 * 1024;
constexpr int QS_OFF_A2 = QS_OFF_RING + 64 * 1024;   // FOLD: the head pair's projected queries (2 x 16 KB) behind a 4 x 16 KB ring
constexpr int QS_OFF_BAR = QS_OFF_RING + QS_RING_BYTES;
constexpr int QS_MAXST = 4;
constexpr int QS_OFF_BIAS = QS_OFF_BAR + 256;       // 512 floats: the unit's bias slice, fetched while the main product runs
constexpr int QS_SMEM = QS_OFF_BIAS + 2048;
static_assert(QS_SMEM <= 227 * 1024, "qside: shared memory budget");
// ATTN epilogue scratch inside the dead A region
constexpr int QS_KV_ROW = 256;                       // 64 fp32 per row
constexpr int QS_OFF_K = 0, QS_OFF_V = 33 * 1024, QS_OFF_P = 66 * 1024, QS_OFF_CTX = 74 * 1024;
constexpr float kMaskFill = -1e9f;                   // modules.py:60

template <int MODE> struct QsCfg;
template <> struct QsCfg<QS_ATTN> { static constexpr int NB = 192, STAGES = 4; };
template <> struct QsCfg<QS_FOLD> { static constexpr int NB = 128, STAGES = 4; };
template <> struct QsCfg<QS_FFN1> { static constexpr int NB = 256, STAGES = 3; };

struct QsArgs {
  const float* ln_a[2][2];   // [direction][which] LayerNorm gain / bias (FOLD: which = 0, 1; others: [u][0])
  const float* ln_b[2][2];
  const float* bias[2];      // per direction: ATTN [3][d] (q | k | v); FOLD [3][d] (q1 | q2 | raw-x value term); FFN1 [dff]
  const uint8_t* qmask;      // [B, Lq] (ATTN)
  int nd, B, Lq, d, h, dff, nks, tpd;   // tpd: tiles per direction = ceil(B / 8)
  int nlong;                 // FOLD: units that run the fold (the rest are the short raw-x units)
  float eps, scale;          // scale = 1 / sqrt(d_k)
  float inv_d, inv_dm1;      // 1 / d, 1 / (d - 1)
  int trace;
};
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define QTRACE(ev) do { if (p.trace && tslot >= 0) g_qtrace[MODE][tslot][ev] = gtimer(); } while (0)

// ---- PTX ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read3() { asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void st_shared_v4(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
// 32 accumulator columns (+ bias) of one row -> bf16 -> four 16-byte chunks c0 .. c0+3 of a SWIZZLE_128B row (128-byte rows)
__device__ __forceinline__ void stage_bf16_32(uint32_t row_addr, int row, int c0, const uint32_t* r, const float* bias, bool relu) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; i += 4) {
      const float4 b = *reinterpret_cast<const float4*>(bias + 8 * c + i);
      v[i] = __uint_as_float(r[8 * c + i]) + b.x; v[i + 1] = __uint_as_float(r[8 * c + i + 1]) + b.y;
      v[i + 2] = __uint_as_float(r[8 * c + i + 2]) + b.z; v[i + 3] = __uint_as_float(r[8 * c + i + 3]) + b.w;
    }
    if (relu) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
    }
    st_shared_v4(row_addr + (uint32_t)((((c0 + c) ^ row) & 7) << 4), pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
  }
}

template <int MODE>
__global__ void __launch_bounds__(320, 1) qside_kernel(const __grid_constant__ CUtensorMap map_b0, const __grid_constant__ CUtensorMap map_b1,
                                                       const __grid_constant__ CUtensorMap map_c0, const __grid_constant__ CUtensorMap map_c1,
                                                       const __grid_constant__ CUtensorMap map_o0, const __grid_constant__ CUtensorMap map_o1,
                                                       const __grid_constant__ CUtensorMap map_o2, const __grid_constant__ CUtensorMap map_a,
                                                       const QsArgs p) {
  constexpr int NB = QsCfg<MODE>::NB, ST = QsCfg<MODE>::STAGES, SLAB = NB * 128;
  static_assert(ST * SLAB <= QS_RING_BYTES && ST <= QS_MAXST, "qside: ring");
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = (uint64_t*)(smem + QS_OFF_BAR);
  uint64_t* bfull = bars;                     // [4]
  uint64_t* bempty = bfull + QS_MAXST;        // [4]
  uint64_t* a_full = bempty + QS_MAXST;       // normalised rows are in shared memory (8 warp arrivals)
  uint64_t* acc_full = a_full + 1;
  uint64_t* a2_full = acc_full + 1;           // FOLD: projected queries of the head pair staged (8 warp arrivals)
  uint64_t* d2_full = a2_full + 1;            // [3] FOLD: a 128-column chunk of folded queries is in TMEM
  uint64_t* d2_empty = d2_full + 3;           // [3] (8 warp arrivals)
  uint64_t* a_land = d2_empty + 3;            // the bf16 rows of the tile have landed (TMA)
  uint32_t* tmem_slot = (uint32_t*)(a_land + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nks = p.nks, d = p.d;
  const int tslot = (threadIdx.x & 31) != 0 ? -1 : blockIdx.x < 2 ? (int)blockIdx.x : blockIdx.x + 2 >= gridDim.x ? (int)(blockIdx.x + 4 - gridDim.x) : -1;
  if (warp == 2) QTRACE(0);

  // ---- unit -> tile (direction u, first sample b0) and the mode's own coordinates ---------------------------------------------
  int tile, which = 0, sub = 0;
  bool fold = false;
  if (MODE == QS_ATTN) { tile = blockIdx.x / p.h; sub = blockIdx.x % p.h; }                       // sub = head
  else if (MODE == QS_FFN1) { const int nc = p.dff >> 8; tile = blockIdx.x / nc; sub = blockIdx.x % nc; }   // sub = 256-column chunk
  else {
    const int hp = p.h >> 1, nq = d >> 7;
    if ((int)blockIdx.x < p.nlong) { fold = true; tile = blockIdx.x / (2 * hp); which = (blockIdx.x / hp) & 1; sub = blockIdx.x % hp; }   // sub = head pair
    else { tile = blockIdx.x - p.nlong; which = 2; (void)nq; }                                    // all d / 128 column chunks of x . W_v2^T
  }
  const int u = tile / p.tpd, b0 = (tile - u * p.tpd) * 8;
  const CUtensorMap* mb = u ? &map_b1 : &map_b0;
  const CUtensorMap* mc = u ? &map_c1 : &map_c0;

  if ((smem_u32(smem) & 1023u) != 0u) __trap();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < QS_MAXST; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
    mbar_init(a_full, 8); mbar_init(acc_full, 1); mbar_init(a2_full, 8); mbar_init(a_land, 1);
    for (int s = 0; s < 3; ++s) { mbar_init(&d2_full[s], 1); mbar_init(&d2_empty[s], 8); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 2) QTRACE(1);
  pdl_wait();
  if (warp == 2) QTRACE(2);

  if (warp == 0) {
    // ---- A tile (8 samples x 16 token slots x d of the bf16 rows; slots >= Lq and samples >= B are zero-filled), then the B ring ----
    if (elect_one()) {
      mbar_expect_tx(a_land, (uint32_t)(nks * QS_ASLAB));
      for (int ks = 0; ks < nks; ++ks) tma_load_4d(smem + QS_OFF_A + ks * QS_ASLAB, &map_a, a_land, ks * 64, 0, b0, u);
    }
    __syncwarp();
    int it = 0;
    for (int ks = 0; ks < nks; ++ks, ++it) {
      const int s = it % ST;
      mbar_wait(&bempty[s], ((it / ST) & 1) ^ 1);
      if (elect_one()) {
        uint8_t* dst = smem + QS_OFF_RING + s * SLAB;
        mbar_expect_tx(&bfull[s], SLAB);
        if (MODE == QS_ATTN) {         // rows q_h | k_h | v_h of the packed [3][d][d] weight
          for (int j = 0; j < 3; ++j) tma_load_2d(dst + j * 64 * 128, mb, &bfull[s], ks * 64, j * d + sub * 64);
        } else if (MODE == QS_FFN1) {
          tma_load_2d(dst, mb, &bfull[s], ks * 64, sub * 256);
        } else {                       // W_q1 | W_q2 | W_v2 stacked [3][d][d]: 128 rows of matrix `which`
          tma_load_2d(dst, mb, &bfull[s], ks * 64, which * d + sub * 128);
        }
      }
      __syncwarp();
    }
    if (MODE == QS_FOLD && !fold) {
      for (int c = 1; c < (d >> 7); ++c)      // the other column chunks of the raw-x value term reuse the resident rows
        for (int ks = 0; ks < nks; ++ks, ++it) {
          const int s = it % ST;
          mbar_wait(&bempty[s], ((it / ST) & 1) ^ 1);
          if (elect_one()) {
            mbar_expect_tx(&bfull[s], SLAB);
            tma_load_2d(smem + QS_OFF_RING + s * SLAB, mb, &bfull[s], ks * 64, 2 * d + c * 128);
          }
          __syncwarp();
        }
    }
    if (MODE == QS_FOLD && fold) {
      for (int i = 0; i < 8; ++i, ++it) {   // (W_k / sqrt(dk))^T pieces: 128 output features x the head's 64 k
        const int s = it % ST;
        mbar_wait(&bempty[s], ((it / ST) & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&bfull[s], SLAB);
          tma_load_2d(smem + QS_OFF_RING + s * SLAB, mc, &bfull[s], (2 * sub + (i >> 2)) * 64, which * d + (i & 3) * 128);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ---- UMMA issuer ---------------------------------------------------------------------------------------------------------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NB >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(a_full, 0);
    tc_fence_after();
    int it = 0;
    for (int ks = 0; ks < nks; ++ks, ++it) {
      const int s = it % ST;
      mbar_wait(&bfull[s], (it / ST) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t ad = make_desc(smem_u32(smem + QS_OFF_A + ks * QS_ASLAB)), bd = make_desc(smem_u32(smem + QS_OFF_RING + s * SLAB));
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma<false>(tmem_base, ad + 2 * k, bd + 2 * k, idesc, (ks | k) ? 1u : 0u);
        tc_commit(&bempty[s]);
        if (ks == nks - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    if (MODE == QS_FOLD && !fold) {
      for (int c = 1; c < (d >> 7); ++c)
        for (int ks = 0; ks < nks; ++ks, ++it) {
          const int s = it % ST;
          mbar_wait(&bfull[s], (it / ST) & 1);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t ad = make_desc(smem_u32(smem + QS_OFF_A + ks * QS_ASLAB)), bd = make_desc(smem_u32(smem + QS_OFF_RING + s * SLAB));
#pragma unroll
            for (int k = 0; k < 4; ++k) tc_mma<false>(tmem_base + (uint32_t)(c * 128), ad + 2 * k, bd + 2 * k, idesc, (ks | k) ? 1u : 0u);
            tc_commit(&bempty[s]);
            if (ks == nks - 1) tc_commit(&d2_full[c - 1]);
          }
          __syncwarp();
        }
    }
    if (MODE == QS_FOLD && fold) {
      mbar_wait(a2_full, 0);
      tc_fence_after();
      for (int i = 0; i < 8; ++i, ++it) {
        const int s = it % ST, buf = i % 3;
        mbar_wait(&bfull[s], (it / ST) & 1);
        mbar_wait(&d2_empty[buf], ((i / 3) & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t ad = make_desc(smem_u32(smem + QS_OFF_A2 + (i >> 2) * QS_ASLAB)), bd = make_desc(smem_u32(smem + QS_OFF_RING + s * SLAB));
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma<false>(tmem_base + 128u + (uint32_t)(buf * 128), ad + 2 * k, bd + 2 * k, idesc, k ? 1u : 0u);
          tc_commit(&bempty[s]);
          tc_commit(&d2_full[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    // ================================ worker warps ==================================================================================
    const int w8 = warp - 2;                      // 0..7: during the LayerNorm, warp w8 owns sample slot w8 (16 token rows)
    {
      // LayerNorm IN PLACE on the landed tile (modules.py:28-31: unbiased std, eps on the std; statistics in fp32): warp w8 owns
      // sample slot w8; a lane holds chunk (lane & 7) of k-slabs (lane >> 3) and (lane >> 3) + 4 of a row = 16 values.
      const bool do_ln = MODE == QS_FFN1 || (MODE == QS_FOLD && which < 2);
      mbar_wait(a_land, 0);
      if (warp == 2) QTRACE(8);
      if (do_ln && b0 + w8 < p.B) {
        const int lw = MODE == QS_FOLD ? (which & 1) : 0;
        const int chunk = lane & 7, sl0 = lane >> 3;
        const bool on0 = sl0 < nks, on1 = sl0 + 4 < nks;
        float ga[16], gb[16];
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const int col = (sl0 + 4 * t) * 64 + chunk * 8;
#pragma unroll
          for (int i = 0; i < 8; i += 4) {
            float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f), b4 = a4;
            if (t ? on1 : on0) { a4 = __ldg(reinterpret_cast<const float4*>(p.ln_a[u][lw] + col + i)); b4 = __ldg(reinterpret_cast<const float4*>(p.ln_b[u][lw] + col + i)); }
            ga[8 * t + i] = a4.x; ga[8 * t + i + 1] = a4.y; ga[8 * t + i + 2] = a4.z; ga[8 * t + i + 3] = a4.w;
            gb[8 * t + i] = b4.x; gb[8 * t + i + 1] = b4.y; gb[8 * t + i + 2] = b4.z; gb[8 * t + i + 3] = b4.w;
          }
        }
        const uint32_t abase = smem_u32(smem + QS_OFF_A) + (uint32_t)(sl0 * QS_ASLAB);
        if (warp == 2) QTRACE(9);
#pragma unroll 1
        for (int l0 = 0; l0 < p.Lq; l0 += 4) {     // four rows at a time: their reductions interleave
          float x[4][16], s[4], ss[4];
          uint32_t ra[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int r = w8 * 16 + ((l0 + i) & 15);   // (slots past Lq hold zeros: computed with the others, never stored)
            ra[i] = abase + (uint32_t)(r * 128 + ((chunk ^ (r & 7)) << 4));
            uint32_t w[8];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              w[4 * t] = w[4 * t + 1] = w[4 * t + 2] = w[4 * t + 3] = 0u;
              if (t ? on1 : on0)
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[4 * t]), "=r"(w[4 * t + 1]), "=r"(w[4 * t + 2]), "=r"(w[4 * t + 3])
                             : "r"(ra[i] + (uint32_t)(4 * t * QS_ASLAB)) : "memory");
            }
            s[i] = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              x[i][2 * j] = __uint_as_float(w[j] << 16); x[i][2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
              s[i] += x[i][2 * j] + x[i][2 * j + 1];
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int i = 0; i < 4; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], o);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float mean = s[i] * p.inv_d;
            ss[i] = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              x[i][j] -= mean;
              if (j < 8 ? on0 : on1) ss[i] = fmaf(x[i][j], x[i][j], ss[i]);
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int i = 0; i < 4; ++i) ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], o);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (l0 + i >= p.Lq) continue;         // zero slots stay zero
            // (MUFU-based square root and reciprocal: the IEEE routines are subroutine calls that serialise the four rows; the
            //  result is rounded to bf16 anyway)
            const float var = ss[i] * p.inv_dm1;
            const float inv = __fdividef(1.f, var * rsqrtf(fmaxf(var, 1e-30f)) + p.eps);
#pragma unroll
            for (int t = 0; t < 2; ++t)
              if (t ? on1 : on0)
                st_shared_v4(ra[i] + (uint32_t)(4 * t * QS_ASLAB),
                             pack2(fmaf(ga[8 * t] * inv, x[i][8 * t], gb[8 * t]), fmaf(ga[8 * t + 1] * inv, x[i][8 * t + 1], gb[8 * t + 1])),
                             pack2(fmaf(ga[8 * t + 2] * inv, x[i][8 * t + 2], gb[8 * t + 2]), fmaf(ga[8 * t + 3] * inv, x[i][8 * t + 3], gb[8 * t + 3])),
                             pack2(fmaf(ga[8 * t + 4] * inv, x[i][8 * t + 4], gb[8 * t + 4]), fmaf(ga[8 * t + 5] * inv, x[i][8 * t + 5], gb[8 * t + 5])),
                             pack2(fmaf(ga[8 * t + 6] * inv, x[i][8 * t + 6], gb[8 * t + 6]), fmaf(ga[8 * t + 7] * inv, x[i][8 * t + 7], gb[8 * t + 7])));
          }
        }
        if (warp == 2) QTRACE(10);
        fence_proxy_async_smem();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full);
      if (warp == 2) QTRACE(3);
    }
    // ---- the unit's bias slice -> shared memory (the epilogue would otherwise wait for global loads after the product is done) ----
    float* sbias = reinterpret_cast<float*>(smem + QS_OFF_BIAS);
    {
      const int t = threadIdx.x - 64;   // 0..255
      if (MODE == QS_ATTN) { if (t < 192) sbias[t] = __ldg(p.bias[u] + (t >> 6) * d + sub * 64 + (t & 63)); }
      else if (MODE == QS_FFN1) sbias[t] = __ldg(p.bias[u] + sub * 256 + t);
      else if (fold) { if (t < 128) sbias[t] = __ldg(p.bias[u] + which * d + sub * 128 + t); }
      else { for (int i = t; i < d; i += 256) sbias[i] = __ldg(p.bias[u] + 2 * d + i); }
      bar_sync(3, 256);
    }
    uint32_t mask_bits = 0;   // ATTN: key mask of this thread's sample (query_mask[b, j] == 0 -> -1e9), one bit per token slot
    if (MODE == QS_ATTN) {
      const int mrow = (warp & 3) * 32 + lane, mb_ = b0 + (mrow >> 4);
      const bool keep = mb_ < p.B && (mrow & 15) < p.Lq && p.qmask[(long long)mb_ * p.Lq + (mrow & 15)] != 0;
      mask_bits = (__ballot_sync(0xffffffffu, keep) >> (16 * (lane >> 4))) & 0xffffu;
    }
    // ---- epilogues: two warps per TMEM lane quarter -------------------------------------------------------------------------------
    const int q = warp & 3, ch = w8 >> 2;
    const int row = q * 32 + lane;                // tile row = TMEM lane: sample slot row >> 4, token row & 15
    const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t stg = smem_u32(smem + QS_OFF_A);
    const bool leader = (w8 & 3) == 0 && elect_one();   // one thread per column half issues (and waits for) the TMA stores
    mbar_wait(acc_full, 0);
    tc_fence_after();
    if (warp == 2) QTRACE(4);
    if (MODE == QS_FFN1) {
      // relu(acc + b_1) -> bf16 -> four [128 x 64] boxes -> ffh[(u, b, l), sub*256 ..)                       modules.py:113
#pragma unroll 1
      for (int bx = 0; bx < 2; ++bx) {
        const int box = ch * 2 + bx;
        const uint32_t ra = stg + (uint32_t)(box * QS_ASLAB + row * 128);
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t r[32];
          tmem_ld32(tlane + (uint32_t)(box * 64 + hf * 32), r);
          stage_bf16_32(ra, row, hf * 4, r, sbias + box * 64 + hf * 32, true);
        }
      }
      fence_proxy_async_smem();
      bar_sync(1 + ch, 128);
      if (leader)
        for (int bx = 0; bx < 2; ++bx)
          tma_store_4d(&map_o0, stg + (uint32_t)((ch * 2 + bx) * QS_ASLAB), sub * 256 + (ch * 2 + bx) * 64, 0, b0, u);
    }
    if (MODE == QS_FOLD && !fold) {
      // raw-x value term, 128 columns per pass: acc + (b_v2 + W_v2 b_o1) -> bf16 -> xv2[(u, b, l), c*128 ..).  The rows stay resident
      // for the later passes, so the staging boxes are the (unused) projected-query buffers.
      const uint32_t box = smem_u32(smem + QS_OFF_A2) + (uint32_t)(ch * QS_ASLAB);
#pragma unroll 1
      for (int c = 0; c < (d >> 7); ++c) {
        if (c > 0) {
          mbar_wait(&d2_full[c - 1], 0);
          tc_fence_after();
          if (leader) tma_store_wait_read0();
          bar_sync(1 + ch, 128);
        }
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t r[32];
          tmem_ld32(tlane + (uint32_t)(c * 128 + ch * 64 + hf * 32), r);
          stage_bf16_32(box + (uint32_t)(row * 128), row, hf * 4, r, sbias + c * 128 + ch * 64 + hf * 32, false);
        }
        fence_proxy_async_smem();
        bar_sync(1 + ch, 128);
        if (leader) tma_store_4d(&map_o2, box, c * 128 + ch * 64, 0, b0, u);
      }
    }
    if (MODE == QS_FOLD && fold) {
      // (1) projected queries of head 2*sub + ch (+ bias) -> bf16 A operand of the fold                       encoder.py:121 / :148
      {
        const uint32_t ra = smem_u32(smem + QS_OFF_A2) + (uint32_t)(ch * QS_ASLAB + row * 128);
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t r[32];
          tmem_ld32(tlane + (uint32_t)(ch * 64 + hf * 32), r);
          stage_bf16_32(ra, row, hf * 4, r, sbias + ch * 64 + hf * 32, false);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(a2_full);
      }
      // (2) folded queries, 128 output features at a time: this warp takes 64 of them -> one [128 x 64] box per chunk
      const CUtensorMap* mo = which ? &map_o1 : &map_o0;
#pragma unroll 1
      for (int i = 0; i < 8; ++i) {
        const int buf = i % 3, slot = i & 3;
        const uint32_t box = stg + (uint32_t)((ch * 4 + slot) * QS_ASLAB);
        mbar_wait(&d2_full[buf], (i / 3) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(tlane + 128u + (uint32_t)(buf * 128 + ch * 64), r0);
        tmem_ld32(tlane + 128u + (uint32_t)(buf * 128 + ch * 64 + 32), r1);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&d2_empty[buf]);
        if (i >= 4) {                         // the store that last used this staging box has read it
          if (leader) tma_store_wait_read3();
          bar_sync(1 + ch, 128);
        }
        const uint32_t ra = box + (uint32_t)(row * 128);
#pragma unroll
        for (int c = 0; c < 4; ++c)
          st_shared_v4(ra + (uint32_t)(((c ^ row) & 7) << 4), pack2(__uint_as_float(r0[8 * c]), __uint_as_float(r0[8 * c + 1])),
                       pack2(__uint_as_float(r0[8 * c + 2]), __uint_as_float(r0[8 * c + 3])), pack2(__uint_as_float(r0[8 * c + 4]), __uint_as_float(r0[8 * c + 5])),
                       pack2(__uint_as_float(r0[8 * c + 6]), __uint_as_float(r0[8 * c + 7])));
#pragma unroll
        for (int c = 0; c < 4; ++c)
          st_shared_v4(ra + (uint32_t)((((4 + c) ^ row) & 7) << 4), pack2(__uint_as_float(r1[8 * c]), __uint_as_float(r1[8 * c + 1])),
                       pack2(__uint_as_float(r1[8 * c + 2]), __uint_as_float(r1[8 * c + 3])), pack2(__uint_as_float(r1[8 * c + 4]), __uint_as_float(r1[8 * c + 5])),
                       pack2(__uint_as_float(r1[8 * c + 6]), __uint_as_float(r1[8 * c + 7])));
        fence_proxy_async_smem();
        bar_sync(1 + ch, 128);
        // qf[which][u][b][l][head][feature]: features (i & 3)*128 + ch*64 .. +64 of head 2*sub + (i >> 2)
        if (leader) tma_store_5d(mo, box, (i & 3) * 128 + ch * 64, 2 * sub + (i >> 2), 0, b0, u);
      }
    }
    if (MODE == QS_ATTN) {
      // accumulator columns [0,64) q_h, [64,128) k_h, [128,192) v_h of this row's token.  Keys / values go to shared memory as
      // fp32 rows (sample slots 16 bytes apart in bank space: the two samples of a warp read different banks); a thread then
      // attends from its token over the Lq tokens of its own sample.                                        modules.py:54-64
      const int sl = row >> 4;
      const uint32_t kv_row = (uint32_t)(row * QS_KV_ROW + sl * 16);
      const uint32_t kbase = stg + QS_OFF_K, vbase = stg + QS_OFF_V, pbase = stg + QS_OFF_P + (uint32_t)(row * 64);
      const float* bq = sbias;
      float qv[64];             // both warps of a lane quarter hold the token's scaled query: each scores half of the keys
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t r[32];
        tmem_ld32(tlane + (uint32_t)(hf * 32), r);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 b4 = *reinterpret_cast<const float4*>(bq + hf * 32 + 4 * c);
          qv[hf * 32 + 4 * c] = (__uint_as_float(r[4 * c]) + b4.x) * p.scale; qv[hf * 32 + 4 * c + 1] = (__uint_as_float(r[4 * c + 1]) + b4.y) * p.scale;
          qv[hf * 32 + 4 * c + 2] = (__uint_as_float(r[4 * c + 2]) + b4.z) * p.scale; qv[hf * 32 + 4 * c + 3] = (__uint_as_float(r[4 * c + 3]) + b4.w) * p.scale;
        }
      }
#pragma unroll 1
      for (int hf = 0; hf < 2; ++hf) {          // ch 0 stages k, ch 1 stages v
        uint32_t r[32];
        tmem_ld32(tlane + (uint32_t)(64 + ch * 64 + hf * 32), r);
        const float* bb = bq + (1 + ch) * 64 + hf * 32;
        const uint32_t dst = (ch ? vbase : kbase) + kv_row + (uint32_t)(hf * 128);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 b4 = *reinterpret_cast<const float4*>(bb + 4 * c);
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16 * c), "f"(__uint_as_float(r[4 * c]) + b4.x),
                       "f"(__uint_as_float(r[4 * c + 1]) + b4.y), "f"(__uint_as_float(r[4 * c + 2]) + b4.z), "f"(__uint_as_float(r[4 * c + 3]) + b4.w) : "memory");
        }
      }
      const uint32_t bits = mask_bits;
      if (warp == 2) QTRACE(11);
      bar_sync(1, 256);
      if (warp == 2) QTRACE(12);
      {
        // raw scores of keys ch*8 .. ch*8+7 (four independent partial sums per key)
        float sc[8];
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          const int j = ch * 8 + jj;
          sc[jj] = -INFINITY;
          if (j < p.Lq) {
            const uint32_t kr = kbase + (uint32_t)((sl * 16 + j) * QS_KV_ROW + sl * 16);
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const float4 k4 = lds128(kr + 16 * c);
              a0 = fmaf(qv[4 * c], k4.x, a0); a1 = fmaf(qv[4 * c + 1], k4.y, a1);
              a2 = fmaf(qv[4 * c + 2], k4.z, a2); a3 = fmaf(qv[4 * c + 3], k4.w, a3);
            }
            sc[jj] = ((bits >> j) & 1u) ? (a0 + a1) + (a2 + a3) : kMaskFill;
          }
        }
#pragma unroll
        for (int c = 0; c < 2; ++c)
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(pbase + (uint32_t)(ch * 32 + 16 * c)), "f"(sc[4 * c]), "f"(sc[4 * c + 1]),
                       "f"(sc[4 * c + 2]), "f"(sc[4 * c + 3]) : "memory");
      }
      if (warp == 2) QTRACE(13);
      bar_sync(1, 256);
      {
        // softmax over the sample's keys (both warps, redundantly), then ctx[token, ch*32 .. +32) = sum_j p_j v_j
        float pj[16];
        float mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const float4 t = lds128(pbase + 16 * c);
          pj[4 * c] = t.x; pj[4 * c + 1] = t.y; pj[4 * c + 2] = t.z; pj[4 * c + 3] = t.w;
          mx = fmaxf(fmaxf(mx, fmaxf(t.x, t.y)), fmaxf(t.z, t.w));
        }
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) { pj[j] = j < p.Lq ? __expf(pj[j] - mx) : 0.f; sum += pj[j]; }
        const float inv = __fdividef(1.f, sum);
        float o[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          if (j < p.Lq) {
            const uint32_t vr = vbase + (uint32_t)((sl * 16 + j) * QS_KV_ROW + sl * 16 + ch * 128);
            const float pw = pj[j] * inv;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const float4 v4 = lds128(vr + 16 * c);
              o[4 * c] = fmaf(pw, v4.x, o[4 * c]); o[4 * c + 1] = fmaf(pw, v4.y, o[4 * c + 1]);
              o[4 * c + 2] = fmaf(pw, v4.z, o[4 * c + 2]); o[4 * c + 3] = fmaf(pw, v4.w, o[4 * c + 3]);
            }
          }
        }
        const uint32_t ra = stg + QS_OFF_CTX + (uint32_t)(row * 128);
#pragma unroll
        for (int c = 0; c < 4; ++c)
          st_shared_v4(ra + (uint32_t)((((ch * 4 + c) ^ row) & 7) << 4), pack2(o[8 * c], o[8 * c + 1]), pack2(o[8 * c + 2], o[8 * c + 3]),
                       pack2(o[8 * c + 4], o[8 * c + 5]), pack2(o[8 * c + 6], o[8 * c + 7]));
      }
      if (warp == 2) QTRACE(14);
      fence_proxy_async_smem();
      bar_sync(1, 256);
      if (leader && ch == 0) tma_store_4d(&map_o0, stg + QS_OFF_CTX, sub * 64, 0, b0, u);   // ctx[(u, b, l), head*64 ..)
    }
    if (warp == 2) QTRACE(5);
    if (leader) tma_store_wait_all();
    if (warp == 2) QTRACE(6);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
  if (warp == 2) QTRACE(7);
}

// rank-n bf16 tensor map, 128-byte swizzle; dims / box innermost first, strides (bytes) for dims 1..rank-1
int make_map_nd(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides, const cuuint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return BIST_ERR_CUDA; }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("qside: cuTensorMapEncodeTiled (rank %d) failed (%d)", rank, (int)r); return BIST_ERR_CUDA; }
  return 0;
}
// rows [nd][B][Lq] x cols: the view (cols, Lq, B, nd) with a box of 64 columns x 16 token slots x 8 samples
int make_map_rows(CUtensorMap* m, const void* base, int64_t cols, int64_t Lq, int64_t B, int64_t nd) {
  const cuuint64_t dims[4] = {(cuuint64_t)cols, (cuuint64_t)Lq, (cuuint64_t)B, (cuuint64_t)nd};
  const cuuint64_t strides[3] = {(cuuint64_t)cols * 2, (cuuint64_t)Lq * cols * 2, (cuuint64_t)B * Lq * cols * 2};
  const cuuint32_t box[4] = {64, 16, 8, 1};
  return make_map_nd(m, base, 4, dims, strides, box);
}

template <int MODE>
int qside_launch(const CUtensorMap* maps, const QsArgs& p, int grid, const char* tag, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(qside_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, QS_SMEM));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = QS_SMEM; cfg.stream = st;
  cfg.attrs = at; cfg.numAttrs = 1;
  ProfScope ps(tag, st);
  BIST_CUDA(cudaLaunchKernelEx(&cfg, qside_kernel<MODE>, maps[0], maps[1], maps[2], maps[3], maps[4], maps[5], maps[6], maps[7], p));
  BIST_LAUNCHED();
  return 0;
}

int qs_common(QsArgs& p, int nd, int64_t B, int64_t Lq, int64_t d, int64_t dff, int64_t h, float eps) {
  BIST_CHECK_ARG(qside_supported(B, Lq, d, dff, h) && (nd == 1 || nd == 2), "qside: unsupported shape");
  p.nd = nd; p.B = (int)B; p.Lq = (int)Lq; p.d = (int)d; p.h = (int)h; p.dff = (int)dff; p.nks = (int)(d / 64);
  p.tpd = (int)((B + 7) / 8); p.eps = eps; p.scale = 1.f / sqrtf((float)(d / h));
  p.inv_d = 1.f / (float)d; p.inv_dm1 = 1.f / (float)(d - 1);
  static const bool trace = getenv("BIST_QSIDE_TRACE") != nullptr;
  p.trace = trace ? 1 : 0;
  return 0;
}

}  // namespace

static int g_qside_switch = -1;   // -1: BIST_QSIDE_UNFUSED decides; 0 / 1: bist_debug_qside (A/B tests)

bool qside_supported(int64_t B, int64_t Lq, int64_t d, int64_t dff, int64_t h) {
  static const bool off = getenv("BIST_QSIDE_UNFUSED") != nullptr;
  if (g_qside_switch == 0 || (g_qside_switch < 0 && off)) return false;
  if (h != 2 && h != 4 && h != 6 && h != 8) return false;
  if (d != 64 * h || Lq < 1 || Lq > 16 || B < 1 || dff < 256 || dff % 256 != 0) return false;
  if (2 * ((B + 7) / 8) * (h > dff / 256 ? h : dff / 256) * 4 >= (1LL << 30)) return false;
  return true;
}

// QKV projection + masked self-attention of the query tokens -> ctx [nd][B*Lq][d] bf16 (encoder.py:176 / :184).
// xn: bf16 [nd][B*Lq][d] normalised layer inputs; wqkv[u]: packed bf16 [3][d][d]; bqkv[u]: fp32 [3][d].
int qside_attn(const void* xn, const void* const* wqkv, const float* const* bqkv, const uint8_t* qmask, void* ctx, int nd, int64_t B,
               int64_t Lq, int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st) {
  QsArgs p = {};
  BIST_TRY(qs_common(p, nd, B, Lq, d, dff, h, eps));
  CUtensorMap maps[8];
  for (int u = 0; u < 2; ++u) {
    const int v = u < nd ? u : 0;
    BIST_CHECK_ARG(xn && wqkv[v] && bqkv[v] && qmask && ctx, "qside_attn: null pointer");
    p.bias[u] = bqkv[v];
    BIST_TRY(make_map_2d(&maps[u], wqkv[v], false, d, 3 * d, 64, 64, 128));
  }
  p.qmask = qmask;
  BIST_TRY(make_map_rows(&maps[4], ctx, d, Lq, B, nd));
  BIST_TRY(make_map_rows(&maps[7], xn, d, Lq, B, nd));
  maps[2] = maps[3] = maps[0]; maps[5] = maps[6] = maps[4];
  return qside_launch<QS_ATTN>(maps, p, nd * p.tpd * (int)h, "qside_attn", st);
}

// LN1 / LN2 + stage-1 / stage-2 query projections + key folds -> qf [2][nd][B*Lq][h][d] bf16, and the raw-x value term
// x . W_v2^T + b -> xv2 [nd][B*Lq][d] bf16.  xb: bf16 copy of x [nd][B*Lq][d] (the statistics are taken in fp32 from the bf16
// values); wq[u]: packed bf16 [3][d][d] (W_q1 | W_q2 | W_v2); bq[u]: fp32 [3][d]; wkt[u]: packed bf16 [2][d][d], row n = output
// feature of the fold, column = head*64 + k (pre-scaled by 1/sqrt(dk)).
int qside_fold(const void* xb, const float* const* ln1_a, const float* const* ln1_b, const float* const* ln2_a, const float* const* ln2_b,
               const void* const* wq, const float* const* bq, const void* const* wkt, void* qf, void* xv2, int nd, int64_t B, int64_t Lq,
               int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st) {
  QsArgs p = {};
  BIST_TRY(qs_common(p, nd, B, Lq, d, dff, h, eps));
  CUtensorMap maps[8];
  for (int u = 0; u < 2; ++u) {
    const int v = u < nd ? u : 0;
    BIST_CHECK_ARG(xb && ln1_a[v] && ln1_b[v] && ln2_a[v] && ln2_b[v] && wq[v] && bq[v] && wkt[v] && qf && xv2, "qside_fold: null pointer");
    p.ln_a[u][0] = ln1_a[v]; p.ln_b[u][0] = ln1_b[v]; p.ln_a[u][1] = ln2_a[v]; p.ln_b[u][1] = ln2_b[v];
    p.bias[u] = bq[v];
    BIST_TRY(make_map_2d(&maps[u], wq[v], false, d, 3 * d, 64, 128, 128));
    BIST_TRY(make_map_2d(&maps[2 + u], wkt[v], false, d, 2 * d, 64, 128, 128));
  }
  for (int w = 0; w < 2; ++w) {   // qf[w]: the view (d, h, Lq, B, nd); box 64 features x 1 head x 16 token slots x 8 samples
    const cuuint64_t dims[5] = {(cuuint64_t)d, (cuuint64_t)h, (cuuint64_t)Lq, (cuuint64_t)B, (cuuint64_t)nd};
    const cuuint64_t strides[4] = {(cuuint64_t)d * 2, (cuuint64_t)h * d * 2, (cuuint64_t)Lq * h * d * 2, (cuuint64_t)B * Lq * h * d * 2};
    const cuuint32_t box[5] = {64, 1, 16, 8, 1};
    BIST_TRY(make_map_nd(&maps[4 + w], (const char*)qf + (size_t)w * nd * B * Lq * h * d * 2, 5, dims, strides, box));
  }
  BIST_TRY(make_map_rows(&maps[6], xv2, d, Lq, B, nd));
  BIST_TRY(make_map_rows(&maps[7], xb, d, Lq, B, nd));
  p.nlong = nd * p.tpd * 2 * (int)(h / 2);
  return qside_launch<QS_FOLD>(maps, p, p.nlong + nd * p.tpd, "qside_fold", st);
}

// LN3 + first feed-forward layer: ffh [nd][B*Lq][dff] bf16 = relu(LN(ts) . W_1^T + b_1) (modules.py:113); tsb: bf16 copy of ts.
int qside_ffn1(const void* tsb, const float* const* ln_a, const float* const* ln_b, const void* const* w1, const float* const* b1, void* ffh,
               int nd, int64_t B, int64_t Lq, int64_t d, int64_t dff, int64_t h, float eps, cudaStream_t st) {
  QsArgs p = {};
  BIST_TRY(qs_common(p, nd, B, Lq, d, dff, h, eps));
  CUtensorMap maps[8];
  for (int u = 0; u < 2; ++u) {
    const int v = u < nd ? u : 0;
    BIST_CHECK_ARG(tsb && ln_a[v] && ln_b[v] && w1[v] && b1[v] && ffh, "qside_ffn1: null pointer");
    p.ln_a[u][0] = ln_a[v]; p.ln_b[u][0] = ln_b[v]; p.bias[u] = b1[v];
    BIST_TRY(make_map_2d(&maps[u], w1[v], false, d, dff, 64, 256, 128));
  }
  BIST_TRY(make_map_rows(&maps[4], ffh, dff, Lq, B, nd));
  BIST_TRY(make_map_rows(&maps[7], tsb, d, Lq, B, nd));
  maps[2] = maps[3] = maps[0]; maps[5] = maps[6] = maps[4];
  return qside_launch<QS_FFN1>(maps, p, nd * p.tpd * (int)(dff / 256), "qside_ffn1", st);
}

}  // namespace bist

// Test switch: 0 = the query-side chain as separate launches, 1 = the fused kernels where supported, -1 = environment default.
extern "C" int bist_debug_qside_trace(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, bist::g_qtrace, sizeof(bist::g_qtrace)) == cudaSuccess ? 0 : 1;
}
extern "C" int bist_debug_qside(int on) {
  bist::g_qside_switch = on;
  return 0;
}
