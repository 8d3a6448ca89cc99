// Stage 1 of temporal2spatial / spatial2temporal for BOTH directions as ONE kernel (bf16 mode) —
// /root/reference/model/encoder.py:109-121 (t2s: per region s, the query tokens attend over the T clips, temporal mask
// on the keys) and :143-148 (s2t: per clip t, over the S regions, no mask), attention() of modules.py:54-64.
//
// Per (sample b, group g, head hh) the reference computes  softmax_k(q_hh . K_k / sqrt(dk)) . V_k  with K, V linear
// projections of the group's video rows.  With the key projection folded into the query the video rows
// x_k enter through ONE product against [qf_hh ; W_v,hh]:
//     D[k, :] = x_k . [qf[b,:,hh,:] (16 rows) ; W_v,hh (64 rows)]^T      (K = d)
// whose first 16 columns are the head's folded scores and whose last 64 columns are its projected values.  Neither the
// projected values [N, 2d] nor the scores [B*Lq*h, T*S] ever go to global memory.
//
// The kernel is shaped like a GEMM with an attention epilogue; its shape follows three measurements on B200
// (scripts/ubench.cu): an SM takes in at most ~59 bytes per clock through TMA whatever the ring depth, so every SM must
// see its video rows ONCE; tcgen05.mma with M = 128 runs at 52 clocks for N = 80 (full rate from N = 128); and mma.sync
// is starved (10 -> 180+ clocks per instruction) while a tcgen05 stream keeps the tensor pipe busy, so the small P.V
// products are tcgen05 instructions too.
//   * One CTA per SM owns a tile of 128 video rows of one sample — a run of whole groups: 8 clips x 16 regions for s2t,
//     3 regions x 40 clips for t2s at config 2; the t2s tile is gathered by the TMA unit itself from rows at stride S*d
//     (a box of the (d, S, T, B) view), so the softmax over T completes inside the tile like the softmax over S does.
//     The tile stays RESIDENT in shared memory (A operand, d/64 k-slabs of 16 KB) for all heads.
//   * [qf_h0 ; qf_h1 ; W_v,h0 ; W_v,h1] (160 rows: two heads per pass, so that the UMMA runs at N = 160 instead of the slower
//     N = 80) streams through a 5-slot ring of 10 KB half-slabs as the B operand: 4 UMMAs (M = 128, N = 160, K = 16) per k-slab
//     into one of two TMEM accumulators.  Accumulator columns: [0,32) scores (2*l + head), [32,96) / [96,160) the heads' values.
//   * Two epilogue groups of four warps take one head of the pass each: TMEM -> registers; values -> bf16 tile in the UMMA
//     MN-major SWIZZLE_128B layout, scores -> shared memory as fp32 [row][16] (XOR-swizzled float4s), key mask / padding
//     applied while staging; masked softmax over the keys of each (group, query token) (modules.py:57-61; a team of `parts`
//     threads per (group, 4 query tokens), float4 loads, warp-shuffle max / sum); probabilities -> bf16 [key][2 groups x 16 l]
//     in the MN-major SWIZZLE_64B layout; the P.V issuer warp then multiplies two groups per UMMA chain (A = staged values of
//     groups r, r+1 as the two 64-row atoms, B = their probabilities, N = 32: lanes 0-63 x columns 0-15 and lanes 64-127 x
//     columns 16-31 are the two results); the group reads them back, adds b_v and assembles the [l][group][64] output tile in
//     place of the dead staged values; ONE TMA store writes it into c1[b, l, g, hh*64 .. +64) — the layout stage 2 reads.
// Warps: 0 video-tile producer, 1 UMMA issuer, 2 TMEM allocation + B producer, 3 P.V issuer, 4-7 / 8-11 epilogue groups.
// Issue warps stay converged and elect one lane (elect.sync): a divergent single-lane region makes the compiler wrap every
// UMMA / TMA instruction in a uniform-register waterfall loop.
#include <stdlib.h>

#include "tc_ptx.cuh"

namespace bist {
__device__ long long g_s1trace[6][256];   // optional timeline of CTA 0 (bist_debug_s1_trace): [role][event] clock64 stamps
namespace {
using namespace tc;

constexpr int S1_MAXKS = 8;                  // d <= 512
constexpr int S1_XSLAB = 128 * 128;          // 128 video rows x 64 k bf16 (SWIZZLE_128B)
constexpr int S1_LQP = 16;                   // folded-query rows per head (Lq <= 16)
constexpr int S1_BROWS = 2 * (S1_LQP + 64);  // B rows of a pass (two heads): 32 qf rows (l-major, head-minor) | 2 x 64 W_v rows
constexpr int S1_BSLAB = S1_BROWS * 64;      // half-slab: 160 rows x 32 k bf16 (SWIZZLE_64B) = 10 KB
constexpr int S1_BSTAGES = 5;
constexpr int S1_XV_BYTES = 128 * 128;       // staged values of one head [row][64] bf16 (UMMA MN-major SWIZZLE_128B)
constexpr int S1_SC_BYTES = 128 * S1_LQP * 4;   // scores [row][l] fp32 (chunk-swizzled); the bf16 probabilities take their place
constexpr int S1_MAXR = 8;                   // groups per tile
constexpr float kMaskFill = -1e9f;           // modules.py:60

constexpr int OFF_X = 0;
constexpr int OFF_B = OFF_X + S1_MAXKS * S1_XSLAB;            // 131072
constexpr int OFF_XV = OFF_B + S1_BSTAGES * S1_BSLAB;         // 182272
constexpr int OFF_SC = OFF_XV + 2 * S1_XV_BYTES;              // 215040
constexpr int OFF_MB = OFF_SC + 2 * S1_SC_BYTES;              // 231424: temporal-mask bits, 2 x 16 bytes
constexpr int OFF_BAR = OFF_MB + 64;
constexpr int S1_NBAR = 2 * S1_MAXKS + 2 * S1_BSTAGES + 9;
constexpr int S1_SMEM = OFF_BAR + S1_NBAR * 8 + 16;
static_assert(S1_SMEM <= 227 * 1024, "stage1_fused: shared memory budget");
static_assert(OFF_B % 1024 == 0 && OFF_XV % 1024 == 0 && S1_BSLAB % 1024 == 0, "swizzled tiles must be 1024-byte aligned");
constexpr uint32_t TM_ACC = 0, TM_ACC_STRIDE = S1_BROWS, TM_PV = 2 * S1_BROWS;   // TMEM columns: two accumulators, one P.V window (128)

// float4 (query tokens 4*lq .. 4*lq+3) of tile row `row` in the score tile [row][16 l]: 64-byte rows, 16-byte chunks
// XOR-swizzled by (row >> 1) & 3 (conflict-free row-wise writes, at most 2-way conflicts on column-wise reads)
__device__ __forceinline__ int sc_f4(int row, int lq) { return row * 4 + (lq ^ ((row >> 1) & 3)); }

#define S1FINE(k) do { if (p.trace && blockIdx.x == 0 && e == 0 && wt == 0 && pc < 25) g_s1trace[5][pc * 10 + (k)] = clock64(); } while (0)
#define S1TRACE(role, idx) do { if (p.trace && blockIdx.x == 0 && (idx) < 256) g_s1trace[role][idx] = clock64(); } while (0)

struct S1Dir {
  const float* bv;        // [d] value-projection bias
  int G, Kn;              // groups per sample, keys per group
  int Kp8, Kp16, R;       // keys padded to 8 (tile rows) / 16 (UMMA K of P.V); groups per tile
  int tpb, units;         // tiles per sample; B * tpb
  int kclips;             // keys are clips (t2s): temporal_mask[b][k]; the tile is gathered region by region (one TMA box each)
  int parts, kpp;         // softmax team: `parts` threads per (group, 4 query tokens), kpp keys each (key = part + j * parts)
};
struct S1Args {
  S1Dir dir[2];
  const uint8_t* tmask;   // [B, T]
  float* dbg;             // optional [unit][head][128][80] raw accumulator dump (tests)
  int B, T, S, Lq, d, nh, nks, trace;
  int total_units;
};

// ---- PTX -------------------------------------------------------------------------------------------------------------------
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
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}
// sixteen 32-bit columns of 32 lanes; no wait (pair with tmem_ld_wait)
__device__ __forceinline__ void tmem_ld16_nw(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// MN-major SWIZZLE_128B: LBO = bytes between 64-element MN atoms, SBO = bytes between 8-row K groups (both >> 4)
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t addr, uint32_t lbo16, uint32_t sbo16) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)lbo16 << 16) | ((uint64_t)sbo16 << 32) | (1ull << 46) | (2ull << 61);
}
// K-major, no swizzle: core matrices of 8 rows x 16 bytes; LBO = bytes between the two K halves of one UMMA,
// SBO = bytes between 8-row groups (both >> 4)
__device__ __forceinline__ uint64_t desc_k_none(uint32_t addr, uint32_t lbo16, uint32_t sbo16) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)lbo16 << 16) | ((uint64_t)sbo16 << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void st_shared_v4(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
// true in exactly one lane of a converged warp (the compiler keeps issue instructions under it on the uniform datapath)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// unit (tile) -> direction, sample, first group, valid groups
struct Unit { int di, b, g0, rv; };
__device__ __forceinline__ Unit decode_unit(const S1Args& p, int u) {
  // sample-major order: all tiles of sample b (direction 0, then direction 1) are adjacent, so CTAs that run at the same time
  // read the same video rows and the second direction finds them in L2
  Unit x;
  const int per_b = p.dir[0].tpb + p.dir[1].tpb;
  x.b = u / per_b;
  int j = u - x.b * per_b;
  x.di = j >= p.dir[0].tpb ? 1 : 0;
  if (x.di) j -= p.dir[0].tpb;
  const S1Dir& D = p.dir[x.di];
  x.g0 = j * D.R;
  const int left = D.G - x.g0;
  x.rv = left < D.R ? left : D.R;
  return x;
}

// MN-major SWIZZLE_64B operand (32 bf16 of N per 64-byte row, K along the rows): SBO = bytes between 8-row K groups (>> 4)
__device__ __forceinline__ uint64_t desc_mn_sw64(uint32_t addr, uint32_t sbo16) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)sbo16 << 32) | (1ull << 46) | (4ull << 61);
}

__global__ void __launch_bounds__(384, 1) stage1_fused_kernel(const __grid_constant__ CUtensorMap map_x0,
                                                              const __grid_constant__ CUtensorMap map_x1,
                                                              const __grid_constant__ CUtensorMap map_w0,
                                                              const __grid_constant__ CUtensorMap map_w1,
                                                              const __grid_constant__ CUtensorMap map_q0,
                                                              const __grid_constant__ CUtensorMap map_q1,
                                                              const __grid_constant__ CUtensorMap map_c0,
                                                              const __grid_constant__ CUtensorMap map_c1, const S1Args p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
  uint64_t* xfull = bars;                       // [8]  k-slab of the video tile landed
  uint64_t* xempty = xfull + S1_MAXKS;          // [8]  the last head pass has consumed the k-slab
  uint64_t* bfull = xempty + S1_MAXKS;          // [5]  a [qf ; W_v] half-slab landed
  uint64_t* bempty = bfull + S1_BSTAGES;        // [5]
  uint64_t* acc_full = bempty + S1_BSTAGES;     // [2]
  uint64_t* acc_empty = acc_full + 2;           // [2]  8 epilogue-warp arrivals
  uint64_t* stg_full = acc_empty + 2;           // [2]  staged values + probabilities of a head are in shared memory
  uint64_t* pv_full = stg_full + 2;             // [2]
  uint64_t* pv_empty = pv_full + 2;             // [1]  the P.V window of TMEM has been read out (4 warp arrivals)
  uint32_t* tmem_slot = (uint32_t*)(pv_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int u_begin = (int)((long long)p.total_units * blockIdx.x / gridDim.x);
  const int u_end = (int)((long long)p.total_units * (blockIdx.x + 1) / gridDim.x);
  const int nks = p.nks, nh = p.nh, npass = p.nh >> 1;   // k-slabs per tile, heads, passes (two heads each)

  if ((smem_u32(smem) & 1023u) != 0u) __trap();   // SWIZZLE_128B operands need 1024-byte aligned tiles
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < S1_MAXKS; ++s) { mbar_init(&xfull[s], 1); mbar_init(&xempty[s], 1); }
    for (int s = 0; s < S1_BSTAGES; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 8);
      mbar_init(&stg_full[a], 1); mbar_init(&pv_full[a], 1);
    }
    mbar_init(pv_empty, 4);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  // staged values: rows nobody writes (past the tile's keys; the last group's K padding reads up to 8 of them) must be finite
  for (int i = threadIdx.x; i < (2 * S1_XV_BYTES + 2 * S1_SC_BYTES) / 16; i += blockDim.x)
    *reinterpret_cast<uint4*>(smem + OFF_XV + i * 16) = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ---- video-tile producer (whole warp converged; the TMA issue sits under elect.sync) ----------------------------------
    for (int u = u_begin, ui = 0; u < u_end; ++u, ++ui) {
      const Unit U = decode_unit(p, u);
      const S1Dir& D = p.dir[U.di];
      const CUtensorMap* mx = U.di ? &map_x1 : &map_x0;
      const uint32_t slab_bytes = (uint32_t)(D.R * D.Kp8 * 128);
      for (int ks = 0; ks < nks; ++ks) {
        mbar_wait(&xempty[ks], (ui & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&xfull[ks], slab_bytes);
          // one box = R groups x Kp8 keys x 64 k; out-of-range groups / keys are zero-filled
          // tile rows are group-major (group * Kp8 + key) in both directions: s2t groups (clips) are adjacent in memory, one box
          // R clips x Kp8 regions; t2s groups (regions) are gathered one box of Kp8 clips at stride S*d each
          if (D.kclips) {
            for (int g = 0; g < D.R; ++g)
              tma_load_4d(smem + OFF_X + ks * S1_XSLAB + g * D.Kp8 * 128, mx, &xfull[ks], ks * 64, U.g0 + g, 0, U.b);
          } else {
            tma_load_4d(smem + OFF_X + ks * S1_XSLAB, mx, &xfull[ks], ks * 64, 0, U.g0, U.b);
          }
          S1TRACE(0, ui * nks + ks);
        }
        __syncwarp();
      }
    }
  } else if (warp == 2) {
    // ---- [qf ; W_v] producer: two heads per pass, half-slabs of 32 k ----------------------------------------------------------
    int it = 0;
    for (int u = u_begin; u < u_end; ++u) {
      const Unit U = decode_unit(p, u);
      const CUtensorMap* mw = U.di ? &map_w1 : &map_w0;
      const CUtensorMap* mq = U.di ? &map_q1 : &map_q0;
      for (int pr = 0; pr < npass; ++pr) {
        for (int kh = 0; kh < 2 * nks; ++kh, ++it) {
          const int s = it % S1_BSTAGES;
          mbar_wait(&bempty[s], ((it / S1_BSTAGES) & 1) ^ 1);
          if (elect_one()) {
            S1TRACE(2, it);
            mbar_expect_tx(&bfull[s], S1_BSLAB);
            uint8_t* dst = smem + OFF_B + s * S1_BSLAB;
            tma_load_3d(dst, mq, &bfull[s], kh * 32, 2 * pr, U.b * p.Lq);              // 16 folded-query rows x 2 heads: row l*2 + head
            tma_load_2d(dst + 2 * S1_LQP * 64, mw, &bfull[s], kh * 32, pr * 128);      // the two heads' 2 x 64 value-weight rows
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    if (u_end > u_begin) {
      // ---- UMMA issuer (whole warp converged; issue under elect.sync) --------------------------------------------------------
      const uint32_t idesc_main = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(S1_BROWS >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      int it = 0, pc = 0;
      for (int u = u_begin, ui = 0; u < u_end; ++u, ++ui) {
        for (int pr = 0; pr < npass; ++pr, ++pc) {
          const int a = pc & 1;
          mbar_wait(&acc_empty[a], ((pc >> 1) & 1) ^ 1);
          tc_fence_after();
          for (int ks = 0; ks < nks; ++ks, it += 2) {
            const int s0 = it % S1_BSTAGES, s1 = (it + 1) % S1_BSTAGES;
            if (pr == 0) mbar_wait(&xfull[ks], ui & 1);
            mbar_wait(&bfull[s0], (it / S1_BSTAGES) & 1);
            mbar_wait(&bfull[s1], ((it + 1) / S1_BSTAGES) & 1);
            tc_fence_after();
            if (elect_one()) {
              S1TRACE(1, it >> 1);
              const uint64_t ad = make_desc(smem_u32(smem + OFF_X + ks * S1_XSLAB));
              const uint64_t b0 = make_desc_sw64(smem_u32(smem + OFF_B + s0 * S1_BSLAB)), b1 = make_desc_sw64(smem_u32(smem + OFF_B + s1 * S1_BSLAB));
              const uint32_t d = tmem_base + TM_ACC + (uint32_t)a * TM_ACC_STRIDE;
              tc_mma<false>(d, ad, b0, idesc_main, ks ? 1u : 0u);
              tc_mma<false>(d, ad + 2, b0 + 2, idesc_main, 1u);
              tc_commit(&bempty[s0]);
              tc_mma<false>(d, ad + 4, b1, idesc_main, 1u);
              tc_mma<false>(d, ad + 6, b1 + 2, idesc_main, 1u);
              tc_commit(&bempty[s1]);
              if (pr == npass - 1) tc_commit(&xempty[ks]);   // the video slab may be replaced by the next tile's
              if (ks == nks - 1) tc_commit(&acc_full[a]);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 3) {
    // ---- P.V issuer: as soon as an epilogue group has staged values and probabilities of a head, multiply two groups per UMMA
    //      (A = staged values of groups r, r+1 as the two 64-row atoms, MN-major; B = their probabilities [key][2 x 16 l], MN-major).
    //      The two heads of a pass share ONE window of tensor memory: each product waits until the previous one has been read out.
    const uint32_t idesc_pv = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(2 * S1_LQP >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    int pvc = 0, pc = 0;
    for (int u = u_begin; u < u_end; ++u) {
      const Unit U = decode_unit(p, u);
      const S1Dir& D = p.dir[U.di];
      for (int pr = 0; pr < npass; ++pr, ++pc) {
        for (int e = 0; e < 2; ++e, ++pvc) {
          mbar_wait(&stg_full[e], pc & 1);
          mbar_wait(pv_empty, (pvc & 1) ^ 1);
          tc_fence_after();
          const uint32_t xv = smem_u32(smem + OFF_XV + e * S1_XV_BYTES), pp = smem_u32(smem + OFF_SC + e * S1_SC_BYTES);
          if (elect_one()) {
            for (int r = 0; r < U.rv; r += 2) {
              for (int kk = 0; kk < D.Kp16; kk += 16) {
                // (an odd last group has no partner: its second atom aliases the first, lanes 64-127 are ignored)
                const uint64_t ad = desc_mn_sw128(xv + (uint32_t)((r * D.Kp8 + kk) * 128), r + 1 < U.rv ? (uint32_t)(D.Kp8 * 128) >> 4 : 0u, 1024u >> 4);
                const uint64_t bd = desc_mn_sw64(pp + (uint32_t)(((r >> 1) * D.Kp16 + kk) * 64), 512u >> 4);
                tc_mma<false>(tmem_base + TM_PV + (uint32_t)((r >> 1) * 32), ad, bd, idesc_pv, kk ? 1u : 0u);
              }
            }
            tc_commit(&pv_full[e]);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4) {
    // ---- epilogue groups: group e handles head 2*pass + e of every pass ----------------------------------------------------
    const int e = (warp - 4) >> 2;
    const int q = warp & 3;                        // TMEM lane quarter
    const int row = q * 32 + lane;                 // tile row = TMEM lane
    const int wt = (warp - 4 - 4 * e) * 32 + lane; // thread index inside the group
    uint8_t* xvp = smem + OFF_XV + e * S1_XV_BYTES;
    const uint32_t xv = smem_u32(xvp), pp = smem_u32(smem + OFF_SC + e * S1_SC_BYTES);
    float* scT = reinterpret_cast<float*>(smem + OFF_SC + e * S1_SC_BYTES);   // [l][row] swizzled; later the probabilities
    uint32_t* mbits = reinterpret_cast<uint32_t*>(smem + OFF_MB + e * 16);
    int pc = 0;
    bool store_pending = false;
    const bool store_leader = elect_one() && wt < 32;   // the one thread of the group that issues (and waits for) its TMA stores
    for (int u = u_begin; u < u_end; ++u) {
      const Unit U = decode_unit(p, u);
      const S1Dir& D = p.dir[U.di];
      const CUtensorMap* mc = U.di ? &map_c1 : &map_c0;
      const bool kclips = D.kclips != 0;
      // this row's group / key inside the tile (group-major rows); the values are staged in the same order
      const int grp = row / D.Kp8, kin = row - grp * D.Kp8;
      const bool row_used = row < D.R * D.Kp8;
      const int srow = grp * D.Kp8 + kin;
      const uint32_t my_xv = xv + (uint32_t)(srow * 128);
      // softmax role: `parts` threads per (group, 4 query tokens); this one takes keys part, part + parts, ...
      const int team = wt / D.parts, part = wt - team * D.parts;
      const int sgrp = team >> 2, slq = team & 3;
      for (int pr = 0; pr < npass; ++pr, ++pc) {
        const int a = pc & 1, hd = 2 * pr + e;
        if (wt == 0) S1TRACE(3 + e, (pc % 40) * 6 + 0);
        mbar_wait(&acc_full[a], (pc >> 1) & 1);
        if (wt == 0) S1TRACE(3 + e, (pc % 40) * 6 + 1);
        if (store_pending) { if (store_leader) tma_store_wait_read(); store_pending = false; }   // the previous output tile has left the staging buffer
        if (kclips && pr == 0) {   // temporal-mask bits of this sample (once per tile)
          const uint32_t bits = __ballot_sync(0xffffffffu, wt < p.T && p.tmask[(long long)U.b * p.T + wt] != 0);
          if (lane == 0) mbits[wt >> 5] = bits;
        }
        S1FINE(0);
        bar_sync(1 + e, 128);
        S1FINE(1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + TM_ACC + (uint32_t)a * TM_ACC_STRIDE;
        float* dbg = p.dbg ? p.dbg + (((long long)u * nh + hd) * 128 + row) * 80 : nullptr;
        {
          // accumulator columns: [0,32) scores (column 2*l + head of the pass), [32,96) values of the even head, [96,160) of the odd one
          uint32_t r[32];
          tmem_ld16_nw(tacc, r);
          tmem_ld16_nw(tacc + 16, r + 16);
          tmem_ld_wait();
          S1FINE(2);
          // keys past Kn (zero-filled padding rows) never take part; masked clips get exactly -1e9 (modules.py:60)
          const bool key_pad = kin >= D.Kn;
          const bool key_masked = kclips && kin < D.Kn && ((mbits[kin >> 5] >> (kin & 31)) & 1u) == 0u;
          float sv[S1_LQP];   // this key's scores against the 16 query tokens of head 2*pass + e
#pragma unroll
          for (int l = 0; l < S1_LQP; ++l) {
            const float x = __uint_as_float(e ? r[2 * l + 1] : r[2 * l]);
            sv[l] = key_pad ? -INFINITY : (key_masked ? kMaskFill : x);
          }
          {
            float4* s4 = reinterpret_cast<float4*>(scT);
#pragma unroll
            for (int c = 0; c < 4; ++c) s4[sc_f4(row, c)] = make_float4(sv[4 * c], sv[4 * c + 1], sv[4 * c + 2], sv[4 * c + 3]);
          }
          if (dbg) for (int i = 0; i < 16; ++i) dbg[i] = __uint_as_float(e ? r[2 * i + 1] : r[2 * i]);
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r[32];
          tmem_ld16_nw(tacc + (uint32_t)(32 + 64 * e + 32 * half), r);
          tmem_ld16_nw(tacc + (uint32_t)(48 + 64 * e + 32 * half), r + 16);
          tmem_ld_wait();
          if (half == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[a]);   // this warp has read its part of the accumulator
          }
          if (row_used) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              st_shared_v4(my_xv + (uint32_t)((((4 * half + c) ^ srow) & 7) << 4),
                           pack2(__uint_as_float(r[8 * c]), __uint_as_float(r[8 * c + 1])),
                           pack2(__uint_as_float(r[8 * c + 2]), __uint_as_float(r[8 * c + 3])),
                           pack2(__uint_as_float(r[8 * c + 4]), __uint_as_float(r[8 * c + 5])),
                           pack2(__uint_as_float(r[8 * c + 6]), __uint_as_float(r[8 * c + 7])));
          }
          if (dbg) for (int i = 0; i < 32; ++i) dbg[16 + 32 * half + i] = __uint_as_float(r[i]);
        }
        S1FINE(3);
        bar_sync(1 + e, 128);
        if (wt == 0) S1TRACE(3 + e, (pc % 40) * 6 + 2);
        // ---- masked softmax over the keys of group sgrp for query tokens 4*slq .. 4*slq+3 — modules.py:57-61 ------------------
        {
          const bool act = sgrp < U.rv;
          const float4* s4 = reinterpret_cast<const float4*>(scT);
          float4 x[8];
          float4 mx = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int k = part + j * D.parts;
            x[j] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            if (j < D.kpp && act && k < D.Kp8) x[j] = s4[sc_f4(sgrp * D.Kp8 + k, slq)];
            mx.x = fmaxf(mx.x, x[j].x); mx.y = fmaxf(mx.y, x[j].y); mx.z = fmaxf(mx.z, x[j].z); mx.w = fmaxf(mx.w, x[j].w);
          }
          for (int o = D.parts >> 1; o > 0; o >>= 1) {
            mx.x = fmaxf(mx.x, __shfl_xor_sync(0xffffffffu, mx.x, o)); mx.y = fmaxf(mx.y, __shfl_xor_sync(0xffffffffu, mx.y, o));
            mx.z = fmaxf(mx.z, __shfl_xor_sync(0xffffffffu, mx.z, o)); mx.w = fmaxf(mx.w, __shfl_xor_sync(0xffffffffu, mx.w, o));
          }
          float4 sm = make_float4(0.f, 0.f, 0.f, 0.f);
          if (act) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              x[j].x = __expf(x[j].x - mx.x); x[j].y = __expf(x[j].y - mx.y); x[j].z = __expf(x[j].z - mx.z); x[j].w = __expf(x[j].w - mx.w);
              sm.x += x[j].x; sm.y += x[j].y; sm.z += x[j].z; sm.w += x[j].w;
            }
          }
          for (int o = D.parts >> 1; o > 0; o >>= 1) {
            sm.x += __shfl_xor_sync(0xffffffffu, sm.x, o); sm.y += __shfl_xor_sync(0xffffffffu, sm.y, o);
            sm.z += __shfl_xor_sync(0xffffffffu, sm.z, o); sm.w += __shfl_xor_sync(0xffffffffu, sm.w, o);
          }
          S1FINE(4);
          bar_sync(1 + e, 128);   // all scores are in registers: the probabilities take their place
          S1FINE(5);
          if (act) {
            // (MUFU reciprocals: four IEEE divisions are four subroutine calls on the critical path of every pass)
            const float ix = __fdividef(1.f, sm.x), iy = __fdividef(1.f, sm.y), iz = __fdividef(1.f, sm.z), iw = __fdividef(1.f, sm.w);
            // probabilities of a pair of groups: [key][2 x 16 query tokens] bf16, 64-byte rows, 16-byte chunks XOR-swizzled by (key >> 1) & 3
            const uint32_t pbase = pp + (uint32_t)((sgrp >> 1) * D.Kp16 * 64);
            const int ch = (sgrp & 1) * 2 + (slq >> 1);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int k = part + j * D.parts;
              if (j < D.kpp && k < D.Kp16) {   // keys Kp8 .. Kp16 are written as zeros (exp(-inf))
                const uint32_t a = pbase + (uint32_t)(k * 64 + ((ch ^ ((k >> 1) & 3)) << 4) + (slq & 1) * 8);
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a), "r"(pack2(x[j].x * ix, x[j].y * iy)), "r"(pack2(x[j].z * iz, x[j].w * iw)) : "memory");
              }
            }
          }
        }
        S1FINE(6);
        fence_proxy_async_smem();   // generic-proxy writes of values and probabilities -> visible to the tensor core
        S1FINE(7);
        bar_sync(1 + e, 128);
        S1FINE(8);
        if (wt == 0) { mbar_arrive(&stg_full[e]); S1TRACE(3 + e, (pc % 40) * 6 + 3); }
        // ---- P.V results: lanes 0-63 = features of the even group of a pair, lanes 64-127 = of the odd group.  The staged
        //      values are dead now: the [l][group][64 features] output tile is assembled in their place ---------------------
        mbar_wait(&pv_full[e], pc & 1);
        if (wt == 0) S1TRACE(3 + e, (pc % 40) * 6 + 4);
        tc_fence_after();
        {
          const float bias_v = __ldg(D.bv + hd * 64 + (q & 1) * 32 + lane);
          __nv_bfloat16* ot = reinterpret_cast<__nv_bfloat16*>(xvp) + (q & 1) * 32 + lane;
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {   // two pairs of groups at a time
            uint32_t v[2][16];
#pragma unroll
            for (int j = 0; j < 2; ++j)
              if (2 * (2 * jj + j) < U.rv) tmem_ld16_nw(tmem_base + ((uint32_t)(q * 32) << 16) + TM_PV + (uint32_t)((2 * jj + j) * 32 + (q >> 1) * S1_LQP), v[j]);
            tmem_ld_wait();
            if (jj == 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(pv_empty);   // the P.V window may be overwritten by the other head's product
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const int r = 2 * (2 * jj + j) + (q >> 1);
              if (r < U.rv) {
#pragma unroll
                for (int l = 0; l < S1_LQP; ++l) ot[(l * D.R + r) * 64] = __float2bfloat16_rn(__uint_as_float(v[j][l]) + bias_v);
              }
            }
          }
        }
        fence_proxy_async_smem();
        bar_sync(1 + e, 128);
        if (store_leader) {
          tma_store_4d(mc, xv, hd * 64, U.g0, 0, U.b);   // rows beyond G / Lq are clipped by the tensor map
          S1TRACE(3 + e, (pc % 40) * 6 + 5);
        }
        store_pending = true;
      }
    }
    if (store_leader) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

int make_map_x(CUtensorMap* m, const void* vft, int64_t B, int64_t T, int64_t S, int64_t d, bool keys_are_clips, int Kp8, int R) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return BIST_ERR_CUDA; }
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)S, (cuuint64_t)T, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)d * 2, (cuuint64_t)S * d * 2, (cuuint64_t)T * S * d * 2};
  // s2t (keys = regions): one box = R clips x Kp8 regions = the whole tile slab; t2s (keys = clips): one box = Kp8 clips of ONE
  // region (rows at stride S*d), R boxes per slab; out-of-range regions / clips are zero-filled
  cuuint32_t box[4] = {64, (cuuint32_t)(keys_are_clips ? 1 : Kp8), (cuuint32_t)(keys_are_clips ? Kp8 : R), 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(vft), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (4d video view) failed (%d)", (int)r); return BIST_ERR_CUDA; }
  return 0;
}

// c1 [B, Lq, G, d] bf16 as (d, G, Lq, B); box = 64 features x R groups x 16 query tokens
int make_map_c(CUtensorMap* m, void* c1, int64_t B, int64_t Lq, int64_t G, int64_t d, int R) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return BIST_ERR_CUDA; }
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)G, (cuuint64_t)Lq, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)d * 2, (cuuint64_t)G * d * 2, (cuuint64_t)Lq * G * d * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)R, (cuuint32_t)S1_LQP, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, c1, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (c1 view) failed (%d)", (int)r); return BIST_ERR_CUDA; }
  return 0;
}

}  // namespace

bool stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h) {
  static const bool off = getenv("BIST_STAGE1_UNFUSED") != nullptr;
  if (off) return false;
  if (h != 2 && h != 4 && h != 6 && h != 8) return false;   // two heads per pass
  if (d != 64 * h || Lq < 1 || Lq > S1_LQP) return false;
  if (T < 1 || S < 1 || T > 128 || S > 128 || B < 1) return false;
  if (B * (T + S) >= (1LL << 29)) return false;
  return true;
}

// dirs: up to two directions; kind 0 = t2s (groups = regions, keys = clips, temporal mask), 1 = s2t.
int stage1_fused(const void* vft, const uint8_t* tmask, int nd, const int* kinds, const void* const* qf, const void* const* wv,
                 const float* const* bv, void* const* c1, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h,
                 float* dbg, cudaStream_t st) {
  BIST_CHECK_ARG(stage1_fused_supported(B, T, S, Lq, d, h), "stage1_fused: unsupported shape");
  BIST_CHECK_ARG(nd == 1 || nd == 2, "stage1_fused: nd must be 1 or 2");
  BIST_CHECK_ARG(((uintptr_t)vft & 15) == 0, "stage1_fused: vft must be 16-byte aligned");
  S1Args p = {};
  p.tmask = tmask; p.dbg = dbg;
  static const bool trace = getenv("BIST_S1_TRACE") != nullptr;
  p.trace = trace ? 1 : 0;
  p.B = (int)B; p.T = (int)T; p.S = (int)S; p.Lq = (int)Lq; p.d = (int)d; p.nh = (int)h; p.nks = (int)(d / 64);
  CUtensorMap mx[2], mw[2], mq[2], mc[2];
  for (int u = 0; u < nd; ++u) {
    S1Dir& D = p.dir[u];
    const bool t2s = kinds[u] == 0;
    BIST_CHECK_ARG(qf[u] && wv[u] && bv[u] && c1[u], "stage1_fused: null pointer");
    BIST_CHECK_ARG((((uintptr_t)qf[u] | (uintptr_t)wv[u] | (uintptr_t)c1[u]) & 15) == 0, "stage1_fused: operands must be 16-byte aligned");
    BIST_CHECK_ARG(((uintptr_t)bv[u] & 7) == 0, "stage1_fused: bias must be 8-byte aligned");
    BIST_CHECK_ARG(!t2s || tmask, "stage1_fused: t2s needs the temporal mask");
    D.bv = bv[u];
    D.G = (int)(t2s ? S : T); D.Kn = (int)(t2s ? T : S);
    D.Kp8 = (D.Kn + 7) & ~7; D.Kp16 = (D.Kn + 15) & ~15;
    D.R = 128 / D.Kp8; if (D.R > S1_MAXR) D.R = S1_MAXR;
    while (D.R > 1 && D.R * D.Kp16 > 160) --D.R;   // probability rows of a tile fit their buffer
    D.tpb = (D.G + D.R - 1) / D.R; D.units = (int)(B * D.tpb);
    D.kclips = t2s ? 1 : 0;
    D.parts = 1;
    while (D.parts < 32 && D.R * 4 * D.parts * 2 <= 128) D.parts *= 2;
    D.kpp = (D.Kp16 + D.parts - 1) / D.parts;
    BIST_CHECK_ARG(D.kpp <= 8, "stage1_fused: softmax key chunk too long");
    BIST_TRY(make_map_x(&mx[u], vft, B, T, S, d, t2s, D.Kp8, D.R));
    BIST_TRY(make_map_c(&mc[u], c1[u], B, Lq, D.G, d, D.R));
    BIST_TRY(make_map_2d(&mw[u], wv[u], false, d, d, 32, 128, 64));   // half-slab: 32 k x the two heads' 128 value-weight rows
    // folded queries [B*Lq, h, d]: (d, h, B*Lq) view, box 32 k x 2 heads x 16 query tokens (rows l*2 + head in shared memory)
    BIST_TRY(make_map_3d(&mq[u], qf[u], false, d, h, B * Lq, d * 2, h * d * 2, 32, 2, S1_LQP, 64));
  }
  if (nd == 1) { p.dir[1] = p.dir[0]; p.dir[1].units = 0; p.dir[1].tpb = 0; mx[1] = mx[0]; mw[1] = mw[0]; mq[1] = mq[0]; mc[1] = mc[0]; }
  p.total_units = p.dir[0].units + p.dir[1].units;

  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(stage1_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S1_SMEM));
    configured = true;
  }
  int ctas = num_sms();
  if (ctas > p.total_units) ctas = p.total_units;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(384); cfg.dynamicSmemBytes = S1_SMEM; cfg.stream = st;
  cfg.attrs = at; cfg.numAttrs = 1;
  if (trace) fprintf(stderr, "[stage1_fused] %d CTAs, tiles %d + %d\n", ctas, p.dir[0].units, p.dir[1].units);
  ProfScope ps("stage1_fused", st);
  BIST_CUDA(cudaLaunchKernelEx(&cfg, stage1_fused_kernel, mx[0], mx[1], mw[0], mw[1], mq[0], mq[1], mc[0], mc[1], p));
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist

// Test / bring-up entry: stage 1 alone.  qf [B*Lq, h, d] bf16 (folded, pre-scaled queries), wv [d, d] bf16 (rows = output
// features), bv fp32 [d], c1 out bf16 [B, Lq, G, d]; kind 0 = t2s, 1 = s2t.
extern "C" int bist_stage1_fused(const void* vft, const uint8_t* temporal_mask, int nd, const int* kinds, const void* const* qf,
                                 const void* const* wv, const float* const* bv, void* const* c1, int64_t B, int64_t T, int64_t S,
                                 int64_t Lq, int64_t d, int64_t h, float* dbg, void* stream) {
  return bist::stage1_fused(vft, temporal_mask, nd, kinds, qf, wv, bv, c1, B, T, S, Lq, d, h, dbg, (cudaStream_t)stream);
}
extern "C" int bist_debug_s1_trace(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, bist::g_s1trace, sizeof(bist::g_s1trace)) == cudaSuccess ? 0 : 1;
}
extern "C" int bist_stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h) {
  return bist::stage1_fused_supported(B, T, S, Lq, d, h) ? 1 : 0;
}
