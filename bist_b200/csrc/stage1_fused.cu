// Stage 1 of temporal2spatial / spatial2temporal for BOTH directions as ONE kernel (bf16 mode) —
// /root/reference/model/encoder.py:109-121 (t2s: per region s, the query tokens attend over the T clips, temporal mask
// on the keys) and :143-148 (s2t: per clip t, over the S regions, no mask), attention() of modules.py:54-64.
//
// Per (sample b, group g, head hh) the reference computes  softmax_k(q_hh . K_k / sqrt(dk)) . V_k  with K, V linear
// projections of the group's video rows.  With the key projection folded into the query the video rows
// x_k enter through ONE product against [W_v,hh ; qf_hh]:
//     D[k, :] = x_k . [W_v,hh (64 rows) ; qf[b,:,hh,:] (Lq rows)]^T      (K = d)
// whose first 64 columns are the head's projected values and whose last Lq columns are its folded scores.  Neither the
// projected values [N, 2d] nor the scores [B*Lq*h, T*S] ever go to global memory.
//
// Layout of the machine: a cluster of h CTAs (one per head) works on one direction; all of them need the same video
// rows, so every [128 keys x 64 k] slab is fetched from L2/HBM ONCE per cluster and multicast by TMA into the h shared
// memories (each CTA issues 1/h of the pieces).  The head's W_v slice (64 x d bf16) stays resident in shared memory for
// the whole kernel, the folded queries of the current sample sit beside it (double buffered), so one UMMA per 16 k
// (M = 128 keys, N = 64 + 16) produces values and scores of a tile together, fp32 in tensor memory.
// A tile is a run of whole groups (R groups x Kp8 keys <= 128 lanes): 8 clips x 16 regions for s2t, 3 regions x 40 clips
// for t2s at config 2 — the t2s tile is gathered by the TMA unit itself from rows at stride S*d (box 1 x 8 of the
// (d, S, T, B) view), so the softmax over T completes inside the tile like the softmax over S does.
//
// Warp roles (384 threads):
//   warp 0      producer: W_v slice once, folded queries per sample, this CTA's share of the multicast video pieces
//   warp 1      UMMA issuer: 4 x d/64 main UMMAs per tile into one of two TMEM accumulators, then the P.V UMMAs of
//               the previous tile (A = staged values, MN-major; B = staged probabilities, K-major; N = 2 groups x 16)
//   warp 2      TMEM allocation
//   warps 4-7 / 8-11   two epilogue groups on alternating tiles: TMEM -> registers, values -> bf16 staging (UMMA
//               MN-major SW128 layout), scores -> smem, masked softmax over the keys of each (group, query token)
//               (modules.py:60-61), probabilities -> bf16 staging, wait for P.V, + b_v, bf16 stores of
//               c1[b, l, g, hh*64 .. +64) (the layout stage 2 reads: the G summaries of one (b,l) contiguous).
#include <stdlib.h>

#include "tc_ptx.cuh"

namespace bist {
namespace {
using namespace tc;

constexpr int S1_STAGES = 4;                 // video slab ring
constexpr int S1_XSLAB = 128 * 128;          // 128 keys x 64 k bf16
constexpr int S1_LQP = 16;                   // folded-query rows per head (Lq <= 16)
constexpr int S1_BROWS = 64 + 2 * S1_LQP;    // rows of the B operand per k slab: qf buffer 0 | W_v | qf buffer 1
constexpr int S1_BSLAB = S1_BROWS * 128;
constexpr int S1_MAXKS = 8;                  // d <= 512
constexpr int S1_XV_ROWS = 160;              // staged value rows per epilogue group (max of R * Kp16)
constexpr int S1_XV_BYTES = S1_XV_ROWS * 128;
constexpr int S1_SC_BYTES = 128 * S1_LQP * 4;
constexpr int S1_P_BYTES = S1_LQP * S1_XV_ROWS * 2;
constexpr int S1_MAXR = 8;                   // groups per tile (TMEM columns of the P.V accumulator: 8 x 16)
constexpr float kMaskFill = -1e9f;           // modules.py:60

constexpr int OFF_B = 0;
constexpr int OFF_X = OFF_B + S1_MAXKS * S1_BSLAB;          //  98304
constexpr int OFF_XV = OFF_X + S1_STAGES * S1_XSLAB;        // 163840
constexpr int OFF_P = OFF_XV + 2 * S1_XV_BYTES;             // 204800
constexpr int OFF_SC = OFF_P + 2 * S1_P_BYTES;              // 215040
constexpr int OFF_BAR = OFF_SC + 2 * S1_SC_BYTES;           // 231424
constexpr int S1_NBAR = 2 * S1_STAGES + 13;
constexpr int S1_SMEM = OFF_BAR + S1_NBAR * 8 + 16;

struct S1Dir {
  const float* bv;        // [d] value-projection bias
  __nv_bfloat16* c1;      // [B, Lq, G, d]
  int G, Kn;              // groups per sample, keys per group
  int Kp8, Kp16, R;       // keys padded to 8 (tile lanes) / 16 (staging, UMMA K); groups per tile
  int tiles_per_b, units; // ceil(G / R); B * tiles_per_b
  int use_mask;           // keys are clips: temporal_mask[b][k]
  int cl_begin, cl_end;   // clusters [cl_begin, cl_end) work on this direction
};
struct S1Args {
  S1Dir dir[2];
  const uint8_t* tmask;   // [B, T]
  float* dbg;             // optional [dir][unit][head][128][80] raw accumulator dump (tests)
  int B, T, S, Lq, d, h, nks;
};

__device__ __forceinline__ void tma_load_4d_mc(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3,
                                               uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
      ::"r"(dst), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
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
__device__ __forceinline__ float ld_shared_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// scores of tile lane `row`, query token l: 16-byte chunks of the 64-byte row XOR-swizzled by (row >> 1) & 3
__device__ __forceinline__ uint32_t sc_addr(uint32_t base, int row, int l) {
  return base + (uint32_t)(row * (S1_LQP * 4) + ((((l >> 2) ^ (row >> 1)) & 3) << 4) + ((l & 3) << 2));
}

struct Unit { int b, g0, rv; };
__device__ __forceinline__ Unit decode_unit(const S1Dir& D, int u) {
  Unit x;
  x.b = u / D.tiles_per_b;
  x.g0 = (u - x.b * D.tiles_per_b) * D.R;
  x.rv = min(D.R, D.G - x.g0);
  return x;
}

__global__ void __launch_bounds__(384, 1) stage1_fused_kernel(const __grid_constant__ CUtensorMap map_x0,
                                                              const __grid_constant__ CUtensorMap map_x1,
                                                              const __grid_constant__ CUtensorMap map_w0,
                                                              const __grid_constant__ CUtensorMap map_w1,
                                                              const __grid_constant__ CUtensorMap map_q0,
                                                              const __grid_constant__ CUtensorMap map_q1, const S1Args p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
  uint64_t* full = bars;                       // [STAGES] video slab landed (TMA bytes from all producers of the cluster)
  uint64_t* empty = full + S1_STAGES;          // [STAGES] slab consumed by the UMMAs of ALL CTAs (h multicast commits)
  uint64_t* acc_full = empty + S1_STAGES;      // [2]
  uint64_t* acc_empty = acc_full + 2;          // [2] 4 epilogue-warp arrivals
  uint64_t* stg_full = acc_empty + 2;          // [2] staged values + probabilities of a tile are in shared memory
  uint64_t* pv_full = stg_full + 2;            // [2]
  uint64_t* qf_full = pv_full + 2;             // [2]
  uint64_t* qf_empty = qf_full + 2;            // [2]
  uint64_t* w_full = qf_empty + 2;             // [1]
  uint32_t* tmem_slot = (uint32_t*)(w_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = p.h, nks = p.nks;
  const int hh = (int)cluster_ctarank();       // this CTA's head
  const int ci = (int)blockIdx.x / h;          // cluster index
  const int di = (p.dir[0].cl_end > ci) ? 0 : 1;
  const S1Dir& D = p.dir[di];
  const CUtensorMap* map_x = di == 0 ? &map_x0 : &map_x1;
  const CUtensorMap* map_w = di == 0 ? &map_w0 : &map_w1;
  const CUtensorMap* map_q = di == 0 ? &map_q0 : &map_q1;
  const int ncl = D.cl_end - D.cl_begin, lc = ci - D.cl_begin;
  const int u_begin = (int)((long long)D.units * lc / ncl), u_end = (int)((long long)D.units * (lc + 1) / ncl);
  const int n_units = u_end - u_begin;
  const uint16_t mc_mask = (uint16_t)((1u << h) - 1u);
  const int ppg = D.Kp8 >> 3;                  // 8-key pieces per group
  const int npieces = D.R * ppg;               // per slab

  if ((smem_u32(smem) & 1023u) != 0u) __trap();   // SWIZZLE_128B operands need 1024-byte aligned tiles
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < S1_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], (uint32_t)h); }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 4);
      mbar_init(&stg_full[a], 1); mbar_init(&pv_full[a], 1);
      mbar_init(&qf_full[a], 1); mbar_init(&qf_empty[a], 1);
    }
    mbar_init(w_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  // staged-value rows that no tile lane maps to (keys Kp8..Kp16 of a group) must be finite zeros: P is zero there
  for (int i = threadIdx.x; i < (2 * S1_XV_BYTES + 2 * S1_P_BYTES) / 16; i += blockDim.x)
    *reinterpret_cast<uint4*>(smem + OFF_XV + i * 16) = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // every CTA's barriers exist before a peer multicasts into it
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0 && n_units > 0) {
      // ---- producer -------------------------------------------------------------------------------------------------
      mbar_expect_tx(w_full, (uint32_t)(nks * 64 * 128));
      for (int j = 0; j < nks; ++j) tma_load_2d(smem + OFF_B + j * S1_BSLAB + S1_LQP * 128, map_w, w_full, j * 64, hh * 64);
      int it = 0, cur_b = -1, nsamp = 0;
      for (int ui = 0; ui < n_units; ++ui) {
        const Unit U = decode_unit(D, u_begin + ui);
        if (U.b != cur_b) {
          const int qb = nsamp & 1;
          mbar_wait(&qf_empty[qb], ((nsamp >> 1) & 1) ^ 1);
          mbar_expect_tx(&qf_full[qb], (uint32_t)(nks * S1_LQP * 128));
          for (int j = 0; j < nks; ++j)
            tma_load_3d(smem + OFF_B + j * S1_BSLAB + (qb ? (S1_LQP + 64) * 128 : 0), map_q, &qf_full[qb], j * 64, hh, U.b * p.Lq);
          cur_b = U.b; ++nsamp;
        }
        for (int ks = 0; ks < nks; ++ks, ++it) {
          const int s = it % S1_STAGES;
          mbar_wait_cluster(&empty[s], ((it / S1_STAGES) & 1) ^ 1);
          mbar_expect_tx(&full[s], (uint32_t)(npieces * 1024));
          const uint32_t dst = smem_u32(smem + OFF_X + s * S1_XSLAB);
          for (int pc = hh; pc < npieces; pc += h) {
            const int grp = pc / ppg, kk = (pc - grp * ppg) * 8;
            if (D.use_mask)   // keys are clips (t2s): (d, s = group, t = key, b)
              tma_load_4d_mc(dst + pc * 1024, map_x, &full[s], ks * 64, U.g0 + grp, kk, U.b, mc_mask);
            else                                                     // keys are regions (s2t): (d, s = key, t = group, b)
              tma_load_4d_mc(dst + pc * 1024, map_x, &full[s], ks * 64, kk, U.g0 + grp, U.b, mc_mask);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && n_units > 0) {
      // ---- UMMA issuer ----------------------------------------------------------------------------------------------
      const uint32_t idesc_main = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)((64 + S1_LQP) >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t idesc_pv = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(2 * S1_LQP >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t b_base = smem_u32(smem + OFF_B);
      const int ncore = D.Kp16 >> 3;                       // 8-key cores per probability row
      auto issue_pv = [&](int ui) {
        const int e = ui & 1;
        const Unit U = decode_unit(D, u_begin + ui);
        mbar_wait(&stg_full[e], (ui >> 1) & 1);
        tc_fence_after();
        const uint32_t xv = smem_u32(smem + OFF_XV + e * S1_XV_BYTES), pp = smem_u32(smem + OFF_P + e * S1_P_BYTES);
        for (int r = 0; r < U.rv; r += 2) {
          for (int kk = 0; kk < D.Kp16; kk += 16) {
            // A: values of groups r (feature atom 0) and r+1 (atom 1, LBO = one group of staged rows), 16 keys = two 8-row groups
            const uint64_t ad = desc_mn_sw128(xv + (uint32_t)((r * D.Kp16 + kk) * 128), (uint32_t)(D.Kp16 * 128) >> 4, 1024u >> 4);
            // B: probabilities, rows (group r, l = 0..15), (group r+1, l = 0..15): 8-row groups ncore*128 bytes apart
            const uint64_t bd = desc_k_none(pp + (uint32_t)(2 * r * ncore * 128 + (kk >> 3) * 128), 128u >> 4, (uint32_t)(ncore * 128) >> 4);
            tc_mma<false>(tmem_base + 256u + (uint32_t)(e * 128 + (r >> 1) * 32), ad, bd, idesc_pv, kk ? 1u : 0u);
          }
        }
        tc_commit(&pv_full[e]);
      };
      mbar_wait(w_full, 0);
      int it = 0, cur_b = -1, nsamp = 0, qb = 0;
      for (int ui = 0; ui < n_units; ++ui) {
        const Unit U = decode_unit(D, u_begin + ui);
        if (U.b != cur_b) {
          qb = nsamp & 1;
          mbar_wait(&qf_full[qb], (nsamp >> 1) & 1);
          cur_b = U.b; ++nsamp;
        }
        const int a = ui & 1;
        mbar_wait(&acc_empty[a], ((ui >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int ks = 0; ks < nks; ++ks, ++it) {
          const int s = it % S1_STAGES;
          mbar_wait(&full[s], (it / S1_STAGES) & 1);
          tc_fence_after();
          const uint64_t ad = make_desc(smem_u32(smem + OFF_X + s * S1_XSLAB));
          const uint64_t bd = make_desc(b_base + (uint32_t)(ks * S1_BSLAB + (qb ? S1_LQP * 128 : 0)));
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma<false>(tmem_base + (uint32_t)(a * 128), ad + 2 * k, bd + 2 * k, idesc_main, (ks | k) ? 1u : 0u);
          tc_commit_mc(&empty[s], mc_mask);   // the slot is free once EVERY CTA of the cluster has consumed it
        }
        tc_commit(&acc_full[a]);
        const bool last_of_sample = (ui + 1 == n_units) || (decode_unit(D, u_begin + ui + 1).b != U.b);
        if (last_of_sample) tc_commit(&qf_empty[qb]);
        if (ui >= 1) issue_pv(ui - 1);
      }
      issue_pv(n_units - 1);
    }
  } else if (warp >= 4) {
    // ---- epilogue groups ----------------------------------------------------------------------------------------------
    const int e = (warp - 4) >> 2;                 // group 0: warps 4-7, group 1: warps 8-11
    const int q = warp & 3;                        // TMEM lane quarter
    const int row = q * 32 + lane;                 // tile lane = key slot
    const int wt = (warp - 4 - 4 * e) * 32 + lane; // thread index inside the group
    const uint32_t xv = smem_u32(smem + OFF_XV + e * S1_XV_BYTES), pp = smem_u32(smem + OFF_P + e * S1_P_BYTES);
    const uint32_t sc = smem_u32(smem + OFF_SC + e * S1_SC_BYTES);
    const int grp = row / D.Kp8, kin = row - grp * D.Kp8;          // this lane's group / key within the tile
    const bool lane_used = grp < D.R;
    const uint32_t my_xv = xv + (uint32_t)((grp * D.Kp16 + kin) * 128);
    const int sw = (grp * D.Kp16 + kin) & 7;
    const int sgrp = wt >> 4, sl = wt & 15;                         // softmax role: (group, query token)
    const int ncore = D.Kp16 >> 3;
    const float bias_v = __ldg(D.bv + hh * 64 + (q & 1) * 32 + lane);
    int cur_b = -1, nsamp = 0, qb = 0;
    for (int ui = 0; ui < n_units; ++ui) {
      const Unit U = decode_unit(D, u_begin + ui);
      if (U.b != cur_b) { qb = nsamp & 1; cur_b = U.b; ++nsamp; }
      if ((ui & 1) != e) continue;
      mbar_wait(&acc_full[e], (ui >> 1) & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(e * 128);
      float* dbg = p.dbg ? p.dbg + ((((long long)di * p.dir[0].units + (u_begin + ui)) * h + hh) * 128 + row) * 80 : nullptr;
#pragma unroll
      for (int j = 0; j < 5; ++j) {   // five 16-column chunks: one of scores, four of projected values
        uint32_t r[16];
        tmem_ld16(tacc + (uint32_t)(16 * j), r);
        if (j == 4) {   // accumulator fully read by this warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[e]);
        }
        const bool is_sc = qb ? (j == 4) : (j == 0);
        if (is_sc) {
          if (lane_used) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              st_shared_v4(sc + (uint32_t)(row * 64 + (((c ^ (row >> 1)) & 3) << 4)), r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
          }
          if (dbg) for (int i = 0; i < 16; ++i) dbg[i] = __uint_as_float(r[i]);
        } else {
          const int vc = qb ? j : j - 1;   // 16-feature chunk of the head's 64 projected values
          if (lane_used) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              const int chunk = vc * 2 + c;
              st_shared_v4(my_xv + (uint32_t)(((chunk ^ sw) & 7) << 4),
                           pack2(__uint_as_float(r[8 * c]), __uint_as_float(r[8 * c + 1])),
                           pack2(__uint_as_float(r[8 * c + 2]), __uint_as_float(r[8 * c + 3])),
                           pack2(__uint_as_float(r[8 * c + 4]), __uint_as_float(r[8 * c + 5])),
                           pack2(__uint_as_float(r[8 * c + 6]), __uint_as_float(r[8 * c + 7])));
            }
          }
          if (dbg) for (int i = 0; i < 16; ++i) dbg[16 + vc * 16 + i] = __uint_as_float(r[i]);
        }
      }
      bar_sync(1 + e, 128);
      // ---- masked softmax over the keys of (group sgrp, query token sl) — modules.py:57-61 -------------------------
      if (sgrp < U.rv && sl < p.Lq) {
        const int row0 = sgrp * D.Kp8;
        const uint8_t* mk = D.use_mask ? p.tmask + (long long)U.b * p.T : nullptr;
        float mx = -INFINITY;
        for (int k = 0; k < D.Kn; ++k) {
          float v = ld_shared_f32(sc_addr(sc, row0 + k, sl));
          if (mk && mk[k] == 0) v = kMaskFill;
          mx = fmaxf(mx, v);
        }
        float sum = 0.f;
        for (int k = 0; k < D.Kn; ++k) {
          float v = ld_shared_f32(sc_addr(sc, row0 + k, sl));
          if (mk && mk[k] == 0) v = kMaskFill;
          sum += __expf(v - mx);
        }
        const float inv = 1.f / sum;
        const uint32_t prow = pp + (uint32_t)(((2 * sgrp + (sl >> 3)) * ncore) * 128 + (sl & 7) * 16);
        for (int c = 0; c < ncore; ++c) {
          float pv[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int k = c * 8 + j;
            float v = 0.f;
            if (k < D.Kn) {
              v = ld_shared_f32(sc_addr(sc, row0 + k, sl));
              if (mk && mk[k] == 0) v = kMaskFill;
              v = __expf(v - mx) * inv;
            }
            pv[j] = v;
          }
          st_shared_v4(prow + (uint32_t)(c * 128), pack2(pv[0], pv[1]), pack2(pv[2], pv[3]), pack2(pv[4], pv[5]), pack2(pv[6], pv[7]));
        }
      }
      fence_proxy_async_smem();   // generic-proxy writes of values and probabilities -> visible to the tensor core
      bar_sync(1 + e, 128);
      if (wt == 0) mbar_arrive(&stg_full[e]);
      // ---- P.V result: lanes 0-63 = features of the even group of a pair, lanes 64-127 = of the odd group ------------
      mbar_wait(&pv_full[e], (ui >> 1) & 1);
      tc_fence_after();
      const int feat = hh * 64 + (q & 1) * 32 + lane;
      for (int r0 = 0; r0 < U.rv; r0 += 2) {
        const int r = r0 + (q >> 1);
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + 256u + (uint32_t)(e * 128 + (r0 >> 1) * 32 + (q >> 1) * S1_LQP), v);
        if (r < U.rv) {
          __nv_bfloat16* o = D.c1 + (((long long)U.b * p.Lq) * D.G + (U.g0 + r)) * p.d + feat;
          const long long ol = (long long)D.G * p.d;
#pragma unroll
          for (int l = 0; l < S1_LQP; ++l)
            if (l < p.Lq) o[l * ol] = __float2bfloat16_rn(__uint_as_float(v[l]) + bias_v);
        }
      }
      tc_fence_before();
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // peers may still arrive on this CTA's barriers / multicast into it until here
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

int make_map_x(CUtensorMap* m, const void* vft, int64_t B, int64_t T, int64_t S, int64_t d, bool keys_are_clips) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return BIST_ERR_CUDA; }
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)S, (cuuint64_t)T, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)d * 2, (cuuint64_t)S * d * 2, (cuuint64_t)T * S * d * 2};
  cuuint32_t box[4] = {64, keys_are_clips ? 1u : 8u, keys_are_clips ? 8u : 1u, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(vft), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (4d video view) failed (%d)", (int)r); return BIST_ERR_CUDA; }
  return 0;
}

}  // namespace

bool stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h) {
  static const bool off = getenv("BIST_STAGE1_UNFUSED") != nullptr;
  if (off) return false;
  if (h != 1 && h != 2 && h != 4 && h != 8) return false;
  if (d != 64 * h || Lq < 1 || Lq > S1_LQP) return false;
  if (T < 1 || S < 1 || T > 128 || S > 128 || B < 1) return false;
  if (B * (T + S) >= (1LL << 30)) return false;
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
  p.B = (int)B; p.T = (int)T; p.S = (int)S; p.Lq = (int)Lq; p.d = (int)d; p.h = (int)h; p.nks = (int)(d / 64);
  CUtensorMap mx[2], mw[2], mq[2];
  long long tiles[2] = {0, 0};
  for (int u = 0; u < nd; ++u) {
    S1Dir& D = p.dir[u];
    const bool t2s = kinds[u] == 0;
    BIST_CHECK_ARG(qf[u] && wv[u] && bv[u] && c1[u], "stage1_fused: null pointer");
    BIST_CHECK_ARG((((uintptr_t)qf[u] | (uintptr_t)wv[u]) & 15) == 0, "stage1_fused: operands must be 16-byte aligned");
    BIST_CHECK_ARG(!t2s || tmask, "stage1_fused: t2s needs the temporal mask");
    D.bv = bv[u]; D.c1 = (__nv_bfloat16*)c1[u];
    D.G = (int)(t2s ? S : T); D.Kn = (int)(t2s ? T : S);
    D.Kp8 = (D.Kn + 7) & ~7; D.Kp16 = (D.Kn + 15) & ~15;
    D.R = 128 / D.Kp8; if (D.R > S1_MAXR) D.R = S1_MAXR;
    while (D.R > 1 && D.R * D.Kp16 > S1_XV_ROWS) --D.R;   // staged value rows of a tile fit one staging buffer
    D.tiles_per_b = (D.G + D.R - 1) / D.R; D.units = (int)(B * D.tiles_per_b);
    D.use_mask = t2s ? 1 : 0;
    tiles[u] = D.units;
    BIST_TRY(make_map_x(&mx[u], vft, B, T, S, d, t2s));
    BIST_TRY(make_map_2d(&mw[u], wv[u], false, d, d, 64, 64, 128));
    // folded queries [B*Lq, h, d]: (d, h, B*Lq) view, box 64 x 1 x 16
    BIST_TRY(make_map_3d(&mq[u], qf[u], false, d, h, B * Lq, d * 2, h * d * 2, 64, 1, S1_LQP));
  }
  if (nd == 1) { p.dir[1] = p.dir[0]; mx[1] = mx[0]; mw[1] = mw[0]; mq[1] = mq[0]; }

  static bool configured = false;
  static int max_clusters[9] = {0};
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(stage1_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S1_SMEM));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)h; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.blockDim = dim3(384); cfg.dynamicSmemBytes = S1_SMEM; cfg.stream = st; cfg.attrs = at; cfg.numAttrs = 2;
  if (max_clusters[h] == 0) {
    cfg.gridDim = dim3((unsigned)(num_sms() / h * h));
    int n = 0;
    cfg.numAttrs = 1;
    BIST_CUDA(cudaOccupancyMaxActiveClusters(&n, stage1_fused_kernel, &cfg));
    cfg.numAttrs = 2;
    max_clusters[h] = n > 0 ? n : 1;
  }
  long long total = tiles[0] + (nd == 2 ? tiles[1] : 0);
  int C = (int)(total < max_clusters[h] ? total : max_clusters[h]);
  if (nd == 2 && C < 2) C = 2;
  if (nd == 2) {
    int c0 = (int)((tiles[0] * C + total / 2) / total);
    c0 = c0 < 1 ? 1 : (c0 > C - 1 ? C - 1 : c0);
    p.dir[0].cl_begin = 0; p.dir[0].cl_end = c0;
    p.dir[1].cl_begin = c0; p.dir[1].cl_end = C;
  } else {
    p.dir[0].cl_begin = 0; p.dir[0].cl_end = C;
    p.dir[1].cl_begin = C; p.dir[1].cl_end = C;
  }
  cfg.gridDim = dim3((unsigned)(C * h));
  ProfScope ps("stage1_fused", st);
  BIST_CUDA(cudaLaunchKernelEx(&cfg, stage1_fused_kernel, mx[0], mx[1], mw[0], mw[1], mq[0], mq[1], p));
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
extern "C" int bist_stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h) {
  return bist::stage1_fused_supported(B, T, S, Lq, d, h) ? 1 : 0;
}
