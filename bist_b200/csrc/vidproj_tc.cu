// VidEncoder8.forward, video branch — /root/reference/model/encoder.py:72-80, as ONE kernel (bf16 mode):
//   out = LayerNorm(relu(fts . W^T + bias))
// fts is read from HBM exactly once as fp32 (TMA, 128B-swizzled [128 rows x 32 k] slabs), converted to bf16
// inside shared memory by the worker warps (64B-swizzled K-major tile), multiplied on the tensor cores
// (tcgen05.mma kind::f16, M = 128, N = d <= 512 as one or two instructions, fp32 accumulation in all 512
// TMEM columns) against the bf16 weight slab streamed by TMA, and normalised in the epilogue straight out of
// TMEM: a 128 x d fp32 tile is exactly the full row that the reference's LayerNorm (modules.py:28-31,
// unbiased std, eps on the std) needs, so the un-normalised projection never goes to memory.
//
// Persistent CTAs (one per SM), 352 threads, two rings: the fp32 feature slabs live in their own ring whose slots are
// released as soon as they have been CONVERTED (not when the MMA that reads the bf16 copy retires), so the HBM stream
// runs up to NA slabs ahead of the tensor core; the bf16 A slab and the W slab share the MMA ring.
//   warp 0      A producer      fts slab (fp32, TMA) -> NA-stage ring (4 at d = 512, more for smaller d)
//   warp 1      MMA issuer      waits for the converted A slab and the W slab, issues the UMMAs, commits slots / accumulator
//   warps 2-9   workers         main loop: fp32 -> bf16 conversion of the A slab (swizzle-aware, conflict-free);
//                               epilogue: bias + ReLU + single-pass (Chan) row statistics, normalise, smem
//                               transpose, 16-byte coalesced stores.  Two warps per TMEM lane quarter.
//   warp 10     W producer      bf16 weight slab (TMA, L2-resident) -> 3-stage MMA ring
#include "tc_ptx.cuh"

namespace bist {
namespace {
using namespace tc;

constexpr int VP_BK = 32;                      // K elements per ring slot
constexpr int VP_STAGES = 6;                   // MMA ring (bf16 A slab + W slab): barrier slots; L.ms of them are used
constexpr int VP_MAX_A = 8;                    // fp32 A ring: at most this many slabs
constexpr int VP_AF = 128 * VP_BK * 4;         // fp32 A slab   16 KB (SWIZZLE_128B)
constexpr int VP_AB = 128 * VP_BK * 2;         // bf16 A slab    8 KB (SWIZZLE_64B)
constexpr int VP_STG_LD = 36;

#ifdef BIST_GEMM_TRACE
__device__ long long g_vtrace[4][64];   // CTA 0 only: [role][event]
#define VTRACE(role, idx) do { if (blockIdx.x == 0 && (idx) < 64) g_vtrace[role][idx] = clock64(); } while (0)
#else
#define VTRACE(role, idx) do {} while (0)
#endif

#ifndef VP_PAIR_STAGES
#define VP_PAIR_STAGES 5
#endif
__host__ __device__ inline int vp_pair_stages() { return VP_PAIR_STAGES; }
struct VpSmem {
  int wb_bytes, na, ms, a16_off, w_off, stg_off, red_off, bar_off, total;
};
__host__ __device__ inline VpSmem vp_smem(int d, int ncta, int ms_override = 0) {
  VpSmem s;
  s.wb_bytes = d * VP_BK * 2 / ncta;            // bf16 W slab (SWIZZLE_64B): d rows, or this CTA's half of each UMMA's N
  // a CTA pair hands slots back and forth across two SMs (multicast commit, remote arrive): its MMA ring is deeper
  s.ms = ms_override > 0 ? ms_override : (ncta == 2 ? vp_pair_stages() : 3);
  // The epilogue's transposition buffers (8 warps x [32][36] floats) live IN the bf16 A ring when it is large enough: when the
  // epilogue of a tile runs, every UMMA that read the ring has retired (acc_full) and the warps that would refill it are the ones
  // running the epilogue.  The 36 KB this saves are two more fp32 slabs in flight.
  const int stg_bytes = 8 * 32 * VP_STG_LD * 4;
  const bool alias = s.ms * VP_AB >= stg_bytes;
  const int fixed = s.ms * (VP_AB + s.wb_bytes) + (alias ? 0 : stg_bytes) + 2 * 128 * 3 * 4 + (2 * VP_MAX_A + 3 * VP_STAGES + 2) * 8 + 16 + 1024;
  int na = (227 * 1024 - fixed) / VP_AF;        // every byte left over goes to fp32 slabs in flight
  s.na = na > VP_MAX_A ? VP_MAX_A : na;
  s.a16_off = s.na * VP_AF;                     // bf16 A ring: ms slabs of 8 KB
  s.w_off = s.a16_off + s.ms * VP_AB;           // W ring: ms slabs
  const int end = s.w_off + s.ms * s.wb_bytes;
  s.stg_off = alias ? s.a16_off : end;
  s.red_off = alias ? end : end + stg_bytes;    // row statistics exchange: [2][128] x (n, mean, M2)
  s.bar_off = s.red_off + 2 * 128 * 3 * 4;
  s.total = s.bar_off + (2 * VP_MAX_A + 3 * VP_STAGES + 2) * 8 + 16 + 1024;
  return s;
}

// 32 accumulator columns -> registers without waiting (pair with vp_tmem_wait): the next chunk's load overlaps this chunk's math
__device__ __forceinline__ void vp_tmem_ld32_nw(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void vp_tmem_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// true in exactly one lane of a converged warp
__device__ __forceinline__ bool vp_elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

struct VpArgs {
  const float* bias;
  const float* a2;
  const float* b2;
  void* out;
  int N, D, d, out_bf16, tiles, nkb, ms, grp;
  float inv_dm1;   // 1 / (d - 1)
  int rpc;   // > 0 (single-CTA form): CTA c owns rows [c*rpc, (c+1)*rpc) — equal shares, the last tile of a share is partial
  float eps;
};

template <int NCTA>
__global__ void __launch_bounds__(352, 1) vidproj_kernel(const __grid_constant__ CUtensorMap map_a,
                                                         const __grid_constant__ CUtensorMap map_w,
                                                         const __grid_constant__ CUtensorMap map_a32, const VpArgs p) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const VpSmem L = vp_smem(p.d, NCTA, p.ms);
  uint64_t* afull = (uint64_t*)(smem + L.bar_off);  // fp32 A slab landed
  uint64_t* aempty = afull + VP_MAX_A;              // fp32 A slab converted (4 worker-warp arrivals: one conversion group)
  uint64_t* wfull = aempty + VP_MAX_A;              // bf16 W slab landed
  uint64_t* conv = wfull + VP_STAGES;               // bf16 A written (4 worker-warp arrivals per CTA, counted at the leader)
  uint64_t* empty = conv + VP_STAGES;               // UMMAs that read the MMA-ring slot have completed
  uint64_t* acc_full = empty + VP_STAGES;
  uint64_t* acc_empty = acc_full + 1;
  uint32_t* tmem_slot = (uint32_t*)(acc_empty + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.d;
  const int nhalf = d > 256 ? 2 : 1, nmma = d / nhalf;   // UMMA N <= 256
  // NCTA == 2: the two CTAs of a cluster work on a PAIR of row tiles with one UMMA of M = 256; each holds its own 128
  // rows of A and HALF of each W slab (the B operand of a cta_group::2 UMMA is split over the pair), so every SM pulls
  // half the weight bytes from L2 and the tensor core reads half the B bytes from each shared memory.
  const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
  const int unit0 = NCTA == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ustep = NCTA == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int units = NCTA == 2 ? (p.tiles + 1) >> 1 : p.tiles;
  const int wrows = nmma / NCTA;                         // W rows per UMMA held by this CTA
  // This CTA's tiles: tile i starts at row m_first + i * m_step; rows >= row_hi are not its own.
  //   round-robin form: whole 128-row tiles dealt out to the CTAs (pairs); 320 tiles on 148 SMs are three rounds for 2.16 of work
  //   equal-share form (rpc > 0, NCTA == 1): every CTA owns the same number of consecutive rows; the last tile of a share is
  //   partial and fetches only its own rows (32-row boxes), so the third round costs its weight stream, not a full tile
  int n_my, m_first, m_step, row_hi;
  if (NCTA == 1 && p.rpc > 0) {
    const long long lo = (long long)blockIdx.x * p.rpc, hi = lo + p.rpc < p.N ? lo + p.rpc : p.N;
    n_my = lo < hi ? (int)((hi - lo + 127) / 128) : 0; m_first = (int)lo; m_step = 128; row_hi = (int)hi;
  } else {
    n_my = unit0 < units ? (units - unit0 + ustep - 1) / ustep : 0;
    m_first = (NCTA * unit0 + rank) * 128; m_step = NCTA * ustep * 128; row_hi = p.N;
  }

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < VP_MAX_A; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 4); }
    for (int s = 0; s < VP_STAGES; ++s) { mbar_init(&wfull[s], 1); mbar_init(&conv[s], 4 * NCTA); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 8 * NCTA);
    fence_barrier_init();
  }
  if (warp == 1) {
    if (NCTA == 2) tmem_alloc2(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();   // the peer's barriers exist before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the tail of the previous kernel

  if (warp == 0) {
    if (lane == 0) {
      int sa = 0, pa = 0, it = 0;
      // (Measured and dropped: prefetching the next 8-64 slabs into L2 with cp.async.bulk.prefetch.tensor makes the kernel
      //  SLOWER, 141.6 -> 144.9 .. 173.4 us — the prefetch boxes go through the same TMA unit as the ring's loads.)
      for (int ti = 0; ti < n_my; ++ti) {
        const int m0 = m_first + ti * m_step;     // a tile past the end (odd tile count) loads zeros
        const int own = row_hi - m0;              // equal-share form: rows of this tile that belong to the CTA
        const int nb32 = (NCTA == 1 && p.rpc > 0 && own < 128) ? (own + 31) / 32 : 4;
        // two slabs (adjacent 128-byte segments of the same 128 feature rows) are requested back to back, so the second
        // finds the DRAM pages the first opened: 154.7 -> 144.9 us at config 2
        const int grp = p.grp > 0 ? p.grp : (L.na >= 4 ? 2 : 1);
        for (int kb = 0; kb < p.nkb; kb += grp) {
          const int n = p.nkb - kb < grp ? p.nkb - kb : grp;
          int s2 = sa, p2 = pa;
          for (int j = 0; j < n; ++j) {
            mbar_wait(&aempty[s2], p2 ^ 1);
            if (++s2 == L.na) { s2 = 0; p2 ^= 1; }
          }
          for (int j = 0; j < n; ++j, ++it) {
            if (nb32 == 4) {
              mbar_expect_tx(&afull[sa], VP_AF);
              tma_load_2d(smem + sa * VP_AF, &map_a, &afull[sa], (kb + j) * VP_BK, m0);
            } else {   // partial tile: only the CTA's own rows (the other rows of the slab keep stale, finite data nobody stores)
              mbar_expect_tx(&afull[sa], (uint32_t)(nb32 * (VP_AF / 4)));
              for (int q = 0; q < nb32; ++q) tma_load_2d(smem + sa * VP_AF + q * (VP_AF / 4), &map_a32, &afull[sa], (kb + j) * VP_BK, m0 + 32 * q);
            }
            VTRACE(0, it);
            if (++sa == L.na) { sa = 0; pa ^= 1; }
          }
        }
      }
    }
  } else if (warp == 10) {
    if (lane == 0) {
      int it = 0;
      for (int ti = 0; ti < n_my; ++ti) {
        for (int kb = 0; kb < p.nkb; ++kb, ++it) {
          const int s = it % L.ms;
          mbar_wait(&empty[s], ((it / L.ms) & 1) ^ 1);
          mbar_expect_tx(&wfull[s], L.wb_bytes);
          uint8_t* st = smem + L.w_off + s * L.wb_bytes;
          for (int hN = 0; hN < nhalf; ++hN)
            tma_load_2d(st + hN * wrows * VP_BK * 2, &map_w, &wfull[s], kb * VP_BK, hN * nmma + rank * wrows);
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {   // the leader CTA issues for the pair.  The whole warp stays converged and one elected lane issues: in a
                       // divergent single-lane region the compiler wraps every tcgen05 instruction in a uniform-register waterfall
                       // loop, and the issue stream (~820 clk per slab in the timeline) was slower than the UMMAs it issues (512)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nmma >> 3) << 17) | ((uint32_t)((128 * NCTA) >> 4) << 24);
      int it = 0, tc_ = 0;
      for (int ti = 0; ti < n_my; ++ti, ++tc_) {
        if (NCTA == 2) mbar_wait_cluster(acc_empty, (tc_ & 1) ^ 1);
        else mbar_wait(acc_empty, (tc_ & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < p.nkb; ++kb, ++it) {
          const int s = it % L.ms;
          mbar_wait(&wfull[s], (it / L.ms) & 1);
          if (NCTA == 2) mbar_wait_cluster(&conv[s], (it / L.ms) & 1);   // both A slabs converted, the peer's W half landed
          else mbar_wait(&conv[s], (it / L.ms) & 1);
          tc_fence_after();
          if (vp_elect_one()) {
            const uint32_t sa = smem_u32(smem + L.a16_off + s * VP_AB), sw = smem_u32(smem + L.w_off + s * L.wb_bytes);
            const uint64_t ad = make_desc_sw64(sa);
#pragma unroll
            for (int k = 0; k < VP_BK / 16; ++k) {
              for (int hN = 0; hN < nhalf; ++hN) {
                const uint64_t wd = make_desc_sw64(sw + hN * wrows * VP_BK * 2);
                if (NCTA == 2) tc_mma2(tmem_base + (uint32_t)(hN * nmma), ad + 2 * k, wd + 2 * k, idesc, (kb | k) ? 1u : 0u);
                else tc_mma<false>(tmem_base + (uint32_t)(hN * nmma), ad + 2 * k, wd + 2 * k, idesc, (kb | k) ? 1u : 0u);
              }
            }
            if (NCTA == 2) tc_commit2(&empty[s]);
            else tc_commit(&empty[s]);
            if (kb == p.nkb - 1) {
              if (NCTA == 2) tc_commit2(acc_full);
              else tc_commit(acc_full);
            }
            VTRACE(1, it);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ================================ worker warps =====================================================
    const int wt = threadIdx.x - 64;                 // 0..255
    const int crow = wt & 127, cgrp = wt >> 7;       // conversion: row of the slab; the group takes slabs it % 2 == cgrp
    const int lg = warp & 3, ch = (warp - 2) >> 2;   // epilogue: TMEM lane quarter, column half
    const int hcols = d >> 1;                        // columns per half (multiple of 32)
    const uint32_t stg = smem_u32(smem + L.stg_off) + (uint32_t)((warp - 2) * 32 * VP_STG_LD * 4);
    const int sub_r = lane >> 3, sub_c = (lane & 7) * 4;
    const uint32_t stg_w = stg + (uint32_t)(lane * VP_STG_LD * 4);
    const uint32_t stg_r = stg + (uint32_t)((sub_r * VP_STG_LD + sub_c) * 4);
    float* red = (float*)(smem + L.red_off);         // [2][128][3]
    int it = 0, tc_ = 0;
    for (int ti = 0; ti < n_my; ++ti, ++tc_) {
      // ---- main loop: fp32 slab -> bf16 slab.  The two groups of four warps take alternate slabs (a thread converts one row of
      //      its slab), so a slab's wait -> load -> convert -> store -> fence -> arrive chain has two slab periods to complete in
      //      instead of one: with all eight warps on every slab that chain, not the tensor core, set the pace.
      for (int kb = 0; kb < p.nkb; ++kb, ++it) {
        if ((it & 1) != cgrp) continue;
        const int s = it % L.ms, sa = it % L.na, pa = (it / L.na) & 1;
        mbar_wait(&afull[sa], pa);
        if (warp == 2 && lane == 0) VTRACE(2, it);
        const uint32_t af = smem_u32(smem + sa * VP_AF) + (uint32_t)(crow * 128);
        const uint32_t ab = smem_u32(smem + L.a16_off + s * VP_AB) + (uint32_t)(crow * 64);
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {   // logical 16-byte chunk j of a SWIZZLE_128B row
          float4 v;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "r"(af + (uint32_t)((j ^ (crow & 7)) << 4)) : "memory");
          __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
          pk[2 * j] = *reinterpret_cast<uint32_t*>(&lo);
          pk[2 * j + 1] = *reinterpret_cast<uint32_t*>(&hi);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&aempty[sa]);     // the fp32 slot is in registers: the next slab may land
        mbar_wait(&empty[s], ((it / L.ms) & 1) ^ 1);   // the UMMAs that read this bf16 slot have retired
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {  // logical 16-byte chunk jj of a SWIZZLE_64B row
          const uint32_t a = ab + (uint32_t)((jj ^ ((crow >> 1) & 3)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pk[4 * jj]), "r"(pk[4 * jj + 1]),
                       "r"(pk[4 * jj + 2]), "r"(pk[4 * jj + 3])
                       : "memory");
        }
        fence_proxy_async_smem();   // make the generic-proxy writes visible to the tensor core's async proxy
        __syncwarp();
        if (lane == 0) {
          if (rank == 0) {
            mbar_arrive(&conv[s]);
          } else {
            mbar_wait(&wfull[s], (it / L.ms) & 1);   // this CTA's half of the W slab has landed too
            mbar_arrive_cta(&conv[s], 0);
          }
        }
        if (warp == 2 && lane == 0) VTRACE(3, it);
      }
      // ---- epilogue -------------------------------------------------------------------------------------------
      const int m0 = m_first + ti * m_step;
      const int row_w0 = m0 + lg * 32;
      mbar_wait(acc_full, tc_ & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * hcols);
      // pass 1: row statistics of relu(acc + bias) over this warp's column half (Chan's parallel update per 32 columns)
      float cnt = 0.f, mean = 0.f, m2 = 0.f;
      uint32_t rn[32];                       // the next chunk's columns, in flight while this chunk is reduced
      vp_tmem_ld32_nw(tacc, rn);
#pragma unroll 1
      for (int c0 = 0; c0 < hcols; c0 += 32) {
        uint32_t r[32];
        vp_tmem_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = rn[j];
        if (c0 + 32 < hcols) vp_tmem_ld32_nw(tacc + c0 + 32, rn);
        const float* bp = p.bias + ch * hcols + c0;
        float sum = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(bp + 4 * q));
          float x0 = fmaxf(__uint_as_float(r[4 * q]) + b.x, 0.f), x1 = fmaxf(__uint_as_float(r[4 * q + 1]) + b.y, 0.f);
          float x2 = fmaxf(__uint_as_float(r[4 * q + 2]) + b.z, 0.f), x3 = fmaxf(__uint_as_float(r[4 * q + 3]) + b.w, 0.f);
          r[4 * q] = __float_as_uint(x0); r[4 * q + 1] = __float_as_uint(x1);
          r[4 * q + 2] = __float_as_uint(x2); r[4 * q + 3] = __float_as_uint(x3);
          sum += (x0 + x1) + (x2 + x3);
        }
        const float mc = sum * (1.f / 32.f);
        float qc = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) { const float t = __uint_as_float(r[j]) - mc; qc = fmaf(t, t, qc); }
        // (MUFU reciprocal: an IEEE division is a subroutine call per chunk; the ratio is 1 / (chunks so far))
        const float ncnt = cnt + 32.f, delta = mc - mean, w32 = __fdividef(32.f, ncnt);
        mean += delta * w32;
        m2 += qc + delta * delta * (cnt * w32);
        cnt = ncnt;
      }
      {
        float* mine = red + (ch * 128 + lg * 32 + lane) * 3;
        mine[0] = cnt; mine[1] = mean; mine[2] = m2;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float* oth = red + ((ch ^ 1) * 128 + lg * 32 + lane) * 3;
        const float on = oth[0], om = oth[1], oq = oth[2];
        const float ncnt = cnt + on, delta = om - mean, won = __fdividef(on, ncnt);
        mean += delta * won;
        m2 += oq + delta * delta * (cnt * won);
      }
      vp_tmem_ld32_nw(tacc, rn);             // pass 2's first chunk
      const float var = m2 * p.inv_dm1;
      const float inv = __fdividef(1.f, var * rsqrtf(fmaxf(var, 1e-30f)) + p.eps);   // modules.py:30-31: unbiased std, eps on the std
      // pass 2: normalise, transpose through smem, gains + coalesced stores
#pragma unroll 1
      for (int c0 = 0; c0 < hcols; c0 += 32) {
        uint32_t r[32];
        vp_tmem_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = rn[j];
        if (c0 + 32 < hcols) vp_tmem_ld32_nw(tacc + c0 + 32, rn);
        if (c0 + 32 >= hcols) {  // accumulator fully read by this warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (rank == 0) mbar_arrive(acc_empty);
            else mbar_arrive_cta(acc_empty, 0);
          }
        }
        const int col = ch * hcols + c0;
        const float* bp = p.bias + col;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(bp + 4 * q));
          const float x0 = (fmaxf(__uint_as_float(r[4 * q]) + b.x, 0.f) - mean) * inv;
          const float x1 = (fmaxf(__uint_as_float(r[4 * q + 1]) + b.y, 0.f) - mean) * inv;
          const float x2 = (fmaxf(__uint_as_float(r[4 * q + 2]) + b.z, 0.f) - mean) * inv;
          const float x3 = (fmaxf(__uint_as_float(r[4 * q + 3]) + b.w, 0.f) - mean) * inv;
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stg_w + 16 * q), "f"(x0), "f"(x1), "f"(x2), "f"(x3) : "memory");
        }
        __syncwarp();
        const int cc = col + sub_c;
        const float4 ga = __ldg(reinterpret_cast<const float4*>(p.a2 + cc)), gb = __ldg(reinterpret_cast<const float4*>(p.b2 + cc));
        float4 y[8];
#pragma unroll
        for (int rr = 0; rr < 8; ++rr)
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(y[rr].x), "=f"(y[rr].y), "=f"(y[rr].z), "=f"(y[rr].w)
                       : "r"(stg_r + (uint32_t)(rr * 4 * VP_STG_LD * 4))
                       : "memory");
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          const int row = row_w0 + rr * 4 + sub_r;
          if (row < row_hi) {
            const float4 o = make_float4(fmaf(y[rr].x, ga.x, gb.x), fmaf(y[rr].y, ga.y, gb.y), fmaf(y[rr].z, ga.z, gb.z),
                                         fmaf(y[rr].w, ga.w, gb.w));
            if (p.out_bf16) st4(reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)row * d + cc, o);
            else st4(reinterpret_cast<float*>(p.out) + (long long)row * d + cc, o);
          }
        }
        __syncwarp();
      }
      // the transposition buffers alias the bf16 A ring: no warp may start converting the next tile's slabs into it while another
      // warp is still in its epilogue
      if (L.stg_off == L.a16_off) asm volatile("bar.sync 1, 256;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();   // the peer may still be signalled by / read through this CTA until here
  if (warp == 1) {
    __syncwarp();
    if (NCTA == 2) tmem_dealloc2(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

#ifdef BIST_GEMM_TRACE
extern "C" int bist_debug_vp_trace(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_vtrace, sizeof(g_vtrace)) == cudaSuccess ? 0 : 1;
}
#endif

// Fused bf16-mode projection; *taken = false when the shape is left to the GEMM + LayerNorm pair (vidproj.cu).
// wb: bf16 copy of W [d, D] (workspace).
int vidproj_fused(const float* fts, const float* W, const float* bias, const float* a2, const float* b2, void* out, int64_t N,
                  int64_t D, int64_t d, float eps, void* wb, int out_dtype, cudaStream_t st, bool* taken) {
  *taken = false;
  if (d % 64 != 0 || d > 512 || D % VP_BK != 0 || N < 1 || N >= (1LL << 31) - 128) return 0;
  if ((reinterpret_cast<uintptr_t>(fts) & 15) || (reinterpret_cast<uintptr_t>(wb) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (reinterpret_cast<uintptr_t>(bias) & 15) || (reinterpret_cast<uintptr_t>(a2) & 15) || (reinterpret_cast<uintptr_t>(b2) & 15))
    return 0;
  static const int pair = getenv("BIST_VP_2CTA") ? atoi(getenv("BIST_VP_2CTA")) : 1;   // 0: one CTA per tile
  const int ncta = (pair && (d / (d > 256 ? 2 : 1)) % 32 == 0) ? 2 : 1;   // cta_group::2: UMMA N a multiple of 16 per CTA
  static const int ms_env = getenv("BIST_VP_STAGES") ? atoi(getenv("BIST_VP_STAGES")) : 0;   // MMA-ring depth (tuning aid)
  const VpSmem L = vp_smem((int)d, ncta, ms_env);
  if (L.total > 227 * 1024 || L.na < 2) return 0;
  *taken = true;
  BIST_TRY(pack_matrix(W, wb, (int)d, (int)D, 0, 1.f, BIST_BF16, st));
  const int nmma = (int)(d > 256 ? d / 2 : d);
  CUtensorMap ma, mw, ma32;
  BIST_TRY(make_map_2d(&ma, fts, true, D, N, VP_BK, 128, 128));
  BIST_TRY(make_map_2d(&ma32, fts, true, D, N, VP_BK, 32, 128));
  BIST_TRY(make_map_2d(&mw, wb, false, D, d, VP_BK, nmma / ncta, 64));
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(vidproj_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    BIST_CUDA(cudaFuncSetAttribute(vidproj_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  VpArgs p;
  p.bias = bias; p.a2 = a2; p.b2 = b2; p.out = out;
  p.N = (int)N; p.D = (int)D; p.d = (int)d; p.out_bf16 = out_dtype == BIST_BF16;
  p.tiles = (int)((N + 127) / 128); p.nkb = (int)(D / VP_BK); p.eps = eps;
  p.rpc = 0;
  p.inv_dm1 = 1.f / (float)(d - 1);
  p.ms = ms_env;
  static const int grp_env = getenv("BIST_VP_GRP") ? atoi(getenv("BIST_VP_GRP")) : 0;
  p.grp = grp_env;
  ProfScope ps("vidproj_fused", st);
  if (ncta == 2) {
    const int pairs = (p.tiles + 1) / 2, max_pairs = num_sms() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (pairs < max_pairs ? pairs : max_pairs)); cfg.blockDim = dim3(352);
    cfg.dynamicSmemBytes = (size_t)L.total; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = 2; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 2;
    BIST_CUDA(cudaLaunchKernelEx(&cfg, vidproj_kernel<2>, ma, mw, ma32, p));
  } else {
    int grid = p.tiles < num_sms() ? p.tiles : num_sms();
    if (p.tiles > grid && p.tiles % grid != 0) {   // uneven rounds: equal shares of rows instead (multiples of 32 for the partial boxes)
      p.rpc = (int)(((N + grid - 1) / grid + 31) / 32 * 32);
      grid = (int)((N + p.rpc - 1) / p.rpc);
    }
    BIST_LAUNCH(vidproj_kernel<1>, grid, 352, L.total, st, ma, mw, ma32, p);
  }
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist
