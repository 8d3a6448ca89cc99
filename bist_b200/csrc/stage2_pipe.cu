// Stage 2 of t2s / s2t (encoder.py:125-132 / :152-161), bf16, persistent and TMA-pipelined.  For every (b,l) unit
//   S[hh, g] = qf[hh, :] . x1[g, :],  P = softmax_g(mask(S)) (o dropout),  U[hh, :] = sum_g P[hh, g] x1[g, :]
// One CTA per SM walks over the units; a producer warp keeps the next tiles in flight with TWO TMA requests per unit (the x1 tile
// [G, d] and the query rows [h, d], each as a 3-D box [d/64 slabs][rows][64 columns] so that they land as SWIZZLE_128B slabs), and
// two groups of eight consumer warps work on alternate units so that their short dependent mma.sync / shuffle chains interleave.
//   * one unit per CTA (attn_mma.cu stage2_mma_kernel) loads, computes and stores in sequence: 25-35 us at G = 40;
//   * one 1 KB bulk copy per row (48 requests per unit) was bound by the TMA request rate: 22 us.
// Both products run transposed -- keys (scores) and feature columns (P.X) are the M dimension of mma.m16n8k16, the 8 heads are N --
// so no half-empty 16-head tile is multiplied.
//   warps 0-7, 8-15  consumer groups      warp 16  producer
// smem: NS stages of { X: d/64 slabs of [G][128 B] | Q, later U[h][d]: 8 KB | key mask 256 B }, per group S (K-split planes) and P
#include <stdlib.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace bist {
namespace {
using namespace tc;

constexpr float kMaskFill = -1e9f;  // modules.py:60

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// SWIZZLE_128B: the 16-byte chunk index (address bits 4-6) is XORed with the 128-byte row index (bits 7-9); stage bases are 1 KB aligned
__device__ __forceinline__ uint32_t sw128(uint32_t a) { return a ^ ((a >> 3) & 0x70u); }
__device__ __forceinline__ void group_sync(int group) { asm volatile("bar.sync %0, 256;" ::"r"(group + 1) : "memory"); }   // barrier 0 = __syncthreads
__device__ __forceinline__ void bulk_store(void* gdst, uint32_t ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

struct S2Args {
  const uint8_t* mask;
  __nv_bfloat16* out;
  float* rsum;
  DropCfg drop;
  int units, G, Gp, Lq, h, d, NS, KS;
  uint32_t x_bytes, q_bytes, stage_bytes, sp_bytes, s_floats;
};

template <int NG>   // consumer groups
__global__ void __launch_bounds__(32 * (8 * NG + 1)) stage2_pipe_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_q,
                                                          const S2Args p) {
  extern __shared__ uint8_t smem_dyn[];
  pdl_trigger();
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  const int G = p.G, Gp = p.Gp, h = p.h, d = p.d, NS = p.NS, KS = p.KS;
  const int ldp = Gp + 8;
  const int lane = threadIdx.x & 31, warp_all = threadIdx.x >> 5;
  const int group = warp_all >> 3, warp = warp_all & 7;                   // group NG = the producer warp
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NS * p.stage_bytes + NG * p.sp_bytes);
  uint64_t* empty = full + NS;
  float* Ss = reinterpret_cast<float*>(smem + (size_t)NS * p.stage_bytes + (size_t)(group % NG) * p.sp_bytes);
  __nv_bfloat16* Ps = reinterpret_cast<__nv_bfloat16*>(Ss + p.s_floats);
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_q)) : "memory");
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 2); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();
  if (group == NG) {
    // ---- producer: the x1 tile and the query rows of a unit as two TMA requests, the key mask of the sample beside them
    int it = 0;
    for (int unit = blockIdx.x; unit < p.units; unit += gridDim.x, ++it) {
      const int s = it % NS;
      uint8_t mk[8];   // (G <= 256: eight mask bytes per lane, requested before the wait for a free stage)
      if (p.mask) {
        const int b = unit / p.Lq;
#pragma unroll
        for (int e = 0; e < 8; ++e) mk[e] = lane + 32 * e < G ? p.mask[(long long)b * G + lane + 32 * e] : (uint8_t)0;
      }
      mbar_wait(&empty[s], ((it / NS) & 1) ^ 1);
      uint8_t* st = smem + (size_t)s * p.stage_bytes;
      if (lane == 0) {
        mbar_expect_tx(&full[s], p.x_bytes + p.q_bytes);
        tma_load_3d(st, &map_x, &full[s], 0, unit * G, 0);
        tma_load_3d(st + p.x_bytes, &map_q, &full[s], 0, unit * h, 0);
      }
      if (p.mask) {
        uint8_t* sm = st + p.x_bytes + p.q_bytes;
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (lane + 32 * e < G) sm[lane + 32 * e] = mk[e];
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    }
    return;
  }
  // ---- consumers: group g takes the CTA's units g, g + NG, ..
  const uint32_t xpitch = (uint32_t)G * 128u, qpitch = (uint32_t)h * 128u;   // bytes between 64-column slabs
  for (int it = group, unit = blockIdx.x + group * gridDim.x; unit < p.units; unit += NG * gridDim.x, it += NG) {
    const int s = it % NS;
    uint8_t* st = smem + (size_t)s * p.stage_bytes;
    const uint32_t xs = smem_u32(st), qs = xs + p.x_bytes;
    const uint8_t* s_mask = st + p.x_bytes + p.q_bytes;
    mbar_wait(&full[s], (it / NS) & 1);
    // ---- scores, transposed: S^T[g, hh] = X[g, :] . Q[hh, :].  Work item = (16-key tile, K split); the KS partial sums of a
    // score land in KS planes of S and are added by the softmax.  Rows g >= G of the last tile read whatever follows the slab
    // (finite values of this stage): their scores are never used.
    {
      const int mtiles = Gp >> 4;
      const int dks = d / KS;                                   // columns per K split (multiple of 64)
      for (int item = warp; item < mtiles * KS; item += 8) {
        const int mt = item % mtiles, ks = item / mtiles;
        const uint32_t arow = (uint32_t)(mt * 16 + (lane & 15)) * 128u, brow = (uint32_t)(lane & 7) * 128u;
        float acc[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k0 = ks * dks; k0 < (ks + 1) * dks; k0 += 32) {   // 32 columns never straddle a slab
          const uint32_t slab = (uint32_t)k0 >> 6, kin = ((uint32_t)k0 & 63u) * 2u;
          const uint32_t xa = xs + slab * xpitch + arow + kin + (uint32_t)(lane >> 4) * 16u;
          const uint32_t qb = qs + slab * qpitch + brow + kin + (uint32_t)(lane >> 3) * 16u;
          uint32_t a0[4], a1[4], b0, b1, b2, b3;
          ldsm_x4(sw128(xa), a0[0], a0[1], a0[2], a0[3]);
          ldsm_x4(sw128(xa + 32u), a1[0], a1[1], a1[2], a1[3]);
          ldsm_x4(sw128(qb), b0, b1, b2, b3);
          mma16816(acc, a0, b0, b1);
          mma16816(acc1, a1, b2, b3);
        }
        float* Sp = Ss + (size_t)ks * 8 * Gp;
        const int g = mt * 16 + (lane >> 2), hh = (lane & 3) * 2;
        Sp[hh * Gp + g] = acc[0] + acc1[0]; Sp[(hh + 1) * Gp + g] = acc[1] + acc1[1];
        Sp[hh * Gp + g + 8] = acc[2] + acc1[2]; Sp[(hh + 1) * Gp + g + 8] = acc[3] + acc1[3];
      }
    }
    group_sync(group);
    // ---- masked softmax over g: warp = head; rows >= h and columns >= G of P are zero
    {
      const int hh = warp;
      if (hh < h) {
        float rs = 0.f;
        if (G <= 64) {   // both of a lane's scores stay in registers: one read of S, one write of P
          const int g0 = lane, g1 = lane + 32;
          float s0 = -INFINITY, s1 = -INFINITY;
          if (g0 < G) { s0 = 0.f; for (int ks = 0; ks < KS; ++ks) s0 += Ss[(ks * 8 + hh) * Gp + g0]; }
          if (g1 < G) { s1 = 0.f; for (int ks = 0; ks < KS; ++ks) s1 += Ss[(ks * 8 + hh) * Gp + g1]; }
          if (p.mask) {
            if (g0 < G && s_mask[g0] == 0) s0 = kMaskFill;
            if (g1 < G && s_mask[g1] == 0) s1 = kMaskFill;
          }
          const float mx = warp_max(fmaxf(s0, s1));
          const float e0 = g0 < G ? __expf(s0 - mx) : 0.f, e1 = g1 < G ? __expf(s1 - mx) : 0.f;
          const float inv = __fdividef(1.f, warp_sum(e0 + e1));
          const unsigned long long base = ((unsigned long long)unit * h + hh) * G;
          const float p0 = g0 < G ? e0 * inv * drop_mul(p.drop, base + g0) : 0.f;
          const float p1 = g1 < G ? e1 * inv * drop_mul(p.drop, base + g1) : 0.f;
          const __nv_bfloat16 pb0 = __float2bfloat16_rn(p0), pb1 = __float2bfloat16_rn(p1);
          if (g0 < Gp) Ps[hh * ldp + g0] = pb0;
          if (g1 < Gp) Ps[hh * ldp + g1] = pb1;
          rs = __bfloat162float(pb0) + __bfloat162float(pb1);
        } else {
          float mx = -INFINITY;
          for (int g = lane; g < G; g += 32) {
            float sc = 0.f;
            for (int ks = 0; ks < KS; ++ks) sc += Ss[(ks * 8 + hh) * Gp + g];
            if (p.mask && s_mask[g] == 0) sc = kMaskFill;
            Ss[hh * Gp + g] = sc;
            mx = fmaxf(mx, sc);
          }
          mx = warp_max(mx);
          float sum = 0.f;
          for (int g = lane; g < G; g += 32) {
            const float e = __expf(Ss[hh * Gp + g] - mx);
            Ss[hh * Gp + g] = e;
            sum += e;
          }
          sum = warp_sum(sum);
          const float inv = __fdividef(1.f, sum);
          for (int g = lane; g < Gp; g += 32) {
            float pv = 0.f;
            if (g < G) pv = Ss[hh * Gp + g] * inv * drop_mul(p.drop, ((unsigned long long)unit * h + hh) * G + g);
            const __nv_bfloat16 pb = __float2bfloat16_rn(pv);
            Ps[hh * ldp + g] = pb;
            rs += __bfloat162float(pb);
          }
        }
        if (p.rsum) {
          rs = warp_sum(rs);
          if (lane == 0) p.rsum[(long long)unit * h + hh] = rs;
        }
      } else {
        for (int g = lane; g < Gp; g += 32) Ps[hh * ldp + g] = __float2bfloat16_rn(0.f);
      }
    }
    group_sync(group);
    // ---- U^T[c, hh] = sum_g X[g, c] P[hh, g]: the feature columns are the M dimension (X through ldmatrix.trans), the heads N.
    // Warp w owns columns [w * d / 8, (w + 1) * d / 8).  Keys g >= G of the last 16-key step meet P = 0 (and finite X).  The result
    // is staged as U[hh][c] (dense rows) in the stage's query region, dead since the scores.
    {
      const int ncol = d >> 3;
      const uint32_t pb = smem_u32(Ps) + (uint32_t)(((lane & 7) * ldp + (lane >> 3) * 8) * 2);
      const int mi = lane >> 3;   // matrix mi of an A^T fragment: keys (mi / 2) * 8 .., columns (mi % 2) * 8 .. of the 16 x 16 block
      const uint32_t xrow = (uint32_t)((lane & 7) + (mi >> 1) * 8) * 128u;
      __nv_bfloat16* Us = reinterpret_cast<__nv_bfloat16*>(st + p.x_bytes);
      for (int m0 = 0; m0 < ncol; m0 += 32) {   // two 16-column blocks per pass
        const int c0 = warp * ncol + m0 + (mi & 1) * 8;
        const uint32_t xa0 = xs + ((uint32_t)c0 >> 6) * xpitch + xrow + ((uint32_t)c0 & 63u) * 2u;
        const uint32_t xa1 = xs + ((uint32_t)(c0 + 16) >> 6) * xpitch + xrow + ((uint32_t)(c0 + 16) & 63u) * 2u;
        float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k0 = 0; k0 < Gp; k0 += 16) {
          uint32_t a0[4], a1[4], b0, b1, b2, b3;
          ldsm_x4(pb + (uint32_t)(k0 * 2), b0, b1, b2, b3);                    // (b2, b3: the next 16 keys, unused)
          ldsm_x4_trans(sw128(xa0 + (uint32_t)k0 * 128u), a0[0], a0[1], a0[2], a0[3]);
          ldsm_x4_trans(sw128(xa1 + (uint32_t)k0 * 128u), a1[0], a1[1], a1[2], a1[3]);
          mma16816(acc0, a0, b0, b1);
          mma16816(acc1, a1, b0, b1);
        }
        const int c = warp * ncol + m0 + (lane >> 2), hh = (lane & 3) * 2;
        Us[(size_t)hh * d + c] = __float2bfloat16_rn(acc0[0]); Us[(size_t)(hh + 1) * d + c] = __float2bfloat16_rn(acc0[1]);
        Us[(size_t)hh * d + c + 8] = __float2bfloat16_rn(acc0[2]); Us[(size_t)(hh + 1) * d + c + 8] = __float2bfloat16_rn(acc0[3]);
        Us[(size_t)hh * d + c + 16] = __float2bfloat16_rn(acc1[0]); Us[(size_t)(hh + 1) * d + c + 16] = __float2bfloat16_rn(acc1[1]);
        Us[(size_t)hh * d + c + 24] = __float2bfloat16_rn(acc1[2]); Us[(size_t)(hh + 1) * d + c + 24] = __float2bfloat16_rn(acc1[3]);
      }
    }
    fence_proxy_async_smem();      // U was written through the generic proxy, the store below reads it through the async proxy
    group_sync(group);
    if (warp == 0 && lane == 0) {
      bulk_store(p.out + (long long)unit * h * d, qs, (uint32_t)h * (uint32_t)d * 2u);   // out[unit] = U[0..h): h * d contiguous bf16
      tma_store_wait_read<0>();    // the stage may be refilled once the store has read it
      mbar_arrive(&empty[s]);
    }
  }
}

}  // namespace

int stage2_pipe(const void* q, const void* x1, const uint8_t* mask, void* out, int B, int G, int Lq, int h, int d, cudaStream_t st,
                bool* taken, const Stage2Extra* ex) {
  *taken = false;
  static const bool off = getenv("BIST_STAGE2_UNPIPED") != nullptr;   // debugging aid: the one-unit-per-CTA kernel of attn_mma.cu
  if (off || h != 8 || d % 256 != 0 || d > 1024 || G < 1 || G > 256) return 0;
  const long long units = (long long)B * Lq;
  if (units * G >= (1ll << 31) || units * h >= (1ll << 31)) return 0;
  S2Args p;
  p.Gp = (G + 15) / 16 * 16;
  p.x_bytes = (uint32_t)G * (uint32_t)d * 2u;
  // (rows g in [G, Gp) of the last slab are read up to 15 * 128 bytes past the tile: into the query region, always loaded)
  p.q_bytes = (uint32_t)h * (uint32_t)d * 2u;
  if (p.x_bytes % 1024u != 0) p.x_bytes = (p.x_bytes + 1023u) & ~1023u;   // keeps the query region 1 KB aligned (G % 8 != 0)
  if ((uint32_t)G * 128u % 1024u != 0) return 0;                          // slab pitch must keep the swizzle phase: G % 8 == 0
  p.stage_bytes = (p.x_bytes + p.q_bytes + 256u + 1023u) & ~1023u;
  p.s_floats = 8u * (uint32_t)(p.Gp > 128 ? p.Gp : 128);
  p.sp_bytes = p.s_floats * 4u + 16u * (uint32_t)(p.Gp + 8) * 2u;
  static const int ng_env = getenv("BIST_STAGE2_GROUPS") ? atoi(getenv("BIST_STAGE2_GROUPS")) : 0;
  int ng = ng_env == 2 || ng_env == 3 ? ng_env : 3;
  size_t fixed = ng * (size_t)p.sp_bytes + 2 * 4 * 8 + 64 + 1024;
  int ns = (int)((225 * 1024 - fixed) / p.stage_bytes);
  if (ns < ng + 1 && ng == 3) {   // three groups want four stages (one in flight beside the three in use)
    ng = 2;
    fixed = ng * (size_t)p.sp_bytes + 2 * 4 * 8 + 64 + 1024;
    ns = (int)((225 * 1024 - fixed) / p.stage_bytes);
  }
  ns = ns > ng + 2 ? ng + 2 : ns;
  if (ns < 2) return 0;
  int ksplit = 1;                               // K splits of the score product: (Gp / 16) * ksplit <= 8 work items for 8 warps
  while (ksplit < 8 && (p.Gp / 16) * ksplit * 2 <= 8 && d / (ksplit * 2) % 64 == 0) ksplit *= 2;
  if ((reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(x1) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) return 0;
  CUtensorMap mx, mq;
  // [d/64 slabs][rows][64 columns]: the slab index is the OUTER box dimension with a 128-byte stride, so one request lands as
  // d/64 SWIZZLE_128B slabs of [rows][128 B]
  if (make_map_3d(&mx, x1, false, 64, units * G, d / 64, (long long)d * 2, 128, 64, G, d / 64) != 0) return 0;
  if (make_map_3d(&mq, q, false, 64, units * h, d / 64, (long long)d * 2, 128, 64, h, d / 64) != 0) return 0;
  *taken = true;
  p.mask = mask; p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.rsum = ex ? ex->rsum : nullptr; p.drop = ex ? ex->drop : DropCfg();
  p.units = (int)units; p.G = G; p.Lq = Lq; p.h = h; p.d = d; p.NS = ns; p.KS = ksplit;
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(stage2_pipe_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    BIST_CUDA(cudaFuncSetAttribute(stage2_pipe_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    configured = true;
  }
  const size_t smem = (size_t)ns * p.stage_bytes + fixed;
  const int grid = units < num_sms() ? (int)units : num_sms();
  ProfScope ps("stage2_mma", st);
  if (ng == 3) BIST_LAUNCH(stage2_pipe_kernel<3>, dim3((unsigned)grid), 800, smem, st, mx, mq, p);
  else BIST_LAUNCH(stage2_pipe_kernel<2>, dim3((unsigned)grid), 544, smem, st, mx, mq, p);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist
