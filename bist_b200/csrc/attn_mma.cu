// Tensor-core attention cores of the bf16 path.
//   softmax_pv_mma : masked softmax over precomputed fp32 scores + P.V with bf16 mma.sync (m16n8k16),
//                    one CTA per attention group, one warp per head.  V rows are staged once in shared
//                    memory with cp.async (16-byte coalesced), P never leaves registers, the context
//                    tile is staged in shared memory and written with 16-byte coalesced stores.
//                    (attention(), /root/reference/model/modules.py:59-64, after the score GEMM)
//   stage2_v2      : stage 2 of temporal2spatial / spatial2temporal with key and value projections folded to
//                    the query side (encoder.py:125-132, 152-161): the G stage-1 summaries of one (b,l) are
//                    read from HBM/L2 exactly once into shared memory.
#include <algorithm>
#include <stdlib.h>

#include "common.cuh"

namespace bist {
namespace {

constexpr float kMaskFill = -1e9f;  // modules.py:60

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------------
// grid (G, ceil(Lq/16)); block 256 (8 warps; warp w owns heads w, w+8, ...).
// smem: V tile [16*KS][d+8] bf16 (rows >= nk zero-filled) + context tile [16][d+8] bf16.
template <int KS, int NT>  // KS = 16-key steps (nk <= 16*KS); NT = 8-column tiles per head (dk = 8*NT, NT even)
__global__ void __launch_bounds__(256, (KS <= 1 ? 4 : KS <= 3 ? 3 : 1)) softmax_pv_mma_kernel(SoftmaxPvArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  const int d = a.h * a.dk, ldv = d + 8;  // +8 bf16 = 16 bytes of row padding: ldmatrix rows hit distinct banks
  __nv_bfloat16* Vs = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Os = Vs + (size_t)16 * KS * ldv;
  const int g = blockIdx.x, l0 = blockIdx.y * 16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nk = a.nk;

  // ---- stage the group's V rows (cp.async, 16 B per request; one warp per row), zero the padding rows ------------
  {
    const __nv_bfloat16* Vg = reinterpret_cast<const __nv_bfloat16*>(a.V) + a.vg.off(g);
    for (int k = warp; k < 16 * KS; k += 8) {
      __nv_bfloat16* drow = Vs + (size_t)k * ldv;
      if (k < nk) {
        const __nv_bfloat16* srow = Vg + a.v.off(k);
        for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), srow + c);
      } else {
        for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  }

  const float* Sg = a.S + a.sg.off(g);
  const uint8_t* Mg = a.mask ? a.mask + (long long)((g / a.mask_gdiv) % a.mask_gmod) * a.m_b : nullptr;
  const int r0 = l0 + (lane >> 2), r1 = r0 + 8;  // the two query rows this thread owns in the fragments
  const int kq = (lane & 3) * 2;
  const bool vec2 = (a.s_k == 1) && ((a.s_l | a.s_h) % 2 == 0) && (a.sg.so % 2 == 0) && (a.sg.si % 2 == 0) &&
                    ((reinterpret_cast<uintptr_t>(a.S) & 7) == 0) &&
                    (!a.mask || (a.m_b % 2 == 0 && (reinterpret_cast<uintptr_t>(a.mask) & 1) == 0));

  for (int hh = warp; hh < a.h; hh += 8) {
    // ---- scores of this head in A-fragment order: s[ks][0..1] row r0 k, [2..3] row r1 k, [4..5] r0 k+8, [6..7] r1 k+8
    float s[KS][8];
    if (vec2 && a.m_l == 0) {
      // fast path: contiguous keys, key mask shared by all query rows
      const float* p0 = Sg + (long long)r0 * a.s_l + (long long)hh * a.s_h;
      const float* p1 = p0 + 8 * a.s_l;
      const bool ok0 = r0 < a.Lq, ok1 = r1 < a.Lq;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int k = ks * 16 + half * 8 + kq;
          float2 t0 = make_float2(-INFINITY, -INFINITY), t1 = t0;
          if (k + 1 < nk) {
            if (ok0) t0 = __ldg(reinterpret_cast<const float2*>(p0 + k));
            if (ok1) t1 = __ldg(reinterpret_cast<const float2*>(p1 + k));
            if (Mg) {
              const uchar2 m = *reinterpret_cast<const uchar2*>(Mg + k);  // k even
              if (m.x == 0) { t0.x = ok0 ? kMaskFill : t0.x; t1.x = ok1 ? kMaskFill : t1.x; }
              if (m.y == 0) { t0.y = ok0 ? kMaskFill : t0.y; t1.y = ok1 ? kMaskFill : t1.y; }
            }
          } else if (k < nk) {
            if (ok0) t0.x = __ldg(p0 + k);
            if (ok1) t1.x = __ldg(p1 + k);
            if (Mg && Mg[k] == 0) { t0.x = ok0 ? kMaskFill : t0.x; t1.x = ok1 ? kMaskFill : t1.x; }
          }
          s[ks][half * 4 + 0] = t0.x; s[ks][half * 4 + 1] = t0.y;
          s[ks][half * 4 + 2] = t1.x; s[ks][half * 4 + 3] = t1.y;
        }
      }
    } else {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int k = ks * 16 + half * 8 + kq;
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            const int row = rr ? r1 : r0;
            float v0 = -INFINITY, v1 = -INFINITY;
            if (row < a.Lq && k < nk) {
              const float* p = Sg + (long long)row * a.s_l + (long long)hh * a.s_h + (long long)k * a.s_k;
              v0 = __ldg(p);
              if (k + 1 < nk) v1 = __ldg(p + a.s_k);
              if (Mg) {
                const uint8_t* mrow = Mg + (long long)row * a.m_l;
                if (mrow[k] == 0) v0 = kMaskFill;
                if (k + 1 < nk && mrow[k + 1] == 0) v1 = kMaskFill;
              }
            }
            s[ks][half * 4 + rr * 2 + 0] = v0;
            s[ks][half * 4 + rr * 2 + 1] = v1;
          }
        }
      }
    }
    // ---- softmax per row (a row lives in the 4 lanes of a quad) ----------------------------------------------
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      mx0 = fmaxf(mx0, fmaxf(fmaxf(s[ks][0], s[ks][1]), fmaxf(s[ks][4], s[ks][5])));
      mx1 = fmaxf(mx1, fmaxf(fmaxf(s[ks][2], s[ks][3]), fmaxf(s[ks][6], s[ks][7])));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    if (mx0 == -INFINITY) mx0 = 0.f;  // row beyond Lq: all -inf, keep exp() finite
    if (mx1 == -INFINITY) mx1 = 0.f;
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool row1 = (j & 2) != 0;
        const float e = __expf(s[ks][j] - (row1 ? mx1 : mx0));
        s[ks][j] = e;
        if (row1) sum1 += e; else sum0 += e;
      }
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = sum0 > 0.f ? 1.f / sum0 : 0.f, inv1 = sum1 > 0.f ? 1.f / sum1 : 0.f;
    if (a.drop.thresh != 0u) {  // dropout on the probabilities: element ((g*h + hh)*Lq + row)*nk + key
      const unsigned long long base0 = (((unsigned long long)g * a.h + hh) * a.Lq + r0) * nk;
      const unsigned long long base1 = (((unsigned long long)g * a.h + hh) * a.Lq + r1) * nk;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int key = ks * 16 + (j >> 2) * 8 + kq + (j & 1);
          s[ks][j] *= drop_mul(a.drop, ((j & 2) ? base1 : base0) + key);
        }
    }
    uint32_t pa[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      pa[ks][0] = pack_bf16(s[ks][0] * inv0, s[ks][1] * inv0);  // (row r0, k..k+1)
      pa[ks][1] = pack_bf16(s[ks][2] * inv1, s[ks][3] * inv1);  // (row r1, k..k+1)
      pa[ks][2] = pack_bf16(s[ks][4] * inv0, s[ks][5] * inv0);  // (row r0, k+8..)
      pa[ks][3] = pack_bf16(s[ks][6] * inv1, s[ks][7] * inv1);  // (row r1, k+8..)
    }
    if (hh == warp) {  // first head of this warp: V must have landed
      cp_async_wait_all();
      __syncthreads();
    }
    // ---- ctx[l, hh*dk + j] = sum_k P[l,k] V[k, hh*dk + j] -----------------------------------------------------
    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
    // ldmatrix.x4.trans lane addresses: lanes 0-7 rows k..k+7 @ col n0, 8-15 rows k+8.. @ n0, 16-23 rows k.. @ n0+8, 24-31 rows k+8.. @ n0+8
    const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lcol = (lane >> 4) * 8;
    const uint32_t vbase = smem_addr(Vs) + (uint32_t)((lrow * ldv + hh * a.dk + lcol) * 2);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int np = 0; np < NT / 2; ++np) {
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(vbase + (uint32_t)((ks * 16 * ldv + np * 16) * 2), b0, b1, b2, b3);
        mma_bf16_16816(acc[2 * np], pa[ks], b0, b1);
        mma_bf16_16816(acc[2 * np + 1], pa[ks], b2, b3);
      }
    }
    // ---- accumulators -> context tile in smem (bf16) ----------------------------------------------------------
    const int orow = lane >> 2, ocol = hh * a.dk + kq;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      *reinterpret_cast<uint32_t*>(Os + (size_t)orow * ldv + ocol + n * 8) = pack_bf16(acc[n][0], acc[n][1]);
      *reinterpret_cast<uint32_t*>(Os + (size_t)(orow + 8) * ldv + ocol + n * 8) = pack_bf16(acc[n][2], acc[n][3]);
    }
  }
  if (warp >= a.h) {  // warps without a head still take part in the V barrier
    cp_async_wait_all();
    __syncthreads();
  }
  __syncthreads();
  // ---- coalesced 16-byte stores of the context rows (one warp per row) -----------------------------------------------
  {
    __nv_bfloat16* Cg = reinterpret_cast<__nv_bfloat16*>(a.ctx) + a.og.off(g);
    const int nl = min(16, a.Lq - l0);
    for (int li = warp; li < nl; li += 8) {
      __nv_bfloat16* crow = Cg + (long long)(l0 + li) * a.o_l;
      for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(crow + c) = *reinterpret_cast<const uint4*>(Os + (size_t)li * ldv + c);
    }
  }
}

template <int KS, int NT>
int launch_pv(const SoftmaxPvArgs& a, cudaStream_t st) {
  const int d = a.h * a.dk;
  const size_t smem = (size_t)(16 * KS + 16) * (d + 8) * 2;
  auto kern = softmax_pv_mma_kernel<KS, NT>;
  static bool configured = false;
  if (!configured && smem > 48 * 1024) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  dim3 grid(a.G, (a.Lq + 15) / 16);
  ProfScope ps("softmax_pv_mma", st);
  BIST_LAUNCH(kern, grid, 256, smem, st, a);
  BIST_LAUNCHED();
  return 0;
}

int pv_num_sms() {
  static const int n = [] {
    int dev = 0, v = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    return v;
  }();
  return n;
}


template <int KS>
int pick_nt(const SoftmaxPvArgs& a, cudaStream_t st) {
  switch (a.dk / 8) {
    case 2: return launch_pv<KS, 2>(a, st);
    case 4: return launch_pv<KS, 4>(a, st);
    case 8: return launch_pv<KS, 8>(a, st);
    case 16: return launch_pv<KS, 16>(a, st);
    default: return -1;
  }
}

// ---------------------------------------------------------------------------------------------
// stage 2 on tensor cores (bf16): grid B*Lq; block 256 (8 warps).  One (b,l) per CTA:
//   S[hh, g] = q[hh,:] . x1[g,:]        mma.sync, heads padded to one 16-row tile, warp w owns key tiles w, w+8, ..
//   P = softmax_g(mask(S))              one warp per head
//   U[hh, :] = sum_g P[hh,g] x1[g,:]    mma.sync, warp w owns columns [w*d/8, (w+1)*d/8)
// smem: X [Gp][d+8] bf16 (Gp = G rounded up to 16, padding rows zero) | Q/U [16][d+8] bf16 | S [16][Gp] f32 | P [16][Gp+8] bf16
__global__ void __launch_bounds__(256) stage2_mma_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ x1,
                                                         const uint8_t* __restrict__ mask, __nv_bfloat16* __restrict__ out,
                                                         int G, int Gp, int Lq, int h, int d, DropCfg drop, float* __restrict__ rsum) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  const int ldx = d + 8, ldp = Gp + 8;
  __nv_bfloat16* Xs = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Qs = Xs + (size_t)Gp * ldx;
  float* Ss = reinterpret_cast<float*>(Qs + (size_t)16 * ldx);
  __nv_bfloat16* Ps = reinterpret_cast<__nv_bfloat16*>(Ss + (size_t)16 * Gp);
  const int bl = blockIdx.x, b = bl / Lq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cpr = d >> 3;
  __shared__ uint8_t s_mask[256];   // key mask of the sample, fetched beside the tile (G <= 256 by the host check)
  if (mask) { for (int g = threadIdx.x; g < G; g += 256) s_mask[g] = mask[(long long)b * G + g]; }
  // ---- stage X (G rows) and Q (h rows); zero the padding rows ---------------------------------------------------
  {
    const __nv_bfloat16* xg = x1 + (long long)bl * G * d;
    for (int r = warp; r < Gp; r += 8) {
      __nv_bfloat16* drow = Xs + (size_t)r * ldx;
      if (r < G) for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), xg + (long long)r * d + c);
      else for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = make_uint4(0u, 0u, 0u, 0u);
    }
    const __nv_bfloat16* qg = q + (long long)bl * h * d;
    for (int r = warp; r < 16; r += 8) {
      __nv_bfloat16* drow = Qs + (size_t)r * ldx;
      if (r < h) for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), qg + (long long)r * d + c);
      else for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = make_uint4(0u, 0u, 0u, 0u);
    }
    cp_async_wait_all();
    __syncthreads();
  }
  // ---- scores: warp per 8-key tile -----------------------------------------------------------------------------------
  {
    const uint32_t qa = smem_addr(Qs) + (uint32_t)(((lane & 15) * ldx + (lane >> 4) * 8) * 2);  // A: rows lane%16, k-half lane/16
    for (int nt = warp; nt < Gp / 8; nt += 8) {
      // B[k][n] = X[n][k]: non-transposed 8x8 tiles of X rows nt*8.., four consecutive 8-wide k slices per ldmatrix.x4
      const uint32_t xb = smem_addr(Xs) + (uint32_t)(((nt * 8 + (lane & 7)) * ldx + (lane >> 3) * 8) * 2);
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int k0 = 0; k0 < d; k0 += 32) {
        uint32_t a0[4], a1[4], b0, b1, b2, b3;
        ldmatrix_x4(qa + (uint32_t)(k0 * 2), a0[0], a0[1], a0[2], a0[3]);
        ldmatrix_x4(xb + (uint32_t)(k0 * 2), b0, b1, b2, b3);
        mma_bf16_16816(acc, a0, b0, b1);
        if (k0 + 16 < d) {
          ldmatrix_x4(qa + (uint32_t)((k0 + 16) * 2), a1[0], a1[1], a1[2], a1[3]);
          mma_bf16_16816(acc, a1, b2, b3);
        }
      }
      const int r = lane >> 2, c = nt * 8 + (lane & 3) * 2;
      Ss[r * Gp + c] = acc[0]; Ss[r * Gp + c + 1] = acc[1];
      Ss[(r + 8) * Gp + c] = acc[2]; Ss[(r + 8) * Gp + c + 1] = acc[3];
    }
  }
  __syncthreads();
  // ---- masked softmax over g: warp per head row; rows >= h and columns >= G of P are zero -------------------------------
  for (int hh = warp; hh < 16; hh += 8) {
    if (hh < h) {
      float mx = -INFINITY;
      for (int g = lane; g < G; g += 32) {
        float sc = Ss[hh * Gp + g];
        if (mask && s_mask[g] == 0) sc = kMaskFill;
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
      float rs = 0.f;
      for (int g = lane; g < Gp; g += 32) {
        float pv = 0.f;
        if (g < G) pv = Ss[hh * Gp + g] * inv * drop_mul(drop, ((unsigned long long)bl * h + hh) * G + g);
        const __nv_bfloat16 pb = __float2bfloat16_rn(pv);
        Ps[hh * ldp + g] = pb;
        rs += __bfloat162float(pb);
      }
      if (rsum) {
        rs = warp_sum(rs);
        if (lane == 0) rsum[(long long)bl * h + hh] = rs;
      }
    } else {
      for (int g = lane; g < Gp; g += 32) Ps[hh * ldp + g] = __float2bfloat16_rn(0.f);
    }
  }
  __syncthreads();
  // ---- U = P . X: warp owns d/8 columns; result rows < h go to the Q tile (reused as the output stage) ---------------------
  {
    const int ncol = d / 8;                       // columns per warp (multiple of 16 by the host check)
    const uint32_t pa = smem_addr(Ps) + (uint32_t)(((lane & 15) * ldp + (lane >> 4) * 8) * 2);
    const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lcol = (lane >> 4) * 8;
    const uint32_t xb = smem_addr(Xs) + (uint32_t)((lrow * ldx + warp * ncol + lcol) * 2);
    for (int n0 = 0; n0 < ncol; n0 += 16) {
      float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
      for (int k0 = 0; k0 < Gp; k0 += 16) {
        uint32_t a[4], b0, b1, b2, b3;
        ldmatrix_x4(pa + (uint32_t)(k0 * 2), a[0], a[1], a[2], a[3]);
        ldmatrix_x4_trans(xb + (uint32_t)((k0 * ldx + n0) * 2), b0, b1, b2, b3);
        mma_bf16_16816(acc0, a, b0, b1);
        mma_bf16_16816(acc1, a, b2, b3);
      }
      const int r = lane >> 2, c = warp * ncol + n0 + (lane & 3) * 2;
      // Q tile is dead after the score phase (barrier above): stage the 16 x d result there
      *reinterpret_cast<uint32_t*>(Qs + (size_t)r * ldx + c) = pack_bf16(acc0[0], acc0[1]);
      *reinterpret_cast<uint32_t*>(Qs + (size_t)(r + 8) * ldx + c) = pack_bf16(acc0[2], acc0[3]);
      *reinterpret_cast<uint32_t*>(Qs + (size_t)r * ldx + c + 8) = pack_bf16(acc1[0], acc1[1]);
      *reinterpret_cast<uint32_t*>(Qs + (size_t)(r + 8) * ldx + c + 8) = pack_bf16(acc1[2], acc1[3]);
    }
  }
  __syncthreads();
  {
    __nv_bfloat16* og = out + (long long)bl * h * d;
    for (int r = warp; r < h; r += 8)
      for (int c = lane * 8; c < d; c += 256)
        *reinterpret_cast<uint4*>(og + (long long)r * d + c) = *reinterpret_cast<const uint4*>(Qs + (size_t)r * ldx + c);
  }
}

// ---------------------------------------------------------------------------------------------
// stage 2: grid B*Lq; block 256.  smem: x1 tile [G][d] (input dtype) + qf [h][d] fp32 + p [h][G] fp32.
template <typename TX, typename TO, int H>
__global__ void __launch_bounds__(256) stage2_v2_kernel(const float* __restrict__ qf, const TX* __restrict__ x1,
                                                        const uint8_t* __restrict__ mask, TO* __restrict__ ctxx,
                                                        int G, int Lq, int d) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  TX* Xs = reinterpret_cast<TX*>(smem_raw);                                        // [G][d]
  float* Qs = reinterpret_cast<float*>(smem_raw + (((size_t)G * d * sizeof(TX) + 15) & ~(size_t)15));  // [H][d]
  float* Ps = Qs + (size_t)H * d;                                                  // [H][G]
  const int bl = blockIdx.x, b = bl / Lq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  // ---- one coalesced pass over the (b,l) block of stage-1 summaries and the folded queries ------------------------
  {
    const uint4* src = reinterpret_cast<const uint4*>(x1 + (long long)bl * G * d);
    uint4* dst = reinterpret_cast<uint4*>(Xs);
    const int n16 = (int)((size_t)G * d * sizeof(TX) / 16);
    for (int i = threadIdx.x; i < n16; i += blockDim.x) cp_async16(smem_addr(dst + i), src + i);
    const uint4* qsrc = reinterpret_cast<const uint4*>(qf + (long long)bl * H * d);
    uint4* qdst = reinterpret_cast<uint4*>(Qs);
    for (int i = threadIdx.x; i < H * d / 4; i += blockDim.x) cp_async16(smem_addr(qdst + i), qsrc + i);
    cp_async_wait_all();
    __syncthreads();
  }
  // ---- scores[hh][g] = qf[hh,:] . x1[g,:]   (one warp per key, lanes stride the columns by 4) ------------------------
  for (int g = warp; g < G; g += nwarp) {
    float part[H];
#pragma unroll
    for (int i = 0; i < H; ++i) part[i] = 0.f;
    const TX* xr = Xs + (size_t)g * d;
    for (int c = lane * 4; c < d; c += 128) {
      const float4 xv = ld4(xr + c);
#pragma unroll
      for (int i = 0; i < H; ++i) {
        const float4 q = *reinterpret_cast<const float4*>(Qs + (size_t)i * d + c);
        part[i] = fmaf(q.x, xv.x, fmaf(q.y, xv.y, fmaf(q.z, xv.z, fmaf(q.w, xv.w, part[i]))));
      }
    }
    const bool masked = mask && mask[(long long)b * G + g] == 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const float sc = warp_sum(part[i]);
      if (lane == 0) Ps[i * G + g] = masked ? kMaskFill : sc;
    }
  }
  __syncthreads();
  // ---- softmax over g: one warp per head -------------------------------------------------------------------------------
  for (int hh = warp; hh < H; hh += nwarp) {
    float mx = -INFINITY;
    for (int g = lane; g < G; g += 32) mx = fmaxf(mx, Ps[hh * G + g]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int g = lane; g < G; g += 32) {
      const float e = sizeof(TX) == 4 ? expf(Ps[hh * G + g] - mx) : __expf(Ps[hh * G + g] - mx);  // fp32 mode: exact exp
      Ps[hh * G + g] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int g = lane; g < G; g += 32) Ps[hh * G + g] *= inv;
  }
  __syncthreads();
  // ---- out[hh][c] = sum_g p[hh][g] x1[g][c]: each thread owns 2 columns for all heads -------------------------------------
  TO* out = ctxx + (long long)bl * H * d;
  for (int c = threadIdx.x * 2; c < d; c += blockDim.x * 2) {
    float acc[H][2];
#pragma unroll
    for (int i = 0; i < H; ++i) acc[i][0] = acc[i][1] = 0.f;
    for (int g = 0; g < G; ++g) {
      const float x0 = to_f32(Xs[(size_t)g * d + c]), x1v = to_f32(Xs[(size_t)g * d + c + 1]);
#pragma unroll
      for (int i = 0; i < H; ++i) {
        const float p = Ps[i * G + g];
        acc[i][0] = fmaf(p, x0, acc[i][0]);
        acc[i][1] = fmaf(p, x1v, acc[i][1]);
      }
    }
#pragma unroll
    for (int i = 0; i < H; ++i) {
      out[(long long)i * d + c] = from_f32<TO>(acc[i][0]);
      out[(long long)i * d + c + 1] = from_f32<TO>(acc[i][1]);
    }
  }
}

template <typename TX, typename TO, int H>
int launch_s2(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int d, size_t smem,
              cudaStream_t st) {
  auto kern = stage2_v2_kernel<TX, TO, H>;
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  ProfScope ps("stage2_v2", st);
  BIST_LAUNCH(kern, dim3((unsigned)(B * Lq)), 256, smem, st, qf, reinterpret_cast<const TX*>(x1), mask, reinterpret_cast<TO*>(ctxx), G, Lq, d);
  BIST_LAUNCHED();
  return 0;
}

template <int H>
int pick_s2(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int d, size_t smem,
            bool xb, bool ob, cudaStream_t st) {
  if (xb && ob) return launch_s2<__nv_bfloat16, __nv_bfloat16, H>(qf, x1, mask, ctxx, B, G, Lq, d, smem, st);
  if (xb) return launch_s2<__nv_bfloat16, float, H>(qf, x1, mask, ctxx, B, G, Lq, d, smem, st);
  if (ob) return launch_s2<float, __nv_bfloat16, H>(qf, x1, mask, ctxx, B, G, Lq, d, smem, st);
  return launch_s2<float, float, H>(qf, x1, mask, ctxx, B, G, Lq, d, smem, st);
}

}  // namespace

// Tensor-core softmax.PV; *taken = false when the shape is left to the CUDA-core kernel (attn.cu).
int softmax_pv_mma(const SoftmaxPvArgs& a, cudaStream_t st, bool* taken) {
  *taken = false;
  const int d = a.h * a.dk;
  if (a.v_dtype != BIST_BF16 || a.ctx_dtype != BIST_BF16 || a.p_out) return 0;
  if (a.nk > 128 || a.dk % 16 != 0 || a.dk > 128 || d % 8 != 0) return 0;
  if ((reinterpret_cast<uintptr_t>(a.V) & 15) || (reinterpret_cast<uintptr_t>(a.ctx) & 15)) return 0;
  if ((a.v.so | a.v.si | a.vg.so | a.vg.si | a.og.so | a.og.si | a.o_l) % 8 != 0) return 0;  // 16-byte row starts
  if (a.v.inner != kNoSplit && a.v.inner <= 0) return 0;
  const int ks = (a.nk + 15) / 16;
  if ((size_t)(16 * ks + 16) * (d + 8) * 2 > 200 * 1024) return 0;
  int rc = -1;
  if (ks == 1) rc = pick_nt<1>(a, st);
  else if (ks == 2) rc = pick_nt<2>(a, st);
  else if (ks == 3) rc = pick_nt<3>(a, st);
  else if (ks == 4) rc = pick_nt<4>(a, st);
  else rc = pick_nt<8>(a, st);
  if (rc == -1) return 0;
  *taken = true;
  return rc;
}

// ---------------------------------------------------------------------------------------------
// Fused self-attention over the query tokens (encoder.py:176,184 -> modules.py:81-100 without the output
// projection): one CTA per (direction, b); the [Lq, 3d] q|k|v rows are staged once; warp w owns heads w, w+8, ..
//   S = Q_h K_h^T / sqrt(dk) -> key mask -> softmax -> P V_h, all in registers (C fragments of two adjacent
//   8-key tiles are the A fragment of the next mma).  KT = ceil(Lq/16) <= 4.
template <int KT, int DKS>  // DKS = dk / 16
__global__ void __launch_bounds__(256) self_attn_mma_kernel(const __nv_bfloat16* __restrict__ qkv, const uint8_t* __restrict__ mask,
                                                            int mask_mod, __nv_bfloat16* __restrict__ ctx, int Lq, int h, float scale) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  constexpr int LP = 16 * KT, dk = 16 * DKS;
  const int d = h * dk, ld = 3 * d + 8;
  __nv_bfloat16* Ts = reinterpret_cast<__nv_bfloat16*>(smem_raw);  // [LP][3d+8]; the q columns are reused as the output stage
  const int g = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {
    const __nv_bfloat16* src = qkv + (long long)g * Lq * 3 * d;
    const int cpr = (3 * d) >> 3;
    for (int i = threadIdx.x; i < LP * cpr; i += blockDim.x) {
      const int r = i / cpr, c = (i - r * cpr) << 3;
      if (r < Lq) cp_async16(smem_addr(Ts + (size_t)r * ld + c), src + (long long)r * 3 * d + c);
      else *reinterpret_cast<uint4*>(Ts + (size_t)r * ld + c) = make_uint4(0u, 0u, 0u, 0u);
    }
    cp_async_wait_all();
    __syncthreads();
  }
  const uint8_t* mrow = mask ? mask + (long long)(g % mask_mod) * Lq : nullptr;
  const int kq = (lane & 3) * 2;
  for (int hh = warp; hh < h; hh += 8) {
    const uint32_t qa = smem_addr(Ts) + (uint32_t)(((lane & 15) * ld + hh * dk + (lane >> 4) * 8) * 2);
    const uint32_t kb = smem_addr(Ts) + (uint32_t)(((lane & 7) * ld + d + hh * dk + (lane >> 3) * 8) * 2);
    const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lcol = (lane >> 4) * 8;
    const uint32_t vb = smem_addr(Ts) + (uint32_t)((lrow * ld + 2 * d + hh * dk + lcol) * 2);
    float o[KT][2 * DKS][4];
#pragma unroll
    for (int mt = 0; mt < KT; ++mt) {
      // ---- scores for query rows mt*16.., all LP keys -------------------------------------------------------
      float sc[2 * KT][4];
#pragma unroll
      for (int nt = 0; nt < 2 * KT; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < DKS; ks += 2) {  // one ldmatrix.x4 of K covers two 16-wide k steps
        uint32_t a0[4], a1[4];
        ldmatrix_x4(qa + (uint32_t)((mt * 16 * ld + ks * 16) * 2), a0[0], a0[1], a0[2], a0[3]);
        if (ks + 1 < DKS) ldmatrix_x4(qa + (uint32_t)((mt * 16 * ld + (ks + 1) * 16) * 2), a1[0], a1[1], a1[2], a1[3]);
#pragma unroll
        for (int nt = 0; nt < 2 * KT; ++nt) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(kb + (uint32_t)((nt * 8 * ld + ks * 16) * 2), b0, b1, b2, b3);
          mma_bf16_16816(sc[nt], a0, b0, b1);
          if (ks + 1 < DKS) mma_bf16_16816(sc[nt], a1, b2, b3);
        }
      }
      // ---- scale, key mask, softmax (rows r = lane/4 and r+8 of this tile) -----------------------------------------
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 2 * KT; ++nt) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int key = nt * 8 + kq + j;
          const bool valid = key < Lq;
          const bool masked = valid && mrow && mrow[key] == 0;
          float v0 = valid ? (masked ? kMaskFill : sc[nt][j] * scale) : -INFINITY;
          float v1 = valid ? (masked ? kMaskFill : sc[nt][2 + j] * scale) : -INFINITY;
          sc[nt][j] = v0; sc[nt][2 + j] = v1;
          mx0 = fmaxf(mx0, v0); mx1 = fmaxf(mx1, v1);
        }
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 2 * KT; ++nt) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float e0 = __expf(sc[nt][j] - mx0), e1 = __expf(sc[nt][2 + j] - mx1);
          sc[nt][j] = e0; sc[nt][2 + j] = e1;
          s0 += e0; s1 += e1;
        }
      }
      s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
      const float i0 = 1.f / s0, i1 = 1.f / s1;
      // ---- P V --------------------------------------------------------------------------------------------------------
#pragma unroll
      for (int n = 0; n < 2 * DKS; ++n) o[mt][n][0] = o[mt][n][1] = o[mt][n][2] = o[mt][n][3] = 0.f;
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) {
        uint32_t pa[4];
        pa[0] = pack_bf16(sc[2 * kt][0] * i0, sc[2 * kt][1] * i0);
        pa[1] = pack_bf16(sc[2 * kt][2] * i1, sc[2 * kt][3] * i1);
        pa[2] = pack_bf16(sc[2 * kt + 1][0] * i0, sc[2 * kt + 1][1] * i0);
        pa[3] = pack_bf16(sc[2 * kt + 1][2] * i1, sc[2 * kt + 1][3] * i1);
#pragma unroll
        for (int np = 0; np < DKS; ++np) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(vb + (uint32_t)((kt * 16 * ld + np * 16) * 2), b0, b1, b2, b3);
          mma_bf16_16816(o[mt][2 * np], pa, b0, b1);
          mma_bf16_16816(o[mt][2 * np + 1], pa, b2, b3);
        }
      }
    }
    // every warp only ever reads its own heads' q columns, so they can be overwritten with the output now
    __syncwarp();
#pragma unroll
    for (int mt = 0; mt < KT; ++mt) {
      const int r = mt * 16 + (lane >> 2);
#pragma unroll
      for (int n = 0; n < 2 * DKS; ++n) {
        *reinterpret_cast<uint32_t*>(Ts + (size_t)r * ld + hh * dk + n * 8 + kq) = pack_bf16(o[mt][n][0], o[mt][n][1]);
        *reinterpret_cast<uint32_t*>(Ts + (size_t)(r + 8) * ld + hh * dk + n * 8 + kq) = pack_bf16(o[mt][n][2], o[mt][n][3]);
      }
    }
  }
  __syncthreads();
  {
    __nv_bfloat16* dst = ctx + (long long)g * Lq * d;
    const int cpr = d >> 3;
    for (int i = threadIdx.x; i < Lq * cpr; i += blockDim.x) {
      const int r = i / cpr, c = (i - r * cpr) << 3;
      *reinterpret_cast<uint4*>(dst + (long long)r * d + c) = *reinterpret_cast<const uint4*>(Ts + (size_t)r * ld + c);
    }
  }
}

template <int KT>
int launch_self(const void* qkv, const uint8_t* mask, int mask_mod, void* ctx, int G, int Lq, int h, int dk, cudaStream_t st) {
  const int d = h * dk;
  const size_t smem = (size_t)16 * KT * (3 * d + 8) * 2;
  const float scale = 1.f / sqrtf((float)dk);
  auto go = [&](auto kern) -> int {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    ProfScope ps("self_attn_mma", st);
    BIST_LAUNCH(kern, dim3((unsigned)G), 256, smem, st, reinterpret_cast<const __nv_bfloat16*>(qkv), mask, mask_mod,
                                              reinterpret_cast<__nv_bfloat16*>(ctx), Lq, h, scale);
    BIST_LAUNCHED();
    return 0;
  };
  switch (dk / 16) {
    case 1: return go(self_attn_mma_kernel<KT, 1>);
    case 2: return go(self_attn_mma_kernel<KT, 2>);
    case 4: return go(self_attn_mma_kernel<KT, 4>);
    default: return -1;
  }
}

int self_attn_mma(const void* qkv, const uint8_t* mask, int mask_mod, void* ctx, int G, int Lq, int h, int dk, cudaStream_t st,
                  bool* taken) {
  *taken = false;
  const int d = h * dk, kt = (Lq + 15) / 16;
  if (kt > 4 || (dk != 16 && dk != 32 && dk != 64) || (size_t)16 * kt * (3 * d + 8) * 2 > 200 * 1024) return 0;
  if ((reinterpret_cast<uintptr_t>(qkv) & 15) || (reinterpret_cast<uintptr_t>(ctx) & 15)) return 0;
  *taken = true;
  if (kt == 1) return launch_self<1>(qkv, mask, mask_mod, ctx, G, Lq, h, dk, st);
  if (kt == 2) return launch_self<2>(qkv, mask, mask_mod, ctx, G, Lq, h, dk, st);
  if (kt == 3) return launch_self<3>(qkv, mask, mask_mod, ctx, G, Lq, h, dk, st);
  return launch_self<4>(qkv, mask, mask_mod, ctx, G, Lq, h, dk, st);
}

bool stage2_mma_supported(int G, int h, int d) {
  const int Gp = (G + 15) / 16 * 16;
  const size_t smem = (size_t)(Gp + 16) * (d + 8) * 2 + (size_t)16 * Gp * 4 + (size_t)16 * (Gp + 8) * 2;
  return h <= 16 && d % 128 == 0 && smem <= 200 * 1024 && G <= 256;
}

// bf16 tensor-core stage 2 (q, x1, out all bf16); *taken = false when the shape is left to the other kernels.
int stage2_mma(const void* q, const void* x1, const uint8_t* mask, void* out, int B, int G, int Lq, int h, int d,
               cudaStream_t st, bool* taken, const Stage2Extra* ex) {
  *taken = false;
  const int Gp = (G + 15) / 16 * 16;
  const size_t smem = (size_t)(Gp + 16) * (d + 8) * 2 + (size_t)16 * Gp * 4 + (size_t)16 * (Gp + 8) * 2;
  if (h > 16 || d % 128 != 0 || smem > 200 * 1024 || G > 256) return 0;
  if ((reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(x1) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) return 0;
  *taken = true;
  {
    // pipelined persistent form (stage2_pipe.cu) when the shape fits it
    bool piped = false;
    BIST_TRY(stage2_pipe(q, x1, mask, out, B, G, Lq, h, d, st, &piped, ex));
    if (piped) return 0;
  }
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(stage2_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  ProfScope ps("stage2_mma", st);
  BIST_LAUNCH(stage2_mma_kernel, dim3((unsigned)(B * Lq)), 256, smem, st, reinterpret_cast<const __nv_bfloat16*>(q),
                                                                reinterpret_cast<const __nv_bfloat16*>(x1), mask,
                                                                reinterpret_cast<__nv_bfloat16*>(out), G, Gp, Lq, h, d,
                                                                ex ? ex->drop : DropCfg(), ex ? ex->rsum : nullptr);
  BIST_LAUNCHED();
  return 0;
}

// Single-pass stage 2 over x1 laid out [B, Lq, G, d]; *taken = false when the tile does not fit in shared memory.
int stage2_v2(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int h, int d,
              int x1_dtype, int ctxx_dtype, cudaStream_t st, bool* taken) {
  *taken = false;
  const bool xb = x1_dtype == BIST_BF16, ob = ctxx_dtype == BIST_BF16;
  const size_t xbytes = align_up((size_t)G * d * (xb ? 2 : 4), 16);
  const size_t smem = xbytes + (size_t)h * d * 4 + (size_t)h * G * 4;
  if (smem > 200 * 1024 || d % 8 != 0 || (h != 8 && h != 4 && h != 2)) return 0;
  if ((reinterpret_cast<uintptr_t>(x1) & 15) || (reinterpret_cast<uintptr_t>(qf) & 15)) return 0;
  *taken = true;
  if (h == 8) return pick_s2<8>(qf, x1, mask, ctxx, B, G, Lq, d, smem, xb, ob, st);
  if (h == 4) return pick_s2<4>(qf, x1, mask, ctxx, B, G, Lq, d, smem, xb, ob, st);
  return pick_s2<2>(qf, x1, mask, ctxx, B, G, Lq, d, smem, xb, ob, st);
}

}  // namespace bist
