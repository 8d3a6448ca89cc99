// Tensor-core (bf16 mma.sync m16n8k16) backward kernels of the two attention stages of the BiST video path.
//   stage1_bwd_mma : per attention group (b,g): P = softmax(mask(S)) recomputed from the saved fp32 scores,
//                    dP = dC_h XV_h^T, dS = P o (dP - rowsum(P o dP)), dXV_h = P^T dC_h; one warp per head,
//                    XV rows staged once with cp.async, dXV written back over the staged tile and stored with
//                    16-byte coalesced rows, dS^T staged in shared memory and stored as whole [Lq*h] rows.
//   stage2_bwd_mma : per (b,l): S/dP = [qf; du] X1^T as ONE 16-row product (h <= 8), softmax backward in registers /
//                    shared memory, dX1 = [P^T | dS^T] [du; qf], dqf = dS X1.
// (attention() backward, /root/reference/model/modules.py:54-64, in the folded formulation of st_train.cu.)
#include "common.cuh"

namespace bist {
namespace {

constexpr float kMaskFill = -1e9f;  // modules.py:60

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ------------------------------------------------------------------------------------------------ stage 1
// grid G groups; block 256 (8 warps; warp w owns heads w, w+8, ..).  Lq <= 16.
// smem (bf16): Vs [16*KS][ldv] (XV rows, later dXV) | Cs [16][ldv] (dC rows) | Ps [8 warps][16][ldp] | Ts [16*KS][ldr] (dS^T)
template <int KS, int NT>  // nk <= 16*KS keys; dk = 8*NT (NT even)
__global__ void __launch_bounds__(256) stage1_bwd_mma_kernel(Stage1BwdArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  const SoftmaxPvArgs& f = a.f;
  const int h = f.h, dk = f.dk, d = h * dk, ldv = d + 8, nk = f.nk, Lq = f.Lq;
  const int R = Lq * h, ldr = ((R + 7) & ~7) + 8, ldp = 16 * KS + 8;
  __nv_bfloat16* Vs = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Cs = Vs + (size_t)16 * KS * ldv;
  __nv_bfloat16* Ps = Cs + (size_t)16 * ldv;
  __nv_bfloat16* Ts = Ps + (size_t)8 * 16 * ldp;
  const int g = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  // ---- stage XV rows and dC rows (cp.async 16 B; one warp per row), zero the padding rows
  {
    const __nv_bfloat16* Vg = reinterpret_cast<const __nv_bfloat16*>(f.V) + f.vg.off(g);
    for (int k = warp; k < 16 * KS; k += 8) {
      __nv_bfloat16* drow = Vs + (size_t)k * ldv;
      if (k < nk) {
        const __nv_bfloat16* srow = Vg + f.v.off(k);
        for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), srow + c);
      } else {
        for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    const __nv_bfloat16* Cg = reinterpret_cast<const __nv_bfloat16*>(a.dctx) + f.og.off(g);
    for (int l = warp; l < 16; l += 8) {
      __nv_bfloat16* drow = Cs + (size_t)l * ldv;
      if (l < Lq) {
        const __nv_bfloat16* srow = Cg + (long long)l * f.o_l;
        for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), srow + c);
      } else {
        for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  }

  const float* Sg = f.S + f.sg.off(g);
  const uint8_t* Mg = f.mask ? f.mask + (long long)((g / f.mask_gdiv) % f.mask_gmod) * f.m_b : nullptr;
  const int r0 = lane >> 2, r1 = r0 + 8;  // the two query rows this thread owns in C fragments
  const int kq = (lane & 3) * 2;
  __nv_bfloat16* Pw = Ps + (size_t)warp * 16 * ldp;
  bool staged = false;

  for (int hh = warp; hh < h; hh += 8) {
    // ---- probabilities of this head, C-fragment order: s[ks][half*4 + rr*2 + {0,1}] = (row rr ? r1 : r0, key ks*16 + half*8 + kq + {0,1})
    float s[KS][8];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int k = ks * 16 + half * 8 + kq;
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int row = rr ? r1 : r0;
          float v0 = -INFINITY, v1 = -INFINITY;
          if (row < Lq && k < nk) {
            const float* p = Sg + (long long)row * f.s_l + (long long)hh * f.s_h + (long long)k * f.s_k;
            v0 = __ldg(p);
            if (k + 1 < nk) v1 = __ldg(p + f.s_k);
            if (Mg) {
              const uint8_t* mrow = Mg + (long long)row * f.m_l;
              if (mrow[k] == 0) v0 = kMaskFill;
              if (k + 1 < nk && mrow[k + 1] == 0) v1 = kMaskFill;
            }
          }
          s[ks][half * 4 + rr * 2 + 0] = v0;
          s[ks][half * 4 + rr * 2 + 1] = v1;
        }
      }
    }
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      mx0 = fmaxf(mx0, fmaxf(fmaxf(s[ks][0], s[ks][1]), fmaxf(s[ks][4], s[ks][5])));
      mx1 = fmaxf(mx1, fmaxf(fmaxf(s[ks][2], s[ks][3]), fmaxf(s[ks][6], s[ks][7])));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    if (mx0 == -INFINITY) mx0 = 0.f;
    if (mx1 == -INFINITY) mx1 = 0.f;
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool row1 = (j & 2) != 0;
        const float e = expf(s[ks][j] - (row1 ? mx1 : mx0));
        s[ks][j] = e;
        if (row1) sum1 += e; else sum0 += e;
      }
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = sum0 > 0.f ? 1.f / sum0 : 0.f, inv1 = sum1 > 0.f ? 1.f / sum1 : 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s[ks][j] *= (j & 2) ? inv1 : inv0;
      // P' = m o P (bf16; m = the forward's dropout multipliers) into this warp's [16 l][keys] tile for the transposed product below
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int k = ks * 16 + half * 8 + kq;
        float m[4] = {1.f, 1.f, 1.f, 1.f};
        if (f.drop.thresh != 0u) {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            m[e] = drop_mul(f.drop, (((unsigned long long)g * h + hh) * Lq + ((e & 2) ? r1 : r0)) * nk + k + (e & 1));
        }
        *reinterpret_cast<uint32_t*>(Pw + (size_t)r0 * ldp + k) = pack2(s[ks][half * 4 + 0] * m[0], s[ks][half * 4 + 1] * m[1]);
        *reinterpret_cast<uint32_t*>(Pw + (size_t)r1 * ldp + k) = pack2(s[ks][half * 4 + 2] * m[2], s[ks][half * 4 + 3] * m[3]);
      }
    }
    if (!staged) {  // first head of this warp: the staged tiles must have landed
      cp_async_wait_all();
      __syncthreads();
      staged = true;
    }
    __syncwarp();

    // ---- dP[l, key] = sum_j dC[l, hh*dk + j] XV[key, hh*dk + j]
    float dp[KS][8];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int j = 0; j < 8; ++j) dp[ks][j] = 0.f;
    {
      // A (dC, row-major [l][j]): matrices (l 0-7, j 0-7), (l 8-15, j 0-7), (l 0-7, j 8-15), (l 8-15, j 8-15)
      const int arow = (lane & 7) + ((lane >> 3) & 1) * 8, acol = (lane >> 4) * 8;
      const uint32_t abase = smem_addr(Cs) + (uint32_t)((arow * ldv + hh * dk + acol) * 2);
      // B (XV as [key][j], j contiguous): matrices (keys 0-7, j 0-7), (keys 0-7, j 8-15), (keys 8-15, j 0-7), (keys 8-15, j 8-15)
      const int brow = (lane & 7) + (lane >> 4) * 8, bcol = ((lane >> 3) & 1) * 8;
      const uint32_t bbase = smem_addr(Vs) + (uint32_t)((brow * ldv + hh * dk + bcol) * 2);
#pragma unroll
      for (int kk = 0; kk < NT / 2; ++kk) {  // 16 columns of the head per step
        uint32_t af[4];
        ldsm_x4(abase + (uint32_t)(kk * 16 * 2), af[0], af[1], af[2], af[3]);
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4(bbase + (uint32_t)((ks * 16 * ldv + kk * 16) * 2), b0, b1, b2, b3);
          mma16816(*reinterpret_cast<float(*)[4]>(&dp[ks][0]), af, b0, b1);
          mma16816(*reinterpret_cast<float(*)[4]>(&dp[ks][4]), af, b2, b3);
        }
      }
    }
    // ---- dS = P o (dP - rowsum(P o dP)), dP = m o dP'
    if (f.drop.thresh != 0u) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int key = ks * 16 + (j >> 2) * 8 + kq + (j & 1);
          dp[ks][j] *= drop_mul(f.drop, (((unsigned long long)g * h + hh) * Lq + ((j & 2) ? r1 : r0)) * nk + key);
        }
    }
    float dot0 = 0.f, dot1 = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j & 2) dot1 = fmaf(s[ks][j], dp[ks][j], dot1);
        else dot0 = fmaf(s[ks][j], dp[ks][j], dot0);
      }
    dot0 += __shfl_xor_sync(0xffffffffu, dot0, 1); dot0 += __shfl_xor_sync(0xffffffffu, dot0, 2);
    dot1 += __shfl_xor_sync(0xffffffffu, dot1, 1); dot1 += __shfl_xor_sync(0xffffffffu, dot1, 2);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int k = ks * 16 + half * 8 + kq;
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int row = rr ? r1 : r0;
          if (row < Lq) {
            const float dd0 = s[ks][half * 4 + rr * 2] * (dp[ks][half * 4 + rr * 2] - (rr ? dot1 : dot0));
            const float dd1 = s[ks][half * 4 + rr * 2 + 1] * (dp[ks][half * 4 + rr * 2 + 1] - (rr ? dot1 : dot0));
            Ts[(size_t)k * ldr + row * h + hh] = __float2bfloat16_rn(dd0);
            Ts[(size_t)(k + 1) * ldr + row * h + hh] = __float2bfloat16_rn(dd1);
          }
        }
      }
    }

    // ---- dXV[key, hh*dk + j] = sum_l P[l, key] dC[l, hh*dk + j]   (overwrites this head's columns of the staged XV)
    {
      // A = P^T via .trans from Pw[l][key]: matrices (keys 0-7, l 0-7), (keys 8-15, l 0-7), (keys 0-7, l 8-15), (keys 8-15, l 8-15)
      const int prow = (lane & 7) + (lane >> 4) * 8, pcol = ((lane >> 3) & 1) * 8;
      const uint32_t pbase = smem_addr(Pw) + (uint32_t)((prow * ldp + pcol) * 2);
      // B = dC via .trans from Cs[l][j]: matrices (l 0-7, j 0-7), (l 8-15, j 0-7), (l 0-7, j 8-15), (l 8-15, j 8-15)
      const int crow = (lane & 7) + ((lane >> 3) & 1) * 8, ccol = (lane >> 4) * 8;
      const uint32_t cbase = smem_addr(Cs) + (uint32_t)((crow * ldv + hh * dk + ccol) * 2);
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t pf[4];
        ldsm_x4_t(pbase + (uint32_t)(ks * 16 * 2), pf[0], pf[1], pf[2], pf[3]);
        float acc[NT][4];
#pragma unroll
        for (int n = 0; n < NT; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
#pragma unroll
        for (int np = 0; np < NT / 2; ++np) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4_t(cbase + (uint32_t)(np * 16 * 2), b0, b1, b2, b3);
          mma16816(acc[2 * np], pf, b0, b1);
          mma16816(acc[2 * np + 1], pf, b2, b3);
        }
        __nv_bfloat16* o0 = Vs + (size_t)(ks * 16 + r0) * ldv + hh * dk + kq;
        __nv_bfloat16* o1 = Vs + (size_t)(ks * 16 + r1) * ldv + hh * dk + kq;
#pragma unroll
        for (int n = 0; n < NT; ++n) {
          *reinterpret_cast<uint32_t*>(o0 + n * 8) = pack2(acc[n][0], acc[n][1]);
          *reinterpret_cast<uint32_t*>(o1 + n * 8) = pack2(acc[n][2], acc[n][3]);
        }
      }
    }
    __syncwarp();
  }
  if (!staged) {
    cp_async_wait_all();
    __syncthreads();
  }
  __syncthreads();

  // ---- coalesced stores: dXV rows (16 B) and dS^T rows
  {
    __nv_bfloat16* dVg = reinterpret_cast<__nv_bfloat16*>(a.dV) + f.vg.off(g);
    for (int k = warp; k < nk; k += 8) {
      __nv_bfloat16* drow = dVg + f.v.off(k);
      for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = *reinterpret_cast<const uint4*>(Vs + (size_t)k * ldv + c);
    }
    const int b = g / a.gdiv, gi = g % a.gdiv;
    const long long ldt = a.dsT_ld > 0 ? a.dsT_ld : R;
    __nv_bfloat16* dsT = reinterpret_cast<__nv_bfloat16*>(a.dsT) + (long long)b * a.TS * ldt;
    const bool v8 = (R % 8) == 0 && (ldt % 8) == 0;
    for (int k = warp; k < nk; k += 8) {
      __nv_bfloat16* drow = dsT + ((long long)gi * a.ts_g + (long long)k * a.ts_k) * ldt;
      if (v8) {
        for (int c = lane * 8; c < R; c += 256) *reinterpret_cast<uint4*>(drow + c) = *reinterpret_cast<const uint4*>(Ts + (size_t)k * ldr + c);
      } else {
        for (int c = lane; c < R; c += 32) drow[c] = Ts[(size_t)k * ldr + c];
      }
    }
  }
}

template <int KS, int NT>
int launch_s1(const Stage1BwdArgs& a, size_t smem, cudaStream_t st) {
  auto kern = stage1_bwd_mma_kernel<KS, NT>;
  static bool configured = false;
  if (!configured && smem > 48 * 1024) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  ProfScope ps("stage1_bwd_mma", st);
  BIST_LAUNCH(kern, (unsigned)a.f.G, 256, smem, st, a);
  BIST_LAUNCHED();
  return 0;
}

template <int KS>
int pick_s1(const Stage1BwdArgs& a, size_t smem, cudaStream_t st) {
  switch (a.f.dk / 8) {
    case 2: return launch_s1<KS, 2>(a, smem, st);
    case 4: return launch_s1<KS, 4>(a, smem, st);
    case 8: return launch_s1<KS, 8>(a, smem, st);
    default: return -1;
  }
}

}  // namespace

// *taken = false when the shape is left to the CUDA-core kernel (bwd.cu).  Writes dV and dsT only (a.ds is ignored).
int stage1_bwd_mma(const Stage1BwdArgs& a, cudaStream_t st, bool* taken) {
  *taken = false;
  static const bool disabled = getenv("BIST_DISABLE_TC") != nullptr;
  if (disabled) return 0;
  const SoftmaxPvArgs& f = a.f;
  const int d = f.h * f.dk, R = f.Lq * f.h;
  if (f.v_dtype != BIST_BF16 || a.ds_dtype != BIST_BF16 || a.ds != nullptr) return 0;
  if (f.Lq > 16 || f.nk > 128 || f.dk % 16 != 0 || f.dk > 64 || d % 8 != 0) return 0;
  if ((reinterpret_cast<uintptr_t>(f.V) & 15) || (reinterpret_cast<uintptr_t>(a.dV) & 15) || (reinterpret_cast<uintptr_t>(a.dctx) & 15) ||
      (reinterpret_cast<uintptr_t>(a.dsT) & 15))
    return 0;
  if ((f.v.so | f.v.si | f.vg.so | f.vg.si | f.og.so | f.og.si | f.o_l) % 8 != 0) return 0;
  const int ks = (f.nk + 15) / 16;
  const int ksT = ks <= 4 ? ks : 8;
  const size_t smem = ((size_t)(16 * ksT + 16) * (d + 8) + (size_t)8 * 16 * (16 * ksT + 8) + (size_t)16 * ksT * (((R + 7) & ~7) + 8)) * 2;
  if (smem > 200 * 1024) return 0;
  int rc = -1;
  if (ks == 1) rc = pick_s1<1>(a, smem, st);
  else if (ks == 2) rc = pick_s1<2>(a, smem, st);
  else if (ks == 3) rc = pick_s1<3>(a, smem, st);
  else if (ks == 4) rc = pick_s1<4>(a, smem, st);
  else rc = pick_s1<8>(a, smem, st);
  if (rc == -1) return 0;
  *taken = true;
  return rc;
}


// ------------------------------------------------------------------------------------------------ stage 2
namespace {

// grid B*Lq; block 256 (8 warps); h <= 8, d % 128 == 0.
// smem: Xs [16*KS][ldx] bf16 (x1 rows, later dx1) | QD [16][ldx] bf16 (rows 0..7 qf, 8..15 du) | S [16][lds] f32 | W2 [16][ldw] bf16 (rows 0..7 dS, 8..15 P)
template <int KS>
__global__ void __launch_bounds__(256) stage2_bwd_mma_kernel(Stage2BwdArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_trigger();
  pdl_wait();
  const int h = a.h, d = a.d, G = a.G, Gp = 16 * KS, ldx = d + 8, lds = Gp + 1, ldw = Gp + 8;
  __nv_bfloat16* Xs = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* QD = Xs + (size_t)Gp * ldx;
  float* S = reinterpret_cast<float*>(QD + (size_t)16 * ldx);
  __nv_bfloat16* W2 = reinterpret_cast<__nv_bfloat16*>(S + 16 * lds);
  const int q = blockIdx.x, b = q / a.Lq;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const __nv_bfloat16* x1 = reinterpret_cast<const __nv_bfloat16*>(a.x1) + (long long)q * G * d;
  const __nv_bfloat16* qf = reinterpret_cast<const __nv_bfloat16*>(a.qf) + (long long)q * h * d;
  const __nv_bfloat16* du = reinterpret_cast<const __nv_bfloat16*>(a.du) + (long long)q * h * d;
  const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
  for (int g = warp; g < Gp; g += 8) {
    __nv_bfloat16* drow = Xs + (size_t)g * ldx;
    if (g < G) for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), x1 + (long long)g * d + c);
    else for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = z4;
  }
  for (int r = warp; r < 16; r += 8) {
    __nv_bfloat16* drow = QD + (size_t)r * ldx;
    const int hh = r & 7;
    if (hh < h) {
      const __nv_bfloat16* src = (r < 8 ? qf : du) + (long long)hh * d;
      for (int c = lane * 8; c < d; c += 256) cp_async16(smem_addr(drow + c), src + c);
    } else {
      for (int c = lane * 8; c < d; c += 256) *reinterpret_cast<uint4*>(drow + c) = z4;
    }
  }
  for (int i = t; i < 16 * lds; i += 256) S[i] = 0.f;
  cp_async_wait_all();
  __syncthreads();

  const int gr = lane >> 2, kq = (lane & 3) * 2;
  // ---- A: S[0..7, g] = qf . x1[g], S[8..15, g] = du . x1[g]; warp w contracts columns [w*d/8, (w+1)*d/8)
  {
    float acc[2 * KS][4];
#pragma unroll
    for (int n = 0; n < 2 * KS; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
    const int arow = (lane & 7) + ((lane >> 3) & 1) * 8, acol = (lane >> 4) * 8;
    const int brow = (lane & 7) + (lane >> 4) * 8, bcol = ((lane >> 3) & 1) * 8;
    const int c_begin = warp * (d / 8);
    for (int c0 = c_begin; c0 < c_begin + d / 8; c0 += 16) {
      uint32_t af[4];
      ldsm_x4(smem_addr(QD + (size_t)arow * ldx + c0 + acol), af[0], af[1], af[2], af[3]);
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(smem_addr(Xs + (size_t)(ks * 16 + brow) * ldx + c0 + bcol), b0, b1, b2, b3);
        mma16816(acc[2 * ks], af, b0, b1);
        mma16816(acc[2 * ks + 1], af, b2, b3);
      }
    }
#pragma unroll
    for (int n = 0; n < 2 * KS; ++n) {
      atomicAdd(&S[gr * lds + n * 8 + kq], acc[n][0]);
      atomicAdd(&S[gr * lds + n * 8 + kq + 1], acc[n][1]);
      atomicAdd(&S[(gr + 8) * lds + n * 8 + kq], acc[n][2]);
      atomicAdd(&S[(gr + 8) * lds + n * 8 + kq + 1], acc[n][3]);
    }
  }
  __syncthreads();
  // ---- B: one warp per head: p = softmax_g(mask(S[hh])), dS = p o (dP - sum p dP)
  for (int hh = warp; hh < 8; hh += 8) {
    float* srow = S + hh * lds;
    const float* drow = S + (8 + hh) * lds;
    float mx = -INFINITY;
    for (int g = lane; g < G; g += 32) {
      float v = srow[g];
      if (a.mask && a.mask[(long long)b * G + g] == 0) v = kMaskFill;
      srow[g] = v;
      mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int g = lane; g < G; g += 32) {
      const float e = expf(srow[g] - mx);
      srow[g] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float dot = 0.f;
    const float dr = (a.drsum && hh < h) ? a.drsum[(long long)q * h + hh] : 0.f;
    const unsigned long long ibase = ((unsigned long long)q * h + hh) * G;
    for (int g = lane; g < G; g += 32) dot = fmaf(srow[g] * inv, (drow[g] + dr) * drop_mul(a.drop, ibase + g), dot);
    dot = warp_sum(dot);
    for (int g = lane; g < Gp; g += 32) {
      float p = 0.f, ds = 0.f;
      if (g < G && hh < h) {
        const float m = drop_mul(a.drop, ibase + g);
        p = srow[g] * inv;
        ds = p * ((drow[g] + dr) * m - dot);
        p *= m;                                    // p' multiplies du in dx1
      }
      W2[(size_t)hh * ldw + g] = __float2bfloat16_rn(ds);
      W2[(size_t)(8 + hh) * ldw + g] = __float2bfloat16_rn(p);
    }
  }
  __syncthreads();
  // ---- D: dqf[hh, col] = sum_g dS[hh, g] x1[g, col]; warp w owns 64-column slabs
  {
    const int arow = (lane & 7) + ((lane >> 3) & 1) * 8, acol = (lane >> 4) * 8;   // W2 [16][g] row-major
    const int xrow = (lane & 7) + ((lane >> 3) & 1) * 8, xcol = (lane >> 4) * 8;   // x1 via .trans: (g 0-7,c 0-7),(g 8-15,c 0-7),(g 0-7,c 8-15),(g 8-15,c 8-15)
    __nv_bfloat16* dqf = reinterpret_cast<__nv_bfloat16*>(a.dqf) + (long long)q * h * d;
    for (int cs = warp * 64; cs < d; cs += 512) {
      float acc[8][4];
#pragma unroll
      for (int n = 0; n < 8; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t af[4];
        ldsm_x4(smem_addr(W2 + (size_t)arow * ldw + ks * 16 + acol), af[0], af[1], af[2], af[3]);
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4_t(smem_addr(Xs + (size_t)(ks * 16 + xrow) * ldx + cs + np * 16 + xcol), b0, b1, b2, b3);
          mma16816(acc[2 * np], af, b0, b1);
          mma16816(acc[2 * np + 1], af, b2, b3);
        }
      }
      if (gr < h) {
#pragma unroll
        for (int n = 0; n < 8; ++n)
          *reinterpret_cast<uint32_t*>(dqf + (long long)gr * d + cs + n * 8 + kq) = pack2(acc[n][0], acc[n][1]);
      }
    }
  }
  __syncthreads();   // every warp is done reading x1 before it is overwritten with dx1
  // ---- C: dx1[g, col] = sum_hh dS[hh,g] qf[hh,col] + P[hh,g] du[hh,col] = (W2^T [g x 16]) . QD [16 x col]
  {
    const int prow = (lane & 7) + (lane >> 4) * 8, pcol = ((lane >> 3) & 1) * 8;    // W2^T via .trans from W2[c16][g]
    const int crow = (lane & 7) + ((lane >> 3) & 1) * 8, ccol = (lane >> 4) * 8;    // QD via .trans from QD[c16][col]
    for (int cs = warp * 64; cs < d; cs += 512) {
      uint32_t bf[4][4];
#pragma unroll
      for (int np = 0; np < 4; ++np)
        ldsm_x4_t(smem_addr(QD + (size_t)crow * ldx + cs + np * 16 + ccol), bf[np][0], bf[np][1], bf[np][2], bf[np][3]);
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t pf[4];
        ldsm_x4_t(smem_addr(W2 + (size_t)prow * ldw + ks * 16 + pcol), pf[0], pf[1], pf[2], pf[3]);
        float acc[8][4];
#pragma unroll
        for (int n = 0; n < 8; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          mma16816(acc[2 * np], pf, bf[np][0], bf[np][1]);
          mma16816(acc[2 * np + 1], pf, bf[np][2], bf[np][3]);
        }
        __nv_bfloat16* o0 = Xs + (size_t)(ks * 16 + gr) * ldx + cs + kq;
        __nv_bfloat16* o1 = o0 + (size_t)8 * ldx;
#pragma unroll
        for (int n = 0; n < 8; ++n) {
          *reinterpret_cast<uint32_t*>(o0 + n * 8) = pack2(acc[n][0], acc[n][1]);
          *reinterpret_cast<uint32_t*>(o1 + n * 8) = pack2(acc[n][2], acc[n][3]);
        }
      }
    }
  }
  __syncthreads();
  __nv_bfloat16* dx1 = reinterpret_cast<__nv_bfloat16*>(a.dx1) + (long long)q * G * d;
  for (int g = warp; g < G; g += 8)
    for (int c = lane * 8; c < d; c += 256)
      *reinterpret_cast<uint4*>(dx1 + (long long)g * d + c) = *reinterpret_cast<const uint4*>(Xs + (size_t)g * ldx + c);
}

template <int KS>
int launch_s2b(const Stage2BwdArgs& a, size_t smem, cudaStream_t st) {
  auto kern = stage2_bwd_mma_kernel<KS>;
  static bool configured = false;
  if (!configured && smem > 48 * 1024) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  ProfScope ps("stage2_bwd_mma", st);
  BIST_LAUNCH(kern, (unsigned)(a.B * a.Lq), 256, smem, st, a);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace

int stage2_bwd_mma(const Stage2BwdArgs& a, cudaStream_t st, bool* taken) {
  *taken = false;
  static const bool disabled = getenv("BIST_DISABLE_TC") != nullptr;
  if (disabled) return 0;
  if (a.dtype != BIST_BF16 || a.h > 8 || a.d % 128 != 0 || a.G > 128) return 0;
  const void* ptrs[5] = {a.qf, a.x1, a.du, a.dx1, a.dqf};
  for (const void* p : ptrs)
    if (reinterpret_cast<uintptr_t>(p) & 15) return 0;
  const int ks = (a.G + 15) / 16, ksT = ks <= 4 ? ks : 8, Gp = 16 * ksT;
  const size_t smem = (size_t)(Gp + 16) * (a.d + 8) * 2 + (size_t)16 * (Gp + 1) * 4 + (size_t)16 * (Gp + 8) * 2 + 16;
  if (smem > 200 * 1024) return 0;
  *taken = true;
  if (ks == 1) return launch_s2b<1>(a, smem, st);
  if (ks == 2) return launch_s2b<2>(a, smem, st);
  if (ks == 3) return launch_s2b<3>(a, smem, st);
  if (ks == 4) return launch_s2b<4>(a, smem, st);
  return launch_s2b<8>(a, smem, st);
}

}  // namespace bist
