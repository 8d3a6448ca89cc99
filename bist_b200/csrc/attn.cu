// Attention cores of the BiST video path (CUDA-core, fp32 math).
//   softmax_pv : masked softmax over precomputed scores + weighted sum of projected values
//                (attention(), /root/reference/model/modules.py:54-64, after the score GEMM)
//   stage2_attn: stage 2 of temporal2spatial / spatial2temporal with the key and value
//                projections folded to the query side (encoder.py:125-132, 152-161)
#include "common.cuh"

namespace bist {
namespace {

constexpr float kMaskFill = -1e9f;  // modules.py:60

// ---------------------------------------------------------------------------------------------
// grid (G, ceil(Lq/LT)); block 256.  smem: P[h][nk][LTP] floats (l fastest).
template <typename TV, typename TC>
__global__ void __launch_bounds__(256) softmax_pv_kernel(SoftmaxPvArgs a, int LT, int LTP) {
  extern __shared__ __align__(16) float Ps[];
  pdl_trigger();
  pdl_wait();
  const int g = blockIdx.x;
  const int l0 = blockIdx.y * LT;
  const int nl = min(LT, a.Lq - l0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int h = a.h, nk = a.nk;

  // phase 1: softmax rows -> smem
  const float* Sg = a.S + a.sg.off(g);
  const uint8_t* Mg = a.mask ? a.mask + (long long)((g / a.mask_gdiv) % a.mask_gmod) * a.m_b : nullptr;
  for (int r = warp; r < nl * h; r += nwarp) {
    const int li = r / h, hh = r % h;
    const int l = l0 + li;
    const float* srow = Sg + (long long)l * a.s_l + (long long)hh * a.s_h;
    const uint8_t* mrow = Mg ? Mg + (long long)l * a.m_l : nullptr;
    float mx = -INFINITY;
    for (int k = lane; k < nk; k += 32) {
      float s = srow[(long long)k * a.s_k];
      if (mrow && mrow[k] == 0) s = kMaskFill;
      Ps[(hh * nk + k) * LTP + li] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < nk; k += 32) {
      float e = expf(Ps[(hh * nk + k) * LTP + li] - mx);
      Ps[(hh * nk + k) * LTP + li] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int k = lane; k < nk; k += 32) {
      float p = Ps[(hh * nk + k) * LTP + li] * inv;
      p *= drop_mul(a.drop, (((unsigned long long)g * h + hh) * a.Lq + l) * nk + k);
      Ps[(hh * nk + k) * LTP + li] = p;
      if (a.p_out) a.p_out[(((long long)g * h + hh) * a.Lq + l) * nk + k] = p;
    }
  }
  // zero the padding rows so that the float4 reads below see zeros
  for (int i = threadIdx.x; i < h * nk * (LTP - nl); i += blockDim.x) {
    int hk = i / (LTP - nl), li = nl + i % (LTP - nl);
    Ps[hk * LTP + li] = 0.f;
  }
  __syncthreads();

  // phase 2: ctx[l, c] = sum_k P[l, head(c), k] * V[k, c]
  const int d = h * a.dk;
  const TV* Vg = reinterpret_cast<const TV*>(a.V) + a.vg.off(g);
  TC* Cg = reinterpret_cast<TC*>(a.ctx) + a.og.off(g) + (long long)l0 * a.o_l;
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const int hh = c / a.dk;
    for (int lb = 0; lb < nl; lb += 16) {
      float acc[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = 0.f;
      const int cnt = min(16, LTP - lb);  // multiple of 4
      for (int k = 0; k < nk; ++k) {
        const float v = to_f32(Vg[a.v.off(k) + c]);
        const float* pp = &Ps[(hh * nk + k) * LTP + lb];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (q * 4 < cnt) {
            float4 p4 = *reinterpret_cast<const float4*>(pp + q * 4);
            acc[q * 4 + 0] = fmaf(p4.x, v, acc[q * 4 + 0]);
            acc[q * 4 + 1] = fmaf(p4.y, v, acc[q * 4 + 1]);
            acc[q * 4 + 2] = fmaf(p4.z, v, acc[q * 4 + 2]);
            acc[q * 4 + 3] = fmaf(p4.w, v, acc[q * 4 + 3]);
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 16; ++i)
        if (lb + i < nl) Cg[(long long)(lb + i) * a.o_l + c] = from_f32<TC>(acc[i]);
    }
  }
}

template <typename TV, typename TC>
int launch_softmax_pv(const SoftmaxPvArgs& a, cudaStream_t st) {
  // rows per CTA so that P fits in 64 KB
  const long long per_l = (long long)a.h * a.nk * 4;
  int LT = (int)min((long long)a.Lq, max(1LL, (64 * 1024) / per_l));
  if (LT >= 4) LT &= ~3;
  const int LTP = (LT + 3) & ~3;
  const size_t smem = (size_t)a.h * a.nk * LTP * 4;
  if (smem > 200 * 1024) {
    set_error("softmax_pv: h*Lk = %d*%d too large for one CTA", a.h, a.nk);
    return BIST_ERR_UNSUPPORTED;
  }
  auto kern = softmax_pv_kernel<TV, TC>;
  if (smem > 48 * 1024)
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  dim3 grid(a.G, (a.Lq + LT - 1) / LT);
  ProfScope ps("softmax_pv", st);
  BIST_LAUNCH(kern, grid, 256, smem, st, a, LT, LTP);
  BIST_LAUNCHED();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// grid B*Lq; block 256.  smem: qf[h][d] + sc[h][G].
template <typename TX, typename TO, int HC>
__global__ void __launch_bounds__(256) stage2_kernel(const float* __restrict__ qf, const TX* __restrict__ x1,
                                                     const uint8_t* __restrict__ mask, TO* __restrict__ ctxx,
                                                     int B, int G, int Lq, int h, int d, DropCfg drop, float* __restrict__ rsum) {
  extern __shared__ __align__(16) float sm[];
  pdl_trigger();
  pdl_wait();
  float* q = sm;           // [h][d]
  float* sc = sm + h * d;  // [h][G]
  const int b = blockIdx.x / Lq, l = blockIdx.x % Lq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float* qrow = qf + (long long)blockIdx.x * h * d;
  for (int i = threadIdx.x; i < h * d; i += blockDim.x) q[i] = qrow[i];
  __syncthreads();

  // scores: one warp per key g
  for (int g = warp; g < G; g += nwarp) {
    const TX* xr = x1 + (((long long)b * Lq + l) * G + g) * d;
    for (int h0 = 0; h0 < h; h0 += HC) {
      float part[HC];
#pragma unroll
      for (int i = 0; i < HC; ++i) part[i] = 0.f;
      for (int c = lane; c < d; c += 32) {
        const float xv = to_f32(xr[c]);
#pragma unroll
        for (int i = 0; i < HC; ++i)
          if (h0 + i < h) part[i] = fmaf(q[(h0 + i) * d + c], xv, part[i]);
      }
#pragma unroll
      for (int i = 0; i < HC; ++i) {
        float s = warp_sum(part[i]);
        if (lane == 0 && h0 + i < h) sc[(h0 + i) * G + g] = (mask && mask[(long long)b * G + g] == 0) ? kMaskFill : s;
      }
    }
  }
  __syncthreads();
  // softmax over g: one warp per head
  for (int hh = warp; hh < h; hh += nwarp) {
    float mx = -INFINITY;
    for (int g = lane; g < G; g += 32) mx = fmaxf(mx, sc[hh * G + g]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int g = lane; g < G; g += 32) {
      float e = expf(sc[hh * G + g] - mx);
      sc[hh * G + g] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float rs = 0.f;
    for (int g = lane; g < G; g += 32) {
      const float pv = sc[hh * G + g] * inv * drop_mul(drop, ((unsigned long long)blockIdx.x * h + hh) * G + g);
      sc[hh * G + g] = pv;
      rs += pv;
    }
    if (rsum) {
      rs = warp_sum(rs);
      if (lane == 0) rsum[(long long)blockIdx.x * h + hh] = rs;
    }
  }
  __syncthreads();
  // weighted sums: thread per column
  TO* out = ctxx + (long long)blockIdx.x * h * d;
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    for (int h0 = 0; h0 < h; h0 += HC) {
      float acc[HC];
#pragma unroll
      for (int i = 0; i < HC; ++i) acc[i] = 0.f;
      for (int g = 0; g < G; ++g) {
        const float xv = to_f32(x1[(((long long)b * Lq + l) * G + g) * d + c]);
#pragma unroll
        for (int i = 0; i < HC; ++i)
          if (h0 + i < h) acc[i] = fmaf(sc[(h0 + i) * G + g], xv, acc[i]);
      }
#pragma unroll
      for (int i = 0; i < HC; ++i)
        if (h0 + i < h) out[(long long)(h0 + i) * d + c] = from_f32<TO>(acc[i]);
    }
  }
}

template <typename TD>
__global__ void pack_kernel(const float* __restrict__ src, TD* __restrict__ dst, int rows, int cols,
                            int transpose, float scale) {
  pdl_trigger();
  pdl_wait();
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)rows * cols) return;
  int r = (int)(i / cols), c = (int)(i % cols);
  float v = transpose ? src[(long long)c * rows + r] : src[i];
  dst[i] = from_f32<TD>(v * scale);
}

struct PackTable {
  PackItem it[kMaxPackItems];
  int n;
};

// grid = total 64x64 destination tiles over all items; block 256.  Both directions of the copy are coalesced (transposes go
// through a padded shared-memory tile) and, where the shapes allow, 16 bytes wide on the fp32 side and 8 bytes on the bf16 side:
// the weights are re-packed after every optimizer step (68 M parameters into two bf16 orientations), so this is an HBM stream.
constexpr int PK_T = 64;
template <typename TD>
__device__ __forceinline__ void pack_store4(TD* base, float* fbase, int to_f32, long long o, float a, float b, float c, float d) {
  if (to_f32) *reinterpret_cast<float4*>(fbase + o) = make_float4(a, b, c, d);
  else st4(base + o, make_float4(a, b, c, d));
}
template <typename TD>
__global__ void __launch_bounds__(256) pack_many_kernel(const __grid_constant__ PackTable tab, TD* __restrict__ base, float* __restrict__ fbase) {
  __shared__ float tile[PK_T][PK_T + 1];
  pdl_trigger();
  pdl_wait();
  int lo = 0, hi = tab.n - 1;   // item of this tile (tile_begin is increasing)
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if ((int)blockIdx.x >= tab.it[mid].tile_begin) lo = mid; else hi = mid - 1;
  }
  const PackItem& p = tab.it[lo];
  const int t = blockIdx.x - p.tile_begin;
  const int tiles_c = (p.cols + PK_T - 1) / PK_T;
  const int r0 = (t / tiles_c) * PK_T, c0 = (t % tiles_c) * PK_T;
  const int q = threadIdx.x & 15, rr = threadIdx.x >> 4;   // 16 threads x 4 elements per 64-wide row, 16 rows per pass
  const TD* bchk = base + p.dst_off;
  const float* fchk = fbase + p.dst_off;
  const bool dst_vec = p.cols % 4 == 0 && p.dst_off % 4 == 0 &&
                       ((p.to_f32 ? reinterpret_cast<uintptr_t>(fchk) % 16 : reinterpret_cast<uintptr_t>(bchk) % (4 * sizeof(TD))) == 0);
  if (p.transpose) {  // src is [cols, rows]: read rows c0.. of src (contiguous in r), write rows r0.. of dst (contiguous in c)
    const bool src_vec = p.rows % 4 == 0 && reinterpret_cast<uintptr_t>(p.src) % 16 == 0;
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) {
      const int j = ps * 16 + rr, c = c0 + j, r = r0 + 4 * q;
      if (c < p.cols) {
        if (src_vec && r + 3 < p.rows) {
          const float4 v = *reinterpret_cast<const float4*>(p.src + (long long)c * p.rows + r);
          tile[j][4 * q] = v.x; tile[j][4 * q + 1] = v.y; tile[j][4 * q + 2] = v.z; tile[j][4 * q + 3] = v.w;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) if (r + e < p.rows) tile[j][4 * q + e] = p.src[(long long)c * p.rows + r + e];
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) {
      const int i = ps * 16 + rr, r = r0 + i, c = c0 + 4 * q;
      if (r < p.rows) {
        const long long o = p.dst_off + (long long)r * p.cols + c;
        if (dst_vec && c + 3 < p.cols) {
          pack_store4<TD>(base, fbase, p.to_f32, o, tile[4 * q][i] * p.scale, tile[4 * q + 1][i] * p.scale, tile[4 * q + 2][i] * p.scale,
                          tile[4 * q + 3][i] * p.scale);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (c + e < p.cols) {
              const float v = tile[4 * q + e][i] * p.scale;
              if (p.to_f32) fbase[o + e] = v; else base[o + e] = from_f32<TD>(v);
            }
        }
      }
    }
  } else {
    const bool src_vec = p.cols % 4 == 0 && reinterpret_cast<uintptr_t>(p.src) % 16 == 0;
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) {
      const int r = r0 + ps * 16 + rr, c = c0 + 4 * q;
      if (r < p.rows) {
        const long long si = (long long)r * p.cols + c, o = p.dst_off + si;
        if (src_vec && dst_vec && c + 3 < p.cols) {
          const float4 v = *reinterpret_cast<const float4*>(p.src + si);
          pack_store4<TD>(base, fbase, p.to_f32, o, v.x * p.scale, v.y * p.scale, v.z * p.scale, v.w * p.scale);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (c + e < p.cols) {
              const float v = p.src[si + e] * p.scale;
              if (p.to_f32) fbase[o + e] = v; else base[o + e] = from_f32<TD>(v);
            }
        }
      }
    }
  }
}

}  // namespace

int pack_many(const PackItem* items, int n, void* base, float* f32_base, int dtype, cudaStream_t st) {
  BIST_CHECK_ARG(items && n > 0 && n <= kMaxPackItems && base && f32_base, "pack_many: bad argument");
  PackTable tab;
  tab.n = n;
  int tiles = 0;
  for (int i = 0; i < n; ++i) {
    tab.it[i] = items[i];
    BIST_CHECK_ARG(items[i].src && items[i].rows > 0 && items[i].cols > 0, "pack_many: bad item");
    tab.it[i].tile_begin = tiles;
    tiles += ((items[i].rows + PK_T - 1) / PK_T) * ((items[i].cols + PK_T - 1) / PK_T);
  }
  ProfScope ps("pack", st);
  if (dtype == BIST_BF16) BIST_LAUNCH(pack_many_kernel<__nv_bfloat16>, dim3((unsigned)tiles), 256, 0, st, tab, (__nv_bfloat16*)base, f32_base);
  else BIST_LAUNCH(pack_many_kernel<float>, dim3((unsigned)tiles), 256, 0, st, tab, (float*)base, f32_base);
  BIST_LAUNCHED();
  return 0;
}

int softmax_pv(const SoftmaxPvArgs& a, cudaStream_t st) {
  BIST_CHECK_ARG(a.S && a.V && a.ctx, "softmax_pv: null pointer");
  BIST_CHECK_ARG(a.G > 0 && a.Lq > 0 && a.h > 0 && a.dk > 0 && a.nk > 0, "softmax_pv: bad shape");
  {
    bool taken = false;
    int rc = softmax_pv_mma(a, st, &taken);
    if (taken) return rc;
  }
  if (a.v_dtype == BIST_F32 && a.ctx_dtype == BIST_F32) return launch_softmax_pv<float, float>(a, st);
  if (a.v_dtype == BIST_BF16 && a.ctx_dtype == BIST_BF16)
    return launch_softmax_pv<__nv_bfloat16, __nv_bfloat16>(a, st);
  if (a.v_dtype == BIST_BF16 && a.ctx_dtype == BIST_F32) return launch_softmax_pv<__nv_bfloat16, float>(a, st);
  return launch_softmax_pv<float, __nv_bfloat16>(a, st);
}

template <typename TX, typename TO>
int launch_stage2(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int h, int d,
                  size_t smem, cudaStream_t st, const Stage2Extra* ex) {
  auto kern = stage2_kernel<TX, TO, 8>;
  if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  ProfScope ps("stage2_attn", st);
  BIST_LAUNCH(kern, dim3((unsigned)(B * Lq)), 256, smem, st, qf, reinterpret_cast<const TX*>(x1), mask, reinterpret_cast<TO*>(ctxx), B, G,
                                                   Lq, h, d, ex ? ex->drop : DropCfg(), ex ? ex->rsum : nullptr);
  BIST_LAUNCHED();
  return 0;
}

int stage2_attn(const float* qf, const void* x1, const uint8_t* mask, void* ctxx, int B, int G, int Lq, int h, int d,
                int x1_dtype, int ctxx_dtype, cudaStream_t st, const Stage2Extra* ex) {
  BIST_CHECK_ARG(qf && x1 && ctxx, "stage2: null pointer");
  BIST_CHECK_ARG(B > 0 && G > 0 && Lq > 0 && h > 0 && d > 0, "stage2: bad shape");
  if (!ex) {  // the training extras (dropout, rsum) live in the plain kernel below
    bool taken = false;
    int rc = stage2_v2(qf, x1, mask, ctxx, B, G, Lq, h, d, x1_dtype, ctxx_dtype, st, &taken);
    if (taken) return rc;
  }
  const size_t smem = ((size_t)h * d + (size_t)h * G) * 4;
  if (smem > 200 * 1024) {
    set_error("stage2: h*(d+G) too large");
    return BIST_ERR_UNSUPPORTED;
  }
  const bool xb = x1_dtype == BIST_BF16, ob = ctxx_dtype == BIST_BF16;
  if (xb && ob) return launch_stage2<__nv_bfloat16, __nv_bfloat16>(qf, x1, mask, ctxx, B, G, Lq, h, d, smem, st, ex);
  if (xb) return launch_stage2<__nv_bfloat16, float>(qf, x1, mask, ctxx, B, G, Lq, h, d, smem, st, ex);
  if (ob) return launch_stage2<float, __nv_bfloat16>(qf, x1, mask, ctxx, B, G, Lq, h, d, smem, st, ex);
  return launch_stage2<float, float>(qf, x1, mask, ctxx, B, G, Lq, h, d, smem, st, ex);
}

int pack_matrix(const float* src, void* dst, int rows, int cols, int transpose, float scale, int dst_dtype,
                cudaStream_t st) {
  BIST_CHECK_ARG(src && dst && rows > 0 && cols > 0, "pack: bad argument");
  long long n = (long long)rows * cols;
  unsigned grid = (unsigned)((n + 255) / 256);
  ProfScope ps("pack", st);
  if (dst_dtype == BIST_BF16)
    BIST_LAUNCH((pack_kernel<__nv_bfloat16>), grid, 256, 0, st, src, reinterpret_cast<__nv_bfloat16*>(dst), rows, cols, transpose, scale);
  else
    BIST_LAUNCH((pack_kernel<float>), grid, 256, 0, st, src, reinterpret_cast<float*>(dst), rows, cols, transpose, scale);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist
