// LayerNorm of /root/reference/model/modules.py:28-31:
//   y = a_2 * (x - mean) / (std + eps) + b_2,  std = sqrt(sum((x-mean)^2) / (d-1))  (unbiased),
//   eps added to the std (NOT nn.LayerNorm).  One warp per row, two-pass statistics in fp32.
#include "common.cuh"

namespace bist {
namespace {

template <typename TX, typename TY, int MAXV>  // MAXV float4 per lane cached in registers
__global__ void __launch_bounds__(256) ln_rows_vec(const TX* __restrict__ x, const float* __restrict__ a2,
                                                   const float* __restrict__ b2, TY* __restrict__ y,
                                                   long long rows, int d, float eps,
                                                   long long group_rows, long long pstride, long long x_gstride) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long long grp = row / group_rows;
  a2 += grp * pstride;
  b2 += grp * pstride;
  const TX* xr = x + grp * x_gstride + (row - grp * group_rows) * d;
  TY* yr = y + row * d;
  const int nv = d >> 2;  // float4 count
  float4 v[MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      v[i] = ld4(xr + c * 4);
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(s) / (float)d;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      float a = v[i].x - mean, b = v[i].y - mean, e = v[i].z - mean, f = v[i].w - mean;
      q += (a * a + b * b) + (e * e + f * f);
    }
  }
  const float sd = sqrtf(warp_sum(q) / (float)(d - 1));
  const float inv = 1.f / (sd + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      float4 g = *reinterpret_cast<const float4*>(a2 + c * 4);
      float4 o = *reinterpret_cast<const float4*>(b2 + c * 4);
      float4 r;
      r.x = g.x * (v[i].x - mean) * inv + o.x;
      r.y = g.y * (v[i].y - mean) * inv + o.y;
      r.z = g.z * (v[i].z - mean) * inv + o.z;
      r.w = g.w * (v[i].w - mean) * inv + o.w;
      st4(yr + c * 4, r);
    }
  }
}

template <typename TX, typename TY>
__global__ void __launch_bounds__(256) ln_rows_any(const TX* __restrict__ x, const float* __restrict__ a2,
                                                   const float* __restrict__ b2, TY* __restrict__ y,
                                                   long long rows, int d, float eps,
                                                   long long group_rows, long long pstride, long long x_gstride) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long long grp = row / group_rows;
  a2 += grp * pstride;
  b2 += grp * pstride;
  const TX* xr = x + grp * x_gstride + (row - grp * group_rows) * d;
  TY* yr = y + row * d;
  float s = 0.f;
  for (int c = lane; c < d; c += 32) s += to_f32(xr[c]);
  const float mean = warp_sum(s) / (float)d;
  float q = 0.f;
  for (int c = lane; c < d; c += 32) {
    float a = to_f32(xr[c]) - mean;
    q += a * a;
  }
  const float sd = sqrtf(warp_sum(q) / (float)(d - 1));
  const float inv = 1.f / (sd + eps);
  for (int c = lane; c < d; c += 32) yr[c] = from_f32<TY>(a2[c] * (to_f32(xr[c]) - mean) * inv + b2[c]);
}

// Two LayerNorms of the same fp32 rows with different gains (the stage-1 and stage-2 query norms of
// VidEncoderLayer4, encoder.py:121,130 / :148,161 both normalise the same x) plus a raw copy of x in the
// activation dtype, in one pass.  One warp per row; rows [k*group_rows, (k+1)*group_rows) use the parameter
// sets at +k*pstride.
template <typename TY, int MAXV>
__global__ void __launch_bounds__(256) ln2_rows_vec(const float* __restrict__ x, const float* __restrict__ a2a,
                                                    const float* __restrict__ b2a, const float* __restrict__ a2b,
                                                    const float* __restrict__ b2b, TY* __restrict__ ya,
                                                    TY* __restrict__ yb, TY* __restrict__ xc, long long rows, int d,
                                                    float eps, long long group_rows, long long pstride) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long long po = (row / group_rows) * pstride;
  const float* xr = x + row * d;
  const int nv = d >> 2;
  float4 v[MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      v[i] = ld4(xr + c * 4);
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(s) / (float)d;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      float a = v[i].x - mean, b = v[i].y - mean, e = v[i].z - mean, f = v[i].w - mean;
      q += (a * a + b * b) + (e * e + f * f);
    }
  }
  const float inv = 1.f / (sqrtf(warp_sum(q) / (float)(d - 1)) + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    int c = lane + i * 32;
    if (c < nv) {
      const float4 n = make_float4((v[i].x - mean) * inv, (v[i].y - mean) * inv, (v[i].z - mean) * inv, (v[i].w - mean) * inv);
      float4 g = *reinterpret_cast<const float4*>(a2a + po + c * 4), o = *reinterpret_cast<const float4*>(b2a + po + c * 4);
      st4(ya + row * d + c * 4, make_float4(g.x * n.x + o.x, g.y * n.y + o.y, g.z * n.z + o.z, g.w * n.w + o.w));
      g = *reinterpret_cast<const float4*>(a2b + po + c * 4); o = *reinterpret_cast<const float4*>(b2b + po + c * 4);
      st4(yb + row * d + c * 4, make_float4(g.x * n.x + o.x, g.y * n.y + o.y, g.z * n.z + o.z, g.w * n.w + o.w));
      if (xc) st4(xc + row * d + c * 4, v[i]);
    }
  }
}

template <typename TX, typename TY>
int launch(const void* x, const float* a2, const float* b2, void* y, int64_t rows, int64_t d, float eps,
           cudaStream_t st, long long group_rows, long long pstride, long long x_gstride) {
  const int wpb = 8;
  dim3 grid((unsigned)((rows + wpb - 1) / wpb));
  const TX* xp = reinterpret_cast<const TX*>(x);
  TY* yp = reinterpret_cast<TY*>(y);
  bool aligned = (d % 4 == 0) && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) &&
                 ((uintptr_t)a2 % 16 == 0) && ((uintptr_t)b2 % 16 == 0);
  ProfScope ps("layer_norm", st);
  if (aligned && d <= 512)
    BIST_LAUNCH((ln_rows_vec<TX, TY, 4>), grid, wpb * 32, 0, st, xp, a2, b2, yp, rows, (int)d, eps, group_rows, pstride, x_gstride);
  else if (aligned && d <= 2048)
    BIST_LAUNCH((ln_rows_vec<TX, TY, 16>), grid, wpb * 32, 0, st, xp, a2, b2, yp, rows, (int)d, eps, group_rows, pstride, x_gstride);
  else
    BIST_LAUNCH((ln_rows_any<TX, TY>), grid, wpb * 32, 0, st, xp, a2, b2, yp, rows, (int)d, eps, group_rows, pstride, x_gstride);
  BIST_LAUNCHED();
  return 0;
}

}  // namespace

int layer_norm(const void* x, const float* a2, const float* b2, void* y, int64_t rows, int64_t d, float eps,
               int x_dtype, int y_dtype, cudaStream_t st, int64_t group_rows, int64_t pstride, int64_t x_gstride, bool x_strided) {
  if (group_rows <= 0) { group_rows = rows; pstride = 0; }
  if (!x_strided) x_gstride = group_rows * d;
  BIST_CHECK_ARG(x_gstride % 4 == 0, "ln: group stride of x must be a multiple of 4 elements");
  BIST_CHECK_ARG(pstride % 4 == 0, "ln: parameter stride must be a multiple of 4");
  BIST_CHECK_ARG(x && a2 && b2 && y, "ln: null pointer");
  BIST_CHECK_ARG(rows > 0 && d > 1 && d <= (1 << 20), "ln: bad shape rows=%lld d=%lld", (long long)rows,
                 (long long)d);
  BIST_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(y_dtype), "ln: bad dtype");
  if (x_dtype == BIST_F32 && y_dtype == BIST_F32) return launch<float, float>(x, a2, b2, y, rows, d, eps, st, group_rows, pstride, x_gstride);
  if (x_dtype == BIST_F32 && y_dtype == BIST_BF16)
    return launch<float, __nv_bfloat16>(x, a2, b2, y, rows, d, eps, st, group_rows, pstride, x_gstride);
  if (x_dtype == BIST_BF16 && y_dtype == BIST_F32)
    return launch<__nv_bfloat16, float>(x, a2, b2, y, rows, d, eps, st, group_rows, pstride, x_gstride);
  return launch<__nv_bfloat16, __nv_bfloat16>(x, a2, b2, y, rows, d, eps, st, group_rows, pstride, x_gstride);
}

// Fused pair of LayerNorms + raw copy (see ln2_rows_vec).  x fp32 [rows, d]; ya, yb, xc of y_dtype; d % 4 == 0, d <= 2048.
int layer_norm2(const float* x, const float* a2a, const float* b2a, const float* a2b, const float* b2b, void* ya, void* yb,
                void* xc, int64_t rows, int64_t d, float eps, int y_dtype, cudaStream_t st, int64_t group_rows, int64_t pstride) {
  BIST_CHECK_ARG(x && a2a && b2a && a2b && b2b && ya && yb, "ln2: null pointer");
  BIST_CHECK_ARG(rows > 0 && d > 1 && d % 4 == 0 && d <= 2048 && pstride % 4 == 0 && dtype_ok(y_dtype), "ln2: bad argument");
  if (group_rows <= 0) { group_rows = rows; pstride = 0; }
  const int wpb = 8;
  dim3 grid((unsigned)((rows + wpb - 1) / wpb));
  ProfScope ps("layer_norm", st);
  if (y_dtype == BIST_BF16) {
    auto Y = [](void* p) { return reinterpret_cast<__nv_bfloat16*>(p); };
    if (d <= 512) BIST_LAUNCH((ln2_rows_vec<__nv_bfloat16, 4>), grid, wpb * 32, 0, st, x, a2a, b2a, a2b, b2b, Y(ya), Y(yb), Y(xc), rows, (int)d, eps, group_rows, pstride);
    else BIST_LAUNCH((ln2_rows_vec<__nv_bfloat16, 16>), grid, wpb * 32, 0, st, x, a2a, b2a, a2b, b2b, Y(ya), Y(yb), Y(xc), rows, (int)d, eps, group_rows, pstride);
  } else {
    auto Y = [](void* p) { return reinterpret_cast<float*>(p); };
    if (d <= 512) BIST_LAUNCH((ln2_rows_vec<float, 4>), grid, wpb * 32, 0, st, x, a2a, b2a, a2b, b2b, Y(ya), Y(yb), Y(xc), rows, (int)d, eps, group_rows, pstride);
    else BIST_LAUNCH((ln2_rows_vec<float, 16>), grid, wpb * 32, 0, st, x, a2a, b2a, a2b, b2b, Y(ya), Y(yb), Y(xc), rows, (int)d, eps, group_rows, pstride);
  }
  BIST_LAUNCHED();
  return 0;
}

}  // namespace bist
