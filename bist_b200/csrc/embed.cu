// Token embedding + positional encoding + dropout in one pass (/root/reference/model/modules.py:115-141:
// Embeddings.forward = lut(x) * sqrt(d_model); PositionalEncoding.forward = dropout(x + pe[:, :L])), and its backward
// (a scatter-add of the masked, scaled gradient rows into the embedding table).
#include "common.cuh"

namespace bist {
namespace {

template <typename T>
__global__ void __launch_bounds__(256) embed_fwd_kernel(const int64_t* __restrict__ idx, const float* __restrict__ lut,
                                                        const float* __restrict__ pe, T* __restrict__ out, int64_t rows, int L,
                                                        int d, float scale, DropCfg drop) {
  pdl_trigger();
  pdl_wait();
  const int64_t total = rows * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / d;
    const int c = (int)(i - r * d);
    const float v = fmaf(lut[idx[r] * d + c], scale, pe[(int64_t)(r % L) * d + c]);
    out[i] = from_f32<T>(v * drop_mul(drop, (unsigned long long)i));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) embed_bwd_kernel(const int64_t* __restrict__ idx, const T* __restrict__ dy,
                                                        float* __restrict__ dlut, int64_t rows, int d, float scale, DropCfg drop) {
  pdl_trigger();
  pdl_wait();
  const int64_t total = rows * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / d;
    const int c = (int)(i - r * d);
    const float g = to_f32(dy[i]) * scale * drop_mul(drop, (unsigned long long)i);
    if (g != 0.f) atomicAdd(&dlut[idx[r] * d + c], g);
  }
}

unsigned grid_for(int64_t n) { return (unsigned)std::min<int64_t>((n + 255) / 256, 148 * 8); }

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

int bist_embed_fwd(const int64_t* idx, const float* lut, const float* pe, void* out, int64_t rows, int64_t L, int64_t d,
                   int64_t vocab, float p_drop, uint64_t seed, uint32_t site, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  (void)vocab;
  BIST_CHECK_ARG(idx && lut && pe && out && rows > 0 && L > 0 && d > 0 && rows % L == 0, "embed: bad argument");
  BIST_CHECK_ARG(dtype_ok(out_dtype) && p_drop >= 0.f && p_drop < 1.f, "embed: bad dtype or dropout probability");
  const DropCfg drop = make_drop(p_drop, seed, site);
  const float scale = sqrtf((float)d);
  ProfScope ps("embed", st);
  if (out_dtype == BIST_BF16)
    BIST_LAUNCH(embed_fwd_kernel<__nv_bfloat16>, grid_for(rows * d), 256, 0, st, idx, lut, pe, (__nv_bfloat16*)out, rows, (int)L, (int)d, scale, drop);
  else
    BIST_LAUNCH(embed_fwd_kernel<float>, grid_for(rows * d), 256, 0, st, idx, lut, pe, (float*)out, rows, (int)L, (int)d, scale, drop);
  BIST_LAUNCHED();
  return 0;
}

int bist_embed_bwd(const int64_t* idx, const void* dy, float* dlut, int64_t rows, int64_t d, float p_drop, uint64_t seed,
                   uint32_t site, int dy_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(idx && dy && dlut && rows > 0 && d > 0, "embed_bwd: bad argument");
  BIST_CHECK_ARG(dtype_ok(dy_dtype) && p_drop >= 0.f && p_drop < 1.f, "embed_bwd: bad dtype or dropout probability");
  const DropCfg drop = make_drop(p_drop, seed, site);
  const float scale = sqrtf((float)d);
  ProfScope ps("embed_bwd", st);
  if (dy_dtype == BIST_BF16)
    BIST_LAUNCH(embed_bwd_kernel<__nv_bfloat16>, grid_for(rows * d), 256, 0, st, idx, (const __nv_bfloat16*)dy, dlut, rows, (int)d, scale, drop);
  else
    BIST_LAUNCH(embed_bwd_kernel<float>, grid_for(rows * d), 256, 0, st, idx, (const float*)dy, dlut, rows, (int)d, scale, drop);
  BIST_LAUNCHED();
  return 0;
}

}  // extern "C"
