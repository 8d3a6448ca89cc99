// Adam step of the training loop (the reference trains with torch.optim.Adam behind NoamOpt, /root/reference/model/optimize.py:9-34,
// train.py) for ALL parameters in one launch: a device table of {p, g, m, v, n} rows and a chunk table {row, first element}; a CTA
// takes one chunk of 16 K elements with 16-byte accesses.  Same arithmetic as torch.optim.Adam (no weight decay, no amsgrad):
//   m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;  p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The stock fused implementation needs ~25 launches and 0.78 ms for the 68 M parameters of config 3 (nothing can overlap it: the
// next forward needs the new weights); this one streams the four arrays once at HBM rate.
#include "common.cuh"

namespace bist {
namespace {

struct AdamRow { float* p; const float* g; float* m; float* v; long long n; };
constexpr int kAdamChunk = 16384;

__global__ void __launch_bounds__(256) adam_kernel(const AdamRow* __restrict__ rows, const int2* __restrict__ chunks, float step_size,
                                                   float b1, float b2, float eps, float inv_sqrt_bc2) {
  pdl_trigger();
  pdl_wait();
  const int2 ck = chunks[blockIdx.x];
  const AdamRow r = rows[ck.x];
  const long long e0 = (long long)ck.y * kAdamChunk;
  const long long e1 = e0 + kAdamChunk < r.n ? e0 + kAdamChunk : r.n;
  const bool vec = ((reinterpret_cast<uintptr_t>(r.p) | reinterpret_cast<uintptr_t>(r.g) | reinterpret_cast<uintptr_t>(r.m) |
                     reinterpret_cast<uintptr_t>(r.v)) & 15) == 0;
  const float c1 = 1.f - b1, c2 = 1.f - b2;
  auto upd = [&](float& p, float g, float& m, float& v) {
    m = fmaf(b1, m, c1 * g);
    v = fmaf(b2, v, c2 * g * g);
    p -= step_size * m / (sqrtf(v) * inv_sqrt_bc2 + eps);
  };
  if (vec) {
    const long long n4 = (e1 - e0) >> 2;
    for (long long i = threadIdx.x; i < n4; i += 256) {
      const long long o = e0 + 4 * i;
      float4 p = *reinterpret_cast<const float4*>(r.p + o), m = *reinterpret_cast<const float4*>(r.m + o), v = *reinterpret_cast<const float4*>(r.v + o);
      const float4 g = *reinterpret_cast<const float4*>(r.g + o);
      upd(p.x, g.x, m.x, v.x); upd(p.y, g.y, m.y, v.y); upd(p.z, g.z, m.z, v.z); upd(p.w, g.w, m.w, v.w);
      *reinterpret_cast<float4*>(r.p + o) = p; *reinterpret_cast<float4*>(r.m + o) = m; *reinterpret_cast<float4*>(r.v + o) = v;
    }
    for (long long o = e0 + 4 * n4 + threadIdx.x; o < e1; o += 256) { float p = r.p[o], m = r.m[o], v = r.v[o]; upd(p, r.g[o], m, v); r.p[o] = p; r.m[o] = m; r.v[o] = v; }
  } else {
    for (long long o = e0 + threadIdx.x; o < e1; o += 256) { float p = r.p[o], m = r.m[o], v = r.v[o]; upd(p, r.g[o], m, v); r.p[o] = p; r.m[o] = m; r.v[o] = v; }
  }
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

int bist_adam_chunk_elems(void) { return kAdamChunk; }

// rows: device array of n_rows {p, g, m, v (device pointers, fp32), n (int64)} records (40 bytes each); chunks: device array of
// n_chunks {row, chunk index} int pairs covering every row in pieces of bist_adam_chunk_elems() elements.
int bist_adam_step(const void* rows, const void* chunks, int n_chunks, float lr, float beta1, float beta2, float eps, int64_t step,
                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(rows && chunks && n_chunks > 0 && step >= 1 && lr >= 0.f && beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f,
                 "adam: bad argument");
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  ProfScope ps("adam", st);
  BIST_LAUNCH(adam_kernel, (unsigned)n_chunks, 256, 0, st, reinterpret_cast<const AdamRow*>(rows), reinterpret_cast<const int2*>(chunks),
              (float)((double)lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)));
  BIST_LAUNCHED();
  return 0;
}

}  // extern "C"
