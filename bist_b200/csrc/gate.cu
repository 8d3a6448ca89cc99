// Softmax gate of MultimodalDecoder8 (/root/reference/model/decoder.py:150-181): after the gate Linear has produced
// K logits per query position, encoded_ft[r,:] = sum_k softmax(logits[r,:])_k * src_k[r,:] over K feature streams
// (temporal_ft, spatial_ft, cap_ft, audio_ft).  One warp per row; rows are tiny ([B*Lq, d]), the kernels are
// launch-latency bound and exist so that the decoder's arithmetic stays inside the library.
#include "common.cuh"

namespace bist {
namespace {

constexpr int kMaxGate = 4;

struct GateArgs {
  const void* src[kMaxGate];
  void* dsrc[kMaxGate];
  const float* logits;   // [R, K]
  float* dlogits;        // [R, K]
  const void* dout;      // [R, d]   (backward)
  void* out;             // [R, d]   (forward)
  int64_t R;
  int d, K;
};

template <typename T>
__global__ void __launch_bounds__(256) gate_fwd_kernel(GateArgs a) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= a.R) return;
  float p[kMaxGate], mx = -INFINITY, sum = 0.f;
#pragma unroll
  for (int k = 0; k < kMaxGate; ++k) {
    p[k] = k < a.K ? a.logits[r * a.K + k] : -INFINITY;
    mx = fmaxf(mx, p[k]);
  }
#pragma unroll
  for (int k = 0; k < kMaxGate; ++k) {
    p[k] = k < a.K ? expf(p[k] - mx) : 0.f;
    sum += p[k];
  }
  const float inv = 1.f / sum;
  T* o = reinterpret_cast<T*>(a.out) + r * a.d;
  for (int c = lane; c < a.d; c += 32) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxGate; ++k)
      if (k < a.K) acc = fmaf(p[k] * inv, to_f32(reinterpret_cast<const T*>(a.src[k])[r * a.d + c]), acc);
    o[c] = from_f32<T>(acc);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) gate_bwd_kernel(GateArgs a) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= a.R) return;
  float p[kMaxGate], dp[kMaxGate], mx = -INFINITY, sum = 0.f;
#pragma unroll
  for (int k = 0; k < kMaxGate; ++k) {
    p[k] = k < a.K ? a.logits[r * a.K + k] : -INFINITY;
    mx = fmaxf(mx, p[k]);
    dp[k] = 0.f;
  }
#pragma unroll
  for (int k = 0; k < kMaxGate; ++k) {
    p[k] = k < a.K ? expf(p[k] - mx) : 0.f;
    sum += p[k];
  }
  const float inv = 1.f / sum;
  const T* go = reinterpret_cast<const T*>(a.dout) + r * a.d;
  for (int c = lane; c < a.d; c += 32) {
    const float g = to_f32(go[c]);
#pragma unroll
    for (int k = 0; k < kMaxGate; ++k)
      if (k < a.K) {
        dp[k] = fmaf(g, to_f32(reinterpret_cast<const T*>(a.src[k])[r * a.d + c]), dp[k]);
        reinterpret_cast<T*>(a.dsrc[k])[r * a.d + c] = from_f32<T>(p[k] * inv * g);
      }
  }
  float dot = 0.f;
#pragma unroll
  for (int k = 0; k < kMaxGate; ++k) {
    dp[k] = warp_sum(dp[k]);
    dot = fmaf(p[k] * inv, dp[k], dot);
  }
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < kMaxGate; ++k)
      if (k < a.K) a.dlogits[r * a.K + k] = p[k] * inv * (dp[k] - dot);
  }
}

int check(const GateArgs& a, int dtype, bool bwd) {
  BIST_CHECK_ARG(a.K >= 1 && a.K <= kMaxGate && a.R > 0 && a.d > 0, "gate: 1 <= K <= 4 streams, non-empty rows");
  BIST_CHECK_ARG(dtype_ok(dtype) && a.logits, "gate: bad dtype or null logits");
  for (int k = 0; k < a.K; ++k) BIST_CHECK_ARG(a.src[k] && (!bwd || a.dsrc[k]), "gate: null stream %d", k);
  BIST_CHECK_ARG(bwd ? (a.dout && a.dlogits) : (a.out != nullptr), "gate: null output");
  return 0;
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

int bist_gate_combine_fwd(const void* const* src, const float* logits, void* out, int64_t R, int64_t d, int K, int dtype,
                          void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(src, "gate: null stream list");
  GateArgs a{};
  for (int k = 0; k < K && k < kMaxGate; ++k) a.src[k] = src[k];
  a.logits = logits; a.out = out; a.R = R; a.d = (int)d; a.K = K;
  BIST_TRY(check(a, dtype, false));
  ProfScope ps("gate_combine", st);
  const unsigned grid = (unsigned)((R + 7) / 8);
  if (dtype == BIST_BF16) BIST_LAUNCH(gate_fwd_kernel<__nv_bfloat16>, grid, 256, 0, st, a);
  else BIST_LAUNCH(gate_fwd_kernel<float>, grid, 256, 0, st, a);
  BIST_LAUNCHED();
  return 0;
}

int bist_gate_combine_bwd(const void* d_out, const void* const* src, const float* logits, void* const* d_src, float* d_logits,
                          int64_t R, int64_t d, int K, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(src && d_src, "gate: null stream list");
  GateArgs a{};
  for (int k = 0; k < K && k < kMaxGate; ++k) { a.src[k] = src[k]; a.dsrc[k] = d_src[k]; }
  a.logits = logits; a.dlogits = d_logits; a.dout = d_out; a.R = R; a.d = (int)d; a.K = K;
  BIST_TRY(check(a, dtype, true));
  ProfScope ps("gate_combine_bwd", st);
  const unsigned grid = (unsigned)((R + 7) / 8);
  if (dtype == BIST_BF16) BIST_LAUNCH(gate_bwd_kernel<__nv_bfloat16>, grid, 256, 0, st, a);
  else BIST_LAUNCH(gate_bwd_kernel<float>, grid, 256, 0, st, a);
  BIST_LAUNCHED();
  return 0;
}

}  // extern "C"
