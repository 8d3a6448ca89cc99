// Pointer attention of the generators (/root/reference/model/generator.py:62-63,68-69 and :104-105,112-114): a
// single-head attention whose PROBABILITIES are the product — pointer_attn(logits, text, text, mask).attn — and whose
// weighted sum is taken over the raw encoded text (text_vec), not over a value projection:
//   P[b,l,:]   = softmax_j( mask( q[b,l,:] . k[b,j,:] / sqrt(d) ) )        q, k: the module's linears[0], linears[1]
//   tvec[b,l,:] = sum_j P[b,l,j] v[b,j,:]                                   v = encoded_text
// Both outputs carry gradients (P into the pointer scatter and the loss, tvec into the switch Linear).
// Sizes are tiny (B*Lt rows, Ls <= a few hundred keys): plain SIMT, one CTA per target row forward, one per batch
// element backward (dk / dv accumulate over the target positions without atomics).
#include "common.cuh"

namespace bist {
namespace {

constexpr float kMaskFill = -1e9f;  // modules.py:60

struct PtrArgs {
  const void *q, *k, *v;     // [B,Lt,d], [B,Ls,d], [B,Ls,d]
  const uint8_t* mask;       // [B, Ls] (0 = masked) or null
  float* P;                  // [B,Lt,Ls]
  void* tvec;                // [B,Lt,d]
  const float* dP;           // backward inputs
  const void* dtvec;
  void *dq, *dk, *dv;        // backward outputs
  int B, Lt, Ls, d;
};

template <typename T>
__global__ void __launch_bounds__(256) ptr_fwd_kernel(PtrArgs a) {
  extern __shared__ __align__(16) float sm[];   // [Ls]
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x, b = r / a.Lt;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nw = blockDim.x >> 5;
  const T* q = reinterpret_cast<const T*>(a.q) + (int64_t)r * a.d;
  const T* k = reinterpret_cast<const T*>(a.k) + (int64_t)b * a.Ls * a.d;
  const T* v = reinterpret_cast<const T*>(a.v) + (int64_t)b * a.Ls * a.d;
  const float scale = rsqrtf((float)a.d);
  for (int j = warp; j < a.Ls; j += nw) {
    float s = 0.f;
    for (int c = lane; c < a.d; c += 32) s = fmaf(to_f32(q[c]), to_f32(k[(int64_t)j * a.d + c]), s);
    s = warp_sum(s) * scale;
    if (a.mask && a.mask[(int64_t)b * a.Ls + j] == 0) s = kMaskFill;
    if (lane == 0) sm[j] = s;
  }
  __syncthreads();
  if (warp == 0) {
    float mx = -INFINITY;
    for (int j = lane; j < a.Ls; j += 32) mx = fmaxf(mx, sm[j]);
    mx = warp_max(mx);
    float s = 0.f;
    for (int j = lane; j < a.Ls; j += 32) {
      const float e = expf(sm[j] - mx);
      sm[j] = e;
      s += e;
    }
    s = warp_sum(s);
    const float inv = 1.f / s;
    for (int j = lane; j < a.Ls; j += 32) {
      const float p = sm[j] * inv;
      sm[j] = p;
      a.P[(int64_t)r * a.Ls + j] = p;
    }
  }
  __syncthreads();
  T* o = reinterpret_cast<T*>(a.tvec) + (int64_t)r * a.d;
  for (int c = t; c < a.d; c += blockDim.x) {
    float acc = 0.f;
    for (int j = 0; j < a.Ls; ++j) acc = fmaf(sm[j], to_f32(v[(int64_t)j * a.d + c]), acc);
    o[c] = from_f32<T>(acc);
  }
}

// grid B; smem floats: dS[Lt][Ls]
template <typename T>
__global__ void __launch_bounds__(256) ptr_bwd_kernel(PtrArgs a) {
  extern __shared__ __align__(16) float dS[];
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nw = blockDim.x >> 5;
  const int Lt = a.Lt, Ls = a.Ls, d = a.d;
  const T* q = reinterpret_cast<const T*>(a.q) + (int64_t)b * Lt * d;
  const T* k = reinterpret_cast<const T*>(a.k) + (int64_t)b * Ls * d;
  const T* v = reinterpret_cast<const T*>(a.v) + (int64_t)b * Ls * d;
  const T* dtv = reinterpret_cast<const T*>(a.dtvec) + (int64_t)b * Lt * d;
  const float* P = a.P + (int64_t)b * Lt * Ls;
  const float* dP = a.dP + (int64_t)b * Lt * Ls;
  const float scale = rsqrtf((float)d);
  // dP_total[l,j] = dP[l,j] + dtvec[l,:] . v[j,:]
  for (int i = warp; i < Lt * Ls; i += nw) {
    const int l = i / Ls, j = i % Ls;
    float s = 0.f;
    for (int c = lane; c < d; c += 32) s = fmaf(to_f32(dtv[(int64_t)l * d + c]), to_f32(v[(int64_t)j * d + c]), s);
    s = warp_sum(s);
    if (lane == 0) dS[i] = s + dP[i];
  }
  __syncthreads();
  // softmax backward per row l
  for (int l = warp; l < Lt; l += nw) {
    float dot = 0.f;
    for (int j = lane; j < Ls; j += 32) dot = fmaf(P[l * Ls + j], dS[l * Ls + j], dot);
    dot = warp_sum(dot);
    for (int j = lane; j < Ls; j += 32) dS[l * Ls + j] = P[l * Ls + j] * (dS[l * Ls + j] - dot) * scale;
  }
  __syncthreads();
  T* dq = reinterpret_cast<T*>(a.dq) + (int64_t)b * Lt * d;
  T* dk = reinterpret_cast<T*>(a.dk) + (int64_t)b * Ls * d;
  T* dv = reinterpret_cast<T*>(a.dv) + (int64_t)b * Ls * d;
  for (int i = t; i < Lt * d; i += blockDim.x) {
    const int l = i / d, c = i % d;
    float acc = 0.f;
    for (int j = 0; j < Ls; ++j) acc = fmaf(dS[l * Ls + j], to_f32(k[(int64_t)j * d + c]), acc);
    dq[i] = from_f32<T>(acc);
  }
  for (int i = t; i < Ls * d; i += blockDim.x) {
    const int j = i / d, c = i % d;
    float ak = 0.f, av = 0.f;
    for (int l = 0; l < Lt; ++l) {
      ak = fmaf(dS[l * Ls + j], to_f32(q[(int64_t)l * d + c]), ak);
      av = fmaf(P[l * Ls + j], to_f32(dtv[(int64_t)l * d + c]), av);
    }
    dk[i] = from_f32<T>(ak);
    dv[i] = from_f32<T>(av);
  }
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

int bist_ptr_attn_fwd(const void* q, const void* k, const void* v, const uint8_t* mask, float* P, void* tvec, int64_t B,
                      int64_t Lt, int64_t Ls, int64_t d, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(q && k && v && P && tvec && B > 0 && Lt > 0 && Ls > 0 && d > 0, "ptr_attn: bad argument");
  BIST_CHECK_ARG(dtype_ok(dtype) && Ls * 4 <= 48 * 1024, "ptr_attn: bad dtype or too many source tokens");
  PtrArgs a{};
  a.q = q; a.k = k; a.v = v; a.mask = mask; a.P = P; a.tvec = tvec;
  a.B = (int)B; a.Lt = (int)Lt; a.Ls = (int)Ls; a.d = (int)d;
  ProfScope ps("ptr_attn", st);
  if (dtype == BIST_BF16) BIST_LAUNCH(ptr_fwd_kernel<__nv_bfloat16>, (unsigned)(B * Lt), 256, (size_t)Ls * 4, st, a);
  else BIST_LAUNCH(ptr_fwd_kernel<float>, (unsigned)(B * Lt), 256, (size_t)Ls * 4, st, a);
  BIST_LAUNCHED();
  return 0;
}

int bist_ptr_attn_bwd(const float* dP, const void* dtvec, const void* q, const void* k, const void* v, const float* P, void* dq,
                      void* dk, void* dv, int64_t B, int64_t Lt, int64_t Ls, int64_t d, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(dP && dtvec && q && k && v && P && dq && dk && dv && B > 0 && Lt > 0 && Ls > 0 && d > 0, "ptr_attn_bwd: bad argument");
  const size_t smem = (size_t)Lt * Ls * 4;
  BIST_CHECK_ARG(dtype_ok(dtype) && smem <= 200 * 1024, "ptr_attn_bwd: bad dtype or Lt*Ls too large");
  PtrArgs a{};
  a.q = q; a.k = k; a.v = v; a.P = const_cast<float*>(P); a.dP = dP; a.dtvec = dtvec; a.dq = dq; a.dk = dk; a.dv = dv;
  a.B = (int)B; a.Lt = (int)Lt; a.Ls = (int)Ls; a.d = (int)d;
  ProfScope ps("ptr_attn_bwd", st);
  if (dtype == BIST_BF16) {
    auto kern = ptr_bwd_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)B, 256, smem, st, a);
  } else {
    auto kern = ptr_bwd_kernel<float>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)B, 256, smem, st, a);
  }
  BIST_LAUNCHED();
  return 0;
}

}  // extern "C"
