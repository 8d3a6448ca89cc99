// Generator + label-smoothed loss, forward and backward in ONE pass over the vocabulary logits
// (/root/reference/model/generator.py:77-127 MultiPointerGenerator / :29-75 PointerGenerator / :11-27 Generator,
//  /root/reference/model/label_smoothing.py:20-30, /root/reference/model/optimize.py:46-53).
//
// Per target position r (one CTA): with Z[r,:] = decoded_text[r,:] . E^T already computed by the GEMM,
//   pv      = softmax(Z[r,:])                                                     generator.py:82-83
//   ptr_i   = scatter_add over the source tokens of pointer i of P_i[r,:]           generator.py:108-110
//   a       = softmax(switch logits [a_0 .. a_{n-1}, a_V])                          generator.py:121
//   p_out   = a_V pv + sum_i a_i ptr_i ;  logp = log p_out                         generator.py:122-126
//   loss_r  = sum_v t[v] (log t[v] - logp[v]),  t = smoothing/(V-2), t[target] = 1 - smoothing, t[pad] = 0,
//             whole row 0 when target == pad                                        label_smoothing.py:22-30
// and, in the same pass, the gradient of loss_r with respect to every input:
//   g[v] = -t[v] / p_out[v];  G_V = <pv, g>;  dZ[v] = a_V pv[v] (g[v] - G_V)  (written over Z in place),
//   dP_i[j] = a_i g[text_i[j]];  G_i = <P_i, g[text_i]>;  dsw_k = a_k (G_k - sum_m a_m G_m).
// The [R, V] probabilities are never written to memory: the row lives in shared memory between the softmax, the
// pointer scatter, the loss and the gradient.  With n_ptr = 0 (and no switch) this is log_softmax + label-smoothed
// KL, the auto-encoder heads of optimize.py:55-82.
#include "common.cuh"

namespace bist {
namespace {

constexpr int kMaxPtr = 2;

struct GenLossArgs {
  void* Z;                       // [R, V] logits in, dZ out (when want_grad)
  float* logp;                   // optional [R, V] log p_out
  const float* P[kMaxPtr];       // [R, Ls_i] pointer attention
  float* dP[kMaxPtr];
  const int64_t* text[kMaxPtr];  // [B, Ls_i] source token ids
  int Ls[kMaxPtr];
  const float* sw;               // [R, n_ptr + 1] switch logits (null when n_ptr == 0)
  float* dsw;
  const int64_t* target;         // [R] (null: inference, no loss)
  float* row_loss;               // [R]
  int64_t R;
  int V, Lt, n_ptr, pad, want_grad;
  float smoothing;
};

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = 0.f;
  for (int i = 0; i < nw; ++i) t += red[i];
  return t;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = -INFINITY;
  for (int i = 0; i < nw; ++i) t = fmaxf(t, red[i]);
  return t;
}

template <typename T>
__global__ void __launch_bounds__(256) genloss_kernel(GenLossArgs a) {
  extern __shared__ __align__(16) float sm[];
  pdl_trigger();
  pdl_wait();
  const int V = a.V;
  float* pv = sm;          // [V] softmax of the logits
  float* po = sm + V;      // [V] p_out
  float* red = po + V;     // [32]
  const int64_t r = blockIdx.x;
  const int b = (int)(r / a.Lt);
  const int t = threadIdx.x, nt = blockDim.x;
  T* z = reinterpret_cast<T*>(a.Z) + r * V;

  // switch: a = softmax(sw[r, :])  (a_V last)
  float aw[kMaxPtr + 1];
  {
    float mx = -INFINITY, s = 0.f;
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k) {
      aw[k] = (a.sw && k <= a.n_ptr) ? a.sw[r * (a.n_ptr + 1) + k] : -INFINITY;
      mx = fmaxf(mx, aw[k]);
    }
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k) {
      aw[k] = (a.sw && k <= a.n_ptr) ? expf(aw[k] - mx) : 0.f;
      s += aw[k];
    }
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k) aw[k] = a.sw ? aw[k] / s : 0.f;
  }
  float aV = 1.f;
  if (a.sw) {
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k)
      if (k == a.n_ptr) aV = aw[k];
  }

  // softmax over the vocabulary
  float mx = -INFINITY;
  for (int v = t; v < V; v += nt) {
    const float x = to_f32(z[v]);
    pv[v] = x;
    mx = fmaxf(mx, x);
  }
  mx = block_max(mx, red);
  float s = 0.f;
  for (int v = t; v < V; v += nt) {
    const float e = expf(pv[v] - mx);
    pv[v] = e;
    s += e;
  }
  s = block_sum(s, red);
  const float inv = 1.f / s;
  for (int v = t; v < V; v += nt) {
    const float p = pv[v] * inv;
    pv[v] = p;
    po[v] = aV * p;
  }
  __syncthreads();
  // pointer distributions scattered onto the vocabulary
#pragma unroll
  for (int i = 0; i < kMaxPtr; ++i)
    if (i < a.n_ptr)
      for (int j = t; j < a.Ls[i]; j += nt) {
        const int64_t tok = a.text[i][(int64_t)b * a.Ls[i] + j];
        atomicAdd(&po[tok], aw[i] * a.P[i][r * a.Ls[i] + j]);
      }
  __syncthreads();
  if (a.logp)
    for (int v = t; v < V; v += nt) a.logp[r * V + v] = logf(po[v]);
  if (!a.target) return;

  const int64_t y = a.target[r];
  const bool live = y != a.pad;
  const float tsm = a.smoothing / (float)(V - 2), conf = 1.f - a.smoothing;
  // loss and G_V = <pv, g>
  float loss = 0.f, gv = 0.f;
  if (live)
    for (int v = t; v < V; v += nt) {
      const float tv = v == a.pad ? 0.f : (v == y ? conf : tsm);
      if (tv > 0.f) {
        loss += tv * (logf(tv) - logf(po[v]));
        gv -= pv[v] * tv / po[v];
      }
    }
  loss = block_sum(loss, red);
  if (t == 0) a.row_loss[r] = loss;
  if (!a.want_grad) return;
  gv = block_sum(gv, red);

  // pointer side: dP_i and G_i
  float G[kMaxPtr + 1];
#pragma unroll
  for (int i = 0; i < kMaxPtr; ++i) {
    G[i] = 0.f;
    if (i < a.n_ptr) {
      float acc = 0.f;
      for (int j = t; j < a.Ls[i]; j += nt) {
        const int64_t tok = a.text[i][(int64_t)b * a.Ls[i] + j];
        const float tv = (!live || tok == a.pad) ? 0.f : (tok == y ? conf : tsm);
        const float g = tv > 0.f ? -tv / po[tok] : 0.f;
        a.dP[i][r * a.Ls[i] + j] = aw[i] * g;
        acc += a.P[i][r * a.Ls[i] + j] * g;
      }
      G[i] = block_sum(acc, red);
    }
  }
  if (a.sw && t == 0) {
    float Gk[kMaxPtr + 1], dot = 0.f;
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k) {
      Gk[k] = k < a.n_ptr ? G[k < kMaxPtr ? k : 0] : (k == a.n_ptr ? gv : 0.f);
      dot += aw[k] * Gk[k];
    }
#pragma unroll
    for (int k = 0; k <= kMaxPtr; ++k)
      if (k <= a.n_ptr) a.dsw[r * (a.n_ptr + 1) + k] = aw[k] * (Gk[k] - dot);
  }
  // dZ over the logits, in place
  for (int v = t; v < V; v += nt) {
    float d = 0.f;
    if (live) {
      const float tv = v == a.pad ? 0.f : (v == y ? conf : tsm);
      const float g = tv > 0.f ? -tv / po[v] : 0.f;
      d = aV * pv[v] * (g - gv);
    }
    z[v] = from_f32<T>(d);
  }
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

int bist_genloss_fwd_bwd(void* Z, float* logp, int n_ptr, const float* const* P, float* const* dP,
                         const int64_t* const* text, const int64_t* Ls, const float* sw, float* dsw,
                         const int64_t* target, float* row_loss, int64_t R, int64_t V, int64_t Lt, int64_t pad,
                         float smoothing, int want_grad, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(Z && R > 0 && V > 2 && Lt > 0 && R % Lt == 0, "genloss: bad shape");
  BIST_CHECK_ARG(dtype_ok(dtype), "genloss: bad dtype");
  BIST_CHECK_ARG(n_ptr >= 0 && n_ptr <= kMaxPtr, "genloss: at most %d pointer sources", kMaxPtr);
  BIST_CHECK_ARG((n_ptr == 0) || (P && text && Ls && sw), "genloss: pointer sources need P, text, Ls and switch logits");
  BIST_CHECK_ARG(!target || row_loss, "genloss: target without row_loss");
  BIST_CHECK_ARG(!want_grad || (target && (n_ptr == 0 || dP) && (!sw || dsw)), "genloss: gradient outputs missing");
  BIST_CHECK_ARG(pad >= 0 && pad < V && smoothing >= 0.f && smoothing < 1.f, "genloss: bad pad id or smoothing");
  GenLossArgs a{};
  a.Z = Z; a.logp = logp; a.sw = sw; a.dsw = dsw; a.target = target; a.row_loss = row_loss;
  a.R = R; a.V = (int)V; a.Lt = (int)Lt; a.n_ptr = n_ptr; a.pad = (int)pad; a.want_grad = want_grad; a.smoothing = smoothing;
  for (int i = 0; i < n_ptr; ++i) {
    BIST_CHECK_ARG(P[i] && text[i] && Ls[i] > 0 && (!want_grad || dP[i]), "genloss: null pointer source %d", i);
    a.P[i] = P[i]; a.dP[i] = dP ? dP[i] : nullptr; a.text[i] = text[i]; a.Ls[i] = (int)Ls[i];
  }
  const size_t smem = ((size_t)2 * V + 32) * sizeof(float);
  if (smem > 200 * 1024) { set_error("genloss: vocabulary of %lld does not fit one CTA's shared memory", (long long)V); return BIST_ERR_UNSUPPORTED; }
  ProfScope ps("genloss", st);
  if (dtype == BIST_BF16) {
    auto kern = genloss_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)R, 256, smem, st, a);
  } else {
    auto kern = genloss_kernel<float>;
    if (smem > 48 * 1024) BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    BIST_LAUNCH(kern, (unsigned)R, 256, smem, st, a);
  }
  BIST_LAUNCHED();
  return 0;
}

}  // extern "C"
