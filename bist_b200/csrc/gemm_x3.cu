// fp32 products on the tensor cores at fp32 accuracy (fp32 mode).
//   A = A_hi + A_lo,  W = W_hi + W_lo   with  x_hi = x with its 13 low mantissa bits cleared (exactly a tf32 value, so the
//   tensor core's fp32 -> tf32 conversion cannot change it) and x_lo = x - x_hi (exact in fp32; 2^-11 of x at most)
//   A W^T  ~=  A_hi W_hi^T + A_lo W_hi^T + A_hi W_lo^T          (the dropped A_lo W_lo^T term is 2^-22 relative)
// as three tcgen05 kind::tf32 products summed in the TMEM accumulator of ONE launch (GemmArgs::A2 / W2: the K loop runs three
// times over different operand pairs), so C is written once, with bias, ReLU and the residual, by the ordinary epilogue.  The splits of the activations are written by one elementwise launch into caller-provided scratch that
// mirrors the operand's memory range; an operand that several products share (the video tensor) is split once.
#include "common.cuh"

namespace bist {
namespace {

__global__ void __launch_bounds__(256) split_tf32_kernel(const float4* __restrict__ x, float4* __restrict__ hi, float4* __restrict__ lo,
                                                         long long n4) {
  pdl_trigger();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    float4 h, l;
    h.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l.x = v.x - h.x;
    h.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l.y = v.y - h.y;
    h.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l.z = v.z - h.z;
    h.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l.w = v.w - h.w;
    hi[i] = h;
    lo[i] = l;
  }
}

}  // namespace

// hi / lo of n fp32 values (n % 4 == 0, 16-byte aligned pointers)
int split_tf32(const float* x, float* hi, float* lo, long long n, cudaStream_t st) {
  BIST_CHECK_ARG(x && hi && lo && n > 0 && n % 4 == 0, "split_tf32: bad argument");
  BIST_CHECK_ARG((((uintptr_t)x | (uintptr_t)hi | (uintptr_t)lo) & 15) == 0, "split_tf32: misaligned pointer");
  const long long n4 = n / 4;
  long long blocks = (n4 + 255) / 256;
  const long long cap = 8LL * 148;
  if (blocks > cap) blocks = cap;
  ProfScope ps("split_tf32", st);
  BIST_LAUNCH(split_tf32_kernel, (unsigned)blocks, 256, 0, st, reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(hi),
              reinterpret_cast<float4*>(lo), n4);
  BIST_LAUNCHED();
  return 0;
}

bool fp32x3_enabled() {
  static const bool off = getenv("BIST_FP32_SIMT") != nullptr;   // debugging aid: fp32 products on the CUDA cores as before
  return !off;
}

// C = act(alpha A W^T + bias) + R for fp32 A, W, C through three tf32 products.  `sa` / `sw` describe the splits of the memory
// ranges that hold A and W: src = first element of the range, n = its length; hi / lo mirror it.  ready != 0: the split has
// been written already (shared operand); else it is written here.  Falls back to the CUDA-core kernel when the shape does not
// fit the tensor-core kernel.
int gemm_fp32x3(const GemmArgs& g, const SplitRef& sa, const SplitRef& sw, cudaStream_t st) {
  const bool ok = fp32x3_enabled() && g.a_dtype == BIST_F32 && g.w_dtype == BIST_F32 && g.c_dtype == BIST_F32 && sa.hi && sa.lo && sw.hi &&
                  sw.lo && sa.n % 4 == 0 && sw.n % 4 == 0 && g.K % 32 == 0 && g.N >= 8 && g.wr_T == 0 &&
                  (long long)g.M * g.N * g.Z >= 64 * 64 && g.a.inner == kNoSplit && g.w.inner == kNoSplit &&
                  ((((uintptr_t)g.A | (uintptr_t)g.W | (uintptr_t)sa.src | (uintptr_t)sw.src) & 15) == 0) && (g.a.si % 4 == 0) && (g.w.si % 4 == 0);
  if (!ok) return gemm(g, st);
  const long long a_off = (const float*)g.A - sa.src, w_off = (const float*)g.W - sw.src;
  BIST_CHECK_ARG(a_off >= 0 && a_off < sa.n && w_off >= 0 && w_off < sw.n, "gemm_fp32x3: operand outside its split range");
  if (!sa.ready) BIST_TRY(split_tf32(sa.src, sa.hi, sa.lo, sa.n, st));
  if (!sw.ready) BIST_TRY(split_tf32(sw.src, sw.hi, sw.lo, sw.n, st));
  GemmArgs p = g;
  p.allow_tf32 = 1;
  p.A = sa.hi + a_off; p.A2 = sa.lo + a_off; p.W = sw.hi + w_off; p.W2 = sw.lo + w_off;
  bool taken = false;
  BIST_TRY(gemm_tc(p, st, &taken));      // one launch: the three products meet in the TMEM accumulator
  if (!taken) return gemm(g, st);
  return 0;
}

}  // namespace bist
