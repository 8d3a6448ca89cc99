// Micro-benchmarks behind the design of csrc/stage1_fused.cu (numbers quoted in DESIGN.md):
//   1. TMA ingest per SM: every SM streams 16 KB boxes of an L2-resident matrix into a shared-memory ring; bytes / clock / SM
//      and issue-to-landed latency of one box.
//   2. tcgen05.mma rate by shape: M = 128 (cta_group::1) and M = 256 (cta_group::2), several N, operands in shared memory.
//   3. mma.sync (legacy tensor path) latency with and without a saturating tcgen05 stream on the same SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I bist_b200/csrc scripts/ubench.cu -o scripts/ubench -lcudart
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "tc_ptx.cuh"

namespace bist {
void set_error(const char* fmt, ...) { (void)fmt; }
}  // namespace bist
using namespace bist::tc;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

// ---------------------------------------------------------------------------------------------------------- 1. ingest
constexpr int ING_BOX = 16384;
template <int STAGES>
__global__ void __launch_bounds__(128, 1) ingest_kernel(const __grid_constant__ CUtensorMap map, int nloads, int rows_total, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = (uint64_t*)(smem + STAGES * ING_BOX);
  uint64_t* empty = full + STAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  long long t0 = clock64();
  if (warp == 0 && lane == 0) {
    int row = (blockIdx.x * 977) % (rows_total / 128);
    for (int i = 0; i < nloads; ++i) {
      const int s = i % STAGES;
      mbar_wait(&empty[s], ((i / STAGES) & 1) ^ 1);
      mbar_expect_tx(&full[s], ING_BOX);
      tma_load_2d(smem + s * ING_BOX, &map, &full[s], 0, row * 128);
      row = (row + 1) % (rows_total / 128);
    }
  } else if (warp == 1 && lane == 0) {
    for (int i = 0; i < nloads; ++i) {
      const int s = i % STAGES;
      mbar_wait(&full[s], (i / STAGES) & 1);
      mbar_arrive(&empty[s]);
    }
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  } else if (warp == 2 && lane == 0 && blockIdx.x == 0) {
    // (latency is measured by a separate launch with nloads = 1)
  }
}

// ---------------------------------------------------------------------------------------------------------- 2. UMMA rate
// smem: A [128 x 64] bf16 SW128 (16 KB) + B [256 x 64] (32 KB), contents irrelevant (zeros)
template <int NCTA>
__global__ void __launch_bounds__(128, 1) umma_kernel(int N, int nmma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* done = (uint64_t*)(smem + 49152);
  uint32_t* slot = (uint32_t*)(done + 1);
  for (int i = threadIdx.x; i < 49152 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(done, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { if (NCTA == 2) tmem_alloc2(slot, 512); else tmem_alloc(slot, 512); }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tm = *slot;
  const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)((128 * NCTA) >> 4) << 24);
    const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
    long long t0 = clock64();
    for (int i = 0; i < nmma; ++i) {
      if (NCTA == 2) tc_mma2(tm, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
      else tc_mma<false>(tm, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
    }
    if (NCTA == 2) tc_commit2(done); else tc_commit(done);
    mbar_wait(done, 0);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  if (NCTA == 2 && rank == 1 && threadIdx.x == 0) mbar_wait(done, 0);
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();
  if (threadIdx.x < 32) { if (NCTA == 2) tmem_dealloc2(tm, 512); else tmem_dealloc(tm, 512); }
}

// ---------------------------------------------------------------------------------------------------------- 3. mma.sync under UMMA load
__device__ __forceinline__ void hmma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__global__ void __launch_bounds__(256, 1) hmma_kernel(int with_umma, int N, int nmma, int nh, long long* out, float* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* done = (uint64_t*)(smem + 49152);
  uint32_t* slot = (uint32_t*)(done + 1);
  for (int i = threadIdx.x; i < 49152 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(done, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    if (threadIdx.x == 0 && with_umma) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
      for (int i = 0; i < nmma; ++i) tc_mma<false>(tm, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
      tc_commit(done);
      mbar_wait(done, 0);
    }
  } else if (warp >= 4) {
    // 4 warps: chains of 8 independent accumulators, nh rounds
    float acc[8][4];
    for (int n = 0; n < 8; ++n) for (int j = 0; j < 4; ++j) acc[n][j] = 0.f;
    const uint32_t a = threadIdx.x * 2654435761u;
    long long t0 = clock64();
    for (int i = 0; i < nh; ++i) {
#pragma unroll
      for (int n = 0; n < 8; ++n) hmma(acc[n], a, a + 1, a + 2, a + 3, a + n, a + 7);
    }
    long long t1 = clock64();
    float s = 0.f;
    for (int n = 0; n < 8; ++n) s += acc[n][0] + acc[n][3];
    if (s == 12345.f) sink[0] = s;
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * 4 + (warp - 4)] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}


// ---------------------------------------------------------------------------------------------------------- 4. issue loop
// One thread: per iteration  [try_wait on a completed barrier] + NM UMMAs + commit (+ optional test_wait poll), like the main loop
// of stage1_fused.  mode bit0: try_wait, bit1: commit per iteration, bit2: test_wait poll, bit3: a SECOND thread (another warp)
// issues small UMMAs (N = 32) into other TMEM columns at the same time.
__device__ __forceinline__ bool mbar_test_(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return done != 0;
}
__global__ void __launch_bounds__(128, 1) issue_kernel(int N, int NM, int iters, int mode, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* done = (uint64_t*)(smem + 49152);
  uint64_t* ready = done + 1;
  uint64_t* ring = done + 2;     // [4] commit targets
  uint64_t* done2 = done + 6;
  uint32_t* slot = (uint32_t*)(done + 8);
  for (int i = threadIdx.x; i < 49152 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    mbar_init(done, 1); mbar_init(ready, 1); mbar_init(done2, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&ring[i], 1);
    fence_barrier_init();
    mbar_arrive(ready);   // phase 0 complete: waits on parity 0 succeed immediately
  }
  if (threadIdx.x < 32) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
    long long t0 = clock64();
    int polls = 0;
    for (int i = 0; i < iters; ++i) {
      if (mode & 1) { mbar_wait(ready, 0); tc_fence_after(); }
      for (int k = 0; k < NM; ++k) tc_mma<false>(tm, ad + 2 * (k & 3), bd + 2 * (k & 3), idesc, 1u);
      if (mode & 2) tc_commit(&ring[i & 3]);
      if (mode & 4) polls += mbar_test_(&ring[(i + 2) & 3], 1) ? 1 : 0;
    }
    tc_commit(done);
    mbar_wait(done, 0);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0 + (polls == -1);
  } else if (threadIdx.x == 64 && (mode & 8)) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
    for (int i = 0; i < iters; ++i)
      for (int k = 0; k < 2; ++k) tc_mma<false>(tm + 256, ad + 2 * k, bd + 2 * k, idesc, 1u);
    tc_commit(done2);
    mbar_wait(done2, 0);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}


// ---------------------------------------------------------------------------------------------------------- 5. converged issue loop
// Same loop, but the WHOLE warp runs it (barrier waits are warp-uniform) and only the issue instructions sit under elect.sync.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__global__ void __launch_bounds__(128, 1) issue2_kernel(int N, int NM, int iters, int mode, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* done = (uint64_t*)(smem + 49152);
  uint64_t* ready = done + 1;
  uint64_t* ring = done + 2;
  uint32_t* slot = (uint32_t*)(done + 8);
  for (int i = threadIdx.x; i < 49152 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    mbar_init(done, 1); mbar_init(ready, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&ring[i], 1);
    fence_barrier_init();
    mbar_arrive(ready);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 64) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
    long long t0 = clock64();
    int polls = 0;
    for (int i = 0; i < iters; ++i) {
      if (mode & 1) { mbar_wait(ready, 0); tc_fence_after(); }
      if (elect_one()) {
        if (NM == 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma<false>(tm, ad + 2 * k, bd + 2 * k, idesc, 1u);
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) tc_mma<false>(tm, ad + 2 * (k & 3), bd + 2 * (k & 3), idesc, 1u);
        }
        if (mode & 2) tc_commit(&ring[i & 3]);
      }
      __syncwarp();
      if (mode & 4) polls += mbar_test_(&ring[(i + 2) & 3], 1) ? 1 : 0;
    }
    if (elect_one()) tc_commit(done);
    __syncwarp();
    mbar_wait(done, 0);
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0 + (polls == -1);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x >= 32 && threadIdx.x < 64) tmem_dealloc(tm, 512);
}


// ---------------------------------------------------------------------------------------------------------- 6. CUDA-core work under UMMA load
// 4 warps run (a) a dependent FFMA chain, (b) dependent shared-memory loads, (c) dependent shuffles, (d) independent 16-byte
// shared-memory stores, with and without a saturating tcgen05 stream (M = 128, N = 160 from shared memory) on the same SM.
__global__ void __launch_bounds__(256, 1) mix_kernel(int with_umma, int N, int nmma, int what, int iters, long long* out, float* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* done = (uint64_t*)(smem + 65536);
  uint32_t* slot = (uint32_t*)(done + 1);
  for (int i = threadIdx.x; i < 65536 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(done, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    if (with_umma) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint64_t ad = make_desc(smem_u32(smem)), bd = make_desc(smem_u32(smem + 16384));
      if (elect_one()) {
        for (int i = 0; i < nmma; ++i) tc_mma<false>(tm, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
        tc_commit(done);
      }
      __syncwarp();
      mbar_wait(done, 0);
    }
  } else if (warp >= 4) {
    volatile uint32_t* sm32 = reinterpret_cast<volatile uint32_t*>(smem + 49152);
    float a = (float)threadIdx.x, acc = 0.f;
    uint32_t idx = threadIdx.x & 31;
    long long t0 = clock64();
    if (what == 0) {
      for (int i = 0; i < iters; ++i) { a = fmaf(a, 1.000001f, 0.5f); }
      acc = a;
    } else if (what == 1) {
      for (int i = 0; i < iters; ++i) { idx = sm32[idx] + (threadIdx.x & 31); }
      acc = (float)idx;
    } else if (what == 2) {
      for (int i = 0; i < iters; ++i) { a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, 1 + (i & 15))); }
      acc = a;
    } else {
      uint4 v = make_uint4(idx, idx, idx, idx);
      uint4* dst = reinterpret_cast<uint4*>(smem + 49152) + threadIdx.x;
      for (int i = 0; i < iters; ++i) { dst[(i & 3) * 256] = v; }
      __syncwarp();
      acc = 1.f;
    }
    long long t1 = clock64();
    if (acc == 12345.678f) sink[0] = acc;
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * 4 + (warp - 4)] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

template <typename T>
static double mean_of(const std::vector<T>& v) { double s = 0; for (auto x : v) s += (double)x; return s / v.size(); }

int main() {
  int dev = 0, sms = 0, clk = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev));
  printf("SMs %d, clock attr %d kHz\n", sms, clk);
  long long* d_out;
  CK(cudaMalloc(&d_out, sizeof(long long) * 4096));
  float* d_sink;
  CK(cudaMalloc(&d_sink, 16));
  std::vector<long long> h(4096);

  // ---- 1. ingest: 8 MB bf16 matrix [65536 rows x 64 cols] (128-byte rows), boxes of 128 rows = 16 KB
  {
    const int rows = 65536;
    void* mat;
    CK(cudaMalloc(&mat, (size_t)rows * 128));
    CK(cudaMemset(mat, 0, (size_t)rows * 128));
    CUtensorMap map;
    if (make_map_2d(&map, mat, false, 64, rows, 64, 128, 128) != 0) { printf("tensor map failed\n"); return 1; }
    auto run = [&](auto kern, int stages, int grid, int nloads, const char* what) {
      const int smem = stages * ING_BOX + 2 * stages * 8 + 64;
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      for (int rep = 0; rep < 2; ++rep) {
        kern<<<grid, 128, smem>>>(map, nloads, rows, d_out);
        CK(cudaDeviceSynchronize());
      }
      CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * grid, cudaMemcpyDeviceToHost));
      std::vector<long long> v(h.begin(), h.begin() + grid);
      const double cyc = mean_of(v);
      printf("ingest %-28s grid %3d stages %2d loads %5d: %.0f cycles/CTA, %.1f B/clk/SM, %.0f cycles per 16 KB box\n", what, grid, stages, nloads,
             cyc, (double)nloads * ING_BOX / cyc, cyc / nloads);
    };
    run(ingest_kernel<1>, 1, 1, 1, "latency (1 SM, 1 box)");
    run(ingest_kernel<1>, 1, 1, 64, "serial (1 SM, depth 1)");
    run(ingest_kernel<2>, 2, sms, 512, "all SMs, ring 2");
    run(ingest_kernel<4>, 4, sms, 512, "all SMs, ring 4");
    run(ingest_kernel<8>, 8, sms, 512, "all SMs, ring 8");
    run(ingest_kernel<12>, 12, sms, 512, "all SMs, ring 12");
    run(ingest_kernel<8>, 8, 1, 512, "one SM, ring 8");
    run(ingest_kernel<8>, 8, 16, 512, "16 SMs, ring 8");
    CK(cudaFree(mat));
  }
  // ---- 2. UMMA rate
  {
    const int smem = 49152 + 64;
    CK(cudaFuncSetAttribute(umma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CK(cudaFuncSetAttribute(umma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int nmma = 2048;
    for (int N : {16, 32, 64, 80, 96, 128, 160, 192, 256}) {
      umma_kernel<1><<<sms, 128, smem>>>(N, nmma, d_out);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
      std::vector<long long> v(h.begin(), h.begin() + sms);
      const double c = mean_of(v) / nmma;
      printf("umma cta_group::1 M=128 N=%3d K=16: %.1f cycles/MMA (%.0f MAC/clk/SM)\n", N, c, 128.0 * N * 16 / c);
    }
    for (int N : {32, 64, 96, 128, 160, 192, 256}) {
      cudaLaunchConfig_t cfg = {};
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.gridDim = dim3(sms / 2 * 2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.attrs = at; cfg.numAttrs = 1;
      CK(cudaLaunchKernelEx(&cfg, umma_kernel<2>, N, nmma, d_out));
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
      std::vector<long long> v;
      for (int i = 0; i < sms / 2 * 2; i += 2) v.push_back(h[i]);
      const double c = mean_of(v) / nmma;
      printf("umma cta_group::2 M=256 N=%3d K=16: %.1f cycles/MMA (%.0f MAC/clk/SM)\n", N, c, 256.0 * N * 16 / c / 2);
    }
  }
  // ---- 3. mma.sync with / without a tcgen05 stream
  {
    const int smem = 49152 + 64;
    CK(cudaFuncSetAttribute(hmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int nh = 256;
    for (int with : {0, 1}) {
      for (int N : {160, 256}) {
        hmma_kernel<<<sms, 256, smem>>>(with, N, 4096, nh, d_out, d_sink);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms * 4, cudaMemcpyDeviceToHost));
        std::vector<long long> v(h.begin(), h.begin() + sms * 4);
        printf("mma.sync m16n8k16 x8 independent per round, 4 warps/SM, tcgen05 stream %s (N=%d): %.1f cycles per mma.sync per warp\n",
               with ? "ON " : "off", N, mean_of(v) / (nh * 8));
      }
    }
  }
  // ---- 4. issue loop
  {
    const int smem = 49152 + 128;
    CK(cudaFuncSetAttribute(issue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int iters = 512;
    for (int N : {80, 160}) {
      for (int NM : {4, 8}) {
        for (int mode : {0, 1, 2, 3, 7, 11}) {
          issue_kernel<<<sms, 128, smem>>>(N, NM, iters, mode, d_out);
          CK(cudaDeviceSynchronize());
          CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
          std::vector<long long> v(h.begin(), h.begin() + sms);
          printf("issue loop N=%3d, %d UMMAs per iteration, mode %2d (1 try_wait, 2 commit, 4 poll, 8 second issuer): %.0f cycles per iteration (pipe time %.0f)\n",
                 N, NM, mode, mean_of(v) / iters, NM * (N <= 80 ? 52.2 : N / 2.0));
        }
      }
    }
  }
  // ---- 5. converged issue loop
  {
    const int smem = 49152 + 128;
    CK(cudaFuncSetAttribute(issue2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int iters = 512;
    for (int N : {80, 160}) {
      for (int NM : {4, 8}) {
        for (int mode : {0, 1, 3, 7}) {
          issue2_kernel<<<sms, 128, smem>>>(N, NM, iters, mode, d_out);
          CK(cudaDeviceSynchronize());
          CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
          std::vector<long long> v(h.begin(), h.begin() + sms);
          printf("converged issue loop N=%3d, %d UMMAs per iteration, mode %2d (1 try_wait, 2 commit, 4 poll): %.0f cycles per iteration (pipe time %.0f)\n",
                 N, NM, mode, mean_of(v) / iters, NM * (N <= 80 ? 52.2 : N / 2.0));
        }
      }
    }
  }
  // ---- 6. CUDA-core work under UMMA load
  {
    const int smem = 65536 + 64;
    CK(cudaFuncSetAttribute(mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const char* names[4] = {"dependent FFMA", "dependent LDS.32", "dependent SHFL", "independent STS.128"};
    const int iters = 2048;
    for (int what = 0; what < 4; ++what) {
      for (int with : {0, 1}) {
        mix_kernel<<<sms, 256, smem>>>(with, 160, 16384, what, iters, d_out, d_sink);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h.data(), d_out, sizeof(long long) * sms * 4, cudaMemcpyDeviceToHost));
        std::vector<long long> v(h.begin(), h.begin() + sms * 4);
        printf("mix %-20s tcgen05 stream (N=160) %s: %.1f cycles per op per warp\n", names[what], with ? "ON " : "off", mean_of(v) / iters);
      }
    }
  }
  printf("done\n");
  return 0;
}
