// tcgen05 / TMEM / TMA GEMM for sm_100a:  C = act(alpha * A . W^T + bias) + R
//   A [M,K] and W [N,K] both K-major (dense rows), bf16 (kind::f16) or fp32 read as tf32
//   (kind::tf32); fp32 accumulation in tensor memory.
// Persistent kernel: each CTA walks 128 x BN output tiles (n fastest, so concurrently resident CTAs
// share A rows and the whole of W in L2).  320 threads, warp-specialised:
//   warp 0    TMA producer   cp.async.bulk.tensor (128B-swizzled, 128-byte-wide K slabs) -> smem ring
//   warp 1    MMA issuer     one elected lane issues tcgen05.mma (UMMA 128 x BN x 16|8) into one of TWO
//                            TMEM accumulator stages; tcgen05.commit releases ring slots and publishes
//                            the accumulator; the warp also owns tcgen05.alloc / dealloc
//   warps 2-9 epilogue       tcgen05.ld (32 lanes x 32 columns) -> per-warp smem transpose -> alpha / bias /
//                            ReLU / residual and 16-byte coalesced row stores.  Two warps per TMEM lane
//                            quarter (half of the columns each).  The epilogue of tile i overlaps the
//                            MMAs of tile i+1 (other accumulator stage).
// Barriers: full/empty per ring slot (TMA complete_tx / tcgen05.commit), acc_full/acc_empty per
// accumulator stage (tcgen05.commit / one arrive per epilogue warp).
#include "tc_ptx.cuh"

#include <stdlib.h>

#include <mutex>

namespace bist {
namespace {
using namespace tc;

struct TcArgs {
  GemmArgs g;
  long long a_ld, w_ld;  // leading dimensions (elements) of the 2-D tensors the TMA maps describe
  int nkb;               // K / (elements per 128-byte slab)
  int vec_c;             // 1 = 16-byte aligned output rows (vector stores)
  int lean;              // 1 = one batch, dense C (and R) rows, offsets below 2^31, no accumulate: the short epilogue
  int npass;             // passes of the K loop into one accumulator: 3 = split-tf32 product A W^T + A2 W^T + A W2^T (GemmArgs::A2 /
                         // W2, split = 1), or kcat parts of a K-concatenated product (split = 0)
  int split, a_cat, w_cat_rows;
  int tma_c;             // 1 = lean, and no residual / second output / regions: the output leaves through TMA stores (map_c)
  int tiles_m, tiles_n, total_tiles;
  int bn_valid;          // output columns per N tile (== BN except in region mode)
  int w3d, s_box;        // region mode: W tile = 3-D box [s_box regions][wr_T rows][128 B]; w_stage_bytes of it land per slot
  int w_stage_bytes;
  long long w_bt_elems;  // elements per bt row group (wr_S * w_ld)
};

#ifdef BIST_GEMM_TRACE
__device__ long long g_trace[3][64];   // CTA 0 only: [role][event]
#define TRACE(role, idx) do { if (blockIdx.x == 0 && (idx) < 64) g_trace[role][idx] = clock64(); } while (0)
#else
#define TRACE(role, idx) do {} while (0)
#endif

constexpr int STG_LD = 36;  // floats per staging row: 16-byte aligned, conflict-free for float4 in both phases

template <int BN, int STAGES, int NCTA = 1>
struct Smem {
  static constexpr int A_BYTES = BM * SLAB_BYTES, W_BYTES = BN * SLAB_BYTES / NCTA;   // a CTA pair splits the W tile
  static constexpr int STAGE_BYTES = A_BYTES + W_BYTES;
  static constexpr int STG_OFF = STAGES * STAGE_BYTES;           // 8 epilogue warps x [32][STG_LD] floats
  static constexpr int BAR_OFF = STG_OFF + 8 * 32 * STG_LD * 4;
  static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 4) * 8 + 16 + 1024;  // + alignment slack
};

// NCTA == 2 (bf16, plain W tiles, BN = 256): the two CTAs of a cluster compute two vertically adjacent 128 x 256 tiles
// with ONE cta_group::2 UMMA of M = 256.  Each CTA loads its own 128 rows of A and 128 of the 256 W rows, so an SM pulls
// 32 KB instead of 48 KB per K slab from L2 and the tensor core reads half of B from each shared memory; the ring is 5
// slots deep instead of 3.  All TMA bytes are counted on the leader's `full` barriers, tcgen05.commit multicasts `empty`
// and `acc_full` to both CTAs, the peer's epilogue warps hand the accumulator back with remote arrives.
template <int BN, int STAGES, bool TF32, int NCTA = 1>
__global__ void __launch_bounds__(320, 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                         const __grid_constant__ CUtensorMap map_w,
                                                         const __grid_constant__ CUtensorMap map_c,
                                                         const __grid_constant__ CUtensorMap map_a2,
                                                         const __grid_constant__ CUtensorMap map_w2, const TcArgs t) {
  using S = Smem<BN, STAGES, NCTA>;
  constexpr int ESZ = TF32 ? 4 : 2;
  constexpr int BKE = SLAB_BYTES / ESZ;        // elements of K per slab: 64 bf16 / 32 tf32
  constexpr int UMMA_K = 32 / ESZ;             // 16 bf16 / 8 tf32 (32 bytes of K per instruction)
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + S::BAR_OFF);
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;   // [2]
  uint64_t* acc_empty = acc_full + 2;    // [2]
  uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);

  const GemmArgs& g = t.g;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    // the first TMA instruction otherwise waits for its descriptor to be fetched: ask for both now, while the rest of the set-up runs
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_w)) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 8 * NCTA); }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (NCTA == 2) tmem_alloc2(tmem_slot, 2 * BN);
    else tmem_alloc(tmem_slot, 2 * BN);
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the tail of the previous kernel
  const int per_z = t.tiles_m * t.tiles_n;   // NCTA == 2: tiles_m counts PAIRS of row tiles
  const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
  const int unit0 = NCTA == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ustep = NCTA == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0) {
    {
      int it = 0;
      for (int tile = unit0; tile < t.total_tiles; tile += ustep) {
        // (unbatched products skip the batch arithmetic: its 64-bit divisions were ~1.7 k clk in front of the first TMA request)
        const int z = g.Z == 1 ? 0 : tile / per_z, rem = tile - z * per_z;
        const int m0 = (NCTA * (rem / t.tiles_n) + rank) * BM, nt = rem % t.tiles_n, n0 = nt * BN + rank * (BN / 2);
        // batch offsets become (row, column) coordinates inside the 2-D tensors
        long long ao = 0, wo = 0;
        int a_row = m0, a_col = 0, w_row = n0, w_col = 0;
        if (g.Z != 1) {
          ao = g.az.off(z, g.zin, g.zmid); wo = g.wz.off(z, g.zin, g.zmid);
          a_row = (int)(ao / t.a_ld) + m0; a_col = (int)(ao % t.a_ld);
          w_row = (int)(wo / t.w_ld) + n0; w_col = (int)(wo % t.w_ld);
        }
        for (int pass = 0; pass < t.npass; ++pass) {
        // split-tf32 products: pass 1 reads the low halves of A, pass 2 the low halves of W, into the same accumulator
        const CUtensorMap* const pa = (t.split && pass == 1) ? &map_a2 : &map_a;
        const CUtensorMap* const pw = (t.split && pass == 2) ? &map_w2 : &map_w;
        // K concatenation (GemmArgs::kcat): part `pass` of A sits a_cat columns further right, its W block w_cat_rows rows further down
        const int a_colp = a_col + (t.split ? 0 : pass * t.a_cat), w_rowp = w_row + (t.split ? 0 : pass * t.w_cat_rows);
        for (int kb = 0; kb < t.nkb; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          if (elect_one_lane()) {
            uint8_t* st = smem + s * S::STAGE_BYTES;
            if (NCTA == 2) {
              if (rank == 0) mbar_expect_tx(&full[s], 2 * S::STAGE_BYTES);   // both CTAs' A and W bytes land on the leader's barrier
              tma_load_2d_pair(st, &map_a, &full[s], a_colp + kb * BKE, a_row);
              tma_load_2d_pair(st + S::A_BYTES, &map_w, &full[s], w_col + kb * BKE, w_rowp);
            } else {
              mbar_expect_tx(&full[s], S::A_BYTES + t.w_stage_bytes);
              tma_load_2d(st, pa, &full[s], a_colp + kb * BKE, a_row);
              if (t.w3d) tma_load_3d(st + S::A_BYTES, &map_w, &full[s], kb * BKE, (int)(wo / t.w_bt_elems), nt * t.s_box);
              else tma_load_2d(st + S::A_BYTES, pw, &full[s], w_col + kb * BKE, w_rowp);
            }
            TRACE(0, it);
          }
          __syncwarp();
        }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // The whole warp stays converged and ONE elected lane issues (elect.sync): in a divergent single-lane region the compiler wraps
      // every tcgen05 instruction in a uniform-register waterfall loop (~110 clk per UMMA issued; an N = 64 UMMA executes in 32).
      // instruction descriptor (cute::UMMA::InstrDescriptor): D = f32, A/B = bf16 | tf32, K-major both
      const uint32_t fmt = TF32 ? 2u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((BM * NCTA) >> 4) << 24);
      int it = 0, tc_ = 0;
      for (int tile = unit0; tile < t.total_tiles; tile += ustep, ++tc_) {
        const int a = tc_ & 1;
        if (NCTA == 2) mbar_wait_cluster(&acc_empty[a], ((tc_ >> 1) & 1) ^ 1);
        else mbar_wait(&acc_empty[a], ((tc_ >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(a * BN);
        const int nkb_all = t.nkb * t.npass;
        for (int kb = 0; kb < nkb_all; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          if (elect_one_lane()) {
            const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
            const uint64_t ad = make_desc(sa), wd = make_desc(sa + S::A_BYTES);
#pragma unroll
            for (int k = 0; k < BKE / UMMA_K; ++k) {  // advance 32 bytes of K inside the swizzle atom: +2 in the address field
              if (NCTA == 2) tc_mma2(tacc, ad + 2 * k, wd + 2 * k, idesc, (kb | k) ? 1u : 0u);
              else tc_mma<TF32>(tacc, ad + 2 * k, wd + 2 * k, idesc, (kb | k) ? 1u : 0u);
            }
            if (NCTA == 2) tc_commit2(&empty[s]);
            else tc_commit(&empty[s]);
            if (kb == nkb_all - 1) {
              if (NCTA == 2) tc_commit2(&acc_full[a]);
              else tc_commit(&acc_full[a]);
            }
            TRACE(1, it);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---- epilogue: warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32); two warps per lane quarter,
    // each owning half of the tile's columns.  All per-element arithmetic happens after the smem transpose,
    // where a lane's 4 columns are fixed (one bias load per chunk) and rows are addressed through offsets
    // computed once per tile (no integer division in the loop).
    const int lg = warp & 3, ch = (warp - 2) >> 2;
    const uint32_t stg = smem_u32(smem + S::STG_OFF) + (uint32_t)((warp - 2) * 32 * STG_LD * 4);
    const int sub_r = lane >> 3, sub_c = (lane & 7) * 4;  // phase-2 mapping: 4 rows x 8 float4 per instruction
    const uint32_t stg_w = stg + (uint32_t)(lane * STG_LD * 4);                 // phase 1: thread = row
    const uint32_t stg_r = stg + (uint32_t)((sub_r * STG_LD + sub_c) * 4);      // phase 2: + rr * 4 rows
    constexpr int CH_COLS = BN / 2;
    const float alpha = g.alpha;
    const int relu = g.relu, vec_c = t.vec_c, Mtot = g.M;
    const bool c_bf16 = g.c_dtype == BIST_BF16, r_bf16 = g.r_dtype == BIST_BF16, has_r = g.R != nullptr;
    int tc_ = 0;
    if (t.tma_c) {
      // ---- plain products (bias / ReLU only): each lane keeps its accumulator row, adds the bias from a shared-memory copy of the
      // tile's slice, converts, and writes 64 bytes of the row into a SWIZZLE_64B box that one TMA store per 32 x 64-byte box sends
      // out (clipped at the matrix edge).  No transposition round trip, no global store instructions, no row predicates.
      constexpr int NCH = CH_COLS / 32;
      const uint32_t sbias = stg + 4096u;                      // behind the two 2 KB boxes of this warp (4608 B per warp)
      const uint32_t my_box_row = (uint32_t)lane * 64u, swz = (uint32_t)((lane >> 1) & 3);
      const float* bias = g.bias;
      const int Ncols = g.N;
      uint32_t nbox = 0;                                       // boxes sent so far: box buffer = nbox & 1
      for (int tile = unit0; tile < t.total_tiles; tile += ustep, ++tc_) {
        const int tm = tile / t.tiles_n, nt = tile - tm * t.tiles_n;
        const int m0 = (NCTA * tm + rank) * BM, n0 = nt * BN;
        const int a = tc_ & 1;
        const int colw = n0 + ch * CH_COLS, row_w0 = m0 + lg * 32;
        if (4 * lane < CH_COLS) {                              // this warp's bias slice -> shared memory (zeros without a bias)
          float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
          const int c = colw + 4 * lane;
          if (bias) {
            if (c + 3 < Ncols) b = __ldg(reinterpret_cast<const float4*>(bias + c));
            else {
              if (c < Ncols) b.x = bias[c];
              if (c + 1 < Ncols) b.y = bias[c + 1];
              if (c + 2 < Ncols) b.z = bias[c + 2];
            }
          }
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sbias + 16u * lane), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
        }
        __syncwarp();
        mbar_wait(&acc_full[a], (tc_ >> 1) & 1);
        tc_fence_after();
        if (warp == 2 && lane == 0) TRACE(2, 2 * tc_);
        const uint32_t tacc = tmem_base + (uint32_t)(a * BN) + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * CH_COLS);
        uint32_t r[32];
        if (colw < Ncols) tmem_ld32_nowait(tacc, r);
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          const int col = colw + 32 * k;
          const bool live = col < Ncols;                       // warp-uniform
          if (live) tmem_ld_wait_all();
          if (k == NCH - 1) {   // every TMEM load of this warp has landed: hand the accumulator stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (rank == 0) mbar_arrive(&acc_empty[a]);
              else mbar_arrive_cta(&acc_empty[a], 0);
            }
          }
          if (!live) continue;
          float o[32];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 b;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "r"(sbias + (uint32_t)((32 * k + 4 * q) * 4)) : "memory");
            o[4 * q] = fmaf(__uint_as_float(r[4 * q]), alpha, b.x);
            o[4 * q + 1] = fmaf(__uint_as_float(r[4 * q + 1]), alpha, b.y);
            o[4 * q + 2] = fmaf(__uint_as_float(r[4 * q + 2]), alpha, b.z);
            o[4 * q + 3] = fmaf(__uint_as_float(r[4 * q + 3]), alpha, b.w);
          }
          if (k + 1 < NCH && col + 32 < Ncols) tmem_ld32_nowait(tacc + 32 * (k + 1), r);   // r is free: overlaps the stores below
          if (relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j] = fmaxf(o[j], 0.f);
          }
          if (c_bf16) {
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const __nv_bfloat162 h = __floats2bfloat162_rn(o[2 * j], o[2 * j + 1]);
              pk[j] = *reinterpret_cast<const uint32_t*>(&h);
            }
            const uint32_t box = stg + (nbox & 1u) * 2048u;
            if (lane == 0) tma_store_wait_read<1>();           // the store that last read this box has finished with it
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 4; ++q)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(box + my_box_row + (((uint32_t)q ^ swz) << 4)), "r"(pk[4 * q]),
                           "r"(pk[4 * q + 1]), "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                           : "memory");
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) tma_store_2d(&map_c, box, col, row_w0);
            ++nbox;
          } else {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              if (col + 16 * hh < Ncols) {
                const uint32_t box = stg + (nbox & 1u) * 2048u;
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(box + my_box_row + (((uint32_t)q ^ swz) << 4)),
                               "f"(o[16 * hh + 4 * q]), "f"(o[16 * hh + 4 * q + 1]), "f"(o[16 * hh + 4 * q + 2]), "f"(o[16 * hh + 4 * q + 3])
                               : "memory");
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) tma_store_2d(&map_c, box, col + 16 * hh, row_w0);
                ++nbox;
              }
            }
          }
        }
        if (warp == 2 && lane == 0) TRACE(2, 2 * tc_ + 1);
      }
      if (lane == 0) tma_store_wait_read<0>();                 // shared memory must outlive the last reads
    } else if (t.lean) {
      // ---- short form for one batch with dense output (and residual) rows: the general form below spends ~300 instructions per
      // 32-column chunk on row maps, 64-bit offsets and predicates, which made K <= 512 products epilogue-bound (8.4 k clk per
      // 128 x 256 tile against 4.1 k clk of UMMAs).  Here offsets are 32-bit, the tile's bias slices are fetched before the
      // accumulator is waited for, and the next chunk's TMEM load is in flight while this chunk is transposed and stored.
      constexpr int NCH = CH_COLS / 32;
      const uint32_t ldc = (uint32_t)g.c.si, ldr = (uint32_t)g.r.si;
      char* const Cb = reinterpret_cast<char*>(g.C);
      const char* const Rb = reinterpret_cast<const char*>(g.R);
      __nv_bfloat16* const C2 = reinterpret_cast<__nv_bfloat16*>(g.C2);
      for (int tile = unit0; tile < t.total_tiles; tile += ustep, ++tc_) {
        int z = 0, rem = tile;
        uint32_t cz0 = 0, rz0 = 0;
        const float* bias = g.bias;
        if (g.Z != 1) {                                 // batch offsets once per tile (host check: all below 2^31)
          z = tile / per_z; rem = tile - z * per_z;
          cz0 = (uint32_t)g.cz.off(z, g.zin, g.zmid);
          if (has_r) rz0 = (uint32_t)g.rz.off(z, g.zin, g.zmid);
          if (bias) bias += g.bz.off(z, g.zin, g.zmid);
        }
        const int tm = rem / t.tiles_n, nt = rem - tm * t.tiles_n;
        const int m0 = (NCTA * tm + rank) * BM, n0 = nt * t.bn_valid;
        const int Ntot = min(g.N, n0 + t.bn_valid);
        const int a = tc_ & 1;
        const int colw = n0 + ch * CH_COLS;            // first column of this warp
        const int row_w0 = m0 + lg * 32, row0 = row_w0 + sub_r;   // phase-2 rows of this lane: row0 + 4 * rr
        float4 bvs[NCH];
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          bvs[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias && colw + 32 * k + 32 <= Ntot) bvs[k] = __ldg(reinterpret_cast<const float4*>(bias + colw + 32 * k + sub_c));
        }
        const uint32_t co0 = cz0 + (uint32_t)row0 * ldc + (uint32_t)(colw + sub_c);
        const uint32_t ro0 = rz0 + (uint32_t)row0 * ldr + (uint32_t)(colw + sub_c);
        mbar_wait(&acc_full[a], (tc_ >> 1) & 1);
        tc_fence_after();
        if (warp == 2 && lane == 0) TRACE(2, 2 * tc_);
        const uint32_t tacc = tmem_base + (uint32_t)(a * BN) + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * CH_COLS);
        uint32_t r[32];
        if (colw < Ntot) tmem_ld32_nowait(tacc, r);
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          const int col = colw + 32 * k;
          const bool live = col < Ntot;                // warp-uniform
          const bool vec_chunk = col + 32 <= Ntot;
          // this chunk's residual rows are requested before its accumulator columns are waited for
          uint4 xr[8];
          if (has_r && vec_chunk) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              xr[rr] = make_uint4(0u, 0u, 0u, 0u);
              if (row0 + 4 * rr < Mtot) {
                const uint32_t ro = ro0 + (uint32_t)(4 * rr) * ldr + (uint32_t)(32 * k);
                if (r_bf16) {
                  const uint2 u = *reinterpret_cast<const uint2*>(Rb + (size_t)ro * 2);
                  xr[rr].x = u.x; xr[rr].y = u.y;
                } else {
                  xr[rr] = *reinterpret_cast<const uint4*>(Rb + (size_t)ro * 4);
                }
              }
            }
          }
          if (live) tmem_ld_wait_all();
          if (k == NCH - 1) {   // every TMEM load of this warp has landed: hand the accumulator stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (rank == 0) mbar_arrive(&acc_empty[a]);
              else mbar_arrive_cta(&acc_empty[a], 0);
            }
          }
          if (!live) continue;
#pragma unroll
          for (int q = 0; q < 8; ++q)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stg_w + 16 * q), "r"(r[4 * q]), "r"(r[4 * q + 1]),
                         "r"(r[4 * q + 2]), "r"(r[4 * q + 3])
                         : "memory");
          if (k + 1 < NCH && col + 32 < Ntot) tmem_ld32_nowait(tacc + 32 * (k + 1), r);   // r was consumed by the stores above
          __syncwarp();
          if (vec_chunk) {
            float4 v[8];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                           : "=f"(v[rr].x), "=f"(v[rr].y), "=f"(v[rr].z), "=f"(v[rr].w)
                           : "r"(stg_r + (uint32_t)(rr * 4 * STG_LD * 4))
                           : "memory");
            const float4 bv = bvs[k];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              v[rr].x = fmaf(v[rr].x, alpha, bv.x); v[rr].y = fmaf(v[rr].y, alpha, bv.y);
              v[rr].z = fmaf(v[rr].z, alpha, bv.z); v[rr].w = fmaf(v[rr].w, alpha, bv.w);
            }
            if (relu) {
#pragma unroll
              for (int rr = 0; rr < 8; ++rr) {
                v[rr].x = fmaxf(v[rr].x, 0.f); v[rr].y = fmaxf(v[rr].y, 0.f);
                v[rr].z = fmaxf(v[rr].z, 0.f); v[rr].w = fmaxf(v[rr].w, 0.f);
              }
            }
            if (has_r) {
              if (r_bf16) {
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) {
                  v[rr].x += __uint_as_float(xr[rr].x << 16); v[rr].y += __uint_as_float(xr[rr].x & 0xffff0000u);
                  v[rr].z += __uint_as_float(xr[rr].y << 16); v[rr].w += __uint_as_float(xr[rr].y & 0xffff0000u);
                }
              } else {
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) {
                  v[rr].x += __uint_as_float(xr[rr].x); v[rr].y += __uint_as_float(xr[rr].y);
                  v[rr].z += __uint_as_float(xr[rr].z); v[rr].w += __uint_as_float(xr[rr].w);
                }
              }
            }
            if (c_bf16) {
#pragma unroll
              for (int rr = 0; rr < 8; ++rr)
                if (row0 + 4 * rr < Mtot)
                  st4(reinterpret_cast<__nv_bfloat16*>(Cb) + (co0 + (uint32_t)(4 * rr) * ldc + (uint32_t)(32 * k)), v[rr]);
            } else {
#pragma unroll
              for (int rr = 0; rr < 8; ++rr)
                if (row0 + 4 * rr < Mtot)
                  st4(reinterpret_cast<float*>(Cb) + (co0 + (uint32_t)(4 * rr) * ldc + (uint32_t)(32 * k)), v[rr]);
            }
            if (C2) {
#pragma unroll
              for (int rr = 0; rr < 8; ++rr)
                if (row0 + 4 * rr < Mtot) st4(C2 + (co0 + (uint32_t)(4 * rr) * ldc + (uint32_t)(32 * k)), v[rr]);
            }
          } else {   // ragged last chunk of the matrix: one column per lane, row by row
            const int cc = col + lane;
            const float bs = (bias && cc < Ntot) ? bias[cc] : 0.f;
            for (int lr = 0; lr < 32; ++lr) {
              const int row = row_w0 + lr;
              if (row < Mtot && cc < Ntot) {
                float sv;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(sv) : "r"(stg + (uint32_t)((lr * STG_LD + lane) * 4)) : "memory");
                float o = fmaf(sv, alpha, bs);
                if (relu) o = fmaxf(o, 0.f);
                const size_t ro = (size_t)(rz0 + (uint32_t)row * ldr + (uint32_t)cc), co = (size_t)(cz0 + (uint32_t)row * ldc + (uint32_t)cc);
                if (has_r)
                  o += r_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(Rb)[ro]) : reinterpret_cast<const float*>(Rb)[ro];
                if (c_bf16) reinterpret_cast<__nv_bfloat16*>(Cb)[co] = __float2bfloat16_rn(o);
                else reinterpret_cast<float*>(Cb)[co] = o;
                if (C2) C2[co] = __float2bfloat16_rn(o);
              }
            }
          }
          __syncwarp();
        }
        if (warp == 2 && lane == 0) TRACE(2, 2 * tc_ + 1);
      }
    } else
    for (int tile = unit0; tile < t.total_tiles; tile += ustep, ++tc_) {
      const int z = tile / per_z, rem = tile - z * per_z;
      const int m0 = (NCTA * (rem / t.tiles_n) + rank) * BM, n0 = (rem % t.tiles_n) * t.bn_valid;
      const int Ntot = min(g.N, n0 + t.bn_valid);  // first column this tile does not own
      const int a = tc_ & 1;
      const float* bias = g.bias ? g.bias + g.bz.off(z, g.zin, g.zmid) : nullptr;
      const int row_w0 = m0 + lg * 32;  // first row of this warp
      // lane i computes the element offsets of row row_w0 + i once per tile; the phase-2 owner of a row
      // fetches them with shuffles (no integer division inside the chunk loop)
      const int my_row = row_w0 + lane;
      long long co_l = 0, ro_l = 0;
      if (my_row < Mtot) {
        co_l = g.cz.off(z, g.zin, g.zmid) + g.c.off(my_row);
        if (has_r) ro_l = g.rz.off(z, g.zin, g.zmid) + g.r.off(my_row);
      }
      long long co_r[8], ro_r[8];
      bool ok_r[8];
#pragma unroll
      for (int rr = 0; rr < 8; ++rr) {
        const int lr = rr * 4 + sub_r;
        co_r[rr] = __shfl_sync(0xffffffffu, co_l, lr);
        ro_r[rr] = __shfl_sync(0xffffffffu, ro_l, lr);
        ok_r[rr] = row_w0 + lr < Mtot;
      }
      mbar_wait(&acc_full[a], (tc_ >> 1) & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) TRACE(2, 2 * tc_);
      const uint32_t tacc = tmem_base + (uint32_t)(a * BN) + ((uint32_t)(lg * 32) << 16);
#pragma unroll 1
      for (int cj = 0; cj < CH_COLS; cj += 32) {
        const int c0 = ch * CH_COLS + cj;
        const int col = n0 + c0;
        const bool live = col < Ntot;  // warp-uniform
        uint32_t r[32];
        if (live) tmem_ld32(tacc + c0, r);
        if (cj + 32 >= CH_COLS) {  // last chunk of this warp: hand the accumulator stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (rank == 0) mbar_arrive(&acc_empty[a]);
            else mbar_arrive_cta(&acc_empty[a], 0);
          }
        }
        if (!live) continue;
#pragma unroll
        for (int q = 0; q < 8; ++q)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stg_w + 16 * q), "r"(r[4 * q]), "r"(r[4 * q + 1]),
                       "r"(r[4 * q + 2]), "r"(r[4 * q + 3])
                       : "memory");
        __syncwarp();
        if (vec_c && col + 32 <= Ntot) {
          const int cc = col + sub_c;
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias) bv = __ldg(reinterpret_cast<const float4*>(bias + cc));
          float4 v[8], x[8];
          if (has_r) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              x[rr] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ok_r[rr])
                x[rr] = r_bf16 ? ld4(reinterpret_cast<const __nv_bfloat16*>(g.R) + ro_r[rr] + cc)
                               : ld4(reinterpret_cast<const float*>(g.R) + ro_r[rr] + cc);
            }
          }
#pragma unroll
          for (int rr = 0; rr < 8; ++rr)
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v[rr].x), "=f"(v[rr].y), "=f"(v[rr].z), "=f"(v[rr].w)
                         : "r"(stg_r + (uint32_t)(rr * 4 * STG_LD * 4))
                         : "memory");
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            v[rr].x = fmaf(v[rr].x, alpha, bv.x); v[rr].y = fmaf(v[rr].y, alpha, bv.y);
            v[rr].z = fmaf(v[rr].z, alpha, bv.z); v[rr].w = fmaf(v[rr].w, alpha, bv.w);
            if (g.accumulate && ok_r[rr]) {   // earlier partial products of a split fp32 product (fp32 C)
              const float4 c0 = ld4(reinterpret_cast<const float*>(g.C) + co_r[rr] + cc);
              v[rr].x += c0.x; v[rr].y += c0.y; v[rr].z += c0.z; v[rr].w += c0.w;
            }
            if (relu) {
              v[rr].x = fmaxf(v[rr].x, 0.f); v[rr].y = fmaxf(v[rr].y, 0.f);
              v[rr].z = fmaxf(v[rr].z, 0.f); v[rr].w = fmaxf(v[rr].w, 0.f);
            }
            if (has_r) { v[rr].x += x[rr].x; v[rr].y += x[rr].y; v[rr].z += x[rr].z; v[rr].w += x[rr].w; }
          }
          if (c_bf16) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              if (ok_r[rr]) st4(reinterpret_cast<__nv_bfloat16*>(g.C) + co_r[rr] + cc, v[rr]);
          } else {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              if (ok_r[rr]) st4(reinterpret_cast<float*>(g.C) + co_r[rr] + cc, v[rr]);
          }
          if (g.C2) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              if (ok_r[rr]) st4(reinterpret_cast<__nv_bfloat16*>(g.C2) + co_r[rr] + cc, v[rr]);
          }
        } else {
          const int cc = col + lane;
          const float bv = (bias && cc < Ntot) ? bias[cc] : 0.f;
          for (int lr = 0; lr < 32; ++lr) {
            const long long co = __shfl_sync(0xffffffffu, co_l, lr), ro = __shfl_sync(0xffffffffu, ro_l, lr);
            if (row_w0 + lr < Mtot && cc < Ntot) {
              float sv;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(sv) : "r"(stg + (uint32_t)((lr * STG_LD + lane) * 4)) : "memory");
              float v = fmaf(sv, alpha, bv);
              if (g.accumulate) v += reinterpret_cast<const float*>(g.C)[co + cc];
              if (relu) v = fmaxf(v, 0.f);
              if (has_r)
                v += r_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g.R)[ro + cc])
                            : reinterpret_cast<const float*>(g.R)[ro + cc];
              if (c_bf16) reinterpret_cast<__nv_bfloat16*>(g.C)[co + cc] = __float2bfloat16_rn(v);
              else reinterpret_cast<float*>(g.C)[co + cc] = v;
              if (g.C2) reinterpret_cast<__nv_bfloat16*>(g.C2)[co + cc] = __float2bfloat16_rn(v);
            }
          }
        }
        __syncwarp();
      }
      if (warp == 2 && lane == 0) TRACE(2, 2 * tc_ + 1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (NCTA == 2) cluster_sync_all();
  if (warp == 1) {
    __syncwarp();
    if (NCTA == 2) tmem_dealloc2(tmem_base, 2 * BN);
    else tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------------------------- host side
template <int BN, int STAGES, bool TF32>
int launch(const TcArgs& t, const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, const CUtensorMap& ma2,
           const CUtensorMap& mw2, cudaStream_t st) {
  auto kern = gemm_tc_kernel<BN, STAGES, TF32>;
  constexpr int smem = Smem<BN, STAGES>::TOTAL;
  static bool configured = false;  // per instantiation
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  TcArgs tt = t;
  tt.tiles_m = (t.g.M + BM - 1) / BM;
  if (!tt.w3d) { tt.bn_valid = BN; tt.w_stage_bytes = BN * SLAB_BYTES; }
  tt.tiles_n = (t.g.N + tt.bn_valid - 1) / tt.bn_valid;
  tt.total_tiles = tt.tiles_m * tt.tiles_n * t.g.Z;
  int grid = tt.total_tiles < num_sms() ? tt.total_tiles : num_sms();
  if (t.g.max_ctas > 0 && grid > t.g.max_ctas) grid = t.g.max_ctas;
  ProfScope ps(t.g.tag, st);
  BIST_LAUNCH(kern, grid, 320, smem, st, ma, mw, mc, ma2, mw2, tt);
  BIST_LAUNCHED();
  return 0;
}

// CTA-pair form (bf16, BN = 256, plain W tiles): units are pairs of vertically adjacent row tiles, clusters of two CTAs.
int launch_pair(const TcArgs& t, const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, cudaStream_t st) {
  auto kern = gemm_tc_kernel<256, 5, false, 2>;
  constexpr int smem = Smem<256, 5, 2>::TOTAL;
  static bool configured = false;
  if (!configured) {
    BIST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  TcArgs tt = t;
  tt.tiles_m = ((t.g.M + BM - 1) / BM + 1) / 2;
  tt.bn_valid = 256; tt.w_stage_bytes = 128 * SLAB_BYTES;
  tt.tiles_n = (t.g.N + 255) / 256;
  tt.total_tiles = tt.tiles_m * tt.tiles_n * t.g.Z;
  int pairs = tt.total_tiles < num_sms() / 2 ? tt.total_tiles : num_sms() / 2;
  if (t.g.max_ctas > 1 && pairs > t.g.max_ctas / 2) pairs = t.g.max_ctas / 2;
  ProfScope ps(t.g.tag, st);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  at[1].id = cudaLaunchAttributeClusterDimension;
  at[1].val.clusterDim.x = 2; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 2;
  BIST_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mw, mc, ma, mw, tt));
  BIST_LAUNCHED();
  return 0;
}

// Can the (row offset, column offset) decomposition of every batch offset be used?  Returns the
// number of tensor rows needed, or -1.
long long batch_rows(const BatchMap& bm, int Z, int zin, int zmid, long long ld, int K, int rows_per_batch) {
  long long max_row = 0;
  for (int z = 0; z < Z; ++z) {
    const long long off = bm.off(z, zin, zmid);
    if (off < 0 || off % ld + K > ld) return -1;
    max_row = off / ld > max_row ? off / ld : max_row;
  }
  return max_row + rows_per_batch;
}

}  // namespace

#ifdef BIST_GEMM_TRACE
extern "C" int bist_debug_gemm_trace(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_trace, sizeof(g_trace)) == cudaSuccess ? 0 : 1;
}
#endif

int gemm_tc(const GemmArgs& g, cudaStream_t st, bool* taken) {
  *taken = false;
  static const bool disabled = getenv("BIST_DISABLE_TC") != nullptr;  // debugging aid: force the CUDA-core GEMM
  if (disabled) return 0;
  const bool bf = g.a_dtype == BIST_BF16 && g.w_dtype == BIST_BF16;
  const bool tf = g.a_dtype == BIST_F32 && g.w_dtype == BIST_F32 && g.allow_tf32;
  if (!bf && !tf) return 0;
  const int esz = bf ? 2 : 4, bke = SLAB_BYTES / esz;
  if (g.K % bke != 0 || g.M < 1 || g.N < 8 || g.Z > 65535) return 0;
  if (g.a.inner != kNoSplit || g.w.inner != kNoSplit) return 0;  // operands must be dense row-major
  const long long a_ld = g.a.si, w_ld = g.w.si;
  if (a_ld < g.K || w_ld < g.K || (a_ld * esz) % 16 || (w_ld * esz) % 16) return 0;
  if (((uintptr_t)g.A % 16) || ((uintptr_t)g.W % 16)) return 0;
  if ((long long)g.M * g.N * g.Z < 64 * 64) return 0;  // tiny products stay on the CUDA cores
  const bool regions = g.wr_T > 0;
  if (regions && (!bf || g.wr_T > 256 || g.wr_S < 1 || g.N != g.wr_T * g.wr_S || w_ld != g.K)) return 0;
  const long long a_rows = batch_rows(g.az, g.Z, g.zin, g.zmid, a_ld, g.K, g.M);
  long long w_rows = 0;
  if (regions) {  // every batch offset must address a whole [wr_T, wr_S] block
    const long long blk = (long long)g.wr_T * g.wr_S * w_ld;
    for (int z = 0; z < g.Z; ++z) {
      const long long off = g.wz.off(z, g.zin, g.zmid);
      if (off < 0 || off % ((long long)g.wr_S * w_ld) != 0) return 0;
      w_rows = (off + blk) / w_ld > w_rows ? (off + blk) / w_ld : w_rows;
    }
  } else {
    w_rows = batch_rows(g.wz, g.Z, g.zin, g.zmid, w_ld, g.K, g.N);
  }
  if (a_rows < 0 || w_rows < 0) return 0;
  *taken = true;

  TcArgs t;
  t.g = g;
  t.w3d = 0; t.s_box = 0; t.bn_valid = 0; t.w_stage_bytes = 0; t.w_bt_elems = 1;
  t.a_ld = a_ld; t.w_ld = w_ld; t.nkb = g.K / bke;
  const int csz = g.c_dtype == BIST_BF16 ? 2 : 4;
  auto al = [&](long long v, int sz) { return (v * sz) % 16 == 0; };
  t.vec_c = ((uintptr_t)g.C % 16 == 0) && al(g.c.si, csz) && al(g.c.so, csz) && al(g.cz.s0, csz) && al(g.cz.s1, csz) && al(g.cz.s2, csz) &&
            (!g.bias || (((uintptr_t)g.bias % 16 == 0) && g.bz.s0 % 4 == 0 && g.bz.s1 % 4 == 0 && g.bz.s2 % 4 == 0));
  if (g.R) {
    const int rsz = g.r_dtype == BIST_BF16 ? 2 : 4;
    t.vec_c = t.vec_c && ((uintptr_t)g.R % 16 == 0) && al(g.r.si, rsz) && al(g.r.so, rsz) && al(g.rz.s0, rsz) && al(g.rz.s1, rsz) && al(g.rz.s2, rsz);
  }
  {
    static const bool lean_off = getenv("BIST_GEMM_GENERAL_EPILOGUE") != nullptr;   // debugging aid: A/B against the general epilogue
    auto dense = [](const RowMap& m) { return m.inner == kNoSplit && m.mod == kNoSplit; };
    const long long lim = 1ll << 31;
    // batches: the largest batch offset of C (and R) must keep element offsets below 2^31 as well
    long long cmax = 0, rmax = 0;
    bool zok = g.Z <= 4096;
    for (int z = 0; zok && z < g.Z; ++z) {
      const long long co = g.cz.off(z, g.zin, g.zmid), ro = g.R ? g.rz.off(z, g.zin, g.zmid) : 0;
      if (co < 0 || ro < 0) zok = false;
      cmax = co > cmax ? co : cmax; rmax = ro > rmax ? ro : rmax;
    }
    t.lean = !lean_off && t.vec_c && zok && !g.accumulate && dense(g.c) && g.c.si >= g.N && cmax + (long long)g.M * g.c.si < lim &&
             (!g.R || (dense(g.r) && g.r.si >= g.N && rmax + (long long)g.M * g.r.si < lim));
  }
  {
    static const bool tma_off = getenv("BIST_GEMM_NO_TMA_STORE") != nullptr;   // debugging aid: A/B against the transposing epilogues
    t.tma_c = !tma_off && t.lean && g.Z == 1 && !regions && !g.R && !g.C2 && ((uintptr_t)g.C % 16 == 0) && (g.c.si * csz) % 16 == 0;
  }
  // tile width: minimise (waves over the SMs) x (per-tile cost ~ BN + fixed overhead)
  int bn = 64;
  {
    const long long tm = (g.M + BM - 1) / BM, sms = num_sms();
    long long best = -1;
    for (int cand : {256, 128, 64}) {
      if (cand > 64 && g.N <= cand / 2) continue;
      const long long tiles = tm * ((g.N + cand - 1) / cand) * g.Z;
      const long long cost = ((tiles + sms - 1) / sms) * (cand + 64);
      if (best < 0 || cost < best) { best = cost; bn = cand; }
    }
  }
  CUtensorMap ma, mw, mc, ma2, mw2;
  t.npass = 1; t.split = 0; t.a_cat = 0; t.w_cat_rows = 0;
  if (g.kcat > 1) {   // C = A_0 W_0^T + A_1 W_1^T + ..: parts side by side in A's rows, W blocks one below the other
    if (g.A2 || g.W2 || regions || g.w_cat % w_ld != 0 || g.a_cat * (g.kcat - 1) + g.K > a_ld || g.a_cat % bke != 0) {
      *taken = false;
      return 0;
    }
    t.npass = g.kcat; t.a_cat = (int)g.a_cat; t.w_cat_rows = (int)(g.w_cat / w_ld);
    w_rows += (long long)(g.kcat - 1) * t.w_cat_rows;
  }
  if (g.A2 || g.W2) {   // split-tf32: the low halves mirror the layout of A and W
    if (!tf || !g.A2 || !g.W2 || regions || (((uintptr_t)g.A2 | (uintptr_t)g.W2) % 16)) { *taken = false; return 0; }
    t.npass = 3; t.split = 1;
  }
  BIST_TRY(make_map_2d(&ma, g.A, !bf, a_ld, a_rows, SLAB_BYTES / esz, BM, 128));
  ma2 = ma;
  if (t.npass == 3) BIST_TRY(make_map_2d(&ma2, g.A2, true, a_ld, a_rows, SLAB_BYTES / esz, BM, 128));
  if (t.tma_c) BIST_TRY(make_map_c(&mc, g.C, g.c_dtype != BIST_BF16, g.N, g.c.si, g.M));
  else mc = ma;   // never used, but a valid descriptor
  if (regions) {
    // N tiles own whole regions: s_box regions x wr_T rows each (<= BN rows of the MMA tile, the rest is never stored)
    const long long tm = (g.M + BM - 1) / BM, sms = num_sms();
    long long best = -1;
    for (int cand : {256, 128}) {
      const int sb = cand / g.wr_T < g.wr_S ? cand / g.wr_T : g.wr_S;
      if (sb < 1) continue;
      const long long tiles = tm * ((g.wr_S + sb - 1) / sb) * g.Z;
      const long long cost = ((tiles + sms - 1) / sms) * (cand + 64);
      if (best < 0 || cost < best) { best = cost; bn = cand; t.s_box = sb; }
    }
    t.w3d = 1;
    t.bn_valid = t.s_box * g.wr_T;
    t.w_stage_bytes = t.bn_valid * SLAB_BYTES;
    t.w_bt_elems = (long long)g.wr_S * w_ld;
    if (t.bn_valid % 4 != 0) t.vec_c = t.lean = 0;  // N tiles must start on a 16-byte boundary for the vector epilogue
    BIST_TRY(make_map_3d(&mw, g.W, false, w_ld, w_rows / g.wr_S, g.wr_S, (long long)g.wr_S * w_ld * esz, w_ld * esz, SLAB_BYTES / esz,
                         g.wr_T, t.s_box));
  } else {
    // min M for the CTA-pair form (0 = off).  With the TMA-store epilogue the single-CTA form is bound by what an SM can pull from
    // L2 (48 KB per K slab: ~7 k clk per K = 512 tile against 4.1 k clk of UMMAs); the pair pulls 32 KB: 44 -> 38 us at M = 40960,
    // N = 1024, K = 512, no difference at M = 16384
    static const int pair_rows = getenv("BIST_GEMM_2CTA") ? atoi(getenv("BIST_GEMM_2CTA")) : 24576;
    if (bf && bn == 256 && pair_rows > 0 && g.M >= pair_rows) {
      BIST_TRY(make_map_2d(&mw, g.W, false, w_ld, w_rows, SLAB_BYTES / esz, 128, 128));
      return launch_pair(t, ma, mw, mc, st);
    }
    BIST_TRY(make_map_2d(&mw, g.W, !bf, w_ld, w_rows, SLAB_BYTES / esz, bn, 128));
  }
  mw2 = mw;
  if (t.npass == 3) BIST_TRY(make_map_2d(&mw2, g.W2, true, w_ld, w_rows, SLAB_BYTES / esz, bn, 128));
  if (bf) {
    if (bn == 256) return launch<256, 3, false>(t, ma, mw, mc, ma2, mw2, st);
    if (bn == 128) return launch<128, 5, false>(t, ma, mw, mc, ma2, mw2, st);
    return launch<64, 7, false>(t, ma, mw, mc, ma2, mw2, st);
  }
  if (bn == 256) return launch<256, 3, true>(t, ma, mw, mc, ma2, mw2, st);
  if (bn == 128) return launch<128, 5, true>(t, ma, mw, mc, ma2, mw2, st);
  return launch<64, 7, true>(t, ma, mw, mc, ma2, mw2, st);
}

}  // namespace bist
