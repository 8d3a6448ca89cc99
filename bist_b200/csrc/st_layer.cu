// VidEncoderLayer4.forward (enc_st_combine == 'none') — /root/reference/model/encoder.py:172-201 with
// temporal2spatial (:109-139) and spatial2temporal (:141-170), eval mode.
//
// Formulation.  The reference copies the query block S (or T) times and projects every video
// row to keys and values, per direction and per block.  Here
//   * the query-side work (LayerNorm, Q projection, residual) is done once per (b,l) instead of once per copy;
//   * the key projection is folded into the query:  q_h.(W_k,h x + b_k,h) = (q_h W_k,h).x + const, and the
//     constant cancels in the softmax, so stage-1 scores are a product of the folded queries [Lq*h, d] with the
//     raw video rows, and the key bias never appears;
//   * the stage-1 output projection W_o1 (applied by the reference to every (b,g,l) row) is folded past stage 2:
//     with c1 = stage-1 context (before W_o1), out1 = x + b_o1 + c1 W_o1^T, so
//       stage-2 scores  = (q2_h W_k2,h W_o1) . c1  + terms constant over the keys (cancel in the softmax)
//       stage-2 context = (x + b_o1) W_v2^T + b_v2 + (sum_g p_g c1_g) (W_v2 W_o1)^T       (sum_g p_g = 1)
//     W_k2 W_o1 and W_v2 W_o1 are formed once at pack time in fp32.
// Only the stage-1 value projection touches every video row (one GEMM for both directions); the stage-1 and
// stage-2 query norms / projections / folds of both directions run as three batched launches.
#include <math.h>

#include "common.cuh"

namespace bist {
namespace {

struct PackedLayout {
  // matrix section, element offsets relative to the direction base
  size_t wqkv0, wo0, wq12x, wkt12, wv2f, wo2, w1, w2, dir_stride;
  size_t wv1cat, mat_elems;
  // fp32 section, float offsets relative to the direction base
  size_t bqkv0, bo0, bq12x, bo2, b1, b2, ln_a, ln_b, fdir_stride;
  size_t bv1cat, f_elems;
  size_t f_byte_off, scratch_byte_off, total_bytes;
};

PackedLayout packed_layout(int64_t d, int64_t dff, int n_dirs, int dtype) {
  PackedLayout L;
  size_t o = 0;
  auto m = [&](size_t n) { size_t r = o; o += n; return r; };
  const size_t dd = (size_t)d * d;
  L.wqkv0 = m(3 * dd); L.wo0 = m(dd);
  L.wq12x = m(3 * dd);  // W_q1 | W_q2 | W_v2          (one batched launch: q1, q2, x W_v2^T)
  L.wkt12 = m(2 * dd);  // (W_k1)^T/sqrt(dk) | (W_k2 W_o1)^T/sqrt(dk)   (one batched fold)
  L.wv2f = m(dd);       // W_v2 W_o1
  L.wo2 = m(dd);
  L.w1 = m((size_t)dff * d); L.w2 = m((size_t)d * dff);
  L.dir_stride = o;
  L.wv1cat = L.dir_stride * n_dirs;
  L.mat_elems = L.wv1cat + (size_t)n_dirs * dd;
  o = 0;
  L.bqkv0 = m(3 * d); L.bo0 = m(d);
  L.bq12x = m(3 * d);   // b_q1 | b_q2 | b_v2 + W_v2 b_o1
  L.bo2 = m(d); L.b1 = m(dff); L.b2 = m(d); L.ln_a = m(4 * d); L.ln_b = m(4 * d);
  L.fdir_stride = o;
  L.bv1cat = L.fdir_stride * n_dirs;
  L.f_elems = L.bv1cat + (size_t)n_dirs * d;
  L.f_byte_off = align_up(L.mat_elems * dtype_size(dtype), 256);
  L.scratch_byte_off = align_up(L.f_byte_off + L.f_elems * 4, 256);
  L.total_bytes = L.scratch_byte_off + 2 * dd * 4;  // pack-time scratch: W_o1^T and one fp32 product
  return L;
}

struct Workspace {
  // `act` buffers are GEMM operands (bf16 in bf16 mode), the rest fp32
  void *nbuf, *qkv0, *cbuf, *n12x, *q12x, *qf12, *u2, *ffh;  // act dtype
  float *sc0, *x, *ts, *sc1, *sc1b, *qf2f;                    // fp32
  void* xv;
  void* ctx1[2];
  // fp32 mode: hi / lo splits for the tensor-core products (gemm_x3.cu): the video tensor, the layer's packed matrices, and the
  // activation operand of the product at hand
  float *v_hi, *v_lo, *w_hi, *w_lo, *a_hi, *a_lo;
  size_t a_cap;
  size_t bytes;
  bool ok;
};

// Can stage 2 of both active directions run on the bf16 tensor-core kernel?
bool use_stage2_mma(int dtype, int64_t T, int64_t S, int64_t h, int64_t d, int nd, const int* kinds) {
  if (dtype != BIST_BF16) return false;
  for (int u = 0; u < nd; ++u)
    if (!stage2_mma_supported((int)(kinds[u] == 0 ? S : T), (int)h, (int)d)) return false;
  return true;
}

Workspace carve(void* base, size_t cap, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff,
                int64_t h, int nd, const int* kinds, int dtype) {
  Arena A(base, cap);
  Workspace W;
  const size_t Q = (size_t)B * Lq, e = dtype_size(dtype);
  auto f = [&](size_t n) { return (float*)A.take(n * 4); };
  auto a = [&](size_t n) { return A.take(n * e); };
  W.nbuf = a(nd * Q * d); W.qkv0 = a(nd * Q * 3 * d); W.cbuf = a(nd * Q * d);
  W.n12x = a(3 * nd * Q * d); W.q12x = a(3 * nd * Q * d); W.qf12 = a(2 * nd * Q * h * d); W.u2 = a(nd * Q * h * d);
  W.ffh = a(nd * Q * dff);
  W.sc0 = f(nd * Q * h * Lq); W.x = f(nd * Q * d); W.ts = f(nd * Q * d);
  W.sc1 = f(Q * h * T * S);  // one direction at a time ...
  W.sc1b = nd == 2 ? f(Q * h * T * S) : nullptr;  // ... or the second direction's, when it runs beside the first
  W.qf2f = (dtype == BIST_BF16 && !use_stage2_mma(dtype, T, S, h, d, nd, kinds)) ? f(nd * Q * h * d) : nullptr;
  W.xv = A.take((size_t)B * T * S * nd * d * e);
  for (int u = 0; u < 2; ++u) W.ctx1[u] = nullptr;
  for (int u = 0; u < nd; ++u) W.ctx1[u] = A.take((size_t)B * (kinds[u] == 0 ? S : T) * Lq * d * e);
  W.v_hi = W.v_lo = W.w_hi = W.w_lo = W.a_hi = W.a_lo = nullptr;
  W.a_cap = 0;
  if (dtype == BIST_F32 && fp32x3_enabled()) {
    const size_t nv = (size_t)B * T * S * d, nw = packed_layout(d, dff, 2, dtype).mat_elems;
    W.a_cap = nd * Q * (size_t)(h * d > dff ? (h * d > 3 * d ? h * d : 3 * d) : (dff > 3 * d ? dff : 3 * d));   // largest activation range one product reads
    W.v_hi = f(nv); W.v_lo = f(nv); W.w_hi = f(nw); W.w_lo = f(nw); W.a_hi = f(W.a_cap); W.a_lo = f(W.a_cap);
  }
  W.bytes = A.used + 256;
  W.ok = A.ok;
  return W;
}

int active_dirs(int n_dirs, int dir_mask, int* kinds) {
  int nd = 0;
  if (dir_mask & 1) kinds[nd++] = 0;
  if (dir_mask & 2) kinds[nd++] = 1;
  if (n_dirs == 1 && nd != 1) return -1;
  return nd;
}

// C[M,N] (fp32) = A[M,K] (fp32) . W[N,K]^T (fp32) (+ bias) on the CUDA-core GEMM (pack-time weight products).
int fp32_product(const float* A, const float* Wm, const float* bias, float* C, int M, int N, int K, cudaStream_t st) {
  GemmArgs g;
  g.A = A; g.W = Wm; g.C = C; g.bias = bias;
  g.M = M; g.N = N; g.K = K;
  g.a = rows_dense(K); g.w = rows_dense(K); g.c = rows_dense(N);
  g.tag = "pack";
  return gemm_simt(g, st);
}

// CTAs of the forked stage-1 value GEMM: the rest of the SMs stay free for the query-side chain it runs beside.
int xv_side_ctas() {
  static const int n = [] {
    const char* e = getenv("BIST_XV_CTAS");
    return e ? atoi(e) : 100;   // measured optimum at config 2 (sweep 56..148: 96-112 best)
  }();
  return n;
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

size_t bist_st_layer_packed_bytes(int64_t d, int64_t d_ff, int n_dirs, int dtype) {
  if (d <= 0 || d_ff <= 0 || (n_dirs != 1 && n_dirs != 2) || !dtype_ok(dtype)) return 0;
  return packed_layout(d, d_ff, n_dirs, dtype).total_bytes;
}

int bist_st_layer_pack(const bist_st_layer_params_t* P, void* packed, size_t packed_bytes, int64_t d, int64_t dff,
                       int64_t h, int n_dirs, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(P && packed, "pack: null pointer");
  BIST_CHECK_ARG(d > 0 && dff > 0 && h > 0 && d % h == 0 && (n_dirs == 1 || n_dirs == 2) && dtype_ok(dtype),
                 "pack: bad argument");
  BIST_CHECK_ARG(d % 4 == 0 && dff % 4 == 0, "pack: d and d_ff must be multiples of 4");
  const PackedLayout L = packed_layout(d, dff, n_dirs, dtype);
  if (packed_bytes < L.total_bytes) { set_error("pack: buffer too small"); return BIST_ERR_WORKSPACE; }
  const size_t e = dtype_size(dtype), dd = (size_t)d * d;
  char* mat = (char*)packed;
  float* fs = (float*)((char*)packed + L.f_byte_off);
  float* tmpT = (float*)((char*)packed + L.scratch_byte_off);  // W_o1^T, then products
  float* tmpP = tmpT + dd;
  const float rs = 1.f / sqrtf((float)(d / h));
  const int di = (int)d, fi = (int)dff;
  auto M = [&](size_t off) { return (void*)(mat + off * e); };
  for (int u = 0; u < n_dirs; ++u) {
    const int ab = 3 * u;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 4; ++j) BIST_CHECK_ARG(P->attn_w[ab + i][j] && P->attn_b[ab + i][j], "pack: null attn parameter");
    BIST_CHECK_ARG(P->ff_w1[u] && P->ff_b1[u] && P->ff_w2[u] && P->ff_b2[u], "pack: null ff parameter");
    const size_t mb = u * L.dir_stride, fb = u * L.fdir_stride;
    const float *Wk2 = P->attn_w[ab + 2][1], *Wv2 = P->attn_w[ab + 2][2], *Wo1 = P->attn_w[ab + 1][3];
    auto V = [&](const float* src, size_t off, int n) { return pack_matrix(src, fs + off, 1, n, 0, 1.f, BIST_F32, st); };
    // query self-attention (encoder.py:176 / :184)
    for (int j = 0; j < 3; ++j) {
      BIST_TRY(pack_matrix(P->attn_w[ab][j], M(mb + L.wqkv0 + j * dd), di, di, 0, 1.f, dtype, st));
      BIST_TRY(V(P->attn_b[ab][j], fb + L.bqkv0 + (size_t)j * d, di));
    }
    BIST_TRY(pack_matrix(P->attn_w[ab][3], M(mb + L.wo0), di, di, 0, 1.f, dtype, st));
    BIST_TRY(V(P->attn_b[ab][3], fb + L.bo0, di));
    // stage 1 / stage 2 query projections and the raw-x value term
    BIST_TRY(pack_matrix(P->attn_w[ab + 1][0], M(mb + L.wq12x), di, di, 0, 1.f, dtype, st));
    BIST_TRY(pack_matrix(P->attn_w[ab + 2][0], M(mb + L.wq12x + dd), di, di, 0, 1.f, dtype, st));
    BIST_TRY(pack_matrix(Wv2, M(mb + L.wq12x + 2 * dd), di, di, 0, 1.f, dtype, st));
    BIST_TRY(V(P->attn_b[ab + 1][0], fb + L.bq12x, di));
    BIST_TRY(V(P->attn_b[ab + 2][0], fb + L.bq12x + d, di));
    BIST_TRY(fp32_product(P->attn_b[ab + 1][3], Wv2, P->attn_b[ab + 2][2], fs + fb + L.bq12x + 2 * d, 1, di, di, st));  // b_v2 + W_v2 b_o1
    // folds: (W_k1)^T / sqrt(dk) and (W_k2 W_o1)^T / sqrt(dk);  W_v2 W_o1
    BIST_TRY(pack_matrix(P->attn_w[ab + 1][1], M(mb + L.wkt12), di, di, 1, rs, dtype, st));
    BIST_TRY(pack_matrix(Wo1, tmpT, di, di, 1, 1.f, BIST_F32, st));              // tmpT[i][o] = W_o1[o][i]
    BIST_TRY(fp32_product(Wk2, tmpT, nullptr, tmpP, di, di, di, st));             // tmpP = W_k2 W_o1
    BIST_TRY(pack_matrix(tmpP, M(mb + L.wkt12 + dd), di, di, 1, rs, dtype, st));
    BIST_TRY(fp32_product(Wv2, tmpT, nullptr, tmpP, di, di, di, st));             // tmpP = W_v2 W_o1
    BIST_TRY(pack_matrix(tmpP, M(mb + L.wv2f), di, di, 0, 1.f, dtype, st));
    BIST_TRY(pack_matrix(P->attn_w[ab + 1][2], M(L.wv1cat + (size_t)u * dd), di, di, 0, 1.f, dtype, st));
    BIST_TRY(V(P->attn_b[ab + 1][2], L.bv1cat + (size_t)u * d, di));
    BIST_TRY(pack_matrix(P->attn_w[ab + 2][3], M(mb + L.wo2), di, di, 0, 1.f, dtype, st));
    BIST_TRY(V(P->attn_b[ab + 2][3], fb + L.bo2, di));
    BIST_TRY(pack_matrix(P->ff_w1[u], M(mb + L.w1), fi, di, 0, 1.f, dtype, st));
    BIST_TRY(pack_matrix(P->ff_w2[u], M(mb + L.w2), di, fi, 0, 1.f, dtype, st));
    BIST_TRY(V(P->ff_b1[u], fb + L.b1, fi));
    BIST_TRY(V(P->ff_b2[u], fb + L.b2, di));
    for (int i = 0; i < 4; ++i) {
      BIST_CHECK_ARG(P->ln_a[4 * u + i] && P->ln_b[4 * u + i], "pack: null norm parameter");
      BIST_TRY(V(P->ln_a[4 * u + i], fb + L.ln_a + (size_t)i * d, di));
      BIST_TRY(V(P->ln_b[4 * u + i], fb + L.ln_b + (size_t)i * d, di));
    }
  }
  return 0;
}

size_t bist_st_layer_workspace_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t d_ff,
                                     int64_t h, int n_dirs, int dtype) {
  if (B <= 0 || T <= 0 || S <= 0 || Lq <= 0 || d <= 0 || d_ff <= 0 || h <= 0 || !dtype_ok(dtype)) return 0;
  int kinds[2] = {0, 1};
  const int nd = n_dirs == 2 ? 2 : 1;
  size_t best = 0;
  for (int k = 0; k < (nd == 1 ? 2 : 1); ++k) {  // a single-direction layer may be applied to either direction
    if (nd == 1) kinds[0] = k;
    Workspace W = carve(nullptr, (size_t)-1, B, T, S, Lq, d, d_ff, h, nd, kinds, dtype);
    best = W.bytes > best ? W.bytes : best;
  }
  return best;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------- forward
extern "C" int bist_st_layer_fwd(const float* in_t2s, const float* in_s2t, const void* vft,
                                 const uint8_t* query_mask, const uint8_t* temporal_mask, const void* packed,
                                 float* out_t2s, float* out_s2t, int64_t B, int64_t T, int64_t S, int64_t Lq,
                                 int64_t d, int64_t dff, int64_t h, int n_dirs, int dir_mask, float eps, void* ws,
                                 size_t ws_bytes, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(vft && query_mask && temporal_mask && packed && ws, "st_layer: null pointer");
  BIST_CHECK_ARG(B > 0 && T > 0 && S > 0 && Lq > 0 && d > 0 && dff > 0 && h > 0 && d % h == 0,
                 "st_layer: bad shape B=%lld T=%lld S=%lld Lq=%lld d=%lld h=%lld", (long long)B, (long long)T,
                 (long long)S, (long long)Lq, (long long)d, (long long)h);
  BIST_CHECK_ARG(d % 4 == 0 && dff % 4 == 0, "st_layer: d and d_ff must be multiples of 4");
  BIST_CHECK_ARG((n_dirs == 1 || n_dirs == 2) && dtype_ok(dtype) && (dir_mask & ~3) == 0 && dir_mask != 0,
                 "st_layer: bad n_dirs/dir_mask/dtype");
  BIST_CHECK_ARG(B * T * S < (1LL << 31) / 2 && B * Lq * h < (1LL << 31), "st_layer: problem too large");
  BIST_CHECK_ARG(B * S <= 65535 && 2 * B * h <= 65535, "st_layer: batch too large for one call (split B)");
  int kinds[2];
  const int nd = active_dirs(n_dirs, dir_mask, kinds);
  BIST_CHECK_ARG(nd > 0, "st_layer: a single-direction layer needs exactly one bit in dir_mask");
  const float* x_in[2];
  float* x_out[2];
  for (int u = 0; u < nd; ++u) {
    x_in[u] = kinds[u] == 0 ? in_t2s : in_s2t;
    x_out[u] = kinds[u] == 0 ? out_t2s : out_s2t;
    BIST_CHECK_ARG(x_in[u] && x_out[u], "st_layer: null stream pointer for an active direction");
  }
  const Workspace W = carve(ws, ws_bytes, B, T, S, Lq, d, dff, h, nd, kinds, dtype);
  if (!W.ok) { set_error("st_layer: workspace too small (%zu bytes given)", ws_bytes); return BIST_ERR_WORKSPACE; }

  const PackedLayout L = packed_layout(d, dff, n_dirs, dtype);
  const int pu0 = n_dirs == 2 ? kinds[0] : 0;  // packed index of slot 0; slot 1 is pu0 + 1
  const size_t e = dtype_size(dtype);
  const int adt = dtype;  // dtype of the activations that feed GEMMs
  const char* mat = (const char*)packed + (size_t)pu0 * L.dir_stride * e;
  const float* fs = (const float*)((const char*)packed + L.f_byte_off) + (size_t)pu0 * L.fdir_stride;
  auto Mw = [&](size_t off) { return (const void*)(mat + off * e); };
  auto act = [&](void* p, long long elems) { return (void*)((char*)p + elems * (long long)e); };
  const long long ms = (long long)L.dir_stride, fsd = (long long)L.fdir_stride;
  const int Q = (int)(B * Lq), dk = (int)(d / h), di = (int)d, hi = (int)h;
  const long long Qd = (long long)Q * d, dd = (long long)d * d;
  const bool s2mma = use_stage2_mma(dtype, T, S, h, d, nd, kinds);
  // fp32 mode: the products run on the tensor cores as three tf32 products over hi / lo splits (gemm_x3.cu).  The video tensor and
  // the layer's packed matrices are split once per call; an activation operand is split right before the product that reads it.
  const bool x3 = dtype == BIST_F32 && W.v_hi != nullptr;
  SplitRef sv, swt;
  if (x3) {
    sv.src = (const float*)vft; sv.n = B * T * S * d; sv.hi = W.v_hi; sv.lo = W.v_lo; sv.ready = 1;
    swt.src = (const float*)packed; swt.n = (long long)L.mat_elems; swt.hi = W.w_hi; swt.lo = W.w_lo; swt.ready = 1;
    BIST_TRY(split_tf32(sv.src, sv.hi, sv.lo, sv.n, st));
    BIST_TRY(split_tf32(swt.src, swt.hi, swt.lo, swt.n, st));
  }
  auto mm = [&](const GemmArgs& g, const void* a_base, long long a_n, cudaStream_t s_, int a_ready = 0) -> int {
    if (!x3) return gemm(g, s_);
    SplitRef sa;
    if (a_base == vft) sa = sv;
    else { sa.src = (const float*)a_base; sa.n = a_n; sa.hi = W.a_hi; sa.lo = W.a_lo; sa.ready = a_ready; }
    const bool w_is_video = (const char*)g.W >= (const char*)vft && (const char*)g.W < (const char*)vft + sv.n * 4;
    return gemm_fp32x3(g, sa, w_is_video ? sv : swt, s_);
  };

  // Batched (over the active directions) linear layer  C[u] = act(A[u] . W[u]^T + b[u]) (+ R[u]).
  auto lin = [&](const void* Ain, long long a_dir, size_t w_off, size_t b_off, void* Cout, long long c_dir, int c_dtype,
                 int N, int K, int relu, const float* R, long long r_dir, void* C2 = nullptr) {
    GemmArgs g;
    g.A = Ain; g.W = Mw(w_off); g.C = Cout; g.bias = fs + b_off; g.R = R; g.C2 = C2;
    g.M = Q; g.N = N; g.K = K;
    g.a = rows_dense(K); g.w = rows_dense(K); g.c = rows_dense(N); g.r = rows_dense(N);
    g.Z = nd; g.zin = 1;
    g.az = {a_dir, 0}; g.wz = {ms, 0}; g.cz = {c_dir, 0}; g.rz = {r_dir, 0}; g.bz = {fsd, 0};
    g.relu = relu; g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = c_dtype; g.tag = "query_side_gemm";
    return mm(g, Ain, (long long)(nd - 1) * a_dir + (long long)Q * K, st);
  };

  // The stage-1 values XV = vft . [Wv1]^T + bv1 (the one big GEMM of the block) depend only on the video tensor: launch
  // them on the side stream now, so they run beside the serial query-side chain below; joined before the first P.XV.
  const long long TS = T * S, xvld = (long long)nd * d;
  auto launch_xv = [&](cudaStream_t s) {
    GemmArgs g;
    g.A = vft; g.W = (const char*)packed + (L.wv1cat + (size_t)pu0 * d * d) * e; g.C = W.xv;
    g.bias = (const float*)((const char*)packed + L.f_byte_off) + L.bv1cat + (size_t)pu0 * d;
    g.M = (int)(B * TS); g.N = nd * di; g.K = di;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(xvld);
    g.a_dtype = dtype; g.w_dtype = dtype; g.c_dtype = dtype; g.tag = "xv_gemm";
    if (s != st) g.max_ctas = xv_side_ctas();
    return mm(g, vft, 0, s);
  };
  // bf16 mode, d_k = 64, Lq <= 16, T and S <= 128: stage 1 of both directions is ONE kernel that reads the video tensor once
  // and keeps projected values and scores on chip (stage1_fused.cu); otherwise the value GEMM + score GEMM + softmax.V chain.
  const bool fused1 = dtype == BIST_BF16 && stage1_fused_supported(B, T, S, Lq, d, h);
  SideStream* side = nullptr;
  bool xv_forked = false, xv_joined = false;
  if (overlap_enabled() && !fused1) {
    BIST_TRY(side_stream(&side));
    BIST_CUDA(cudaEventRecord(side->fork, st));
    BIST_CUDA(cudaStreamWaitEvent(side->s, side->fork, 0));
    const int rc = launch_xv(side->s);
    BIST_CUDA(cudaEventRecord(side->join, side->s));
    if (rc != 0) { cudaStreamWaitEvent(st, side->join, 0); return rc; }
    xv_forked = true;
  }

  // ---- self-attention over the query tokens (encoder.py:176 / :184) -------------------------------
  // the layer inputs of the two directions are separate (possibly aliased) tensors: address them as a batch
  // whose stride is the pointer difference (zero when both are ft['encoded_query'], decoder.py:109-110)
  const long long in_dir = nd == 2 ? (long long)(x_in[1] - x_in[0]) : 0;
  const long long out_dir = nd == 2 ? (long long)(x_out[1] - x_out[0]) : 0;
  // bf16 mode at the shapes the fused stage 1 takes: the query-side chain runs as three kernels that normalise their fp32 rows
  // themselves and finish with the attention / fold / ReLU epilogue (qside.cu) instead of ten launches.
  const bool fusedq = fused1 && s2mma && qside_supported(B, Lq, d, dff, h);
  const float *lnA[4][2], *lnB[4][2], *bq0[2], *bq12[2], *bf1[2];
  const void *wq0[2], *wq12[2], *wkt[2], *wf1[2];
  for (int u = 0; u < nd; ++u) {
    const float* fu = fs + (size_t)u * L.fdir_stride;
    for (int i = 0; i < 4; ++i) { lnA[i][u] = fu + L.ln_a + (size_t)i * d; lnB[i][u] = fu + L.ln_b + (size_t)i * d; }
    bq0[u] = fu + L.bqkv0; bq12[u] = fu + L.bq12x; bf1[u] = fu + L.b1;
    const char* mu = mat + (size_t)u * L.dir_stride * e;
    wq0[u] = mu + L.wqkv0 * e; wq12[u] = mu + L.wq12x * e; wkt[u] = mu + L.wkt12 * e; wf1[u] = mu + L.w1 * e;
  }
  BIST_TRY(layer_norm(x_in[0], fs + L.ln_a, fs + L.ln_b, W.nbuf, (int64_t)nd * Q, d, eps, BIST_F32, adt, st, Q, fsd, in_dir, true));
  if (fusedq) {
    BIST_TRY(qside_attn(W.nbuf, wq0, bq0, query_mask, W.cbuf, nd, B, Lq, d, dff, h, eps, st));
  } else {
  BIST_TRY(lin(W.nbuf, Qd, L.wqkv0, L.bqkv0, W.qkv0, 3 * Qd, adt, 3 * di, di, 0, nullptr, 0));
  {
    bool fused = false;
    if (dtype == BIST_BF16) BIST_TRY(self_attn_mma(W.qkv0, query_mask, (int)B, W.cbuf, (int)(nd * B), (int)Lq, hi, dk, st, &fused));
    if (!fused) {
      GemmArgs g;  // scores[u,b,l,hh,k] = q.k / sqrt(dk)
      g.A = W.qkv0; g.W = act(W.qkv0, d); g.C = W.sc0;
      g.M = (int)Lq; g.N = (int)Lq; g.K = dk;
      g.a = rows_dense(3 * d); g.w = rows_dense(3 * d); g.c = rows_dense(h * Lq);
      g.Z = (int)(nd * B * h); g.zin = hi;
      g.az = {Lq * 3 * d, dk}; g.wz = {Lq * 3 * d, dk}; g.cz = {Lq * h * Lq, Lq};
      g.alpha = 1.f / sqrtf((float)dk);
      g.a_dtype = adt; g.w_dtype = adt; g.tag = "scores0_gemm";
      BIST_TRY(gemm_simt(g, st));  // 16x16x64 products: CUDA cores
      SoftmaxPvArgs a;
      a.S = W.sc0; a.sg = rows_dense(Lq * h * Lq); a.s_l = h * Lq; a.s_h = Lq;
      a.mask = query_mask; a.mask_gmod = (int)B; a.m_b = Lq; a.m_l = 0;
      a.V = act(W.qkv0, 2 * d); a.v = rows_dense(3 * d); a.vg = rows_dense(Lq * 3 * d);
      a.ctx = W.cbuf; a.og = rows_dense(Lq * d); a.o_l = d;
      a.G = (int)(nd * B); a.Lq = (int)Lq; a.h = hi; a.dk = dk; a.nk = (int)Lq;
      a.v_dtype = adt; a.ctx_dtype = adt;
      BIST_TRY(softmax_pv(a, st));
    }
  }
  }
  // output projection + residual with the layer inputs (modules.py:44)
  // (the fused chain also takes a bf16 copy of x: its kernels load their rows by TMA and normalise them in shared memory)
  BIST_TRY(lin(W.cbuf, Qd, L.wo0, L.bo0, W.x, Qd, BIST_F32, di, di, 0, x_in[0], in_dir, fusedq ? act(W.n12x, 2 * nd * Qd) : nullptr));

  // ---- query side of stage 1 and stage 2 in three launches ------------------------------------------------------
  // n12x = [LN1(x) | LN2(x) | x] (encoder.py:121,130 / :148,161 norm the same x with different gains)
  void* xcopy = act(W.n12x, 2 * nd * Qd);
  if (fusedq) {
    BIST_TRY(qside_fold(xcopy, lnA[1], lnB[1], lnA[2], lnB[2], wq12, bq12, wkt, W.qf12, act(W.q12x, 2 * nd * Qd), nd, B, Lq, d, dff, h, eps, st));
  } else {
  BIST_TRY(layer_norm2(W.x, fs + L.ln_a + d, fs + L.ln_b + d, fs + L.ln_a + 2 * d, fs + L.ln_b + 2 * d, W.n12x,
                       act(W.n12x, nd * Qd), xcopy, (int64_t)nd * Q, d, eps, adt, st, Q, fsd));
  {
    GemmArgs g;  // q12x[which][u] = n12x[which][u] . {W_q1 | W_q2 | W_v2}^T + {b_q1 | b_q2 | b_v2 + W_v2 b_o1}
    g.A = W.n12x; g.W = Mw(L.wq12x); g.C = W.q12x; g.bias = fs + L.bq12x;
    g.M = Q; g.N = di; g.K = di;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(d);
    g.Z = 3 * nd; g.zin = nd;  // z = which * nd + u
    g.az = {nd * Qd, Qd}; g.wz = {dd, ms}; g.cz = {nd * Qd, Qd}; g.bz = {d, fsd};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "query_side_gemm";
    BIST_TRY(mm(g, W.n12x, 3 * nd * Qd, st));
  }
  {
    // Query-side folds of the key projections: qf[which][u][q,hh,:] = q[which][u][q, hh*dk:(hh+1)*dk] . WkT[:, hh*dk:(hh+1)*dk]^T
    GemmArgs g;
    g.A = W.q12x; g.W = Mw(L.wkt12); g.C = W.qf12;
    g.M = Q; g.N = di; g.K = dk;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(h * d);
    g.zin = hi; g.zmid = nd;  // z = (which * nd + u) * h + hh
    g.az = {nd * Qd, dk, Qd}; g.wz = {dd, dk, ms}; g.cz = {nd * Qd * h, d, Qd * h};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "fold_gemm";
    if (dtype == BIST_BF16 && !s2mma) {  // the CUDA-core stage-2 kernels want fp32 folded queries
      g.Z = nd * hi;
      BIST_TRY(gemm(g, st));
      g.A = act(W.q12x, nd * Qd); g.W = Mw(L.wkt12 + dd); g.C = W.qf2f; g.c_dtype = BIST_F32;
      BIST_TRY(gemm(g, st));
    } else {
      g.Z = 2 * nd * hi;
      BIST_TRY(mm(g, W.q12x, 2 * nd * Qd, st));
    }
  }
  }

  // ---- stage 1 (encoder.py:111-121 / :143-148) ------------------------------------------------------
  if (fused1) {
    const void *qfp[2], *wvp[2];
    const float* bvp[2];
    void* c1p[2];
    for (int u = 0; u < nd; ++u) {
      qfp[u] = act(W.qf12, (long long)u * Qd * h);
      wvp[u] = (const char*)packed + (L.wv1cat + (size_t)(pu0 + u) * d * d) * e;
      bvp[u] = (const float*)((const char*)packed + L.f_byte_off) + L.bv1cat + (size_t)(pu0 + u) * d;
      c1p[u] = W.ctx1[u];
    }
    BIST_TRY(stage1_fused(vft, temporal_mask, nd, kinds, qfp, wvp, bvp, c1p, B, T, S, Lq, d, h, nullptr, st));
    if (overlap_enabled() && nd == 2 && dir_overlap_enabled()) BIST_TRY(side_stream(&side));
  } else if (!xv_forked) {
    BIST_TRY(launch_xv(st));
  }
  // The two directions are independent from here to the stage-2 outputs: the second one's kernels can run on the
  // side stream (behind the value GEMM it needs anyway) beside the first one's, with its own score buffer.
  // fp32 tensor-core mode: the folded stage-1 queries of both directions are split once, before the second direction forks
  if (x3 && !fused1) BIST_TRY(split_tf32((const float*)W.qf12, W.a_hi, W.a_lo, (long long)nd * Qd * h, st));
  const bool dir_ovl = (fused1 ? side != nullptr : xv_forked) && nd == 2 && dir_overlap_enabled();
  if (dir_ovl) {
    BIST_CUDA(cudaEventRecord(side->fork2, st));  // the folded queries are ready
    BIST_CUDA(cudaStreamWaitEvent(side->s, side->fork2, 0));
  }
  auto stream_of = [&](int u) { return dir_ovl && u == 1 ? side->s : st; };
  auto join_dirs = [&]() {
    if (dir_ovl) { cudaEventRecord(side->join2, side->s); cudaStreamWaitEvent(st, side->join2, 0); }
  };
  for (int u = 0; u < nd && !fused1; ++u) {
    const bool t2s = kinds[u] == 0;
    cudaStream_t su = stream_of(u);
    float* sc1 = dir_ovl && u == 1 ? W.sc1b : W.sc1;
    const int G = (int)(t2s ? S : T), nk = (int)(t2s ? T : S);
    bool t2s_regions = false;
    {
      // Folded stage-1 scores, video rows in place, keys of a group contiguous in the output:
      //   t2s: one product per (b,s): [Lq*h, d] x [T rows at stride S*d]^T -> sc[b,s,(l,hh),t]
      //   s2t: one product per b:     [Lq*h, d] x [T*S rows]^T            -> sc[b,(l,hh),(t,s)]
      GemmArgs g;
      g.A = act(W.qf12, (long long)u * Qd * h); g.W = vft; g.C = sc1;
      g.M = (int)(Lq * h); g.K = di;
      g.a = rows_dense(d);
      g.a_dtype = adt; g.w_dtype = dtype; g.tag = "scores1_gemm";
      bool done = false;
      if (t2s) {
        if (dtype == BIST_BF16) {
          // tensor-core path: one product per b with region-major columns n = s*T + t (whole regions per N tile,
          // the folded queries of b are fetched once per tile instead of once per region) -> sc[b,(l,hh),s,t]
          GemmArgs r = g;
          r.N = (int)TS; r.w = rows_dense(d); r.c = rows_dense(TS); r.wr_T = (int)T; r.wr_S = (int)S;
          r.Z = (int)B; r.zin = 1;
          r.az = {Lq * h * d, 0}; r.wz = {TS * d, 0}; r.cz = {Lq * h * TS, 0};
          { const int rc = gemm_tc(r, su, &done); if (rc != 0) { join_dirs(); return rc; } }
          t2s_regions = done;
        }
        g.N = (int)T; g.w = rows_dense(S * d); g.c = rows_dense(T);
        g.Z = (int)(B * S); g.zin = (int)S;
        g.az = {Lq * h * d, 0}; g.wz = {TS * d, d}; g.cz = {S * Lq * h * T, Lq * h * T};
      } else {
        g.N = (int)TS; g.w = rows_dense(d); g.c = rows_dense(TS);
        g.Z = (int)B; g.zin = 1;
        g.az = {Lq * h * d, 0}; g.wz = {TS * d, 0}; g.cz = {Lq * h * TS, 0};
      }
      if (!done) { const int rc = mm(g, W.qf12, (long long)nd * Qd * h, su, 1); if (rc != 0) { join_dirs(); return rc; } }
    }
    SoftmaxPvArgs a;
    a.S = sc1;
    if (t2s) {  // group (b,s): keys are its T clips (encoder.py:117-121: temporal mask on the keys)
      if (t2s_regions) { a.sg = rows_2level((int)S, Lq * h * TS, T); a.s_l = h * TS; a.s_h = TS; a.s_k = 1; }
      else { a.sg = rows_dense(Lq * h * T); a.s_l = h * T; a.s_h = T; a.s_k = 1; }
      a.mask = temporal_mask; a.mask_gdiv = (int)S; a.m_b = T; a.m_l = 0;
    } else {    // group (b,t): keys are its S regions, no mask (encoder.py:148)
      a.sg = rows_2level((int)T, Lq * h * TS, S); a.s_l = h * TS; a.s_h = TS; a.s_k = 1;
    }
    a.V = (const char*)W.xv + (size_t)u * d * e;
    a.vg = t2s ? rows_2level((int)S, TS * xvld, xvld) : rows_dense(S * xvld);
    a.v = t2s ? rows_dense(S * xvld) : rows_dense(xvld);
    a.ctx = W.ctx1[u];  // c1 [B, Lq, G, d]: the G summaries of one (b,l) are contiguous for stage 2
    a.og = rows_2level(G, Lq * G * d, d); a.o_l = (long long)G * d;
    a.G = (int)(B * G); a.Lq = (int)Lq; a.h = hi; a.dk = dk; a.nk = nk;
    a.v_dtype = dtype; a.ctx_dtype = dtype;
    if (xv_forked && !xv_joined && su == st) {  // (the side stream is ordered behind the value GEMM by itself)
      BIST_CUDA(cudaStreamWaitEvent(st, side->join, 0));
      xv_joined = true;
    }
    { const int rc = softmax_pv(a, su); if (rc != 0) { join_dirs(); return rc; } }
  }
  if (xv_forked && !xv_joined) BIST_CUDA(cudaStreamWaitEvent(st, side->join, 0));

  // ---- stage 2 (encoder.py:125-132 / :152-161) with W_o1 folded through it ------------------------------------------
  for (int u = 0; u < nd; ++u) {
    const bool t2s = kinds[u] == 0;
    const int G = (int)(t2s ? S : T);
    const uint8_t* m2 = t2s ? nullptr : temporal_mask;  // :130 no mask / :157-161 temporal mask
    if (s2mma) {
      bool taken = false;
      const int rc = stage2_mma(act(W.qf12, (nd + u) * Qd * h), W.ctx1[u], m2, act(W.u2, u * Qd * h), (int)B, G, (int)Lq, hi, di,
                                stream_of(u), &taken);
      if (rc != 0 || !taken) join_dirs();
      BIST_TRY(rc);
      BIST_CHECK_ARG(taken, "st_layer: stage2_mma declined a shape it reported as supported");
    } else {
      const float* qf = dtype == BIST_BF16 ? W.qf2f + u * Qd * h : (const float*)W.qf12 + (nd + u) * Qd * h;
      const int rc = stage2_attn(qf, W.ctx1[u], m2, act(W.u2, u * Qd * h), (int)B, G, (int)Lq, hi, di, dtype, adt, stream_of(u));
      if (rc != 0) join_dirs();
      BIST_TRY(rc);
    }
  }
  join_dirs();
  {
    GemmArgs g;  // ctx2[q, hh*dk+j] = u2[q,hh,:] . (W_v2 W_o1)[hh*dk+j,:] + (x W_v2^T + b_v2 + W_v2 b_o1)[q, hh*dk+j]
    g.A = W.u2; g.W = Mw(L.wv2f); g.C = W.cbuf; g.R = act(W.q12x, 2 * nd * Qd);
    g.M = Q; g.N = dk; g.K = di;
    g.a = rows_dense(h * d); g.w = rows_dense(d); g.c = rows_dense(d); g.r = rows_dense(d);
    g.Z = nd * hi; g.zin = hi;
    g.az = {Qd * h, d}; g.wz = {ms, (long long)dk * d}; g.cz = {Qd, dk}; g.rz = {Qd, dk};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.r_dtype = adt; g.tag = "vfold2_gemm";
    BIST_TRY(mm(g, W.u2, (long long)nd * Qd * h, st));
  }
  BIST_TRY(lin(W.cbuf, Qd, L.wo2, L.bo2, W.ts, Qd, BIST_F32, di, di, 0, W.x, Qd, fusedq ? W.nbuf : nullptr));

  // ---- feed-forward (encoder.py:134-135 / :165-166) -----------------------------------------------
  if (fusedq) {
    BIST_TRY(qside_ffn1(W.nbuf, lnA[3], lnB[3], wf1, bf1, W.ffh, nd, B, Lq, d, dff, h, eps, st));
  } else {
    BIST_TRY(layer_norm(W.ts, fs + L.ln_a + 3 * d, fs + L.ln_b + 3 * d, W.nbuf, (int64_t)nd * Q, d, eps, BIST_F32, adt, st, Q, fsd));
    BIST_TRY(lin(W.nbuf, Qd, L.w1, L.b1, W.ffh, (long long)Q * dff, adt, (int)dff, di, 1, nullptr, 0));
  }
  BIST_TRY(lin(W.ffh, (long long)Q * dff, L.w2, L.b2, x_out[0], out_dir, BIST_F32, di, (int)dff, 0, W.ts, Qd));
  return 0;
}
