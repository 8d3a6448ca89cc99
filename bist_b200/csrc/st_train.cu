// Training path of VidEncoderLayer4 (encoder.py:172-201, enc_st_combine == 'none'): a forward that keeps what the
// backward needs on a caller-owned tape, and the backward itself.
//
// Formulation.  Same query-side de-duplication and key-projection folds as the inference forward (st_layer.cu,
// DESIGN.md §3), except that the stage-1 output projection W_o1 is applied explicitly (out1 = x + c1 W_o1^T + b_o1 is
// materialised), because the reference's sublayer dropout sits between W_o1 and stage 2 (modules.py:44).  The backward
// is the chain rule through exactly those operations:
//   stage 2 : s2 = qf2 . out1, u2 = sum_g p2 out1, c2 = u2_h W_v2,h^T + b_v2        (qf2 = q2_h W_k2,h / sqrt(dk))
//   stage 1 : s1 = qf1 . vft,  c1 = sum_k p1 XV,   XV = vft W_v1^T + b_v1           (qf1 = q1_h W_k1,h / sqrt(dk))
// Key biases get a zero gradient (their score term is constant over the keys and cancels in the softmax).
// Gradients are accumulated (+=) into one fp32 block per layer laid out in the reference's parameter order
// (bist_st_layer_grad_elems / include/bist_b200.h); d(vft) is accumulated into a caller-owned fp32 [B,T,S,d] tensor.
#include <math.h>

#include "common.cuh"

namespace bist {
namespace {

struct TrainLayout {
  // matrices (compute dtype), element offsets relative to the direction base
  size_t wqkv0, wo0, wq12, wkt12, wv2, wo1, wo2, w1, w2;             // forward operands  [out, in]
  size_t wqkv0T, wo0T, wq12T, wks12, wv2T, wo1T, wo2T, w1T, w2T;     // backward operands [in, out]
  size_t dir_stride, wv1, wv1T, mat_elems;                           // wv1 / wv1T: [nd][d,d] after the directions
  // fp32 vectors, float offsets relative to the direction base
  size_t bqkv0, bo0, bq12, bv2, bo1, bo2, b1, b2, ln_a, ln_b, fdir_stride, bv1, f_elems;
  size_t f_byte_off, total_bytes;
};

TrainLayout train_layout(int64_t d, int64_t dff, int n_dirs, int dtype) {
  TrainLayout L;
  size_t o = 0;
  auto m = [&](size_t n) { size_t r = o; o += n; return r; };
  const size_t dd = (size_t)d * d, df = (size_t)d * dff;
  L.wqkv0 = m(3 * dd); L.wo0 = m(dd); L.wq12 = m(2 * dd); L.wkt12 = m(2 * dd); L.wv2 = m(dd); L.wo1 = m(dd); L.wo2 = m(dd);
  L.w1 = m(df); L.w2 = m(df);
  L.wqkv0T = m(3 * dd); L.wo0T = m(dd); L.wq12T = m(2 * dd); L.wks12 = m(2 * dd); L.wv2T = m(dd); L.wo1T = m(dd); L.wo2T = m(dd);
  L.w1T = m(df); L.w2T = m(df);
  L.dir_stride = o;
  L.wv1 = L.dir_stride * n_dirs;
  L.wv1T = L.wv1 + (size_t)n_dirs * dd;
  L.mat_elems = L.wv1T + (size_t)n_dirs * dd;
  o = 0;
  L.bqkv0 = m(3 * d); L.bo0 = m(d); L.bq12 = m(2 * d); L.bv2 = m(d); L.bo1 = m(d); L.bo2 = m(d); L.b1 = m(dff); L.b2 = m(d);
  L.ln_a = m(4 * d); L.ln_b = m(4 * d);
  L.fdir_stride = o;
  L.bv1 = L.fdir_stride * n_dirs;
  L.f_elems = L.bv1 + (size_t)n_dirs * d;
  L.f_byte_off = align_up(L.mat_elems * dtype_size(dtype), 256);
  L.total_bytes = L.f_byte_off + L.f_elems * 4 + 256;
  return L;
}

// Gradient block (fp32), per direction: attn{0,1,2}: W[4][d,d] b[4][d] ; ff: w1 b1 w2 b2 ; norms{0..3}: a b
struct GradLayout {
  size_t attn_stride, ff, ln, dir_stride, total;
  size_t W(int ai, int j, size_t dd) const { return ai * attn_stride + j * dd; }
  size_t b(int ai, int j, size_t dd, size_t d) const { return ai * attn_stride + 4 * dd + j * d; }
};
GradLayout grad_layout(int64_t d, int64_t dff, int n_dirs) {
  GradLayout G;
  const size_t dd = (size_t)d * d;
  G.attn_stride = 4 * dd + 4 * d;
  G.ff = 3 * G.attn_stride;
  G.ln = G.ff + 2 * (size_t)d * dff + dff + d;
  G.dir_stride = G.ln + 8 * d;
  G.total = G.dir_stride * n_dirs;
  return G;
}

bool stage2_on_mma(int dtype, int64_t T, int64_t S, int64_t h, int64_t d, int nd, const int* kinds) {
  if (dtype != BIST_BF16) return false;
  for (int u = 0; u < nd; ++u)
    if (!stage2_mma_supported((int)(kinds[u] == 0 ? S : T), (int)h, (int)d)) return false;
  return true;
}

struct Tape {
  void *n0, *qkv0, *c0, *n12, *q12, *qf12, *xv, *u2, *c2, *n3, *ffh;  // compute dtype
  float *x, *ts, *qf2f, *sc0, *rsum;
  float* sc1[2];
  void* ctx1[2];
  void* out1[2];
  size_t bytes;
  bool ok;
};

Tape carve_tape(void* base, size_t cap, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff, int64_t h, int nd,
                const int* kinds, int dtype) {
  Arena A(base, cap);
  Tape W;
  const size_t Q = (size_t)B * Lq, e = dtype_size(dtype);
  auto f = [&](size_t n) { return (float*)A.take(n * 4); };
  auto a = [&](size_t n) { return A.take(n * e); };
  W.n0 = a(nd * Q * d); W.qkv0 = a(nd * Q * 3 * d); W.c0 = a(nd * Q * d);
  W.n12 = a(2 * nd * Q * d); W.q12 = a(2 * nd * Q * d); W.qf12 = a(2 * nd * Q * h * d);
  W.xv = a((size_t)B * T * S * nd * d);
  W.u2 = a(nd * Q * h * d); W.c2 = a(nd * Q * d); W.n3 = a(nd * Q * d); W.ffh = a(nd * Q * dff);
  W.x = f(nd * Q * d); W.ts = f(nd * Q * d);
  W.qf2f = (dtype == BIST_BF16 && !stage2_on_mma(dtype, T, S, h, d, nd, kinds)) ? f(nd * Q * h * d) : nullptr;
  W.sc0 = f(nd * Q * h * Lq);
  W.rsum = f(nd * Q * h);
  for (int u = 0; u < 2; ++u) { W.sc1[u] = nullptr; W.ctx1[u] = nullptr; W.out1[u] = nullptr; }
  for (int u = 0; u < nd; ++u) {
    const size_t G = kinds[u] == 0 ? S : T;
    W.sc1[u] = f(Q * h * T * S);
    W.ctx1[u] = a(Q * G * d);
    W.out1[u] = a(Q * G * d);
  }
  W.bytes = A.used + 256;
  W.ok = A.ok;
  return W;
}

struct BwdWs {
  void *a_dout, *a_dts, *a_dx, *dffh, *dn, *dc, *du2, *dqf12, *dq12, *dxv, *qf1T, *dsT, *dqkv0;  // compute dtype
  void* dout1[2];
  void* dc1[2];
  float *dts, *dqf1, *drsum;
  size_t bytes;
  bool ok;
};

BwdWs carve_bwd(void* base, size_t cap, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff, int64_t h, int nd,
                const int* kinds, int dtype) {
  Arena A(base, cap);
  BwdWs W;
  const size_t Q = (size_t)B * Lq, e = dtype_size(dtype);
  auto f = [&](size_t n) { return (float*)A.take(n * 4); };
  auto a = [&](size_t n) { return A.take(n * e); };
  W.a_dout = a(nd * Q * d); W.a_dts = a(nd * Q * d); W.a_dx = a(nd * Q * d);
  W.dffh = a(nd * Q * dff); W.dn = a(2 * nd * Q * d); W.dc = a(nd * Q * d); W.du2 = a(nd * Q * h * d);
  W.dqf12 = a(2 * nd * Q * h * d); W.dq12 = a(2 * nd * Q * d);
  W.dxv = a((size_t)B * T * S * nd * d);
  W.qf1T = a(nd * Q * h * d); W.dsT = a(nd * Q * h * T * S);   // both directions side by side (one product into d_vft)
  W.dqkv0 = a(nd * Q * 3 * d);
  for (int u = 0; u < 2; ++u) { W.dout1[u] = nullptr; W.dc1[u] = nullptr; }
  for (int u = 0; u < nd; ++u) {
    const size_t G = kinds[u] == 0 ? S : T;
    W.dout1[u] = a(Q * G * d);
    W.dc1[u] = a(Q * G * d);
  }
  W.dts = f(nd * Q * d); W.dqf1 = f(nd * Q * h * d); W.drsum = f(nd * Q * h);
  W.bytes = A.used + 256;
  W.ok = A.ok;
  return W;
}

// Dropout sites of one layer call (DropCfg::site); u = slot of the active direction.
enum : uint32_t { S_SELF_P = 0, S_SELF_OUT = 1, S_ST1_P = 2 /* +u */, S_ST1_OUT = 4 /* +u */, S_ST2_P = 6 /* +u */, S_ST2_OUT = 8,
                  S_FF_HID = 9, S_FF_OUT = 10 };

int active(int n_dirs, int dir_mask, int* kinds) {
  int nd = 0;
  if (dir_mask & 1) kinds[nd++] = 0;
  if (dir_mask & 2) kinds[nd++] = 1;
  if (n_dirs == 1 && nd != 1) return -1;
  return nd;
}

// Stage-1 argument block shared by the forward (softmax_pv) and the backward (stage1_bwd).
SoftmaxPvArgs stage1_args(bool t2s, bool regions, const float* sc1, const uint8_t* temporal_mask, const void* xv_u, void* ctx,
                          int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h, int nd, int dtype, const DropCfg& drop) {
  const long long TS = T * S, xvld = (long long)nd * d;
  const int G = (int)(t2s ? S : T);
  SoftmaxPvArgs a;
  a.S = sc1;
  if (t2s) {  // group (b,s): keys are its T clips, temporal mask on the keys (encoder.py:117-121)
    if (regions) { a.sg = rows_2level((int)S, Lq * h * TS, T); a.s_l = h * TS; a.s_h = TS; a.s_k = 1; }
    else { a.sg = rows_dense(Lq * h * T); a.s_l = h * T; a.s_h = T; a.s_k = 1; }
    a.mask = temporal_mask; a.mask_gdiv = (int)S; a.m_b = T; a.m_l = 0;
  } else {    // group (b,t): keys are its S regions, no mask (encoder.py:148)
    a.sg = rows_2level((int)T, Lq * h * TS, S); a.s_l = h * TS; a.s_h = TS; a.s_k = 1;
  }
  a.V = xv_u;
  a.vg = t2s ? rows_2level((int)S, TS * xvld, xvld) : rows_dense(S * xvld);
  a.v = t2s ? rows_dense(S * xvld) : rows_dense(xvld);
  a.ctx = ctx;  // c1 [B, Lq, G, d]
  a.og = rows_2level(G, Lq * G * d, d); a.o_l = (long long)G * d;
  a.G = (int)(B * G); a.Lq = (int)Lq; a.h = (int)h; a.dk = (int)(d / h); a.nk = (int)(t2s ? T : S);
  a.v_dtype = dtype; a.ctx_dtype = dtype;
  a.drop = drop;
  return a;
}

// t2s scores on the tensor-core GEMM use the region-major layout; decide once so forward and backward agree.
bool t2s_region_scores(int dtype, int64_t T, int64_t S, int64_t d) {
  if (dtype != BIST_BF16 || getenv("BIST_DISABLE_TC") != nullptr) return false;
  return d % 64 == 0 && T <= 256 && S >= 1 && (T * S) >= 8;
}
bool region_scores_ok(bool base, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t h) {
  return base && B * Lq * h * T * S >= 64 * 64;  // gemm_tc leaves tiny products to the CUDA cores
}

}  // namespace
}  // namespace bist

using namespace bist;

extern "C" {

size_t bist_st_layer_train_packed_bytes(int64_t d, int64_t d_ff, int n_dirs, int dtype) {
  if (d <= 0 || d_ff <= 0 || (n_dirs != 1 && n_dirs != 2) || !dtype_ok(dtype)) return 0;
  return train_layout(d, d_ff, n_dirs, dtype).total_bytes;
}

size_t bist_st_layer_grad_elems(int64_t d, int64_t d_ff, int n_dirs) {
  if (d <= 0 || d_ff <= 0 || (n_dirs != 1 && n_dirs != 2)) return 0;
  return grad_layout(d, d_ff, n_dirs).total;
}

int bist_st_layer_train_pack(const bist_st_layer_params_t* P, void* packed, size_t packed_bytes, int64_t d, int64_t dff,
                             int64_t h, int n_dirs, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(P && packed, "train_pack: null pointer");
  BIST_CHECK_ARG(d > 0 && dff > 0 && h > 0 && d % h == 0 && (n_dirs == 1 || n_dirs == 2) && dtype_ok(dtype), "train_pack: bad argument");
  BIST_CHECK_ARG(d % 4 == 0 && dff % 4 == 0, "train_pack: d and d_ff must be multiples of 4");
  const TrainLayout L = train_layout(d, dff, n_dirs, dtype);
  if (packed_bytes < L.total_bytes) { set_error("train_pack: buffer too small"); return BIST_ERR_WORKSPACE; }
  const size_t dd = (size_t)d * d;
  float* fs = (float*)((char*)packed + L.f_byte_off);
  const float rs = 1.f / sqrtf((float)(d / h));
  const int di = (int)d, fi = (int)dff;
  for (int u = 0; u < n_dirs; ++u) {   // one launch per direction (the item table travels as a kernel argument)
    const int ab = 3 * u;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 4; ++j) BIST_CHECK_ARG(P->attn_w[ab + i][j] && P->attn_b[ab + i][j], "train_pack: null attn parameter");
    BIST_CHECK_ARG(P->ff_w1[u] && P->ff_b1[u] && P->ff_w2[u] && P->ff_b2[u], "train_pack: null ff parameter");
    for (int i = 0; i < 4; ++i) BIST_CHECK_ARG(P->ln_a[4 * u + i] && P->ln_b[4 * u + i], "train_pack: null norm parameter");
    const size_t mb = u * L.dir_stride, fb = u * L.fdir_stride;
    PackItem items[kMaxPackItems];
    int n = 0;
    // dst [rows, cols] at matrix offset `off`: W (transpose = 0, src [rows, cols]) or W^T (transpose = 1, src [cols, rows])
    auto Mx = [&](const float* src, size_t off, int rows, int cols, int transpose, float scale) {
      items[n++] = PackItem{src, (long long)off, rows, cols, transpose, 0, scale, 0};
    };
    auto V = [&](const float* src, size_t off, int len) { items[n++] = PackItem{src, (long long)off, 1, len, 0, 1, 1.f, 0}; };
    auto both = [&](const float* src, size_t off, size_t offT, int rows, int cols) {
      Mx(src, mb + off, rows, cols, 0, 1.f);
      Mx(src, mb + offT, cols, rows, 1, 1.f);
    };
    for (int j = 0; j < 3; ++j) {
      Mx(P->attn_w[ab][j], mb + L.wqkv0 + j * dd, di, di, 0, 1.f);
      // W_j^T of the self-attention q/k/v projections: dn0 = sum_j dqkv0[:, j-th block] . W_j runs as three accumulating products
      Mx(P->attn_w[ab][j], mb + L.wqkv0T + j * dd, di, di, 1, 1.f);
      V(P->attn_b[ab][j], fb + L.bqkv0 + (size_t)j * d, di);
    }
    both(P->attn_w[ab][3], L.wo0, L.wo0T, di, di);
    V(P->attn_b[ab][3], fb + L.bo0, di);
    both(P->attn_w[ab + 1][0], L.wq12, L.wq12T, di, di);
    both(P->attn_w[ab + 2][0], L.wq12 + dd, L.wq12T + dd, di, di);
    V(P->attn_b[ab + 1][0], fb + L.bq12, di);
    V(P->attn_b[ab + 2][0], fb + L.bq12 + d, di);
    // key folds: forward operand (W_k)^T / sqrt(dk), backward operand W_k / sqrt(dk)
    Mx(P->attn_w[ab + 1][1], mb + L.wkt12, di, di, 1, rs);
    Mx(P->attn_w[ab + 2][1], mb + L.wkt12 + dd, di, di, 1, rs);
    Mx(P->attn_w[ab + 1][1], mb + L.wks12, di, di, 0, rs);
    Mx(P->attn_w[ab + 2][1], mb + L.wks12 + dd, di, di, 0, rs);
    both(P->attn_w[ab + 2][2], L.wv2, L.wv2T, di, di);
    V(P->attn_b[ab + 2][2], fb + L.bv2, di);
    both(P->attn_w[ab + 1][3], L.wo1, L.wo1T, di, di);
    V(P->attn_b[ab + 1][3], fb + L.bo1, di);
    both(P->attn_w[ab + 2][3], L.wo2, L.wo2T, di, di);
    V(P->attn_b[ab + 2][3], fb + L.bo2, di);
    both(P->ff_w1[u], L.w1, L.w1T, fi, di);
    both(P->ff_w2[u], L.w2, L.w2T, di, fi);
    V(P->ff_b1[u], fb + L.b1, fi);
    V(P->ff_b2[u], fb + L.b2, di);
    Mx(P->attn_w[ab + 1][2], L.wv1 + (size_t)u * dd, di, di, 0, 1.f);
    Mx(P->attn_w[ab + 1][2], L.wv1T + (size_t)u * dd, di, di, 1, 1.f);
    V(P->attn_b[ab + 1][2], L.bv1 + (size_t)u * d, di);
    for (int i = 0; i < 4; ++i) {
      V(P->ln_a[4 * u + i], fb + L.ln_a + (size_t)i * d, di);
      V(P->ln_b[4 * u + i], fb + L.ln_b + (size_t)i * d, di);
    }
    BIST_TRY(pack_many(items, n, packed, fs, dtype, st));
  }
  return 0;
}

size_t bist_st_layer_tape_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t d_ff, int64_t h, int n_dirs,
                                int dtype) {
  if (B <= 0 || T <= 0 || S <= 0 || Lq <= 0 || d <= 0 || d_ff <= 0 || h <= 0 || !dtype_ok(dtype)) return 0;
  int kinds[2] = {0, 1};
  const int nd = n_dirs == 2 ? 2 : 1;
  size_t best = 0;
  for (int k = 0; k < (nd == 1 ? 2 : 1); ++k) {
    if (nd == 1) kinds[0] = k;
    Tape W = carve_tape(nullptr, (size_t)-1, B, T, S, Lq, d, d_ff, h, nd, kinds, dtype);
    best = W.bytes > best ? W.bytes : best;
  }
  return best;
}

size_t bist_st_layer_bwd_workspace_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t d_ff, int64_t h,
                                         int n_dirs, int dtype) {
  if (B <= 0 || T <= 0 || S <= 0 || Lq <= 0 || d <= 0 || d_ff <= 0 || h <= 0 || !dtype_ok(dtype)) return 0;
  int kinds[2] = {0, 1};
  const int nd = n_dirs == 2 ? 2 : 1;
  size_t best = 0;
  for (int k = 0; k < (nd == 1 ? 2 : 1); ++k) {
    if (nd == 1) kinds[0] = k;
    BwdWs W = carve_bwd(nullptr, (size_t)-1, B, T, S, Lq, d, d_ff, h, nd, kinds, dtype);
    best = W.bytes > best ? W.bytes : best;
  }
  return best;
}

}  // extern "C"

namespace {

struct Ctx {
  int64_t B, T, S, Lq, d, dff, h;
  int n_dirs, nd, kinds[2], pu0, dtype;
  float eps;
  TrainLayout L;
  const char* mat;
  const float* fs;
  cudaStream_t st;
  size_t e;
  int Q, dk, di, hi;
  long long Qd, dd, ms, fsd, TS;
  bool s2mma, regions;
  float p_attn = 0.f, p_sub = 0.f, p_ffn = 0.f;
  unsigned long long seed = 0;
  DropCfg drop(uint32_t site, float p) const { return make_drop(p, seed, site); }
  const void* Mw(size_t off) const { return mat + off * e; }
  const void* Mshared(size_t off) const { return mat - (size_t)pu0 * L.dir_stride * e + off * e; }
  const float* Fshared(size_t off) const { return fs - (size_t)pu0 * L.fdir_stride + off; }
  void* act(void* p, long long elems) const { return (char*)p + elems * (long long)e; }
  const void* act(const void* p, long long elems) const { return (const char*)p + elems * (long long)e; }
};

int make_ctx(Ctx& c, const void* packed, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff, int64_t h, int n_dirs,
             int dir_mask, float eps, int dtype, cudaStream_t st, const bist_dropout_t* dropout) {
  if (dropout) {
    BIST_CHECK_ARG(dropout->p_attn >= 0.f && dropout->p_attn < 1.f && dropout->p_sublayer >= 0.f && dropout->p_sublayer < 1.f &&
                       dropout->p_ffn >= 0.f && dropout->p_ffn < 1.f, "st_layer(train): dropout probabilities must be in [0, 1)");
    c.p_attn = dropout->p_attn; c.p_sub = dropout->p_sublayer; c.p_ffn = dropout->p_ffn; c.seed = dropout->seed;
  }
  BIST_CHECK_ARG(B > 0 && T > 0 && S > 0 && Lq > 0 && d > 0 && dff > 0 && h > 0 && d % h == 0, "st_layer(train): bad shape");
  BIST_CHECK_ARG(d % 4 == 0 && dff % 4 == 0, "st_layer(train): d and d_ff must be multiples of 4");
  BIST_CHECK_ARG((n_dirs == 1 || n_dirs == 2) && dtype_ok(dtype) && (dir_mask & ~3) == 0 && dir_mask != 0, "st_layer(train): bad n_dirs/dir_mask/dtype");
  BIST_CHECK_ARG(B * T * S < (1LL << 31) / 2 && B * Lq * h < (1LL << 31), "st_layer(train): problem too large");
  BIST_CHECK_ARG(B * S <= 65535 && 2 * B * h <= 65535, "st_layer(train): batch too large for one call (split B)");
  c.B = B; c.T = T; c.S = S; c.Lq = Lq; c.d = d; c.dff = dff; c.h = h; c.n_dirs = n_dirs; c.dtype = dtype; c.eps = eps; c.st = st;
  c.nd = active(n_dirs, dir_mask, c.kinds);
  BIST_CHECK_ARG(c.nd > 0, "st_layer(train): a single-direction layer needs exactly one bit in dir_mask");
  c.L = train_layout(d, dff, n_dirs, dtype);
  c.pu0 = n_dirs == 2 ? c.kinds[0] : 0;
  c.e = dtype_size(dtype);
  c.mat = (const char*)packed + (size_t)c.pu0 * c.L.dir_stride * c.e;
  c.fs = (const float*)((const char*)packed + c.L.f_byte_off) + (size_t)c.pu0 * c.L.fdir_stride;
  c.Q = (int)(B * Lq); c.dk = (int)(d / h); c.di = (int)d; c.hi = (int)h;
  c.Qd = (long long)c.Q * d; c.dd = (long long)d * d; c.ms = (long long)c.L.dir_stride; c.fsd = (long long)c.L.fdir_stride;
  c.TS = T * S;
  c.s2mma = stage2_on_mma(dtype, T, S, h, d, c.nd, c.kinds);
  c.regions = region_scores_ok(t2s_region_scores(dtype, T, S, d), B, T, S, Lq, h);
  return 0;
}

// Batched (over the active directions) product  C[u] = act(A[u] . W[u]^T + b[u]) (+ R[u]),  A [Q, K] dense.
int lin(const Ctx& c, const void* A, long long a_dir, int a_dtype, size_t w_off, const float* bias, void* C, long long c_dir,
        int c_dtype, int N, int K, int relu, const void* R, long long r_dir, int r_dtype, const char* tag) {
  GemmArgs g;
  g.A = A; g.W = c.Mw(w_off); g.C = C; g.bias = bias; g.R = R;
  g.M = c.Q; g.N = N; g.K = K;
  g.a = rows_dense(K); g.w = rows_dense(K); g.c = rows_dense(N); g.r = rows_dense(N);
  g.Z = c.nd; g.zin = 1;
  g.az = {a_dir, 0}; g.wz = {c.ms, 0}; g.cz = {c_dir, 0}; g.rz = {r_dir, 0}; g.bz = {c.fsd, 0};
  g.relu = relu; g.a_dtype = a_dtype; g.w_dtype = c.dtype; g.c_dtype = c_dtype; g.r_dtype = r_dtype; g.tag = tag;
  return gemm(g, c.st);
}

}  // namespace

// ------------------------------------------------------------------------------------------- training forward
extern "C" int bist_st_layer_train_fwd(const float* in_t2s, const float* in_s2t, const void* vft, const uint8_t* query_mask,
                                       const uint8_t* temporal_mask, const void* packed, float* out_t2s, float* out_s2t,
                                       int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff, int64_t h,
                                       int n_dirs, int dir_mask, float eps, void* tape, size_t tape_bytes, int dtype,
                                       const bist_dropout_t* dropout, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(vft && query_mask && temporal_mask && packed && tape, "st_layer_train_fwd: null pointer");
  Ctx c;
  BIST_TRY(make_ctx(c, packed, B, T, S, Lq, d, dff, h, n_dirs, dir_mask, eps, dtype, st, dropout));
  const int nd = c.nd;
  const float* x_in[2];
  float* x_out[2];
  for (int u = 0; u < nd; ++u) {
    x_in[u] = c.kinds[u] == 0 ? in_t2s : in_s2t;
    x_out[u] = c.kinds[u] == 0 ? out_t2s : out_s2t;
    BIST_CHECK_ARG(x_in[u] && x_out[u], "st_layer_train_fwd: null stream pointer for an active direction");
  }
  const Tape W = carve_tape(tape, tape_bytes, B, T, S, Lq, d, dff, h, nd, c.kinds, dtype);
  if (!W.ok) { set_error("st_layer_train_fwd: tape too small (%zu bytes given)", tape_bytes); return BIST_ERR_WORKSPACE; }
  const TrainLayout& L = c.L;
  const int adt = dtype, Q = c.Q, di = c.di, hi = c.hi, dk = c.dk;
  const long long Qd = c.Qd, dd = c.dd, ms = c.ms, fsd = c.fsd, TS = c.TS;
  const float* fs = c.fs;
  const long long in_dir = nd == 2 ? (long long)(x_in[1] - x_in[0]) : 0;
  const long long out_dir = nd == 2 ? (long long)(x_out[1] - x_out[0]) : 0;

  // ---- query self-attention (encoder.py:176 / :184)
  BIST_TRY(layer_norm(x_in[0], fs + L.ln_a, fs + L.ln_b, W.n0, (int64_t)nd * Q, d, eps, BIST_F32, adt, st, Q, fsd, in_dir, true));
  BIST_TRY(lin(c, W.n0, Qd, adt, L.wqkv0, fs + L.bqkv0, W.qkv0, 3 * Qd, adt, 3 * di, di, 0, nullptr, 0, adt, "query_side_gemm"));
  {
    bool fused = false;
    if (dtype == BIST_BF16 && c.p_attn == 0.f)
      BIST_TRY(self_attn_mma(W.qkv0, query_mask, (int)B, W.c0, (int)(nd * B), (int)Lq, hi, dk, st, &fused));
    if (!fused) {
      GemmArgs g;
      g.A = W.qkv0; g.W = c.act(W.qkv0, d); g.C = W.sc0;
      g.M = (int)Lq; g.N = (int)Lq; g.K = dk;
      g.a = rows_dense(3 * d); g.w = rows_dense(3 * d); g.c = rows_dense(h * Lq);
      g.Z = (int)(nd * B * h); g.zin = hi;
      g.az = {Lq * 3 * d, dk}; g.wz = {Lq * 3 * d, dk}; g.cz = {Lq * h * Lq, Lq};
      g.alpha = 1.f / sqrtf((float)dk);
      g.a_dtype = adt; g.w_dtype = adt; g.tag = "scores0_gemm";
      BIST_TRY(gemm_simt(g, st));
      SoftmaxPvArgs a;
      a.S = W.sc0; a.sg = rows_dense(Lq * h * Lq); a.s_l = h * Lq; a.s_h = Lq;
      a.mask = query_mask; a.mask_gmod = (int)B; a.m_b = Lq; a.m_l = 0;
      a.V = c.act(W.qkv0, 2 * d); a.v = rows_dense(3 * d); a.vg = rows_dense(Lq * 3 * d);
      a.ctx = W.c0; a.og = rows_dense(Lq * d); a.o_l = d;
      a.G = (int)(nd * B); a.Lq = (int)Lq; a.h = hi; a.dk = dk; a.nk = (int)Lq;
      a.v_dtype = adt; a.ctx_dtype = adt;
      a.drop = c.drop(S_SELF_P, c.p_attn);
      BIST_TRY(softmax_pv(a, st));
    }
  }
  // x = x_in + dropout(c0 W_o0^T + b_o0)   (modules.py:44)
  if (c.p_sub > 0.f) {
    BIST_TRY(lin(c, W.c0, Qd, adt, L.wo0, fs + L.bo0, W.x, Qd, BIST_F32, di, di, 0, nullptr, 0, BIST_F32, "query_side_gemm"));
    for (int u = 0; u < nd; ++u)
      BIST_TRY(dropout_add(W.x + u * Qd, x_in[u], Q, d, 1, c.drop(S_SELF_OUT, c.p_sub), 0, 0, BIST_F32, st, (int64_t)u * Q));
  } else {
    BIST_TRY(lin(c, W.c0, Qd, adt, L.wo0, fs + L.bo0, W.x, Qd, BIST_F32, di, di, 0, x_in[0], in_dir, BIST_F32, "query_side_gemm"));
  }

  // ---- query side of stage 1 / stage 2: n12 = [LN1(x) | LN2(x)], q12 = n12 . {W_q1 | W_q2}^T + b, folds
  BIST_TRY(layer_norm2(W.x, fs + L.ln_a + d, fs + L.ln_b + d, fs + L.ln_a + 2 * d, fs + L.ln_b + 2 * d, W.n12, c.act(W.n12, nd * Qd),
                       nullptr, (int64_t)nd * Q, d, eps, adt, st, Q, fsd));
  {
    GemmArgs g;
    g.A = W.n12; g.W = c.Mw(L.wq12); g.C = W.q12; g.bias = fs + L.bq12;
    g.M = Q; g.N = di; g.K = di;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(d);
    g.Z = 2 * nd; g.zin = nd;  // z = which * nd + u
    g.az = {nd * Qd, Qd}; g.wz = {dd, ms}; g.cz = {nd * Qd, Qd}; g.bz = {d, fsd};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "query_side_gemm";
    BIST_TRY(gemm(g, st));
  }
  {
    GemmArgs g;  // qf[which][u][q,hh,:] = q[which][u][q, hh-slice] . WkT[:, hh-slice]^T
    g.A = W.q12; g.W = c.Mw(L.wkt12); g.C = W.qf12;
    g.M = Q; g.N = di; g.K = dk;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(h * d);
    g.zin = hi; g.zmid = nd; g.Z = 2 * nd * hi;  // z = (which * nd + u) * h + hh
    g.az = {nd * Qd, dk, Qd}; g.wz = {dd, dk, ms}; g.cz = {nd * Qd * h, d, Qd * h};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "fold_gemm";
    BIST_TRY(gemm(g, st));
    if (W.qf2f) {  // the CUDA-core stage-2 kernels want fp32 folded queries
      g.Z = nd * hi;
      g.A = c.act(W.q12, nd * Qd); g.W = c.Mw(L.wkt12 + dd); g.C = W.qf2f; g.c_dtype = BIST_F32;
      BIST_TRY(gemm(g, st));
    }
  }

  // ---- stage 1 (encoder.py:111-121 / :143-148)
  const long long xvld = (long long)nd * d;
  {
    GemmArgs g;  // XV = vft . [W_v1]^T + b_v1 for the active directions
    g.A = vft; g.W = c.Mshared(L.wv1 + (size_t)c.pu0 * dd); g.C = W.xv;
    g.bias = c.Fshared(L.bv1 + (size_t)c.pu0 * d);
    g.M = (int)(B * TS); g.N = nd * di; g.K = di;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(xvld);
    g.a_dtype = dtype; g.w_dtype = dtype; g.c_dtype = dtype; g.tag = "xv_gemm";
    BIST_TRY(gemm(g, st));
  }
  for (int u = 0; u < nd; ++u) {
    const bool t2s = c.kinds[u] == 0;
    const int G = (int)(t2s ? S : T);
    bool regions = false;
    {
      GemmArgs g;
      g.A = c.act(W.qf12, (long long)u * Qd * h); g.W = vft; g.C = W.sc1[u];
      g.M = (int)(Lq * h); g.K = di;
      g.a = rows_dense(d);
      g.a_dtype = adt; g.w_dtype = dtype; g.tag = "scores1_gemm";
      bool done = false;
      if (t2s) {
        if (c.regions) {
          GemmArgs r = g;
          r.N = (int)TS; r.w = rows_dense(d); r.c = rows_dense(TS); r.wr_T = (int)T; r.wr_S = (int)S;
          r.Z = (int)B; r.zin = 1;
          r.az = {Lq * h * d, 0}; r.wz = {TS * d, 0}; r.cz = {Lq * h * TS, 0};
          BIST_TRY(gemm_tc(r, st, &done));
          BIST_CHECK_ARG(done, "st_layer_train_fwd: the region-major score product was declined");
          regions = true;
        }
        g.N = (int)T; g.w = rows_dense(S * d); g.c = rows_dense(T);
        g.Z = (int)(B * S); g.zin = (int)S;
        g.az = {Lq * h * d, 0}; g.wz = {TS * d, d}; g.cz = {S * Lq * h * T, Lq * h * T};
      } else {
        g.N = (int)TS; g.w = rows_dense(d); g.c = rows_dense(TS);
        g.Z = (int)B; g.zin = 1;
        g.az = {Lq * h * d, 0}; g.wz = {TS * d, 0}; g.cz = {Lq * h * TS, 0};
      }
      if (!done) BIST_TRY(gemm(g, st));
    }
    SoftmaxPvArgs a = stage1_args(t2s, regions, W.sc1[u], temporal_mask, c.act(W.xv, (long long)u * d), W.ctx1[u], B, T, S, Lq, d, h, nd, dtype,
                                  c.drop(S_ST1_P + u, c.p_attn));
    BIST_TRY(softmax_pv(a, st));
    {
      GemmArgs g;  // out1[b,l,g,:] = x[b,l,:] + c1[b,l,g,:] W_o1^T + b_o1   (modules.py:44 around encoder.py:121 / :148)
      g.A = W.ctx1[u]; g.W = c.Mw(L.wo1 + (size_t)u * ms); g.C = W.out1[u]; g.bias = fs + L.bo1 + u * fsd;
      g.R = c.p_sub > 0.f ? nullptr : W.x + u * Qd;
      g.M = Q * G; g.N = di; g.K = di;
      g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(d); g.r = rows_2level(G, d, 0);
      g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.r_dtype = BIST_F32; g.tag = "o1_gemm";
      BIST_TRY(gemm(g, st));
      if (c.p_sub > 0.f)  // the reference drops elements of its [B*G, Lq, d] tensor: index through the (b,l,g) -> (b,g,l) row map
        BIST_TRY(dropout_add(W.out1[u], W.x + u * Qd, (int64_t)Q * G, d, G, c.drop(S_ST1_OUT + u, c.p_sub), (int)Lq, G, adt, st));
    }
  }

  // ---- stage 2 (encoder.py:125-132 / :152-161) over out1
  for (int u = 0; u < nd; ++u) {
    const bool t2s = c.kinds[u] == 0;
    const int G = (int)(t2s ? S : T);
    const uint8_t* m2 = t2s ? nullptr : temporal_mask;
    Stage2Extra ex;
    ex.drop = c.drop(S_ST2_P + u, c.p_attn); ex.rsum = W.rsum + (long long)u * Q * h;
    const Stage2Extra* exp = c.p_attn > 0.f ? &ex : nullptr;
    if (c.s2mma) {
      bool taken = false;
      BIST_TRY(stage2_mma(c.act(W.qf12, (nd + u) * Qd * h), W.out1[u], m2, c.act(W.u2, u * Qd * h), (int)B, G, (int)Lq, hi, di, st, &taken, exp));
      BIST_CHECK_ARG(taken, "st_layer_train_fwd: stage2_mma declined a shape it reported as supported");
    } else {
      const float* qf = dtype == BIST_BF16 ? W.qf2f + u * Qd * h : (const float*)W.qf12 + (nd + u) * Qd * h;
      BIST_TRY(stage2_attn(qf, W.out1[u], m2, c.act(W.u2, u * Qd * h), (int)B, G, (int)Lq, hi, di, dtype, adt, st, exp));
    }
  }
  {
    GemmArgs g;  // c2[q, hh-slice] = u2[q,hh,:] . W_v2[hh-slice,:]^T + b_v2[hh-slice]
    g.A = W.u2; g.W = c.Mw(L.wv2); g.C = W.c2; g.bias = fs + L.bv2;
    g.M = Q; g.N = dk; g.K = di;
    g.a = rows_dense(h * d); g.w = rows_dense(d); g.c = rows_dense(d);
    g.Z = nd * hi; g.zin = hi;
    g.az = {Qd * h, d}; g.wz = {ms, (long long)dk * d}; g.cz = {Qd, dk}; g.bz = {fsd, dk};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "vfold2_gemm";
    BIST_TRY(gemm(g, st));
    if (c.p_attn > 0.f)  // dropped probabilities do not sum to one: the value bias enters with weight rsum (modules.py:62-64)
      for (int u = 0; u < nd; ++u)
        BIST_TRY(bias_rsum_fwd(c.act(W.c2, u * Qd), fs + u * fsd + L.bv2, W.rsum + (long long)u * Q * h, Q, hi, dk, adt, st));
  }
  if (c.p_sub > 0.f) {
    BIST_TRY(lin(c, W.c2, Qd, adt, L.wo2, fs + L.bo2, W.ts, Qd, BIST_F32, di, di, 0, nullptr, 0, BIST_F32, "query_side_gemm"));
    BIST_TRY(dropout_add(W.ts, W.x, (int64_t)nd * Q, d, 1, c.drop(S_ST2_OUT, c.p_sub), 0, 0, BIST_F32, st));
  } else {
    BIST_TRY(lin(c, W.c2, Qd, adt, L.wo2, fs + L.bo2, W.ts, Qd, BIST_F32, di, di, 0, W.x, Qd, BIST_F32, "query_side_gemm"));
  }

  // ---- feed-forward (encoder.py:134-135 / :165-166)
  BIST_TRY(layer_norm(W.ts, fs + L.ln_a + 3 * d, fs + L.ln_b + 3 * d, W.n3, (int64_t)nd * Q, d, eps, BIST_F32, adt, st, Q, fsd));
  BIST_TRY(lin(c, W.n3, Qd, adt, L.w1, fs + L.b1, W.ffh, (long long)Q * dff, adt, (int)dff, di, 1, nullptr, 0, adt, "query_side_gemm"));
  if (c.p_ffn > 0.f)   // modules.py:113
    BIST_TRY(dropout_add(W.ffh, nullptr, (int64_t)nd * Q, dff, 1, c.drop(S_FF_HID, c.p_ffn), 0, 0, adt, st));
  if (c.p_sub > 0.f) {
    BIST_TRY(lin(c, W.ffh, (long long)Q * dff, adt, L.w2, fs + L.b2, x_out[0], out_dir, BIST_F32, di, (int)dff, 0, nullptr, 0, BIST_F32, "query_side_gemm"));
    for (int u = 0; u < nd; ++u)
      BIST_TRY(dropout_add(x_out[u], W.ts + u * Qd, Q, d, 1, c.drop(S_FF_OUT, c.p_sub), 0, 0, BIST_F32, st, (int64_t)u * Q));
  } else {
    BIST_TRY(lin(c, W.ffh, (long long)Q * dff, adt, L.w2, fs + L.b2, x_out[0], out_dir, BIST_F32, di, (int)dff, 0, W.ts, Qd, BIST_F32, "query_side_gemm"));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------- backward
extern "C" int bist_st_layer_bwd(const float* d_out_t2s, const float* d_out_s2t, const float* in_t2s, const float* in_s2t,
                                 const void* vft, const uint8_t* query_mask, const uint8_t* temporal_mask, const void* packed,
                                 const void* tape, size_t tape_bytes, float* d_in_t2s, float* d_in_s2t, float* d_vft,
                                 float* grads, int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t dff, int64_t h,
                                 int n_dirs, int dir_mask, float eps, void* ws, size_t ws_bytes, int dtype,
                                 const bist_dropout_t* dropout, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BIST_CHECK_ARG(vft && query_mask && temporal_mask && packed && tape && ws && d_vft && grads, "st_layer_bwd: null pointer");
  Ctx c;
  BIST_TRY(make_ctx(c, packed, B, T, S, Lq, d, dff, h, n_dirs, dir_mask, eps, dtype, st, dropout));
  const int nd = c.nd;
  const float *x_in[2], *g_out[2];
  float* g_in[2];
  for (int u = 0; u < nd; ++u) {
    x_in[u] = c.kinds[u] == 0 ? in_t2s : in_s2t;
    g_out[u] = c.kinds[u] == 0 ? d_out_t2s : d_out_s2t;
    g_in[u] = c.kinds[u] == 0 ? d_in_t2s : d_in_s2t;
    BIST_CHECK_ARG(x_in[u] && g_out[u] && g_in[u], "st_layer_bwd: null stream pointer for an active direction");
  }
  const Tape W = carve_tape(const_cast<void*>(tape), tape_bytes, B, T, S, Lq, d, dff, h, nd, c.kinds, dtype);
  if (!W.ok) { set_error("st_layer_bwd: tape too small (%zu bytes given)", tape_bytes); return BIST_ERR_WORKSPACE; }
  const BwdWs D = carve_bwd(ws, ws_bytes, B, T, S, Lq, d, dff, h, nd, c.kinds, dtype);
  if (!D.ok) { set_error("st_layer_bwd: workspace too small (%zu bytes given)", ws_bytes); return BIST_ERR_WORKSPACE; }
  const TrainLayout& L = c.L;
  const GradLayout GL = grad_layout(d, dff, n_dirs);
  float* G0 = grads + (size_t)c.pu0 * GL.dir_stride;  // gradient block of slot 0; slot 1 at + gd
  const long long gd = (long long)GL.dir_stride;
  const size_t sdd = (size_t)d * d;
  const int adt = dtype, Q = c.Q, di = c.di, hi = c.hi, dk = c.dk, fi = (int)dff;
  const long long Qd = c.Qd, dd = c.dd, ms = c.ms, fsd = c.fsd, TS = c.TS, Qf = (long long)Q * dff;
  const float* fs = c.fs;
  const bool bf = dtype == BIST_BF16;
  const float rs = 1.f / sqrtf((float)dk);
  const long long go_dir = nd == 2 ? (long long)(g_out[1] - g_out[0]) : 0;
  const long long gi_dir = nd == 2 ? (long long)(g_in[1] - g_in[0]) : 0;
  const long long in_dir = nd == 2 ? (long long)(x_in[1] - x_in[0]) : 0;
  BIST_CHECK_ARG(nd == 1 || (gi_dir != 0 && go_dir != 0), "st_layer_bwd: the two directions need separate gradient tensors");

  // Batched (over directions) weight gradient: C[u] += alpha * A[u]^T B[u], bias[u] += colsum(A[u]).
  auto wg = [&](const void* A, long long a_dir, long long lda, int a_dt, const void* Bm, long long b_dir, long long ldb, int b_dt,
                size_t c_off, size_t bias_off, bool has_bias, int N, int K, const char* tag) {
    WgradArgs w;
    w.A = A; w.B = Bm; w.C = G0 + c_off; w.colsum = has_bias ? G0 + bias_off : nullptr;
    w.M = Q; w.N = N; w.K = K; w.lda = lda; w.ldb = ldb; w.ldc = K;
    w.Z = nd; w.zin = 1;
    w.az = {a_dir, 0}; w.bz = {b_dir, 0}; w.cz = {gd, 0}; w.sz = {gd, 0};
    w.a_dtype = a_dt; w.b_dtype = b_dt; w.tag = tag;
    return wgrad(w, st);
  };
  // Batched (over directions) LayerNorm backward into the fp32 residual-stream gradient.
  auto lnb = [&](const void* dy, const float* x, long long /*x rows are [nd*Q, d] dense*/, int ln_idx, float* dx, int accumulate) {
    LnBwdArgs a;
    a.dy = dy; a.x = x; a.a2 = fs + L.ln_a + (size_t)ln_idx * d; a.dx = dx;
    a.da = G0 + GL.ln + (size_t)ln_idx * 2 * d; a.db = a.da + d;
    a.rows = (int64_t)nd * Q; a.d = d; a.group_rows = Q; a.pstride = fsd; a.gstride = gd; a.eps = eps;
    a.dy_dtype = adt; a.x_dtype = BIST_F32; a.dx_dtype = BIST_F32; a.accumulate = accumulate;
    return layer_norm_bwd(a, st);
  };

  // ---- feed-forward backward: out = ts + relu(n3 W1^T + b1) W2^T + b2
  // gradient of the dropped sublayer output, in the compute dtype: a_dout = m o d_out   (plain cast/copy without dropout)
  for (int u = 0; u < nd; ++u)
    BIST_TRY(dropout_scale(g_out[u], c.act(D.a_dout, u * Qd), Q, d, c.drop(S_FF_OUT, c.p_sub), 0, 0, BIST_F32, adt, st, (int64_t)u * Q));
  const void* a_dout = D.a_dout;
  const long long a_dout_dir = Qd;
  BIST_TRY(lin(c, a_dout, a_dout_dir, adt, L.w2T, nullptr, D.dffh, Qf, adt, fi, di, 0, nullptr, 0, adt, "bwd_dx_gemm"));
  BIST_TRY(relu_gate(D.dffh, W.ffh, (int64_t)nd * Qf, adt, st, c.p_ffn > 0.f ? 1.f / (1.f - c.p_ffn) : 1.f));
  BIST_TRY(wg(a_dout, a_dout_dir, d, adt, W.ffh, Qf, dff, adt, GL.ff + (size_t)d * dff + dff, GL.ff + 2 * (size_t)d * dff + dff, true, di, fi, "wgrad_q"));
  BIST_TRY(lin(c, D.dffh, Qf, adt, L.w1T, nullptr, D.dn, Qd, adt, di, fi, 0, nullptr, 0, adt, "bwd_dx_gemm"));
  BIST_TRY(wg(D.dffh, Qf, dff, adt, W.n3, Qd, d, adt, GL.ff, GL.ff + (size_t)d * dff, true, fi, di, "wgrad_q"));
  // dts = d_out + dLN3(dn3; ts)
  for (int u = 0; u < nd; ++u) BIST_CUDA(cudaMemcpyAsync(D.dts + u * Qd, g_out[u], Qd * 4, cudaMemcpyDeviceToDevice, st));
  BIST_TRY(lnb(D.dn, W.ts, 0, 3, D.dts, 1));

  // ---- ts = x + c2 W_o2^T + b_o2
  BIST_TRY(dropout_scale(D.dts, D.a_dts, (int64_t)nd * Q, d, c.drop(S_ST2_OUT, c.p_sub), 0, 0, BIST_F32, adt, st));
  const void* a_dts = D.a_dts;
  BIST_TRY(lin(c, a_dts, Qd, adt, L.wo2T, nullptr, D.dc, Qd, adt, di, di, 0, nullptr, 0, adt, "bwd_dx_gemm"));   // dc2
  BIST_TRY(wg(a_dts, Qd, d, adt, W.c2, Qd, d, adt, GL.W(2, 3, sdd), GL.b(2, 3, sdd, d), true, di, di, "wgrad_q"));
  // from here on D.dts is the running gradient of x (dx): x feeds ts directly (residual)

  // ---- c2[q, hh-slice] = u2[q,hh,:] . W_v2[hh-slice,:]^T + b_v2
  {
    GemmArgs g;  // du2[q,hh,:] = dc2[q, hh-slice] . W_v2[hh-slice, :]  ->  NT product against W_v2^T[:, hh-slice]
    g.A = D.dc; g.W = c.Mw(L.wv2T); g.C = D.du2;
    g.M = Q; g.N = di; g.K = dk;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(h * d);
    g.zin = hi; g.Z = nd * hi;
    g.az = {Qd, dk}; g.wz = {ms, dk}; g.cz = {Qd * h, d};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "bwd_fold_gemm";
    BIST_TRY(gemm(g, st));
    WgradArgs w;  // dW_v2[hh-slice, :] += dc2[:, hh-slice]^T u2[:, hh, :] ; db_v2 += colsum(dc2)
    w.A = D.dc; w.B = W.u2; w.C = G0 + GL.W(2, 2, sdd); w.colsum = c.p_attn > 0.f ? nullptr : G0 + GL.b(2, 2, sdd, d);
    w.M = Q; w.N = dk; w.K = di; w.lda = d; w.ldb = h * d; w.ldc = d;
    w.Z = nd * hi; w.zin = hi;
    w.az = {Qd, dk}; w.bz = {Qd * h, d}; w.cz = {gd, (long long)dk * d}; w.sz = {gd, dk};
    w.a_dtype = adt; w.b_dtype = adt; w.tag = "wgrad_q";
    BIST_TRY(wgrad(w, st));
    if (c.p_attn > 0.f)  // c2 += b_v2 * rsum: db_v2 += sum_q dc2 * rsum, drsum = dc2 . b_v2
      for (int u = 0; u < nd; ++u)
        BIST_TRY(bias_rsum_bwd(c.act(D.dc, u * Qd), fs + u * fsd + L.bv2, W.rsum + (long long)u * Q * h, D.drsum + (long long)u * Q * h,
                               G0 + u * gd + GL.b(2, 2, sdd, d), Q, hi, dk, adt, st));
  }

  // ---- stage 2 backward, then out1 = x + dropout(c1 W_o1^T + b_o1)
  for (int u = 0; u < nd; ++u) {
    const bool t2s = c.kinds[u] == 0;
    const int G = (int)(t2s ? S : T);
    Stage2BwdArgs a;
    a.qf = c.act(W.qf12, (nd + u) * Qd * h); a.x1 = W.out1[u]; a.du = c.act(D.du2, u * Qd * h);
    a.mask = t2s ? nullptr : temporal_mask;
    a.dx1 = D.dout1[u]; a.dqf = c.act(D.dqf12, (nd + u) * Qd * h);
    a.B = (int)B; a.G = G; a.Lq = (int)Lq; a.h = hi; a.d = di; a.dtype = adt;
    a.drop = c.drop(S_ST2_P + u, c.p_attn);
    a.drsum = c.p_attn > 0.f ? D.drsum + (long long)u * Q * h : nullptr;
    BIST_TRY(stage2_bwd(a, st));
    BIST_TRY(group_sum(D.dout1[u], D.dts + u * Qd, Q, G, d, adt, st));  // dx += sum_g dout1 (residual branch, not dropped)
    if (c.p_sub > 0.f)
      BIST_TRY(dropout_scale(D.dout1[u], D.dout1[u], (int64_t)Q * G, d, c.drop(S_ST1_OUT + u, c.p_sub), (int)Lq, G, adt, adt, st));
    {
      GemmArgs g;  // dc1 = dout1 . W_o1
      g.A = D.dout1[u]; g.W = c.Mw(L.wo1T + (size_t)u * ms); g.C = D.dc1[u];
      g.M = Q * G; g.N = di; g.K = di;
      g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(d);
      g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "bwd_o1_gemm";
      BIST_TRY(gemm(g, st));
      WgradArgs w;  // dW_o1 += dout1^T c1 ; db_o1 += colsum(dout1)
      w.A = D.dout1[u]; w.B = W.ctx1[u]; w.C = G0 + u * gd + GL.W(1, 3, sdd); w.colsum = G0 + u * gd + GL.b(1, 3, sdd, d);
      w.M = Q * G; w.N = di; w.K = di; w.lda = d; w.ldb = d; w.ldc = d;
      w.a_dtype = adt; w.b_dtype = adt; w.tag = "wgrad_o1";
      BIST_TRY(wgrad(w, st));
    }
  }

  // ---- stage 1 backward
  const long long xvld = (long long)nd * d;
  BIST_CUDA(cudaMemsetAsync(D.dqf1, 0, (size_t)nd * Qd * h * 4, st));
  for (int u = 0; u < nd; ++u) {
    const bool t2s = c.kinds[u] == 0;
    const int G = (int)(t2s ? S : T);
    Stage1BwdArgs a;
    a.f = stage1_args(t2s, t2s && c.regions, W.sc1[u], temporal_mask, c.act(W.xv, (long long)u * d), nullptr, B, T, S, Lq, d, h, nd, dtype,
                      c.drop(S_ST1_P + u, c.p_attn));
    a.dctx = D.dc1[u];
    a.dV = c.act(D.dxv, (long long)u * d);
    const int R = (int)(Lq * h);
    a.ds = nullptr; a.dsT = c.act(D.dsT, (long long)u * R); a.dsT_ld = nd * R; a.gdiv = G; a.TS = (int)TS;   // dsT [B][TS][nd * R]
    a.ts_g = t2s ? 1 : S; a.ts_k = t2s ? S : 1;
    a.ds_dtype = adt;
    BIST_TRY(stage1_bwd(a, st));
    {
      // qf1^T [B][d][R] of this direction for the product into d_vft below
      BIST_TRY(transpose_batched(c.act(W.qf12, (long long)u * Qd * h), c.act(D.qf1T, (long long)u * B * d * R), B, R, d, adt, st));
      WgradArgs v;  // dqf1[b][r, :] += sum_ts dsT[b][ts, r] vft[b][ts, :]
      v.A = c.act(D.dsT, (long long)u * R); v.B = vft; v.C = D.dqf1 + u * Qd * h;
      v.M = (int)TS; v.N = R; v.K = di; v.lda = nd * R; v.ldb = d; v.ldc = d;
      v.Z = (int)B; v.zin = 1;
      v.az = {TS * nd * R, 0}; v.bz = {TS * d, 0}; v.cz = {(long long)R * d, 0};
      v.a_dtype = adt; v.b_dtype = dtype; v.tag = "bwd_scores_dqf";
      BIST_TRY(wgrad(v, st));
    }
  }
  {
    // d_vft[b][ts, :] += sum_u sum_r dsT_u[b][ts, r] qf1_u[b][r, :]: NT products against qf1_u^T [d, R], the directions K-concatenated
    const int R = (int)(Lq * h);
    GemmArgs g;
    g.A = D.dsT; g.W = D.qf1T; g.C = d_vft; g.R = d_vft;
    g.M = (int)TS; g.N = di; g.K = R;
    g.kcat = nd; g.a_cat = R; g.w_cat = (long long)B * d * R;
    g.a = rows_dense((long long)nd * R); g.w = rows_dense(R); g.c = rows_dense(d); g.r = rows_dense(d);
    g.Z = (int)B; g.zin = 1;
    g.az = {TS * nd * R, 0}; g.wz = {(long long)d * R, 0}; g.cz = {TS * d, 0}; g.rz = {TS * d, 0};
    g.a_dtype = adt; g.w_dtype = adt; g.c_dtype = BIST_F32; g.r_dtype = BIST_F32; g.tag = "bwd_scores_dvft";
    BIST_TRY(gemm(g, st));
  }
  {
    // d_vft += sum_u dXV[:, u-slice] . W_v1[u]: the directions' value gradients sit side by side in dXV's rows and their weights one
    // after the other, so the sum over u is ONE K-concatenated product and d_vft (fp32, 84 MB at config 3) is read and written once
    GemmArgs g;
    g.A = c.act(D.dxv, 0); g.W = c.Mshared(L.wv1T + (size_t)c.pu0 * dd); g.C = d_vft; g.R = d_vft;
    g.M = (int)(B * TS); g.N = di; g.K = di;
    g.kcat = nd; g.a_cat = d; g.w_cat = dd;
    g.a = rows_dense(xvld); g.w = rows_dense(d); g.c = rows_dense(d); g.r = rows_dense(d);
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = BIST_F32; g.r_dtype = BIST_F32; g.tag = "bwd_xv_gemm";
    BIST_TRY(gemm(g, st));
  }
  {
    WgradArgs w;  // dW_v1[u] += dXV[:, u-slice]^T vft ; db_v1[u] += colsum
    w.A = D.dxv; w.B = vft; w.C = G0 + GL.W(1, 2, sdd); w.colsum = G0 + GL.b(1, 2, sdd, d);
    w.M = (int)(B * TS); w.N = di; w.K = di; w.lda = xvld; w.ldb = d; w.ldc = d;
    w.Z = nd; w.zin = 1;
    w.az = {d, 0}; w.bz = {0, 0}; w.cz = {gd, 0}; w.sz = {gd, 0};
    w.a_dtype = adt; w.b_dtype = dtype; w.tag = "wgrad_xv";
    BIST_TRY(wgrad(w, st));
  }
  // dqf1 (fp32 accumulator) -> compute dtype beside dqf2: dqf12 = [dqf1 | dqf2]
  BIST_TRY(cast_copy(D.dqf1, D.dqf12, (int64_t)nd * Qd * h, BIST_F32, adt, st));

  // ---- fold backward: qf[which][u][q,hh,:] = q[which][u][q, hh-slice] . WkT[which][:, hh-slice]^T
  {
    GemmArgs g;  // dq[q, hh-slice] = dqf[q,hh,:] . (W_k/sqrt(dk))[hh-slice, :]^T
    g.A = D.dqf12; g.W = c.Mw(L.wks12); g.C = D.dq12;
    g.M = Q; g.N = dk; g.K = di;
    g.a = rows_dense(h * d); g.w = rows_dense(d); g.c = rows_dense(d);
    g.zin = hi; g.zmid = nd; g.Z = 2 * nd * hi;  // z = (which * nd + u) * h + hh
    g.az = {nd * Qd * h, d, Qd * h}; g.wz = {dd, (long long)dk * d, ms}; g.cz = {nd * Qd, dk, Qd};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "bwd_fold_gemm";
    BIST_TRY(gemm(g, st));
    for (int which = 0; which < 2; ++which) {
      WgradArgs w;  // dW_k[hh-slice, :] += (1/sqrt(dk)) q[:, hh-slice]^T dqf[:, hh, :]
      w.A = c.act(W.q12, which * nd * Qd); w.B = c.act(D.dqf12, which * nd * Qd * h); w.C = G0 + GL.W(1 + which, 1, sdd);
      w.M = Q; w.N = dk; w.K = di; w.lda = d; w.ldb = h * d; w.ldc = d;
      w.Z = nd * hi; w.zin = hi;
      w.az = {Qd, dk}; w.bz = {Qd * h, d}; w.cz = {gd, (long long)dk * d};
      w.alpha = rs; w.a_dtype = adt; w.b_dtype = adt; w.tag = "wgrad_q";
      BIST_TRY(wgrad(w, st));
    }
  }
  // ---- q12 = n12 . {W_q1 | W_q2}^T + b
  {
    GemmArgs g;  // dn12[which][u] = dq12[which][u] . W_q{1,2}
    g.A = D.dq12; g.W = c.Mw(L.wq12T); g.C = D.dn;
    g.M = Q; g.N = di; g.K = di;
    g.a = rows_dense(d); g.w = rows_dense(d); g.c = rows_dense(d);
    g.Z = 2 * nd; g.zin = nd;
    g.az = {nd * Qd, Qd}; g.wz = {dd, ms}; g.cz = {nd * Qd, Qd};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.tag = "bwd_dx_gemm";
    BIST_TRY(gemm(g, st));
    for (int which = 0; which < 2; ++which)
      BIST_TRY(wg(c.act(D.dq12, which * nd * Qd), Qd, d, adt, c.act(W.n12, which * nd * Qd), Qd, d, adt, GL.W(1 + which, 0, sdd),
                  GL.b(1 + which, 0, sdd, d), true, di, di, "wgrad_q"));
  }
  // dx += dLN1(dnA; x) + dLN2(dnB; x)
  BIST_TRY(lnb(D.dn, W.x, 0, 1, D.dts, 1));
  BIST_TRY(lnb(c.act(D.dn, nd * Qd), W.x, 0, 2, D.dts, 1));

  // ---- x = x0 + c0 W_o0^T + b_o0
  BIST_TRY(dropout_scale(D.dts, D.a_dx, (int64_t)nd * Q, d, c.drop(S_SELF_OUT, c.p_sub), 0, 0, BIST_F32, adt, st));
  const void* a_dx = D.a_dx;
  BIST_TRY(lin(c, a_dx, Qd, adt, L.wo0T, nullptr, D.dc, Qd, adt, di, di, 0, nullptr, 0, adt, "bwd_dx_gemm"));  // dc0
  BIST_TRY(wg(a_dx, Qd, d, adt, W.c0, Qd, d, adt, GL.W(0, 3, sdd), GL.b(0, 3, sdd, d), true, di, di, "wgrad_q"));
  {
    AttnBwdArgs a;  // query self-attention backward (modules.py:54-64)
    a.q = W.qkv0; a.k = c.act(W.qkv0, d); a.v = c.act(W.qkv0, 2 * d); a.dctx = D.dc;
    a.dq = D.dqkv0; a.dk_ = c.act(D.dqkv0, d); a.dv = c.act(D.dqkv0, 2 * d);
    a.q_b = Lq * 3 * d; a.q_l = 3 * d; a.k_b = Lq * 3 * d; a.k_l = 3 * d; a.c_b = Lq * d; a.c_l = d;
    a.mask = query_mask; a.mask_mod = (int)B; a.m_b = Lq; a.m_l = 0;
    a.Bp = (int)(nd * B); a.Lq = (int)Lq; a.Lk = (int)Lq; a.h = hi; a.dk = dk; a.dtype = adt;
    a.drop = c.drop(S_SELF_P, c.p_attn);
    BIST_TRY(attn_core_bwd(a, st));
  }
  // dn0 = sum_j dqkv0[:, j-block] . W_j   (three accumulating products against the transposed weights)
  for (int j = 0; j < 3; ++j) {
    GemmArgs g;
    g.A = c.act(D.dqkv0, (long long)j * d); g.W = c.Mw(L.wqkv0T + (size_t)j * dd); g.C = D.dn; g.R = j ? D.dn : nullptr;
    g.M = Q; g.N = di; g.K = di;
    g.a = rows_dense(3 * d); g.w = rows_dense(d); g.c = rows_dense(d); g.r = rows_dense(d);
    g.Z = nd; g.zin = 1;
    g.az = {3 * Qd, 0}; g.wz = {ms, 0}; g.cz = {Qd, 0}; g.rz = {Qd, 0};
    g.a_dtype = adt; g.w_dtype = dtype; g.c_dtype = adt; g.r_dtype = adt; g.tag = "bwd_dx_gemm";
    BIST_TRY(gemm(g, st));
  }
  BIST_TRY(wg(D.dqkv0, 3 * Qd, 3 * d, adt, W.n0, Qd, d, adt, GL.W(0, 0, sdd), GL.b(0, 0, sdd, d), true, 3 * di, di, "wgrad_q"));
  // d_in = dx + dLN0(dn0; x0)
  for (int u = 0; u < nd; ++u) {
    BIST_CUDA(cudaMemcpyAsync(g_in[u], D.dts + u * Qd, Qd * 4, cudaMemcpyDeviceToDevice, st));
    LnBwdArgs a;
    a.dy = c.act(D.dn, u * Qd); a.x = x_in[u]; a.a2 = fs + u * fsd + L.ln_a; a.dx = g_in[u];
    a.da = G0 + u * gd + GL.ln; a.db = a.da + d;
    a.rows = Q; a.d = d; a.eps = eps;
    a.dy_dtype = adt; a.x_dtype = BIST_F32; a.dx_dtype = BIST_F32; a.accumulate = 1;
    BIST_TRY(layer_norm_bwd(a, st));
  }
  (void)in_dir; (void)bf; (void)go_dir;
  return 0;
}

/* The multipliers (0 or 1/(1-p)) that the kernels apply at one dropout site, for tests that inject the same masks into
 * the oracle.  out fp32 [n]. */
extern "C" int bist_dropout_mask(uint64_t seed, uint32_t site, float p, int64_t n, float* out, void* stream) {
  BIST_CHECK_ARG(p >= 0.f && p < 1.f, "dropout_mask: p must be in [0, 1)");
  return dropout_mask(out, n, make_drop(p, seed, site), (cudaStream_t)stream);
}
