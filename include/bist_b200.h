/*
 * bist_b200.h — C ABI of libbist_b200.so: the B200 (sm_100a) implementation of the BiST
 * video hot path.  The reference (salesforce/BiST) has no FFI of its own:
 * its boundary is the Python nn.Module protocol of model/modules.py and model/encoder.py.
 * Each entry point below therefore replaces the *body* of one reference forward(); the
 * Python mirror in bist_b200/{modules,encoder}.py keeps the reference's constructor and
 * forward signatures and calls these functions through ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless it is named host_*; tensors are dense,
 *     row-major, innermost dimension contiguous; 16-byte aligned base addresses
 *   - the library never allocates or frees device memory and keeps no pointer after a
 *     call returns; scratch is passed in (`ws`, sized by the *_workspace_bytes helpers)
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing
 *     synchronises the host
 *   - return value: 0 = success, non-zero = bist_status_t error; the message is available
 *     from bist_last_error() (thread-local).  Nothing throws across the ABI.  There is no
 *     CPU or eager fallback: unsupported shapes/dtypes are errors.
 *   - dtype arguments take bist_dtype_t values.  "bf16 mode" (BIST_BF16) stores the large
 *     activations and the weights in bf16 and accumulates/normalises/softmaxes in fp32;
 *     LayerNorm gains/offsets and biases are always fp32.
 */
#ifndef BIST_B200_H_
#define BIST_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BIST_B200_VERSION 100 /* 0.1.0 */

typedef enum { BIST_F32 = 0, BIST_BF16 = 1 } bist_dtype_t;

typedef enum {
  BIST_OK = 0,
  BIST_ERR_ARG = 1,       /* null pointer, bad shape, bad enum            */
  BIST_ERR_UNSUPPORTED = 2, /* legal in the reference, not built here     */
  BIST_ERR_WORKSPACE = 3, /* ws too small                                 */
  BIST_ERR_CUDA = 4,      /* a CUDA runtime/driver call failed            */
  BIST_ERR_DEVICE = 5     /* not an sm_100 device                         */
} bist_status_t;

/* mask_mode of bist_mha_fwd */
typedef enum {
  BIST_MASK_NONE = 0, /* mask == NULL                                            */
  BIST_MASK_KEY = 1,  /* uint8 [Bp, Lk]      (reference mask [Bp,1,Lk])          */
  BIST_MASK_FULL = 2  /* uint8 [Bp, Lq, Lk]  (reference mask [Bp,Lq,Lk])         */
} bist_mask_mode_t;

int bist_version(void);
const char* bist_last_error(void);
/* 0 when the current CUDA device is compute capability 10.x, BIST_ERR_DEVICE otherwise. */
int bist_device_check(void);
/* Number of kernels this library has launched in this process (bench.py "gpu_launches"). */
int64_t bist_launch_count(void);
/* Per-launch CUDA-event timing for bench.py's roofline line (off by default, adds two event
 * records per launch).  bist_prof_collect synchronises the device and writes one
 * "tag,launches,total_ms\n" line per kernel tag into the HOST buffer `host_buf`. */
int bist_prof_enable(int on);
int bist_prof_collect(char* host_buf, size_t cap);

/* ---------------------------------------------------------------------------------------
 * LayerNorm.forward — /root/reference/model/modules.py:28-31
 *   y = a_2 * (x - mean) / (std + eps) + b_2, std unbiased (divide by d-1), eps on the std.
 *   x,y [rows, d]; a_2,b_2 fp32 [d].
 */
int bist_ln_fwd(const void* x, const float* a_2, const float* b_2, void* y, int64_t rows,
                int64_t d, float eps, int x_dtype, int y_dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * nn.Linear (+ReLU) (+residual) — the building block of modules.py:74-77,100,106-113
 *   y[M,N] = act(x[M,K] . w[N,K]^T + bias[N]) (+ residual[M,N]);  bias/residual may be NULL.
 */
int bist_linear_fwd(const void* x, const void* w, const float* bias, const void* residual,
                    void* y, int64_t M, int64_t N, int64_t K, int relu, int x_dtype,
                    int w_dtype, int y_dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * MultiHeadedAttention.forward — modules.py:81-100 (+ attention(), modules.py:54-64), eval
 * mode (dropout identity).
 *   query [Bp,Lq,d_in], key/value [Bp,Lk,d_in] (dtype `dtype`); w[0..2] [d,d_in], w[3]
 *   [d_in,d] (dtype `dtype`), b[0..3] fp32; mask per mask_mode (0 = masked -> -1e9);
 *   out [Bp,Lq,d_in]; attn_out (may be NULL) fp32 [Bp,h,Lq,Lk] = the reference's `.attn`.
 */
typedef struct {
  const void* w[4];
  const float* b[4];
} bist_mha_params_t;

size_t bist_mha_workspace_bytes(int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int64_t d_in,
                                int64_t h);
int bist_mha_fwd(const void* query, const void* key, const void* value, const uint8_t* mask,
                 int mask_mode, const bist_mha_params_t* host_params, void* out,
                 float* attn_out, int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int64_t d_in,
                 int64_t h, void* ws, size_t ws_bytes, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * PositionwiseFeedForward.forward — modules.py:112-113 (eval): y = w_2(relu(w_1 x)).
 *   x [rows,d]; w1 [d_ff,d]; w2 [d_out,d_ff]; b1,b2 fp32; y [rows,d_out].
 *   ws >= rows*d_ff*sizeof(dtype).
 */
int bist_ffn_fwd(const void* x, const void* w1, const float* b1, const void* w2,
                 const float* b2, void* y, int64_t rows, int64_t d, int64_t d_ff,
                 int64_t d_out, void* ws, size_t ws_bytes, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * VidEncoder8.forward, video branch — /root/reference/model/encoder.py:72-80
 *   out[n,:] = LayerNorm(relu(fts[n,:] . W^T + bias))       (vid_position is always None)
 *   fts fp32 [N, D] (N = B*T*S rows of the [B,T,S,D] tensor); W fp32 [d, D]; bias, a_2, b_2
 *   fp32 [d]; out [N, d] of out_dtype.  compute_dtype BIST_F32: fp32-accurate contraction;
 *   BIST_BF16: bf16 operands, fp32 accumulate.  The same call serves the audio branch
 *   (encoder.py:83-91) with D = 128.
 */
size_t bist_vidproj_workspace_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype);
int bist_vidproj_fwd(const float* fts, const float* W, const float* bias, const float* a_2,
                     const float* b_2, void* out, int64_t N, int64_t D, int64_t d, float eps,
                     void* ws, size_t ws_bytes, int compute_dtype, int out_dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * VidEncoderLayer4.forward with enc_st_combine == 'none' — encoder.py:172-201, including
 * temporal2spatial (encoder.py:109-139) and spatial2temporal (encoder.py:141-170), eval mode.
 *
 * Parameters are handed over once per weight update as fp32 device pointers in the
 * reference's clone order (state_dict names attn.{i}.linears.{j}, ff.{i}.w_{1,2},
 * sublayer.{i}.norm) and packed by bist_st_layer_pack into `packed` (layout private to the
 * library: fused QKV rows, pre-scaled transposed key weights for the query-side fold, one
 * direction after the other at a constant stride).  n_dirs = 2 -> 6 attn / 2 ff / 8 norms
 * (direction 0 = t2s, 1 = s2t); n_dirs = 1 -> 3 attn / 1 ff / 4 norms (mtn.py:123-128).
 */
typedef struct {
  const float* attn_w[6][4]; /* attn.{i}.linears.{j}.weight [d,d] */
  const float* attn_b[6][4]; /* attn.{i}.linears.{j}.bias   [d]   */
  const float* ff_w1[2];     /* ff.{i}.w_1.weight [d_ff,d]        */
  const float* ff_b1[2];
  const float* ff_w2[2];     /* ff.{i}.w_2.weight [d,d_ff]        */
  const float* ff_b2[2];
  const float* ln_a[8];      /* sublayer.{i}.norm.a_2 [d]         */
  const float* ln_b[8];
} bist_st_layer_params_t;

size_t bist_st_layer_packed_bytes(int64_t d, int64_t d_ff, int n_dirs, int dtype);
int bist_st_layer_pack(const bist_st_layer_params_t* host_params, void* packed,
                       size_t packed_bytes, int64_t d, int64_t d_ff, int64_t h, int n_dirs,
                       int dtype, void* stream);

size_t bist_st_layer_workspace_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d,
                                     int64_t d_ff, int64_t h, int n_dirs, int dtype);
/*
 *   in_t2s / in_s2t   fp32 [B,Lq,d]   in_ft['t2s'], in_ft['s2t'] (may alias each other)
 *   vft               dtype [B,T,S,d]  ft['spatiotemporal_ft'] (read-only)
 *   query_mask        uint8 [B,Lq]     b.query_mask [B,1,Lq]
 *   temporal_mask     uint8 [B,T]      b.temporal_mask [B,1,T]
 *   out_t2s / out_s2t fp32 [B,Lq,d]    new in_ft['t2s'], in_ft['s2t'] (must not alias inputs)
 *   dir_mask          bit0 = run t2s, bit1 = run s2t.  With n_dirs == 1 the single packed
 *                     direction is applied to whichever one bit is set.
 */
int bist_st_layer_fwd(const float* in_t2s, const float* in_s2t, const void* vft,
                      const uint8_t* query_mask, const uint8_t* temporal_mask,
                      const void* packed, float* out_t2s, float* out_s2t, int64_t B, int64_t T,
                      int64_t S, int64_t Lq, int64_t d, int64_t d_ff, int64_t h, int n_dirs,
                      int dir_mask, float eps, void* ws, size_t ws_bytes, int dtype,
                      void* stream);

/* =======================================================================================
 * Training step: forward with a tape + backward.
 *
 * The reference trains through torch autograd over the modules of model/modules.py and
 * model/encoder.py; these entry points replace the autograd graph of the same forward()s.
 * Gradients are fp32 and ACCUMULATED (+=) into caller-owned buffers (zero them for a fresh
 * gradient).  Parity is against torch.autograd over the oracle (tests/test_gpu_backward.py).
 * ======================================================================================= */

/* VidEncoder8.forward for training — encoder.py:72-80.  Same output as bist_vidproj_fwd; `tape`
 * (bist_vidproj_tape_bytes) receives relu(fts W^T + b) as fp32 [N, d] and, in bf16 mode, bf16
 * copies of fts and W that the forward contraction and the weight gradient both read. */
size_t bist_vidproj_tape_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype);
int bist_vidproj_train_fwd(const float* fts, const float* W, const float* bias, const float* a_2,
                           const float* b_2, void* out, void* tape, size_t tape_bytes, int64_t N, int64_t D,
                           int64_t d, float eps, int compute_dtype, int out_dtype, void* stream);
size_t bist_vidproj_bwd_workspace_bytes(int64_t N, int64_t D, int64_t d, int compute_dtype);
/*   d_out fp32 [N, d]: gradient w.r.t. ft['spatiotemporal_ft'] (summed over blocks/directions);
 *   dW [d, D], dbias [d], da_2 [d], db_2 [d] fp32, accumulated.  fts is data: no gradient. */
int bist_vidproj_bwd(const float* d_out, const float* fts, const void* tape, size_t tape_bytes,
                     const float* a_2, float* dW, float* dbias, float* da_2, float* db_2, int64_t N,
                     int64_t D, int64_t d, float eps, void* ws, size_t ws_bytes, int compute_dtype,
                     void* stream);

/* VidEncoderLayer4 for training — encoder.py:172-201 (+ :109-170), enc_st_combine == 'none'.
 *
 * bist_st_layer_train_pack: like bist_st_layer_pack, for the training formulation (the stage-1
 * output projection is applied explicitly because the reference's sublayer dropout,
 * modules.py:44, sits between it and stage 2), plus the transposed operands of the backward.
 * `tape` (bist_st_layer_tape_bytes) is written by the forward and must reach the backward
 * unchanged together with the same inputs, masks and packed weights.
 *
 * Gradient block `grads` (fp32, bist_st_layer_grad_elems floats), per packed direction u
 * (0 = t2s, 1 = s2t for n_dirs == 2), in the reference's parameter order:
 *     attn.{3u+i}, i = 0..2 :  linears.{0..3}.weight [4][d,d], then linears.{0..3}.bias [4][d]
 *     ff.{u}                :  w_1.weight [d_ff,d], w_1.bias [d_ff], w_2.weight [d,d_ff], w_2.bias [d]
 *     sublayer.{4u+i}.norm, i = 0..3 : a_2 [d], b_2 [d]
 * Key-projection biases (linears.1.bias of the stage-1 / stage-2 attentions) get exactly zero:
 * their score contribution is constant over the keys and cancels in the softmax.
 */
/* Dropout of the training forward (model.train()): p_attn on the attention probabilities (modules.py:62-63, fixed 0.1
 * by mtn.py:77), p_sublayer on every sublayer output (modules.py:44), p_ffn on the FFN hidden layer (modules.py:113).
 * Masks are counter-based — keep(seed, site, element index in the REFERENCE's tensor at that site) — regenerated by
 * the backward from the same struct, never stored.  Sites of one layer call (u = slot of the active direction, 0 or 1):
 *   0 self-attention probabilities [nd*B, h, Lq, Lq]     1 self-attention sublayer output [nd*B*Lq, d]
 *   2+u stage-1 probabilities [B*G, h, Lq, nk]            4+u stage-1 sublayer output [B*G, Lq, d]
 *   6+u stage-2 probabilities [B*Lq, h, 1, G]             8 stage-2 sublayer output [nd*B*Lq, d]
 *   9 FFN hidden [nd*B*Lq, d_ff]                          10 FFN sublayer output [nd*B*Lq, d]
 * NULL (or all p == 0) = eval mode.  The stream of random numbers differs from torch's; the distribution is the same. */
typedef struct {
  float p_attn, p_sublayer, p_ffn;
  uint64_t seed;
} bist_dropout_t;

/* Dropout under CUDA-graph replay: when a device counter is registered, every dropout site hashes with
 * seed + *device_counter, read by the kernels at run time — a captured training step that increments the counter as
 * its first node draws fresh masks on every replay (bist_b200/graph.py).  NULL unregisters.  Process-wide. */
int bist_set_dropout_step(const uint64_t* device_counter);
/* out[i] = multiplier (0 or 1/(1-p)) of element i of `site`, fp32 [n] (tests inject these into the oracle). */
int bist_dropout_mask(uint64_t seed, uint32_t site, float p, int64_t n, float* out, void* stream);

size_t bist_st_layer_train_packed_bytes(int64_t d, int64_t d_ff, int n_dirs, int dtype);
int bist_st_layer_train_pack(const bist_st_layer_params_t* host_params, void* packed, size_t packed_bytes,
                             int64_t d, int64_t d_ff, int64_t h, int n_dirs, int dtype, void* stream);
size_t bist_st_layer_grad_elems(int64_t d, int64_t d_ff, int n_dirs);
size_t bist_st_layer_tape_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t d_ff,
                                int64_t h, int n_dirs, int dtype);
size_t bist_st_layer_bwd_workspace_bytes(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d,
                                         int64_t d_ff, int64_t h, int n_dirs, int dtype);
/* Arguments as bist_st_layer_fwd; `packed` from bist_st_layer_train_pack. */
int bist_st_layer_train_fwd(const float* in_t2s, const float* in_s2t, const void* vft,
                            const uint8_t* query_mask, const uint8_t* temporal_mask, const void* packed,
                            float* out_t2s, float* out_s2t, int64_t B, int64_t T, int64_t S, int64_t Lq,
                            int64_t d, int64_t d_ff, int64_t h, int n_dirs, int dir_mask, float eps,
                            void* tape, size_t tape_bytes, int dtype, const bist_dropout_t* dropout, void* stream);
/*   d_out_*   fp32 [B,Lq,d]  gradients of the layer outputs (separate tensors per direction)
 *   d_in_*    fp32 [B,Lq,d]  gradients of the layer inputs (written, not accumulated)
 *   d_vft     fp32 [B,T,S,d] gradient of ft['spatiotemporal_ft'], ACCUMULATED
 *   grads     the gradient block described above, ACCUMULATED */
int bist_st_layer_bwd(const float* d_out_t2s, const float* d_out_s2t, const float* in_t2s,
                      const float* in_s2t, const void* vft, const uint8_t* query_mask,
                      const uint8_t* temporal_mask, const void* packed, const void* tape,
                      size_t tape_bytes, float* d_in_t2s, float* d_in_s2t, float* d_vft, float* grads,
                      int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t d_ff, int64_t h,
                      int n_dirs, int dir_mask, float eps, void* ws, size_t ws_bytes, int dtype,
                      const bist_dropout_t* dropout, void* stream);

/* LayerNorm backward — modules.py:28-31.  dx (dx_dtype) written; da_2/db_2 fp32 accumulated
 * (both may be NULL). */
int bist_ln_bwd(const void* dy, const void* x, const float* a_2, void* dx, float* da_2, float* db_2,
                int64_t rows, int64_t d, float eps, int dy_dtype, int x_dtype, int dx_dtype, void* stream);

/* nn.Linear backward: dx[M,K] = dy[M,N] . w[N,K] (dx may be NULL), dw[N,K] += dy^T x,
 * db[N] += colsum(dy) (dw/db fp32, accumulated, may be NULL together).  wT is w transposed
 * ([K,N], dtype w_dtype), needed only when dx != NULL. */
int bist_linear_bwd(const void* dy, const void* x, const void* wT, void* dx, float* dw, float* db,
                    int64_t M, int64_t N, int64_t K, int dy_dtype, int x_dtype, int w_dtype, int dx_dtype,
                    void* stream);

/* ---- stand-alone modules, training (modules.py:33-113 used outside VidEncoderLayer4: the decoder, caption and audio
 * layers of decoder.py:11-60 / encoder.py:203-235 are compositions of these) ----------------------------------------
 * Dropout masks are counter based, as on the video path: element i of the reference's tensor at a site is kept iff
 * hash(seed, site, i) >= p * 2^32; the backward regenerates them from the same (seed, site).
 *
 * bist_dropout_fwd: y = x o m / (1 - p) over a dense [rows, d] tensor (nn.Dropout of SublayerConnection,
 * modules.py:44); its own backward is the same call on dy. */
int bist_dropout_fwd(const void* x, void* y, int64_t rows, int64_t d, float p, uint64_t seed, uint32_t site,
                     int dtype, void* stream);

/* Refresh many weight copies at once: item i writes dst[dst_off + r*cols + c] = transpose ? src[c*rows + r] : src[r*cols + c]
 * (rows x cols = shape of the DESTINATION; src fp32 parameters, dst of `dtype`), 72 items per launch. */
typedef struct {
  const float* src;
  int64_t dst_off;
  int32_t rows, cols;
  int32_t transpose;
  int32_t reserved;
} bist_pack_item_t;
int bist_pack_params(const bist_pack_item_t* items, int64_t n, void* base, int dtype, void* stream);

/* SublayerConnection's residual fused with its dropout: out = x + y o m / (1 - p) (modules.py:44), dense [rows, d]. */
int bist_dropout_add_fwd(const void* y, const void* x, void* out, int64_t rows, int64_t d, float p, uint64_t seed,
                         uint32_t site, int dtype, void* stream);

/* MultiHeadedAttention.forward in train() mode (modules.py:81-100): as bist_mha_fwd, with dropout p_drop on the
 * probabilities (modules.py:62-63; attn_out receives the dropped probabilities, as self.attn does) and the workspace
 * kept as the tape of bist_mha_bwd (tape_bytes = bist_mha_workspace_bytes). */
int bist_mha_train_fwd(const void* query, const void* key, const void* value, const uint8_t* mask, int mask_mode,
                       const bist_mha_params_t* params, void* out, float* attn_out, int64_t Bp, int64_t Lq,
                       int64_t Lk, int64_t d, int64_t d_in, int64_t h, void* tape, size_t tape_bytes, int dtype,
                       float p_drop, uint64_t seed, uint32_t site, void* stream);
size_t bist_mha_bwd_workspace_bytes(int64_t Bp, int64_t Lq, int64_t Lk, int64_t d, int dtype);
/* Backward of the same.  params_T->w[i] = linears[i].weight TRANSPOSED ([in, out], dtype); params_T->b is not read.
 * dW[i] ([out, in]) and db[i] are fp32 and accumulated into.  d_query / d_key / d_value ([Bp, L, d_in], dtype) may
 * each be NULL.  Limits: (2 Lq + 2 Lk)(d_k + 1) + 2 Lq (Lk + 1) floats must fit 200 KB of shared memory. */
int bist_mha_bwd(const void* d_out, const void* query, const void* key, const void* value, const uint8_t* mask,
                 int mask_mode, const bist_mha_params_t* params_T, const void* tape, size_t tape_bytes,
                 void* d_query, void* d_key, void* d_value, float* const* dW, float* const* db, int64_t Bp,
                 int64_t Lq, int64_t Lk, int64_t d, int64_t d_in, int64_t h, void* ws, size_t ws_bytes, int dtype,
                 float p_drop, uint64_t seed, uint32_t site, void* stream);

/* PositionwiseFeedForward in train() mode (modules.py:112-113).  hidden [rows, d_ff] (dtype) is written by the
 * forward (the ReLU output after dropout) and read by the backward.  w1T [d, d_ff], w2T [d_ff, d_out]. */
int bist_ffn_train_fwd(const void* x, const void* w1, const float* b1, const void* w2, const float* b2, void* y,
                       void* hidden, int64_t rows, int64_t d, int64_t d_ff, int64_t d_out, int dtype, float p_drop,
                       uint64_t seed, uint32_t site, void* stream);
int bist_ffn_bwd(const void* dy, const void* x, const void* w1T, const void* w2T, const void* hidden, void* dx,
                 float* dW1, float* db1, float* dW2, float* db2, int64_t rows, int64_t d, int64_t d_ff,
                 int64_t d_out, void* ws, size_t ws_bytes, int dtype, float p_drop, void* stream);

/* Softmax gate of MultimodalDecoder8 (decoder.py:150-181): out[r,:] = sum_k softmax(logits[r,0:K])_k * src[k][r,:],
 * K <= 4 streams of [R, d] (dtype); logits fp32 [R, K] come from the gate Linear (bist_linear_fwd with an fp32
 * output).  Backward: d_src[k] = p_k * d_out, d_logits = softmax backward of dp_k = <d_out, src[k]>. */
int bist_gate_combine_fwd(const void* const* src, const float* logits, void* out, int64_t R, int64_t d, int K,
                          int dtype, void* stream);
int bist_gate_combine_bwd(const void* d_out, const void* const* src, const float* logits, void* const* d_src,
                          float* d_logits, int64_t R, int64_t d, int K, int dtype, void* stream);

/* ---- feature ingest ------------------------------------------------------------------------
 * VidEncoder8's projection (encoder.py:72-80) for features already held in bf16 (bist_b200/ingest.py keeps the feature
 * cache in 2 bytes per value on the host and the device): fts16 bf16 [N, D], W16 bf16 [d, D], ws >= N*d*4 bytes. */
int bist_vidproj_fwd_bf16in(const void* fts16, const void* W16, const float* bias, const float* a_2, const float* b_2,
                            void* out, int64_t N, int64_t D, int64_t d, float eps, void* ws, size_t ws_bytes,
                            int out_dtype, void* stream);

/* No-allocation variant of bist_vidproj_fwd (bf16 compute).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).
 * Same inputs, outputs and result as bist_vidproj_fwd(..., compute_dtype = BIST_BF16, ...), but the scratch (the bf16 copy
 * of W) is a static device array inside the library: d*D*2 <= BIST_VIDPROJ_NOALLOC_MAX_BYTES, else BIST_ERR_WORKSPACE;
 * shapes the fused kernel does not take (d % 64 != 0, d > 512, D % 32 != 0) return BIST_ERR_UNSUPPORTED.  Calls no
 * cudaMalloc / cudaFree.  All calls on a device share the array: calls that overlap on the GPU race. */
#define BIST_VIDPROJ_NOALLOC_MAX_BYTES (2 * 1024 * 1024)
int bist_vidproj_fwd_noalloc(const float* fts, const float* W, const float* bias, const float* a_2, const float* b_2,
                             void* out, int64_t N, int64_t D, int64_t d, float eps, int out_dtype, void* stream);

/* Adam step for every parameter in ONE launch — the optimizer of the reference's training loop (torch.optim.Adam behind NoamOpt,
 * model/optimize.py:9-34; no weight decay, no amsgrad), same arithmetic as torch.optim.Adam.  `rows`: DEVICE array of
 * {float* p, const float* g, float* m, float* v, int64 n} records (40 bytes each); `chunks`: DEVICE array of {int row, int piece}
 * pairs, one per piece of bist_adam_chunk_elems() elements of a row; step >= 1 is the 1-based step count (bias correction). */
int bist_adam_chunk_elems(void);
int bist_adam_step(const void* rows, const void* chunks, int n_chunks, float lr, float beta1, float beta2, float eps, int64_t step,
                   void* stream);

/* n ragged host->device copies (src[i]: page-locked host rows, nbytes[i] to dst_base + dst_off[i]) each followed by the
 * zero fill of that video's padding (pad_nbytes[i] at dst_base + pad_off[i]), all on `stream`: the collate of
 * data/dataset.py:164-174 without a padded host copy. */
int bist_ingest_gather(const void* const* src, void* dst_base, const int64_t* dst_off, const int64_t* nbytes,
                       const int64_t* pad_off, const int64_t* pad_nbytes, int64_t n, void* stream);

/* ---- text side and generator ------------------------------------------------------------
 * Embeddings + PositionalEncoding (modules.py:115-141): out[r,:] = dropout(lut[idx[r],:] * sqrt(d) + pe[r % L,:]),
 * rows = B*L token positions, lut fp32 [vocab, d], pe fp32 [>= L, d], out in out_dtype.  Backward scatter-adds
 * sqrt(d) * mask * dy[r,:] into dlut[idx[r],:] (fp32, accumulated). */
int bist_embed_fwd(const int64_t* idx, const float* lut, const float* pe, void* out, int64_t rows, int64_t L,
                   int64_t d, int64_t vocab, float p_drop, uint64_t seed, uint32_t site, int out_dtype, void* stream);
int bist_embed_bwd(const int64_t* idx, const void* dy, float* dlut, int64_t rows, int64_t d, float p_drop,
                   uint64_t seed, uint32_t site, int dy_dtype, void* stream);

/* Pointer attention of the generators (generator.py:62-63,104-114): one head over projected q [B,Lt,d] / k [B,Ls,d]
 * (scale 1/sqrt(d)), key mask [B,Ls] (0 = masked -> -1e9) or NULL; P [B,Lt,Ls] fp32 are the pointer probabilities,
 * tvec [B,Lt,d] = P . v with v the RAW encoded text.  Backward takes dP and dtvec, returns dq, dk, dv. */
int bist_ptr_attn_fwd(const void* q, const void* k, const void* v, const uint8_t* mask, float* P, void* tvec,
                      int64_t B, int64_t Lt, int64_t Ls, int64_t d, int dtype, void* stream);
int bist_ptr_attn_bwd(const float* dP, const void* dtvec, const void* q, const void* k, const void* v, const float* P,
                      void* dq, void* dk, void* dv, int64_t B, int64_t Lt, int64_t Ls, int64_t d, int dtype,
                      void* stream);

/* Generator + label-smoothed loss, forward AND gradient in one pass over the logits (generator.py:11-127,
 * label_smoothing.py:20-30, optimize.py:46-82).  Z [R, V] (dtype) = decoded_text . E^T on entry; when want_grad it is
 * overwritten with d(sum_r row_loss[r]) / dZ.  n_ptr in {0,1,2} pointer sources: P[i] [R, Ls[i]] fp32 (from
 * bist_ptr_attn_fwd), text[i] [R / Lt, Ls[i]] token ids, dP[i] out; sw [R, n_ptr+1] fp32 switch logits (softmax inside,
 * the vocabulary distribution's weight LAST), dsw out; both NULL with n_ptr == 0 (plain log_softmax: the auto-encoder
 * heads).  target [R] (pad rows contribute nothing; NULL = inference), row_loss [R] fp32 un-normalised KL per row
 * with smoothing mass smoothing/(V-2) and the pad column zeroed; logp (optional) [R, V] fp32 = log p_out. */
int bist_genloss_fwd_bwd(void* Z, float* logp, int n_ptr, const float* const* P, float* const* dP,
                         const int64_t* const* text, const int64_t* Ls, const float* sw, float* dsw,
                         const int64_t* target, float* row_loss, int64_t R, int64_t V, int64_t Lt, int64_t pad,
                         float smoothing, int want_grad, int dtype, void* stream);

/* Test switch for the fused query-side kernels of bist_st_layer_fwd (csrc/qside.cu: LN0 + QKV + self-attention, LN1/LN2 + Q
 * projections + key folds, LN3 + first FFN layer — encoder.py:176-201 on the query rows): 0 = the separate launches,
 * 1 = fused where supported (bf16, d_k = 64, h even, Lq <= 16, d_ff % 256 == 0), -1 = default (on unless BIST_QSIDE_UNFUSED). */
int bist_debug_qside(int on);

/* Stage 1 of temporal2spatial / spatial2temporal for up to two directions as ONE launch (bf16; the video tensor is read
 * once per launch, projected values and scores stay on chip) — encoder.py:109-121 (t2s) and :143-148 (s2t), attention()
 * of modules.py:54-64.  bist_st_layer_fwd calls it when bist_stage1_fused_supported(); exported for the kernel-level tests.
 *   vft [B,T,S,d] bf16; temporal_mask uint8 [B,T] (0 = padded clip; used by t2s);  per direction u < nd:
 *   kinds[u]  0 = t2s (groups = S regions, keys = T clips, key mask), 1 = s2t (groups = T clips, keys = S regions)
 *   qf[u]     bf16 [B*Lq, h, d]  folded queries (q_hh W_k,hh / sqrt(dk)), wv[u] bf16 [d,d] value weight (rows = outputs),
 *   bv[u]     fp32 [d], c1[u] out bf16 [B, Lq, G, d] (G = S for t2s, T for s2t): sum_k softmax_k(qf.x_k) (W_v x_k + b_v)
 *   kinds / qf / wv / bv / c1 are HOST arrays (of device pointers).  dbg: NULL, or fp32 [units, h, 128, 80] raw tiles. */
int bist_stage1_fused_supported(int64_t B, int64_t T, int64_t S, int64_t Lq, int64_t d, int64_t h);
int bist_stage1_fused(const void* vft, const uint8_t* temporal_mask, int nd, const int* kinds, const void* const* qf,
                      const void* const* wv, const float* const* bv, void* const* c1, int64_t B, int64_t T, int64_t S,
                      int64_t Lq, int64_t d, int64_t h, float* dbg, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BIST_B200_H_ */
