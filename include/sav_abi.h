/* sav_abi.h -- C ABI of the B200-native transformer-block kernels (libsav_b200.so).
 *
 * The reference (NZ99/self-attention-experiments-vision) has NO native interface: its only seams
 * are Python call signatures of Flax modules whose arithmetic XLA lowers.  Each entry point below
 * therefore cites the reference call site whose XLA-lowered computation it replaces; these are
 * exactly the symbols an XLA-FFI / ctypes binding for this path binds.
 *
 * Conventions (all entry points):
 *   - plain C, POD args; every pointer is a DEVICE pointer owned by the caller
 *   - the callee never allocates, frees, synchronises or switches device; all work is enqueued on
 *     `stream` (a cudaStream_t passed as void*); CUDA-graph-capture safe; thread-safe
 *   - return 0 on success, negative sav_status otherwise; sav_last_error() is thread-local
 *   - bf16 = IEEE bfloat16 (2 bytes); "row-major [R,C] ld" = element (r,c) at r*ld + c
 */
#ifndef SAV_ABI_H_
#define SAV_ABI_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SAV_ABI_VERSION 3

enum sav_status {
  SAV_OK = 0,
  SAV_ERR_SHAPE = -1,
  SAV_ERR_ALIGN = -2,
  SAV_ERR_ARCH = -3,
  SAV_ERR_CUDA = -4,
  SAV_ERR_ARG = -5
};

int sav_abi_version(void);
const char* sav_last_error(void);
/* 0 when the current device is sm_100 (B200); SAV_ERR_ARCH otherwise. */
int sav_device_check(void);
/* number of kernels launched by this library in this process (bench.py: "gpu_launches") */
int64_t sav_launch_count(void);

/* ---------------------------------------------------------------- GEMM (tcgen05 / TMEM / TMA)
 * D[M,N] = epilogue( sum_k A[m,k] * B[n,k] )           bf16 x bf16 -> fp32 accumulate
 * Replaces nn.Dense / nn.DenseGeneral dot_generals and their reverse-mode products:
 *   attention.py:29-37,60-63 (queries/keys/values/out projections), ff.py:24,27 (Dense_0/Dense_1),
 *   patch_embed.py:23-25, vit.py:96-98 (head), and jax.value_and_grad of those (train.py:94).
 * Operand storage:
 *   a_mn=0: A is row-major [M,K] (K contiguous);  a_mn=1: A is row-major [K,M] (M contiguous)
 *   b_mn=0: B is row-major [N,K] (K contiguous);  b_mn=1: B is row-major [K,N] (N contiguous)
 * Output / epilogue modes (selected from the arguments; anything else is SAV_ERR_ARG):
 *   bf16 D           : v = acc + bias[n]; v *= col_scale[n]; v *= row_scale[m / rows_per_scale];
 *                      D = bf16(act ? gelu_tanh(v) : v)
 *   bf16 D + aux_out : (act=1) aux_out = bf16(acc + bias) [pre-activation saved for backward],
 *                      D = bf16(gelu_tanh(acc + bias))                       (ff.py:24-25)
 *   bf16 D + aux_in  : D = bf16(acc * gelu_tanh'(aux_in[m,n]))               (reverse of ff.py:25)
 *                      (+ dgelu_colsum[n] += sum_m of that product, the reverse of ff.py:24's bias)
 *   fp32 D           : v = acc + bias; v *= col_scale; v *= row_scale; v += residual[m,n] (fp32);
 *                      D = v        -- the fp32 residual stream (vit.py:24,31; layerscale.py:20-23;
 *                      stochastic_depth.py:16-27)
 *   fp32 D, accumulate or split_k != 1 : D += acc, K partitioned over split_k CTAs groups whose
 *                      partial tiles are combined with TMA reduce-add in L2 (dW = X^T dY where
 *                      K = tokens).  split_k = 0 picks the split automatically.  With
 *                      accumulate=0 the callee zero-fills D first.
 * All D / aux / residual traffic goes through TMA (row pitches must be multiples of 16 bytes).
 * Alignment: all base pointers 16 B aligned; lda/ldb/ldd(bf16) multiples of 8, ldd(fp32)/ldr
 * multiples of 4 elements; N % 8 == 0.
 */
typedef struct {
  const void* A;
  const void* B;
  void* D;
  int64_t M, N, K;
  int64_t lda, ldb, ldd;
  int32_t a_mn, b_mn, d_fp32, split_k;
  const float* bias;       /* [N] fp32 or NULL */
  int32_t act;             /* 0 none, 1 tanh-GELU (ff.py:12, jax.nn.gelu approximate=True) */
  int32_t accumulate;      /* d_fp32 only: D += result (gradient accumulation) */
  void* aux_out;           /* bf16 [M,N] ld=ldd: pre-activation, or NULL */
  const void* aux_in;      /* bf16 [M,N] ld=ldd: multiply by gelu'(aux_in), or NULL */
  const float* col_scale;  /* [N] LayerScale (layerscale.py:20-23) or NULL */
  const float* row_scale;  /* [ceil(M/rows_per_scale)] drop-path (stochastic_depth.py:16-27) */
  int64_t rows_per_scale;
  const void* residual;    /* fp32 [M,N] ld=ldr or NULL (vit.py:24,31); needs d_fp32 */
  int64_t ldr;
  void* workspace;         /* unused (kept for ABI stability); may be NULL */
  int64_t workspace_bytes;
  float* dgelu_colsum;     /* aux_in mode only: fp32 [N], += column sums of acc * gelu'(aux_in) -- the
                            * bias gradient of the Dense that produced aux_in (reverse of ff.py:24),
                            * accumulated by the epilogue instead of a separate pass; or NULL */
  /* Fused LayerNorm (vit.py:19,26 applied to the stream this GEMM produces; fp32 D + residual mode,
   * no scales, sav_gemm_ln_tiles(N) != 0): besides D the epilogue writes ln_out = bf16(LayerNorm(D))
   * with flax statistics (var = E[x^2] - E[x]^2, eps) and saves mean / rstd fp32 [M] for the
   * reverse mode -- the rows never make a second trip through HBM.  ln_out NULL = off. */
  void* ln_out;            /* bf16 [M,N] ld=ldy */
  int64_t ldy;
  const float* ln_gamma;   /* [N] */
  const float* ln_beta;    /* [N] */
  float* ln_mean;          /* [M] */
  float* ln_rstd;          /* [M] */
  float ln_eps;
  int32_t reserved;
} sav_gemm_args;
/* 1 or 2 when a LayerNorm over N columns can be fused into sav_gemm's fp32-stream epilogue (the row
 * fits the two TMEM accumulators at the tile width chosen for N: N = 128, 192, 256, 384, 512), else 0 */
int32_t sav_gemm_ln_tiles(int64_t N);
int64_t sav_gemm_workspace_bytes(const sav_gemm_args* a);
int sav_gemm(const sav_gemm_args* a, void* stream);

/* ---------------------------------------------------------------- batched GEMM (same kernel)
 * D[o,i][M,N] = sum_k A[o,i][m,k] * B[o,i][n,k] for every (outer o, inner i) batch; operand
 * element (o, i, row, col) at ptr + o*stride_outer + i*stride_inner + row*ld + col (elements; all
 * byte strides multiples of 16).  a_mn / b_mn as in sav_gemm.  D bf16 or fp32, no epilogue.
 * Out-of-range rows / columns / K of a batch are zero-filled on load and clipped on store, so one
 * batch never reads or writes its neighbour (M, N, K need not be tile multiples).
 * Replaces the einsums of the materialised-logits attention path: attention.py:41-42,57-58 when
 * talking heads (attention.py:44-52) sit between them, cait.py:10-15 (class attention, Nq = 1),
 * botnet.py:187-197 (+ relative logits) and any head dim the fused kernel does not cover. */
typedef struct {
  const void* A;
  const void* B;
  void* D;
  int64_t M, N, K;
  int64_t batch_outer, batch_inner;
  int64_t lda, a_stride_inner, a_stride_outer;
  int64_t ldb, b_stride_inner, b_stride_outer;
  int64_t ldd, d_stride_inner, d_stride_outer;
  int32_t a_mn, b_mn, d_fp32, reserved;
} sav_bgemm_args;
int sav_bgemm(const sav_bgemm_args* a, void* stream);

/* column sums: out[n] (+)= sum_m X[m,n]   (bias gradients; ff.py:24,27 reverse-mode) */
int sav_colsum_bf16(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                    int32_t accumulate, void* stream);
/* Same result without same-address atomics: row-split partial sums go to `workspace` (device,
 * >= sav_colsum_workspace_bytes(M, N) bytes, 16 B aligned, contents don't matter), a second small
 * kernel folds them into out.  Deterministic summation order. */
int64_t sav_colsum_workspace_bytes(int64_t M, int64_t N);
int sav_colsum_bf16_ws(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                       int32_t accumulate, void* workspace, int64_t workspace_bytes, void* stream);
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  sav_colsum_bf16_ws with its scratch in a file-scope `static __device__` array of
 * sav_colsum_static_max_bytes() = 4 MiB inside the library (every GPU that loads the library holds
 * it until exit); the call allocates nothing and takes no workspace.  Shapes whose
 * sav_colsum_workspace_bytes(M, N) exceeds the array are rejected with SAV_ERR_SHAPE (every
 * bias-gradient shape of the models here needs <= 1.9 MiB).  All calls on a device share the
 * array: calls that overlap on the GPU (different streams, parallel graph branches) race -- keep
 * them on one stream or order them with events.  The array's previous contents never matter. */
int64_t sav_colsum_static_max_bytes(void);
int sav_colsum_bf16_static(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                           int32_t accumulate, void* stream);

/* ---------------------------------------------------------------- LayerNorm (vit.py:19,26,57)
 * Flax nn.LayerNorm: fp32 stats, var = E[x^2]-E[x]^2, eps inside rsqrt, y = xhat*scale + bias.
 * x is the fp32 residual stream, row r at x + r*ldx (ldx lets the final LayerNorm run on the
 * class-token rows only, vit.py:57,95); y bf16, row r at y + r*ldy; scale,bias fp32 [C];
 * mean,rstd fp32 [rows] (saved for backward).  C % 8 == 0, C <= 1024.
 */
int sav_layernorm_fwd(const float* x, int64_t ldx, const float* scale, const float* bias, void* y,
                      int64_t ldy, float* mean, float* rstd, int64_t rows, int64_t C, float eps,
                      void* stream);
/* dx = LNbwd(dy) (+ dx_add if non-NULL: the residual-branch gradient, vit.py:24,31), written as the
 * fp32 stream gradient `dx` and, if dx_bf16 != NULL, also as a bf16 copy (the MMA operand of the
 * preceding block's dX / dW GEMMs).  dx_add, dx, dx_bf16 share the row pitch ldx with x; dy (bf16)
 * has row pitch ldy.  dscale/dbias [C] fp32 are ACCUMULATED (red.add); the caller zeroes them. */
int sav_layernorm_bwd(const void* dy, int64_t ldy, const float* x, int64_t ldx, const float* scale,
                      const float* mean, const float* rstd, const float* dx_add, float* dx,
                      void* dx_bf16, float* dscale, float* dbias, int64_t rows, int64_t C,
                      void* stream);
/* Same, and additionally dx_bf16_colsum[c] += sum_rows dx_bf16[row, c] when non-NULL: dx_bf16 is the
 * dY of the Linear that wrote this residual stream (out-proj / FC2, vit.py:24,31), so this is that
 * Linear's bias gradient (attention.py:60-63, ff.py:27 reverse mode) without a pass over dY. */
int sav_layernorm_bwd_colsum(const void* dy, int64_t ldy, const float* x, int64_t ldx,
                             const float* scale, const float* mean, const float* rstd,
                             const float* dx_add, float* dx, void* dx_bf16, float* dscale,
                             float* dbias, float* dx_bf16_colsum, int64_t rows, int64_t C,
                             void* stream);

/* ---------------------------------------------------------------- patch embedding stem
 * train.py:80-81 + patch_embed.py:19-22: images fp32 [H,W,C,N] (reference batch layout) ->
 * patches bf16 [N*(H/ph)*(W/pw), ph*pw*C] with feature order (ph, pw, c).  hwcn=0: input is
 * [N,H,W,C] instead. */
int sav_patchify(const float* images, void* patches, int64_t N, int64_t H, int64_t W, int64_t C,
                 int64_t ph, int64_t pw, int32_t hwcn, void* stream);
/* vit.py:81-85 + position_embed.py:54-56: x[b,0,:] = cls + pos[0]; x[b,1+i,:] = tok[b,i,:] +
 * pos[1+i].  has_cls=0: x[b,i,:] = tok[b,i,:] + pos[i] (cait.py:69).  tok bf16; x fp32 (the
 * residual stream); cls,pos fp32. */
int sav_embed_fwd(const void* tok, const float* cls, const float* pos, void* x, int64_t B,
                  int64_t n_tok, int64_t C, int32_t has_cls, void* stream);
/* reverse-mode of the above: dx fp32 -> dtok (bf16), dcls [C], dpos [n_tok+has_cls, C] fp32 */
int sav_embed_bwd(const void* dx, void* dtok, float* dcls, float* dpos, int64_t B, int64_t n_tok,
                  int64_t C, int32_t has_cls, int32_t accumulate, void* stream);

/* ---------------------------------------------------------------- fused attention core
 * attention.py:39-58: q/sqrt(d); S = q k^T; softmax over keys; O = P v.   One pass, S never
 * leaves the SM (TMEM).  q,k,v,o are bf16 with element (b, token, h, :) at
 *   ptr + b*batch_stride + token*row_stride + h*D           (strides in elements)
 * so they can alias slices of a fused [B,N,3,H,D] QKV buffer.  lse fp32 [B,H,Nq] (natural log of
 * the softmax denominator incl. max) is saved for backward.
 * D = 64 (D=48/128 and BoTNet relative logits: see DESIGN.md status table).
 */
typedef struct {
  const void* q;
  const void* k;
  const void* v;
  void* o;
  void* o_lo;  /* optional (training): bf16(o_fp32 - bf16(o_fp32)), same strides as o; NULL to skip */
  float* lse;
  int64_t B, H, Nq, Nk, D;
  int64_t q_batch_stride, q_row_stride;
  int64_t k_batch_stride, k_row_stride;
  int64_t v_batch_stride, v_row_stride;
  int64_t o_batch_stride, o_row_stride;
  float scale; /* 1/sqrt(D) (attention.py:39) */
} sav_attn_fwd_args;
int sav_attn_fwd(const sav_attn_fwd_args* a, void* stream);

typedef struct {
  const void* q;
  const void* k;
  const void* v;
  const void* o;
  const void* o_lo; /* optional low half of o saved by sav_attn_fwd (sharper delta), or NULL */
  const void* d_o;
  const float* lse;
  void* dq;
  void* dk;
  void* dv;
  float* delta;      /* workspace fp32 [B,H,Nq] */
  float* dq_accum;   /* workspace, only for Nq > 256 (else may be NULL): fp32 [2][B,Nk,H,D], zeroed by the callee --
                        the dK / dV accumulators of the query groups (TMA reduce-add), converted to bf16 at the end */
  int64_t B, H, Nq, Nk, D;
  int64_t q_batch_stride, q_row_stride; /* shared by q and dq */
  int64_t k_batch_stride, k_row_stride; /* shared by k and dk */
  int64_t v_batch_stride, v_row_stride; /* shared by v and dv */
  int64_t o_batch_stride, o_row_stride; /* shared by o and d_o */
  float scale;
} sav_attn_bwd_args;
int sav_attn_bwd(const sav_attn_bwd_args* a, void* stream);

/* ---------------------------------------------------------------- materialised-logits attention rows
 * Between S = q k^T and O = P v (both sav_bgemm) for the attention variants that couple heads or
 * add position logits.  S fp32 [B,H,Nq,ld] holds the RAW q.k products (ld >= Nk, ld % 8 == 0);
 *   logits[b,i,q,k] = scale * ( sum_h mix_pre[h,i] * S[b,h,q,k]      (talking_heads.py:13, pre)
 *                               + rel[b,q,i, kx-qx+W-1] + rel[b,q,i, (2W-1)+ky-qy+Hh-1] )
 *                                                       (botnet.py:78-141; q = qy*W+qx, k = ky*W+kx)
 *   p = softmax_k(logits)                                             (attention.py:48)
 *   P[b,j,q,k] = sum_i mix_post[i,j] * p[b,i,q,k]   -> bf16 [B,H,Nq,ld], columns >= Nk zeroed
 * mix_pre/mix_post [H,H] fp32 (both or neither; NULL = identity; H in {2,3,4,6,8,12,16} when given);
 * rel fp32 [B,Nq,H,ldr] = unscaled q.r^W (first 2W-1 columns) | q.r^H (next 2Hh-1), or NULL.
 * Backward takes dP = dO v^T (fp32, same layout), recomputes p from S, and writes dS (bf16, the
 * gradient of the raw q.k products, scale folded in), drel (fp32 [B,Nq,H,ldr], overwritten) and
 * ACCUMULATES dmix_pre / dmix_post [H,H]. */
typedef struct {
  const float* S;
  void* P;             /* fwd out */
  const float* dP;     /* bwd in */
  void* dS;            /* bwd out */
  const float* mix_pre;
  const float* mix_post;
  float* dmix_pre;     /* bwd, accumulated */
  float* dmix_post;
  const float* rel;
  float* drel;         /* bwd out */
  int64_t B, H, Nq, Nk, ld, ldr;
  int32_t rel_w, rel_h;
  float scale;
  int32_t reserved;
} sav_thsoftmax_args;
int sav_thsoftmax_fwd(const sav_thsoftmax_args* a, void* stream);
int sav_thsoftmax_bwd(const sav_thsoftmax_args* a, void* stream);

/* ---------------------------------------------------------------- CvT convolutional projection
 * cvt_attention.py:12-40: depthwise 3x3 conv (stride 1|2, padding SAME, no bias) -> BatchNorm
 * (batch statistics) -> 1x1 conv (a sav_gemm).  Channel-last tensors, C % 4 == 0.
 *   x fp32 [B,H,W,C] (the fp32 residual stream -- CvT feeds the attention without a LayerNorm,
 *   cvt.py:50-61); w fp32 [3,3,C] (flax depthwise kernel [3,3,1,C]); y bf16 [B,Ho,Wo,C] with
 *   Ho = ceil(H/stride) (sav_dwconv3x3_out_hw).  sum / sumsq fp32 [C] receive the per-channel sum
 *   and sum of squares of the conv output (zeroed by the callee) = the BatchNorm statistics. */
int sav_dwconv3x3_out_hw(int64_t H, int64_t W, int32_t stride, int64_t* Ho, int64_t* Wo);
int sav_dwconv3x3_fwd(const float* x, const float* w, void* y, float* sum, float* sumsq, int64_t B,
                      int64_t H, int64_t W, int64_t C, int32_t stride, void* stream);
/* dx fp32 [B,H,W,C] (+= if accumulate_dx: q, k and v projections read the same input);
 * dw fp32 [3,3,C] ACCUMULATED. */
int sav_dwconv3x3_bwd(const float* x, const float* w, const void* dy, float* dx, float* dw, int64_t B,
                      int64_t H, int64_t W, int64_t C, int32_t stride, int32_t accumulate_dx,
                      void* stream);
/* flax nn.BatchNorm in training mode over rows = B*Ho*Wo: mean = sum/rows, var = sumsq/rows -
 * mean^2 (biased), z = (y - mean) rsqrt(var + eps) gamma + beta; writes mean, rstd (saved for
 * backward) and var (NULL to skip; for the caller's running averages).  y, z bf16 [rows, C]. */
int sav_batchnorm_fwd(const void* y, const float* sum, const float* sumsq, const float* gamma,
                      const float* beta, float eps, void* z, float* mean, float* rstd, float* var,
                      int64_t rows, int64_t C, void* stream);
/* dy = gamma rstd (dz - mean_rows(dz) - yhat mean_rows(dz yhat)); dgamma, dbeta ACCUMULATED;
 * scratch fp32 [2C]. */
int sav_batchnorm_bwd(const void* dz, const void* y, const float* mean, const float* rstd,
                      const float* gamma, float* dgamma, float* dbeta, float* scratch, void* dy,
                      int64_t rows, int64_t C, void* stream);

/* LayerNorm INTO the fp32 residual stream (the LayerNorm that closes the CvT conv token embedding,
 * cvt.py:34): fp32 in, fp32 out; reverse mode with an fp32 upstream gradient; otherwise as
 * sav_layernorm_fwd / sav_layernorm_bwd. */
int sav_layernorm_fwd_f32(const float* x, int64_t ldx, const float* scale, const float* bias, float* y,
                          int64_t ldy, float* mean, float* rstd, int64_t rows, int64_t C, float eps,
                          void* stream);
int sav_layernorm_bwd_f32(const float* dy, int64_t ldy, const float* x, int64_t ldx, const float* scale,
                          const float* mean, const float* rstd, const float* dx_add, float* dx,
                          void* dx_bf16, float* dscale, float* dbias, int64_t rows, int64_t C,
                          void* stream);
/* dst fp32 [rows, cols] ldd (+)= src bf16 [rows, cols] lds */
int sav_cast2d_bf16_f32(const void* src, int64_t lds, float* dst, int64_t ldd, int64_t rows,
                        int64_t cols, int32_t accumulate, void* stream);
/* Batched transpose with zero padding -- rearrange '... l d -> ... d l' and back around the
 * token-mixing FFBlock of MLP-Mixer (models/mlp_mixer.py:18-23).  Per batch b (in: R rows of pitch
 * ld_in, out: C rows of pitch ld_out, residual: C rows of pitch ld_res):
 *   out[b, c, r] = in[b, r, c] (+ residual[b, c, r])   r < R, c < C;   out[b, c, R..out_cols-1] = 0.
 * in / out are bf16 unless in_fp32 / out_fp32; residual (optional) is fp32. */
int sav_transpose_pad(const void* in, int32_t in_fp32, int64_t ld_in, void* out, int32_t out_fp32,
                      int64_t ld_out, const float* residual, int64_t ld_res, int64_t B, int64_t R,
                      int64_t C, int64_t out_cols, void* stream);

/* ---------------------------------------------------------------- dense convolutions (im2col + GEMM)
 * nn.Conv of the CvT token embedding (cvt.py:27-31: 7x7 / 3x3, stride 4 / 2, SAME, bias) and of the
 * BoTNet ResNet stem / bottleneck convs (botnet.py:47-64,291-305): the patch matrix is built by
 * sav_im2col, the contraction is sav_gemm (forward [M,K]x[K,Cout]; dW = cols^T dY; dcols = dY W^T),
 * and sav_col2im folds dcols back onto the input (gather form, no atomics).
 * x is addressed through element strides, so both [B,H,W,C] and the loader's [H,W,C,N] work. */
typedef struct sav_conv2d_geom {
  int64_t B, H, W, C;            /* input */
  int64_t Ho, Wo;                /* output spatial size */
  int32_t kh, kw, sh, sw;        /* kernel, strides */
  int32_t pad_top, pad_left;     /* low-side zero padding (high side implied by Ho / Wo) */
  int64_t xs_b, xs_h, xs_w, xs_c; /* element strides of x */
  int64_t ld;                    /* row pitch of the patch matrix: >= kh*kw*C, multiple of 8 */
} sav_conv2d_geom;
/* flax / XLA padding='SAME': out = ceil(in / stride), total pad = max((out-1)*stride + k - in, 0),
 * low side = total / 2. */
int sav_conv_same_geom(int64_t in, int32_t k, int32_t stride, int64_t* out, int32_t* pad_lo);
/* cols bf16 [B*Ho*Wo, ld], feature order (kh, kw, cin); columns >= kh*kw*C are zero */
int sav_im2col(const void* x, int32_t x_is_fp32, void* cols, const sav_conv2d_geom* g, void* stream);
/* dx [B,H,W,C] contiguous (fp32 or bf16) (+)= fold(dcols) */
int sav_col2im(const void* dcols, void* dx, int32_t dx_is_fp32, int32_t accumulate,
               const sav_conv2d_geom* g, void* stream);
/* BatchNorm pieces beyond sav_batchnorm_fwd/bwd: batch sums of a bf16 [rows,C] tensor (callee
 * zero-fills sum / sumsq), the running-average update ra = momentum*ra + (1-momentum)*stat
 * (flax nn.BatchNorm, cvt_attention.py:33-36), and the eval-mode apply from running averages. */
int sav_bn_stats(const void* y, int64_t rows, int64_t C, float* sum, float* sumsq, void* stream);
/* mean, rstd = rsqrt(var + eps), biased var (flax nn.BatchNorm) from the sums over `rows` rows */
int sav_bn_finalize(const float* sum, const float* sumsq, int64_t rows, float eps, float* mean,
                    float* rstd, float* var, int64_t C, void* stream);
int sav_bn_update_running(const float* mean, const float* var, float* ra_mean, float* ra_var,
                          float momentum, int64_t C, void* stream);
int sav_batchnorm_eval(const void* y, const float* ra_mean, const float* ra_var, const float* gamma,
                       const float* beta, float eps, void* z, int64_t rows, int64_t C, void* stream);

/* ---------------------------------------------------------------- BoTNet conv-stage pieces
 * Channel-last bf16 [rows, C] tensors around the conv GEMMs of botnet.py:17-67,202-331.
 * kind: 0 none, 1 swish (nn.swish, the BoTNet activation_fn), 2 sigmoid (SE gate), 3 relu,
 * 4 tanh-GELU (CeiT's LeFF block, leff.py:43,55,61).
 *   sav_bn_act_fwd / bwd : z = act(BatchNorm(y)) from given mean / rstd (sav_bn_stats +
 *       sav_batchnorm_fwd produce them); the reverse mode recomputes act' from y, accumulates
 *       dgamma / dbeta, scratch fp32 [2C]                          (botnet.py:44-50,225-228)
 *   sav_gate_res_act_fwd / bwd : out = act(residual + x * gate[b, c]) -- squeeze-excite gate, skip
 *       connection and activation of a bottleneck block in one pass; gate bf16 [B, C] and residual
 *       optional; reverse mode writes dx, dresidual (optional) and dgate fp32 [B, C] (zero-filled by
 *       the callee)                                                 (botnet.py:54-67,244-252)
 *   sav_spatial_mean_fwd / bwd : [B, rows_per_sample, C] <-> [B, C]  (jnp.mean(.., (1, 2)):
 *       squeeze_excite.py:32, botnet.py:322)
 *   sav_maxpool3x3s2_fwd / bwd : nn.max_pool((3,3), (2,2), 'SAME') (botnet.py:297), arg-max uint8
 *   sav_act_fwd / bwd : elementwise activation on n bf16 elements (SE hidden swish, SE sigmoid) */
int sav_act_fwd(const void* x, void* y, int64_t n, int32_t kind, void* stream);
int sav_act_bwd(const void* x, const void* dy, void* dx, int64_t n, int32_t kind, void* stream);
int sav_bn_act_fwd(const void* y, const float* mean, const float* rstd, const float* gamma,
                   const float* beta, void* z, int64_t rows, int64_t C, int32_t kind, void* stream);
int sav_bn_act_bwd(const void* dz, const void* y, const float* mean, const float* rstd,
                   const float* gamma, const float* beta, float* dgamma, float* dbeta, float* scratch,
                   void* dy, int64_t rows, int64_t C, int32_t kind, void* stream);
int sav_gate_res_act_fwd(const void* x, const void* gate, const void* residual, void* out, int64_t B,
                         int64_t rows_per_sample, int64_t C, int32_t kind, void* stream);
int sav_gate_res_act_bwd(const void* dout, const void* x, const void* gate, const void* residual,
                         void* dx, void* dresidual, float* dgate, int64_t B, int64_t rows_per_sample,
                         int64_t C, int32_t kind, void* stream);
int sav_spatial_mean_fwd(const void* x, void* out, int64_t B, int64_t rows_per_sample, int64_t C,
                         void* stream);
int sav_spatial_mean_bwd(const void* dout, void* dx, int64_t B, int64_t rows_per_sample, int64_t C,
                         void* stream);
int sav_maxpool3x3s2_fwd(const void* x, void* y, void* argmax, int64_t B, int64_t H, int64_t W,
                         int64_t C, void* stream);
int sav_maxpool3x3s2_bwd(const void* dy, const void* argmax, void* dx, int64_t B, int64_t H, int64_t W,
                         int64_t C, void* stream);

/* ---------------------------------------------------------------- rotary helpers
 * models/layers/position_embed.py:8-45 (unused by every reference model and unfinished there:
 * parity unpinned; the intended RoFormer rotation with the reference's literal float32 frequency
 * table inv_freq[i] = 1 / (10e4^(2i) / dim), position_embed.py:29).
 * sav_rotary: x, y [outer, n, inner, d] contiguous (bf16, or fp32 when is_fp32), rotation of the
 * (2i, 2i+1) pairs of the last axis by t * inv_freq[i], t = index on the n axis
 * (apply_rotary_pos_emb + rotate_every_two, position_embed.py:8-20); inverse != 0 rotates back
 * (the reverse mode).  sav_rotary_table: FixedPositionalEmbedding (position_embed.py:23-34),
 * emb fp32 [n, dim] = concat(freqs, freqs). */
int sav_rotary(const void* x, void* y, int64_t outer, int64_t n, int64_t inner, int64_t d,
               int32_t is_fp32, int32_t inverse, void* stream);
int sav_rotary_table(float* emb, int64_t n, int64_t dim, void* stream);

/* ---------------------------------------------------------------- loss head (train.py:83-91)
 * one_hot -> smooth_labels(alpha) -> softmax_cross_entropy -> mean; writes dlogits = dloss/dlogits
 * (bf16 [B,classes]) and accumulates loss into loss_out[0] (zeroed by the callee), top1 count into
 * loss_out[1].  logits bf16 [B, classes] ld. loss_scale multiplies the gradient (1/B/world). */
int sav_softmax_xent(const void* logits, int64_t ld, const int32_t* labels, int64_t B,
                     int64_t classes, float label_smoothing, float grad_scale, void* dlogits,
                     float* loss_out, void* stream);

/* Same with the mixup / cutmix branch of train.py:84-87: y = ratio[b]*one_hot(labels[b]) +
 * (1-ratio[b])*one_hot(mix_labels[b]) before smooth_labels.  mix_labels (int32 [B]) and ratio
 * (fp32 [B]) are both NULL (plain labels) or both given. */
int sav_softmax_xent_mix(const void* logits, int64_t ld, const int32_t* labels,
                         const int32_t* mix_labels, const float* ratio, int64_t B, int64_t classes,
                         float label_smoothing, float grad_scale, void* dlogits, float* loss_out,
                         void* stream);

/* utils.py:20-31 topk_correct (train.py:98-99): out[i*B + b] = 1.0 if the label of example b is
 * among the k_i highest logits (argsort tie order: larger class index ranks higher), else 0.0;
 * 1 <= nk <= 4.  logits bf16 [B, classes] ld; out fp32 [nk, B]. */
int sav_topk_correct(const void* logits, int64_t ld, const int32_t* labels, int64_t B,
                     int64_t classes, int32_t nk, int32_t k0, int32_t k1, int32_t k2, int32_t k3,
                     float* out, void* stream);

/* ---------------------------------------------------------------- parameters / optimizer
 * fp32 -> bf16 shadow copy of the flat parameter arena */
int sav_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);
/* out[0] += sum(g^2) over n elements (global-norm clip, train.py:25, optax.clip_by_global_norm) */
int sav_sumsq(const float* g, int64_t n, float* out, void* stream);
/* optax chain of train.py:25-27 / simple_train.py:25-27 as one pass over the arena:
 * g *= min(1, max_norm/||g||) (sumsq[0] holds ||g||^2; max_norm<=0 disables) -> Adam(b1,b2,eps,
 * bias-corrected with step t) -> + wd*p -> * (-lr) -> p += update; also refreshes bf16 shadow.
 * hyper (optional, device fp32[3] = {lr, 1-b1^t, 1-b2^t}) overrides lr/t so that the step can be
 * replayed from a CUDA graph while the schedule advances.  b1 / b2 are doubles so that
 * (1 - decay) is formed as optax forms it: in double from the Python float, then rounded. */
int sav_adam_step(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n,
                  const float* sumsq, float max_norm, float lr, double b1, double b2, float eps,
                  float wd, int64_t t, const float* hyper, void* stream);
/* optax.scale_by_schedule + the Adam bias corrections as a device-resident counter (train.py:25-27,
 * 214-220): step[0] (device int64, the optax `count`) is read, incremented, and hyper (device
 * fp32[3]) is set to {lr(count), 1-b1^(count+1), 1-b2^(count+1)} for the sav_adam_step enqueued
 * after it, so a captured CUDA graph advances the schedule by itself on every replay.
 * mode 0: lr = peak_value (constant; simple_train.py:27);  mode 1:
 * optax.warmup_cosine_decay_schedule(init_value, peak_value, warmup_steps, decay_steps, end_value);
 * mode 2: hyper[0] is left as the host wrote it (arbitrary Python schedule). */
int sav_schedule_step(int64_t* step, float* hyper, int32_t mode, float init_value, float peak_value,
                      int64_t warmup_steps, int64_t decay_steps, float end_value, double b1, double b2,
                      void* stream);
/* stochastic_depth.py:16-27 mask source: out[b] = floor(keep + U[b]) / keep with U from
 * Philox4x32-10 keyed by seed, counter (b, call_index, step); the step comes from step_dev (device
 * int64, e.g. the counter of sav_schedule_step) when non-NULL so graph replays resample, else from
 * step_host.  out fp32 [B]. */
int sav_drop_path_factors(uint64_t seed, const int64_t* step_dev, int64_t step_host,
                          int64_t call_index, int64_t B, float keep, float* out, void* stream);

/* elementwise helpers (bf16): y = a + b ; used only where no GEMM epilogue can absorb the add */
int sav_add_bf16(const void* a, const void* b, void* y, int64_t n, void* stream);
int sav_fill_f32(float* p, float v, int64_t n, void* stream);
/* strided row copy dst[r, 0:cols] = src[r, 0:cols] (class-token gather/scatter, vit.py:95) */
int sav_copy2d_bf16(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t rows,
                    int64_t cols, void* stream);

/* reverse mode of the LayerScale / drop-path GEMM epilogue (layerscale.py:20-23,
 * stochastic_depth.py:16-27): out[m,n] = bf16(g[m,n] * col_scale[n] * row_scale[m/rows_per_scale])
 * from the fp32 stream gradient g; either scale may be NULL. */
int sav_scale_rows_cols(const float* g, int64_t ldg, const float* col_scale, const float* row_scale,
                        int64_t rows_per_scale, void* out, int64_t ldo, int64_t rows, int64_t cols,
                        void* stream);
/* LayerScale gradient from the finished weight / bias gradients of the scaled Dense (no saved
 * activation): dcol_scale[n] += (sum_k W[k,n] dW[k,n] + bias[n] dbias[n]) / col_scale[n]. */
int sav_layerscale_grad(const float* W, int64_t ldw, const float* dW, int64_t ldg, int64_t K,
                        int64_t N, const float* bias, const float* dbias, const float* col_scale,
                        float* dcol_scale, void* stream);
/* strided fp32 row copy / accumulate: dst[r, 0:cols] (+)= src[r, 0:cols] (class-token concat of
 * cait.py:97,175 and its reverse mode; fan-in of fp32 stream gradients) */
int sav_copy2d_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t rows,
                   int64_t cols, int32_t accumulate, void* stream);

/* ---------------------------------------------------------------- CeiT (models/ceit.py)
 * nn.max_pool(x, (3,3), strides (2,2)) with flax's default 'VALID' padding (image_to_token.py:33-37);
 * same buffers as sav_maxpool3x3s2_*, Ho = (H - 3) / 2 + 1. */
int sav_maxpool3x3s2_valid_fwd(const void* x, void* y, void* argmax, int64_t B, int64_t H, int64_t W,
                               int64_t C, void* stream);
int sav_maxpool3x3s2_valid_bwd(const void* dy, const void* argmax, void* dx, int64_t B, int64_t H,
                               int64_t W, int64_t C, void* stream);
/* 'b (h ph) (w pw) c -> b (h w) (ph pw c)' of a bf16 activation [B,H,W,C] (image_to_token.py:38-43);
 * inverse != 0: the permutation back (src [B, tokens, ph*pw*C] -> dst [B,H,W,C]), i.e. its gradient. */
int sav_patchify_act(const void* src, void* dst, int64_t B, int64_t H, int64_t W, int64_t C, int64_t ph,
                     int64_t pw, int32_t inverse, void* stream);

/* ---------------------------------------------------------------- TNT (models/tnt.py)
 * Pixel-embedding input (tnt.py:19-26): fp32 images ([N,H,W,C], or [H,W,C,N] if hwcn) -> bf16
 * [N*(H/ph)*(W/pw), (ph/th)*(pw/tw), C*th*tw], feature order (c, t1, t2). */
int sav_pixel_patchify(const float* images, void* out, int64_t N, int64_t H, int64_t W, int64_t C,
                       int64_t ph, int64_t pw, int64_t th, int64_t tw, int32_t hwcn, void* stream);
/* attention.py:39-58 for short sequences with narrow heads (TNT's inner transformer, tnt.py:60-69:
 * 16 tokens, head dim 6 / 10): qkv bf16 contiguous [B, N, 3, H, D] (the fused projection output),
 * o / d_o bf16 [B, N, H, D], dqkv like qkv; N <= 32, D <= 16; fp32 CUDA-core math, P recomputed in
 * the backward pass (nothing saved). */
int sav_attn_small_fwd(const void* qkv, void* o, int64_t B, int64_t N, int64_t H, int64_t D, float scale,
                       void* stream);
int sav_attn_small_bwd(const void* qkv, const void* d_o, void* dqkv, int64_t B, int64_t N, int64_t H,
                       int64_t D, float scale, void* stream);

/* ---------------------------------------------------------------- data-parallel gradient exchange
 * lax.pmean(grads) of train.py:96 as one kernel over NVLink / NVSwitch peer memory (csrc/dp.cu):
 * in-place SUM all-reduce of buf[offset, offset + n) where `buf` is the same symmetric allocation on
 * every rank.  The caller (sav/dp.py through torch's symmetric-memory rendezvous, or any cuMem /
 * cuMulticast owner) provides
 *   multicast_ptr        multicast address of the allocation (NVLS: multimem.ld_reduce / multimem.st),
 *                        or 0 -> peer loads / stores through peer_ptrs_dev instead
 *   peer_ptrs_dev        device array of `world` pointers to every rank's copy (own included)
 *   signal_pad_ptrs_dev  device array of `world` pointers to every rank's zero-initialised signal pad
 *                        of signal_pad_bytes; the kernel uses blocks * world 32-bit words of it and
 *                        leaves them zero
 * offset, n in floats, multiples of 4.  Every rank must enqueue the call with identical offset, n,
 * world and blocks (0 = default); the kernel waits for its peers on the device, never on the host,
 * and can be captured in a CUDA graph.  Replaces ncclAllReduce on the path. */
int sav_dp_allreduce_sum_f32(uint64_t multicast_ptr, const uint64_t* peer_ptrs_dev,
                             const uint64_t* signal_pad_ptrs_dev, int64_t signal_pad_bytes,
                             int64_t offset, int64_t n, int32_t rank, int32_t world, int32_t blocks,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SAV_ABI_H_ */
