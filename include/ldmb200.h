/*
 * ldmb200 — C-ABI of the B200-native latent-diffusion denoising path (sm_100a).
 *
 * This is the drop-in boundary.  The reference (GiannisPikoulis/dsml-thesis) is pure
 * Python/PyTorch and has no FFI of its own, so each entry point below replaces the ATen call sites of one
 * kernel family on the hot path; the reference file:line each one stands for is cited per function
 * (paths relative to face_reenactment/; the UNet files are byte-identical under talking_face/).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no torch / C++ types cross the boundary.
 *   - Every pointer is a DEVICE pointer owned by the caller; nothing is allocated, freed or retained.
 *   - All work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream);
 *     calls are stateless, re-entrant and CUDA-Graph capturable.
 *   - Return value: 0 on success, negative LDMB200_ERR_* otherwise; ldmb200_last_error() returns a
 *     thread-local message.  Nothing throws.
 *   - Activations are channels-last: a feature map is a row-major matrix [B*H*W, C] ("pixels x channels").
 *   - dtype codes: LDMB200_F32 = 0, LDMB200_BF16 = 1, LDMB200_F16 = 2 (IEEE half; conversions saturate).
 */
#ifndef LDMB200_H_
#define LDMB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LDMB200_ABI_VERSION 5

#define LDMB200_F32 0
#define LDMB200_BF16 1
#define LDMB200_F16 2

#define LDMB200_ERR_INVALID (-1) /* bad argument / unsupported shape */
#define LDMB200_ERR_CUDA (-2)    /* CUDA runtime / driver error     */

int ldmb200_abi_version(void);
const char* ldmb200_last_error(void);
/* number of SMs of the current device (grid sizing); negative on error */
int ldmb200_sm_count(void);

/* ---------------------------------------------------------------------------------------------------
 * K13  DDIM x_{t-1} update — DDIMSampler.p_sample_ddim, ldm/models/diffusion/ddim.py:173-203
 * (talking_face: ldm/models/diffusion/ddim2cond.py:167-194).  One fused pass:
 *   e      = e_uncond ? e_uncond + cfg_scale*(e_cond - e_uncond) : e_cond                 (ddim.py:177)
 *   pred_x0 = (x - sqrt_one_minus_at*e) / sqrtf(a_t)                                        (ddim.py:194)
 *   x_prev  = sqrtf(a_prev)*pred_x0 + sqrtf(1 - a_prev - sigma_t^2)*e + sigma_t*noise*temperature
 * Coefficients are the fp32-rounded per-step scalars of torch.full (ddim.py:188-191); the three square
 * roots are evaluated in fp32 on the device like the reference.  noise may be NULL (treated as 0).
 * n = number of elements.  x_prev may alias x.
 */
int ldmb200_ddim_step(const float* x, const float* e_cond, const float* e_uncond, const float* noise,
                      float* x_prev, float* pred_x0, int64_t n, float a_t, float a_prev, float sigma_t,
                      float sqrt_one_minus_at, float cfg_scale, float temperature, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * K12  timestep_embedding — ldm/modules/diffusionmodules/util.py:151-171:  out[b] = [cos(t_b f) | sin(t_b f)],
 * f_i = exp(-ln(max_period) * i / (dim/2)), fp32.  t is int64 [B]; out is fp32 [B, dim] (dim even).
 */
int ldmb200_timestep_embedding(const int64_t* t, float* out, int batch, int dim, float max_period, void* stream);

/* Small dense layer for per-sample vectors (time_embed MLP openaimodel.py:506-511,724; ResBlock emb_layers
 * 218-224,264; cross-attention to_k/to_v/to_out on a length-1 context attention.py:175-176,193):
 *   out[r, n] = bias[n] + sum_k act(in[r, k]) * W[n, k]      fp32 throughout, act = SiLU if silu_in else id.
 * in [R, K] (row stride in_ld), W [N, K] row-major (nn.Linear layout), out [R, N] (row stride out_ld).
 */
int ldmb200_small_linear(const float* in, int in_ld, const float* w, const float* bias, float* out, int out_ld,
                         int rows, int k, int n, int silu_in, void* stream);

/* The same layer for a whole table of independent (in, W, bias, out) problems in ONE launch: all 17 ResBlock
 * emb projections of a UNet eval depend only on emb (openaimodel.py:264), all 16 cross-attention value paths only
 * on the context (attention.py:175-176,193), so each group is one launch.  jobs is a DEVICE array; first_block is
 * the exclusive prefix sum of ldmb200_small_linear_blocks(rows, n) over the table, total_blocks its total. */
typedef struct ldmb200_linear_job {
  const float* in;
  const float* w;    /* [n, k] row-major (nn.Linear layout) */
  const float* bias; /* [n] or NULL */
  float* out;
  int32_t in_ld, out_ld;
  int32_t rows, k, n;
  int32_t silu_in;
  int32_t first_block;
  int32_t pad_;
} ldmb200_linear_job;
int ldmb200_small_linear_blocks(int rows, int n);
int ldmb200_small_linear_batched(const ldmb200_linear_job* jobs, int n_jobs, int total_blocks, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Layout changes at the UNet boundary (openaimodel.py:730,738-742: NCHW fp32 in, NCHW fp32 out).
 *   nchw_to_nhwc: x fp32 [B, C, HW] -> y [B, HW, c_pad] (dtype out_dtype), channels >= C zero-filled.
 *   nhwc_to_nchw: y fp32 [B, HW, ld] (first C columns) -> x fp32 [B, C, HW].
 */
int ldmb200_nchw_to_nhwc(const float* x, void* y, int batch, int c, int hw, int c_pad, int out_dtype, void* stream);
int ldmb200_nhwc_to_nchw(const float* y, int ld, float* x, int batch, int c, int hw, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * K7 (+K11)  GroupNorm(32 groups) [+ SiLU] over a channels-last map that may be the virtual channel
 * concat of two maps — GroupNorm32 util.py:214-216 / normalization 199-205 (eps 1e-5) followed by nn.SiLU
 * in ResBlock.in_layers/out_layers (openaimodel.py:201-203,226-227) and UNetModel.out (683-684); Normalize
 * attention.py:76-77 (eps 1e-6, no SiLU); th.cat([h, hs.pop()], 1) openaimodel.py:736.
 * x1 [B, HW, c1], x2 [B, HW, c2] (x2 NULL / c2 = 0 for no concat), both fp32.  Statistics in fp32 with a
 * float64 combine.  y [B, HW, c1+c2] (out_dtype).  raw_copy (nullable) receives the un-normalised concat
 * cast to out_dtype (operand of ResBlock.skip_connection, openaimodel.py:241,275).
 * stats_ws: caller workspace of ldmb200_groupnorm_ws_floats(batch, hw, c) floats.
 */
int64_t ldmb200_groupnorm_ws_floats(int batch, int hw, int c);
int ldmb200_groupnorm(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta,
                      void* y, void* raw_copy, int out_dtype, int batch, int hw, int groups, float eps, int silu,
                      float* stats_ws, void* stream);

/* Allocation-free variant of ldmb200_groupnorm.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and results as ldmb200_groupnorm without stats_ws: the statistics workspace is a
 * file-scope device array inside the library (LDMB200_GN_STATIC_MAX_BATCH x 32 KiB = 2 MiB, held by every GPU that loads
 * the library until exit; its device address is resolved with cudaGetSymbolAddress once per device).  The call
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  batch > LDMB200_GN_STATIC_MAX_BATCH is
 * rejected with LDMB200_ERR_INVALID.  All calls on a GPU share the array: calls that overlap on the GPU (different
 * streams or host threads) RACE — serialise them on one stream.  Nothing needs to start at zero between calls. */
#define LDMB200_GN_STATIC_MAX_BATCH 64
int ldmb200_groupnorm_static(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta,
                             void* y, void* raw_copy, int out_dtype, int batch, int hw, int groups, float eps, int silu,
                             void* stream);

/* K7 with the statistics pass fused into the producers: colstats1 / colstats2 are the ldmb200_gemm_desc.colstats
 * buffers the tensor-core GEMMs that wrote x1 / x2 filled ([B*hw/32][c][2] floats); the statistics never re-read
 * x.  hw % 32 == 0.  Launches: finalize (float64 fold -> stats_ws[b][g] = {mean, rstd}) + apply. */
int ldmb200_groupnorm_fused(const float* x1, int c1, const float* colstats1, const float* x2, int c2,
                            const float* colstats2, const float* gamma, const float* beta, void* y, void* raw_copy,
                            int out_dtype, int batch, int hw, int groups, float eps, int silu, float* stats_ws,
                            void* stream);

/* K9  LayerNorm over the last dim — nn.LayerNorm(dim) attention.py:203-205 (eps 1e-5, affine).
 * x fp32 [M, C] -> y [M, C] (out_dtype). */
int ldmb200_layernorm(const float* x, const float* gamma, const float* beta, void* y, int out_dtype, int64_t rows,
                      int c, float eps, void* stream);

/* fp32 -> out_dtype helpers feeding the strided / upsampled convolutions on the tensor-core path:
 *   space_to_depth2: x fp32 [B, H, W, C] -> y [B, 4, H/2, W/2, C], plane p = (h&1)*2 + (w&1)
 *                    (input re-ordering for Downsample.op, conv 3x3 stride 2, openaimodel.py:151-160)
 *   upsample2x:      x fp32 [B, H, W, C] -> y [B, 2H, 2W, C] nearest  (F.interpolate, openaimodel.py:116)
 */
int ldmb200_space_to_depth2(const float* x, void* y, int out_dtype, int batch, int h, int w, int c, void* stream);
int ldmb200_upsample2x(const float* x, void* y, int out_dtype, int batch, int h, int w, int c, void* stream);
/* x fp32 [n] -> y [n] in out_dtype (n % 4 == 0): the 16-bit operand copy of a residual-stream map that a GEMM reads as it
 * is — the low-res input of the folded Upsample convolution (ldmb200_igemm_tc with upsample = 1). */
int ldmb200_cast_rows(const float* x, void* y, int out_dtype, int64_t n, void* stream);
/* fp16 operand range guard (UNetModel.f16_guard): *counter_u64 += number of the n fp16 values at x that sit at the clamp of
 * the saturating conversions (|v| >= 65504) or are not finite.  Replaces no reference call site: fp16 operands are this
 * implementation's choice, the counter is what keeps that choice from changing results silently. */
int ldmb200_count_f16_saturated(const void* x, int64_t n, void* counter_u64, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * K1-K4 (+K8,K10,K11)  Convolution / linear as (implicit) GEMM with fused epilogue.
 *   conv_nd(2,..,3,padding=1) openaimodel.py:204,230,519,685,107 ; Downsample.op stride 2 151-153 ;
 *   1x1 skip_connection 241, proj_in/proj_out attention.py:233-248 ; nn.Linear attention.py:40,60,161-166.
 *
 *   acc[m, n] = sum_{tap, c} A(m, tap, c) * W[n, tap, c]
 *   v         = acc + bias[n] + rowvec[sample(m), n] + residual[m, n]
 *   out[m, j] = geglu ? v[value col] * gelu_erf(v[gate col]) : v           (attention.py:43-44)
 *
 * A is read from one or two channels-last maps (virtual concat, c1 + c2 channels per tap) of spatial size
 * h_in x w_in; ksize 1 or 3 (padding 1); stride 1 or 2; upsample=1 = Upsample.forward (openaimodel.py:109-119):
 * F.interpolate(x, scale_factor=2, mode="nearest") followed by the 3x3 convolution, h_in x w_in being the LOW-RES map.
 * The SIMT path folds the upsample into its input index map; the tensor-core path runs it as FOUR 2x2 convolutions on
 * the low-res map, one per output parity, with tap-summed weights (w = the kind-4 pack of ldmb200_pack_weights:
 * [4][n][4 taps][c]) — 4/9 of the flops and no materialised 4x larger map.
 */
typedef struct ldmb200_gemm_desc {
  const void* a;         /* first map  [B, h_in, w_in, c1] (plain GEMM: [M, c1] with h_in = M/batch, w_in = 1) */
  const void* a2;        /* second map [B, h_in, w_in, c2] or NULL                                    */
  int32_t c1, c2;
  int32_t a_dtype;       /* LDMB200_F32 | LDMB200_BF16                                                */
  int32_t batch, h_in, w_in;
  int32_t ksize;         /* 1 | 3                                                                     */
  int32_t stride;        /* 1 | 2  (tensor-core path, stride 2: `a` is the space_to_depth2 layout)    */
  int32_t upsample;      /* 0 | 1                                                                     */
  const void* w;         /* SIMT: fp32 native nn layout [n, c1+c2, k, k] ; TC: packed bf16 (ldmb200_pack_weights) */
  int32_t n;             /* GEMM N = rows of w (for geglu: 2 x output width)                          */
  const float* bias;     /* [n] or NULL (TC + geglu: packed order)                                    */
  const float* rowvec;   /* [batch, rowvec_ld] per-sample broadcast add or NULL (K8: emb_out; K6: M=1 cross-attn) */
  int32_t rowvec_ld;
  const float* residual; /* fp32 [M, residual_ld] or NULL                                             */
  int32_t residual_ld;
  int32_t geglu;         /* 0 | 1                                                                     */
  void* out;             /* [M, out_ld]                                                               */
  int32_t out_dtype;
  int32_t out_ld;
  int32_t out_nchw;      /* 1: out is fp32 NCHW [B, n_valid, H_out*W_out]  (UNetModel.out conv)       */
  int32_t n_valid;       /* columns actually stored (<= n; TC path pads n to a multiple of 16)        */
  float* colstats;       /* TC path, optional: [M/32][n][2] per 32-row block {sum, sum of squares} of every stored
                          * column — the GroupNorm statistics of the tensor being written (util.py:214-216), consumed
                          * by ldmb200_groupnorm_fused so the normalisation never re-reads its input for statistics */
  int32_t pad_mode;      /* 0: symmetric padding ksize/2.  1: 3x3 stride 2 over an input padded (0,1,0,1) - one zero
                          * row below and one zero column right, none above / left - the Downsample of the first-stage
                          * Encoder (ldm/modules/diffusionmodules/model.py:72-76)                                    */
  int32_t pad_;
  /* TC path, optional split-K scratch (NULL: never split).  Batch 1-4 sampling leaves GEMMs of 8-74 output tiles with K
   * up to 11520 on 148 SMs; with a workspace the launcher splits such a tile's reduction over up to 8 CTAs, which leave fp32
   * partials here and fold them in split order (bit-reproducible) in whichever CTA arrives last, before the same fused
   * epilogue.  Layout: LDMB200_SPLITK_COUNTER_BYTES of arrival counters — ZERO before the first use, left zero by every
   * launch — followed by the partials.  16-byte aligned; one workspace per stream (launches that may overlap must not
   * share it).  16 MiB + the counters covers every shape the heuristic picks. */
  void* splitk_ws;
  int64_t splitk_ws_bytes;
  /* TC path, optional: a read-only range (the packed weights of a LATER launch) that this launch pulls into L2 with bulk
   * prefetches while it runs.  Small-batch sampling reads each GEMM's weights exactly once per eval, cold, from a handful of
   * CTAs; the launch before it has the HBM idle.  16-byte aligned, multiple of 16 bytes; NULL / 0: none. */
  const void* prefetch;
  int64_t prefetch_bytes;
} ldmb200_gemm_desc;
#define LDMB200_SPLITK_COUNTER_BYTES 1024

/* fp32-exact SIMT path (precision mode "fp32": per-step eps rel-L2 <= 1e-4 vs the fp32 reference). */
int ldmb200_igemm_simt(const ldmb200_gemm_desc* d, void* stream);
/* tcgen05 / TMEM / TMA path: bf16 operands, fp32 accumulation in tensor memory.
 * Requirements: a_dtype bf16 / f16, c1 % 16 == 0, c2 % 16 == 0, w packed; upsample = 1 needs the kind-4 pack. */
int ldmb200_igemm_tc(const ldmb200_gemm_desc* d, void* stream);

/* Weight packing for the tensor-core path (one launch for the whole model, re-run whenever weights change —
 * LitEma.copy_to/restore write param.data in place, ldm/modules/ema.py:46-76).
 * Each job converts one fp32 nn.Parameter into the bf16 K-major layout ldmb200_igemm_tc expects:
 *   kind 0  conv/linear weight [n, c, k, k] -> [n_pad, k, k, c_pad]   (zero padded)
 *   kind 1  same, rows permuted for the GEGLU epilogue (value/gate interleaved in groups of 16)
 *   kind 2  fp32 vector copy with the GEGLU permutation (bias of ff.net.0.proj)
 *   kind 3  data-gradient operand (training): the transposed, tap-flipped matrix
 *           [c_pad_as_rows, k, k, n_pad_as_cols]:  dst[cc, (2-kh, 2-kw), nn] = src[nn, cc, kh, kw], so that
 *           dX = conv(dY, dst) through the same ldmb200_igemm_tc (autograd's conv2d / linear input gradient).
 *           n_pad = padded row count (from c), c_pad = padded column count (from n); row_offset here offsets the
 *           destination COLUMN (fused QKV: the q, k, v blocks sit side by side along K)
 *   kind 4  folded Upsample convolution (K10): [n, c, 3, 3] -> [4 parities][n_pad][4 taps][c_pad], parity = py*2 + px of
 *           the output pixel, tap = a*2 + b of the 2x2 window on the low-res map; each tap is the fp32 sum of the 3x3 taps
 *           that read the same low-res pixel (py=0: a0 <- kh 0, a1 <- kh 1+2; py=1: a0 <- kh 0+1, a1 <- kh 2; kw alike)
 * jobs is a DEVICE array of n_jobs ldmb200_pack_job.
 */
typedef struct ldmb200_pack_job {
  const float* src;
  void* dst;
  int32_t kind;
  int32_t n, c, ksize;   /* source shape [n, c, ksize, ksize] (vectors: n)            */
  int32_t n_pad, c_pad;  /* destination padding (>= n, >= c)                          */
  int32_t row_offset;    /* destination row offset (fused QKV: q rows, k rows, v rows) */
  int32_t block_n;       /* GEGLU permutation tile (kind 1/2)                          */
  int32_t dst_dtype;     /* LDMB200_BF16 | LDMB200_F16 (kinds 0/1)                      */
  int32_t pad_;
} ldmb200_pack_job;
int ldmb200_pack_weights(const ldmb200_pack_job* jobs, int n_jobs, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * K5/K6  Attention — CrossAttention.forward attention.py:170-193: softmax(q k^T * d^-0.5) v per head,
 * head index = outer factor of the channel dim ('b n (h d) -> (b h) n d', attention.py:178), d_head = 32.
 *   q  [B, Nq,  q_ld ] (+q_off ), k [B, Nkv, kv_ld] (+k_off), v [B, Nkv, kv_ld] (+v_off): element (b,i,h,d) of q is
 *   q[(b*Nq + i)*q_ld + q_off + h*32 + d]; out [B, Nq, out_ld].  Self-attention passes the fused QKV matrix
 *   [B*N, 3C] as all three with offsets 0, C, 2C.
 * _simt: fp32 in/out (precision mode fp32, and the general cross-attention path with a short context);
 * _tc:   bf16 q/k/v in, out_dtype (bf16 | f16) out, flash-style on tcgen05 (S and PV in tensor memory, online
 *        softmax in fp32, P in bf16 so the exponent range of fp32 scores is kept).
 */
int ldmb200_attention_simt(const float* q, int q_ld, int q_off, const float* k, int k_off, const float* v, int v_off,
                           int kv_ld, float* out, int out_ld, int batch, int heads, int nq, int nkv, float scale,
                           void* stream);
int ldmb200_attention_tc(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                         int out_dtype, int batch, int heads, int n_tokens, float scale, void* stream);

/* Self-attention forward for training: as ldmb200_attention_tc, plus lse2[b, h, i] = log2 sum_j exp2(s_ij) with
 * s = scale * log2(e) * q.k (fp32 [B, heads, N]) which the backward kernels use to recompute P. */
int ldmb200_attention_tc_lse(const void* qkv, int ld, int q_off, int k_off, int v_off, void* out, int out_ld,
                             int out_dtype, float* lse2, int batch, int heads, int n_tokens, float scale, void* stream);

/* ===================================================================================================
 * Training step (BASELINE config 4): the backward kernels of LatentDiffusion.p_losses
 * (ldm/models/diffusion/ddpm.py:1014-1047; talking_face ddpm2cond.py:965-998) through the UNet.  In the reference
 * these are autograd's derivative formulas of the ATen ops cited per function; operands are bf16, accumulation
 * and the residual-stream gradient are fp32.  Parameter gradients are ACCUMULATED (+=) into fp32 buffers in the
 * nn.Parameter layouts, which the caller zeroes at the start of a step.
 * =================================================================================================== */

/* q_sample — ddpm.py:273-276:  out[b] = sqrt_ac[t_b] * x0[b] + sqrt_1mac[t_b] * noise[b]  (fp32, separately
 * rounded products like the reference's extract_into_tensor arithmetic).  per_sample = C*H*W. */
int ldmb200_q_sample(const float* x0, const float* noise, const int64_t* t, const float* sqrt_ac,
                     const float* sqrt_1mac, float* out, int batch, int64_t per_sample, void* stream);

/* get_loss(pred, target, mean=False).mean([1,2,3]) then .mean() with l_simple_weight = 1, logvar = 0
 * (ddpm.py:1029-1044, loss_type l2):  loss_per_sample[b] = mean_j (pred - target)^2 ;
 * dpred = 2 (pred - target) / (batch * per_sample)  — the gradient of the batch-mean loss. */
int ldmb200_mse_loss_grad(const float* pred, const float* target, float* loss_per_sample, float* dpred, int batch,
                          int64_t per_sample, void* stream);

/* Gradient cast + column sums.  g [rows, n] (in_dtype fp32 | bf16) -> g16 [rows, n] bf16 (nullable);
 * per_sample_sum[b, j] += sum of sample b's rows (nullable; [batch, ps_ld]; autograd of the `h + emb_out[..., None,
 * None]` broadcast openaimodel.py:270 and of the length-1 cross-attention row vector), total_sum[j] += sum over all
 * rows for j < n_valid (nullable; bias gradients); total_sum2 likewise (second bias fed by the same gradient). */
int ldmb200_grad_cast(const void* g, int in_dtype, void* g16, int64_t rows, int n, int rows_per_sample,
                      float* per_sample_sum, int ps_ld, float* total_sum, float* total_sum2, int n_valid, void* stream);

/* Weight gradient of a convolution / linear on the tensor cores (autograd of F.conv2d / F.linear w.r.t. weight):
 *   dw[nn, cc, tap] += sum_m dy[m, nn] * X(m, tap, cc)
 * dy bf16 [M, dy_ld] (M = batch*h*w output pixels, columns [0, n) used); x bf16 channels-last with row pitch c_ld,
 * channels [0, c) used: stride 1 -> [batch, h, w, c_ld]; stride 2 -> the space_to_depth2 planes [batch, 4, h, w, c_ld]
 * of the 2h x 2w input.  dw is addressed dw[nn*n_stride + cc*c_stride + tap*tap_stride] (fp32, atomically
 * accumulated; nn.Conv2d layout: n_stride = c*k*k, c_stride = k*k, tap_stride = 1). */
typedef struct ldmb200_wgrad_desc {
  const void* dy;
  const void* x;
  float* dw;
  int64_t n_stride, c_stride, tap_stride;
  int32_t dy_ld, n;
  int32_t c_ld, c;
  int32_t batch, h, w;
  int32_t ksize, stride;
  int32_t pad_;
} ldmb200_wgrad_desc;
int ldmb200_wgrad(const ldmb200_wgrad_desc* d, void* stream);

/* GroupNorm (+SiLU) backward — autograd of GroupNorm32 util.py:214-216 followed by nn.SiLU.
 * x1/x2, gamma, beta, groups, eps, silu, stats_ws: exactly what the forward call took (stats_ws still holds what it
 * left there; fused = 1 if that call was ldmb200_groupnorm_fused).  dy bf16 [B*hw, C].  addend (nullable; dtype
 * add_dtype fp32 | bf16, [B*hw, C]) is added to the input gradient (the identity skip / the raw-copy branch).
 * dx1 [B*hw, c1], dx2 [B*hw, c2] fp32: written (acc = 0) or accumulated (acc = 1).  dgamma/dbeta [C] accumulated.
 * ws2: workspace of ldmb200_groupnorm_ws_floats floats. */
int ldmb200_groupnorm_bwd(const float* x1, int c1, const float* x2, int c2, const float* gamma, const float* beta,
                          const void* dy, const void* addend, int add_dtype, float* dx1, int acc1, float* dx2,
                          int acc2, float* dgamma, float* dbeta, int batch, int hw, int groups, float eps, int silu,
                          const float* stats_ws, int fused, float* ws2, void* stream);

/* LayerNorm backward (nn.LayerNorm attention.py:203-205): dx = ln'(x)^T dy + addend (fp32, nullable); dx may alias
 * addend.  dy bf16 [rows, c].  dgamma/dbeta accumulated. */
int ldmb200_layernorm_bwd(const float* x, const float* gamma, const void* dy, const float* addend, float* dx,
                          float* dgamma, float* dbeta, int64_t rows, int c, float eps, void* stream);

/* GEGLU (attention.py:37-44) unfused for training: u bf16 [rows, 2*inner] = proj(x) (value | gate halves);
 * fwd: out = value * gelu_erf(gate) (bf16 [rows, inner]); bwd: du from dout. */
int ldmb200_geglu_fwd(const void* u, void* out, int64_t rows, int inner, void* stream);
int ldmb200_geglu_bwd(const void* u, const void* dout, void* du, int64_t rows, int inner, void* stream);

/* Self-attention backward (autograd of attention.py:178-192 with d_head = 32), recomputing P from lse2:
 * qkv bf16 [B*N, ld] (offsets as the forward), dout bf16 [B*N, o_ld], out bf16 [B*N, o_ld] (forward output, for
 * D = rowsum(dout * out)), lse2 from ldmb200_attention_tc_lse -> dqkv bf16 [B*N, ld] at the same offsets.
 * ws: workspace of ldmb200_attention_bwd_ws_floats floats ({lse2, D} pairs) — with it the tcgen05 kernels run;
 * NULL selects the warp-level (mma.sync) kernels. */
int64_t ldmb200_attention_bwd_ws_floats(int batch, int heads, int n_tokens);
int ldmb200_attention_bwd(const void* qkv, int ld, int q_off, int k_off, int v_off, const void* out, const void* dout,
                          int o_ld, const float* lse2, void* dqkv, float* ws, int batch, int heads, int n_tokens,
                          float scale, void* stream);

/* Resampling backward: zero_insert2x — dy bf16 [B, h, w, C] -> z bf16 [B, 2h, 2w, C], z[2i, 2j] = dy[i, j], else 0
 * (input gradient of the stride-2 Downsample.op conv = 3x3 conv of z with the kind-3 weights);
 * downsum2x — du fp32 [B, 2h, 2w, C] -> dx fp32 [B, h, w, C] = sum of each 2x2 block (autograd of F.interpolate
 * nearest 2x, openaimodel.py:116). */
int ldmb200_zero_insert2x(const void* dy, void* z, int batch, int h, int w, int c, void* stream);
int ldmb200_downsum2x(const float* du, float* dx, int batch, int h, int w, int c, void* stream);

/* Backward of ldmb200_small_linear (out = bias + act(in) W^T): dw [n, k] += dout^T act(in); dbias [n] += colsum(dout);
 * din [rows, k] (nullable) = or += (acc_din) (dout W) * act'(in).  Any of dw/dbias/din may be NULL. */
int ldmb200_small_linear_bwd(const float* in, int in_ld, const float* w, const float* dout, int dout_ld, float* dw,
                             float* dbias, float* din, int din_ld, int acc_din, int rows, int k, int n, int silu_in,
                             void* stream);

/* The same for a table of independent layers in ONE launch (DEVICE array of jobs).  din is ACCUMULATED atomically
 * (several jobs may share one input gradient: every ResBlock's emb projection reads emb, openaimodel.py:264).
 * k <= 4096 per job. */
typedef struct ldmb200_linear_bwd_job {
  const float* in;
  const float* w;
  const float* dout;
  float* dw;
  float* dbias;
  float* din;
  int32_t in_ld, dout_ld, din_ld;
  int32_t rows, k, n;
  int32_t silu_in;
  int32_t pad_;
} ldmb200_linear_bwd_job;
int ldmb200_small_linear_bwd_batched(const ldmb200_linear_bwd_job* jobs, int n_jobs, void* stream);

/* Fused AdamW step + LitEma update over flat fp32 arrays (torch.optim.AdamW semantics ddpm.py:1363-1372;
 * LitEma.forward ema.py:25-44: shadow -= (1 - decay) * (shadow - param)).  ema may be NULL. */
int ldmb200_adamw_ema(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* ema, int64_t n,
                      float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1,
                      float bias_corr2, float ema_one_minus_decay, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * First-stage decode - VQModelInterface.decode, ldm/models/autoencoder.py:274-283.
 *
 * vq_lookup_postconv: VectorQuantizer2.forward (taming/modules/vqvae/quantize.py:285-309) + post_quant_conv
 *   (autoencoder.py:43,280) + the channels-last staging of Decoder.conv_in's input, one thread per latent pixel.
 *   z fp32 NCHW [B, e_dim, HW]; codebook fp32 [n_e, e_dim] or NULL (force_not_quantize=True: z passes through);
 *   d = (sum z^2 + sum e^2) - 2 z.e in fp32, first index of the minimum; z_q = z + (e - z);
 *   y[pixel, c] = bias[c] + sum_e w[c, e] z_q[e] for c < z_channels, 0 for z_channels <= c < c_pad (out_dtype);
 *   indices int64 [B*HW] and zq fp32 NCHW are optional outputs (NULL to skip).
 * softmax_rows: p[r, :] = softmax(scale * s[r, :]) - AttnBlock.forward model.py:186-187 (single head, d = C, so
 *   the two contractions are plain GEMMs run through ldmb200_igemm_* with the per-sample K / V^T as "weight").
 *   s fp32, p out_dtype, n <= 16384. */
int ldmb200_vq_lookup_postconv(const float* z, const float* codebook, int n_e, int e_dim, const float* w,
                               const float* bias, int z_channels, void* y, int out_dtype, int c_pad, int64_t* indices,
                               float* zq, int batch, int hw, void* stream);
int ldmb200_softmax_rows(const float* s, int64_t s_ld, void* p, int out_dtype, int64_t p_ld, int rows, int n, float scale,
                         void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Audio conditioning encoder - Conv1DTemporalAttention.forward,
 * talking_face/ldm/modules/encoders/modules.py:76-113 (cond_stage_config_2 of mead-128-ldm-f4.yaml:75-80):
 *   s = attentionConvNet(x^T): n_layers x [Conv1d(k=3, pad=1) + LeakyReLU(slope)], channels[0] = D -> ... -> 1;
 *   a = softmax(Linear(L, L)(s));  out[b, :] = sum_l a[l] * x[b, l, :]        (torch.bmm(x^T, attention))
 * x fp32 [B, L, D], conv_w[i] fp32 [channels[i+1], channels[i], 3], lin_w fp32 [L, L]; out fp32 [B, D].
 * conv_w / conv_b / channels are HOST arrays (of device pointers / ints); L <= 32; one CTA per sample. */
int ldmb200_conv1d_temporal_attention(const float* x, float* out, const float* const* conv_w, const float* const* conv_b,
                                      const int* channels, int n_layers, const float* lin_w, const float* lin_b,
                                      int batch, int seq_len, float leaky_slope, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LDMB200_H_ */
