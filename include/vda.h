/*
 * libvda_sm100a.so -- C ABI of the B200-native Video-Depth-Anything inference hot path.
 *
 * The reference (FriedFeid/Video-Depth-Anything) is pure Python/PyTorch and has no FFI of its own
 * (SURVEY.md section 8b): every entry point below replaces the library kernel(s) that PyTorch eager
 * dispatches for the cited reference lines.  Conventions:
 *   - extern "C", plain pointers and sizes; no torch / C++ types in any signature;
 *   - every pointer is a DEVICE pointer owned by the caller unless the name says `host`;
 *   - every launcher takes a cudaStream_t (as void*) and is asynchronous; no hidden allocation, no hidden sync;
 *   - return value: 0 = ok, otherwise a cudaError_t (>0) or a negative library code; vda_last_error() gives text;
 *   - built only for sm_100a; there is no CPU path and no other architecture.
 *
 * Activation layout everywhere: token-major / channels-last.  A feature map of `frames` images of h x w
 * pixels and C channels is a row-major matrix [frames*h*w, C] ("dense"), or, when it feeds a 3x3
 * convolution, [frames*(h+2)*(w+2), C] with a zero one-pixel border per image ("padded").  In padded
 * form a 3x3 tap is a constant row offset, so convolutions are GEMMs with a 9-step K loop.
 */
#ifndef VDA_H_
#define VDA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VDA_VERSION 1

typedef void* vda_stream_t; /* cudaStream_t */

enum vda_dtype { VDA_F32 = 0, VDA_BF16 = 1 };

/* How a row index m of the GEMM's M space maps to a row of an output / residual matrix.
 * (b, y, x) is the pixel of m in a grid of gh x gw pixels per frame (dense M space: m = (b*gh+y)*gw+x;
 * padded M space: m = (b*(gh+2)+y+1)*(gw+2)+x+1, border rows produce no output). */
enum vda_rowmap {
  VDA_MAP_SAME = 0,    /* row = m                                                                  */
  VDA_MAP_DENSE = 1,   /* row = (b*gh + y)*gw + x                                                  */
  VDA_MAP_PADDED = 2,  /* row = (b*(gh+2) + y+1)*(gw+2) + x+1                                      */
  VDA_MAP_STRIDE2 = 3, /* only even y,x: row = (b*gh2 + y/2)*gw2 + x/2, gh2=(gh+1)/2, gw2=(gw+1)/2 */
  VDA_MAP_TOKENS = 4,  /* row = m + m/(gh*gw) + 1  (one class-token row precedes each frame)       */
  VDA_MAP_CONVT = 5,   /* conv-transpose k=s: row = padded index of (b, y*k+ky, x*k+kx) in a
                          (gh*k) x (gw*k) grid; (ky,kx) come from the column index                 */
  VDA_MAP_S2D = 6,     /* space-to-depth of the PADDED map for a stride-2 3x3 conv: with (Y, X) = (y+1, x+1),
                          row = (((Y&1)*2 + (X&1))*gb + b)*hp*wp + (Y>>1)*wp + (X>>1), hp = (gh+1)/2 + 1,
                          wp = (gw+1)/2 + 1: four parity planes of hp x wp pixels per frame (zero where unwritten) */
  VDA_MAP_CROP = 7     /* dense grid minus its last row and column: valid iff y < gh-1 and x < gw-1,
                          row = (b*(gh-1) + y)*(gw-1) + x  (the outputs of a stride-2 conv over parity planes) */
};

enum vda_act { VDA_ACT_NONE = 0, VDA_ACT_GELU = 1, VDA_ACT_GEGLU = 2, VDA_ACT_RELU = 3 };

/*
 * One dense contraction with a fused epilogue:  acc[m,n] = sum_{t<taps} sum_{k<K} A[m + off(t), k] * B[n, (s*taps+t)*K + k]
 * (summed over s < ksplit), followed by the epilogue described field by field below.
 *
 * Replaces, in the reference:
 *   nn.Linear  : dinov2_layers/attention.py:44-46,57-60 (qkv, proj), dinov2_layers/mlp.py:30-32,35-41 (fc1+GELU, fc2),
 *                layer_scale.py:27-28 + block.py:105-106 (LayerScale + residual, fused as `resid`),
 *                motion_module/motion_module.py:90,106,119,127 (proj_in/out), motion_module/attention.py:81-90 (to_q/k/v, to_out),
 *                attention.py:374,382-384 (GEGLU), :333 (ff out)
 *   nn.Conv2d  : patch_embed.py:66,76 (k14 s14 via vda_im2col_patch), dpt.py:50-58 (projects 1x1), :83-88 (3x3 s2),
 *                util/blocks.py:20-33 (layerN_rn), :56-62,78-91 (ResidualConvUnit), :123-125,160 (out_conv),
 *                dpt.py:117-124 (output_conv1, output_conv2 incl. the fused ReLU -> 1x1 -> ReLU tail)
 *   nn.ConvTranspose2d k=s : dpt.py:61-72 (resize_layers[0], [1])
 */
typedef struct vda_gemm_desc {
  /* ---- operands ---- */
  const void* A;     /* [rowsA, K] row-major, leading dimension lda (elements)                       */
  const void* B;     /* [N, ksplit*taps*K] row-major (K-major packed weights), leading dimension ldb */
  int64_t lda, ldb;
  int64_t rowsA;     /* rows of A that exist; reads outside [0,rowsA) yield 0                        */
  int32_t a_dtype;   /* vda_dtype of A                                                               */
  int32_t b_dtype;   /* vda_dtype of B                                                               */
  int32_t M, N, K;   /* K = contraction length per tap                                               */
  int32_t taps;      /* 1 = plain GEMM; 9 = 3x3 conv over a padded M space, off(t) = (t/3-1)*(gw+2) + (t%3-1) */
  int32_t ksplit;    /* 1, or 2 when B carries a (hi, lo) bf16 split of fp32 weights along K         */
  /* ---- geometry of the M space ---- */
  int32_t gb, gh, gw; /* frames, dense grid height / width (0,0,0 when no map needs them)             */
  int32_t m_padded;  /* 1: m runs over the padded (gh+2)x(gw+2) grid.
                        2 (taps = 9 only): stride-2 3x3 conv (dpt.py:83-88) over the four parity planes written through
                        VDA_MAP_S2D: m runs over ONE plane grid (gb, gh, gw) = (frames, hp, wp), A holds the planes
                        back to back, off(t) = ((dy&1)*2 + (dx&1))*gb*gh*gw + (dy>>1)*gw + (dx>>1) with (dy, dx) =
                        (t/3, t%3): only the kept outputs are computed (a quarter of the padded-grid form)        */
  /* ---- epilogue: v = acc + bias[n]; v = R(v); v = act(v) ... (R = round-to-bf16 when round_bf16) ---- */
  const float* bias; /* [N] (conv-transpose: [ct_cout], indexed by co) or NULL                        */
  int32_t act;       /* vda_act; GEGLU: B rows are packed in groups of 64 = 32 value rows + 32 gate rows,
                        output column count is N/2                                                   */
  int32_t round_bf16; /* 1: reproduce the autocast rounding points; 0: pure fp32 arithmetic            */
  int32_t out_dtype; /* storage type of out / out_relu / res / res2                                  */
  void* out;         /* primary output (may be NULL when only `resid` or `head_out` is produced)     */
  void* out_relu;    /* optional second copy, relu(v)                                                */
  int64_t ldo;
  int32_t out_map;   /* vda_rowmap                                                                   */
  int32_t ct_k;      /* VDA_MAP_CONVT: kernel = stride; N = ct_k*ct_k*ct_stride, n = (ky*ct_k+kx)*ct_stride+co */
  int32_t ct_cout;   /* real output channels: columns with co >= ct_cout are padding (zero weights) and are not stored */
  int32_t ct_stride; /* columns per (ky,kx) block, >= ct_cout; 0 means ct_cout.  Padding the block to a multiple of 64
                        columns keeps narrow transposed convs (48 / 96 channels) on the coalesced epilogue */
  const void* res;   /* optional: v = R(v + res[row, n])                                             */
  const void* res2;  /* optional: v = R(v + res2[row, n])                                            */
  int64_t ldres;
  int32_t res_map;
  /* fp32 residual stream (encoder): resid[row, n] = resid[row, n] + gamma[n]*v,
     or with resid_assign: resid[row, n] = v + pos[(row % pos_rows), n]  (patch embed + pos-embed)    */
  float* resid;
  const float* gamma;
  const float* pos;
  int64_t ldresid;
  int32_t resid_map;
  int32_t resid_assign;
  int32_t pos_rows;
  /* fused output tail (dpt.py:119-123): d = relu(sum_n relu(acc+bias)[n]*head_w[n] + head_b); head_out[row] = d (fp32, never rounded) */
  const float* head_w;
  void* head_out;    /* [rows] fp32 (the island runs outside autocast, dpt_temporal.py:95-97)        */
  float head_b;
  /* ---- execution ---- */
  int32_t use_tc;    /* 1: tcgen05/TMEM/TMA kernel (bf16 operands); 0: SIMT fp32-accumulate kernel   */
  void* workspace;   /* SIMT path only: >= M*N*4 bytes of scratch                                    */
  int64_t workspace_bytes;
} vda_gemm_desc;

const char* vda_last_error(void);
int vda_version(void);
/* 0 when device `dev` is an sm_100 part and the library can run; fills sm_count when non-NULL. */
int vda_device_check(int dev, int* sm_count);

int vda_gemm(const vda_gemm_desc* d, vda_stream_t stream);

/* LayerNorm over the last dim D of x[rows, D] (dinov2.py:95 eps 1e-6; block.py:83-87; dinov2.py:309-310;
 * motion_module.py:162-170,175-183 eps 1e-5).  y = R((x-mean)*rstd*w + b (+ pe[(r / pe_group) % pe_T, :])).
 * skip_first/group_rows: drop row 0 of every group of `group_rows` rows from the output (class token,
 * dinov2.py:312) -- the output then has rows - rows/group_rows rows. */
int vda_layernorm(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w, const float* b,
                  float eps, int64_t rows, int32_t D, int32_t group_rows, int32_t skip_first,
                  const float* pe, int32_t pe_group, int32_t pe_T, vda_stream_t stream);

/* GroupNorm(groups, C, eps) per frame over dense tokens x[frames, hw, C] (motion_module.py:89,116) -> y tokens.
 * stats: scratch of frames*groups*2*34 doubles (mean/rstd, per-block partial sums, float copies for the apply pass;
 *        the reduction is deterministic). */
int vda_groupnorm_tokens(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w, const float* b,
                         float eps, int32_t frames, int32_t hw, int32_t C, int32_t groups, double* stats,
                         vda_stream_t stream);

/* Allocation-free variant of vda_groupnorm_tokens.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  Same kernels and results as vda_groupnorm_tokens; the statistics scratch
 * lives in a file-scope static __device__ array of the library (1.06 MiB, held by every GPU that loads the library until
 * the process exits) whose device address is resolved once per device with cudaGetSymbolAddress.  Limit:
 * frames * groups <= vda_groupnorm_noalloc_max_frames_x_groups() (2048, e.g. 64 frames x 32 groups); larger calls
 * return -1 with a message in vda_last_error().  The call performs no cudaMalloc / cudaFree and no other allocating
 * call.  All calls on a device share the one array (it keeps the last call's contents; nothing in it has to start at
 * zero): calls that overlap on the GPU -- different streams -- race, so serialise them on one stream. */
int vda_groupnorm_tokens_noalloc(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w, const float* b,
                                 float eps, int32_t frames, int32_t hw, int32_t C, int32_t groups, vda_stream_t stream);
int32_t vda_groupnorm_noalloc_max_frames_x_groups(void);

/* Encoder self-attention (dinov2_layers/attention.py:49-59): qkv[frames*ntok, 3*heads*64] -> o[frames*ntok, heads*64];
 * softmax(0.125 * q k^T) v per (frame, head).  head_dim is 64 for every DINOv2 size. */
int vda_attention(const void* qkv, void* o, int32_t dtype, int32_t frames, int32_t ntok, int32_t heads,
                  int32_t round_bf16, int32_t use_tc, vda_stream_t stream);

/* Temporal attention over the frame axis (motion_module.py:253,317-322; motion_module/attention.py:182-211):
 * qkv[(b*T + t)*hw + p, 3*C] -> o[(b*T+t)*hw + p, C]; `heads` heads of C/heads; one problem per (b, p, head). */
int vda_temporal_attention(const void* qkv, void* o, int32_t dtype, int32_t batch, int32_t T, int32_t hw, int32_t C,
                           int32_t heads, int32_t round_bf16, vda_stream_t stream);

/* Rotary position embedding of the temporal attention, pe='rope' (motion_module/attention.py:403-429, applied to query
 * and key at motion_module.py:290-293 before the head split): columns [0, 2C) of qkv[rows, 3C] are rotated in place,
 * pair (2i, 2i+1) of row r by the angle t*theta_i with t = (r / hw) % T, in fp32, result rounded to the activation type.
 * tab[T, C] holds (cos, sin)(t*theta_i) interleaved -- the caller's `freqs_cis` (precompute_freqs_cis, :403-408). */
int vda_rope_qk(void* qkv, int32_t dtype, const float* tab, int64_t rows, int32_t hw, int32_t T, int32_t C,
                int32_t round_bf16, vda_stream_t stream);

/* Bilinear resize, align_corners=True (blocks.py:156-158, dpt_temporal.py:92-94), channels-last.  Under CUDA autocast
 * upsample_bilinear2d is an fp32 op: bf16 in -> fp32 out is the un-rounded form the output head consumes. */
int vda_bilinear(const void* in, int32_t in_dtype, void* out, int32_t out_dtype, int32_t frames, int32_t ih, int32_t iw,
                 int32_t in_padded, int32_t oh, int32_t ow, int32_t out_padded, int32_t C, vda_stream_t stream);

/* Fused output head of the bf16 build (dpt_temporal.py:92-97, dpt.py:118-124): x[frames, ih, iw, C] bf16 (the output_conv1
 * result, dense channels-last) -> bilinear align_corners=True to (oh, ow) in fp32 (never materialised) -> 3x3 conv C->32,
 * fp32 weights w[32, 9*C] (k = (dy*3+dx)*C + ci) on the tensor cores in TF32, + bias -> ReLU -> 1x1 conv 32->1 (head_w,
 * head_b) -> ReLU -> depth[frames, oh, ow] fp32.  C must be a multiple of 32.  Same arithmetic as
 * vda_bilinear (fp32 out, padded) followed by vda_gemm (taps = 9, f32 operands, head_w), without the fp32 round trip. */
int vda_output_head(const void* x, const float* w, const float* bias, const float* head_w, float head_b, float* depth,
                    int32_t frames, int32_t ih, int32_t iw, int32_t C, int32_t oh, int32_t ow, vda_stream_t stream);

/* The same head with the 3x3 conv's operands in IEEE fp16 (tcgen05 kind::f16, K = 16 per MMA: half the MMAs of the TF32
 * form).  w_f16[32, 9*C] is the fp32 weight rounded to nearest fp16 by the caller; the interpolated activations are
 * rounded to nearest fp16 in the kernel (saturating at +-65504).  fp16 carries the 10 explicit mantissa bits TF32 does,
 * so the products are those of the TF32 conv up to round-to-nearest vs. truncation of the operands; accumulation is
 * fp32.  C must be a multiple of 64. */
int vda_output_head_f16(const void* x, const void* w_f16, const float* bias, const float* head_w, float head_b,
                        float* depth, int32_t frames, int32_t ih, int32_t iw, int32_t C, int32_t oh, int32_t ow,
                        vda_stream_t stream);

/* The form the bf16 build runs: fp16 operands where they are safe, TF32 where they are not, decided on the device.
 * A range guard first sets *flag (device int32) when any element of x is beyond +-6e4 or non-finite (the interpolated
 * operands are convex combinations of x, so they stay inside the same range); the fp16-operand kernel returns at once
 * when the flag is set, the TF32 kernel (fp32's exponent range: what the reference's island computes under cuDNN's
 * defaults, dpt_temporal.py:95-97) returns at once when it is clear and otherwise recomputes the whole launch.  No host
 * synchronisation; the caller may read *flag afterwards to learn which form produced `depth`.  The caller checks ONCE
 * that the weights themselves fit fp16 (max |w| <= 6e4) before choosing this entry point. */
int vda_output_head_guarded(const void* x, const void* w_f16, const float* w_f32, const float* bias, const float* head_w,
                            float head_b, float* depth, int32_t frames, int32_t ih, int32_t iw, int32_t C, int32_t oh,
                            int32_t ow, int32_t* flag, vda_stream_t stream);

/* Patch-embed im2col (patch_embed.py:76): x[frames,3,H,W] f32 -> A[frames*(H/14)*(W/14), kpad] (k = c*196+i*14+j; pad = 0). */
int vda_im2col_patch(const float* x, void* A, int32_t a_dtype, int32_t frames, int32_t H, int32_t W, int32_t kpad,
                     vda_stream_t stream);

/* The same im2col with one pointer per frame: frame f is the planar [3, H, W] f32 image at frame_ptrs[f] (8-byte aligned).
 * frame_ptrs is a DEVICE array of `frames` device pointers, read when the kernel runs -- the sliding-window driver rewrites
 * it per window, so a window's 32 inputs (video_depth.py:358-364: new frames + the previous window's key frames) are
 * gathered from the once-pre-processed frame pool by the patch-embed itself, also under a replayed CUDA graph. */
int vda_im2col_patch_frames(const float* const* frame_ptrs, void* A, int32_t a_dtype, int32_t frames, int32_t H, int32_t W,
                            int32_t kpad, vda_stream_t stream);

/* Class-token rows of the residual stream (dinov2.py:220-221): resid[f*ntok, :] = cls + pos[0, :]. */
int vda_cls_rows(float* resid, const float* cls, const float* pos, int32_t frames, int32_t ntok, int32_t D,
                 vda_stream_t stream);

/* Bicubic (a=-0.75, align_corners=False, explicit scale) resample of the 37x37 position grid (dinov2.py:179-210). */
int vda_pos_embed_bicubic(const float* pos, float* out, int32_t D, int32_t src, int32_t oh, int32_t ow,
                          double scale_h, double scale_w, vda_stream_t stream);

/* ---- sliding-window driver kernels (video_depth.py:329-417, util/transform.py, utils/util.py) ---- */

/* uint8 HWC frame -> /255 -> cv2.INTER_CUBIC resize to (oh, ow) -> (x-mean)/std -> planar f32 [3, oh, ow]
 * (transform.py:109-158, video_depth.py:361), `n` frames at once. */
int vda_preprocess_u8(const uint8_t* frames, float* out, int32_t n, int32_t fh, int32_t fw, int32_t oh, int32_t ow,
                      vda_stream_t stream);

/* depth[n, ih, iw] (dtype) -> f32 [n, oh, ow], bilinear align_corners=True on the f32-converted values
 * (video_depth.py:370-371). */
int vda_resize_depth(const void* in, int32_t dtype, float* out, int32_t n, int32_t ih, int32_t iw, int32_t oh,
                     int32_t ow, vda_stream_t stream);

/* Scale/shift least squares (utils/util.py:40-62 with an all-ones mask):
 * st[0..1] = (scale, shift) minimising |scale*pred + shift - target|^2; the five sums are accumulated in f64 in
 * scratch[0..4].  phase bit 0: zero the scratch first; bit 1: solve into st after accumulating.  The reference fits
 * over two frames (video_depth.py:390-395): call once per frame pair with phase 1 then 2. */
int vda_scale_shift(const float* pred, const float* target, int64_t count, double* scratch, float* st, int32_t phase,
                    vda_stream_t stream);

/* The same fit (utils/util.py:40-62, video_depth.py:390-395) over one contiguous (pred, target) pair with a DETERMINISTIC
 * reduction: fixed 4096-element chunks -> partial[chunk][4] (f64), then one block adds the partials in a fixed order and
 * solves.  No atomics; the result does not depend on the grid or the SM count, so every rank of a window-sharded run
 * computes bit-identical (scale, shift).  partial: vda_scale_shift_det_workspace_doubles(count) doubles, fully rewritten. */
int64_t vda_scale_shift_det_workspace_doubles(int64_t count);
int vda_scale_shift_det(const float* pred, const float* target, int64_t count, double* partial, float* st,
                        vda_stream_t stream);

/* out = max(in*st[0] + st[1], 0) over count elements (video_depth.py:399-412). */
int vda_scale_clamp(const float* in, float* out, const float* st, int64_t count, vda_stream_t stream);

/* ---- evaluation aligner + metrics (utils/align.py:151-236, utils/metrics.py:92-192, eval.py:163-181) ---- */

/* Least squares of [pred, 1] c = 1/gt over the pixels with valid != 0 and gt != 0 (the reference inverts the MASKED ground
 * truth, utils/align.py:142-147,152-155; numpy does this in float64): coef[0] = scale = 1/c0, coef[1] = shift = -c1/c0
 * ((inf, 0) when c0 == 0), so that (pred - shift)/scale is the aligned inverse depth.  The five sums accumulate in f64 in scratch[0..4]; phase bit 0 zeroes
 * the scratch first, bit 1 solves after accumulating (several calls may feed one fit).  `valid` is a byte mask. */
int vda_eval_fit(const float* pred, const float* gt, const uint8_t* valid, int64_t count, double* scratch, double* coef,
                 int32_t phase, vda_stream_t stream);

/* aligned = clip(1 / (a == 0 ? 1e-4 : a), 0, max_depth) with a = clip((pred - shift)/scale, 0, 1) for every pixel
 * (utils/align.py:228-236, evaluated in f64 like the reference and stored as f32; `aligned` may be NULL), and over the valid pixels the sums behind the reference's metrics
 * (utils/metrics.py): sums[0..7] = n, sum|d-g|, sum|d-g|/g, sum(d-g)/g, sum(d-g)^2, and the counts of
 * max(d/g, g/d) > 1.25, 1.25^2, 1.25^3 (f64, zeroed by the call). */
int vda_eval_apply(const float* pred, const float* gt, const uint8_t* valid, const double* coef, double max_depth,
                   float* aligned, int64_t count, double* sums, vda_stream_t stream);

/* out = in*st[0] + st[1] over count elements, not clamped: the per-frame alignment of the streaming mode
 * (`infere_single_image`, video_depth.py:310-315). */
int vda_scale_apply(const float* in, float* out, const float* st, int64_t count, vda_stream_t stream);

/* Frame gather / scatter of the streaming mode's feature store (video_depth.py:228-231,244-258,281-296; the `torch.cat`
 * of stored and new features at dpt_temporal.py:188-196 and the prediction subset at :222-227):
 * frame dst_idx[i] of dst <- frame src_idx[i] of src for i < n; a NULL index array means i.  Index arrays are DEVICE
 * int32 arrays so the launch is static under a CUDA graph.  frame_bytes must be a multiple of 16. */
int vda_copy_frames(const void* src, void* dst, const int32_t* src_idx, const int32_t* dst_idx, int32_t n,
                    int64_t frame_bytes, vda_stream_t stream);

/* Depth visualisation of `save_video(..., is_depths=True)` (utils/dc_utils.py:70-86), SURVEY section 8f rank 2.
 * vda_depth_minmax: minmax[0], minmax[1] = min, max over `count` f32 values (`frames.min(), frames.max()`, :74);
 * `partial` is caller-provided scratch of vda_depth_minmax_workspace_floats() floats.
 * vda_depth_vis: u = trunc((depth - min) / (max - min) * 255) in f32 arithmetic like numpy (:77, bit-identical; 0 when
 * max == min, where numpy casts a NaN); out[i] = u (lut == NULL: `grayscale=True`, 1 byte per pixel) or
 * out[3i..3i+2] = lut[3u..3u+2] (lut = DEVICE uint8[256][3], the caller's `(colormap * 255).astype(uint8)`: :78-81). */
int64_t vda_depth_minmax_workspace_floats(void);
int vda_depth_minmax(const float* depth, int64_t count, float* partial, float* minmax, vda_stream_t stream);
int vda_depth_vis(const float* depth, int64_t count, const float* minmax, const uint8_t* lut, uint8_t* out,
                  vda_stream_t stream);

/* 8-frame linear cross-fade (utils/util.py:65-74, video_depth.py:397-402):
 * inout[i] = inout[i]*(1-w_i) + max(post[i]*st[0]+st[1], 0)*w_i, w = i/(n-1). */
int vda_blend(float* inout, const float* post, const float* st, int32_t n, int64_t frame_elems, vda_stream_t stream);

/* Plain device-to-device copy of `bytes` bytes on `stream` (window 0's frames enter the output un-scaled and un-clamped,
 * video_depth.py:382-387; the exchange buffers of the window-sharded driver). */
int vda_copy_device(const void* src, void* dst, int64_t bytes, vda_stream_t stream);

/* Page-lock (cudaHostRegisterPortable) / release a HOST address range: each rank of a window-sharded run registers its
 * mapping of the node-wide shared result buffer, so that its device->host copies of the final frames (video_depth.py:372
 * in the reference: one blocking copy per frame) are asynchronous DMA over its own PCIe link.  On failure the error is
 * returned (text in vda_last_error) and cleared from the CUDA runtime's last-error slot. */
int vda_host_register(void* host_ptr, int64_t bytes);
int vda_host_unregister(void* host_ptr);

/* The same cross-fade from two RAW tails (window-sharded driver: the previous window's last n frames come un-aligned from
 * the rank that owns it): out[i] = P*(1-w_i) + Q*w_i with P = max(pre[i]*st_pre[0]+st_pre[1], 0) (P = pre[i] when st_pre
 * is NULL: window 0 is never scaled or clamped, video_depth.py:382-387) and Q = max(post[i]*st_post[0]+st_post[1], 0).
 * Bit-identical to vda_scale_clamp into a buffer followed by vda_blend on it. */
int vda_blend_raw(const float* pre, const float* st_pre, const float* post, const float* st_post, float* out, int32_t n,
                  int64_t frame_elems, vda_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* VDA_H_ */
