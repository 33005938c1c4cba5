/* b200sd — C ABI of the B200-native Stable-Diffusion-class batch text-to-image hot path.
 *
 * The reference (christianshub/k8s-nvidia-gpus) defines no C ABI / FFI: its compute is the un-vendored
 * diffusers/transformers/torch stack called from cluster-config/apps/sd15-api/configmap.yaml:105-112
 * (`pipe(prompt, num_inference_steps, guidance_scale, width, height, generator)`), reached over HTTP from
 * scripts/batch_generate.py:20-26.  Each entry point below therefore cites the *operation inside that call*
 * which it replaces (SURVEY.md §8(a) row numbers, upstream module names).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; the CALLER owns every buffer (inputs, outputs, scratch).
 *   - all pointers are device pointers unless a field says "host".
 *   - activations and weights are 16-bit floats, written "act16" below: IEEE fp16 by default — the reference's own
 *     CUDA dtype (TORCH_DTYPE = float16, configmap.yaml:31) — or bfloat16 when the library is built with
 *     -DB200SD_ACT_BF16; b200sd_act_dtype() tells which.  Accumulation, statistics and the latents are fp32.
 *   - activations are NHWC ("channels-last"); token tensors [B, HW, C] alias the same memory.
 *   - every call is asynchronous on `stream`, performs no allocation and no synchronisation, and can be
 *     captured into a CUDA graph (exception: b200sd_device_info, which only queries).
 *   - return value 0 = ok, negative = b200sd_status; b200sd_last_error() (thread-local) has the detail.
 *   - sm_100a only.  There is no CPU path and no other backend.
 */
#ifndef B200SD_H_
#define B200SD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200sd_stream; /* cudaStream_t */

enum b200sd_status {
  B200SD_OK = 0,
  B200SD_ERR_BAD_SHAPE = -1,
  B200SD_ERR_BAD_ALIGNMENT = -2,
  B200SD_ERR_UNSUPPORTED_ARCH = -3,
  B200SD_ERR_CUDA = -4,
  B200SD_ERR_BAD_ARGUMENT = -5,
  B200SD_ERR_TOO_LARGE = -6
};

enum b200sd_act {
  B200SD_ACT_NONE = 0,
  B200SD_ACT_SILU = 1,
  B200SD_ACT_QUICK_GELU = 2,
  B200SD_ACT_GEGLU = 3,
  B200SD_ACT_GELU = 4 /* exact (erf) GELU: the MLP of OpenCLIP-bigG, SDXL's second text encoder */
};
enum b200sd_dtype { B200SD_ACT16 = 0, B200SD_F32 = 1 }; /* output type selector: act16 (see above) or f32 */

int b200sd_version(void);
/* 16-bit element type this build computes in: 0 = bfloat16, 1 = IEEE fp16. */
int b200sd_act_dtype(void);
const char* b200sd_last_error(void);
/* out[0]=SM count, out[1]=cc major, out[2]=cc minor, out[3]=max dynamic smem per block. */
int b200sd_device_info(int32_t* out4);

/* ---------------------------------------------------------------------------------------------------------
 * Implicit-GEMM convolution / linear layer on tcgen05 (TMA-fed, fp32 accumulation in TMEM).
 * Replaces: torch conv2d 3x3 s1 p1 / 1x1 (ResnetBlock2D.conv1/conv2/conv_shortcut, Transformer2DModel.proj_in/out,
 * Upsample2D.conv, conv_in/conv_out, VAE decoder convs) and nn.Linear (attention projections, GEGLU, time MLP,
 * CLIP MLP) inside UNet2DConditionModel.forward / AutoencoderKL.decode / CLIPTextModel — SURVEY §8(a) a7,a10,a13;
 * call site configmap.yaml:105.
 *
 *   out[m, n] = act( alpha * sum_{tap,c} A[pixel(m)+tap, c] * W[n, tap, c] + bias[n] + rowbias[img(m), n] )
 *               + residual[m, n]
 * A is the virtual channel-concatenation of up to two NHWC sources (a0: c0 channels, a1: c1 channels; a1 may be
 * NULL with c1 = 0) — the UNet up-path skip concat is never materialised.  ksize = 1 or 3 (stride 1, zero padding
 * ksize/2).  A linear layer over M rows is nimg = 1, h = 1, w = M, ksize = 1.
 * W is act16 [n, ksize*ksize, c0p + c1p] with c0p/c1p = c0/c1 rounded up to 64 (zero padded), K contiguous.
 * GEGLU: W rows/bias packed in 64-row groups [32 value rows | 32 gate rows]; out has n/2 columns.
 * Requirements: c0, c1, ld0, ld1 multiples of 8; pointers 16-byte aligned.
 */
typedef struct {
  const void* a0;       /* act16 NHWC source 0, pixel stride ld0 elements */
  const void* a1;       /* act16 NHWC source 1 or NULL */
  const void* w;        /* act16 packed weights */
  void* out;            /* act16 or f32 [M, ldc] */
  const float* bias;    /* f32 [n] or NULL */
  const float* rowbias; /* f32 [nimg, ld_rowbias] or NULL (time-embedding projection) */
  const void* residual; /* act16 [M, ldr] or NULL */
  int32_t nimg, h, w_, c0, c1, ld0, ld1;
  int32_t n;      /* GEMM N (weight rows) */
  int32_t ksize;  /* 1 or 3 */
  int32_t ldc, ldr, ld_rowbias;
  int32_t act;       /* b200sd_act */
  int32_t out_dtype; /* b200sd_dtype */
  int32_t bn_hint;   /* 0 = auto.  Else the tile width in columns: a multiple of 64 (32 for f32 out, 128 for GEGLU), <= 256,
                        or <= 512 in multiples of 64 (wide single-accumulator tiles); + 1000 forces single-CTA tiles,
                        + 2000 forces CTA pairs.  For tests and tuning sweeps. */
  float alpha;       /* accumulator scale (1.0 for plain layers) */
  /* Optional fused GroupNorm statistics of the OUTPUT (the next layer's GroupNorm, torch.nn.GroupNorm inside
   * ResnetBlock2D / Transformer2DModel): the epilogue adds sum and sum of squares of out[., g * n/gn_groups ...] per
   * (image, group) to gn_stats - the same scratch b200sd_groupnorm_stats fills (b200sd_groupnorm_workspace_bytes;
   * order-independent integer accumulation) - the caller zeroes it; b200sd_groupnorm_apply then needs no
   * b200sd_groupnorm_stats pass.  16-bit, non-GEGLU outputs only; n / gn_groups whole and even.  The sums are
   * taken over the ROUNDED 16-bit outputs (what the apply pass reads).  Every 32-row slab of
   * a tile must lie in one image: gn_rows_per_img = 0 takes images from the NHWC geometry (needs
   * pow2ceil(w) * pow2ceil(h) >= 32); for a flattened geometry (nimg = 1, h = 1) pass the rows per image (% 32 == 0). */
  void* gn_stats;
  int32_t gn_groups;
  int32_t gn_rows_per_img;
  /* Optional filter / output geometry (all 0 = the plain ksize x ksize, stride-1, dense-output layer above).
   *   kh, kw (1..3), off_h, off_w: explicit filter window; tap (i, j) reads input pixel
   *       (y * stride + i + off_h, x * stride + j + off_w).  W is then [n, kh*kw, c0p + c1p].
   *   stride = 2: Downsample2D (conv 3x3, stride 2, padding 1: off = -1).  h, w_ are the OUTPUT size; the sources are
   *       [nimg, 2h, 2w_, c] and the TMA unit reads every other pixel - no patch matrix is materialised.
   *   out_pix_stride / out_row_stride / out_img_stride (elements, multiples of 16 bytes): the output pixels are
   *       scattered into a larger tensor.  Upsample2D (nearest 2x + conv 3x3) runs as four 2x2 convolutions on the
   *       LOW-resolution input (weights pre-summed per output phase (py, px); taps at off = py - 1, px - 1), each
   *       writing its quarter of the pixels: out + ((py * 2w_) + px) * C, pixel stride 2C, row stride 4 w_ C.
   *       No upsampled tensor exists and 16 instead of 36 MACs are spent per input pixel and channel pair. */
  int32_t kh, kw, off_h, off_w, stride;
  int32_t out_pix_stride, out_row_stride, out_img_stride;
  /* Split-K (0 or 1 = off): the reduction of every output tile is cut into split_k slices computed by different
   * thread blocks - for layers with few output tiles and a long reduction (the 8x8-level convolutions: 40 tiles for
   * 148 SMs at batch 8; everything at batch 1).  Each slice parks its fp32 accumulator tile in `workspace`; the block
   * that arrives last adds the slices in INDEX order (bit-identical results whatever the arrival order) and runs the
   * epilogue.  `workspace`: b200sd_gemm_workspace_bytes(p) bytes, 256-byte aligned, zero-filled once before its first
   * use (calls leave its counters at zero, so consecutive calls on one stream may share it).  Single-CTA tiles only. */
  int32_t split_k;
  void* workspace;
  int64_t workspace_bytes;
  /* LayerNorm folded into the layers around it (nn.LayerNorm norm1-3 of BasicTransformerBlock; K9): no normalised
   * tensor is written or re-read.
   *   ln_out_stats: the PRODUCER of a LayerNorm input adds sum and sum of squares of every output row to
   *       ln_out_stats [M, B200SD_GN_STATS_WORDS 4-byte words] (same integer format as the GroupNorm statistics; the
   *       caller zeroes it; sums over the rounded 16-bit outputs, all column tiles of the row).
   *   ln_in_stats + ln_colsum + ln_eps: the CONSUMER (a linear layer whose input x is to be normalised) runs on the
   *       RAW x with weights W' = gamma (.) W and computes
   *           out[m, n] = rstd_m * (x W'^T)[m, n] - rstd_m * mean_m * ln_colsum[n] + bias[n]
   *       with (mean_m, rstd_m) from ln_in_stats (c0 = the normalised width), ln_colsum[n] = sum_k W'[n, k] (of the
   *       16-bit weights as stored) and bias = b + W beta precomputed by the caller.  Exact in real arithmetic. */
  void* ln_out_stats;
  const void* ln_in_stats;
  const float* ln_colsum;
  float ln_eps;
  /* Non-zero: the caller guarantees that nothing still in flight on the device writes `w` when this call is enqueued
   * (model weights, as opposed to an activation used as the B operand).  The kernel is launched with programmatic
   * stream serialisation: its blocks may become resident while the previous kernel of the stream is still running,
   * and with this flag they fetch the first weight tiles (and prefetch the rest of their first tile's weights into
   * L2) before waiting for that kernel; everything else (a0 / a1, residual, biases, statistics, the output) is only
   * touched after it has completed.  0: weights are read after the wait as well. */
  int32_t weights_static;
} b200sd_gemm_params;
int b200sd_gemm(const b200sd_gemm_params* p, b200sd_stream stream);
/* bytes of split-K workspace these parameters need (0 when split_k <= 1 or the parameters are rejected) */
size_t b200sd_gemm_workspace_bytes(const b200sd_gemm_params* p);

/* ---------------------------------------------------------------------------------------------------------
 * Flash attention on tcgen05: out = softmax(Q K^T * scale [+ causal mask]) V, per (batch, head).
 * Replaces: Attention / SlicedAttnProcessor (configmap.yaml:42 enables slicing — identity up to rounding) in
 * BasicTransformerBlock.attn1/attn2 and CLIPAttention — SURVEY §8(a) a7, a10.3.
 * Q: [batch, nq, ldq] bf16, head h at columns [h*d, (h+1)*d); K, V likewise with nk rows.  out: [batch, nq, ldo].
 * d (head dim) multiple of 8, <= 160.  ld* multiples of 8.
 */
typedef struct {
  const void* q;
  const void* k;
  const void* v;
  void* out;
  int32_t batch, heads, nq, nk, d;
  int32_t ldq, ldk, ldv, ldo;
  int32_t causal;
  float scale;
} b200sd_attn_params;
int b200sd_attention(const b200sd_attn_params* p, b200sd_stream stream);

/* ---------------------------------------------------------------------------------------------------------
 * GroupNorm (+ optional SiLU) over the virtual concat of two NHWC sources; two passes.
 * Replaces: torch group_norm + silu in ResnetBlock2D.norm1/norm2, Transformer2DModel.norm, conv_norm_out, VAE
 * norms — SURVEY §8(a) a10.2/a10.3/a13 (K7/K8/K14).
 * stats: caller-owned scratch of b200sd_groupnorm_workspace_bytes(p) bytes (B200SD_GN_STATS_WORDS 4-byte words per
 * (image, group), 16-byte aligned); b200sd_groupnorm_stats ACCUMULATES into it (the caller zeroes it first).
 * The sums are kept as 64-bit fixed-point integers, so the totals - and therefore the images - are bit-identical from
 * run to run and do not depend on the order in which thread blocks finish (fp32 atomics would).
 * Limits: c0, c1 multiples of 8, (c0 + c1) divisible by groups and <= 6144, groups <= 1024 (B200SD_ERR_BAD_SHAPE /
 * B200SD_ERR_TOO_LARGE otherwise); nimg <= 65535 (one grid row per image).
 */
#define B200SD_GN_STATS_WORDS 8
typedef struct {
  const void* x0;
  const void* x1; /* or NULL */
  void* stats;
  const float* gamma; /* f32 [c0+c1] */
  const float* beta;  /* f32 [c0+c1] */
  void* out;          /* act16 [nimg, hw, c0+c1] */
  int32_t nimg, hw, c0, c1, groups;
  int32_t silu;
  float eps;
} b200sd_groupnorm_params;
int b200sd_groupnorm_stats(const b200sd_groupnorm_params* p, b200sd_stream stream);
int b200sd_groupnorm_apply(const b200sd_groupnorm_params* p, b200sd_stream stream);
/* bytes of `stats` scratch for p->nimg images and p->groups groups (0 for bad arguments) */
size_t b200sd_groupnorm_workspace_bytes(const b200sd_groupnorm_params* p);

/* GroupNorm (+ optional SiLU) in ONE call that needs no scratch from the caller and allocates nothing.
 * (No counterpart in the reference: it has no C ABI at all; this exists for hosts of this library.)
 *
 * Who needs it: a caller that is handed only its input and output buffers and has no way to obtain more device
 * memory at call time — for example a normalisation hook running inside another runtime's pre-planned memory arena,
 * or a service pinned to a fixed memory budget (the reference pod runs under a hard memory limit, deployment.yaml:
 * 59-66) where a cudaMalloc in the request path may fail, and in any case synchronises the whole device.
 *
 * Same parameters and same result as b200sd_groupnorm_stats followed by b200sd_groupnorm_apply, except that
 * p->stats is IGNORED: the per-(image, group) sums are kept in a fixed array inside the library,
 * B200SD_GN_NOALLOC_MAX_STATS 4-byte words (64 KiB) that every GPU which loads the library holds until the process
 * exits.
 *   - limit: nimg * groups * B200SD_GN_STATS_WORDS <= B200SD_GN_NOALLOC_MAX_STATS (nimg * groups <= 2048), otherwise B200SD_ERR_TOO_LARGE is returned and nothing
 *     is launched (use the two-call form with your own scratch);
 *   - the array is shared by all calls on a device and still holds the previous call's sums, so each call first
 *     clears the part it uses (a cudaMemsetAsync on `stream`); two calls whose GPU work overlaps (different streams,
 *     or several host threads) race on it — serialise them or use the two-call form;
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
 *     current); make that first call before starting any stream capture;
 *   - no cudaMalloc / cudaFree or other allocating routine is called on any path.
 */
#define B200SD_GN_NOALLOC_MAX_STATS 16384
int b200sd_groupnorm_noalloc(const b200sd_groupnorm_params* p, b200sd_stream stream);

/* LayerNorm over the last dim. Replaces nn.LayerNorm in BasicTransformerBlock.norm1-3 / CLIPEncoderLayer (K9). */
typedef struct {
  const void* x; /* act16 [rows, c] */
  const float* gamma;
  const float* beta;
  void* out; /* act16 [rows, c] */
  int32_t rows, c;
  float eps;
} b200sd_layernorm_params;
int b200sd_layernorm(const b200sd_layernorm_params* p, b200sd_stream stream);

/* ---------------------------------------------------------------------------------------------------------
 * Denoise-loop glue (bandwidth-bound, fused).  A device-resident step table drives them so one captured
 * CUDA graph can be replayed for every step with no host interaction:
 *   step_table: f32 [nsteps, 16]  (see B200SD_ST_* indices), step_counter: int32 [2] on device
 *   ([0] = current step, [1] = internal CTA ticket, both zero-initialised by the caller).
 */
enum {
  B200SD_ST_TIMESTEP = 0,  /* timestep fed to the UNet */
  B200SD_ST_IN_SCALE = 1,  /* scale_model_input factor */
  B200SD_ST_KIND = 2,      /* 0 DDIM, 1 Euler, 2 PNDM/PLMS */
  B200SD_ST_C0 = 3,        /* DDIM: sqrt(1-a_t)   Euler: sigma_next - sigma    PNDM: sample coeff */
  B200SD_ST_C1 = 4,        /* DDIM: sqrt(a_t)                                  PNDM: -model-output coeff */
  B200SD_ST_C2 = 5,        /* DDIM: sqrt(a_prev) */
  B200SD_ST_C3 = 6,        /* DDIM: sqrt(1-a_prev) */
  B200SD_ST_W0 = 7,        /* PNDM: weights on eps history [newest .. oldest], 4 entries (7..10) */
  B200SD_ST_PNDM_MODE = 11,/* PNDM: 0 normal, 1 first step (store cur_sample), 2 second step (average, restore) */
  B200SD_ST_PUSH = 12,     /* PNDM: 1 = push eps into history ring */
  B200SD_ST_GUIDANCE = 13, /* guidance scale of this call; 0 = use b200sd_sched_params.guidance_scale (so one captured
                              graph serves every guidance value) */
  B200SD_ST_WIDTH = 16
};

/* latents f32 NHWC [b, hw, 4] -> UNet input act16 [2b (or b), hw, 8] (channels 4..7 zero), scaled by IN_SCALE.
 * Replaces torch.cat([latents]*2) + scheduler.scale_model_input (pipeline loop). */
typedef struct {
  const float* latents;
  void* out;
  const float* step_table;
  const int32_t* step_counter;
  int32_t b, hw, cfg; /* cfg=1: duplicate batch [uncond; cond] */
} b200sd_prep_params;
int b200sd_prep_unet_input(const b200sd_prep_params* p, b200sd_stream stream);

/* Sinusoidal timestep embedding (flip_sin_to_cos, shift 0): out act16 [rows, dim] = [cos | sin]. (a10.1, K10) */
typedef struct {
  void* out;
  const float* step_table;
  const int32_t* step_counter;
  int32_t rows, dim;
} b200sd_temb_params;
int b200sd_timestep_embedding(const b200sd_temb_params* p, b200sd_stream stream);

/* Classifier-free-guidance combine + scheduler step, in place on latents; increments *step_counter.
 * Replaces eps.chunk(2) / e_u + gs*(e_c-e_u) and scheduler.step (DDIM / Euler / PNDM-PLMS) — a11, a12 (K11/K12).
 * eps f32 [2b (or b), hw, ld_eps] (first 4 columns used).  history: f32 [4, b, hw, 4] ring (PNDM), cur_sample
 * f32 [b, hw, 4] (PNDM). */
typedef struct {
  const float* eps;
  float* latents;
  float* history;
  float* cur_sample;
  const float* step_table;
  int32_t* step_counter;
  int32_t b, hw, ld_eps, cfg;
  float guidance_scale;
} b200sd_sched_params;
int b200sd_cfg_scheduler_step(const b200sd_sched_params* p, b200sd_stream stream);

/* ---------------------------------------------------------------------------------------------------------
 * Layout / gather helpers (bandwidth-bound).
 */
/* CLIP token + position embedding: out act16 [rows, c] = tok[ids[r]] + pos[r % seq]. (CLIPTextEmbeddings; a7) */
typedef struct {
  const int32_t* ids;
  const void* tok; /* act16 [vocab, c] */
  const void* pos; /* act16 [seq, c] */
  void* out;
  int32_t rows, seq, c, vocab;
} b200sd_embed_params;
int b200sd_clip_embed(const b200sd_embed_params* p, b200sd_stream stream);

/* VAE output -> uint8 image: clamp(x/2+0.5, 0, 1)*255, round-half-even. x f32 [npix, ldx] (3 used) -> u8 [npix, 3].
 * Replaces VaeImageProcessor.postprocess + numpy round/astype (a15, K17). */
typedef struct {
  const float* x;
  uint8_t* out;
  int64_t npix;
  int32_t ldx;
} b200sd_denorm_params;
int b200sd_denorm_u8(const b200sd_denorm_params* p, b200sd_stream stream);

/* latents f32 NHWC [b, hw, 4] * scale -> act16 [b, hw, 8] zero padded (VAE input: latents / scaling_factor). */
typedef struct {
  const float* latents;
  void* out;
  int32_t b, hw;
  float scale;
} b200sd_latent_cast_params;
int b200sd_latent_to_act(const b200sd_latent_cast_params* p, b200sd_stream stream);

/* Row softmax: f32 logits [rows, cols] -> act16 probabilities [rows, cols] (VAE mid-block attention, head dim 512,
 * unfused path: the logits stay fp32 until after the exponent). */
typedef struct {
  const float* x;
  void* out;
  int32_t rows, cols;
} b200sd_softmax_params;
int b200sd_softmax_rows(const b200sd_softmax_params* p, b200sd_stream stream);

/* cudaMemsetAsync(ptr, 0, bytes) on `stream` — zeroes caller-owned scratch (GroupNorm statistics) inside a
 * captured step without leaving the C ABI. */
typedef struct {
  void* ptr;
  int64_t bytes;
} b200sd_memset_params;
int b200sd_memset_zero(const b200sd_memset_params* p, b200sd_stream stream);

/* sizeof() of every params struct, in declaration order above (host bindings check their mirrors against it). */
int b200sd_struct_sizes(int32_t* out, int32_t max_entries);

#ifdef __cplusplus
}
#endif
#endif /* B200SD_H_ */
