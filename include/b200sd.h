/* libb200sd -- C ABI of the B200-native (sm_100a) Stable Diffusion 2.1 denoising hot path.
 *
 * The reference (vin2803/scalable-hw-agnostic-inference) has no FFI and no native code: its hot path is
 * one Python call, `pipe(**model_args)` (app/run-sd.py:140), into un-vendored diffusers, and its plugin
 * seam is attribute replacement on the pipeline object (app/run-sd.py:108-134).  Every entry point below
 * therefore cites the diffusers module / torch op it stands in for and the run-sd.py line through which
 * the reference reaches it.  The Python host (b200sd/ops.py) binds these with ctypes and registers them
 * as torch.library ops in namespace `b200sd::`; INTEGRATION.md shows the reference-side binding.
 *
 * Conventions
 *   - C linkage, plain pointers and sizes; no torch / C++ types cross this boundary.
 *   - every compute entry point is  int32_t fn(const params*, b200sd_stream_t)  -> 0 or a negative
 *     b200sd_status; b200sd_last_error() gives the thread-local message.
 *   - entry points never allocate or free device memory, never synchronise the device, never throw.
 *     All buffers (including scratch, see *_workspace_bytes) are caller-owned device memory that must
 *     stay alive until `stream` reaches the op.
 *   - activations are NHWC; "tokens" [n_img, h*w, C] are the same memory as NHWC [n_img, h, w, C].
 *   - tensor-core operands are 16-bit (bf16 or fp16, chosen per tensor by a dtype code); accumulation,
 *     statistics, softmax and scheduler state are fp32.
 */
#ifndef B200SD_H_
#define B200SD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SD_ABI_VERSION 8

typedef struct CUstream_st* b200sd_stream_t; /* == cudaStream_t / CUstream */

typedef enum {
    B200SD_OK = 0,
    B200SD_ERR_BAD_ARG = -1,
    B200SD_ERR_UNSUPPORTED = -2,
    B200SD_ERR_DEVICE = -3,
    B200SD_ERR_WORKSPACE = -4,
    B200SD_ERR_CUDA = -5
} b200sd_status;

typedef enum { B200SD_BF16 = 1, B200SD_F32 = 2, B200SD_F16 = 3 } b200sd_dtype;

/* ---- library / device ------------------------------------------------------------------------ */
/* no reference counterpart (the reference never checks the device: app/run-sd.py:107 hard-codes "cuda") */
int32_t b200sd_abi_version(void);
const char* b200sd_last_error(void);
/* 0 iff device `dev` is compute capability 10.x (sm_100a cubins load there) */
int32_t b200sd_check_device(int32_t dev);
/* number of kernel launches issued by this library in this process (bench.py's gpu_launches) */
int64_t b200sd_launch_count(void);

/* ---- K1/K2: implicit-GEMM conv3x3 / conv1x1 / linear on tcgen05 + TMEM, fed by TMA --------------
 * Replaces torch.nn.Conv2d (3x3 pad 1 stride 1, and 1x1) and torch.nn.Linear inside diffusers'
 * ResnetBlock2D / Transformer2DModel / Attention / FeedForward / Decoder, which the reference reaches
 * through pipe.unet(...) and pipe.vae.decode(...) (app/run-sd.py:110-116,124-134,140).
 *   out[m, n] = epilogue( sum_{tap,c} A[pixel(m)+offset(tap), c] * W[n, tap*cin + c] )
 * A is an NHWC 16-bit activation; a plain [M,K] matrix is n_img=1, img_h=1, img_w=M, cin=K, taps=1.
 * epilogue: + bias[n]  + row_bias[img(m)*row_bias_stride + n]  + residual[m, n]  then optional GEGLU.
 */
typedef struct {
    const void* a;          /* NHWC [n_img, img_h, img_w, cin] of a_dtype; 16-byte aligned; cin % 8 == 0 */
    int32_t a_dtype;        /* B200SD_BF16 or B200SD_F16 */
    int32_t n_img, img_h, img_w, cin;
    int32_t taps;           /* 1 or 9 */
    const void* w;          /* [n_out, taps*cin] of w_dtype, K contiguous, K order = (tap, c)  (KRSC) */
    int32_t w_dtype;        /* B200SD_BF16 or B200SD_F16 */
    int32_t n_out;
    const float* bias;      /* fp32 [n_out] (or [M] when bias_along_m), may be NULL */
    int32_t bias_along_m;
    const float* row_bias;  /* fp32, per-image additive row (timestep-embedding projection), may be NULL */
    int64_t row_bias_stride;/* elements between consecutive images; 0 broadcasts one row */
    const void* residual;   /* [M, ld_res] of residual_dtype, may be NULL; may alias `out` */
    int32_t residual_dtype; /* any b200sd_dtype */
    int64_t ld_res;
    void* out;              /* [M, ld_out] of out_dtype (GEGLU: n_out/2 columns are written) */
    int32_t out_dtype;      /* any b200sd_dtype */
    int64_t ld_out;
    int32_t geglu;          /* 1: out[m,j] = val_j * gelu_erf(gate_j); weight/bias rows interleaved per block_n
                               tile by the host packer, b200sd/layers.py pack_geglu (first half of a tile = values,
                               second half = gates) */
    int32_t block_n;        /* 0 = choose; else one of 64, 128, 160, 256 */
    void* workspace;        /* optional split-K scratch, 256-byte aligned, ZERO-FILLED once by the caller (the kernel
                               leaves its counters at zero again); NULL = never split K */
    int64_t workspace_bytes;
    int32_t split_k;        /* 0 = choose (splits only when the problem has fewer tiles than SMs and a workspace is
                               given), 1 = off, 2..8 = forced */
    int32_t act;            /* 0 = none, 1 = erf-GELU applied to (acc + bias) before the residual is added: the CLIP text
                               encoder's MLP (transformers CLIPMLP, hidden_act "gelu").  Not with geglu. */
    float* tile_sums;       /* optional: the epilogue also writes, per 32-row slice of every 128-row output tile and per
                               output column, (sum, sum of squares) of the values it stores (fp32, before the 16-bit
                               rounding): [n_img][parts][n_out][2] with parts = b200sd_gemm_tile_sum_parts(p).  This is
                               torch.nn.GroupNorm's statistics pass for the NEXT GroupNorm, taken while the tensor is
                               still in registers (b200sd_gn_tile_sums reduces it); NULL = off */
    int32_t rows_per_image; /* only with tile_sums on a plain [M, K] problem (img_h == 1, n_img == 1): rows of `out`
                               that belong to one image; must be a multiple of 128.  0 for NHWC problems. */
    int32_t conv_stride;    /* taps == 9 only.  0 or 1: stride 1.  2: Downsample2D's stride-2 convolution, read straight from
                               the NHWC input by a TMA box that takes every second pixel (no im2col copy).  img_h / img_w
                               stay the INPUT size; `out`, `residual`, `tile_sums` have the output size
                               out_h = (img_h + pad_before + 1 - 3) / stride + 1 (likewise out_w). */
    int32_t conv_pad_mode;  /* 0: Conv2d(padding=1), one zero row / column before and after.  1: nothing before, one after
                               -- the VAE encoder's Downsample2D(padding=0): F.pad(x, (0,1,0,1)) then no padding. */
} b200sd_gemm_params;
int32_t b200sd_gemm(const b200sd_gemm_params* p, b200sd_stream_t stream);
/* slices per image that `tile_sums` would hold for this problem, or 0 if this problem cannot emit them (a tile that
 * spans two images, GEGLU, split K, bias along M, a ragged last column chunk ...): the caller then runs the
 * GroupNorm statistics pass as usual */
int32_t b200sd_gemm_tile_sum_parts(const b200sd_gemm_params* p);
/* bytes of workspace this problem would use if one were supplied (0 = it would not split K) */
int64_t b200sd_gemm_workspace_bytes(const b200sd_gemm_params* p);
/* the block_n the library would choose for this problem (weight packers for GEGLU need it) */
int32_t b200sd_gemm_pick_block_n(int32_t n_out, int32_t geglu);

/* ---- K3: fused flash attention, head_dim 64, tcgen05 QK^T / PV, online softmax in registers -----
 * Replaces diffusers Attention (attn1 self-attention and attn2 cross-attention of BasicTransformerBlock),
 * which the reference runs through SlicedAttnProcessor because of pipe.enable_attention_slicing()
 * (app/run-sd.py:135).  Row r of batch b is q[(b*sq + r)*ld_q + h*64 + d]; same for k/v with skv.
 */
typedef struct {
    const void *q, *k, *v;  /* all of `dtype` */
    int64_t ld_q, ld_k, ld_v;
    void* out;              /* `dtype` [batch*sq, ld_out], head h at columns h*64 */
    int64_t ld_out;
    int32_t dtype;          /* B200SD_BF16 or B200SD_F16 (also the format of the probability tile) */
    int32_t batch, heads, head_dim; /* head_dim must be 64 */
    int32_t sq, skv;
    float scale;            /* softmax(scale * q.k) */
    int32_t causal;         /* 1 = key j is visible to query i only if j <= i (text encoder); needs sq == skv */
} b200sd_attention_params;
int32_t b200sd_attention(const b200sd_attention_params* p, b200sd_stream_t stream);
/* Same parameters for ONE head of dimension 512 (heads == 1, head_dim == 512, causal == 0): the mid-block attention of
 * AutoencoderKL's decoder and encoder (diffusers Attention inside UNetMidBlock2D, reached through pipe.vae.decode at
 * app/run-sd.py:140).  Flash form as well: the score matrix never leaves the SM.  Each CTA computes the scores of its
 * 128 query rows against every key over all 512 dims and one 256-column half of the output. */
int32_t b200sd_attention512(const b200sd_attention_params* p, b200sd_stream_t stream);

/* ---- K1b: GroupNorm(32) (+SiLU) over NHWC, optional channel-concat of two inputs ------------------
 * Replaces torch.nn.GroupNorm + F.silu in ResnetBlock2D.norm1/norm2, Transformer2DModel.norm,
 * conv_norm_out, and the torch.cat([h, skip], dim=1) feeding up-block resnets (via pipe.unet /
 * pipe.vae.decoder, app/run-sd.py:110-116,124-128).  Two kernels: deterministic split statistics into
 * `workspace`, then normalise+affine(+SiLU) -> 16-bit.  `raw` (optional) receives the un-normalised 16-bit
 * concat, the A operand of the 1x1 conv_shortcut.
 */
typedef struct {
    const void* x1;         /* [n_img, hw, c1] of in_dtype */
    int32_t c1;
    const void* x2;         /* [n_img, hw, c2] or NULL */
    int32_t c2;
    int32_t in_dtype;       /* any b200sd_dtype */
    int32_t n_img, hw, groups;
    float eps;
    const float *gamma, *beta; /* fp32 [c1+c2] */
    int32_t silu;
    void* y;                /* out_dtype [n_img, hw, c1+c2] */
    void* raw;              /* out_dtype [n_img, hw, c1+c2] or NULL */
    int32_t out_dtype;      /* B200SD_BF16 or B200SD_F16 */
    float* workspace;
    int64_t workspace_bytes;
    const float* sums;      /* optional fp32 [n_img, groups, 2] = (sum, sum of squares) per image and group: when given,
                               the statistics pass is skipped and these are normalised with (no workspace needed) */
    int64_t stat_hw;        /* with `sums`: pixels per image the sums were taken over (0 = hw).  Larger than hw when the
                               image is split over several GPUs and the sums were all-reduced */
    float* sums_out;        /* optional fp32 [n_img, groups, 2]: when given, ONLY the statistics are computed and
                               written here (y may be NULL) -- the first half of a GroupNorm whose sums are combined
                               across GPUs before b200sd_groupnorm is called again with `sums` */
} b200sd_groupnorm_params;
int64_t b200sd_groupnorm_workspace_bytes(const b200sd_groupnorm_params* p);
int32_t b200sd_groupnorm(const b200sd_groupnorm_params* p, b200sd_stream_t stream);
/* Reduces the per-tile column sums a producing GEMM left in `tile_sums` (see b200sd_gemm_params) to per-(image, group)
 * sums for b200sd_groupnorm's `sums`.  The normalised tensor may be the channel concat of two producers' outputs
 * (torch.cat([h, skip], dim=1) in the up blocks): groups then run over the concatenation.  Fixed reduction order. */
typedef struct {
    const float* tile_sums1; /* [n_img, parts1, c1, 2] */
    int32_t parts1, c1;
    const float* tile_sums2; /* [n_img, parts2, c2, 2] or NULL */
    int32_t parts2, c2;
    int32_t n_img, groups;
    float* sums;             /* out: fp32 [n_img, groups, 2] */
} b200sd_gn_tile_sums_params;
int32_t b200sd_gn_tile_sums(const b200sd_gn_tile_sums_params* p, b200sd_stream_t stream);
/* GroupNorm for hosts that cannot supply scratch memory -- it has no counterpart in the reference.
 * Who needs it: an FFI front end (cgo, JNI, ctypes) that is handed device pointers for its tensors by some other
 * component and links no CUDA runtime of its own, so it has nothing to allocate a workspace with; and a server whose
 * whole device budget went to a tensor arena at start-up, leaving no memory to carve a workspace from.
 * Same params and result as b200sd_groupnorm, but `workspace` / `workspace_bytes` are ignored: the split statistics
 * go to a 256 KiB array built into the library (held by every GPU that loads it, until process exit).
 *   - allocates nothing (no cudaMalloc / cudaFree or allocating library call);
 *   - limit: the statistics (n_img * splits * groups * 2 floats) must fit the array's 65536 floats.  The entry point
 *     lowers its split count, down to one per image, until they do (a large batch then runs its statistics pass in
 *     fewer, larger CTAs: slower than b200sd_groupnorm, same result up to the summation order); only when
 *     n_img * groups * 2 > 65536 does it return B200SD_ERR_WORKSPACE, and the message says to use b200sd_groupnorm;
 *   - the first call on a device looks up the array's device address; issue that first call before any stream
 *     capture on that device;
 *   - every call on a device shares the array, which keeps the previous call's contents; nothing needs clearing
 *     (the statistics kernel overwrites every slot that is read), but calls whose kernels overlap on the GPU race
 *     on it -- keep them ordered on one stream or by events. */
int32_t b200sd_groupnorm_static_scratch(const b200sd_groupnorm_params* p, b200sd_stream_t stream);

/* ---- LayerNorm over the last dim -> 16-bit (BasicTransformerBlock.norm1/2/3) ---------------------- */
typedef struct {
    const void* x;          /* [rows, cols] of in_dtype, contiguous */
    int32_t in_dtype;
    int64_t rows;
    int32_t cols;
    float eps;
    const float *gamma, *beta;
    void* y;                /* out_dtype [rows, cols] */
    int32_t out_dtype;      /* B200SD_BF16 or B200SD_F16 */
} b200sd_layernorm_params;
int32_t b200sd_layernorm(const b200sd_layernorm_params* p, b200sd_stream_t stream);

/* ---- im2col for the convs TMA cannot tile: stride-2 Downsample2D and the 4-channel conv_in ---------
 * out[m, (r*3+s)*cin + c] = x[(n % n_src), oy*stride + r - pad_lo, ox*stride + s - pad_lo, c]
 * (0 outside, 0 for k >= 9*cin); out_h = (in_h + pad_lo + 1 - 3) / stride + 1, likewise out_w.
 * pad_lo = 1: Conv2d(padding=1).  pad_lo = 0: the VAE encoder's Downsample2D(padding=0), which pads (0,1,0,1) --
 * nothing before, one row / column after -- and then convolves with stride 2 (AutoencoderKL.encode).
 * n % n_src lets the CFG pair (uncond, cond) share one latent: n_img = 2B, n_src = B.
 */
typedef struct {
    const void* x;          /* NHWC [n_src, in_h, in_w, cin] of in_dtype */
    int32_t in_dtype;
    int32_t n_img, n_src, in_h, in_w, cin, stride;
    void* out;              /* out_dtype [n_img*out_h*out_w, k_pad], k_pad >= 9*cin, k_pad % 8 == 0 */
    int32_t out_dtype;      /* B200SD_BF16 or B200SD_F16 */
    int32_t k_pad;
    int32_t pad_lo;         /* 1 or 0 (see above) */
} b200sd_im2col_params;
int32_t b200sd_im2col3x3(const b200sd_im2col_params* p, b200sd_stream_t stream);

/* ---- nearest-neighbour 2x upsample NHWC -> 16-bit (Upsample2D's F.interpolate) ---------------------- */
typedef struct {
    const void* x;          /* [n_img, h, w, c] of in_dtype */
    int32_t in_dtype;
    int32_t n_img, h, w, c;
    void* out;              /* out_dtype [n_img, 2h, 2w, c] */
    int32_t out_dtype;      /* B200SD_BF16 or B200SD_F16 */
} b200sd_upsample_params;
int32_t b200sd_upsample2x(const b200sd_upsample_params* p, b200sd_stream_t stream);

/* ---- K4: classifier-free guidance + scheduler step, one pass -----------------------------------------
 * Replaces, per denoising step, `noise_pred_uncond + g*(noise_pred_text - noise_pred_uncond)` and
 * DDIMScheduler.step (forced at app/run-sd.py:108) / DPMSolverMultistep / EulerDiscrete (BASELINE.json)
 * inside StableDiffusionPipeline.__call__ (app/run-sd.py:140):
 *   m = u + g (c - u);  x0 = ax x + am m;  x' = cx x + cm m + cp x0_prev
 * coefficients: fp32[6] = {g, cx, cm, cp, ax, am} read from DEVICE memory at coef[step_index*6] so that a
 * CUDA graph of the step can be replayed for every timestep.  eps2 = [2B, ...] (uncond first), any layout
 * shared by all operands.  x is updated in place; x0_prev (optional) is read then overwritten with x0.
 */
typedef struct {
    const void* eps2;       /* [2*n] of eps_dtype: first n uncond, last n cond */
    int32_t eps_dtype;      /* any b200sd_dtype */
    float* x;               /* fp32 [n], in place */
    float* x0_prev;         /* fp32 [n] or NULL */
    const float* coef;      /* device fp32 [6 * n_steps] */
    const int32_t* step_index; /* device int32 scalar, or NULL for step 0 */
    int64_t n;
} b200sd_cfg_sched_params;
int32_t b200sd_cfg_sched_step(const b200sd_cfg_sched_params* p, b200sd_stream_t stream);

/* ---- K5: small-M linear for the timestep path -----------------------------------------------------------
 * y[m, n] = act_out( sum_k act_in(x[m,k]) * w[n,k] + b[n] ),  m <= 64.  Used for Timesteps (sinusoid, when
 * `sinusoid_dim` > 0 x is a device fp32 [rows] vector of timesteps), TimestepEmbedding.linear_1/2 and the 22
 * ResnetBlock2D.time_emb_proj (batched into one weight), all inside pipe.unet (app/run-sd.py:140).
 */
typedef struct {
    const float* x;         /* fp32 [rows, k]   (or fp32 [rows] timesteps when sinusoid_dim > 0) */
    int32_t rows, k, n;
    int32_t sinusoid_dim;   /* 0, or == k: x[m] -> cat(cos, sin)(x[m] * 10000^(-i/(k/2))) */
    int32_t silu_in, silu_out;
    const void* w;          /* [n, k] of w_dtype */
    int32_t w_dtype;        /* B200SD_BF16 or B200SD_F16 */
    const float* b;         /* fp32 [n] or NULL */
    float* y;               /* fp32 [rows, n] */
} b200sd_small_linear_params;
int32_t b200sd_small_linear(const b200sd_small_linear_params* p, b200sd_stream_t stream);

/* ---- row softmax y = softmax(scale * x) -> 16-bit (VAE mid-block attention, 1 head x 512) ---------------
 * y may alias x when both are 16-bit. */
typedef struct {
    const void* x;          /* [rows, ld_in] of in_dtype, first `cols` of each row are live */
    int32_t in_dtype;
    int64_t ld_in;
    void* y;                /* out_dtype [rows, ld_out] */
    int32_t out_dtype;      /* B200SD_BF16 or B200SD_F16 */
    int64_t ld_out;
    int64_t rows;
    int32_t cols;
    float scale;
} b200sd_softmax_params;
int32_t b200sd_softmax_rows(const b200sd_softmax_params* p, b200sd_stream_t stream);

/* ---- text-encoder embeddings: out[b, t, :] = token_table[ids[b, t], :] + position_table[t, :] ---------------------
 * (transformers CLIPTextEmbeddings; reached from the reference through pipe.text_encoder, app/run-sd.py:118-122).
 * ids int32 [rows] with rows = batch * seq_len; tables fp32; out fp32 [rows, dim]; dim % 4 == 0; an id outside
 * [0, vocab) is an error the kernel cannot report, so the caller clamps or validates. */
typedef struct {
    const int32_t* ids;
    const float* token_table;     /* [vocab, dim] */
    const float* position_table;  /* [seq_len, dim] */
    float* out;
    int64_t rows;
    int32_t seq_len, dim, vocab;
} b200sd_embed_params;
int32_t b200sd_embed_tokens(const b200sd_embed_params* p, b200sd_stream_t stream);

/* ---- K7: VaeImageProcessor.postprocess on device: (x/2+0.5).clamp(0,1)*255 round -> uint8 -------------
 * (the reference does this on the CPU after a fp32 D2H copy; app/run-sd.py:140 `.images[0]`, :177-181) */
typedef struct {
    const void* x;          /* [n] of in_dtype */
    int32_t in_dtype;
    uint8_t* out;           /* [n] */
    int64_t n;
} b200sd_denorm_u8_params;
int32_t b200sd_denorm_u8(const b200sd_denorm_u8_params* p, b200sd_stream_t stream);

/* ---- dtype cast between fp32 / bf16 / fp16 (boundary tensors) --------------------------------------------- */
typedef struct {
    const void* x;
    int32_t in_dtype;
    void* y;
    int32_t out_dtype;
    int64_t n;
} b200sd_cast_params;
int32_t b200sd_cast(const b200sd_cast_params* p, b200sd_stream_t stream);

/* ---- UMMA self-test: one 128 x n x k tile product with the B operand K-major or MN-major -----------------
 * no reference counterpart; used by tests to pin the shared-memory / instruction descriptor encodings
 * (including mixed bf16 x fp16 operands) on real hardware. */
int32_t b200sd_umma_selftest(const void* a /*[128,k] of a_dtype*/, int32_t a_dtype,
                             const void* b /*[n,k] or [k,n] of b_dtype*/, int32_t b_dtype,
                             float* d /*[128,n]*/, int32_t n, int32_t k, int32_t b_mn_major,
                             b200sd_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SD_H_ */
