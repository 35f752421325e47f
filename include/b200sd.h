/*
 * b200sd.h — C-ABI of the B200 (sm_100a) DDPM denoising hot path.
 *
 * The reference (wangze22/StableDiffusion-PyTorch) has no FFI: its boundary is the Python module surface
 * (SURVEY.md §8b). Each entry point below replaces the library kernels that one reference call site launches
 * through PyTorch; the reference file:line is cited per function. Host code (models/, scheduler/, cim_layers/)
 * binds these with ctypes and passes raw device pointers + sizes. No torch types cross
 * this boundary.
 *
 * Conventions
 *   - every function returns 0 on success or a negative B200SD_ERR_* code; b200sd_last_error() has the text
 *   - the caller owns every buffer; the library never allocates or frees device memory and keeps no pointer
 *     past return (it caches only host-side TMA descriptors keyed by pointer+shape)
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises, so every
 *     call is CUDA-graph capturable
 *   - activations are NHWC ("pixel-major") bf16 unless stated: a (N,H,W,C) tensor is a row-major
 *     [N*H*W, C] matrix, which is also the (B, L, C) token layout attention uses
 */
#ifndef B200SD_H
#define B200SD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SD_OK 0
#define B200SD_ERR_ARG (-1)         /* bad shape / null pointer / misalignment */
#define B200SD_ERR_UNSUPPORTED (-2) /* valid request this build has no kernel for (never a CPU fallback) */
#define B200SD_ERR_CUDA (-3)        /* CUDA runtime / driver error; text in b200sd_last_error() */

#define B200SD_ACT_NONE 0
#define B200SD_ACT_SILU 1
#define B200SD_ACT_RELU 2

#define B200SD_DT_BF16 0
#define B200SD_DT_F32 1

const char* b200sd_last_error(void);
int b200sd_version(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches counter) */
int64_t b200sd_launch_count(void);

/* ------------------------------------------------------------------------------------------------ scheduler
 * LinearNoiseScheduler.add_noise — scheduler/linear_noise_scheduler.py:26-48.
 * out[b,i] = sqrt_acp[t[b]] * x0[b,i] + sqrt_1m_acp[t[b]] * noise[b,i]; fp32, products and sum individually
 * rounded (no FMA contraction) so the result is bit-identical to the reference's separate torch ops.
 * t: int64 on device, shape (batch,), or (1,) when t_len == 1 (broadcast). Tables: fp32 on device, length T. */
int b200sd_sched_add_noise(const float* x0, const float* noise, const int64_t* t, int64_t t_len,
                           const float* sqrt_acp, const float* sqrt_1m_acp, int num_timesteps, float* out,
                           int64_t batch, int64_t per_sample, void* stream);

/* LinearNoiseScheduler.sample_prev_timestep — scheduler/linear_noise_scheduler.py:50-78.
 * x0 = clamp((xt - sqrt_1m_acp[t]*eps)/sqrt(acp[t]), -1, 1); mean = (xt - betas[t]*eps/sqrt_1m_acp[t])/sqrt(alphas[t]);
 * x_prev = mean (t == 0) or mean + sqrt((1-acp[t-1])/(1-acp[t])*betas[t]) * z.
 * t is one timestep shared by the batch: read from t_dev (int64 on device) when non-null, else t_host.
 * z: injected N(0,1) noise (parity path) or NULL to draw it in-kernel with Philox4x32-10 (seed, offset).
 * rng_state (optional): device uint64[2] = {seed, call index}; when non-null the Philox key is
 * seed + 0x9E3779B97F4A7C15 * call index read from DEVICE memory (the `seed` argument is ignored), so a CUDA graph
 * that froze its launch arguments still draws a different noise trajectory per sampling call: the host bumps
 * rng_state[1] between calls. A device-side t outside [0, T) is clamped (the host-side t is range-checked).
 * Optional classifier-free guidance fused in (tools/sample_ddpm_text_image_cond.py:84-86):
 * eps = eps_uncond + cfg_scale*(eps - eps_uncond) when eps_uncond is non-null. */
int b200sd_sched_prev_step(const float* xt, const float* eps, const float* eps_uncond, float cfg_scale,
                           const float* z, const int64_t* t_dev, int64_t t_host, const float* betas,
                           const float* alphas, const float* acp, const float* sqrt_1m_acp, int num_timesteps,
                           float* x_prev, float* x0, int64_t n, uint64_t seed, uint64_t offset,
                           const uint64_t* rng_state, void* stream);

/* t[i] += delta on device: the `for i in reversed(range(T))` loop counter of tools/sample_ddpm_text_image_cond.py:79-81
 * kept on the device so a whole sampling step (UNet forward + ancestral update + counter) replays as one CUDA graph. */
int b200sd_advance_t(int64_t* t, int64_t t_len, int64_t delta, void* stream);

/* One update of DDPMSampler.sample_one_step (scheduler/linear_noise_scheduler.py:106-131: c1 = coeff_1[t],
 * c2 = -coeff_2[t], sigma = sqrt(posterior_variance[t])) or DDIMSampler.sample_one_step (:185-204: c1 =
 * sqrt(a_prev / a_t), c2 = sqrt(1 - a_prev - sigma^2) - sqrt(a_prev (1 - a_t) / a_t)), coefficients computed by the
 * caller: out = (c1 * x + c2 * eps) + sigma * z, each operation rounded separately in that order. z may be NULL
 * (term dropped). out may alias x. fp32, n elements. */
int b200sd_sampler_step(const float* x, const float* eps, const float* z, float c1, float c2, float sigma, float* out,
                        int64_t n, void* stream);

/* get_time_embedding — models/blocks.py:5-24 (and models/transformer.py:18-40).
 * out[b, j] = sin(t[b]/f_j), out[b, half+j] = cos(t[b]/f_j), f_j = 10000^(j/half); t int64 (t_len = batch or 1).
 * out is bf16 or fp32 [batch, dim]. */
int b200sd_time_embedding(const int64_t* t, int64_t t_len, void* out, int out_dtype, int64_t batch, int dim,
                          void* stream);

/* ------------------------------------------------------------------------------------ implicit-GEMM core
 * One tcgen05/TMEM kernel serves nn.Conv2d 3x3/1x1/4x4-s2, nn.ConvTranspose2d 4x4-s2 (as 4 sub-pixel phases)
 * and nn.Linear (models/blocks.py:45-109,339,457; models/unet_cond_base.py:67-83,122).
 *
 *   acc[p, n]  = sum over segments s, channels c:  A_s[pixel p shifted by (dy_s, dx_s), c0_s + c] * Wt[n, k(s, c)]
 *   v          = acc + bias[n] + rowbias[img(p), n]          (each optional)
 *   v          = act(v)
 *   v          = gate[img(p), n] * v  (optional)  + residual[p, n] (optional)
 *   out[p, n]  = v   (bf16 or fp32)
 *
 * A sources are 4-D views (C, W, H, N) of NHWC bf16 tensors given as base pointer + byte strides, read by TMA
 * with out-of-bounds zero fill (this is the conv padding). The weight is bf16 [n_out, k_total] row-major
 * (K contiguous) with k running over segments in order, 64-channel chunk by chunk. */
typedef struct {
    const void* base;  /* bf16 */
    int32_t dims[4];   /* C, W, H, N extents */
    int64_t strides[3]; /* byte strides of W, H, N (C is contiguous) */
} b200sd_src;

typedef struct {
    int32_t src;    /* index into srcs[] */
    int32_t dy, dx; /* tap offset added to the output pixel coordinate */
    int32_t c0;     /* first channel read from the source */
    int32_t nchunk; /* number of 64-channel chunks (source channels beyond dims[0] read as zero) */
} b200sd_seg;

typedef struct {
    int32_t n_src, n_seg;
    b200sd_src srcs[4];
    b200sd_seg segs[40];
    const void* weight; /* bf16 [n_out, k_total], k_total = 64 * sum(nchunk) */
    int32_t n_out;
    /* output pixel grid and its addressing (element strides; out/residual share geometry, not pointers) */
    int32_t on, oh, ow;
    void* out;
    int32_t out_dtype;
    int64_t out_sn, out_sy, out_sx; /* element strides of image, row, column in `out` */
    const float* bias;              /* fp32 [n_out] or NULL */
    const void* rowbias;            /* bf16 [on, rowbias_ld] or NULL (time-embedding add, blocks.py:118) */
    int64_t rowbias_ld;
    const void* gate; /* bf16 [on, gate_ld] or NULL (adaLN-Zero gate, transformer_layer.py:90) */
    int64_t gate_ld;
    const void* residual; /* bf16, same pixel geometry as out, or NULL */
    int64_t res_sn, res_sy, res_sx;
    int32_t act;
    /* optional partial GroupNorm statistics of the bf16-rounded output: fp32 [on, stats_p, stats_ld, 2] holding, per
       image and per tile slot, each channel's (sum, sum of squares) over the tile's pixels. This call writes slots
       [stats_slot0, stats_slot0 + b200sd_igemm_stats_slots(oh, ow)) at channels [0, n_out); every element it owns
       is written exactly once (no atomics, no zero-initialisation needed, bit-reproducible). The consumer
       (b200sd_groupnorm_apply) sums the slots in order. */
    float* stats;
    int32_t stats_p, stats_slot0, stats_ld;
    /* precise mode: an fp32 output may take its residual and / or its per-image row bias as fp32 too
       (B200SD_DT_F32; 0 = bf16 as above). With residual == out the call accumulates in place (each thread reads the
       element it then writes): the second launch of a split-operand product A_hi*[W_hi|W_lo] + A_lo*W_hi. */
    int32_t residual_dtype, rowbias_dtype;
    /* split-K for problems with too few output tiles to fill the GPU (the 4x4 / 8x8 layers of a batch-16 training
       step: 12-96 tiles on 148 SMs, each CTA bound by its own L2 read bandwidth over a long reduction). With
       splitk_max > 1 the library may cut the K loop of every tile into up to splitk_max ranges run by different CTAs;
       they park fp32 partials in splitk_ws and the last one to arrive sums them in split order (deterministic) and
       runs the epilogue. The split depth depends on the tile count, hence on the batch: leave splitk_max at 0 where a
       sample's result must not depend on the batch it travels in (inference). splitk_ws: >= splitk_ws_floats floats,
       16-byte aligned; splitk_counters: splitk_counters_len zero-initialised uint32 (returned to zero). Calls that may
       run concurrently must not share either buffer. */
    float* splitk_ws;
    int64_t splitk_ws_floats;
    uint32_t* splitk_counters;
    int32_t splitk_counters_len;
    int32_t splitk_max;
} b200sd_igemm_params;

int b200sd_igemm(const b200sd_igemm_params* p, void* stream);
int b200sd_igemm_stats_slots(int oh, int ow);
/* split-K depth the most recent b200sd_igemm call of this process chose (1 = none; pair kernels always 1): tests, tools */
int b200sd_igemm_last_splits(void);

/* ------------------------------------------------------------------------------------------- weight gradient
 * Backward of the same leaves w.r.t. their weights (loss.backward(), train_ddpm_cond_celebhq_multi_gpu.py:362):
 *   dWp[co, k(s, c)] = sum over output pixels p: dY[p, co] * A_s[p shifted by (dy_s, dx_s), c0_s + c]
 * in the packed [n_out, k_total] layout b200sd_igemm reads (srcs/segs exactly as in the forward call). dY is a
 * (C = n_out, W, H, N) bf16 view like the sources. The pixel reduction is split into `splits` ranges (use
 * b200sd_wgrad_splits); each range stores fp32 partial[split][n_out][k_total] and b200sd_wgrad_reduce sums them in
 * order into `out` (packed fp32; accumulate != 0 adds to what is there). No atomics: bit-reproducible. */
typedef struct {
    int32_t n_src, n_seg;
    b200sd_src srcs[4];
    b200sd_seg segs[20];
    b200sd_src dy;
    int32_t n_out;
    int32_t on, oh, ow;
    float* partial;
    int32_t splits;
} b200sd_wgrad_params;
int b200sd_wgrad_splits(int n_out, int k_total, int on, int oh, int ow);
int b200sd_wgrad(const b200sd_wgrad_params* p, void* stream);
int b200sd_wgrad_reduce(const float* partial, int splits, int64_t elems, float* out, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------- normalisation
 * nn.GroupNorm(G, C) (+ SiLU / ReLU) — models/blocks.py:48-49,67-68,126,136; unet_cond_base.py:179-180.
 * x: NHWC bf16 [n, hw, ld_x] (C contiguous); statistics over (hw, C/G) per image and group, biased variance,
 * eps inside the sqrt, fp32 arithmetic; y = act((x-mean)*rstd*gamma + beta) written as bf16 [n, hw, ld_y]. */
int b200sd_groupnorm(const void* x, int64_t ld_x, void* y, int64_t ld_y, const float* gamma, const float* beta,
                     int64_t n, int64_t hw, int c, int groups, float eps, int act, void* stream);

/* Precise mode. nn.GroupNorm (+ activation) of an fp32 NHWC tensor x [n, hw, ld_x], statistics computed here in fp32
 * (two passes, no E[x^2] - E[x]^2 cancellation), result written as a SPLIT bf16 operand: y [n*hw, ld_y] holds
 * hi = bf16(v) in columns [0, c) and lo = bf16(v - hi) in columns [lo_off, lo_off + c) — together 16 mantissa bits, the
 * A operand of the split-operand GEMM (b200sd_igemm called on the hi and on the lo view). lo_off < 0: hi only. */
int b200sd_groupnorm_split(const float* x, int64_t ld_x, void* y, int64_t ld_y, int32_t lo_off, const float* gamma,
                           const float* beta, int64_t n, int64_t hw, int c, int groups, float eps, int act,
                           void* stream);
/* dst[r, hi_off + j] = hi, dst[r, lo_off + j] = lo of act(src[r, j]) for an fp32 [rows, c] matrix (pitch ld_s): any
 * fp32 tensor that feeds a GEMM directly in precise mode (block input of the 1x1 shortcut, down / up-sample input,
 * time embedding, text context). lo_off < 0: hi only (= the autocast cast). */
int b200sd_split_f32(const float* src, int64_t ld_s, void* dst, int64_t ld_d, int32_t hi_off, int32_t lo_off,
                     int64_t rows, int32_t c, int act, void* stream);

/* The same normalisation when the producer of x already left partial statistics (b200sd_igemm `stats`): x is read
 * once and written once. Channels [0, c_split) take their statistics from stats0 ([n, p0, ld_s0, 2]) and channels
 * [c_split, c) from stats1 ([n, p1, ld_s1, 2]; the skip-connection concat of models/blocks.py:464, whose halves were
 * produced by different kernels); pass c_split = c and stats1 = NULL for a single producer. */
int b200sd_groupnorm_apply(const void* x, int64_t ld_x, void* y, int64_t ld_y, const float* stats0, int p0, int ld_s0,
                           int c_split, const float* stats1, int p1, int ld_s1, const float* gamma,
                           const float* beta, int64_t n, int64_t hw, int c, int groups, float eps, int act,
                           void* stream);

/* Backward of GroupNorm (+activation), given the forward's partial statistics (same stats0/stats1 as
 * b200sd_groupnorm_apply), x, and dy = dL/d act(GN(x)):  dx = dL/dx (+ add, an optional extra gradient of x with the
 * same geometry, e.g. the ResNet shortcut's). part: fp32 scratch [n, b200sd_rowblocks(n, hw, c), c, 2]; ab (optional):
 * fp32 [n, c, 2] receiving per-image (sum dz, sum dz*xhat), whose sums over n are dbeta and dgamma.
 * counter (optional): a run of 64 zero-initialised uint32 in device memory, not shared with a concurrently running
 * launch (the one-pass route keeps one arrival counter per channel chunk, the two-pass route uses the first).
 * With it the last CTA to finish folds ab over the batch into dgamma[c] / dbeta[c] (fixed order) and leaves the
 * counter at 0 again — no second kernel. */
int b200sd_rowblocks(int64_t n, int64_t hw, int c);
/* How b200sd_groupnorm_bwd will run this shape: 1 = the one-pass cluster kernel, with out6 (optional) = {groups per
 * cluster, CTAs per cluster, rows per CTA, rows per pass, threads per CTA, dynamic shared memory in bytes}; 0 = the
 * two-pass kernels (the slab does not fit, or B200SD_GN_BWD_FUSED=0). Host-only: no device call. */
int b200sd_groupnorm_bwd_plan(int64_t n, int64_t hw, int c, int groups, int with_add, int32_t* out6);
int b200sd_groupnorm_bwd(const void* x, int64_t ld_x, const void* dy, int64_t ld_dy, const void* add, int64_t ld_add,
                         void* dx, int64_t ld_dx, const float* stats0, int p0, int ld_s0, int c_split,
                         const float* stats1, int p1, int ld_s1, const float* gamma, const float* beta, int64_t n,
                         int64_t hw, int c, int groups, float eps, int act, float* part, float* ab, float* dgamma,
                         float* dbeta, uint32_t* counter, void* stream);
/* Per-(image, row block) column sums of a bf16 [n, hw, ld_x] tensor: part fp32 [n, b200sd_colsum_blocks(n, hw, c), c]
 * (bias gradients; the per-image time-embedding gradient of models/blocks.py:118). With `counter` — here
 * ceil(c / 64) zero-initialised uint32, one per 64-channel slab — the last CTA of each slab also folds that slab's
 * partials in fixed order into out: fp32 [n, c] if per_image else [c], and leaves the counters at 0. */
int b200sd_colsum_blocks(int64_t n, int64_t hw, int c);
int b200sd_colsum(const void* x, int64_t ld_x, int64_t n, int64_t hw, int c, float* part, float* out, int per_image,
                  uint32_t* counter, void* stream);
/* b200sd_colsum without caller scratch and without any allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it). Same result as b200sd_colsum with `counter` (out: fp32 [n, c] if per_image else
 * [c], bit-identical — same kernel, same row blocks). The row-block partials and tail counters live in a static
 * __device__ array of the library: n * b200sd_colsum_blocks(n, hw, c) * c must be <= b200sd_colsum_static_max_floats()
 * (524,288 floats = 2 MiB, held by every GPU that loads the library until exit); larger requests return
 * B200SD_ERR_ARG with the sizes in b200sd_last_error(). Calls cudaGetSymbolAddress once per device (make the first
 * call on a device outside stream capture), then only cudaMemsetAsync of the counters + one kernel launch: no
 * cudaMalloc / cudaFree / allocating library routine. Every call shares the array, so calls that overlap on the GPU
 * (two streams, parallel graph branches) race; calls ordered on one stream are safe. */
int64_t b200sd_colsum_static_max_floats(void);
int b200sd_colsum_static(const void* x, int64_t ld_x, int64_t n, int64_t hw, int c, float* out, int per_image,
                         void* stream);
/* out[b, j] (+)= sum_r in[(b*rows + r)*cols*col_stride + j*col_stride], fp32, fixed order. */
int b200sd_reduce_f32(const float* in, int64_t batch, int64_t rows, int64_t cols, int64_t col_stride, float* out,
                      int accumulate, void* stream);

/* ----------------------------------------------------------------------------------------------- attention
 * softmax(Q K^T / sqrt(d)) V per (batch, head) — nn.MultiheadAttention call sites models/blocks.py:128,140,242,
 * 254,480,495; CustomMultiheadAttention models/multihead_attention.py:41-80; DiT Attention models/attention.py:33-78.
 * q: bf16 rows of length ld_q, head h at columns [h*d, (h+1)*d); likewise k, v with their own ld; row index =
 * b*L + token. out: bf16 [batch*lq, ld_o]. Scores, softmax and the PV accumulation are fp32. */
int b200sd_attention(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                     void* out, int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale,
                     void* stream);

/* The same product on the tcgen05 / TMEM / TMA kernel (csrc/attention_tc.cu: S = Q K^T and O = P V as tcgen05.mma with
 * TMEM accumulators, Q / K / V tiles by TMA with the head padding done by out-of-bounds zero fill, one softmax thread per
 * query row). b200sd_attention picks it by shape (head dim >= 32, at least 128 queries and keys); this entry point runs
 * it for any head dim that is a multiple of 8 in [8, 64]. lse: optional fp32 [batch, heads, lq] (log2 domain) or NULL.
 * Needs ld_o % 8 == 0, a 16-byte aligned out and scale > 0: B200SD_ERR_UNSUPPORTED otherwise. */
int b200sd_attention_tc(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v, void* out,
                        int64_t ld_o, float* lse, int64_t batch, int heads, int lq, int lk, int d, float scale,
                        void* stream);

/* Training variants. b200sd_attention_lse also writes lse: fp32 [batch, heads, lq], the log2-domain log-sum-exp of the
 * scaled scores. b200sd_attention_bwd recomputes the scores tile by tile from q, k, v and lse and returns dq / dk / dv
 * (bf16, same layout as q / k / v); o and d_o are the forward output and its gradient; dvec is fp32 scratch
 * [batch, heads, lq]. Needs K/V (and Q/dO) of one (batch, head) to fit shared memory: B200SD_ERR_UNSUPPORTED otherwise. */
int b200sd_attention_lse(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                         void* out, int64_t ld_o, float* lse, int64_t batch, int heads, int lq, int lk, int d,
                         float scale, void* stream);
int b200sd_attention_bwd(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                         const void* o, int64_t ld_o, const void* d_o, int64_t ld_do, const float* lse, float* dvec,
                         void* dq, int64_t ld_dq, void* dk, int64_t ld_dk, void* dv, int64_t ld_dv, int64_t batch,
                         int heads, int lq, int lk, int d, float scale, void* stream);

/* ------------------------------------------------------------------------------------- layout / glue
 * NCHW fp32 -> NHWC bf16 channel slice: dst[n, p, c_off + c] = bf16(src[n, c, p]), c < c. */
int b200sd_nchw_to_nhwc(const float* src, void* dst, int64_t ld_dst, int c_off, int64_t n, int c, int64_t hw,
                        void* stream);
/* NHWC (bf16 or fp32) -> NCHW fp32: dst[n, c, p] = src[n, p, c]. */
int b200sd_nhwc_to_nchw(const void* src, int src_dtype, int64_t ld_src, float* dst, int64_t n, int c, int64_t hw,
                        void* stream);
/* Mask conditioning — models/unet_cond_base.py:134-138: nearest-neighbour resize of mask (n, cin, mh, mw) fp32 to
 * (h, w) (src index = floor(dst * mh / h)), 1x1 conv cin->cout without bias (w: fp32 [cout, cin]), written as bf16
 * into channels [c_off, c_off+cout) of the NHWC conv_in input. Reads only the h*w sampled pixels. */
int b200sd_mask_cond(const float* mask, const float* w, void* dst, int64_t ld_dst, int c_off, int64_t n, int cin,
                     int cout, int mh, int mw, int h, int wd, void* stream);
/* dst[r, c_off + c] = act(src[r, c]) for a [rows, c] bf16 matrix (SiLU before the time-embedding Linear,
 * blocks.py:58-61; copy of the skip tensor into the concat buffer, blocks.py:464). */
int b200sd_copy_act(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int c_off, int64_t rows, int c,
                    int act, void* stream);

/* nn.Unfold(k, padding, stride) of an NHWC bf16 tensor x [n*h*w, ld_x] (c channels) into GEMM rows
 * rows[(img, oy, ox), ch*k*k + ky*k + kx] = x[img, oy*stride - pad + ky, ox*stride - pad + kx, ch] (0 outside): the
 * (c, ky, kx) K order of weight.view(Cout, -1), which is the axis the compute-in-memory weight blocks split
 * (cim_layers/layers_qn_lsq_adda_cim.py:92-94,198-215). Columns beyond c*k*k are left untouched.
 * b200sd_col2im is its adjoint (nn.Fold) for the backward pass: a per-input-element gather in fp32, fixed order. */
int b200sd_im2col(const void* x, int64_t ld_x, void* rows, int64_t ld_r, int64_t n, int h, int w, int c, int k,
                  int stride, int pad, void* stream);
int b200sd_col2im(const void* drows, int64_t ld_r, void* dx, int64_t ld_x, int64_t n, int h, int w, int c, int k,
                  int stride, int pad, void* stream);

/* dst[i] = bf16(src[i]) — the autocast cast of fp32 inputs (text context, utils/text_utils.py:32 output). */
int b200sd_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);

/* out[i] = a[i] + b[i] on contiguous bf16 (residual add when a quantised leaf prevents epilogue fusion). */
int b200sd_add_bf16(const void* a, const void* b, void* out, int64_t n, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * DiT token-row operators (csrc/rows.cu). Rows are bf16 [rows, ld] with `c` valid channels (c % 8 == 0, c <= 1024,
 * pitches multiples of 8); rows are grouped into images of `rows_per_img` tokens; per-image operands (shift, scale,
 * gate) are bf16 [n, ld_m] row vectors, typically column slices of the adaLN projection output.
 * ------------------------------------------------------------------------------------------------------------- */

/* y = LayerNorm(x, eps, no affine) * (1 + scale[img]) + shift[img]  — models/transformer_layer.py:88-89,103-104 and
 * models/transformer.py:196-197; shift = scale = NULL gives the un-modulated cross_attn_norm (:94).
 * stats (optional) receives fp32 (mean, rstd) per row for the backward. */
int b200sd_layernorm_mod(const void* x, int64_t ld_x, void* y, int64_t ld_y, const void* shift, const void* scale,
                         int64_t ld_m, int64_t rows, int64_t rows_per_img, int32_t c, float eps, float* stats,
                         void* stream);

/* Backward of b200sd_layernorm_mod: dx (+ dres, the gradient of a second consumer of x, optional). part (optional,
 * required when scale != NULL gradients are wanted) = fp32 [n, slots, 2, c] partial sums over
 * b200sd_rows_per_slot() rows of (dy, dy * xhat) = the gradients of (shift, scale); fold with b200sd_reduce_f32. */
int b200sd_layernorm_mod_bwd(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const float* stats,
                             const void* scale, int64_t ld_m, const void* dres, int64_t ld_r, void* dx, int64_t ld_dx,
                             int64_t n, int64_t rows_per_img, int32_t c, float* part, void* stream);
int b200sd_rows_per_slot(void);

/* y = x + gate[img] * h — the adaLN-Zero gated residual (transformer_layer.py:90,105) as a stand-alone kernel for
 * training (inference fuses it into b200sd_igemm's epilogue), and its backward: dh = gate * dy,
 * part = fp32 [n, slots, c] partial sums of dy * h (= d gate); dx = dy needs no kernel. */
int b200sd_gate_residual(const void* x, int64_t ld_x, const void* h, int64_t ld_h, const void* gate, int64_t ld_g,
                         void* y, int64_t ld_y, int64_t rows, int64_t rows_per_img, int32_t c, void* stream);
int b200sd_gate_residual_bwd(const void* dy, int64_t ld_dy, const void* h, int64_t ld_h, const void* gate,
                             int64_t ld_g, void* dh, int64_t ld_dh, int64_t n, int64_t rows_per_img, int32_t c,
                             float* part, void* stream);

/* dx = dy where y > 0 (nn.ReLU of the MLP block, transformer_layer.py:34-38; y is the saved forward output). */
int b200sd_relu_bwd(const void* dy, int64_t ld_dy, const void* y, int64_t ld_y, void* dx, int64_t ld_dx, int64_t rows,
                    int32_t c, void* stream);

/* tokens [(img, gy, gx)][(py, px, ch)] (bf16, or fp32 when tokens_f32) -> image fp32 NCHW [n, ch, gh*ph, gw*pw]
 * (einops 'b (nh nw) (ph pw c) -> b c (nh ph) (nw pw)', models/transformer.py:207-211) and its adjoint
 * (models/patch_embed.py:84-86), which zero-fills token columns beyond ph*pw*ch. */
int b200sd_unpatchify(const void* tokens, int32_t tokens_f32, int64_t ld, float* image, int64_t n, int32_t channels,
                      int32_t grid_h, int32_t grid_w, int32_t patch_h, int32_t patch_w, void* stream);
int b200sd_patchify(const float* image, void* tokens, int64_t ld, int64_t n, int32_t channels, int32_t grid_h,
                    int32_t grid_w, int32_t patch_h, int32_t patch_w, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * cim_layers quantisers on activations (csrc/quant.cu). Step sizes / ADC scale are read through device pointers
 * (they are nn.Parameters): no host synchronisation. Tensors are [rows, c] with pitch ld (c, ld multiples of 8),
 * bf16 or fp32 as flagged. Codes are bit-exact with the reference's fp32 torch ops on identical inputs.
 * ------------------------------------------------------------------------------------------------------------- */

/* Lazy LSQ step-size initialisation (cim_layers/layers_utils_lsq.py:31-34,40-42,57-59,74-76): while *step == 1.0,
 * *step = 1 / (1 / max|x| * qp). scratch: one uint32 of device memory. */
int b200sd_lsq_init_step(const void* x, int32_t x_f32, int64_t ld_x, int64_t rows, int32_t c, float* step, float qp,
                         uint32_t* scratch, void* stream);

/* data_quant_lsq (cim_layers/quant_noise_utils.py:101-120) with s' = the VALUE of grad_scale(s, grad_scale) =
 * fl(fl(s - s*g) + s*g), g = 1/sqrt(qp * numel) (:51-58): mode 0 -> round(clamp(x / s', -qp, qp)) * s' (isint=False),
 * mode 1 -> that code times s' / s' (isint=True, :117-118 — an integer up to the reference's own final rounding). */
int b200sd_lsq_quant(const void* x, int32_t x_f32, int64_t ld_x, void* y, int32_t y_f32, int64_t ld_y, int64_t rows,
                     int32_t c, const float* step, float qp, float grad_scale, int32_t mode, void* stream);

/* Its backward (straight-through round, clamp mask, LSQ step-size gradient): dx bf16; part[b200sd_lsq_quant_bwd_blocks()]
 * = fp32 partial sums of d(step) before the 1/sqrt(qp * numel) gradient scale (quant_noise_utils.py:51-58).
 * total / counter (both or neither): the last CTA folds the partials in fixed order and writes
 * *total = sum(part) * grad_scale; counter = a zeroed uint32 that is returned to zero. */
int b200sd_lsq_quant_bwd(const void* dy, int64_t ld_dy, const void* x, int32_t x_f32, int64_t ld_x, void* dx,
                         int64_t ld_dx, int64_t rows, int32_t c, const float* step, float qp, float grad_scale,
                         int32_t mode, float* part, float* total, uint32_t* counter, void* stream);
int b200sd_lsq_quant_bwd_blocks(int64_t rows, int32_t c);

/* Weight side of a cim_layers leaf in two launches: weight_quant_noise (cim_layers/layers_utils_lsq.py:48-64) = lazy
 * step init (while *step == 1: *step = 1 / (1 / max|w| * qp)), weight_quant_lsq (quant_noise_utils.py:160-179; isint:
 * q * s' / s') and add_noise (:92-98: + (max wq - min wq) * noise_scale * unit[i]; unit = the caller's N(0,1) draw, required
 * when noise_scale != 0). w, out: fp32 [n]. scratch: b200sd_wq_scratch_floats() floats, ZEROED once by the caller and
 * then kept with the layer: it carries the extrema and tie counts to b200sd_weight_qn_bwd, which returns
 * dw = dL/dw (straight-through round, clamp mask, the noise term's gradient through max / min spread evenly over tied
 * codes as torch does) and *dstep = dL/dstep including the 1/sqrt(qp n) gradient scale. grad_scale = that factor. */
int b200sd_wq_scratch_floats(void);
int b200sd_weight_qn(const float* w, int64_t n, float* step, float qp, float grad_scale, int32_t isint, int32_t lazy_init,
                     float noise_scale, const float* unit, float* out, float* scratch, void* stream);
int b200sd_weight_qn_bwd(const float* gout, const float* w, int64_t n, const float* step, float qp, float grad_scale,
                         int32_t isint, float noise_scale, const float* unit, float* dw, float* dstep, float* scratch,
                         void* stream);

/* The same two passes for MANY layers in one launch per phase (forward: statistics, apply; backward: noise dot product,
 * apply) — the table rides in the kernel parameters, so nothing is uploaded or kept alive and the call is CUDA-graph
 * capturable. Row fields as the arguments of b200sd_weight_qn / b200sd_weight_qn_bwd: `out` is the quantised weight in
 * the forward pass and d(weight) in the backward pass (backward != 0), which also needs gout and dstep. `scratch` is
 * the layer's own b200sd_wq_scratch_floats() buffer, kept from its forward to its backward. */
typedef struct {
    const float* w;
    float* step;
    const float* unit;
    float* out;
    const float* gout;
    float* dstep;
    float* scratch;
    int64_t n;
    float qp, grad_scale, noise_scale;
    int32_t isint, lazy_init;
} b200sd_wq_job;
int b200sd_weight_qn_multi(const b200sd_wq_job* jobs, int32_t n_jobs, int32_t backward, void* stream);

/* bit_split_tensor (cim_layers/layers_utils_adda.py:161-185): codes [rows, c] -> `slices` signed digits of slice_bit
 * bits stacked along rows, bf16 [slices * rows, c] — the DAC input slices. step == NULL: x holds the codes (bf16 or
 * fp32); step != NULL: x holds activations and the LSQ code (mode 1 above, with qp / grad_scale) is computed in the
 * same pass, so the fp32 codes never travel through HBM. */
int b200sd_bit_split(const void* x, int32_t x_f32, int64_t ld_x, void* digits, int64_t ld_o, int64_t rows, int32_t c,
                     int32_t slices, int32_t slice_bit, int32_t x_bit, const float* step, float qp, float grad_scale,
                     void* stream);

/* ADC + bit_concat of one mapped weight block (cim_layers/layers_qn_lsq_adda_cim.py:464-480,
 * layers_utils_adda.py:108-116,144-159): partial = fp32 [slices * rows, ncol] analogue sums of the block;
 * out[r, col0 + j] (+)= sum_k 2^(k * slice_bit) * round(clamp(scale * partial_k * (1 + gain[col0 + j])
 *                                                          + range * offset[col0 + j], -range - 1, range)). */
int b200sd_adc_combine(const float* partial, int64_t ld_p, float* out, int64_t ld_o, int64_t rows, int32_t ncol,
                       int32_t col0, int32_t slices, int32_t slice_bit, const float* adc_scale, float adc_range,
                       const float* gain_noise, const float* offset_noise, int32_t accumulate, void* stream);

/* Backward of b200sd_adc_combine for one block (straight-through round, clamp mask on the noisy scaled value):
 * dpartial bf16 [slices * rows, ld_d] = adc_scale * 2^(k*slice_bit) * dacc[r, col0 + j] * [in range] (the operand of
 * the block's dgrad / wgrad GEMMs); part[b200sd_adc_bwd_blocks()] = fp32 partial sums of d(adc_scale). */
int b200sd_adc_bwd(const float* partial, int64_t ld_p, const float* dacc, int64_t ld_g, void* dpartial, int64_t ld_d,
                   int64_t rows, int32_t ncol, int32_t col0, int32_t slices, int32_t slice_bit, const float* adc_scale,
                   float adc_range, const float* gain_noise, const float* offset_noise, float* part, float* total,
                   uint32_t* counter, void* stream);   /* total / counter as in b200sd_lsq_quant_bwd: *total = sum(part) */
int b200sd_adc_bwd_blocks(int64_t rows, int32_t ncol);

/* Backward of b200sd_bit_split (layers_utils_adda.py:178-183): dcodes = sum_k ddigits_k / (slices * 2^(k*slice_bit)). */
int b200sd_bit_split_bwd(const void* ddigits, int64_t ld_d, void* dcodes, int64_t ld_q, int64_t rows, int32_t c,
                         int32_t slices, int32_t slice_bit, void* stream);

/* Output side of a compute-in-memory layer in one launch (cim_layers/layers_qn_lsq_adda_cim.py:482-500 — the
 * reference's `x = x / w_scale / in_scale / adc_scale`, `+ bias`, `data_quant_lsq(x, step_size_output)`):
 *   y = acc / (1 / *step_w) / (1 / *step_in) / *adc_scale + bias[col]     (true divisions, that order; step_w may be
 *                                                                          NULL: w_scale = 1)
 *   mode 0: out = fp32 y;  mode 1: out = bf16 round(clamp(y / s', -qp, qp)) * s', s' from *step_out and grad_scale as
 *   in b200sd_lsq_quant;  mode 2: out = bf16 y.      acc fp32 [rows, ld_a], c % 8 == 0, bias fp32 [c] or NULL. */
int b200sd_adda_out(const float* acc, int64_t ld_a, const float* bias, const float* step_w, const float* step_in,
                    const float* adc_scale, const float* step_out, float qp, float grad_scale, void* out, int64_t ld_o,
                    int64_t rows, int32_t c, int32_t mode, void* stream);

/* Backward of b200sd_adda_out (step_out NULL: no output quantiser). gout bf16 [rows, ld_g]; dacc fp32 [rows, c];
 * dy bf16 [rows, c] or NULL = the gradient behind the quantiser's clamp mask (its column sums are d bias);
 * res[0] = d(step_out) * grad_scale, res[1] = d(adc_scale) through the division. scratch: zero-initialised floats,
 * b200sd_adda_out_scratch_floats() of them, private to one layer (the last CTA folds the per-CTA partials). */
int b200sd_adda_out_bwd(const void* gout, int64_t ld_g, const float* acc, int64_t ld_a, const float* bias,
                        const float* step_w, const float* step_in, const float* adc_scale, const float* step_out,
                        float qp, float grad_scale, float* dacc, void* dy, int64_t rows, int32_t c, float* scratch,
                        float* res, void* stream);
int b200sd_adda_out_scratch_floats(void);

/* Single-block layers: b200sd_adc_combine + b200sd_adda_out in one pass over the fp32 block product partial
 * [slices * rows, ld_p] (the integer accumulation is never written out); out bf16 [rows, c], c % 4 == 0. step_out NULL:
 * no output quantiser. Same arithmetic in the same order as the two calls it replaces. */
int b200sd_adda_tail(const float* partial, int64_t ld_p, const float* bias, const float* step_w, const float* step_in,
                     const float* adc_scale, const float* step_out, float qp, float grad_scale, float adc_range,
                     int32_t slices, int32_t slice_bit, const float* gain_noise, const float* offset_noise, void* out,
                     int64_t rows, int32_t c, void* stream);
/* Its backward = b200sd_adda_out_bwd + b200sd_adc_bwd: dpartial bf16 [slices * rows, ld_d] (operand of the dgrad / wgrad
 * GEMMs), dy bf16 [rows, c] or NULL (column sums = d bias), res[0] = d(step_out) * grad_scale, res[1] = d(adc_scale)
 * through both the division and the ADC input; scratch as for b200sd_adda_out_bwd. */
int b200sd_adda_tail_bwd(const void* gout, int64_t ld_g, const float* partial, int64_t ld_p, const float* bias,
                         const float* step_w, const float* step_in, const float* adc_scale, const float* step_out,
                         float qp, float grad_scale, float adc_range, int32_t slices, int32_t slice_bit,
                         const float* gain_noise, const float* offset_noise, void* dpartial, int64_t ld_d, void* dy,
                         int64_t rows, int32_t c, float* scratch, float* res, void* stream);

/* get_adc_scale (cim_layers/layers_utils_adda.py:40-49) on the device: out[0] = adc_scale of *gain (clamped to
 * [gain_min, gain_max] straight-through; gain_mode != 0 rounds the gain, 0 rounds the range 1 / gain) * scale1,
 * out[1] = d adc_scale / d gain. */
int b200sd_adc_scale(const float* gain, float gain_min, float gain_max, int32_t gain_mode, float scale1, float* out,
                     void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Enhance wrappers (cim_qn_train/layers_enhance.py), csrc/enhance.cu.
 * ------------------------------------------------------------------------------------------------------------- */

/* Block-diagonal dense form of a GroupedLinear / grouped 1x1 Conv2d parameter (layers_enhance.py:324-365):
 * dense[(g*out_group + r) * ld_d + g*in_group + c] = bf16(w[g, r, c]) and, when dense_t is non-NULL, its transpose
 * dense_t[(g*in_group + c) * ld_t + g*out_group + r]. Only the diagonal blocks are written: the caller keeps both
 * buffers zero-initialised between calls. */
int b200sd_blockdiag_pack(const float* w, int32_t groups, int32_t out_group, int32_t in_group, void* dense, int64_t ld_d,
                          void* dense_t, int64_t ld_t, void* stream);

/* dw[g, r, c] = dense[(g*out_group + r) * ld + g*in_group + c]: the parameter gradient out of a dense weight gradient. */
int b200sd_blockdiag_gather(const float* dense, int64_t ld, int32_t groups, int32_t out_group, int32_t in_group, float* dw,
                            void* stream);

/* merge_feature_map (layers_enhance.py:596-608) in one launch: out = orig * (1 - s) + enh * s, s = sigmoid(*blend) when
 * `sigmoid` else *blend, optional ReLU; bf16 [rows, c] operands, c % 8 == 0. stats (optional) receives std(orig),
 * mean(orig), std(enh), mean(enh) (unbiased std, as Tensor.std()); it needs part[b200sd_blend_part_floats()] and a
 * zeroed uint32 counter (returned to zero). */
int b200sd_blend(const void* orig, int64_t ld_a, const void* enh, int64_t ld_b, const float* blend, int32_t sigmoid,
                 int32_t relu, void* out, int64_t ld_o, int64_t rows, int32_t c, float* part, uint32_t* counter,
                 float* stats, void* stream);
int b200sd_blend_part_floats(void);

/* Its backward: d_orig = g (1 - s), d_enh = g s (bf16 [rows, c] contiguous), *d_blend = sum g (enh - orig) times
 * s (1 - s) when `sigmoid`. out: the forward result when a ReLU was fused (masks g), else NULL. */
int b200sd_blend_bwd(const void* g, int64_t ld_g, const void* orig, int64_t ld_a, const void* enh, int64_t ld_b,
                     const void* out, int64_t ld_o, const float* blend, int32_t sigmoid, void* d_orig, void* d_enh,
                     int64_t rows, int32_t c, float* part, uint32_t* counter, float* d_blend, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Weight-side data movement of a training step (csrc/pack.cu).
 * ------------------------------------------------------------------------------------------------------------- */

/* One job: packed[r, t * cp + c] = bf16(src[r * s_r + c * s_c + tap_off[t]]) for c < cin, 0 for cin <= c < cp; the
 * K-major matrices b200sd_igemm / b200sd_wgrad read, from parameters in the reference layouts (nn.Conv2d
 * [Cout, Cin, kh, kw], nn.ConvTranspose2d [Cin, Cout, 4, 4], nn.Linear [N, K]); forward, transposed and
 * flipped-tap (dgrad) forms differ only in strides. tile0 / tiles_c are the launch geometry: a job owns
 * ceil(rows / 32) * (cp / 32) consecutive tiles starting at tile0, tiles_c = cp / 32. */
typedef struct {
    const float* src;
    void* dst;               /* bf16, already offset to the job's first column */
    int64_t s_r, s_c, dst_ld;
    int32_t rows, cin, cp, taps;
    int32_t tap_off[16];
    int32_t tile0, tiles_c;
} b200sd_pack_job;

/* jobs_dev: DEVICE array of n_jobs jobs sorted by tile0; n_tiles = total tiles. Every weight of a model in one launch. */
int b200sd_pack_weights(const void* jobs_dev, int32_t n_jobs, int32_t n_tiles, void* stream);
int b200sd_pack_job_size(void);

/* Split-K partials of b200sd_wgrad ([splits, n_out, k_total], columns col0 + t * cp + c) -> the gradient in the
 * parameter's own layout: out[co * s_co + c * s_ci + tap_off[t]] (+)= sum over splits (fixed order). */
int b200sd_wgrad_scatter(const float* partial, int32_t splits, int32_t n_out, int32_t k_total, int32_t col0, int32_t cin,
                         int32_t cp, int32_t taps, const int32_t* tap_off, float* out, int64_t s_co, int64_t s_ci,
                         int32_t accumulate, void* stream);

/* ------------------------------------------------------------------------------------------ VQVAE.quantize
 * models/vqvae.py:93-128 (called by VQVAE.encode :130-139 when latents are pre-generated,
 * gen_vqvae_latents_multi_GPU.py:42-188). x: fp32 rows [n*hw, ld_x] (NHWC pixels, c <= 16 channels); codebook fp32
 * [n_codes, c] (nn.Embedding weight). idx[n*hw]: index of the nearest code (first of equal minima); zq_nchw: fp32
 * [n, c, hw] = x + (e - x) as the straight-through expression evaluates; loss[0] = mean((e - x)^2), the value of both
 * codebook_loss and commitment_loss. partial: fp32 scratch [b200sd_vq_blocks(n*hw)]; counter: one zeroed uint32. */
int b200sd_vq_blocks(int64_t rows);
int b200sd_vq_nearest(const float* x, int64_t ld_x, const float* codebook, int32_t n_codes, int32_t c, int64_t n,
                      int64_t hw, int64_t* idx, float* zq_nchw, float* partial, float* loss, uint32_t* counter,
                      void* stream);

/* ------------------------------------------------------------------------------------------ optimizer step
 * Replaces, for one training step, torch.nn.utils.clip_grad_norm_ + the isfinite(grad_norm) check + Adam.step + the
 * per-parameter EMA loop of train_ddpm_cond_celebhq_multi_gpu.py:362-378 (same code in Model_DiT_9L_train.py).
 * Gradients, exp_avg and exp_avg_sq are flat fp32 buffers; a chunk names a run of <=
 * b200sd_opt_chunk_elems() consecutive elements of one parameter: its address, its EMA shadow's (or NULL), the run's
 * offsets in the moment buffers and in the gradient buffer, and its length. vec = 1 when p (and ema) are 16-byte aligned. */
typedef struct {
    float* p;
    float* ema;
    int64_t off;  /* offset of the run in m_base / v_base */
    int64_t goff; /* offset of the run in g_base (the gradient buffer may use another slot order) */
    int32_t n;
    int32_t vec;
} b200sd_opt_chunk;
int b200sd_opt_chunk_size(void);
int b200sd_opt_chunk_elems(void);
/* scalars (device fp32[4]): [0] total L2 norm of the gradients, [1] clip coefficient min(1, max_norm / (norm + 1e-6))
 * (1 when max_norm <= 0), [2] 1 if the norm is finite (or check_finite == 0) else 0, [3] step counter, incremented here
 * when [2] is 1. partial: fp32 scratch [n_chunks]; counter: one zeroed uint32 (left at 0). chunks: DEVICE array. */
int b200sd_grad_sqnorm(const void* chunks, int32_t n_chunks, const float* g_base, float max_norm, int32_t check_finite,
                       float* partial, float* scalars, uint32_t* counter, void* stream);
/* Adam (torch.optim.Adam semantics: L2 weight decay, bias correction from scalars[3]) on the clipped gradient, then
 * ema = ema_decay * ema + (1 - ema_decay) * p. Does nothing when scalars[2] == 0. */
int b200sd_adam_ema_step(const void* chunks, int32_t n_chunks, const float* g_base, float* m_base, float* v_base,
                         const float* scalars, float lr, float beta1, float beta2, float eps, float weight_decay,
                         float ema_decay, void* stream);
/* PercentOptimizerFP.step (cim_layers/IBA_optimizer.py:80-143) for the tensors named by the chunk table: Adam moments
 * (w1 = fp32(1 - beta1), w2 = fp32(1 - beta2)), u = (adj * m) / (sqrt(v) / bc2_sqrt + eps) with adj = sqrt(bc2) / bc1,
 * then p -= u * max(lr * max|p| / (max|u| + eps), lr) per TENSOR. tensor_of_chunk: device int32 [n_chunks], the dense
 * tensor index of each chunk; tensor_max: device uint32 [n_tensors][2], ZEROED by the caller before the call (float
 * bit patterns of max|u|, max|p|). Two launches, fp32 operation order of the reference, IEEE-rounded each step. */
int b200sd_percent_opt_step(const void* chunks, const int32_t* tensor_of_chunk, int32_t n_chunks, const float* g_base,
                            float* m_base, float* v_base, uint32_t* tensor_max, float lr, float beta1, float beta2,
                            float w1, float w2, float adj, float bc2_sqrt, float eps, void* stream);

#ifdef __cplusplus
}
#endif
#endif
