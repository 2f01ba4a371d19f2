/* sdb200 — C ABI of the B200-native (sm_100a) denoising train-step kernels.
 *
 * This is the drop-in boundary for ONE hot path of lodestone-rock/stable_diffusion_training: the data-parallel
 * denoising train step (training_utils.py:504-762) and the operators it runs (schedulers/, lion_quant.py, the
 * diffusers-0.21.4 Flax UNet it calls). The reference has no FFI of its own (pure JAX); each entry point below cites
 * the reference code it replaces, and INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 * Conventions (all entry points):
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends in _host;
 *   - the caller owns every buffer including workspaces; the library allocates nothing persistent;
 *   - return 0 on success, a negative SdbStatus on failure; sdb_last_error_string() gives the thread-local reason;
 *   - never throws, never exits, never synchronises: work is only enqueued on `stream` (CUDA-graph capturable);
 *   - bf16 tensors are passed as `void*` (raw __nv_bfloat16 storage); activations are NHWC ("pixels x channels").
 */
#ifndef SDB200_H
#define SDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDB200_ABI_VERSION 6

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

/* ---- library ---- */
const char* sdb_last_error_string(void);
int sdb_abi_version(void);
int sdb_check_device(int device); /* fails unless the device is compute capability 10.x */
long long sdb_launch_count(void); /* kernels this library has launched in this process (monotonic; bench.py's
                                     gpu_launches is the difference across the timed region) */

/* ---- 8-bit Lion (lion_quant.py:20-156, chain 201-211; optax clip/apply as wired at training_utils.py:355-382, 732)
 * One launch over a flat fp32 segment. Block b of `block_size` (16, 32 or 64; lion_quant.py:66-80 reshapes to any
 * block, the reference config uses 16, lion_8bit's default is 64) consecutive elements has int8
 * codes[bs*b .. bs*b+bs-1] and the RECIPROCAL absmax inv_scales[b]. The scale dtype is fp32
 * (lion_quant.py:47 canonicalize_dtype(None)); other mu_scale_dtype values are not built. `gnorm_dev` (nullable) is
 * the global gradient norm: when it is >= max_norm the gradient is divided by it first (optax.clip_by_global_norm).
 * mode 0: out = p + (-lr) * (sign(c) + wd*p)   (out may alias p_in; optax.apply_updates fused)
 * mode 1: out =     (-lr) * (sign(c) + wd*p)   (the optax `updates` tree, parameters untouched)
 * shadow_bf16 (nullable, mode 0 only): bf16 copy of the new parameters for the GEMM operands.
 * ema (nullable, mode 0 only): fp32 exponential moving average of the parameters, updated in the same pass with the
 *   NEW parameters: ema = ema_rate * ema + (1 - ema_rate) * out  (compute_model_ema, training_utils.py:537-544,
 *   applied at 735-739). */
int sdb_lion8_update(const float* g, const float* p_in, float* out, void* shadow_bf16, int8_t* codes,
                     float* inv_scales, long long n, int block_size, double b1, double b2, double lr, double wd,
                     int wd_flag, const float* gnorm_dev, double max_norm, int mode, float* ema, double ema_rate,
                     cudaStream_t stream);

/* Unquantised-state leaves of the same optimizer (lion_quant.py:109, 145). */
int sdb_lion_fp32_update(const float* g, const float* p_in, float* out, void* shadow_bf16, float* mu, long long n,
                         double b1, double b2, double lr, double wd, int wd_flag, const float* gnorm_dev,
                         double max_norm, int mode, float* ema, double ema_rate, cudaStream_t stream);

/* compute_model_ema as a stand-alone leaf op (training_utils.py:537-544): ema = ema_rate * ema + (1 - ema_rate) * p. */
int sdb_ema_update(float* ema, const float* params, long long n, double ema_rate, cudaStream_t stream);

/* optax.global_norm over a flat fp32 buffer: gnorm_out[0] = sqrt(sum x^2 [+ carried sum if accumulate]).
 * Deterministic (fixed grid, fixed reduction tree). Workspace: sdb_sumsq_workspace_bytes(). */
size_t sdb_sumsq_workspace_bytes(void);
int sdb_global_norm(const float* x, long long n, void* workspace, size_t ws_bytes, float* gnorm_out, int accumulate,
                    cudaStream_t stream);

/* Per-leaf pieces of the optax chain for the generic pytree API (optax.clip_by_global_norm, optax.apply_updates). */
int sdb_clip_by_global_norm_f32(const float* g, float* out, long long n, const float* gnorm_dev, double max_norm,
                                cudaStream_t stream);
int sdb_add_f32(const float* a, const float* b, float* out, long long n, cudaStream_t stream);

/* Reduction step of the peer-memory gradient all-reduce that replaces the all-reduce GSPMD inserts for the reference's
 * data-parallel step (training_utils.py:835-840, 852-931): dst[i] = sum in rank order of the nranks contributions to
 * this rank's slice; the own contribution is dst itself (position self_pos), staged_host is a HOST array of the
 * nranks - 1 device pointers (ascending rank, self skipped) holding the slices pulled from the peers. nranks <= 9. */
int sdb_sum_slices_f32(float* dst, const float* const* staged_host, int nranks, int self_pos, long long n,
                       cudaStream_t stream);
/* The same exchange step through NVSwitch multicast (NVLS): `multicast_slice` is the multicast address of this rank's
 * slice of the bucket (every rank's gradient arena bound to one multicast object); the switch returns the sum over
 * ranks (multimem.ld_reduce) and the kernel stores it to every rank's copy (multimem.st). Call between two cross-rank
 * barriers; every rank calls it on ITS slice, so all ranks end with identical bits. ctas <= 0: 148. */
int sdb_nvls_allreduce_f32(float* multicast_slice, long long n, int ctas, cudaStream_t stream);

/* ---- schedulers (schedulers/scheduling_utils_flax.py:316-343; scheduling_ddpm_flax.py:196-297) ----
 * alphas_cumprod has num_timesteps entries; a timestep outside the table is wrapped once if negative and clamped,
 * as jnp's gather does (it never reads out of bounds). */
int sdb_add_noise(const float* original_samples, const float* noise, const int* timesteps, const float* alphas_cumprod,
                  int num_timesteps, float* noisy, int batch, long long per_sample, cudaStream_t stream);
int sdb_get_velocity(const float* sample, const float* noise, const int* timesteps, const float* alphas_cumprod,
                     int num_timesteps, float* velocity, int batch, long long per_sample, cudaStream_t stream);
/* prediction_type: 0 epsilon, 1 sample, 2 v_prediction. Scalars are computed by the host in fp32 exactly as the
 * reference does (scheduling_ddpm_flax.py:229-264); noise == NULL means t == 0 (no variance term). */
int sdb_ddpm_step(const float* model_output, const float* sample, const float* noise, float* prev_sample, long long n,
                  float sqrt_alpha_prod, float sqrt_beta_prod, float pred_original_coeff, float current_sample_coeff,
                  float sigma, int prediction_type, int clip_sample, cudaStream_t stream);

/* FlaxDDIMScheduler.step (schedulers/scheduling_ddim_flax.py:199-284): prev = sqrt(a_prev) * x0 + direction_coeff * eps
 * with (x0, eps) from the model output per prediction type; direction_coeff = sqrt(1 - a_prev - sigma_t^2). eta = 0
 * (the reference's default and the only use, training_utils.py:997-1004) has no noise term. */
int sdb_ddim_step(const float* model_output, const float* sample, float* prev_sample, long long n,
                  float sqrt_alpha_prod, float sqrt_beta_prod, float sqrt_alpha_prod_prev, float direction_coeff,
                  int prediction_type, cudaStream_t stream);

/* ---- loss (training_utils.py:704-709, min-SNR weight 546-568) ----
 * pred: UNet output NHWC bf16 [B,HW,C]; target: NCHW fp32 [B,C,HW]; sample_weight: nullable fp32 [B].
 * loss[0] = mean(w_b (target-pred)^2); dpred (nullable) = d loss / d pred * loss_grad_scale, NHWC bf16. */
size_t sdb_mse_workspace_bytes(void);
/* min_snr_gamma_loss_rescale (training_utils.py:546-568): weight[b] = min(snr, gamma) / snr  (epsilon) or
 * / (snr + 1) (v_prediction), snr = ac[t_b] / (1 - ac[t_b]); fp32 IEEE division; no host round trip. */
int sdb_min_snr_weight(const float* alphas_cumprod, int num_timesteps, const int* timesteps, int batch, float gamma,
                       int v_prediction, float* weight, cudaStream_t stream);
int sdb_mse_fwd_bwd(const void* pred_nhwc_bf16, const float* target_nchw, const float* sample_weight,
                    void* dpred_nhwc_bf16, float* loss, int B, int C, int HW, float loss_grad_scale, void* workspace,
                    size_t ws_bytes, cudaStream_t stream);

/* ---- in-step sampling (training_utils.py:590-624): jax.random's Threefry-2x32-20 counter-based generator ----
 * Keys are two uint32 words (jax.random.PRNGKey / split, done on the host: jax_random.py). Arrays are drawn as jax
 * draws them — element i of threefry_2x32(key, iota(n_total)) — and a caller may ask for the sub-range
 * [offset, offset + n_local) only (a data-parallel rank's rows of the global batch). Bits and randint are bit-exact
 * with jax's published algorithm; normal = sqrt(2) * erf_inv(uniform) follows XLA's F32 ErfInv polynomial. */
int sdb_threefry_bits(uint32_t key0, uint32_t key1, uint32_t* out, long long n_total, long long offset,
                      long long n_local, cudaStream_t stream);
/* noise[b,c,p] = N(noise_key) (+ N(offset_key)[b,c] * offset_noise_magnitude) (+ perturbation_noise_magnitude *
 * N(perturb_key)[b,c,p]) for rows batch_begin .. batch_begin + batch_local - 1 of a [batch_global, channels, hw]
 * draw; keys are 2-word HOST arrays (offset / perturb keys may be NULL when their magnitude is 0). */
int sdb_sample_noise(const uint32_t* noise_key_host, const uint32_t* offset_key_host, const uint32_t* perturb_key_host,
                     float offset_noise_magnitude, float perturbation_noise_magnitude, float* noise_out,
                     int batch_global, int batch_begin, int batch_local, int channels, int hw, cudaStream_t stream);
/* jax.random.randint(key, (n_total,), minval, maxval) -> int32, sub-range as above. */
int sdb_threefry_randint(uint32_t key0, uint32_t key1, int* out, long long n_total, long long offset, long long n_local,
                         int minval, int maxval, cudaStream_t stream);

/* ---- tensor-core GEMM / implicit-GEMM convolution (tcgen05 + TMEM + TMA) ----
 * Replaces flax nn.Conv / nn.Dense forward and their VJPs inside FlaxUNet2DConditionModel (diffusers 0.21.4,
 * called at training_utils.py:678-684).
 *
 * A 4-D bf16 view: extents (C innermost, then W, H, N) with element strides; C is contiguous. A [rows, cols]
 * matrix is {C = cols, W = rows, H = N = 1, strideW = ld}. Strides must be multiples of 8 elements. */
typedef struct {
  const void* ptr;
  int C, W, H, N;
  long long strideW, strideH, strideN;
} sdb_view4;

#define SDB_MAX_TAPS 9
#define SDB_MAX_SRC 4

/* mode SDB_GEMM_FWD:  D[pixel, n] = sum_taps sum_k A_src(tap)[pixel + (dh,dw), k] * Bop[tap][k, n]   (+ epilogue)
 *      the output pixel space is (outW, outH, outN); a tile is a box (bw, bh, bn) of 128 pixels.
 *      b_mn_major = 1: B memory is [K rows, N cols] (forward weights, Flax HWIO / (in,out));
 *                      tap_brow[t] = first K row of tap t.
 *      b_mn_major = 0: B memory is [N rows, K cols] (the same weights used transposed for dgrad);
 *                      tap_brow[t] = first N row of tap t.
 *      epilogue: + bias[n] (fp32, nullable) + temb[img, n] (bf16, nullable, row stride ldt)
 *                + residual[pixel, n] (bf16, nullable, row stride ldr); output bf16, row stride ldd.
 * mode SDB_GEMM_WGRAD: dW[tap_brow[t] + m, n] (+)= sum_pixels A_src(tap)[pixel + (dh,dw), m] * B[pixel, n]
 *      A and B are both pixel-major activations (B is passed as `b`, a 4-D view over the same pixel space);
 *      a K block is a box (bw, bh, bn) of 64 pixels; output fp32 with row stride ldd; `splits` > 1 splits the pixel
 *      range and accumulates with red.global.add (the call zeroes the rows of D it covers first); splits == 1 stores.  */
enum { SDB_GEMM_FWD = 0, SDB_GEMM_WGRAD = 1 };

typedef struct {
  int mode;
  int num_src;
  sdb_view4 a[SDB_MAX_SRC];
  sdb_view4 b;
  int b_mn_major;
  int ntaps;
  int tap_src[SDB_MAX_TAPS];
  int tap_dh[SDB_MAX_TAPS];
  int tap_dw[SDB_MAX_TAPS];
  int tap_brow[SDB_MAX_TAPS];
  int outW, outH, outN; /* pixel space */
  int bw, bh, bn;       /* pixel box: 128 * m_halves pixels (FWD) / 64 pixels (WGRAD) */
  int M;                /* WGRAD: rows of dW per tap (Cin). FWD: ignored (pixels) */
  int N;                /* output columns */
  int K;                /* FWD: reduction length per tap (channels of A) */
  int block_n;          /* N tile width, multiple of 16, <= 256 */
  int splits;           /* WGRAD: split of the pixel range. FWD: split-K (> 1: D is fp32 scratch, row stride ldd;
                           bias / temb / residual / d_stride* must be unset; finish with sdb_splitk_finish).
                           splitk_slab > 0: K range s stores its partial sums into the slab D + s * splitk_slab
                           (the effective number of ranges is ceil(KB / ceil(KB / splits)), KB = ntaps * ceil(K / 64))
                           — deterministic; splitk_slab == 0: one [pixels, N] buffer that the call zeroes and
                           accumulates into with red.global.add */
  void* d;
  long long ldd;
  long long d_strideW, d_strideH, d_strideN; /* FWD only: element strides of the output pixel grid; all 0 = dense
                                                (ldd, ldd*outW, ldd*outW*outH). Lets a tile write a strided sub-grid
                                                (stride-2 transposed convolution by output parity). */
  const float* bias;
  const void* temb;
  long long ldt;
  const void* residual;
  long long ldr;
  int m_halves;         /* sdb_gemm_tc: M tile = 128 * m_halves rows (1 or 2; 0 = 1). Two halves share every B stage */
  long long splitk_slab; /* FWD split-K: elements between per-range slabs (see splits); 0 = atomic accumulation.
                            WGRAD with splits > 1: > 0 = every pixel range stores its dW-shaped partial result into the
                            slab D + s * splitk_slab (the effective number of ranges is ceil(PB / ceil(PB / splits)),
                            PB = pixel blocks of 64); fold them with sdb_splitk_finish_f32 — run-to-run identical
                            bits. 0 = the call zeroes dW and the ranges accumulate with red.global.add (order varies) */
} sdb_gemm_desc;

int sdb_gemm_tc(const sdb_gemm_desc* desc, cudaStream_t stream);

/* Reference-quality CUDA-core GEMM for shapes the tensor-core tiles do not cover (time embedding M = batch,
 * conv_in K = 4 per tap, conv_out N = 4) and as the on-device cross-check of sdb_gemm_tc in the tests.
 * Same descriptor, same semantics; no alignment or divisibility constraints. */
int sdb_gemm_simt(const sdb_gemm_desc* desc, cudaStream_t stream);

/* 3x3 stride-1 pad-1 convolutions with a 4-channel side — FlaxUNet2DConditionModel.conv_in (4 -> C) and .conv_out
 * (C -> 4) [diffusers 0.21.4; training_utils.py:678-684]: K = 4 per tap / N = 4 cannot fill a tcgen05 tile and the
 * layers are memory-bound, so they run on CUDA cores. NHWC bf16 activations, HWIO bf16 weights, fp32 bias / dW.
 * C: the wide channel count (multiple of 8, <= 904). wgrad zeroes dw itself (cudaMemsetAsync) and accumulates with red. */
int sdb_conv3x3_c4in_fwd(const void* x4, const void* w_hwio, const float* bias, void* y, int B, int H, int W, int C,
                         cudaStream_t stream);
int sdb_conv3x3_c4out_fwd(const void* x, const void* w_hwio, const float* bias, void* y4, int B, int H, int W, int C,
                          cudaStream_t stream);
int sdb_conv3x3_c4out_dgrad(const void* dy4, const void* w_hwio, void* dx, int B, int H, int W, int C,
                            cudaStream_t stream);
/* c4_is_input = 1: conv_in  (c4 = x [B,H,W,4],  wide = dy [B,H,W,C]) -> dw [3,3,4,C]
 * c4_is_input = 0: conv_out (c4 = dy [B,H,W,4], wide = x  [B,H,W,C]) -> dw [3,3,C,4]
 * workspace (sdb_conv3x3_c4_wgrad_workspace_bytes(C) bytes): every CTA stores its partial dw there and a second launch
 * folds them in CTA order — run-to-run identical bits; NULL: red.global.add into the zeroed dw (order varies). */
size_t sdb_conv3x3_c4_wgrad_workspace_bytes(int C);
int sdb_conv3x3_c4_wgrad(const void* wide, const void* c4, float* dw, int B, int H, int W, int C, int c4_is_input,
                         void* workspace, size_t ws_bytes, cudaStream_t stream);

/* ---- normalisation / activation (flax.linen GroupNorm, LayerNorm, nn.gelu as used by diffusers-0.21.4
 * FlaxResnetBlock2D / FlaxTransformer2DModel / FlaxBasicTransformerBlock / FlaxGEGLU [upstream, un-vendored], reached
 * from training_utils.py:678-684). bf16 "pixels x channels" tensors, fp32 parameters and statistics.
 * GroupNorm: 32 groups, fast variance, y = act((x-mean)*rstd*gamma+beta), act = SiLU when `silu`.
 *   stats: fp32 [B][32][2] = (mean, rstd), written by fwd, read by bwd.
 *   bwd: dx = GN'(dy) (+ dx_add, nullable, same shape: gradient accumulation), dgamma/dbeta fp32 [C] (overwritten). */
size_t sdb_groupnorm_workspace_bytes(int B, int HW, int C);
int sdb_groupnorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, int B, int HW, int C,
                      float eps, int silu, void* workspace, size_t ws_bytes, cudaStream_t stream);
int sdb_groupnorm_bwd(const void* x, const void* dy, const float* stats, const float* gamma, const float* beta,
                      const void* dx_add, void* dx, float* dgamma, float* dbeta, int B, int HW, int C, int silu,
                      void* workspace, size_t ws_bytes, cudaStream_t stream);
/* LayerNorm over the last axis; stats fp32 [rows][2] = (mean, rstd). */
int sdb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, long long rows, int C,
                      float eps, cudaStream_t stream);
size_t sdb_layernorm_bwd_workspace_bytes(long long rows, int C);
int sdb_layernorm_bwd(const void* x, const void* dy, const float* stats, const float* gamma, const void* dx_add, void* dx,
                      float* dgamma, float* dbeta, long long rows, int C, void* workspace, size_t ws_bytes,
                      cudaStream_t stream);
/* GEGLU: u [rows, 2H] = [lin | gate] -> out [rows, H] = lin * gelu_tanh(gate); bwd writes du [rows, 2H]. */
int sdb_geglu_fwd(const void* u, void* out, long long rows, int H, cudaStream_t stream);
int sdb_geglu_bwd(const void* u, const void* dout, void* du, long long rows, int H, cudaStream_t stream);

/* ---- attention (diffusers-0.21.4 FlaxAttention + jax_memory_efficient_attention [upstream] as patched by
 * key_chunk_patch.patch:5-6). q/o/dO: [B, Sq, H*d] bf16 with row strides ld* (batch stride = S * ld: the buffers are
 * [B*S, ld] matrices, so column slices of fused projections work in place); k/v: [B, Skv, H*d]. lse2: fp32
 * [B][H][Sq], log2-domain log-sum-exp. kbias2 (nullable): fp32 [Skv] additive per-key bias in the log2 domain —
 * log2(multiplicity) reproduces the patched reference's clamped-chunk double count. `scale` multiplies
 * q.k (the reference pre-divides q by sqrt(d)). */
int sdb_attn_fwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv, void* o,
                 long long ldo, float* lse2, const float* kbias2, int B, int H, int Sq, int Skv, int d, float scale,
                 cudaStream_t stream);
size_t sdb_attn_bwd_workspace_bytes(int B, int H, int Sq);
int sdb_attn_bwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv, const void* o,
                 long long ldo, const void* dO, long long lddo, const float* lse2, const float* kbias2, void* dq,
                 long long lddq, void* dk, long long lddk, void* dv, long long lddv, int B, int H, int Sq, int Skv, int d,
                 float scale, void* workspace, size_t ws_bytes, cudaStream_t stream);

/* ---- text encoder (FlaxCLIPTextModel [transformers, third-party], loaded at training_utils.py:215-217, called at
 * 635-640; context assembly 641-674). GEMMs / LayerNorm / bias gradients reuse the entry points above. ----
 * Embeddings: out[r, :] = bf16(token[ids[r], :] + position[r % seq_len, :]) over bf16 tables; the backward zeroes
 * dtoken (vocab x hidden fp32) itself, adds the rows with fp32 atomics and writes dposition (fixed-order sums). */
int sdb_embed_fwd(const int* input_ids, const void* token_embedding_bf16, const void* position_embedding_bf16,
                  void* out_bf16, int rows, int seq_len, int hidden, int vocab, cudaStream_t stream);
int sdb_embed_bwd(const int* input_ids, const void* dout_bf16, float* dtoken_embedding, float* dposition_embedding,
                  int rows, int seq_len, int hidden, int vocab, cudaStream_t stream);
/* Attention over short sequences (S <= 128, d <= 128, d % 8 == 0; the backward keeps q, k, v, dO, P and dS of one head in
 * shared memory: (4 S (d + 1) + 2 S (S + 1)) * 4 B <= 227 KB, e.g. S = 77 with d = 64), optional causal mask (FlaxCLIPAttention): q / k / v /
 * o are [N * S, ld] bf16 matrices with the H heads side by side in the columns; lse fp32 [N][H][S] (natural log). */
int sdb_attn_small_fwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv, void* o,
                       long long ldo, float* lse, int N, int H, int S, int d, float scale, int causal,
                       cudaStream_t stream);
int sdb_attn_small_bwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                       const void* o, long long ldo, const void* dO, long long lddo, const float* lse, void* dq,
                       long long lddq, void* dk, long long lddk, void* dv, long long lddv, int N, int H, int S, int d,
                       float scale, int causal, cudaStream_t stream);
/* ACT2FN[hidden_act] of FlaxCLIPMLP: kind 0 = quick_gelu (z * sigmoid(1.702 z)), 1 = gelu (erf). bf16 in / out. */
int sdb_act_fwd(const void* x, void* y, long long n, int kind, cudaStream_t stream);
int sdb_act_bwd(const void* x, const void* dy, void* dx, long long n, int kind, cudaStream_t stream);
/* Context assembly (training_utils.py:641-674) as a row gather: scatter_out_rows == 0: out[r, :] = src[row_index[r], :]
 * for r < rows; scatter_out_rows > 0 (the adjoint): out is a [scatter_out_rows, cols] matrix that the call zeroes, then
 * out[row_index[r], :] = src[r, :] (every index used at most once). */
int sdb_gather_rows_bf16(const void* src, const int* row_index, void* out, long long rows, int cols,
                         long long scatter_out_rows, cudaStream_t stream);

/* ---- frozen VAE encode (diffusers-0.21.4 FlaxAutoencoderKL.encode [third-party], training_utils.py:574-586): convs /
 * GroupNorm reuse the entry points above; these are the extra pieces. */
int sdb_nchw_f32_to_nhwc_bf16_pad(const float* src, void* dst, int B, int C, int Cpad, int HW, cudaStream_t stream);
/* out[r, :cols] = softmax(scale * x[r, :cols]) (bf16 rows, row stride ld, fp32 statistics): the encoder mid block's
 * single-head attention over H*W tokens with head dim 512 runs as GEMM -> this -> GEMM. */
int sdb_softmax_rows_bf16(const void* x, void* out, long long rows, int cols, long long ld, float scale,
                          cudaStream_t stream);
/* FlaxDiagonalGaussianDistribution.sample + the 0.18215 scaling: latents NCHW fp32 = (mean + exp(0.5 clip(logvar, -30,
 * 20)) * eps) * scaling; moments NHWC bf16 [B, HW, 2 L] = (mean | logvar), eps fp32 NHWC [B, HW, L] (NULL: the mode). */
int sdb_vae_sample(const void* moments_nhwc_bf16, const float* eps_nhwc, float* latents_nchw, int B, int latent_channels,
                   int HW, float scaling, cudaStream_t stream);

/* ---- validation sampler (models/pipeline_flax_stable_diffusion.py:160-254): the elementwise steps between the UNet
 * forward, sdb_ddim_step and the VAE decode. The decoder itself reuses the conv / GroupNorm / GEMM entry points. */
/* out NHWC bf16 [repeat * B, HW, C] = bf16(latents NCHW fp32 [B, C, HW] * scale), the batch written `repeat` times:
 * `jnp.concatenate([latents] * 2)` + the UNet's entry transpose (211-224; repeat 2, scale 1), and `1 / scaling_factor *
 * latents` + FlaxAutoencoderKL.decode's transpose (246-249; repeat 1). */
int sdb_latents_to_nhwc_bf16(const float* latents_nchw, float scale, int repeat, void* out_nhwc_bf16, int B, int C, int HW,
                             cudaStream_t stream);
/* classifier-free guidance (226-227): pred NHWC bf16 [2B, HW, C] (uncond rows first) -> NCHW fp32 [B, C, HW] =
 * uncond + guidance_scale * bf16(text - uncond). */
int sdb_cfg_combine(const void* pred_nhwc_bf16, float guidance_scale, float* out_nchw, int B, int C, int HW,
                    cudaStream_t stream);
/* (image / 2 + 0.5).clip(0, 1).transpose(0, 2, 3, 1) (250) on the decoder's NHWC bf16 [npix, Cpad] output -> NHWC fp32
 * [npix, C] (bf16 arithmetic, as the reference's bf16 VAE computes it). */
int sdb_image_postprocess(const void* x_nhwc_bf16, float* out_nhwc, long long npix, int Cpad, int C, cudaStream_t stream);

/* ---- data movement / small elementwise pieces of the UNet graph (FlaxUpsample2D nearest resize, skip-connection
 * concatenate, FlaxTimesteps sinusoid, silu(temb), NCHW->NHWC entry transpose [upstream diffusers 0.21.4]) ---- */
int sdb_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream);
int sdb_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int HW, cudaStream_t stream);
int sdb_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int HW, cudaStream_t stream);
/* += 1 on up to three int32 device counters (NULL = skip): TrainState.step, ScaleBy8bitLionState.count
 * (lion_quant.py:151), ScaleByScheduleState.count — one launch instead of three framework kernels. */
int sdb_bump_counters(int* a, int* b, int* c, cudaStream_t stream);
int sdb_upsample2x_fwd(const void* x, void* y, int B, int H, int W, int C, cudaStream_t stream); /* H,W: input */
int sdb_upsample2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, cudaStream_t stream);
int sdb_copy2d_bf16(const void* src, long long lds, void* dst, long long ldd, long long rows, int cols, int accumulate,
                    cudaStream_t stream);
int sdb_add_bf16(const void* a, const void* b, void* out, long long n, cudaStream_t stream);

/* Epilogue of a split-K forward GEMM: out[r, c] = bf16(acc[r, c] + bias[c] + temb[r / rows_per_image, c] + residual[r, c])
 * (bias fp32 [N], temb bf16 [images, ldt], residual bf16 [rows, ldr]; each may be NULL). Replaces the fused epilogue of
 * sdb_gemm_tc for the low-resolution layers of the UNet (flax nn.Conv / nn.Dense bias add, the time-embedding add of
 * FlaxResnetBlock2D and the residual adds, diffusers 0.21.4 resnet_flax.py / attention_flax.py). */
int sdb_splitk_finish(const float* acc, long long lda, int nslabs, long long slab_stride, const float* bias,
                      const void* temb, long long ldt, long long rows_per_image, const void* residual, long long ldr,
                      void* out, long long ldo, long long rows, int N, cudaStream_t stream);
/* out[i] = slab_0[i] + slab_1[i] + ... (fixed order): the fold of a slab-mode split weight gradient (sdb_gemm_tc, WGRAD,
 * splitk_slab > 0); n elements per slab, n % 4 == 0. */
int sdb_splitk_finish_f32(const float* slabs, int nslabs, long long slab_stride, float* out, long long n,
                          cudaStream_t stream);
/* acc: nslabs fp32 [rows, lda] slabs slab_stride elements apart, summed in slab order (nslabs = 1: one buffer). */
size_t sdb_colsum_workspace_bytes(long long rows, int cols);
int sdb_colsum_bf16(const void* x, long long ld, int nseg, long long rows_per_seg, int cols, float* out_total,
                    void* out_seg_bf16, void* workspace, size_t ws_bytes, cudaStream_t stream);
/* column sums of a bf16 [nseg*rows_per_seg, cols] matrix: out_total[c] (fp32, nullable) = bias gradient;
 * out_seg[s][c] (bf16, nullable) = per-image sums = gradient of the broadcast time-embedding add. */
/* Allocation-free variant of sdb_colsum_bf16 (one segment, out_total only). It has NO counterpart in the reference: it was
 * added for callers that cannot allocate (CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). The partial sums go to a file-scope static __device__ array of
 * sdb_colsum_static_max_scratch_bytes() = 2 MiB that every GPU loading the library holds until exit; the call itself
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and returns an error (-3) for shapes whose
 * partial sums exceed the array (ceil(rows / rpb) * cols * 4 bytes, rpb in [32, 256] as sdb_colsum_bf16 picks it: e.g.
 * [32768, 2560] fits, [262144, 2560] does not). The array is shared by all calls on a device and keeps the last call's
 * contents (nothing needs clearing: every partial sum read was written by the same call): calls on one stream are
 * ordered; calls that overlap on the GPU (different streams) race on it. x bf16 [rows, cols], row stride ld (elements),
 * cols % 8 == 0, ld % 8 == 0, 16-byte aligned; out_total fp32 [cols]. */
size_t sdb_colsum_static_max_scratch_bytes(void);
int sdb_colsum_bf16_static(const void* x, long long ld, long long rows, int cols, float* out_total, cudaStream_t stream);
int sdb_silu_fwd(const void* x, void* y, long long n, cudaStream_t stream);
int sdb_silu_bwd(const void* x, const void* dy, void* dx, long long n, cudaStream_t stream);
int sdb_timestep_embedding(const int* timesteps, void* out, int B, int dim, int flip_sin_to_cos, float freq_shift,
                           cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SDB200_H */
