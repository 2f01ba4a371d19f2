// Forward-diffusion noise mix, velocity target, DDPM reverse step and the MSE loss (+ its gradient).
//
// Replaces (reference): schedulers/scheduling_utils_flax.py:316-343 (get_sqrt_alpha_prod, add_noise_common,
// get_velocity_common), schedulers/scheduling_ddpm_flax.py:196-279 (step), training_utils.py:688-709 (target + MSE,
// min-SNR weight 546-568).
//
// add_noise / get_velocity / ddpm_step are bit-exact against oracle/sched.py: individually rounded fp32 ops in the
// reference's order (no FMA contraction), sqrt = IEEE correctly rounded.
#include "sdb_common.cuh"
#include "../../include/sdb200.h"

namespace {

// jnp `table[t]`: a negative index wraps once, then the index is clamped into the table (jax gather semantics: an
// out-of-range timestep never reads outside alphas_cumprod).
__device__ __forceinline__ int gather_index(int t, int n) {
  if (t < 0) t += n;
  return min(max(t, 0), n - 1);
}

// mode 0: out = a*x0 + s*noise (add_noise); mode 1: out = a*noise - s*x0 (velocity)
template <int MODE>
__global__ void __launch_bounds__(256)
noise_mix_kernel(const float* __restrict__ x0, const float* __restrict__ noise, const int* __restrict__ timesteps,
                 const float* __restrict__ alphas_cumprod, int num_timesteps, float* __restrict__ out,
                 int per_sample_vec4, long long total_vec4) {
  sdb::pdl_trigger();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < total_vec4; v += stride) {
    const int b = (int)(v / per_sample_vec4);
    const float ac = __ldg(alphas_cumprod + gather_index(__ldg(timesteps + b), num_timesteps));
    const float a = __fsqrt_rn(ac);
    const float s = __fsqrt_rn(__fsub_rn(1.0f, ac));
    const float4 x = sdb::ld_f4(x0 + 4 * v);
    const float4 e = sdb::ld_f4(noise + 4 * v);
    float4 o;
    if (MODE == 0) {
      o.x = __fadd_rn(__fmul_rn(a, x.x), __fmul_rn(s, e.x));
      o.y = __fadd_rn(__fmul_rn(a, x.y), __fmul_rn(s, e.y));
      o.z = __fadd_rn(__fmul_rn(a, x.z), __fmul_rn(s, e.z));
      o.w = __fadd_rn(__fmul_rn(a, x.w), __fmul_rn(s, e.w));
    } else {
      o.x = __fsub_rn(__fmul_rn(a, e.x), __fmul_rn(s, x.x));
      o.y = __fsub_rn(__fmul_rn(a, e.y), __fmul_rn(s, x.y));
      o.z = __fsub_rn(__fmul_rn(a, e.z), __fmul_rn(s, x.z));
      o.w = __fsub_rn(__fmul_rn(a, e.w), __fmul_rn(s, x.w));
    }
    sdb::st_f4(out + 4 * v, o);
  }
}

struct DdpmCoef {
  float sqrt_alpha_prod, sqrt_beta_prod, x0_coeff, sample_coeff, sigma;
  int pred_type;  // 0 epsilon, 1 sample, 2 v_prediction
  int clip_sample;
};

__global__ void __launch_bounds__(256)
ddpm_step_kernel(const float* __restrict__ model_out, const float* __restrict__ sample,
                 const float* __restrict__ noise, float* __restrict__ prev, long long n, DdpmCoef c) {
  sdb::pdl_trigger();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float mo = model_out[i], x = sample[i];
    float x0;
    if (c.pred_type == 0)
      x0 = __fdiv_rn(__fsub_rn(x, __fmul_rn(c.sqrt_beta_prod, mo)), c.sqrt_alpha_prod);
    else if (c.pred_type == 1)
      x0 = mo;
    else
      x0 = __fsub_rn(__fmul_rn(c.sqrt_alpha_prod, x), __fmul_rn(c.sqrt_beta_prod, mo));
    if (c.clip_sample) x0 = fminf(fmaxf(x0, -1.0f), 1.0f);
    float p = __fadd_rn(__fmul_rn(c.x0_coeff, x0), __fmul_rn(c.sample_coeff, x));
    float var = (noise != nullptr) ? __fmul_rn(c.sigma, noise[i]) : 0.0f;
    prev[i] = __fadd_rn(p, var);
  }
}

// MSE over pred (NHWC bf16, the UNet output) vs target (NCHW fp32). Writes dpred (NHWC bf16) and per-block partial
// sums of weight_b * (target - pred)^2; a second tiny kernel reduces them in a fixed order -> deterministic loss.
constexpr int kMseBlocks = 148 * 2;
__global__ void __launch_bounds__(256)
mse_kernel(const __nv_bfloat16* __restrict__ pred, const float* __restrict__ target,
           const float* __restrict__ sample_weight, __nv_bfloat16* __restrict__ dpred, double* __restrict__ partials,
           int B, int C, int HW, float grad_scale) {
  sdb::pdl_trigger();
  const long long total = (long long)B * C * HW;
  const long long stride = (long long)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    // i indexes NHWC: ((b*HW + p)*C + c)
    const int c = (int)(i % C);
    const long long bp = i / C;
    const int p = (int)(bp % HW);
    const int b = (int)(bp / HW);
    const float t = target[((long long)b * C + c) * HW + p];
    const float y = __bfloat162float(pred[i]);
    const float w = sample_weight ? sample_weight[b] : 1.0f;
    const float d = t - y;
    acc += (double)(w * d * d);
    if (dpred) dpred[i] = __float2bfloat16_rn(-2.0f * w * d * grad_scale);
  }
  acc = sdb::warp_sum(acc);
  __shared__ double s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += s[i];
    partials[blockIdx.x] = t;
  }
}
__global__ void mse_finalize_kernel(const double* __restrict__ partials, int n, double inv_count, float* loss) {
  sdb::pdl_trigger();
  __shared__ double s[256];
  double t = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) t += partials[i];
  s[threadIdx.x] = t;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = (float)(s[0] * inv_count);
}

int grid_cap(long long work, int threads, int per_sm) {
  long long blocks = (work + threads - 1) / threads;
  long long cap = (long long)sdb::kNumSMs * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

static int noise_mix(int mode, const float* x0, const float* noise, const int* timesteps, const float* ac,
                     int num_timesteps, float* out, int batch, long long per_sample, cudaStream_t stream,
                     const char* name) {
  SDB_CHECK_ARG(num_timesteps > 0, "%s: num_timesteps must be positive", name);
  SDB_CHECK_ARG(batch > 0 && per_sample > 0 && per_sample % 4 == 0, "%s: per-sample size (%lld) must be a multiple of 4",
                name, per_sample);
  SDB_CHECK_ARG((((uintptr_t)x0 | (uintptr_t)noise | (uintptr_t)out) & 15) == 0, "%s: 16-byte alignment required", name);
  const long long total_vec4 = (long long)batch * per_sample / 4;
  const int grid = grid_cap(total_vec4, 256, 8);
  if (mode == 0)
    noise_mix_kernel<0><<<grid, 256, 0, stream>>>(x0, noise, timesteps, ac, num_timesteps, out, (int)(per_sample / 4), total_vec4);
  else
    noise_mix_kernel<1><<<grid, 256, 0, stream>>>(x0, noise, timesteps, ac, num_timesteps, out, (int)(per_sample / 4), total_vec4);
  SDB_CUDA_LAUNCH_CHECK(name);
  return SDB_OK;
}

extern "C" int sdb_add_noise(const float* original_samples, const float* noise, const int* timesteps,
                             const float* alphas_cumprod, int num_timesteps, float* noisy, int batch,
                             long long per_sample, cudaStream_t stream) {
  return noise_mix(0, original_samples, noise, timesteps, alphas_cumprod, num_timesteps, noisy, batch, per_sample,
                   stream, "sdb_add_noise");
}

extern "C" int sdb_get_velocity(const float* sample, const float* noise, const int* timesteps,
                                const float* alphas_cumprod, int num_timesteps, float* velocity, int batch,
                                long long per_sample, cudaStream_t stream) {
  return noise_mix(1, sample, noise, timesteps, alphas_cumprod, num_timesteps, velocity, batch, per_sample, stream,
                   "sdb_get_velocity");
}

// min_snr_gamma_loss_rescale (training_utils.py:546-568): snr = ac / (1 - ac); w = min(snr, gamma) / snr
// (epsilon) or / (snr + 1) (v_prediction). fp32, IEEE division, one weight per sample; stays on the device.
namespace {
__global__ void min_snr_weight_kernel(const float* __restrict__ alphas_cumprod, int num_timesteps,
                                      const int* __restrict__ timesteps, int batch, float gamma, int v_pred,
                                      float* __restrict__ weight) {
  sdb::pdl_trigger();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const float ac = alphas_cumprod[gather_index(timesteps[b], num_timesteps)];
  const float snr = __fdiv_rn(ac, __fsub_rn(1.0f, ac));
  const float m = fminf(snr, gamma);
  weight[b] = v_pred ? __fdiv_rn(m, __fadd_rn(snr, 1.0f)) : __fdiv_rn(m, snr);
}
}  // namespace

extern "C" int sdb_min_snr_weight(const float* alphas_cumprod, int num_timesteps, const int* timesteps, int batch,
                                  float gamma, int v_prediction, float* weight, cudaStream_t stream) {
  SDB_CHECK_ARG(batch > 0 && num_timesteps > 0, "sdb_min_snr_weight: bad sizes");
  min_snr_weight_kernel<<<(batch + 127) / 128, 128, 0, stream>>>(alphas_cumprod, num_timesteps, timesteps, batch, gamma,
                                                                v_prediction, weight);
  SDB_CUDA_LAUNCH_CHECK("sdb_min_snr_weight");
  return SDB_OK;
}

extern "C" int sdb_ddpm_step(const float* model_output, const float* sample, const float* noise, float* prev_sample,
                             long long n, float sqrt_alpha_prod, float sqrt_beta_prod, float pred_original_coeff,
                             float current_sample_coeff, float sigma, int prediction_type, int clip_sample,
                             cudaStream_t stream) {
  SDB_CHECK_ARG(prediction_type >= 0 && prediction_type <= 2, "sdb_ddpm_step: prediction_type must be 0/1/2");
  SDB_CHECK_ARG(n > 0, "sdb_ddpm_step: n must be positive");
  DdpmCoef c{sqrt_alpha_prod, sqrt_beta_prod, pred_original_coeff, current_sample_coeff, sigma, prediction_type,
             clip_sample};
  ddpm_step_kernel<<<grid_cap(n, 256, 8), 256, 0, stream>>>(model_output, sample, noise, prev_sample, n, c);
  SDB_CUDA_LAUNCH_CHECK("sdb_ddpm_step");
  return SDB_OK;
}

// FlaxDDIMScheduler.step (schedulers/scheduling_ddim_flax.py:199-284), elementwise part; the scalars (alpha / beta products,
// eta-scaled standard deviation, direction coefficient) are computed by the host in fp32 in the reference's order.
namespace {
__global__ void __launch_bounds__(256)
ddim_step_kernel(const float* __restrict__ model_out, const float* __restrict__ sample, float* __restrict__ prev,
                 long long n, float sa, float sb, float sa_prev, float cdir, int pred_type) {
  sdb::pdl_trigger();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float mo = model_out[i], x = sample[i];
    float x0, eps;
    if (pred_type == 0) {
      x0 = __fdiv_rn(__fsub_rn(x, __fmul_rn(sb, mo)), sa);
      eps = mo;
    } else if (pred_type == 1) {
      x0 = mo;
      eps = __fdiv_rn(__fsub_rn(x, __fmul_rn(sa, x0)), sb);
    } else {
      x0 = __fsub_rn(__fmul_rn(sa, x), __fmul_rn(sb, mo));
      eps = __fadd_rn(__fmul_rn(sa, mo), __fmul_rn(sb, x));
    }
    prev[i] = __fadd_rn(__fmul_rn(sa_prev, x0), __fmul_rn(cdir, eps));
  }
}
}  // namespace

extern "C" int sdb_ddim_step(const float* model_output, const float* sample, float* prev_sample, long long n,
                             float sqrt_alpha_prod, float sqrt_beta_prod, float sqrt_alpha_prod_prev,
                             float direction_coeff, int prediction_type, cudaStream_t stream) {
  SDB_CHECK_ARG(prediction_type >= 0 && prediction_type <= 2, "sdb_ddim_step: prediction_type must be 0/1/2");
  SDB_CHECK_ARG(n > 0, "sdb_ddim_step: n must be positive");
  ddim_step_kernel<<<grid_cap(n, 256, 8), 256, 0, stream>>>(model_output, sample, prev_sample, n, sqrt_alpha_prod,
                                                           sqrt_beta_prod, sqrt_alpha_prod_prev, direction_coeff,
                                                           prediction_type);
  SDB_CUDA_LAUNCH_CHECK("sdb_ddim_step");
  return SDB_OK;
}

extern "C" size_t sdb_mse_workspace_bytes(void) { return sizeof(double) * kMseBlocks; }

extern "C" int sdb_mse_fwd_bwd(const void* pred_nhwc_bf16, const float* target_nchw, const float* sample_weight,
                               void* dpred_nhwc_bf16, float* loss, int B, int C, int HW, float loss_grad_scale,
                               void* workspace, size_t ws_bytes, cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "sdb_mse_fwd_bwd: bad shape");
  if (ws_bytes < sdb_mse_workspace_bytes())
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_mse_fwd_bwd: workspace %zu < %zu", ws_bytes, sdb_mse_workspace_bytes());
  const long long total = (long long)B * C * HW;
  const float gscale = loss_grad_scale / (float)total;  // d(mean)/d(elem)
  const int grid = grid_cap(total, 256, 2);
  mse_kernel<<<grid, 256, 0, stream>>>((const __nv_bfloat16*)pred_nhwc_bf16, target_nchw, sample_weight,
                                       (__nv_bfloat16*)dpred_nhwc_bf16, (double*)workspace, B, C, HW, gscale);
  SDB_CUDA_LAUNCH_CHECK("sdb_mse_fwd_bwd");
  mse_finalize_kernel<<<1, 256, 0, stream>>>((const double*)workspace, grid, 1.0 / (double)total, loss);
  SDB_CUDA_LAUNCH_CHECK("sdb_mse_fwd_bwd(finalize)");
  return SDB_OK;
}
