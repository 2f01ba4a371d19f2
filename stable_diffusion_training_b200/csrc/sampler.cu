// Elementwise pieces of the validation sampler: everything between the UNet forward, the DDIM step
// (sched_loss.cu: sdb_ddim_step) and the VAE decode.
//
// Replaces (reference): /root/reference/models/pipeline_flax_stable_diffusion.py:208-251 —
//   `latents_input = jnp.concatenate([latents] * 2)` + the UNet's NCHW -> NHWC entry transpose and bf16 cast  (211-224),
//   `noise_pred_uncond + guidance_scale * (noise_prediction_text - noise_pred_uncond)`                          (226-227),
//   `1 / scaling_factor * latents` before `vae.decode`                                                        (246-247),
//   `(image / 2 + 0.5).clip(0, 1).transpose(0, 2, 3, 1)`                                                       (250).
// All HBM-bound streaming work on a few hundred KB of latents (and B x 512 x 512 x 3 pixels at the end): one pass each,
// rounding points as the reference's dtypes put them (UNet / VAE outputs are bf16, the scheduler works in fp32).
#include "sdb_common.cuh"
#include "../../include/sdb200.h"

namespace {

__device__ __forceinline__ float bf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

int grid_for(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)sdb::kNumSMs * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// out NHWC bf16 [repeat * B, HW, C] = bf16(latents NCHW fp32 [B, C, HW] * scale), the batch written `repeat` times
__global__ void __launch_bounds__(256)
latents_to_nhwc_kernel(const float* __restrict__ lat, float scale, int repeat, __nv_bfloat16* __restrict__ out, int B, int C,
                       int HW) {
  sdb::pdl_trigger();
  const long long per = (long long)B * HW * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long bp = i / C;
    const int p = (int)(bp % HW);
    const int b = (int)(bp / HW);
    const __nv_bfloat16 v = __float2bfloat16_rn(__fmul_rn(lat[((long long)b * C + c) * HW + p], scale));
    for (int r = 0; r < repeat; ++r) out[r * per + i] = v;
  }
}

// classifier-free guidance: pred NHWC bf16 [2B, HW, C] (first B unconditional, last B text) -> NCHW fp32 [B, C, HW]
//   diff = bf16(text - uncond)   (a bf16 subtraction in the reference), out = f32(uncond) + guidance * f32(diff)
__global__ void __launch_bounds__(256)
cfg_combine_kernel(const __nv_bfloat16* __restrict__ pred, float guidance, float* __restrict__ out, int B, int C, int HW) {
  sdb::pdl_trigger();
  const long long per = (long long)B * HW * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(i % HW);
    const long long bc = i / HW;
    const int c = (int)(bc % C);
    const int b = (int)(bc / C);
    const long long src = ((long long)b * HW + p) * C + c;
    const float u = bf(pred + src), t = bf(pred + per + src);
    const float diff = round_bf16(__fsub_rn(t, u));
    out[i] = __fadd_rn(u, __fmul_rn(guidance, diff));
  }
}

// image NHWC fp32 [npix, C] = clip(bf16(bf16(x / 2) + 0.5), 0, 1) from the decoder's NHWC bf16 [npix, Cpad] output
__global__ void __launch_bounds__(256)
image_postprocess_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, long long npix, int Cpad, int C) {
  sdb::pdl_trigger();
  const long long total = npix * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long px = i / C;
    const int c = (int)(i - px * C);
    const float v = round_bf16(round_bf16(bf(x + px * Cpad + c) * 0.5f) + 0.5f);
    out[i] = fminf(fmaxf(v, 0.0f), 1.0f);
  }
}

}  // namespace

extern "C" int sdb_latents_to_nhwc_bf16(const float* latents_nchw, float scale, int repeat, void* out_nhwc_bf16, int B,
                                        int C, int HW, cudaStream_t stream) {
  SDB_CHECK_ARG(latents_nchw && out_nhwc_bf16, "sdb_latents_to_nhwc_bf16: null pointer");
  SDB_CHECK_ARG(B > 0 && C > 0 && HW > 0 && repeat >= 1, "sdb_latents_to_nhwc_bf16: bad shape");
  latents_to_nhwc_kernel<<<grid_for((long long)B * C * HW), 256, 0, stream>>>(latents_nchw, scale, repeat,
                                                                             (__nv_bfloat16*)out_nhwc_bf16, B, C, HW);
  SDB_CUDA_LAUNCH_CHECK("sdb_latents_to_nhwc_bf16");
  return SDB_OK;
}

extern "C" int sdb_cfg_combine(const void* pred_nhwc_bf16, float guidance_scale, float* out_nchw, int B, int C, int HW,
                               cudaStream_t stream) {
  SDB_CHECK_ARG(pred_nhwc_bf16 && out_nchw, "sdb_cfg_combine: null pointer");
  SDB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "sdb_cfg_combine: bad shape");
  cfg_combine_kernel<<<grid_for((long long)B * C * HW), 256, 0, stream>>>((const __nv_bfloat16*)pred_nhwc_bf16,
                                                                         guidance_scale, out_nchw, B, C, HW);
  SDB_CUDA_LAUNCH_CHECK("sdb_cfg_combine");
  return SDB_OK;
}

extern "C" int sdb_image_postprocess(const void* x_nhwc_bf16, float* out_nhwc, long long npix, int Cpad, int C,
                                     cudaStream_t stream) {
  SDB_CHECK_ARG(x_nhwc_bf16 && out_nhwc, "sdb_image_postprocess: null pointer");
  SDB_CHECK_ARG(npix > 0 && C > 0 && Cpad >= C, "sdb_image_postprocess: bad shape");
  image_postprocess_kernel<<<grid_for(npix * C), 256, 0, stream>>>((const __nv_bfloat16*)x_nhwc_bf16, out_nhwc, npix, Cpad,
                                                                  C);
  SDB_CUDA_LAUNCH_CHECK("sdb_image_postprocess");
  return SDB_OK;
}
