// 3x3 stride-1 pad-1 convolutions with a 4-channel side: conv_in (4 -> C) and conv_out (C -> 4) of the SD UNet
// (FlaxUNet2DConditionModel conv_in / conv_out [diffusers 0.21.4, un-vendored], reached from
// /root/reference/training_utils.py:678-684). K = 4 per tap (conv_in) or N = 4 (conv_out) cannot fill a tcgen05 tile
// (UMMA K = 16, N >= 16), and at 36 MACs per output element the layers are HBM/L2-bound anyway: CUDA-core kernels with
// 128-bit accesses, weights staged once per CTA in shared memory as fp32.
//
//   c4_to_wide   out[pix, n] = sum_tap sum_{k<4} a4[pix + tap, k] * Wt[tap][k][n]     conv_in fwd, conv_out dgrad
//   wide_to_c4   out[pix, j] = sum_tap sum_c   x[pix + tap, c]   * W[tap][c][j]       conv_out fwd
//   wgrad_c4     dW[tap][..] = sum_pix x[pix + tap, .] * dy[pix, .]                   conv_in wgrad, conv_out wgrad
#include "sdb_common.cuh"
#include "../../include/sdb200.h"

namespace {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = sdb::bf16_lo(v.x); f[1] = sdb::bf16_hi(v.x); f[2] = sdb::bf16_lo(v.y); f[3] = sdb::bf16_hi(v.y);
  f[4] = sdb::bf16_lo(v.z); f[5] = sdb::bf16_hi(v.z); f[6] = sdb::bf16_lo(v.w); f[7] = sdb::bf16_hi(v.w);
}
__device__ __forceinline__ void unpack4(const uint2& v, float* f) {
  f[0] = sdb::bf16_lo(v.x); f[1] = sdb::bf16_hi(v.x); f[2] = sdb::bf16_lo(v.y); f[3] = sdb::bf16_hi(v.y);
}

// ---------------------------------------------------------------------------------------------------------------
// out[pix, 8v..8v+7] for thread (pixel lane, v). smem: wt[9][4][C] fp32.
// w_tap / w_k / w_n: element strides of the bf16 weight tensor for (tap, k, n); flip: taps mirrored (dgrad).
__global__ void __launch_bounds__(256)
c4_to_wide_kernel(const __nv_bfloat16* __restrict__ a4, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
                  __nv_bfloat16* __restrict__ out, int B, int H, int W, int C, int w_tap, int w_k, int w_n, int flip) {
  sdb::pdl_trigger();
  extern __shared__ float s_w[];  // [9][4][C]
  for (int i = threadIdx.x; i < 36 * C; i += blockDim.x) {
    const int n = i % C, k = (i / C) & 3, tap = i / (4 * C);
    const int src_tap = flip ? 8 - tap : tap;
    s_w[i] = __bfloat162float(w[(long long)src_tap * w_tap + (long long)k * w_k + (long long)n * w_n]);
  }
  __syncthreads();
  const int nv = C / 8;
  const int lanes = blockDim.x / nv;
  const int v = threadIdx.x % nv, pl = threadIdx.x / nv;
  if (pl >= lanes) return;
  float bv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (bias != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) bv[i] = bias[v * 8 + i];
  }
  const long long npix = (long long)B * H * W;
  for (long long pix = (long long)blockIdx.x * lanes + pl; pix < npix; pix += (long long)gridDim.x * lanes) {
    const int x = (int)(pix % W), y = (int)((pix / W) % H);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = bv[i];
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      float a[4];
      unpack4(*reinterpret_cast<const uint2*>(a4 + (pix + (long long)(tap / 3 - 1) * W + (tap % 3 - 1)) * 4), a);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float4 w0 = *reinterpret_cast<const float4*>(&s_w[(tap * 4 + k) * C + v * 8]);
        const float4 w1 = *reinterpret_cast<const float4*>(&s_w[(tap * 4 + k) * C + v * 8 + 4]);
        acc[0] = fmaf(a[k], w0.x, acc[0]); acc[1] = fmaf(a[k], w0.y, acc[1]);
        acc[2] = fmaf(a[k], w0.z, acc[2]); acc[3] = fmaf(a[k], w0.w, acc[3]);
        acc[4] = fmaf(a[k], w1.x, acc[4]); acc[5] = fmaf(a[k], w1.y, acc[5]);
        acc[6] = fmaf(a[k], w1.z, acc[6]); acc[7] = fmaf(a[k], w1.w, acc[7]);
      }
    }
    uint4 o;
    o.x = sdb::pack_bf16x2(acc[0], acc[1]); o.y = sdb::pack_bf16x2(acc[2], acc[3]);
    o.z = sdb::pack_bf16x2(acc[4], acc[5]); o.w = sdb::pack_bf16x2(acc[6], acc[7]);
    sdb::st_u4(out + pix * C + v * 8, o);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// conv_out forward: one warp per pixel, lanes over 8-channel vectors.
__global__ void __launch_bounds__(256)
wide_to_c4_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
                  __nv_bfloat16* __restrict__ out, int B, int H, int W, int C) {
  sdb::pdl_trigger();
  // smem weights as [tap][i = channel % 8][v = channel / 8][4]: the lanes of a warp (consecutive v) read consecutive
  // float4 -> conflict-free LDS.128
  extern __shared__ float s_w[];
  const int nv = C / 8;
  for (int i = threadIdx.x; i < 36 * C; i += blockDim.x) {
    const int j = i & 3, c = (i >> 2) % C, tap = i / (4 * C);
    s_w[(((tap * 8 + (c & 7)) * nv) + (c >> 3)) * 4 + j] = __bfloat162float(w[i]);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long npix = (long long)B * H * W;
  for (long long pix = (long long)blockIdx.x * 8 + warp; pix < npix; pix += (long long)gridDim.x * 8) {
    const int px = (int)(pix % W), py = (int)((pix / W) % H);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = py + tap / 3 - 1, xx = px + tap % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const __nv_bfloat16* row = x + (pix + (long long)(tap / 3 - 1) * W + (tap % 3 - 1)) * C;
      for (int v = lane; v < nv; v += 32) {
        float xv[8];
        unpack8(sdb::ld_u4(row + v * 8), xv);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 wv = *reinterpret_cast<const float4*>(&s_w[((tap * 8 + i) * nv + v) * 4]);
          acc[0] = fmaf(xv[i], wv.x, acc[0]); acc[1] = fmaf(xv[i], wv.y, acc[1]);
          acc[2] = fmaf(xv[i], wv.z, acc[2]); acc[3] = fmaf(xv[i], wv.w, acc[3]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[j] = sdb::warp_sum(acc[j]);
    if (lane == 0) {
      if (bias != nullptr) { acc[0] += bias[0]; acc[1] += bias[1]; acc[2] += bias[2]; acc[3] += bias[3]; }
      uint2 o;
      o.x = sdb::pack_bf16x2(acc[0], acc[1]);
      o.y = sdb::pack_bf16x2(acc[2], acc[3]);
      *reinterpret_cast<uint2*>(out + pix * 4) = o;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// wgrad: thread (tap, v) accumulates dW for 8 wide channels x 4 narrow channels over this CTA's pixel chunk, then
// red.add into the zeroed fp32 output.
//   C4_IS_INPUT = true  (conv_in):  x4 [pix,4] shifted by the tap, dy [pix,C];  dW[tap][k][8v+i]  (HWIO: [9][4][C])
//   C4_IS_INPUT = false (conv_out): x [pix,C] shifted by the tap, dy4 [pix,4]; dW[tap][8v+i][j]  (HWIO: [9][C][4])
template <bool C4_IS_INPUT>
__global__ void __launch_bounds__(1024)
wgrad_c4_kernel(const __nv_bfloat16* __restrict__ wide, const __nv_bfloat16* __restrict__ c4, float* __restrict__ dw, int B,
                int H, int W, int C, long long pix_per_cta, float* __restrict__ partials) {
  sdb::pdl_trigger();
  const int nv = C / 8;
  const int tap = threadIdx.x / nv, v = threadIdx.x % nv;
  if (tap >= 9) return;
  const int dy_ = tap / 3 - 1, dx_ = tap % 3 - 1;
  const long long npix = (long long)B * H * W;
  const long long p0 = (long long)blockIdx.x * pix_per_cta;
  const long long p1 = min(npix, p0 + pix_per_cta);
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (long long pix = p0; pix < p1; ++pix) {
    const int x = (int)(pix % W), y = (int)((pix / W) % H);
    const int yy = y + dy_, xx = x + dx_;
    if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
    const long long spix = pix + (long long)dy_ * W + dx_;
    float wv[8], cv[4];
    unpack8(sdb::ld_u4(wide + (C4_IS_INPUT ? pix : spix) * C + v * 8), wv);
    unpack4(*reinterpret_cast<const uint2*>(c4 + (C4_IS_INPUT ? spix : pix) * 4), cv);
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i], cv[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long o = C4_IS_INPUT ? ((long long)tap * 4 + j) * C + v * 8 + i : ((long long)tap * C + v * 8 + i) * 4 + j;
      if (partials != nullptr)  // this CTA's own dW-shaped slab, folded afterwards in CTA order (deterministic)
        partials[(long long)blockIdx.x * 36 * C + o] = acc[i][j];
      else
        asm volatile("red.global.add.f32 [%0], %1;" ::"l"(dw + o), "f"(acc[i][j]) : "memory");
    }
}

// dw[i] = sum over CTAs, in CTA order, of partials[cta][i]
__global__ void __launch_bounds__(256) wgrad_c4_fold_kernel(const float* __restrict__ partials, int nparts, int n,
                                                            float* __restrict__ dw) {
  sdb::pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float a = 0.f;
  for (int s = 0; s < nparts; ++s) a += partials[(long long)s * n + i];
  dw[i] = a;
}

int check_c4(const char* name, int B, int H, int W, int C) {
  if (B <= 0 || H <= 0 || W <= 0 || C < 8 || C % 8 != 0 || C > 904)
    return sdb_set_error(SDB_ERR_INVALID, "%s: need B,H,W > 0 and C a multiple of 8 in [8, 904] (got %d,%d,%d,%d)", name,
                         B, H, W, C);
  return SDB_OK;
}

template <typename K>
int ensure_smem(K kernel, size_t bytes, const char* name) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e));
  }
  return SDB_OK;
}

}  // namespace

extern "C" int sdb_conv3x3_c4in_fwd(const void* x4, const void* w, const float* bias, void* y, int B, int H, int W, int C,
                                    cudaStream_t stream) {
  int rc = check_c4("sdb_conv3x3_c4in_fwd", B, H, W, C);
  if (rc) return rc;
  const size_t sm = (size_t)36 * C * sizeof(float);
  if ((rc = ensure_smem(c4_to_wide_kernel, sm, "sdb_conv3x3_c4in_fwd"))) return rc;
  const int lanes = 256 / (C / 8);
  const long long npix = (long long)B * H * W;
  long long blocks = (npix + lanes - 1) / lanes;
  if (blocks > 2 * sdb::kNumSMs) blocks = 2 * sdb::kNumSMs;
  // HWIO [3][3][4][C]: tap stride 4C, k stride C, n stride 1
  c4_to_wide_kernel<<<(int)blocks, 256, sm, stream>>>((const __nv_bfloat16*)x4, (const __nv_bfloat16*)w, bias,
                                                     (__nv_bfloat16*)y, B, H, W, C, 4 * C, C, 1, 0);
  SDB_CUDA_LAUNCH_CHECK("sdb_conv3x3_c4in_fwd");
  return SDB_OK;
}

extern "C" int sdb_conv3x3_c4out_dgrad(const void* dy4, const void* w, void* dx, int B, int H, int W, int C,
                                       cudaStream_t stream) {
  int rc = check_c4("sdb_conv3x3_c4out_dgrad", B, H, W, C);
  if (rc) return rc;
  const size_t sm = (size_t)36 * C * sizeof(float);
  if ((rc = ensure_smem(c4_to_wide_kernel, sm, "sdb_conv3x3_c4out_dgrad"))) return rc;
  const int lanes = 256 / (C / 8);
  const long long npix = (long long)B * H * W;
  long long blocks = (npix + lanes - 1) / lanes;
  if (blocks > 2 * sdb::kNumSMs) blocks = 2 * sdb::kNumSMs;
  // HWIO [3][3][C][4] read transposed with mirrored taps: tap stride 4C, k (= cout) stride 1, n (= cin) stride 4
  c4_to_wide_kernel<<<(int)blocks, 256, sm, stream>>>((const __nv_bfloat16*)dy4, (const __nv_bfloat16*)w, nullptr,
                                                     (__nv_bfloat16*)dx, B, H, W, C, 4 * C, 1, 4, 1);
  SDB_CUDA_LAUNCH_CHECK("sdb_conv3x3_c4out_dgrad");
  return SDB_OK;
}

extern "C" int sdb_conv3x3_c4out_fwd(const void* x, const void* w, const float* bias, void* y4, int B, int H, int W, int C,
                                     cudaStream_t stream) {
  int rc = check_c4("sdb_conv3x3_c4out_fwd", B, H, W, C);
  if (rc) return rc;
  const size_t sm = (size_t)36 * C * sizeof(float);
  if ((rc = ensure_smem(wide_to_c4_kernel, sm, "sdb_conv3x3_c4out_fwd"))) return rc;
  const long long npix = (long long)B * H * W;
  long long blocks = (npix + 7) / 8;
  if (blocks > 4 * sdb::kNumSMs) blocks = 4 * sdb::kNumSMs;
  wide_to_c4_kernel<<<(int)blocks, 256, sm, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, bias,
                                                     (__nv_bfloat16*)y4, B, H, W, C);
  SDB_CUDA_LAUNCH_CHECK("sdb_conv3x3_c4out_fwd");
  return SDB_OK;
}

extern "C" size_t sdb_conv3x3_c4_wgrad_workspace_bytes(int C) { return (size_t)2 * sdb::kNumSMs * 36 * C * sizeof(float); }

extern "C" int sdb_conv3x3_c4_wgrad(const void* wide, const void* c4, float* dw, int B, int H, int W, int C,
                                    int c4_is_input, void* workspace, size_t ws_bytes, cudaStream_t stream) {
  int rc = check_c4("sdb_conv3x3_c4_wgrad", B, H, W, C);
  if (rc) return rc;
  float* partials = reinterpret_cast<float*>(workspace);
  if (partials != nullptr && ws_bytes < sdb_conv3x3_c4_wgrad_workspace_bytes(C))
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_conv3x3_c4_wgrad: workspace %zu < %zu", ws_bytes,
                         sdb_conv3x3_c4_wgrad_workspace_bytes(C));
  if (partials == nullptr) {  // no workspace: the CTAs accumulate into the zeroed dw with red.global.add (order varies)
    cudaError_t e = cudaMemsetAsync(dw, 0, (size_t)36 * C * sizeof(float), stream);
    if (e != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_conv3x3_c4_wgrad: memset: %s", cudaGetErrorString(e));
  }
  const long long npix = (long long)B * H * W;
  int blocks = 2 * sdb::kNumSMs;
  if (blocks > npix) blocks = (int)npix;
  const long long ppc = (npix + blocks - 1) / blocks;
  blocks = (int)((npix + ppc - 1) / ppc);
  const int threads = ((9 * (C / 8) + 31) / 32) * 32;
  if (c4_is_input)
    wgrad_c4_kernel<true><<<blocks, threads, 0, stream>>>((const __nv_bfloat16*)wide, (const __nv_bfloat16*)c4, dw, B, H, W,
                                                         C, ppc, partials);
  else
    wgrad_c4_kernel<false><<<blocks, threads, 0, stream>>>((const __nv_bfloat16*)wide, (const __nv_bfloat16*)c4, dw, B, H,
                                                          W, C, ppc, partials);
  SDB_CUDA_LAUNCH_CHECK("sdb_conv3x3_c4_wgrad");
  if (partials != nullptr) {
    wgrad_c4_fold_kernel<<<(36 * C + 255) / 256, 256, 0, stream>>>(partials, blocks, 36 * C, dw);
    SDB_CUDA_LAUNCH_CHECK("sdb_conv3x3_c4_wgrad(fold)");
  }
  return SDB_OK;
}
