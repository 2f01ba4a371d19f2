// Data-movement and small elementwise kernels of the UNet graph: layout casts, nearest 2x upsample (+ its adjoint),
// strided 2-D copies (skip-connection concat / split), gradient accumulation, bias gradients (column sums), SiLU on
// the time embedding and the sinusoidal timestep embedding.
//
// Replaces [upstream diffusers 0.21.4 Flax, un-vendored]: FlaxUpsample2D (jax.image.resize nearest),
// jnp.concatenate of skip connections in FlaxCrossAttnUpBlock2D / FlaxUpBlock2D, FlaxTimesteps /
// get_sinusoidal_embeddings (embeddings_flax.py), nn.silu on temb in FlaxResnetBlock2D, and the NCHW->NHWC transpose
// at the top of FlaxUNet2DConditionModel.__call__ (call site /root/reference/training_utils.py:678-684).
#include "sdb_common.cuh"
#include "../../include/sdb200.h"
#include <atomic>
#include <mutex>

namespace {

int ew_grid(long long work, int per_sm = 8) {
  long long blocks = (work + 255) / 256;
  long long cap = (long long)sdb::kNumSMs * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = sdb::bf16_lo(v.x); f[1] = sdb::bf16_hi(v.x); f[2] = sdb::bf16_lo(v.y); f[3] = sdb::bf16_hi(v.y);
  f[4] = sdb::bf16_lo(v.z); f[5] = sdb::bf16_hi(v.z); f[6] = sdb::bf16_lo(v.w); f[7] = sdb::bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 o;
  o.x = sdb::pack_bf16x2(f[0], f[1]); o.y = sdb::pack_bf16x2(f[2], f[3]);
  o.z = sdb::pack_bf16x2(f[4], f[5]); o.w = sdb::pack_bf16x2(f[6], f[7]);
  return o;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  sdb::pdl_trigger();
  const long long n4 = n / 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n4; v += stride) {
    const float4 a = sdb::ld_f4(src + 4 * v);
    uint2 o;
    o.x = sdb::pack_bf16x2(a.x, a.y);
    o.y = sdb::pack_bf16x2(a.z, a.w);
    *reinterpret_cast<uint2*>(dst + 4 * v) = o;
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// src NCHW fp32 [B,C,HW] -> dst NHWC bf16 [B,HW,C] (C tiny: 4)
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int B, int C,
                                    int HW) {
  sdb::pdl_trigger();
  const long long total = (long long)B * C * HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long bp = i / C;
    const int p = (int)(bp % HW);
    const int b = (int)(bp / HW);
    dst[i] = __float2bfloat16_rn(src[((long long)b * C + c) * HW + p]);
  }
}

// y[b, 2h+i, 2w+j, :] = x[b, h, w, :]
__global__ void upsample2x_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int B, int H, int W, int CV) {
  sdb::pdl_trigger();
  const long long total = (long long)B * 2 * H * 2 * W * CV;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % CV);
    long long r = i / CV;
    const int ow = (int)(r % (2 * W)); r /= 2 * W;
    const int oh = (int)(r % (2 * H));
    const int b = (int)(r / (2 * H));
    y[i] = x[(((long long)b * H + oh / 2) * W + ow / 2) * CV + v];
  }
}
// dx[b,h,w,:] = sum_{i,j} dy[b,2h+i,2w+j,:]
__global__ void upsample2x_bwd_kernel(const uint4* __restrict__ dy, uint4* __restrict__ dx, int B, int H, int W,
                                      int CV) {
  sdb::pdl_trigger();
  const long long total = (long long)B * H * W * CV;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % CV);
    long long r = i / CV;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int b = (int)(r / H);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int di = 0; di < 2; ++di)
#pragma unroll
      for (int dj = 0; dj < 2; ++dj) {
        float f[8];
        unpack8(dy[(((long long)b * 2 * H + 2 * h + di) * 2 * W + 2 * w + dj) * CV + v], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
    dx[i] = pack8(acc);
  }
}

// dst[r, 0:cols] (=|+=) src[r, 0:cols], bf16, row strides in elements (multiples of 8)
template <bool ACC>
__global__ void copy2d_kernel(const __nv_bfloat16* __restrict__ src, long long lds, __nv_bfloat16* __restrict__ dst,
                              long long ldd, long long rows, int cv) {
  sdb::pdl_trigger();
  const long long total = rows * cv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cv;
    const int v = (int)(i - r * cv);
    const uint4 a = sdb::ld_u4(src + r * lds + v * 8);
    if (!ACC) {
      sdb::st_u4(dst + r * ldd + v * 8, a);
    } else {
      float fa[8], fb[8];
      unpack8(a, fa);
      unpack8(sdb::ld_u4(dst + r * ldd + v * 8), fb);
#pragma unroll
      for (int k = 0; k < 8; ++k) fb[k] += fa[k];
      sdb::st_u4(dst + r * ldd + v * 8, pack8(fb));
    }
  }
}

__global__ void add_bf16_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ out,
                                long long nv) {
  sdb::pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float fa[8], fb[8];
    unpack8(a[i], fa);
    unpack8(b[i], fb);
#pragma unroll
    for (int k = 0; k < 8; ++k) fa[k] += fb[k];
    out[i] = pack8(fa);
  }
}

constexpr int kColsumMinRows = 32;
// column sums of a bf16 [nseg * rows_per_seg, cols] matrix (row stride ld), per segment and in total:
// stage 1 -> part[seg][block][cols]; stage 2 -> out_total[c] (fp32) and/or out_seg[seg][c] (bf16).
// grid: x = seg * bps + row block (rpb rows each), y = 256-column chunk; rpb is chosen by the host so that the grid
// fills the 148 SMs even for short, wide matrices (2048 x 10240 FF gradients).
// Arrival counters of the single-launch column sum: one per 256-column chunk, zero at rest (the last block through
// resets its counter), one slot per launch in flight (main and parameter-gradient streams run column sums concurrently).
constexpr int kColsumSlots = 256;
constexpr int kColsumMaxChunks = 64;
__device__ unsigned int g_colsum_sync[kColsumSlots * kColsumMaxChunks];

// The block that arrives LAST for its column chunk folds all partials of the chunk in a fixed order (deterministic) and
// writes the results — the former second launch (colsum_final_kernel: 185 dependent launches of ~4.6 us per step).
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, long long ld, long long rows_per_seg, int bps, int rpb,
                      int cols, float* __restrict__ part, int nseg, float* __restrict__ out_total,
                      __nv_bfloat16* __restrict__ out_seg, unsigned int* sync) {
  sdb::pdl_trigger();
  // thread (ty, tx): tx over 8-wide column vectors (32 per chunk), ty over 8 row lanes
  __shared__ float s[8][256 + 8];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int seg = blockIdx.x / bps, blk = blockIdx.x - seg * bps;
  const long long r0 = seg * rows_per_seg + (long long)blk * rpb;
  const long long r1 = min((seg + 1) * rows_per_seg, r0 + rpb);
  const int c0 = blockIdx.y * 256;
  const int c = c0 + tx * 8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < cols) {
    long long r = r0 + ty;
    for (; r + 24 < r1; r += 32) {  // 4 independent 128-bit loads in flight per thread
      const uint4 v0 = sdb::ld_u4(x + r * ld + c), v1 = sdb::ld_u4(x + (r + 8) * ld + c);
      const uint4 v2 = sdb::ld_u4(x + (r + 16) * ld + c), v3 = sdb::ld_u4(x + (r + 24) * ld + c);
      float f[8];
      unpack8(v0, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(v1, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(v2, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(v3, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
    for (; r < r1; r += 8) {
      float f[8];
      unpack8(sdb::ld_u4(x + r * ld + c), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) s[ty][tx * 8 + k] = acc[k];
  __syncthreads();
  const int cc = c0 + threadIdx.x;
  if (cc < cols) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += s[j][threadIdx.x];
    part[(size_t)blockIdx.x * cols + cc] = t;
  }
  if (sync == nullptr) return;  // two-launch mode: colsum_final_kernel follows
  __shared__ bool s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(&sync[blockIdx.y], 1u) == gridDim.x - 1u;
    if (s_last) sync[blockIdx.y] = 0u;
    __threadfence();
  }
  __syncthreads();
  if (!s_last || cc >= cols) return;
  float tot = 0.f;
  for (int sgm = 0; sgm < nseg; ++sgm) {
    const float* ps = part + (size_t)sgm * bps * cols + cc;
    float t4[4] = {0.f, 0.f, 0.f, 0.f};
    int j = 0;
    for (; j + 3 < bps; j += 4) {
      t4[0] += __ldcg(ps + (size_t)j * cols);
      t4[1] += __ldcg(ps + (size_t)(j + 1) * cols);
      t4[2] += __ldcg(ps + (size_t)(j + 2) * cols);
      t4[3] += __ldcg(ps + (size_t)(j + 3) * cols);
    }
    for (; j < bps; ++j) t4[0] += __ldcg(ps + (size_t)j * cols);
    const float tseg = (t4[0] + t4[1]) + (t4[2] + t4[3]);
    if (out_seg != nullptr) out_seg[(size_t)sgm * cols + cc] = __float2bfloat16_rn(tseg);
    tot += tseg;
  }
  if (out_total != nullptr) out_total[cc] = tot;
}
// 32 columns per block; 8 row-block slices per column, combined through shared memory (fixed order: deterministic)
__global__ void __launch_bounds__(256)
colsum_final_kernel(const float* __restrict__ part, int nseg, int bps, int cols, float* __restrict__ out_total,
                    __nv_bfloat16* __restrict__ out_seg) {
  sdb::pdl_trigger();
  __shared__ float s[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float tot = 0.f;
  for (int sgm = 0; sgm < nseg; ++sgm) {
    float t = 0.f;
    if (c < cols)
      for (int j = ty; j < bps; j += 8) t += part[((size_t)sgm * bps + j) * cols + c];
    s[ty][tx] = t;
    __syncthreads();
    if (ty == 0) {
#pragma unroll
      for (int j = 1; j < 8; ++j) t += s[j][tx];
      if (c < cols && out_seg != nullptr) out_seg[(size_t)sgm * cols + c] = __float2bfloat16_rn(t);
      tot += t;
    }
    __syncthreads();
  }
  if (ty == 0 && c < cols && out_total != nullptr) out_total[c] = tot;
}

// any (small) column count, no alignment requirement: one block per column
__global__ void __launch_bounds__(256)
colsum_narrow_kernel(const __nv_bfloat16* __restrict__ x, long long ld, long long rows, float* __restrict__ out) {
  sdb::pdl_trigger();
  const int c = blockIdx.x;
  float acc = 0.f;
  for (long long r = threadIdx.x; r < rows; r += 256) acc += __bfloat162float(x[r * ld + c]);
  acc = sdb::warp_sum(acc);
  __shared__ float s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += s[i];
    out[c] = t;
  }
}

__global__ void silu_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n) {
  sdb::pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = __bfloat162float(x[i]);
    y[i] = __float2bfloat16_rn(v / (1.0f + __expf(-v)));
  }
}
// dx = dy * silu'(x) (+ add)
__global__ void silu_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                __nv_bfloat16* __restrict__ dx, long long n) {
  sdb::pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = __bfloat162float(x[i]);
    const float sg = 1.0f / (1.0f + __expf(-v));
    dx[i] = __float2bfloat16_rn(__bfloat162float(dy[i]) * sg * (1.0f + v * (1.0f - sg)));
  }
}

// emb[b, :] = flip ? [cos(t f), sin(t f)] : [sin, cos], f_i = exp(-ln(10000) * i / (half - freq_shift))
__global__ void timestep_embedding_kernel(const int* __restrict__ t, __nv_bfloat16* __restrict__ out, int B, int dim,
                                          int flip_sin_to_cos, float freq_shift) {
  sdb::pdl_trigger();
  const int half = dim / 2;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * half) return;
  const int b = idx / half, i = idx - b * half;
  const float f = expf(-logf(10000.0f) * (float)i / ((float)half - freq_shift));
  const float arg = (float)t[b] * f;
  const float sv = sinf(arg), cv = cosf(arg);
  __nv_bfloat16* o = out + (size_t)b * dim;
  o[i] = __float2bfloat16_rn(flip_sin_to_cos ? cv : sv);
  o[half + i] = __float2bfloat16_rn(flip_sin_to_cos ? sv : cv);
}

}  // namespace

// dst NCHW fp32 [B,C,HW] <- src NHWC bf16 [B,HW,C]: the transpose back at the end of FlaxUNet2DConditionModel.__call__
namespace {
__global__ void nhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int B, int C, int HW) {
  sdb::pdl_trigger();
  const long long total = (long long)B * C * HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {  // i indexes NCHW
    const int p = (int)(i % HW);
    const long long bc = i / HW;
    const int c = (int)(bc % C);
    const int b = (int)(bc / C);
    dst[i] = __bfloat162float(src[((long long)b * HW + p) * C + c]);
  }
}
// TrainState.step, ScaleBy8bitLionState.count, ScaleByScheduleState.count: += 1 (flax TrainState.apply_gradients,
// lion_quant.py:151, optax scale_by_schedule)
__global__ void bump_counters_kernel(int* a, int* b, int* c) {
  sdb::pdl_trigger();
  if (threadIdx.x == 0 && a) *a += 1;
  if (threadIdx.x == 1 && b) *b += 1;
  if (threadIdx.x == 2 && c) *c += 1;
}
}  // namespace

extern "C" int sdb_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int HW, cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "sdb_nhwc_bf16_to_nchw_f32: bad shape");
  nhwc_to_nchw_kernel<<<ew_grid((long long)B * C * HW), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src),
                                                                          dst, B, C, HW);
  SDB_CUDA_LAUNCH_CHECK("sdb_nhwc_bf16_to_nchw_f32");
  return SDB_OK;
}

extern "C" int sdb_bump_counters(int* a, int* b, int* c, cudaStream_t stream) {
  bump_counters_kernel<<<1, 32, 0, stream>>>(a, b, c);
  SDB_CUDA_LAUNCH_CHECK("sdb_bump_counters");
  return SDB_OK;
}

extern "C" int sdb_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
  SDB_CHECK_ARG(n >= 0 && (((uintptr_t)src) & 15) == 0 && (((uintptr_t)dst) & 7) == 0, "sdb_cast_f32_bf16: alignment");
  if (n == 0) return SDB_OK;
  cast_f32_bf16_kernel<<<ew_grid(n / 4 + 1), 256, 0, stream>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  SDB_CUDA_LAUNCH_CHECK("sdb_cast_f32_bf16");
  return SDB_OK;
}

extern "C" int sdb_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int HW, cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "sdb_nchw_f32_to_nhwc_bf16: bad shape");
  nchw_to_nhwc_kernel<<<ew_grid((long long)B * C * HW), 256, 0, stream>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), B,
                                                                          C, HW);
  SDB_CUDA_LAUNCH_CHECK("sdb_nchw_f32_to_nhwc_bf16");
  return SDB_OK;
}

extern "C" int sdb_upsample2x_fwd(const void* x, void* y, int B, int H, int W, int C, cudaStream_t stream) {
  SDB_CHECK_ARG(C % 8 == 0, "sdb_upsample2x_fwd: C=%d must be a multiple of 8", C);
  upsample2x_fwd_kernel<<<ew_grid((long long)B * H * W * 4 * (C / 8)), 256, 0, stream>>>(
      reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(y), B, H, W, C / 8);
  SDB_CUDA_LAUNCH_CHECK("sdb_upsample2x_fwd");
  return SDB_OK;
}

extern "C" int sdb_upsample2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, cudaStream_t stream) {
  SDB_CHECK_ARG(C % 8 == 0, "sdb_upsample2x_bwd: C=%d must be a multiple of 8", C);
  upsample2x_bwd_kernel<<<ew_grid((long long)B * H * W * (C / 8)), 256, 0, stream>>>(
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<uint4*>(dx), B, H, W, C / 8);
  SDB_CUDA_LAUNCH_CHECK("sdb_upsample2x_bwd");
  return SDB_OK;
}

extern "C" int sdb_copy2d_bf16(const void* src, long long lds, void* dst, long long ldd, long long rows, int cols,
                               int accumulate, cudaStream_t stream) {
  SDB_CHECK_ARG(cols % 8 == 0 && lds % 8 == 0 && ldd % 8 == 0, "sdb_copy2d_bf16: cols/strides must be multiples of 8");
  SDB_CHECK_ARG(((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0, "sdb_copy2d_bf16: 16-byte alignment required");
  if (rows <= 0 || cols <= 0) return SDB_OK;
  const int grid = ew_grid(rows * (cols / 8));
  if (accumulate)
    copy2d_kernel<true><<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src), lds,
                                                  reinterpret_cast<__nv_bfloat16*>(dst), ldd, rows, cols / 8);
  else
    copy2d_kernel<false><<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src), lds,
                                                   reinterpret_cast<__nv_bfloat16*>(dst), ldd, rows, cols / 8);
  SDB_CUDA_LAUNCH_CHECK("sdb_copy2d_bf16");
  return SDB_OK;
}

// out[r, c] = bf16(acc[r, c] + bias[c] + temb[r / rows_per_image, c] + residual[r, c]): the epilogue of a split-K forward
// GEMM (sdb_gemm_tc with splits > 1 leaves fp32 partial sums in acc)
__global__ void splitk_finish_kernel(const float* __restrict__ acc, long long lda, int nslabs, long long slab_stride,
                                     const float* __restrict__ bias,
                                     const __nv_bfloat16* __restrict__ temb, long long ldt, long long rows_per_image,
                                     const __nv_bfloat16* __restrict__ residual, long long ldr,
                                     __nv_bfloat16* __restrict__ out, long long ldo, long long rows, int nvec) {
  sdb::pdl_trigger();
  const long long total = rows * nvec;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = (long long)((unsigned int)i / (unsigned int)nvec);  // rows * nvec < 2^31 (checked on the host)
    const int c = (int)(i - r * nvec) * 8;
    const float4 a0 = sdb::ld_f4(acc + r * lda + c), a1 = sdb::ld_f4(acc + r * lda + c + 4);
    float f[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    for (int s = 1; s < nslabs; ++s) {  // fixed order: the result does not depend on the GEMM's schedule
      const float* ps = acc + (long long)s * slab_stride + r * lda + c;
      const float4 s0 = sdb::ld_f4(ps), s1 = sdb::ld_f4(ps + 4);
      f[0] += s0.x; f[1] += s0.y; f[2] += s0.z; f[3] += s0.w; f[4] += s1.x; f[5] += s1.y; f[6] += s1.z; f[7] += s1.w;
    }
    if (bias != nullptr) {
      const float4 b0 = sdb::ld_f4(bias + c), b1 = sdb::ld_f4(bias + c + 4);
      f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w; f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
    }
    if (temb != nullptr) {
      float t[8];
      unpack8(sdb::ld_u4(temb + (long long)((unsigned int)r / (unsigned int)rows_per_image) * ldt + c), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += t[j];
    }
    if (residual != nullptr) {
      float t[8];
      unpack8(sdb::ld_u4(residual + r * ldr + c), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += t[j];
    }
    sdb::st_u4(out + r * ldo + c, pack8(f));
  }
}

extern "C" int sdb_splitk_finish(const float* acc, long long lda, int nslabs, long long slab_stride, const float* bias,
                                 const void* temb, long long ldt, long long rows_per_image, const void* residual,
                                 long long ldr, void* out, long long ldo, long long rows, int N, cudaStream_t stream) {
  SDB_CHECK_ARG(acc != nullptr && out != nullptr && rows > 0 && N > 0, "sdb_splitk_finish: bad arguments");
  SDB_CHECK_ARG(nslabs >= 1 && (nslabs == 1 || (slab_stride >= rows * lda && slab_stride % 4 == 0)),
                "sdb_splitk_finish: bad slab layout (nslabs %d, stride %lld)", nslabs, slab_stride);
  SDB_CHECK_ARG(rows * (N / 8) < (1ll << 31) && rows_per_image < (1ll << 31), "sdb_splitk_finish: too many elements");
  SDB_CHECK_ARG(N % 8 == 0 && lda % 4 == 0 && ldo % 8 == 0 && ldt % 8 == 0 && ldr % 8 == 0 && rows_per_image > 0,
                "sdb_splitk_finish: N %% 8, lda %% 4, ldo/ldt/ldr %% 8 and rows_per_image > 0 required");
  SDB_CHECK_ARG((((uintptr_t)acc | (uintptr_t)out | (uintptr_t)bias | (uintptr_t)temb | (uintptr_t)residual) & 15) == 0,
                "sdb_splitk_finish: 16-byte alignment required");
  splitk_finish_kernel<<<ew_grid(rows * (N / 8)), 256, 0, stream>>>(
      acc, lda, nslabs, slab_stride, bias, reinterpret_cast<const __nv_bfloat16*>(temb), ldt, rows_per_image,
      reinterpret_cast<const __nv_bfloat16*>(residual), ldr, reinterpret_cast<__nv_bfloat16*>(out), ldo, rows, N / 8);
  SDB_CUDA_LAUNCH_CHECK("sdb_splitk_finish");
  return SDB_OK;
}

// out[i] = slab_0[i] + slab_1[i] + ... in that order: the fold of a slab-mode split weight gradient (fp32 in, fp32 out)
__global__ void splitk_finish_f32_kernel(const float* __restrict__ slabs, int nslabs, long long slab_stride,
                                         float* __restrict__ out, long long nvec) {
  sdb::pdl_trigger();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 a = sdb::ld_f4(slabs + 4 * i);
    for (int s = 1; s < nslabs; ++s) {
      const float4 b = sdb::ld_f4(slabs + (long long)s * slab_stride + 4 * i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    sdb::st_f4(out + 4 * i, a);
  }
}

extern "C" int sdb_splitk_finish_f32(const float* slabs, int nslabs, long long slab_stride, float* out, long long n,
                                     cudaStream_t stream) {
  SDB_CHECK_ARG(slabs != nullptr && out != nullptr && n > 0 && n % 4 == 0, "sdb_splitk_finish_f32: bad arguments");
  SDB_CHECK_ARG(nslabs >= 1 && (nslabs == 1 || (slab_stride >= n && slab_stride % 4 == 0)),
                "sdb_splitk_finish_f32: bad slab layout (nslabs %d, stride %lld)", nslabs, slab_stride);
  SDB_CHECK_ARG((((uintptr_t)slabs | (uintptr_t)out) & 15) == 0, "sdb_splitk_finish_f32: 16-byte alignment required");
  splitk_finish_f32_kernel<<<ew_grid(n / 4), 256, 0, stream>>>(slabs, nslabs, slab_stride, out, n / 4);
  SDB_CUDA_LAUNCH_CHECK("sdb_splitk_finish_f32");
  return SDB_OK;
}

extern "C" int sdb_add_bf16(const void* a, const void* b, void* out, long long n, cudaStream_t stream) {
  SDB_CHECK_ARG(n % 8 == 0, "sdb_add_bf16: n must be a multiple of 8");
  if (n == 0) return SDB_OK;
  add_bf16_kernel<<<ew_grid(n / 8), 256, 0, stream>>>(reinterpret_cast<const uint4*>(a),
                                                      reinterpret_cast<const uint4*>(b), reinterpret_cast<uint4*>(out),
                                                      n / 8);
  SDB_CUDA_LAUNCH_CHECK("sdb_add_bf16");
  return SDB_OK;
}

extern "C" size_t sdb_colsum_workspace_bytes(long long rows, int cols) {
  // upper bound for any segmentation of `rows` rows (each segment rounds its block count up)
  const long long blocks = (rows + kColsumMinRows - 1) / kColsumMinRows;
  return (size_t)(blocks + 4096) * cols * sizeof(float);
}

extern "C" int sdb_colsum_bf16(const void* x, long long ld, int nseg, long long rows_per_seg, int cols, float* out_total,
                               void* out_seg_bf16, void* workspace, size_t ws_bytes, cudaStream_t stream) {
  if ((cols % 8 != 0 || ld % 8 != 0) && nseg == 1 && out_seg_bf16 == nullptr && cols <= 64 && out_total != nullptr) {
    colsum_narrow_kernel<<<cols, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), ld, rows_per_seg, out_total);
    SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16(narrow)");
    return SDB_OK;
  }
  SDB_CHECK_ARG(cols % 8 == 0 && ld % 8 == 0 && rows_per_seg > 0 && nseg > 0 && nseg <= 4096,
                "sdb_colsum_bf16: cols/ld must be multiples of 8, 1 <= nseg <= 4096");
  const int col_chunks = (cols + 255) / 256;
  int rpb = 256;  // halve the rows per block until the grid covers ~2 waves (or the block gets too short)
  while (rpb > kColsumMinRows &&
         (long long)nseg * ((rows_per_seg + rpb - 1) / rpb) * col_chunks < 2 * sdb::kNumSMs)
    rpb >>= 1;
  const int bps = (int)((rows_per_seg + rpb - 1) / rpb);
  const size_t need = (size_t)nseg * bps * cols * sizeof(float);
  if (ws_bytes < need) return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_colsum_bf16: workspace %zu < %zu", ws_bytes, need);
  float* part = reinterpret_cast<float*>(workspace);
  if (col_chunks <= kColsumMaxChunks) {  // one launch: the last block of a column chunk folds its partials
    static unsigned int* base = nullptr;
    static std::once_flag once;
    std::call_once(once, [] { cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_colsum_sync); });
    static std::atomic<unsigned int> next{0};
    unsigned int* sync = base + (size_t)kColsumMaxChunks * (next.fetch_add(1u) % (unsigned)kColsumSlots);
    colsum_partial_kernel<<<dim3(nseg * bps, col_chunks), 256, 0, stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), ld, rows_per_seg, bps, rpb, cols, part, nseg, out_total,
        reinterpret_cast<__nv_bfloat16*>(out_seg_bf16), sync);
    SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16");
    return SDB_OK;
  }
  colsum_partial_kernel<<<dim3(nseg * bps, col_chunks), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), ld,
                                                                          rows_per_seg, bps, rpb, cols, part, nseg,
                                                                          nullptr, nullptr, nullptr);
  SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16(partial)");
  colsum_final_kernel<<<(cols + 31) / 32, 256, 0, stream>>>(part, nseg, bps, cols, out_total,
                                                              reinterpret_cast<__nv_bfloat16*>(out_seg_bf16));
  SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16(final)");
  return SDB_OK;
}

extern "C" int sdb_silu_fwd(const void* x, void* y, long long n, cudaStream_t stream) {
  if (n <= 0) return SDB_OK;
  silu_fwd_kernel<<<ew_grid(n), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                  reinterpret_cast<__nv_bfloat16*>(y), n);
  SDB_CUDA_LAUNCH_CHECK("sdb_silu_fwd");
  return SDB_OK;
}

extern "C" int sdb_silu_bwd(const void* x, const void* dy, void* dx, long long n, cudaStream_t stream) {
  if (n <= 0) return SDB_OK;
  silu_bwd_kernel<<<ew_grid(n), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                  reinterpret_cast<const __nv_bfloat16*>(dy),
                                                  reinterpret_cast<__nv_bfloat16*>(dx), n);
  SDB_CUDA_LAUNCH_CHECK("sdb_silu_bwd");
  return SDB_OK;
}

extern "C" int sdb_timestep_embedding(const int* timesteps, void* out, int B, int dim, int flip_sin_to_cos,
                                      float freq_shift, cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && dim > 0 && dim % 2 == 0, "sdb_timestep_embedding: dim must be even");
  timestep_embedding_kernel<<<(B * (dim / 2) + 255) / 256, 256, 0, stream>>>(
      timesteps, reinterpret_cast<__nv_bfloat16*>(out), B, dim, flip_sin_to_cos, freq_shift);
  SDB_CUDA_LAUNCH_CHECK("sdb_timestep_embedding");
  return SDB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free column sum (no counterpart in the reference).
// The partial sums of sdb_colsum_bf16 live in a caller-provided workspace (torch's allocator in this package). Callers
// that must not allocate — code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc — get
// the same result from a scratch array that is part of the library image: every GPU that loads libsdb200.so holds these
// 2 MiB until the process exits.
//   * One array per device, shared by every call: two calls that overlap on the GPU (different streams) RACE on it.
//     Calls on one stream are ordered and safe.
//   * The array keeps the previous call's partial sums; nothing has to start at zero (every element that the second
//     launch reads was written by the first launch of the same call), so nothing is cleared.
//   * In host code `g_colsum_static_scratch` is only a stand-in for the symbol: the device address comes from
//     cudaGetSymbolAddress, once per device, with that device current.
namespace {
constexpr size_t kColsumStaticBytes = (size_t)2 << 20;
static __device__ __align__(256) float g_colsum_static_scratch[kColsumStaticBytes / sizeof(float)];

constexpr int kMaxDevices = 64;
std::mutex g_static_scratch_mutex;
float* g_static_scratch_ptr[kMaxDevices] = {};

// device address of the calling thread's current device's copy of the array (nullptr + error set on failure)
float* colsum_static_scratch_ptr() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    sdb_set_error(SDB_ERR_CUDA, "sdb_colsum_bf16_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      sdb_set_error(SDB_ERR_CUDA, "sdb_colsum_bf16_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return nullptr;
    }
    g_static_scratch_ptr[dev] = p;
  }
  return g_static_scratch_ptr[dev];
}
}  // namespace

extern "C" size_t sdb_colsum_static_max_scratch_bytes(void) { return kColsumStaticBytes; }

extern "C" int sdb_colsum_bf16_static(const void* x, long long ld, long long rows, int cols, float* out_total,
                                      cudaStream_t stream) {
  SDB_CHECK_ARG(x != nullptr && out_total != nullptr, "sdb_colsum_bf16_static: null pointer");
  SDB_CHECK_ARG(cols > 0 && cols % 8 == 0 && ld % 8 == 0 && ld >= cols && rows > 0 && rows < (1ll << 31),
                "sdb_colsum_bf16_static: cols / ld must be multiples of 8, ld >= cols, 0 < rows < 2^31");
  SDB_CHECK_ARG((((uintptr_t)x) & 15) == 0, "sdb_colsum_bf16_static: x needs 16-byte alignment");
  // the grid of sdb_colsum_bf16 for one segment: rows per block halved until ~2 waves of blocks exist
  const int col_chunks = (cols + 255) / 256;
  int rpb = 256;
  while (rpb > kColsumMinRows && ((rows + rpb - 1) / rpb) * col_chunks < 2 * sdb::kNumSMs) rpb >>= 1;
  const int bps = (int)((rows + rpb - 1) / rpb);
  const size_t need = (size_t)bps * cols * sizeof(float);
  if (need > kColsumStaticBytes)
    return sdb_set_error(SDB_ERR_WORKSPACE,
                         "sdb_colsum_bf16_static: [%lld, %d] needs %zu bytes of partial sums, the static scratch holds "
                         "%zu (use sdb_colsum_bf16 with a caller-provided workspace)",
                         rows, cols, need, kColsumStaticBytes);
  float* part = colsum_static_scratch_ptr();
  if (part == nullptr) return SDB_ERR_CUDA;
  // two launches, no arrival counters: no state besides the scratch array
  colsum_partial_kernel<<<dim3(bps, col_chunks), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), ld, rows,
                                                                   bps, rpb, cols, part, 1, nullptr, nullptr, nullptr);
  SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16_static(partial)");
  colsum_final_kernel<<<(cols + 31) / 32, 256, 0, stream>>>(part, 1, bps, cols, out_total, nullptr);
  SDB_CUDA_LAUNCH_CHECK("sdb_colsum_bf16_static(final)");
  return SDB_OK;
}
