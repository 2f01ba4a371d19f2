// GroupNorm(32)+SiLU, LayerNorm and GEGLU, forward and backward — HBM-bound kernels with 128-bit coalesced access.
//
// Replaces flax.linen.GroupNorm / LayerNorm / nn.gelu as used by diffusers-0.21.4 FlaxResnetBlock2D,
// FlaxTransformer2DModel, FlaxBasicTransformerBlock and FlaxGEGLU [upstream, un-vendored; called from
// /root/reference/training_utils.py:678-684]. Semantics restated in SURVEY.md B.4/B.5: statistics in fp32 with the
// fast variance max(0, E[x^2]-E[x]^2), eps 1e-5, affine per channel, SiLU in fp32, tanh-GELU.
//
// Layout: activations are "pixels x channels" bf16 (NHWC), one 16-byte vector = 8 channels. A thread owns a fixed set
// of channel vectors (so gamma/beta and the per-channel accumulators live in registers) and strides over pixel rows.
#include "sdb_common.cuh"
#include "../../include/sdb200.h"
#include <atomic>
#include <cstdlib>
#include <mutex>

namespace {

constexpr int kGroups = 32;
constexpr int kGnUnroll = 4;  // rows in flight per thread

struct GnShape {
  int C, HW, B, cpg, nv, TC, RL, KC, chunks, rows_per_chunk;
};

__host__ bool gn_shape(int B, int HW, int C, GnShape* s) {
  if (C % 8 != 0 || C % kGroups != 0) return false;
  s->C = C; s->HW = HW; s->B = B; s->cpg = C / kGroups; s->nv = C / 8;
  int tc = 1;
  for (int d = 1; d <= s->nv && d <= 256; ++d)
    if (s->nv % d == 0) tc = d;
  s->TC = tc;
  s->KC = s->nv / tc;
  if (s->KC > 2) return false;
  s->RL = 256 / tc;
  if (s->RL > HW) s->RL = HW;
  // enough blocks for ~2 waves, at least RL rows each
  int want = (2 * sdb::kNumSMs + B - 1) / B;
  int max_chunks = (HW + s->RL - 1) / s->RL;
  s->chunks = want < max_chunks ? want : max_chunks;
  if (s->chunks < 1) s->chunks = 1;
  s->rows_per_chunk = (HW + s->chunks - 1) / s->chunks;
  s->chunks = (HW + s->rows_per_chunk - 1) / s->rows_per_chunk;
  return true;
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
// MUFU.EX2 + MUFU.RCP (2 ulp): the IEEE division of `1.0f / (...)` costs ~10 instructions with its fix-up path, and these
// kernels are instruction-bound (profiles/r1n_ncu_full.md); the result is rounded to bf16 anyway
__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
// MUFU.TANH (max relative error 2^-11) instead of the ~25-instruction tanhf
__device__ __forceinline__ float tanh_approx_f(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---------------------------------------------------------------------------------------------------------------
// GroupNorm pass 1: per (sample, chunk, channel) partial sums.
//   BWD = false: (sum x, sum x^2)
//   BWD = true : dz = dy * (SILU ? silu'(z) : 1), z = xhat*gamma+beta; (sum dz, sum dz*xhat)
template <int KC, bool BWD, bool SILU>
__global__ void __launch_bounds__(256)
gn_partial_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                  const float2* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                  float2* __restrict__ part, GnShape s) {
  sdb::pdl_trigger();
  extern __shared__ float smem[];
  float* s_a = smem;                       // [RL][C]
  float* s_b = smem + (size_t)s.RL * s.C;  // [RL][C]
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int tc = threadIdx.x % s.TC, rl = threadIdx.x / s.TC;
  const int r0 = chunk * s.rows_per_chunk;
  const int r1 = min(s.HW, r0 + s.rows_per_chunk);
  float a0[KC][8], a1[KC][8], gm[KC][8], bt[KC][8], mu[KC][8], rs[KC][8];
#pragma unroll
  for (int k = 0; k < KC; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a0[k][i] = 0.f; a1[k][i] = 0.f;
      if (BWD) {
        const int c = (tc + k * s.TC) * 8 + i;
        const float2 st = stats[b * kGroups + c / s.cpg];
        mu[k][i] = st.x; rs[k][i] = st.y; gm[k][i] = gamma[c]; bt[k][i] = beta[c];
      }
    }
  const __nv_bfloat16* xb = x + (size_t)b * s.HW * s.C;
  const __nv_bfloat16* dyb = BWD ? dy + (size_t)b * s.HW * s.C : nullptr;
  if (rl < s.RL) {
    // kGnUnroll rows per trip: all 16-byte loads of a trip are issued before any is consumed (latency hiding; a
    // thread owns only a handful of rows, so one load in flight at a time leaves the kernel latency-bound)
    for (int r = r0 + rl; r < r1; r += kGnUnroll * s.RL) {
      uint4 xr[kGnUnroll][KC], dr[kGnUnroll][KC];
#pragma unroll
      for (int u = 0; u < kGnUnroll; ++u) {
        const int rr = r + u * s.RL;
#pragma unroll
        for (int k = 0; k < KC; ++k) {
          const size_t off = (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
          xr[u][k] = (rr < r1) ? sdb::ld_u4(xb + off) : make_uint4(0u, 0u, 0u, 0u);
          if (BWD) dr[u][k] = (rr < r1) ? sdb::ld_u4(dyb + off) : make_uint4(0u, 0u, 0u, 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < kGnUnroll; ++u) {
#pragma unroll
        for (int k = 0; k < KC; ++k) {
          float xv[8];
          unpack8(xr[u][k], xv);
          if (!BWD) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { a0[k][i] += xv[i]; a1[k][i] = fmaf(xv[i], xv[i], a1[k][i]); }
          } else {
            float dv[8];
            unpack8(dr[u][k], dv);  // rows past the chunk carry dy = 0: they add nothing
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float xh = (xv[i] - mu[k][i]) * rs[k][i];
              float dz = dv[i];
              if (SILU) {
                const float z = fmaf(xh, gm[k][i], bt[k][i]);
                const float sg = sigmoid_f(z);
                dz *= sg * (1.0f + z * (1.0f - sg));
              }
              a0[k][i] += dz; a1[k][i] = fmaf(dz, xh, a1[k][i]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = (tc + k * s.TC) * 8 + i;
        s_a[rl * s.C + c] = a0[k][i];
        s_b[rl * s.C + c] = a1[k][i];
      }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < s.C; c += blockDim.x) {
    float t0 = 0.f, t1 = 0.f;
    for (int j = 0; j < s.RL; ++j) { t0 += s_a[j * s.C + c]; t1 += s_b[j * s.C + c]; }
    part[((size_t)b * s.chunks + chunk) * s.C + c] = make_float2(t0, t1);
  }
}

// GroupNorm pass 2 (one block per (group, sample)): reduce chunks -> per-channel totals -> per-group scalars.
// Warp w owns channels w, w+8, ... of the group; lanes stride over the chunk partials.
//   BWD = false: stats[b][g] = (mean, rstd)
//   BWD = true : chan_tot[b][c] = (sum dz, sum dz*xhat); grp[b][g] = (mean_g(gamma*dz), mean_g(gamma*dz*xhat))
template <bool BWD>
__global__ void __launch_bounds__(256)
gn_finalize_kernel(const float2* __restrict__ part, const float* __restrict__ gamma, float2* __restrict__ out_grp,
                   float2* __restrict__ chan_tot, GnShape s, float eps) {
  sdb::pdl_trigger();
  __shared__ float2 s_w[8];
  const int g = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float g0 = 0.f, g1 = 0.f;
  for (int ci = warp; ci < s.cpg; ci += 8) {
    const int c = g * s.cpg + ci;
    float t0 = 0.f, t1 = 0.f;
    for (int j = lane; j < s.chunks; j += 32) {
      const float2 p = part[((size_t)b * s.chunks + j) * s.C + c];
      t0 += p.x; t1 += p.y;
    }
    t0 = sdb::warp_sum(t0);
    t1 = sdb::warp_sum(t1);
    if (BWD) {
      if (lane == 0) chan_tot[(size_t)b * s.C + c] = make_float2(t0, t1);
      const float gm = gamma[c];
      t0 *= gm; t1 *= gm;
    }
    g0 += t0; g1 += t1;
  }
  if (lane == 0) s_w[warp] = make_float2(g0, g1);
  __syncthreads();
  if (threadIdx.x == 0) {
    float t0 = 0.f, t1 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { t0 += s_w[w].x; t1 += s_w[w].y; }
    const float inv_n = 1.0f / ((float)s.HW * (float)s.cpg);
    if (!BWD) {
      const float mean = t0 * inv_n;
      const float var = fmaxf(0.0f, t1 * inv_n - mean * mean);
      out_grp[b * kGroups + g] = make_float2(mean, rsqrtf(var + eps));
    } else {
      out_grp[b * kGroups + g] = make_float2(t0 * inv_n, t1 * inv_n);
    }
  }
}

// GroupNorm pass 3: apply.
//   BWD = false: y = act(xhat*gamma + beta)
//   BWD = true : dx = rstd * (gamma*dz - A_g - xhat*B_g) (+ dx_add if given: gradient accumulation into a fan-out)
template <int KC, bool BWD, bool SILU>
__global__ void __launch_bounds__(256)
gn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                const float2* __restrict__ stats, const float2* __restrict__ grp, const float* __restrict__ gamma,
                const float* __restrict__ beta, const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ out,
                GnShape s) {
  sdb::pdl_trigger();
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int tc = threadIdx.x % s.TC, rl = threadIdx.x / s.TC;
  if (rl >= s.RL) return;
  const int r0 = chunk * s.rows_per_chunk;
  const int r1 = min(s.HW, r0 + s.rows_per_chunk);
  float gm[KC][8], bt[KC][8], mu[KC][8], rs[KC][8], ga[KC][8], gb[KC][8];
#pragma unroll
  for (int k = 0; k < KC; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = (tc + k * s.TC) * 8 + i;
      const int g = c / s.cpg;
      const float2 st = stats[b * kGroups + g];
      mu[k][i] = st.x; rs[k][i] = st.y; gm[k][i] = gamma[c]; bt[k][i] = beta[c];
      if (BWD) { const float2 gg = grp[b * kGroups + g]; ga[k][i] = gg.x; gb[k][i] = gg.y; }
    }
  const size_t base = (size_t)b * s.HW * s.C;
  for (int r = r0 + rl; r < r1; r += kGnUnroll * s.RL) {
    uint4 xr[kGnUnroll][KC], dr[kGnUnroll][KC], ar[kGnUnroll][KC];
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int rr = r + u * s.RL;
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const size_t off = base + (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
        const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
        xr[u][k] = (rr < r1) ? sdb::ld_u4(x + off) : z4;
        if (BWD) {
          dr[u][k] = (rr < r1) ? sdb::ld_u4(dy + off) : z4;
          ar[u][k] = (rr < r1 && add != nullptr) ? sdb::ld_u4(add + off) : z4;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int rr = r + u * s.RL;
      if (rr >= r1) break;
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const size_t off = base + (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
        float xv[8], o[8];
        unpack8(xr[u][k], xv);
        if (!BWD) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float z = fmaf((xv[i] - mu[k][i]) * rs[k][i], gm[k][i], bt[k][i]);
            o[i] = SILU ? z * sigmoid_f(z) : z;
          }
        } else {
          float dv[8], av[8];
          unpack8(dr[u][k], dv);
          unpack8(ar[u][k], av);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float xh = (xv[i] - mu[k][i]) * rs[k][i];
            float dz = dv[i];
            if (SILU) {
              const float z = fmaf(xh, gm[k][i], bt[k][i]);
              const float sg = sigmoid_f(z);
              dz *= sg * (1.0f + z * (1.0f - sg));
            }
            o[i] = rs[k][i] * (gm[k][i] * dz - ga[k][i] - xh * gb[k][i]) + av[i];
          }
        }
        sdb::st_u4(out + off, pack8(o));
      }
    }
  }
}

// dgamma[c] = sum_b chan_tot[b][c].y ; dbeta[c] = sum_b chan_tot[b][c].x
__global__ void gn_param_grad_kernel(const float2* __restrict__ chan_tot, float* __restrict__ dgamma,
                                     float* __restrict__ dbeta, int B, int C) {
  sdb::pdl_trigger();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float t0 = 0.f, t1 = 0.f;
  for (int b = 0; b < B; ++b) { const float2 p = chan_tot[(size_t)b * C + c]; t0 += p.x; t1 += p.y; }
  dbeta[c] = t0;
  dgamma[c] = t1;
}

// ---------------------------------------------------------------------------------------------------------------
// Fused GroupNorm: ONE launch per direction. All CTAs of a sample are co-resident (grid <= occupancy x SMs, checked on
// the host), so the statistics are exchanged through global memory behind a per-sample arrival counter instead of
// ending the kernel: pass 1 (partial sums) -> counter -> every CTA folds the chunks' group partials in a fixed order
// (deterministic) -> pass 2 (normalise) re-reads the rows it has just read (L1/L2 hits). Replaces the partial /
// finalize / param-grad / apply chain above (3-4 launches whose dependent start-up latency dominated on the L2-sized
// feature maps of this UNet); that chain remains the fallback when the grid cannot be co-resident.
//   sync[b] arrivals of sample b, sync[B+b] CTAs past the wait, sync[2B] arrivals of the whole grid (backward: parameter
//   gradients). The counters live in a static device array, zero at rest: the last CTA through resets them.
__device__ __forceinline__ void gn_arrive_and_wait(unsigned int* counter, unsigned int* passed, unsigned int expected) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    while (*reinterpret_cast<volatile unsigned int*>(counter) < expected) __nanosleep(40);
    __threadfence();
    // the last CTA to get past the wait leaves both counters zeroed for the next launch that uses this slot
    if (atomicAdd(passed, 1u) == expected - 1u) {
      *counter = 0u;
      *passed = 0u;
    }
  }
  __syncthreads();
}

// Arrival counters of the fused kernel: zero at rest (self-resetting), one slot per launch in flight.
constexpr int kGnSyncSlots = 256;
constexpr int kGnSyncInts = 160;   // per slot: [0..B) arrivals, [B..2B) passed, [2B] grid arrivals -> B <= 79
__device__ unsigned int g_gn_sync[kGnSyncSlots * kGnSyncInts];

template <int KC, bool BWD, bool SILU>
__global__ void __launch_bounds__(256, 2)
gn_fused_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float2* stats,
                const float* __restrict__ gamma, const float* __restrict__ beta, const __nv_bfloat16* __restrict__ add,
                __nv_bfloat16* __restrict__ out, float2* part_g, float2* part_c, float2* chan_tot,
                float* __restrict__ dgamma, float* __restrict__ dbeta, unsigned int* sync, GnShape s, float eps) {
  sdb::pdl_trigger();
  extern __shared__ float smem[];
  float* s_a = smem;                       // [RL][C]
  float* s_b = smem + (size_t)s.RL * s.C;  // [RL][C]
  __shared__ float2 s_red[8][kGroups];
  __shared__ float2 s_grp[kGroups];   // fwd: (mean, rstd) ; bwd: (mean_g(gamma dz), mean_g(gamma dz xhat))
  __shared__ int s_last;
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int nthr = blockDim.x;
  const int tc = threadIdx.x % s.TC, rl = threadIdx.x / s.TC;
  const int r0 = chunk * s.rows_per_chunk;
  const int r1 = min(s.HW, r0 + s.rows_per_chunk);
  const size_t base = (size_t)b * s.HW * s.C;

  // ---------------- pass 1: per-channel partial sums of this chunk ----------------
  {
    float a0[KC][8], a1[KC][8], sc[KC][8], sh[KC][8], rs_[KC][8], xb[KC][8];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const int cbase = (tc + k * s.TC) * 8;
      int g = cbase / s.cpg, gend = (g + 1) * s.cpg;  // one division per vector: the group index only steps forward
      float gmv[8], btv[8];
      if (BWD) {
        const float4 g0 = sdb::ld_f4(gamma + cbase), g1 = sdb::ld_f4(gamma + cbase + 4);
        const float4 b0 = sdb::ld_f4(beta + cbase), b1 = sdb::ld_f4(beta + cbase + 4);
        gmv[0] = g0.x; gmv[1] = g0.y; gmv[2] = g0.z; gmv[3] = g0.w; gmv[4] = g1.x; gmv[5] = g1.y; gmv[6] = g1.z; gmv[7] = g1.w;
        btv[0] = b0.x; btv[1] = b0.y; btv[2] = b0.z; btv[3] = b0.w; btv[4] = b1.x; btv[5] = b1.y; btv[6] = b1.z; btv[7] = b1.w;
      }
      float2 st = BWD ? stats[b * kGroups + g] : make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a0[k][i] = 0.f; a1[k][i] = 0.f;
        if (BWD) {
          while (cbase + i >= gend) { ++g; gend += s.cpg; st = stats[b * kGroups + g]; }
          rs_[k][i] = st.y; xb[k][i] = -st.x * st.y;              // xhat = x*rs + xb
          sc[k][i] = st.y * gmv[i]; sh[k][i] = fmaf(-st.x * st.y, gmv[i], btv[i]);  // z = x*sc + sh
        }
      }
    }
    if (rl < s.RL) {
      for (int r = r0 + rl; r < r1; r += kGnUnroll * s.RL) {
        uint4 xr[kGnUnroll][KC], dr[kGnUnroll][KC];
#pragma unroll
        for (int u = 0; u < kGnUnroll; ++u) {
          const int rr = r + u * s.RL;
#pragma unroll
          for (int k = 0; k < KC; ++k) {
            const size_t off = base + (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
            xr[u][k] = (rr < r1) ? sdb::ld_u4(x + off) : make_uint4(0u, 0u, 0u, 0u);
            if (BWD) dr[u][k] = (rr < r1) ? sdb::ld_u4(dy + off) : make_uint4(0u, 0u, 0u, 0u);
          }
        }
#pragma unroll
        for (int u = 0; u < kGnUnroll; ++u) {
#pragma unroll
          for (int k = 0; k < KC; ++k) {
            float xv[8];
            unpack8(xr[u][k], xv);
            if (!BWD) {
#pragma unroll
              for (int i = 0; i < 8; ++i) { a0[k][i] += xv[i]; a1[k][i] = fmaf(xv[i], xv[i], a1[k][i]); }
            } else {
              float dv[8];
              unpack8(dr[u][k], dv);  // rows past the chunk carry dy = 0: they add nothing
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float xh = fmaf(xv[i], rs_[k][i], xb[k][i]);
                float dz = dv[i];
                if (SILU) {
                  const float z = fmaf(xv[i], sc[k][i], sh[k][i]);
                  const float sg = sigmoid_f(z);
                  dz *= sg * (1.0f + z * (1.0f - sg));
                }
                a0[k][i] += dz; a1[k][i] = fmaf(dz, xh, a1[k][i]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int k = 0; k < KC; ++k)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int c = (tc + k * s.TC) * 8 + i;
          s_a[rl * s.C + c] = a0[k][i];
          s_b[rl * s.C + c] = a1[k][i];
        }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < s.C; c += nthr) {  // fold the row lanes into row 0 (a thread only touches column c)
    float t0 = 0.f, t1 = 0.f;
    for (int j = 0; j < s.RL; ++j) { t0 += s_a[j * s.C + c]; t1 += s_b[j * s.C + c]; }
    if (BWD) {
      part_c[((size_t)b * s.chunks + chunk) * s.C + c] = make_float2(t0, t1);
      const float gm = gamma[c];
      t0 *= gm; t1 *= gm;
    }
    s_a[c] = t0; s_b[c] = t1;
  }
  __syncthreads();
  if (threadIdx.x < kGroups) {
    float t0 = 0.f, t1 = 0.f;
    for (int ci = 0; ci < s.cpg; ++ci) { t0 += s_a[threadIdx.x * s.cpg + ci]; t1 += s_b[threadIdx.x * s.cpg + ci]; }
    part_g[((size_t)b * s.chunks + chunk) * kGroups + threadIdx.x] = make_float2(t0, t1);
  }

  // ---------------- exchange: wait for the other chunks of this sample ----------------
  gn_arrive_and_wait(&sync[b], &sync[s.B + b], (unsigned int)s.chunks);
  {
    // full warps only (the block has 129..256 threads): warp `slice` folds chunks slice, slice + nslices, ...
    const int g = threadIdx.x & 31, slice = threadIdx.x >> 5, nslices = nthr >> 5;
    float t0 = 0.f, t1 = 0.f;
    if (slice < nslices) {
      for (int j = slice; j < s.chunks; j += nslices) {
        const float2 q = __ldcg(&part_g[((size_t)b * s.chunks + j) * kGroups + g]);
        t0 += q.x; t1 += q.y;
      }
      s_red[slice][g] = make_float2(t0, t1);
    }
    __syncthreads();
    if (threadIdx.x < kGroups) {
      t0 = 0.f; t1 = 0.f;
      for (int j = 0; j < nslices; ++j) { t0 += s_red[j][g].x; t1 += s_red[j][g].y; }
      const float inv_n = 1.0f / ((float)s.HW * (float)s.cpg);
      if (!BWD) {
        const float mean = t0 * inv_n;
        const float var = fmaxf(0.0f, t1 * inv_n - mean * mean);
        const float2 st = make_float2(mean, rsqrtf(var + eps));
        s_grp[g] = st;
        if (chunk == 0) stats[b * kGroups + g] = st;
      } else {
        s_grp[g] = make_float2(t0 * inv_n, t1 * inv_n);
      }
    }
    if (BWD) {
      // channel slice of this CTA: totals over the sample's chunks; the last CTA of the grid folds the samples
      const int cps = (s.C + s.chunks - 1) / s.chunks;
      for (int t = threadIdx.x; t < cps; t += nthr) {
        const int c = chunk * cps + t;
        if (c < s.C) {
          float u0 = 0.f, u1 = 0.f;
          for (int j = 0; j < s.chunks; ++j) {
            const float2 q = __ldcg(&part_c[((size_t)b * s.chunks + j) * s.C + c]);
            u0 += q.x; u1 += q.y;
          }
          chan_tot[(size_t)b * s.C + c] = make_float2(u0, u1);
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int old = atomicAdd(&sync[2 * s.B], 1u);
        s_last = (old == (unsigned int)(s.B * s.chunks) - 1u) ? 1 : 0;
        if (s_last) sync[2 * s.B] = 0u;  // self-reset
        __threadfence();
      }
    }
    __syncthreads();
    if (BWD && s_last) {
      for (int c = threadIdx.x; c < s.C; c += nthr) {
        float u0 = 0.f, u1 = 0.f;
        for (int bb = 0; bb < s.B; ++bb) {
          const float2 q = __ldcg(&chan_tot[(size_t)bb * s.C + c]);
          u0 += q.x; u1 += q.y;
        }
        dbeta[c] = u0;
        dgamma[c] = u1;
      }
    }
  }

  // ---------------- pass 2: normalise / input gradient ----------------
  if (rl >= s.RL) return;
  float sc[KC][8], sh[KC][8], rs_[KC][8], xb[KC][8], rA[KC][8], rB[KC][8];
#pragma unroll
  for (int k = 0; k < KC; ++k) {
    const int cbase = (tc + k * s.TC) * 8;
    int g = cbase / s.cpg, gend = (g + 1) * s.cpg;
    const float4 g0 = sdb::ld_f4(gamma + cbase), g1 = sdb::ld_f4(gamma + cbase + 4);
    const float4 b0 = sdb::ld_f4(beta + cbase), b1 = sdb::ld_f4(beta + cbase + 4);
    const float gmv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float btv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    float2 st = BWD ? stats[b * kGroups + g] : s_grp[g];
    float2 gg = s_grp[g];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      while (cbase + i >= gend) {
        ++g; gend += s.cpg;
        st = BWD ? stats[b * kGroups + g] : s_grp[g];
        gg = s_grp[g];
      }
      sc[k][i] = st.y * gmv[i]; sh[k][i] = fmaf(-st.x * st.y, gmv[i], btv[i]);
      if (BWD) {
        rs_[k][i] = st.y; xb[k][i] = -st.x * st.y;
        rA[k][i] = st.y * gg.x; rB[k][i] = st.y * gg.y;
      }
    }
  }
  for (int r = r0 + rl; r < r1; r += kGnUnroll * s.RL) {
    uint4 xr[kGnUnroll][KC], dr[kGnUnroll][KC], ar[kGnUnroll][KC];
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int rr = r + u * s.RL;
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const size_t off = base + (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
        const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
        xr[u][k] = (rr < r1) ? sdb::ld_u4(x + off) : z4;
        if (BWD) {
          dr[u][k] = (rr < r1) ? sdb::ld_u4(dy + off) : z4;
          ar[u][k] = (rr < r1 && add != nullptr) ? sdb::ld_u4(add + off) : z4;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int rr = r + u * s.RL;
      if (rr >= r1) break;
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const size_t off = base + (size_t)rr * s.C + (size_t)(tc + k * s.TC) * 8;
        float xv[8], o[8];
        unpack8(xr[u][k], xv);
        if (!BWD) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float z = fmaf(xv[i], sc[k][i], sh[k][i]);
            o[i] = SILU ? z * sigmoid_f(z) : z;
          }
        } else {
          float dv[8], av[8];
          unpack8(dr[u][k], dv);
          unpack8(ar[u][k], av);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float xh = fmaf(xv[i], rs_[k][i], xb[k][i]);
            float dz = dv[i];
            if (SILU) {
              const float z = fmaf(xv[i], sc[k][i], sh[k][i]);
              const float sg = sigmoid_f(z);
              dz *= sg * (1.0f + z * (1.0f - sg));
            }
            // dx = rstd * (gamma*dz - A_g - xhat*B_g) (+ add)
            o[i] = fmaf(sc[k][i], dz, av[i] - fmaf(xh, rB[k][i], rA[k][i]));
          }
        }
        sdb::st_u4(out + off, pack8(o));
      }
    }
  }
}

constexpr int kGnFusedMaxOcc = 4;

// workspace of the fused path: sync [B+1] u32 (padded to 256 B) | part_g [B][chunks][32] | part_c [B][chunks][C] | chan_tot [B][C]
size_t gn_fused_ws_bytes(const GnShape& s) {
  return 256 + (((size_t)s.B + 1) * 4 + 255) / 256 * 256 +
         sizeof(float2) * ((size_t)s.B * s.chunks * kGroups + (size_t)s.B * s.chunks * s.C + (size_t)s.B * s.C);
}

// Co-resident grid for the fused kernel: chunks per sample = min(rows / RL, capacity / B). Returns false when the
// samples alone exceed the capacity (fall back to the multi-launch path).
unsigned int* gn_sync_slot() {
  static unsigned int* base = nullptr;
  static std::once_flag once;
  std::call_once(once, [] { cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_gn_sync); });
  static std::atomic<unsigned int> next{0};
  return base + (size_t)kGnSyncInts * (next.fetch_add(1u) % (unsigned)kGnSyncSlots);
}

// SDB_GN_LEGACY=1 keeps the multi-launch path (A/B timing and a second implementation for the parity tests)
bool gn_force_legacy() {
  static const bool v = getenv("SDB_GN_LEGACY") != nullptr && atoi(getenv("SDB_GN_LEGACY")) != 0;
  return v;
}

template <typename K>
bool gn_fused_plan(K kernel, GnShape* s) {
  const int threads = s->TC * s->RL;
  const size_t sm = (size_t)s->RL * s->C * 2 * sizeof(float);
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, sm) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return false;
  }
  if (occ > kGnFusedMaxOcc) occ = kGnFusedMaxOcc;
  const int capacity = occ * sdb::kNumSMs;
  if (s->B > capacity) return false;
  int chunks = capacity / s->B;
  const int max_chunks = (s->HW + s->RL - 1) / s->RL;
  if (chunks > max_chunks) chunks = max_chunks;
  s->rows_per_chunk = (s->HW + chunks - 1) / chunks;
  s->chunks = (s->HW + s->rows_per_chunk - 1) / s->rows_per_chunk;
  return true;
}

size_t gn_ws_bytes(const GnShape& s) {
  // part [B][chunks][C] float2 | chan_tot [B][C] float2 | grp [B][32] float2
  return sizeof(float2) * ((size_t)s.B * s.chunks * s.C + (size_t)s.B * s.C + (size_t)s.B * kGroups);
}

// ---------------------------------------------------------------------------------------------------------------
// LayerNorm: one warp per row; lane l owns vectors l, l+32, ... (C <= 2048).
__device__ __forceinline__ void opaque_u4(uint4& v) { asm volatile("" : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)); }
constexpr int kLnMaxVec = 8;  // C <= 8*32*8 = 2048
constexpr int kLnRowsPerBlock = 8;

template <int NV>  // vectors per lane
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float2* __restrict__ stats, int rows, int C, float eps) {
  sdb::pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = C / 8;
  // R rows of a warp are in flight together (all their 16-byte loads are issued before the first reduction): one row
  // at a time left the kernel latency-bound (a row is only 640 B .. 2.5 KB)
  constexpr int R = (NV <= 2) ? 4 : (NV <= 4 ? 2 : 1);
  const int wstride = gridDim.x * kLnRowsPerBlock;
  for (int row0 = blockIdx.x * kLnRowsPerBlock + warp; row0 < rows; row0 += R * wstride) {
    uint4 xr[R][NV];
    float s0[R], s1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r * wstride;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = lane + 32 * k;
        xr[r][k] = (row < rows && v < nv) ? sdb::ld_u4(x + (size_t)row * C + v * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      s0[r] = 0.f; s1[r] = 0.f;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        float xv[8];
        unpack8(xr[r][k], xv);  // vectors past the row are zero: they add nothing
#pragma unroll
        for (int i = 0; i < 8; ++i) { s0[r] += xv[i]; s1[r] = fmaf(xv[i], xv[i], s1[r]); }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { s0[r] = sdb::warp_sum(s0[r]); s1[r] = sdb::warp_sum(s1[r]); }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r * wstride;
      if (row >= rows) break;
      const float mean = s0[r] / (float)C;
      const float var = fmaxf(0.0f, s1[r] / (float)C - mean * mean);
      const float rstd = rsqrtf(var + eps);
      if (lane == 0 && stats != nullptr) stats[row] = make_float2(mean, rstd);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = lane + 32 * k;
        if (v < nv) {
          float xv[8], o[8];
          opaque_u4(xr[r][k]);
          unpack8(xr[r][k], xv);
          const float4 g0 = sdb::ld_f4(gamma + v * 8), g1 = sdb::ld_f4(gamma + v * 8 + 4);
          const float4 b0 = sdb::ld_f4(beta + v * 8), b1 = sdb::ld_f4(beta + v * 8 + 4);
          const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
          const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = fmaf((xv[i] - mean) * rstd, g[i], bb[i]);
          sdb::st_u4(y + (size_t)row * C + v * 8, pack8(o));
        }
      }
    }
  }
}

// dx = rstd*(g*dy - mean(g*dy) - xhat*mean(g*dy*xhat)) (+ add); per-block partial dgamma/dbeta -> part[block][2][C]
template <int NV>
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
              const float2* __restrict__ stats, const float* __restrict__ gamma, const __nv_bfloat16* __restrict__ add,
              __nv_bfloat16* __restrict__ dx, float* __restrict__ part, int rows, int C) {
  sdb::pdl_trigger();
  extern __shared__ float smem[];  // [8 warps][2][C]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = C / 8;
  float dg[NV][8], db[NV][8], g[NV][8];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int v = lane + 32 * k;
#pragma unroll
    for (int i = 0; i < 8; ++i) { dg[k][i] = 0.f; db[k][i] = 0.f; g[k][i] = (v < nv) ? gamma[v * 8 + i] : 0.f; }
  }
  // R rows of a warp in flight together; x / dy stay packed in registers and are unpacked again for the second pass
  // (the second unpack is hidden from CSE: keeping the unpacked floats of every row alive would not fit the registers)
  constexpr int R = (NV <= 2) ? 2 : 1;
  const int wstride = gridDim.x * kLnRowsPerBlock;
  for (int row0 = blockIdx.x * kLnRowsPerBlock + warp; row0 < rows; row0 += R * wstride) {
    uint4 xr[R][NV], dr[R][NV];
    float2 st[R];
    float s0[R], s1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r * wstride;
      st[r] = (row < rows) ? stats[row] : make_float2(0.f, 0.f);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = lane + 32 * k;
        const bool in = row < rows && v < nv;
        xr[r][k] = in ? sdb::ld_u4(x + (size_t)row * C + v * 8) : make_uint4(0u, 0u, 0u, 0u);
        dr[r][k] = in ? sdb::ld_u4(dy + (size_t)row * C + v * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      s0[r] = 0.f; s1[r] = 0.f;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        float xv[8], dv[8];
        unpack8(xr[r][k], xv);
        unpack8(dr[r][k], dv);  // dy = 0 past the row / the rows: nothing is added
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (xv[i] - st[r].x) * st[r].y;
          const float gd = g[k][i] * dv[i];
          s0[r] += gd;
          s1[r] = fmaf(gd, xh, s1[r]);
          dg[k][i] = fmaf(dv[i], xh, dg[k][i]);
          db[k][i] += dv[i];
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { s0[r] = sdb::warp_sum(s0[r]) / (float)C; s1[r] = sdb::warp_sum(s1[r]) / (float)C; }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r * wstride;
      if (row >= rows) break;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int v = lane + 32 * k;
        if (v < nv) {
          float xv[8], dv[8], o[8], av[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          opaque_u4(xr[r][k]);
          opaque_u4(dr[r][k]);
          unpack8(xr[r][k], xv);
          unpack8(dr[r][k], dv);
          if (add != nullptr) unpack8(sdb::ld_u4(add + (size_t)row * C + v * 8), av);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float xh = (xv[i] - st[r].x) * st[r].y;
            o[i] = st[r].y * (g[k][i] * dv[i] - s0[r] - xh * s1[r]) + av[i];
          }
          sdb::st_u4(dx + (size_t)row * C + v * 8, pack8(o));
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int v = lane + 32 * k;
    if (v < nv) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        smem[(warp * 2 + 0) * C + v * 8 + i] = dg[k][i];
        smem[(warp * 2 + 1) * C + v * 8 + i] = db[k][i];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    const int which = c / C, cc = c - which * C;
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += smem[(w * 2 + which) * C + cc];
    part[(size_t)blockIdx.x * 2 * C + c] = t;
  }
}

// out[which][c] = sum_blocks part[block][which][c]   (which 0 -> dgamma, 1 -> dbeta)
__global__ void __launch_bounds__(256)
colsum_finalize_kernel(const float* __restrict__ part, int nblocks, int width, float* __restrict__ out0,
                       float* __restrict__ out1, int C) {
  sdb::pdl_trigger();
  // 32 columns per block, 8 row slices per column
  __shared__ float s[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float t = 0.f;
  if (c < width)
    for (int j = ty; j < nblocks; j += 8) t += part[(size_t)j * width + c];
  s[ty][tx] = t;
  __syncthreads();
  if (ty == 0 && c < width) {
#pragma unroll
    for (int j = 1; j < 8; ++j) t += s[j][tx];
    if (c < C) out0[c] = t;
    else out1[c - C] = t;
  }
}

int ln_grid(int rows) {
  int blocks = (rows + kLnRowsPerBlock - 1) / kLnRowsPerBlock;
  const int cap = sdb::kNumSMs * 4;
  return blocks < cap ? blocks : cap;
}

// ---------------------------------------------------------------------------------------------------------------
// GEGLU: u [rows, 2*H] = [lin | gate]; out [rows, H] = lin * gelu_tanh(gate)
__device__ __forceinline__ float gelu_tanh_f(float z) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float t = tanh_approx_f(k0 * (z + k1 * z * z * z));
  return 0.5f * z * (1.0f + t);
}
__device__ __forceinline__ void gelu_tanh_both(float z, float* val, float* grad) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float inner = k0 * (z + k1 * z * z * z);
  const float t = tanh_approx_f(inner);
  *val = 0.5f * z * (1.0f + t);
  *grad = 0.5f * (1.0f + t) + 0.5f * z * (1.0f - t * t) * k0 * (1.0f + 3.0f * k1 * z * z);
}

__global__ void __launch_bounds__(256)
geglu_fwd_kernel(const __nv_bfloat16* __restrict__ u, __nv_bfloat16* __restrict__ out, long long rows, int H) {
  sdb::pdl_trigger();
  const int hv = H / 8;
  const long long total = rows * hv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    // 32-bit division when the index fits (always, for the UNet's shapes): a 64-bit one costs ~40 instructions per vector
    const long long r = (total < (1ll << 32)) ? (long long)((unsigned int)idx / (unsigned int)hv) : idx / hv;
    const int v = (int)(idx - r * hv);
    float lin[8], gate[8], o[8];
    unpack8(sdb::ld_u4(u + r * 2 * H + v * 8), lin);
    unpack8(sdb::ld_u4(u + r * 2 * H + H + v * 8), gate);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = lin[i] * gelu_tanh_f(gate[i]);
    sdb::st_u4(out + r * H + v * 8, pack8(o));
  }
}

__global__ void __launch_bounds__(256)
geglu_bwd_kernel(const __nv_bfloat16* __restrict__ u, const __nv_bfloat16* __restrict__ dout,
                 __nv_bfloat16* __restrict__ du, long long rows, int H) {
  sdb::pdl_trigger();
  const int hv = H / 8;
  const long long total = rows * hv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    // 32-bit division when the index fits (always, for the UNet's shapes): a 64-bit one costs ~40 instructions per vector
    const long long r = (total < (1ll << 32)) ? (long long)((unsigned int)idx / (unsigned int)hv) : idx / hv;
    const int v = (int)(idx - r * hv);
    float lin[8], gate[8], d[8], dl[8], dg[8];
    unpack8(sdb::ld_u4(u + r * 2 * H + v * 8), lin);
    unpack8(sdb::ld_u4(u + r * 2 * H + H + v * 8), gate);
    unpack8(sdb::ld_u4(dout + r * H + v * 8), d);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float gv, gg;
      gelu_tanh_both(gate[i], &gv, &gg);
      dl[i] = d[i] * gv;
      dg[i] = d[i] * lin[i] * gg;
    }
    sdb::st_u4(du + r * 2 * H + v * 8, pack8(dl));
    sdb::st_u4(du + r * 2 * H + H + v * 8, pack8(dg));
  }
}

int ew_grid(long long work, int per_sm = 8) {
  long long blocks = (work + 255) / 256;
  long long cap = (long long)sdb::kNumSMs * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

// ===================================================== C ABI =====================================================
extern "C" size_t sdb_groupnorm_workspace_bytes(int B, int HW, int C) {
  GnShape s;
  if (!gn_shape(B, HW, C, &s)) return 0;
  const size_t legacy = gn_ws_bytes(s);
  GnShape f = s;  // upper bound of the fused path's chunk count
  int chunks = (kGnFusedMaxOcc * sdb::kNumSMs) / (B > 0 ? B : 1);
  const int max_chunks = (HW + s.RL - 1) / s.RL;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  f.chunks = chunks;
  const size_t fused = gn_fused_ws_bytes(f);
  return legacy > fused ? legacy : fused;
}

extern "C" int sdb_groupnorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, int B,
                                 int HW, int C, float eps, int silu, void* workspace, size_t ws_bytes,
                                 cudaStream_t stream) {
  GnShape s;
  SDB_CHECK_ARG(gn_shape(B, HW, C, &s), "sdb_groupnorm_fwd: unsupported C=%d (need C %% 32 == 0, C %% 8 == 0)", C);
  if (ws_bytes < gn_ws_bytes(s))
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_groupnorm_fwd: workspace %zu < %zu", ws_bytes, gn_ws_bytes(s));
  float2* part = reinterpret_cast<float2*>(workspace);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y);
  float2* st = reinterpret_cast<float2*>(stats);
  const int threads = s.TC * s.RL;
  const size_t sm = (size_t)s.RL * s.C * 2 * sizeof(float);
  if (!gn_force_legacy()) {
    GnShape f = s;
    bool ok;
#define GN_PLAN(KC_, SILU_) ok = gn_fused_plan(gn_fused_kernel<KC_, false, SILU_>, &f)
    if (s.KC == 1) { if (silu) GN_PLAN(1, true); else GN_PLAN(1, false); }
    else { if (silu) GN_PLAN(2, true); else GN_PLAN(2, false); }
#undef GN_PLAN
    if (ok && gn_fused_ws_bytes(f) <= ws_bytes && 2 * B + 1 <= kGnSyncInts) {
      unsigned int* sync = gn_sync_slot();
      const size_t sync_bytes = (((size_t)B + 1) * 4 + 255) / 256 * 256;
      float2* part_g = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(workspace) + sync_bytes);
      const dim3 fgrid(f.chunks, B);
#define GN_FUSED(KC_, SILU_)                                                                                        \
  gn_fused_kernel<KC_, false, SILU_><<<fgrid, threads, sm, stream>>>(xp, nullptr, st, gamma, beta, nullptr, yp, part_g, \
                                                                    nullptr, nullptr, nullptr, nullptr, sync, f, eps)
      if (s.KC == 1) { if (silu) GN_FUSED(1, true); else GN_FUSED(1, false); }
      else { if (silu) GN_FUSED(2, true); else GN_FUSED(2, false); }
#undef GN_FUSED
      SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_fwd(fused)");
      return SDB_OK;
    }
  }
  const dim3 grid(s.chunks, B);
  if (s.KC == 1) gn_partial_kernel<1, false, false><<<grid, threads, sm, stream>>>(xp, nullptr, nullptr, nullptr, nullptr, part, s);
  else gn_partial_kernel<2, false, false><<<grid, threads, sm, stream>>>(xp, nullptr, nullptr, nullptr, nullptr, part, s);
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_fwd(partial)");
  gn_finalize_kernel<false><<<dim3(kGroups, B), 256, 0, stream>>>(part, nullptr, st, nullptr, s, eps);
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_fwd(finalize)");
#define GN_APPLY(KC_, SILU_) \
  gn_apply_kernel<KC_, false, SILU_><<<grid, threads, 0, stream>>>(xp, nullptr, st, nullptr, gamma, beta, nullptr, yp, s)
  if (s.KC == 1) { if (silu) GN_APPLY(1, true); else GN_APPLY(1, false); }
  else { if (silu) GN_APPLY(2, true); else GN_APPLY(2, false); }
#undef GN_APPLY
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_fwd(apply)");
  return SDB_OK;
}

extern "C" int sdb_groupnorm_bwd(const void* x, const void* dy, const float* stats, const float* gamma,
                                 const float* beta, const void* dx_add, void* dx, float* dgamma, float* dbeta, int B,
                                 int HW, int C, int silu, void* workspace, size_t ws_bytes, cudaStream_t stream) {
  GnShape s;
  SDB_CHECK_ARG(gn_shape(B, HW, C, &s), "sdb_groupnorm_bwd: unsupported C=%d", C);
  if (ws_bytes < gn_ws_bytes(s))
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_groupnorm_bwd: workspace %zu < %zu", ws_bytes, gn_ws_bytes(s));
  float2* part = reinterpret_cast<float2*>(workspace);
  float2* chan_tot = part + (size_t)s.B * s.chunks * s.C;
  float2* grp = chan_tot + (size_t)s.B * s.C;
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* dyp = reinterpret_cast<const __nv_bfloat16*>(dy);
  const __nv_bfloat16* addp = reinterpret_cast<const __nv_bfloat16*>(dx_add);
  __nv_bfloat16* dxp = reinterpret_cast<__nv_bfloat16*>(dx);
  const float2* st = reinterpret_cast<const float2*>(stats);
  const int threads = s.TC * s.RL;
  const size_t sm = (size_t)s.RL * s.C * 2 * sizeof(float);
  if (!gn_force_legacy()) {
    GnShape f = s;
    bool ok;
#define GN_PLAN(KC_, SILU_) ok = gn_fused_plan(gn_fused_kernel<KC_, true, SILU_>, &f)
    if (s.KC == 1) { if (silu) GN_PLAN(1, true); else GN_PLAN(1, false); }
    else { if (silu) GN_PLAN(2, true); else GN_PLAN(2, false); }
#undef GN_PLAN
    if (ok && gn_fused_ws_bytes(f) <= ws_bytes && 2 * B + 1 <= kGnSyncInts) {
      unsigned int* sync = gn_sync_slot();
      const size_t sync_bytes = (((size_t)B + 1) * 4 + 255) / 256 * 256;
      float2* part_g = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(workspace) + sync_bytes);
      float2* part_c = part_g + (size_t)f.B * f.chunks * kGroups;
      float2* ctot = part_c + (size_t)f.B * f.chunks * f.C;
      const dim3 fgrid(f.chunks, B);
#define GN_FUSED(KC_, SILU_)                                                                                          \
  gn_fused_kernel<KC_, true, SILU_><<<fgrid, threads, sm, stream>>>(xp, dyp, const_cast<float2*>(st), gamma, beta, addp, \
                                                                   dxp, part_g, part_c, ctot, dgamma, dbeta, sync, f, 0.f)
      if (s.KC == 1) { if (silu) GN_FUSED(1, true); else GN_FUSED(1, false); }
      else { if (silu) GN_FUSED(2, true); else GN_FUSED(2, false); }
#undef GN_FUSED
      SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_bwd(fused)");
      return SDB_OK;
    }
  }
  const dim3 grid(s.chunks, B);
#define GN_PART(KC_, SILU_) \
  gn_partial_kernel<KC_, true, SILU_><<<grid, threads, sm, stream>>>(xp, dyp, st, gamma, beta, part, s)
  if (s.KC == 1) { if (silu) GN_PART(1, true); else GN_PART(1, false); }
  else { if (silu) GN_PART(2, true); else GN_PART(2, false); }
#undef GN_PART
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_bwd(partial)");
  gn_finalize_kernel<true><<<dim3(kGroups, B), 256, 0, stream>>>(part, gamma, grp, chan_tot, s, 0.f);
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_bwd(finalize)");
  gn_param_grad_kernel<<<(C + 255) / 256, 256, 0, stream>>>(chan_tot, dgamma, dbeta, B, C);
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_bwd(param)");
#define GN_APPLY(KC_, SILU_) \
  gn_apply_kernel<KC_, true, SILU_><<<grid, threads, 0, stream>>>(xp, dyp, st, grp, gamma, beta, addp, dxp, s)
  if (s.KC == 1) { if (silu) GN_APPLY(1, true); else GN_APPLY(1, false); }
  else { if (silu) GN_APPLY(2, true); else GN_APPLY(2, false); }
#undef GN_APPLY
  SDB_CUDA_LAUNCH_CHECK("sdb_groupnorm_bwd(apply)");
  return SDB_OK;
}

extern "C" int sdb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats,
                                 long long rows, int C, float eps, cudaStream_t stream) {
  SDB_CHECK_ARG(C % 8 == 0 && C <= kLnMaxVec * 256, "sdb_layernorm_fwd: C=%d must be a multiple of 8 and <= 2048", C);
  SDB_CHECK_ARG(rows > 0 && rows < (1ll << 31), "sdb_layernorm_fwd: bad rows");
  const int NV = (C / 8 + 31) / 32;
  const int grid = ln_grid((int)rows);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y);
  float2* st = reinterpret_cast<float2*>(stats);
#define LN_F(NV_) ln_fwd_kernel<NV_><<<grid, 256, 0, stream>>>(xp, gamma, beta, yp, st, (int)rows, C, eps)
  switch (NV) {
    case 1: LN_F(1); break;
    case 2: LN_F(2); break;
    case 3: LN_F(3); break;
    case 4: LN_F(4); break;
    case 5: LN_F(5); break;
    case 6: LN_F(6); break;
    case 7: LN_F(7); break;
    default: LN_F(8); break;
  }
#undef LN_F
  SDB_CUDA_LAUNCH_CHECK("sdb_layernorm_fwd");
  return SDB_OK;
}

extern "C" size_t sdb_layernorm_bwd_workspace_bytes(long long rows, int C) {
  return (size_t)ln_grid((int)rows) * 2 * C * sizeof(float);
}

extern "C" int sdb_layernorm_bwd(const void* x, const void* dy, const float* stats, const float* gamma,
                                 const void* dx_add, void* dx, float* dgamma, float* dbeta, long long rows, int C,
                                 void* workspace, size_t ws_bytes, cudaStream_t stream) {
  SDB_CHECK_ARG(C % 8 == 0 && C <= kLnMaxVec * 256, "sdb_layernorm_bwd: C=%d must be a multiple of 8 and <= 2048", C);
  SDB_CHECK_ARG(rows > 0 && rows < (1ll << 31), "sdb_layernorm_bwd: bad rows");
  if (ws_bytes < sdb_layernorm_bwd_workspace_bytes(rows, C))
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_layernorm_bwd: workspace too small");
  const int NV = (C / 8 + 31) / 32;
  const int grid = ln_grid((int)rows);
  const size_t sm = (size_t)8 * 2 * C * sizeof(float);
  float* part = reinterpret_cast<float*>(workspace);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* dyp = reinterpret_cast<const __nv_bfloat16*>(dy);
  const __nv_bfloat16* addp = reinterpret_cast<const __nv_bfloat16*>(dx_add);
  __nv_bfloat16* dxp = reinterpret_cast<__nv_bfloat16*>(dx);
  const float2* st = reinterpret_cast<const float2*>(stats);
#define LN_B(NV_)                                                                                                  \
  do {                                                                                                             \
    if (sm > 48 * 1024)                                                                                            \
      cudaFuncSetAttribute(ln_bwd_kernel<NV_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);              \
    ln_bwd_kernel<NV_><<<grid, 256, sm, stream>>>(xp, dyp, st, gamma, addp, dxp, part, (int)rows, C);              \
  } while (0)
  switch (NV) {
    case 1: LN_B(1); break;
    case 2: LN_B(2); break;
    case 3: LN_B(3); break;
    case 4: LN_B(4); break;
    case 5: LN_B(5); break;
    case 6: LN_B(6); break;
    case 7: LN_B(7); break;
    default: LN_B(8); break;
  }
#undef LN_B
  SDB_CUDA_LAUNCH_CHECK("sdb_layernorm_bwd");
  colsum_finalize_kernel<<<(2 * C + 31) / 32, 256, 0, stream>>>(part, grid, 2 * C, dgamma, dbeta, C);
  SDB_CUDA_LAUNCH_CHECK("sdb_layernorm_bwd(finalize)");
  return SDB_OK;
}

extern "C" int sdb_geglu_fwd(const void* u, void* out, long long rows, int H, cudaStream_t stream) {
  SDB_CHECK_ARG(H % 8 == 0 && rows > 0, "sdb_geglu_fwd: H=%d must be a multiple of 8", H);
  geglu_fwd_kernel<<<ew_grid(rows * (H / 8)), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(u),
                                                                reinterpret_cast<__nv_bfloat16*>(out), rows, H);
  SDB_CUDA_LAUNCH_CHECK("sdb_geglu_fwd");
  return SDB_OK;
}

extern "C" int sdb_geglu_bwd(const void* u, const void* dout, void* du, long long rows, int H, cudaStream_t stream) {
  SDB_CHECK_ARG(H % 8 == 0 && rows > 0, "sdb_geglu_bwd: H=%d must be a multiple of 8", H);
  geglu_bwd_kernel<<<ew_grid(rows * (H / 8)), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(u),
                                                                reinterpret_cast<const __nv_bfloat16*>(dout),
                                                                reinterpret_cast<__nv_bfloat16*>(du), rows, H);
  SDB_CUDA_LAUNCH_CHECK("sdb_geglu_bwd");
  return SDB_OK;
}
