// Text-encoder specific kernels: embeddings, short-sequence (<= 128 tokens) causal attention, quick-GELU / GELU, and the
// row gather / scatter of the reference's context assembly.
//
// Replaces (reference): the FlaxCLIPTextModel the step trains (transformers modeling_flax_clip.py [third-party,
// un-vendored]; loaded at /root/reference/training_utils.py:215-217, called at 635-640): FlaxCLIPTextEmbeddings,
// FlaxCLIPAttention (query * scale, causal mask, softmax, weights @ value), ACT2FN[hidden_act] of FlaxCLIPMLP; and
// training_utils.py:641-674 (reshape to [B, chunks, 77, D], strip BOS / EOS, concatenate).
//
// The encoder's GEMMs, LayerNorms and bias gradients reuse the UNet's kernels (gemm_tc.cu, norms.cu, elementwise.cu). A
// CLIP sequence is 77 tokens: attention is 12 x 77 x 77 x 64 per sequence — latency, not throughput — so it runs on CUDA
// cores with one CTA per (sequence, head) and the whole problem in shared memory; the tcgen05 flash kernels of
// attention.cu are built for the UNet's 4096-token maps and have no causal mask.
#include "sdb_common.cuh"
#include "../../include/sdb200.h"
#include <mutex>

namespace {

constexpr int kSmallMaxS = 128;

__device__ __forceinline__ float bf(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// out[n, s, :] = bf16(tok[ids[n, s], :] + pos[s, :])   (bf16 operands: flax nn.Embed(dtype=bf16) casts the tables)
__global__ void __launch_bounds__(256)
embed_fwd_kernel(const int* __restrict__ ids, const __nv_bfloat16* __restrict__ tok, const __nv_bfloat16* __restrict__ pos,
                 __nv_bfloat16* __restrict__ out, int rows, int S, int D, int vocab) {
  sdb::pdl_trigger();
  const long long total = (long long)rows * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / D), c = (int)(i - (long long)r * D);
    int id = ids[r];
    id = min(max(id < 0 ? id + vocab : id, 0), vocab - 1);  // jnp.take semantics: never out of bounds
    out[i] = __float2bfloat16_rn(bf(tok + (long long)id * D + c) + bf(pos + (long long)(r % S) * D + c));
  }
}

// dtok[ids[r], :] += d_out[r, :] (fp32 atomics: a token may repeat), dpos[s, :] = sum_n d_out[n * S + s, :]
__global__ void __launch_bounds__(256)
embed_bwd_kernel(const int* __restrict__ ids, const __nv_bfloat16* __restrict__ dout, float* __restrict__ dtok,
                 float* __restrict__ dpos, int rows, int S, int D, int vocab) {
  sdb::pdl_trigger();
  const long long total = (long long)rows * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / D), c = (int)(i - (long long)r * D);
    int id = ids[r];
    id = min(max(id < 0 ? id + vocab : id, 0), vocab - 1);
    atomicAdd(dtok + (long long)id * D + c, bf(dout + i));
  }
  const long long ptotal = (long long)S * D;
  const int nseq = rows / S;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < ptotal; i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;  // fixed order over the sequences: deterministic
    for (int n = 0; n < nseq; ++n) acc += bf(dout + (long long)n * ptotal + i);
    dpos[i] = acc;
  }
}

// ---- short-sequence attention: one CTA per (sequence, head), thread t = query row t (forward) / key row t (2nd half
// of backward); q / k / v / o: [N * S, ld] matrices with the heads side by side in the columns ----
struct SmallAttn {
  const __nv_bfloat16 *q, *k, *v, *o, *dO;
  __nv_bfloat16 *out, *dq, *dk, *dv;
  float* lse;
  long long ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv;
  int S, d, H, causal;
  float scale;
};

__global__ void __launch_bounds__(128) attn_small_fwd_kernel(const SmallAttn p) {
  sdb::pdl_trigger();
  extern __shared__ float sm[];
  const int S = p.S, d = p.d, dp = d + 1;
  float* sQ = sm;
  float* sK = sQ + S * dp;
  float* sV = sK + S * dp;
  float* sP = sV + S * dp;  // [S][S + 1]
  const int h = blockIdx.x, n = blockIdx.y, t = threadIdx.x;
  const long long row0 = (long long)n * S;
  for (int i = t; i < S * d; i += blockDim.x) {
    const int r = i / d, c = i - r * d;
    sQ[r * dp + c] = bf(p.q + (row0 + r) * p.ldq + h * d + c);
    sK[r * dp + c] = bf(p.k + (row0 + r) * p.ldk + h * d + c);
    sV[r * dp + c] = bf(p.v + (row0 + r) * p.ldv + h * d + c);
  }
  __syncthreads();
  if (t >= S) return;
  const int jend = p.causal ? t + 1 : S;
  float* pr = sP + t * (S + 1);
  float mx = -INFINITY;
  for (int j = 0; j < jend; ++j) {
    float acc = 0.f;
    for (int c = 0; c < d; ++c) acc = fmaf(sQ[t * dp + c], sK[j * dp + c], acc);
    acc *= p.scale;
    pr[j] = acc;
    mx = fmaxf(mx, acc);
  }
  float l = 0.f;
  for (int j = 0; j < jend; ++j) {
    const float e = __expf(pr[j] - mx);
    pr[j] = e;
    l += e;
  }
  const float inv = 1.0f / l;
  __nv_bfloat16* orow = p.out + (row0 + t) * p.ldo + h * d;
  for (int c0 = 0; c0 < d; c0 += 8) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int j = 0; j < jend; ++j) {
      const float pj = pr[j];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(pj, sV[j * dp + c0 + e], acc[e]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) orow[c0 + e] = __float2bfloat16_rn(acc[e] * inv);
  }
  p.lse[((long long)n * p.H + h) * S + t] = mx + __logf(l);
}

__global__ void __launch_bounds__(128) attn_small_bwd_kernel(const SmallAttn p) {
  sdb::pdl_trigger();
  extern __shared__ float sm[];
  const int S = p.S, d = p.d, dp = d + 1, sp = S + 1;
  float* sQ = sm;
  float* sK = sQ + S * dp;
  float* sV = sK + S * dp;
  float* sG = sV + S * dp;   // dO
  float* sP = sG + S * dp;   // [S][S + 1] probabilities
  float* sD = sP + S * sp;   // [S][S + 1] dS (scaled)
  const int h = blockIdx.x, n = blockIdx.y, t = threadIdx.x;
  const long long row0 = (long long)n * S;
  for (int i = t; i < S * d; i += blockDim.x) {
    const int r = i / d, c = i - r * d;
    sQ[r * dp + c] = bf(p.q + (row0 + r) * p.ldq + h * d + c);
    sK[r * dp + c] = bf(p.k + (row0 + r) * p.ldk + h * d + c);
    sV[r * dp + c] = bf(p.v + (row0 + r) * p.ldv + h * d + c);
    sG[r * dp + c] = bf(p.dO + (row0 + r) * p.lddo + h * d + c);
  }
  for (int i = t; i < S * sp; i += blockDim.x) {
    sP[i] = 0.f;
    sD[i] = 0.f;
  }
  __syncthreads();
  if (t < S) {  // query row t: P, dS, dQ
    const int jend = p.causal ? t + 1 : S;
    const float lse = p.lse[((long long)n * p.H + h) * S + t];
    float Dt = 0.f;
    for (int c = 0; c < d; ++c) Dt = fmaf(sG[t * dp + c], bf(p.o + (row0 + t) * p.ldo + h * d + c), Dt);
    for (int j = 0; j < jend; ++j) {
      float s = 0.f, dpv = 0.f;
      for (int c = 0; c < d; ++c) {
        s = fmaf(sQ[t * dp + c], sK[j * dp + c], s);
        dpv = fmaf(sG[t * dp + c], sV[j * dp + c], dpv);
      }
      const float pj = __expf(s * p.scale - lse);
      sP[t * sp + j] = pj;
      sD[t * sp + j] = pj * (dpv - Dt) * p.scale;
    }
    __nv_bfloat16* qrow = p.dq + (row0 + t) * p.lddq + h * d;
    for (int c0 = 0; c0 < d; c0 += 8) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int j = 0; j < jend; ++j) {
        const float ds = sD[t * sp + j];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(ds, sK[j * dp + c0 + e], acc[e]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) qrow[c0 + e] = __float2bfloat16_rn(acc[e]);
    }
  }
  __syncthreads();
  if (t < S) {  // key row t: dV = P^T dO, dK = dS^T Q (rows i < t are zero under the causal mask)
    const int i0 = p.causal ? t : 0;
    __nv_bfloat16* krow = p.dk + (row0 + t) * p.lddk + h * d;
    __nv_bfloat16* vrow = p.dv + (row0 + t) * p.lddv + h * d;
    for (int c0 = 0; c0 < d; c0 += 8) {
      float ak[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, av[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int i = i0; i < S; ++i) {
        const float pij = sP[i * sp + t], dij = sD[i * sp + t];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          av[e] = fmaf(pij, sG[i * dp + c0 + e], av[e]);
          ak[e] = fmaf(dij, sQ[i * dp + c0 + e], ak[e]);
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        krow[c0 + e] = __float2bfloat16_rn(ak[e]);
        vrow[c0 + e] = __float2bfloat16_rn(av[e]);
      }
    }
  }
}

// kind 0: quick_gelu z * sigmoid(1.702 z); kind 1: gelu (erf)
__device__ __forceinline__ float act_f(float z, int kind) {
  if (kind == 0) return z / (1.0f + __expf(-1.702f * z));
  return 0.5f * z * (1.0f + erff(z * 0.70710678118654752f));
}
__device__ __forceinline__ float act_grad(float z, int kind) {
  if (kind == 0) {
    const float sg = 1.0f / (1.0f + __expf(-1.702f * z));
    return sg * (1.0f + 1.702f * z * (1.0f - sg));
  }
  return 0.5f * (1.0f + erff(z * 0.70710678118654752f)) + z * 0.3989422804014327f * __expf(-0.5f * z * z);
}
__global__ void __launch_bounds__(256)
act_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n, int kind) {
  sdb::pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16_rn(act_f(bf(x + i), kind));
}
__global__ void __launch_bounds__(256)
act_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
               long long n, int kind) {
  sdb::pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = __float2bfloat16_rn(bf(dy + i) * act_grad(bf(x + i), kind));
}

// out[r, :] = src[idx[r], :]  /  dsrc[idx[r], :] = dout[r, :] (dsrc zeroed by the caller; every source row used <= once)
__global__ void __launch_bounds__(256)
gather_rows_kernel(const __nv_bfloat16* __restrict__ src, const int* __restrict__ idx, __nv_bfloat16* __restrict__ out,
                   long long rows, int cols, int scatter) {
  sdb::pdl_trigger();
  const long long total = rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    if (scatter) out[(long long)idx[r] * cols + c] = src[i];
    else out[i] = src[(long long)idx[r] * cols + c];
  }
}

int grid_for(long long n) {
  long long blocks = (n + 255) / 256;
  long long cap = (long long)sdb::kNumSMs * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

extern "C" int sdb_embed_fwd(const int* input_ids, const void* token_embedding_bf16, const void* position_embedding_bf16,
                             void* out_bf16, int rows, int seq_len, int hidden, int vocab, cudaStream_t stream) {
  SDB_CHECK_ARG(rows > 0 && seq_len > 0 && hidden > 0 && vocab > 0 && rows % seq_len == 0, "sdb_embed_fwd: bad shape");
  embed_fwd_kernel<<<grid_for((long long)rows * hidden), 256, 0, stream>>>(
      input_ids, (const __nv_bfloat16*)token_embedding_bf16, (const __nv_bfloat16*)position_embedding_bf16,
      (__nv_bfloat16*)out_bf16, rows, seq_len, hidden, vocab);
  SDB_CUDA_LAUNCH_CHECK("sdb_embed_fwd");
  return SDB_OK;
}

extern "C" int sdb_embed_bwd(const int* input_ids, const void* dout_bf16, float* dtoken_embedding,
                             float* dposition_embedding, int rows, int seq_len, int hidden, int vocab,
                             cudaStream_t stream) {
  SDB_CHECK_ARG(rows > 0 && seq_len > 0 && hidden > 0 && vocab > 0 && rows % seq_len == 0, "sdb_embed_bwd: bad shape");
  const cudaError_t me = cudaMemsetAsync(dtoken_embedding, 0, (size_t)vocab * hidden * sizeof(float), stream);
  if (me != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_embed_bwd: memset: %s", cudaGetErrorString(me));
  embed_bwd_kernel<<<grid_for((long long)rows * hidden), 256, 0, stream>>>(
      input_ids, (const __nv_bfloat16*)dout_bf16, dtoken_embedding, dposition_embedding, rows, seq_len, hidden, vocab);
  SDB_CUDA_LAUNCH_CHECK("sdb_embed_bwd");
  return SDB_OK;
}

static int small_attn_check(const char* name, int N, int H, int S, int d) {
  SDB_CHECK_ARG(N > 0 && H > 0 && S > 0 && S <= kSmallMaxS && d > 0 && d % 8 == 0 && d <= 128 && H <= 65535 && N <= 65535,
                "%s: needs 1 <= S <= %d, d %% 8 == 0, d <= 128 (got S %d d %d)", name, kSmallMaxS, S, d);
  return SDB_OK;
}

extern "C" int sdb_attn_small_fwd(const void* q, long long ldq, const void* k, long long ldk, const void* v,
                                  long long ldv, void* o, long long ldo, float* lse, int N, int H, int S, int d,
                                  float scale, int causal, cudaStream_t stream) {
  int rc = small_attn_check("sdb_attn_small_fwd", N, H, S, d);
  if (rc) return rc;
  SmallAttn p;
  memset(&p, 0, sizeof(p));
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.out = (__nv_bfloat16*)o; p.lse = lse;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.ldo = ldo;
  p.S = S; p.d = d; p.H = H; p.causal = causal; p.scale = scale;
  const size_t smem = ((size_t)3 * S * (d + 1) + (size_t)S * (S + 1)) * sizeof(float);
  static std::once_flag once;
  static cudaError_t aerr = cudaSuccess;
  std::call_once(once, [] {
    aerr = cudaFuncSetAttribute(attn_small_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  if (aerr != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_small_fwd: %s", cudaGetErrorString(aerr));
  attn_small_fwd_kernel<<<dim3(H, N), 128, smem, stream>>>(p);
  SDB_CUDA_LAUNCH_CHECK("sdb_attn_small_fwd");
  return SDB_OK;
}

extern "C" int sdb_attn_small_bwd(const void* q, long long ldq, const void* k, long long ldk, const void* v,
                                  long long ldv, const void* o, long long ldo, const void* dO, long long lddo,
                                  const float* lse, void* dq, long long lddq, void* dk, long long lddk, void* dv,
                                  long long lddv, int N, int H, int S, int d, float scale, int causal,
                                  cudaStream_t stream) {
  int rc = small_attn_check("sdb_attn_small_bwd", N, H, S, d);
  if (rc) return rc;
  SmallAttn p;
  memset(&p, 0, sizeof(p));
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.o = (const __nv_bfloat16*)o; p.dO = (const __nv_bfloat16*)dO; p.lse = const_cast<float*>(lse);
  p.dq = (__nv_bfloat16*)dq; p.dk = (__nv_bfloat16*)dk; p.dv = (__nv_bfloat16*)dv;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.ldo = ldo; p.lddo = lddo; p.lddq = lddq; p.lddk = lddk; p.lddv = lddv;
  p.S = S; p.d = d; p.H = H; p.causal = causal; p.scale = scale;
  const size_t smem = ((size_t)4 * S * (d + 1) + (size_t)2 * S * (S + 1)) * sizeof(float);
  if (smem > 227 * 1024) return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_attn_small_bwd: S %d x d %d needs %zu B of smem", S, d, smem);
  static std::once_flag once;
  static cudaError_t aerr = cudaSuccess;
  std::call_once(once, [] {
    aerr = cudaFuncSetAttribute(attn_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  if (aerr != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_small_bwd: %s", cudaGetErrorString(aerr));
  attn_small_bwd_kernel<<<dim3(H, N), 128, smem, stream>>>(p);
  SDB_CUDA_LAUNCH_CHECK("sdb_attn_small_bwd");
  return SDB_OK;
}

extern "C" int sdb_act_fwd(const void* x, void* y, long long n, int kind, cudaStream_t stream) {
  SDB_CHECK_ARG(n > 0 && (kind == 0 || kind == 1), "sdb_act_fwd: n > 0 and kind 0 (quick_gelu) / 1 (gelu) required");
  act_fwd_kernel<<<grid_for(n), 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n, kind);
  SDB_CUDA_LAUNCH_CHECK("sdb_act_fwd");
  return SDB_OK;
}

extern "C" int sdb_act_bwd(const void* x, const void* dy, void* dx, long long n, int kind, cudaStream_t stream) {
  SDB_CHECK_ARG(n > 0 && (kind == 0 || kind == 1), "sdb_act_bwd: n > 0 and kind 0 (quick_gelu) / 1 (gelu) required");
  act_bwd_kernel<<<grid_for(n), 256, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n,
                                                  kind);
  SDB_CUDA_LAUNCH_CHECK("sdb_act_bwd");
  return SDB_OK;
}

extern "C" int sdb_gather_rows_bf16(const void* src, const int* row_index, void* out, long long rows, int cols,
                                    long long scatter_out_rows, cudaStream_t stream) {
  SDB_CHECK_ARG(rows > 0 && cols > 0 && scatter_out_rows >= 0, "sdb_gather_rows_bf16: bad shape");
  if (scatter_out_rows > 0) {
    const cudaError_t me = cudaMemsetAsync(out, 0, (size_t)scatter_out_rows * cols * sizeof(__nv_bfloat16), stream);
    if (me != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_gather_rows_bf16: memset: %s", cudaGetErrorString(me));
  }
  gather_rows_kernel<<<grid_for(rows * cols), 256, 0, stream>>>((const __nv_bfloat16*)src, row_index, (__nv_bfloat16*)out,
                                                                rows, cols, scatter_out_rows > 0 ? 1 : 0);
  SDB_CUDA_LAUNCH_CHECK("sdb_gather_rows_bf16");
  return SDB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Frozen VAE encoder pieces (diffusers-0.21.4 FlaxAutoencoderKL.encode [third-party, un-vendored], called at
// /root/reference/training_utils.py:574-586): pixel layout change with channel padding, the single-head 512-wide
// attention of the encoder's mid block as GEMM -> row softmax -> GEMM, and the latent sample.
namespace {

// src NCHW fp32 [B, C, HW] -> dst NHWC bf16 [B, HW, Cpad] (channels C..Cpad-1 zero: conv_in runs on 4-channel pixels)
__global__ void __launch_bounds__(256)
nchw_to_nhwc_pad_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int B, int C, int Cpad, int HW) {
  sdb::pdl_trigger();
  const long long total = (long long)B * Cpad * HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cpad);
    const long long bp = i / Cpad;
    const int p = (int)(bp % HW);
    const int b = (int)(bp / HW);
    dst[i] = __float2bfloat16_rn(c < C ? src[((long long)b * C + c) * HW + p] : 0.0f);
  }
}

// out[r, :] = softmax(scale * x[r, :]) over bf16 rows; one CTA per row, fp32 statistics
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int cols, long long ld,
                    float scale) {
  sdb::pdl_trigger();
  const __nv_bfloat16* xr = x + (long long)blockIdx.x * ld;
  __nv_bfloat16* orow = out + (long long)blockIdx.x * ld;
  __shared__ float red[8];
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < cols; c += 256) mx = fmaxf(mx, bf(xr + c) * scale);
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int c = threadIdx.x; c < cols; c += 256) sum += __expf(bf(xr + c) * scale - mx);
  sum = sdb::warp_sum(sum);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  sum = 0.f;
  for (int i = 0; i < 8; ++i) sum += red[i];
  const float inv = 1.0f / sum;
  for (int c = threadIdx.x; c < cols; c += 256) orow[c] = __float2bfloat16_rn(__expf(bf(xr + c) * scale - mx) * inv);
}

// latents NCHW fp32 [B, L, HW] = (mean + exp(0.5 * clip(logvar, -30, 20)) * eps) * scaling; moments NHWC bf16 [B, HW, 2L]
// (mean | logvar), eps fp32 in NHWC order [B, HW, L] (FlaxDiagonalGaussianDistribution.sample draws in the NHWC shape)
__global__ void __launch_bounds__(256)
vae_sample_kernel(const __nv_bfloat16* __restrict__ moments, const float* __restrict__ eps, float* __restrict__ latents,
                  int B, int L, int HW, float scaling) {
  sdb::pdl_trigger();
  const long long total = (long long)B * L * HW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(i % HW);
    const long long bc = i / HW;
    const int c = (int)(bc % L);
    const int b = (int)(bc / L);
    const long long row = (long long)b * HW + p;
    const float mean = bf(moments + row * 2 * L + c);
    const float logvar = fminf(fmaxf(bf(moments + row * 2 * L + L + c), -30.0f), 20.0f);
    const float e = eps ? eps[row * L + c] : 0.0f;
    latents[i] = (mean + __expf(0.5f * logvar) * e) * scaling;
  }
}

}  // namespace

extern "C" int sdb_nchw_f32_to_nhwc_bf16_pad(const float* src, void* dst, int B, int C, int Cpad, int HW,
                                             cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && C > 0 && Cpad >= C && HW > 0, "sdb_nchw_f32_to_nhwc_bf16_pad: bad shape");
  nchw_to_nhwc_pad_kernel<<<grid_for((long long)B * Cpad * HW), 256, 0, stream>>>(src, (__nv_bfloat16*)dst, B, C, Cpad, HW);
  SDB_CUDA_LAUNCH_CHECK("sdb_nchw_f32_to_nhwc_bf16_pad");
  return SDB_OK;
}

extern "C" int sdb_softmax_rows_bf16(const void* x, void* out, long long rows, int cols, long long ld, float scale,
                                     cudaStream_t stream) {
  SDB_CHECK_ARG(rows > 0 && rows < (1ll << 31) && cols > 0 && ld >= cols, "sdb_softmax_rows_bf16: bad shape");
  softmax_rows_kernel<<<(unsigned)rows, 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, cols, ld, scale);
  SDB_CUDA_LAUNCH_CHECK("sdb_softmax_rows_bf16");
  return SDB_OK;
}

extern "C" int sdb_vae_sample(const void* moments_nhwc_bf16, const float* eps_nhwc, float* latents_nchw, int B,
                              int latent_channels, int HW, float scaling, cudaStream_t stream) {
  SDB_CHECK_ARG(B > 0 && latent_channels > 0 && HW > 0, "sdb_vae_sample: bad shape");
  vae_sample_kernel<<<grid_for((long long)B * latent_channels * HW), 256, 0, stream>>>(
      (const __nv_bfloat16*)moments_nhwc_bf16, eps_nhwc, latents_nchw, B, latent_channels, HW, scaling);
  SDB_CUDA_LAUNCH_CHECK("sdb_vae_sample");
  return SDB_OK;
}
