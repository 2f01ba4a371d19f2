// CUDA-core GEMM / implicit convolution with the same descriptor semantics as sdb_gemm_tc.
//
// Used (1) for the few layers whose shapes do not fill a tensor-core tile — time embedding (M = batch),
// conv_in (K = 4 per tap), conv_out (N = 4) — and (2) as the independent on-device cross-check of the tcgen05
// kernel in tests/test_gemm_gpu.py. fp32 accumulation, bf16 operands, no alignment constraints.
#include "sdb_common.cuh"
#include "../../include/sdb200.h"

namespace {

struct SimtParams {
  sdb_view4 a[SDB_MAX_SRC];
  sdb_view4 b;
  int b_mn, ntaps;
  int tap_src[SDB_MAX_TAPS], tap_dh[SDB_MAX_TAPS], tap_dw[SDB_MAX_TAPS], tap_brow[SDB_MAX_TAPS];
  int W, H, NI, M, N, K;
  void* d;
  long long ldd, dsW, dsH, dsN;
  const float* bias;
  const __nv_bfloat16* temb;
  long long ldt;
  const __nv_bfloat16* residual;
  long long ldr;
  int pixel_chunks;
};

__device__ __forceinline__ const __nv_bfloat16* view_ptr(const sdb_view4& v, int w, int h, int n) {
  if (w < 0 || w >= v.W || h < 0 || h >= v.H || n < 0 || n >= v.N) return nullptr;
  return reinterpret_cast<const __nv_bfloat16*>(v.ptr) + (long long)w * v.strideW + (long long)h * v.strideH +
         (long long)n * v.strideN;
}

// one thread per output element (pixel, n)
__global__ void __launch_bounds__(256) simt_fwd_kernel(const SimtParams p) {
  sdb::pdl_trigger();
  const long long total = (long long)p.W * p.H * p.NI * p.N;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % p.N);
    const long long pix = idx / p.N;
    const int w = (int)(pix % p.W);
    const int h = (int)((pix / p.W) % p.H);
    const int img = (int)(pix / ((long long)p.W * p.H));
    float acc = 0.0f;
    const __nv_bfloat16* bbase = reinterpret_cast<const __nv_bfloat16*>(p.b.ptr);
    for (int t = 0; t < p.ntaps; ++t) {
      const __nv_bfloat16* arow = view_ptr(p.a[p.tap_src[t]], w + p.tap_dw[t], h + p.tap_dh[t], img);
      if (arow == nullptr) continue;
      const int kmax = min(p.K, p.a[p.tap_src[t]].C);
      if (p.b_mn) {
        const __nv_bfloat16* bp = bbase + (long long)p.tap_brow[t] * p.b.strideW + n;
        for (int k = 0; k < kmax; ++k)
          acc = fmaf(__bfloat162float(arow[k]), __bfloat162float(bp[(long long)k * p.b.strideW]), acc);
      } else {
        const __nv_bfloat16* bp = bbase + ((long long)p.tap_brow[t] + n) * p.b.strideW;
        for (int k = 0; k < kmax; ++k) acc = fmaf(__bfloat162float(arow[k]), __bfloat162float(bp[k]), acc);
      }
    }
    if (p.bias) acc += p.bias[n];
    if (p.temb) acc += __bfloat162float(p.temb[(long long)img * p.ldt + n]);
    if (p.residual) acc += __bfloat162float(p.residual[pix * p.ldr + n]);
    reinterpret_cast<__nv_bfloat16*>(p.d)[(long long)img * p.dsN + (long long)h * p.dsH + (long long)w * p.dsW + n] =
        __float2bfloat16_rn(acc);
  }
}

// grid.x over (tap, m, n) elements, grid.y over pixel chunks; atomicAdd into the (zeroed) fp32 output
__global__ void __launch_bounds__(256) simt_wgrad_kernel(const SimtParams p) {
  sdb::pdl_trigger();
  const long long total = (long long)p.ntaps * p.M * p.N;
  const long long npix = (long long)p.W * p.H * p.NI;
  const long long per_chunk = (npix + p.pixel_chunks - 1) / p.pixel_chunks;
  const long long pix_begin = (long long)blockIdx.y * per_chunk;
  const long long pix_end = min(npix, pix_begin + per_chunk);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % p.N);
    const int m = (int)((idx / p.N) % p.M);
    const int t = (int)(idx / ((long long)p.N * p.M));
    const sdb_view4& av = p.a[p.tap_src[t]];
    float acc = 0.0f;
    for (long long pix = pix_begin; pix < pix_end; ++pix) {
      const int w = (int)(pix % p.W);
      const int h = (int)((pix / p.W) % p.H);
      const int img = (int)(pix / ((long long)p.W * p.H));
      const __nv_bfloat16* arow = view_ptr(av, w + p.tap_dw[t], h + p.tap_dh[t], img);
      const __nv_bfloat16* brow = view_ptr(p.b, w, h, img);
      if (arow == nullptr || brow == nullptr) continue;
      acc = fmaf(__bfloat162float(arow[m]), __bfloat162float(brow[n]), acc);
    }
    float* dst = reinterpret_cast<float*>(p.d) + ((long long)p.tap_brow[t] + m) * p.ldd + n;
    if (p.pixel_chunks > 1) atomicAdd(dst, acc);
    else *dst = acc;
  }
}

}  // namespace

extern "C" int sdb_gemm_simt(const sdb_gemm_desc* d, cudaStream_t stream) {
  SDB_CHECK_ARG(d != nullptr, "sdb_gemm_simt: null descriptor");
  SDB_CHECK_ARG(d->mode == SDB_GEMM_FWD || d->mode == SDB_GEMM_WGRAD, "sdb_gemm_simt: bad mode %d", d->mode);
  SDB_CHECK_ARG(d->num_src >= 1 && d->num_src <= SDB_MAX_SRC && d->ntaps >= 1 && d->ntaps <= SDB_MAX_TAPS,
                "sdb_gemm_simt: bad num_src/ntaps");
  SDB_CHECK_ARG(d->d != nullptr && d->N > 0, "sdb_gemm_simt: bad output");
  SimtParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < d->num_src; ++i) p.a[i] = d->a[i];
  p.b = d->b;
  p.b_mn = d->b_mn_major;
  p.ntaps = d->ntaps;
  for (int t = 0; t < d->ntaps; ++t) {
    SDB_CHECK_ARG(d->tap_src[t] >= 0 && d->tap_src[t] < d->num_src, "sdb_gemm_simt: tap %d names source %d", t,
                  d->tap_src[t]);
    p.tap_src[t] = d->tap_src[t];
    p.tap_dh[t] = d->tap_dh[t];
    p.tap_dw[t] = d->tap_dw[t];
    p.tap_brow[t] = d->tap_brow[t];
  }
  p.W = d->outW; p.H = d->outH; p.NI = d->outN;
  p.M = d->M; p.N = d->N; p.K = d->K;
  p.d = d->d; p.ldd = d->ldd;
  if (d->d_strideW || d->d_strideH || d->d_strideN) {
    p.dsW = d->d_strideW; p.dsH = d->d_strideH; p.dsN = d->d_strideN;
  } else {
    p.dsW = d->ldd; p.dsH = d->ldd * d->outW; p.dsN = d->ldd * d->outW * d->outH;
  }
  p.bias = d->bias;
  p.temb = reinterpret_cast<const __nv_bfloat16*>(d->temb); p.ldt = d->ldt;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(d->residual); p.ldr = d->ldr;
  if (d->mode == SDB_GEMM_FWD) {
    SDB_CHECK_ARG(d->K > 0, "sdb_gemm_simt: K must be positive");
    const long long total = (long long)p.W * p.H * p.NI * p.N;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    simt_fwd_kernel<<<(int)blocks, 256, 0, stream>>>(p);
  } else {
    SDB_CHECK_ARG(d->M > 0, "sdb_gemm_simt: WGRAD needs M");
    const long long total = (long long)p.ntaps * p.M * p.N;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 8) blocks = 148LL * 8;
    const long long npix = (long long)p.W * p.H * p.NI;
    int chunks = d->splits < 1 ? 1 : d->splits;
    if (chunks > npix) chunks = (int)npix;
    p.pixel_chunks = chunks;
    if (chunks > 1) {  // same contract as sdb_gemm_tc: the call zeroes the rows it accumulates into
      int last_row = 0;
      for (int t = 0; t < d->ntaps; ++t) last_row = d->tap_brow[t] > last_row ? d->tap_brow[t] : last_row;
      const cudaError_t me = cudaMemset2DAsync(d->d, (size_t)d->ldd * sizeof(float), 0, (size_t)d->N * sizeof(float),
                                               (size_t)last_row + (size_t)d->M, stream);
      if (me != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_simt: memset: %s", cudaGetErrorString(me));
    }
    dim3 grid((unsigned)blocks, (unsigned)chunks);
    simt_wgrad_kernel<<<grid, 256, 0, stream>>>(p);
  }
  SDB_CUDA_LAUNCH_CHECK("sdb_gemm_simt");
  return SDB_OK;
}
