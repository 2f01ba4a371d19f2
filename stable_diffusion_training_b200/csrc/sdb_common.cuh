// sdb200 — shared device/host helpers for the sm_100a kernels.
// Everything here is internal; the public surface is include/sdb200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

// ---- error plumbing (thread-local string, negative enum; never throws, never exits) ----
enum SdbStatus : int {
  SDB_OK = 0,
  SDB_ERR_INVALID = -1,    // bad argument (shape / alignment / enum)
  SDB_ERR_CUDA = -2,       // a CUDA runtime / driver call failed
  SDB_ERR_WORKSPACE = -3,  // workspace too small
  SDB_ERR_UNSUPPORTED = -4 // shape outside what the kernels cover
};

int sdb_set_error(int code, const char* fmt, ...);
void sdb_note_launch();  // bumps the counter behind sdb_launch_count() (api.cu)

#define SDB_CHECK_ARG(cond, ...)                                   \
  do {                                                             \
    if (!(cond)) return sdb_set_error(SDB_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define SDB_CUDA_LAUNCH_CHECK(name)                                                      \
  do {                                                                                   \
    cudaError_t e__ = cudaGetLastError();                                                \
    sdb_note_launch();                                                                   \
    if (e__ != cudaSuccess)                                                              \
      return sdb_set_error(SDB_ERR_CUDA, "%s: launch failed: %s", name, cudaGetErrorString(e__)); \
  } while (0)

// ---- small device utilities ----
namespace sdb {

constexpr int kNumSMs = 148;  // B200

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// 128-bit streaming loads/stores (data touched once: keep it out of L1)
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st_f4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 ld_u4(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st_u4(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
// 32-byte store (STG.256, sm_100): one full DRAM sector per thread instead of two half-sector 16-byte stores
__device__ __forceinline__ void st_u8(void* p, uint4 a, uint4 b) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z),
               "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(h);
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }
// d/dx silu(x) = s + x*s*(1-s), s = sigmoid(x)
__device__ __forceinline__ float silu_grad_f(float x) {
  float s = 1.0f / (1.0f + __expf(-x));
  return s * (1.0f + x * (1.0f - s));
}

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch (PDL) ----
// Every kernel of the library executes pdl_trigger() first: once all CTAs of a grid have started, the NEXT kernel in
// the stream may begin launching — but only if it was launched with launch_pdl(), and such kernels run their prologue
// (barrier init, TMEM allocation, descriptor prefetch) and then block in pdl_wait() until the previous grid has
// completed and flushed. Kernels launched with <<<>>> keep the ordinary full dependency. Inside a CUDA graph the pair
// becomes a programmatic edge: the ~2-4 us of launch latency + prologue per tensor-core kernel overlap the tail of
// its predecessor.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace sdb
