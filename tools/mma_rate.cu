// Microbenchmark: cycles per tcgen05.mma (M=128, K=16, bf16) as a function of N, for A from shared memory (SS) and
// A from tensor memory (TS). Settles how tile widths should be chosen in gemm_tc.cu / attention.cu.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate tools/mma_rate.cu -lcuda && ./mma_rate
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../stable_diffusion_training_b200/csrc/ptx_sm100.cuh"

__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mode: 0 SS (B K-major), 1 SS (B MN-major), 2 TS (B K-major), 3 TS (B MN-major), 4 SS alternating accumulators,
//       5 SS with TWO issuing warps (separate accumulators), 6 SS M=64, 7 TS with two issuing warps
__global__ void __launch_bounds__(128, 1) rate_kernel(int mode, int n, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;              // 128 x 64 bf16 (16 KB)
  uint8_t* sB = smem + 16384;      // 256 x 64 bf16 (32 KB) K-major, or 4 boxes of 64x64 MN-major
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::mbar_init(bar + 1, 1);
    ptx::fence_barrier_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc(tptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tptr;
  if (threadIdx.x == 0 || (threadIdx.x == 32 && (mode == 5 || mode == 7))) {
    const int w = threadIdx.x >> 5;
    uint64_t* mybar = bar + w;
    const bool b_mn = (mode & 1) && mode < 4;
    const bool ts = (mode == 2 || mode == 3 || mode == 7);
    const uint32_t idesc = ptx::umma_idesc_bf16(n, 0, b_mn ? 1 : 0, mode == 6 ? 64 : 128);
    const uint32_t tD = tmem + (uint32_t)w * 128u;  // second issuer: its own accumulator (N <= 128 there)
    const uint64_t dA = ptx::umma_desc_sw128(ptx::smem_u32(sA), 0u, 1024u);
    const uint64_t dB = ptx::umma_desc_sw128(ptx::smem_u32(sB), b_mn ? 8192u : 0u, 1024u);
    const uint32_t tA = tmem + 256u + (uint32_t)w * 64u;  // A operand region for TS (garbage values: timing only)
    // warm-up
    for (int k = 0; k < 4; ++k) ptx::umma_bf16(tD, dA + 2u * k, dB + (b_mn ? 128u * k : 2u * k), idesc, 1u);
    ptx::umma_commit(mybar);
    ptx::mbar_wait(mybar, 0);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t db = dB + (b_mn ? 128u * k : 2u * k);
        const uint32_t d = (mode == 4) ? tmem + (uint32_t)((it * 4 + k) & 1) * 256u : tD;
        if (ts) umma_bf16_ts(d, tA + 8u * k, db, idesc, 1u);
        else ptx::umma_bf16(d, dA + 2u * k, db, idesc, 1u);
      }
    }
    const long long t_issue = clock64();
    ptx::umma_commit(mybar);
    ptx::mbar_wait(mybar, 1);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && w == 0) {
      out[0] = t1 - t0;
      out[1] = t_issue - t0;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}


__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

// Same measurement with the WHOLE warp running the issue loop (all values warp-uniform) and only the instruction itself
// under elect.sync: the compiler keeps descriptors in uniform registers instead of the ELECT / R2UR.BROADCAST / BRA.U.ANY
// sequence it emits around every tcgen05.mma that sits in divergent (`if (lane == 0)`) code.
// mode: 0 SS one issuing warp, 5 SS two issuing warps, 2 TS one warp, 7 TS two warps
__global__ void __launch_bounds__(128, 1) rate_kernel_uniform(int mode, int n, int iters, long long* out, int group) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = smem + 16384;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::mbar_init(bar + 1, 1);
    ptx::fence_barrier_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc(tptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tptr;
  const int w = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (w == 0 || (w == 1 && (mode == 5 || mode == 7))) {
    uint64_t* mybar = bar + w;
    const bool ts = (mode == 2 || mode == 7);
    const uint32_t idesc = ptx::umma_idesc_bf16(n, 0, 0, 128);
    const uint32_t tD = tmem + (uint32_t)w * 128u;
    const uint64_t dA = ptx::umma_desc_sw128(ptx::smem_u32(sA), 0u, 1024u);
    const uint64_t dB = ptx::umma_desc_sw128(ptx::smem_u32(sB), 0u, 1024u);
    const uint32_t tA = tmem + 256u + (uint32_t)w * 64u;
    if (elect_one()) {
      for (int k = 0; k < 4; ++k) ptx::umma_bf16(tD, dA + 2u * k, dB + 2u * k, idesc, 1u);
      ptx::umma_commit(mybar);
    }
    __syncwarp();
    ptx::mbar_wait(mybar, 0);
    const long long t0 = clock64();
    if (group) {  // one elect per group of four MMAs
      for (int it = 0; it < iters; ++it) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (ts) umma_bf16_ts(tD, tA + 8u * k, dB + 2u * k, idesc, 1u);
            else ptx::umma_bf16(tD, dA + 2u * k, dB + 2u * k, idesc, 1u);
          }
        }
        __syncwarp();
      }
    } else {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (elect_one()) {
          if (ts) umma_bf16_ts(tD, tA + 8u * k, dB + 2u * k, idesc, 1u);
          else ptx::umma_bf16(tD, dA + 2u * k, dB + 2u * k, idesc, 1u);
        }
      }
    }
    }
    const long long t_issue = clock64();
    if (elect_one()) ptx::umma_commit(mybar);
    __syncwarp();
    ptx::mbar_wait(mybar, 1);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t_issue - t0;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

int main() {
  long long* out;
  cudaMalloc(&out, 16);
  const int smem = 16384 + 32768 + 1024 + 64;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 2000;
  const char* names[8] = {"SS  B K-major ", "SS  B MN-major", "TS  B K-major ", "TS  B MN-major", "SS  alt-accum ", "SS  2 issuers ", "SS  M=64      ", "TS  2 issuers "};
  const int ns[] = {16, 32, 48, 64, 96, 128, 160, 192, 256};
  for (int grid : {148}) {
    printf("grid %d (cycles per MMA M=128 K=16; total incl. drain / issue-only)\n", grid);
    for (int mode : {0, 2, 5, 6, 7}) {
      for (int n : ns) {
        if ((mode == 5 || mode == 7) && n > 128) continue;
        rate_kernel<<<grid, 128, smem>>>(mode, n, iters, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf("%s N=%d: %s\n", names[mode], n, cudaGetErrorString(e));
          return 1;
        }
        long long h[2];
        cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
        printf("%s N=%3d  %7.1f  issue %6.1f\n", names[mode], n, (double)h[0] / (iters * 4.0), (double)h[1] / (iters * 4.0));
      }
    }
  }
  cudaFuncSetAttribute(rate_kernel_uniform, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  printf("warp-uniform issue loop (elect.sync around the instruction only)\n");
  for (int group : {0, 1})
  for (int mode : {0, 5, 2, 7}) {
    for (int n : ns) {
      if ((mode == 5 || mode == 7) && n > 128) continue;
      rate_kernel_uniform<<<148, 128, smem>>>(mode, n, iters, out, group);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("uniform mode %d N=%d: %s\n", mode, n, cudaGetErrorString(e));
        return 1;
      }
      long long h[2];
      cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
      printf("g%d %s N=%3d  %7.1f  issue %6.1f\n", group, names[mode], n, (double)h[0] / (iters * 4.0), (double)h[1] / (iters * 4.0));
    }
  }
  return 0;
}
