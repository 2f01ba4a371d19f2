// Microbenchmark: bytes per cycle ONE SM takes in through TMA when all 148 SMs stream at once — the quantity that paces
// the GEMM main loop (every tile shape runs at 46-48 B/clk per SM). A persistent CTA per SM runs a
// TMA producer that fills a ring of `stages` shared-memory stages of `stage_kb` KB (boxes of 64 bf16 x 64 rows, like
// the GEMM's operand boxes) and a consumer that releases each stage as soon as it has landed (no MMA). The source is
// either L2-resident (32 MB, re-read by every SM) or far larger than L2 (HBM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ingest_rate tools/ingest_rate.cu -lcuda && ./tools/ingest_rate
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../stable_diffusion_training_b200/csrc/ptx_sm100.cuh"

constexpr int kBoxBytes = 64 * 128;  // 64 rows x 64 bf16 (the unit of the stage sizes below)

// box_mult: rows of one TMA box / 64 (1: 8 KB per instruction, 4: 32 KB per instruction like the GEMM's 256-pixel A box)
__global__ void __launch_bounds__(64, 1) ingest_kernel(const __grid_constant__ CUtensorMap map, int stages, int boxes,
                                                      int iters, int rows_total, long long* out, int box_mult) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * boxes * kBoxBytes);
  uint64_t* empty = full + 16;
  const int warp = ptx::warp_idx_uniform();
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    // every SM walks its own rows first, then wraps: neighbouring SMs do not share lines at the same time
    int row = (int)(((long long)blockIdx.x * 7919 * 64) % rows_total);
    for (int it = 0; it < iters; ++it) {
      ptx::mbar_wait(&empty[stage], phase ^ 1u);
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&full[stage], (uint32_t)(boxes * kBoxBytes));
        for (int j = 0; j < boxes; j += box_mult) {
          ptx::tma_load_4d(smem + ((size_t)stage * boxes + j) * kBoxBytes, &map, &full[stage], 0, row, 0, 0);
          row += 64 * box_mult;
          if (row + 64 * box_mult > rows_total) row = 0;
        }
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
      ptx::mbar_wait(&full[stage], phase);
      if (ptx::elect_one()) ptx::mbar_arrive(&empty[stage]);
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

int main() {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaFree(0);
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || !sym) {
    printf("no cuTensorMapEncodeTiled\n");
    return 1;
  }
  EncodeFn enc = reinterpret_cast<EncodeFn>(sym);
  long long* out;
  cudaMalloc(&out, 148 * sizeof(long long));
  cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  // {source MB, row pitch in bytes}: pitch 128 = every box is 8 KB contiguous; pitch 640 / 2560 = a 64-channel K block of
  // an NHWC activation with 320 / 1280 channels (128-byte pieces, one per pixel), which is what the implicit GEMM reads
  const size_t sizes_mb[5] = {32, 4096, 32, 32, 32};
  const size_t pitches[5] = {128, 128, 640, 2560, 640};
  const int box_mults[5] = {1, 1, 1, 1, 4};
  for (int si = 0; si < 5; ++si) {
    const int box_mult = box_mults[si];
    const size_t bytes = sizes_mb[si] << 20;
    const size_t pitch = pitches[si];
    void* buf;
    if (cudaMalloc(&buf, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(buf, 0, bytes);
    const int rows_total = (int)(bytes / pitch);
    CUtensorMap map;
    cuuint64_t gdim[4] = {64, (cuuint64_t)rows_total, 1, 1};
    cuuint64_t gstr[3] = {pitch, (cuuint64_t)rows_total * pitch, (cuuint64_t)rows_total * pitch};
    cuuint32_t box[4] = {64, (cuuint32_t)(64 * box_mult), 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, buf, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    printf("source %zu MB (%s), row pitch %zu B: B/clk per SM, all 148 SMs streaming (max cycles over SMs)\n",
           sizes_mb[si], sizes_mb[si] <= 64 ? "L2-resident" : "HBM", pitch);
    if (box_mult > 1) printf("  (one TMA instruction per %d KB)\n", 8 * box_mult);
    const int cfgs[][2] = {{2, 8}, {4, 4}, {4, 6}, {8, 2}, {8, 3}, {2, 12}, {7, 4}, {3, 8}};  // {boxes per stage, stages}
    for (auto& c : cfgs) {
      const int boxes = c[0], stages = c[1];
      const size_t smem = (size_t)stages * boxes * kBoxBytes + 1024 + 256;
      if (smem > 227 * 1024 || boxes % box_mult) continue;
      const int iters = 4000 / boxes;
      for (int rep = 0; rep < 2; ++rep) {
        ingest_kernel<<<148, 64, smem>>>(map, stages, boxes, iters, rows_total, out, box_mult);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
      }
      long long h[148], mx = 0;
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("  stage %3d KB x %d stages (%3d KB in flight): %6.1f B/clk/SM\n", boxes * 8, stages, boxes * 8 * stages,
             (double)iters * boxes * kBoxBytes / (double)mx);
    }
    cudaFree(buf);
  }
  return 0;
}
