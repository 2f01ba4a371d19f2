// Tensor-core GEMM / implicit-GEMM 3x3 convolution for sm_100a: TMA -> smem ring -> tcgen05.mma -> TMEM -> epilogue.
//
// Replaces flax nn.Conv / nn.Dense (forward, dgrad, wgrad) inside FlaxUNet2DConditionModel [diffusers 0.21.4,
// un-vendored; call site /root/reference/training_utils.py:678-684] — SURVEY §2.1 K1/K2.
//
// One persistent, warp-specialised kernel (one CTA per SM, 352 threads):
//   warp 0      : TMA producer  — 4-D tiled tensor maps over NHWC activations; a 3x3 tap is a box shifted by
//                 (dh, dw) whose out-of-bounds part TMA zero-fills (= the conv padding); weights are 2-D boxes.
//   warps 1, 6  : tcgen05.mma issuers, one per 128-row half of the tile (UMMA 128 x N x 16, bf16 -> fp32); warp 1 owns TMEM.
//   warps 2..5, 7..10 : two epilogue sets — tcgen05.ld of the accumulator (one row per thread), bias / time-embedding /
//                 residual add, bf16 (or fp32 / red.add for wgrad) stores.
// Two 256-column TMEM accumulators let the epilogue of tile i overlap the main loop of tile i+1.
//
// Operand layouts (all SWIZZLE_128B, produced directly by TMA):
//   activations as A (fwd/dgrad)     K-major : 128 pixel rows x 64 channels (128 B rows)
//   weights [K,N] as B (fwd)         MN-major: per 64-column box, 64 K rows x 128 B; boxes 8 KB apart (LBO)
//   weights [N,K] as B (dgrad)       K-major : N rows x 64 channels
//   activations as A and B (wgrad)   MN-major: per 64-channel box, 64 pixel rows x 128 B
#include "sdb_common.cuh"
#include "ptx_sm100.cuh"
#include "../../include/sdb200.h"
#include <cstdlib>
#include <mutex>

// The SDB_GEMM_DBG timing experiments (skip TMA / stores / MMA; profiles/r1f_gemm_experiments.md) are compiled in only
// with -DSDB_GEMM_EXPERIMENTS: the flag tests sit in the single-thread producer / issuer loops.
#ifdef SDB_GEMM_EXPERIMENTS
#define SDB_DBG(p, bit) ((p).dbg & (bit))
#else
#define SDB_DBG(p, bit) 0
#endif

// -DSDB_GEMM_TIMELINE: CTA 0 records clock64() per K block (its first 512) at the hand-off points of the operand ring —
// producer: stage free / TMA issued; issuers: stage landed / MMAs issued (tools/gemm_timeline.py). Off by default.
#ifdef SDB_GEMM_TIMELINE
__device__ long long g_gemm_tl[3 * 512 * 2];
#define SDB_GTL(role, k, slot)                                                              \
  do {                                                                                      \
    if (blockIdx.x == 0 && (k) < 512 && (threadIdx.x & 31) == 0)                            \
      g_gemm_tl[((role) * 512 + (k)) * 2 + (slot)] = clock64();                             \
  } while (0)
extern "C" int sdb_gemm_timeline(long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_gemm_tl, sizeof(long long) * 3 * 512 * 2);
}
#else
#define SDB_GTL(role, k, slot) do {} while (0)
#endif

namespace {

// Two sets of four epilogue warps (2-5 and 7-10). MT == 2: set e drains half e, which halves the tile-boundary gap the
// operand-ring timeline exposed (5.4 K of ~39 K cycles per 256-row tile, profiles/r3g_gemm_timeline.md); MT == 1: the sets
// alternate 32-column chunks. Measured r4a: conv3x3 320->320 @64x64 fwd 62.5 -> 58.4 us, FF linear 320->2560 fwd
// 80.9 -> 68.6 us, the step -0.6 ms (together with the lean producer -1.4 ms).
constexpr int kThreads = 352;                // warp 0 TMA, warps 1 / 6 MMA issuers (half 0 / 1), warps 2-5 and 7-10 epilogue
constexpr int kEpiSets = 2;
constexpr int kBlockK = 64;                 // bf16 elements per K block (one 128-byte swizzle row)
constexpr uint32_t kAStageBytes = 128 * 128;  // 16 KB: 128 rows x 128 B, or two 64-row MN-major boxes
constexpr uint32_t kBoxBytes = 64 * 128;      // one 64-row box
constexpr int kMaxStages = 8;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kAccCols = 256;

struct alignas(64) TcParams {
  CUtensorMap mapA[SDB_MAX_SRC];
  CUtensorMap mapB;
  int mode, b_mn, ntaps;
  int tap_src[SDB_MAX_TAPS], tap_dh[SDB_MAX_TAPS], tap_dw[SDB_MAX_TAPS], tap_brow[SDB_MAX_TAPS];
  int W, H, NI, bw, bh, bn, tw, th, tn;
  int M, N, K, BN;
  int tiles_m, tiles_n, total_tiles, kblocks;
  int splits, pb_per_split, total_pb;
  long long slab;  // FWD split-K: > 0 = elements between the per-split fp32 slabs (plain stores, deterministic); 0 = red.add
  int stages, dbg, mt;
  int st32;  // every output row segment the epilogue writes is 32-byte aligned: 32-byte stores
  uint32_t a_stage_bytes;
  uint32_t b_stage_bytes, stage_bytes;
  void* d;
  long long ldd, dsW, dsH, dsN;
  const float* bias;
  const __nv_bfloat16* temb;
  long long ldt;
  const __nv_bfloat16* residual;
  long long ldr;
};

struct Tile {
  int n_off, nw, umma_n, nboxes;
  int w0, h0, n0;          // FWD: pixel origin of the M tile
  int tap, m_off;          // WGRAD
  int kb_begin, num_k;     // K iterations
  int split;               // FWD split-K: index of this tile's K range
};

__device__ __forceinline__ Tile decode_tile(const TcParams& p, int t) {
  Tile ti;
  if (p.mode == SDB_GEMM_FWD) {
    // split-K (p.splits > 1): consecutive tile indices are the K ranges of one output tile
    int split = t % p.splits;
    int tt = t / p.splits;
    int n_tile = tt % p.tiles_n;
    int m_tile = tt / p.tiles_n;
    int iw = m_tile % p.tw;
    int r = m_tile / p.tw;
    int ih = r % p.th;
    int in = r / p.th;
    ti.w0 = iw * p.bw;
    ti.h0 = ih * p.bh;
    ti.n0 = in * p.bn;
    ti.n_off = n_tile * p.BN;
    ti.tap = 0;
    ti.m_off = 0;
    ti.kb_begin = split * p.pb_per_split;
    ti.num_k = min(p.total_pb, ti.kb_begin + p.pb_per_split) - ti.kb_begin;
    ti.split = split;
  } else {
    // m tile and tap vary fastest: the CTAs running side by side share one dY block (n tile, pixel range) and read
    // neighbouring x boxes, so dY comes from HBM once instead of once per (tap, m tile). With the n tile fastest the
    // FF weight gradients ([32768, 320]^T x [32768, 2560]: dY = 168 MB > L2) read 0.4 GB at 5.2 TB/s (r4c capture).
    int r = t;
    int m_tile = r % p.tiles_m;
    r /= p.tiles_m;
    ti.tap = r % p.ntaps;
    r /= p.ntaps;
    int split = r % p.splits;
    int n_tile = r / p.splits;
    ti.m_off = m_tile * 128 * p.mt;
    ti.n_off = n_tile * p.BN;
    ti.kb_begin = split * p.pb_per_split;
    int kb_end = min(p.total_pb, ti.kb_begin + p.pb_per_split);
    ti.num_k = kb_end - ti.kb_begin;
    ti.w0 = ti.h0 = ti.n0 = 0;
    ti.split = split;
  }
  ti.nw = min(p.BN, p.N - ti.n_off);
  ti.umma_n = (ti.nw + 15) & ~15;
  ti.nboxes = (ti.umma_n + 63) >> 6;
  return ti;
}

// MT = number of 128-row halves of the M tile (1: 128 x BN with two accumulators double-buffered across tiles;
// 2: 256 x BN, one accumulator and one issuing warp per half — the B stage is fetched once for 256 rows, which is what
// matters when the SM's ingest of TMA bytes, not the tensor pipe, paces the main loop).
template <int MODE, int MT>
__global__ void __launch_bounds__(kThreads, 1) gemm_tc_kernel(const __grid_constant__ TcParams p) {
  sdb::pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * p.stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kMaxStages;
  uint64_t* tmem_full = bars + 2 * kMaxStages;
  uint64_t* tmem_empty = bars + 2 * kMaxStages + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * kMaxStages + 4);

  const int warp = ptx::warp_idx_uniform();  // provably warp-uniform: see ptx::elect_one
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < SDB_MAX_SRC; ++i) ptx::tma_prefetch_desc(&p.mapA[i]);
    ptx::tma_prefetch_desc(&p.mapB);
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], MT);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(&tmem_full[a], 1);
      ptx::mbar_init(&tmem_empty[a], (kEpiSets == 2 && MT == 1) ? 256 : 128);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  sdb::pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  if (warp == 0) {
    // ===================== TMA producer (whole warp, warp-uniform values; the TMA instructions under elect) =========
    {
      int stage = 0;
      uint32_t phase = 0;
      int gk = 0;  // K blocks of this CTA so far (timeline index)
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        const Tile ti = decode_tile(p, t);
        // Everything that depends on the tap or the pixel block is carried incrementally: with an integer division, four
        // tap-indexed parameter loads and the box arithmetic in front of every TMA one producer iteration cost 500-600
        // clk and paced the 128-row GEMMs of the 16x16 / 8x8 levels (MMAs 320 clk, fetch latency 440 clk: r3g timeline).
        // Measured r4a: conv3x3 1280->1280 @16x16 fwd 76.8 -> 68.6 us, dgrad 59.4 -> 54.3 us, the step -0.9 ms.
        int f_tap = 0, f_kb = 0, f_cw = 0, f_ch = 0, f_brow = 0, f_iw = 0, f_ih = 0, f_in = 0;
        const CUtensorMap* f_ma = &p.mapA[0];
        if (MODE == SDB_GEMM_FWD) {
          f_tap = ti.kb_begin / p.kblocks;
          f_kb = ti.kb_begin - f_tap * p.kblocks;
          f_ma = &p.mapA[p.tap_src[f_tap]];
          f_cw = ti.w0 + p.tap_dw[f_tap];
          f_ch = ti.h0 + p.tap_dh[f_tap];
          f_brow = p.tap_brow[f_tap];
        } else {
          f_iw = ti.kb_begin % p.tw;
          const int r = ti.kb_begin / p.tw;
          f_ih = r % p.th;
          f_in = r / p.th;
          f_ma = &p.mapA[p.tap_src[ti.tap]];
          f_cw = p.tap_dw[ti.tap];
          f_ch = p.tap_dh[ti.tap];
        }
        const uint32_t f_bytes = p.a_stage_bytes + ((MODE == SDB_GEMM_FWD && !p.b_mn) ? p.b_stage_bytes
                                                                                        : (uint32_t)ti.nboxes * kBoxBytes);
        for (int it = 0; it < ti.num_k; ++it, ++gk) {
          ptx::mbar_wait_sleep(&empty_bar[stage], phase ^ 1u);
          SDB_GTL(0, gk, 0);
          uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
          uint8_t* sb = sa + p.a_stage_bytes;
          if (ptx::elect_one()) {
            ptx::mbar_expect_tx(&full_bar[stage], f_bytes);
            if (MODE == SDB_GEMM_FWD) {
              ptx::tma_load_4d(sa, f_ma, &full_bar[stage], f_kb * kBlockK, f_cw, f_ch, ti.n0);
              if (p.b_mn) {
                for (int j = 0; j < ti.nboxes; ++j)
                  ptx::tma_load_4d(sb + j * kBoxBytes, &p.mapB, &full_bar[stage], ti.n_off + 64 * j,
                                   f_brow + f_kb * kBlockK, 0, 0);
              } else {
                ptx::tma_load_4d(sb, &p.mapB, &full_bar[stage], f_kb * kBlockK, f_brow + ti.n_off, 0, 0);
              }
            } else {
              const int pw0 = f_iw * p.bw, ph0 = f_ih * p.bh, pn0 = f_in * p.bn;
              for (int i = 0; i < 2 * MT; ++i)
                ptx::tma_load_4d(sa + i * kBoxBytes, f_ma, &full_bar[stage], ti.m_off + 64 * i, pw0 + f_cw, ph0 + f_ch, pn0);
              for (int j = 0; j < ti.nboxes; ++j)
                ptx::tma_load_4d(sb + j * kBoxBytes, &p.mapB, &full_bar[stage], ti.n_off + 64 * j, pw0, ph0, pn0);
            }
          }
          __syncwarp();
          SDB_GTL(0, gk, 1);
          if (MODE == SDB_GEMM_FWD) {
            if (++f_kb == p.kblocks && it + 1 < ti.num_k) {  // next tap
              f_kb = 0;
              ++f_tap;
              f_ma = &p.mapA[p.tap_src[f_tap]];
              f_cw = ti.w0 + p.tap_dw[f_tap];
              f_ch = ti.h0 + p.tap_dh[f_tap];
              f_brow = p.tap_brow[f_tap];
            }
          } else if (++f_iw == p.tw) {  // next pixel block: w fastest, then h, then image
            f_iw = 0;
            if (++f_ih == p.th) { f_ih = 0; ++f_in; }
          }
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1 || warp == 6) {
    // ===================== MMA issuers: warp 1 -> rows 0..127 of the tile, warp 6 -> rows 128..255 (MT == 2) ==========
    const int half = (warp == 6) ? 1 : 0;
    if (half < MT) {  // whole warp; the MMAs and commits under elect with descriptors in uniform registers
      int stage = 0;
      uint32_t phase = 0;
      int tile_count = 0;
      int gk = 0;
      const int a_mn = (MODE == SDB_GEMM_WGRAD) ? 1 : 0;
      const int b_mn = (MODE == SDB_GEMM_WGRAD) ? 1 : p.b_mn;
      const uint32_t a_lbo = a_mn ? kBoxBytes : 0u, a_kadv = a_mn ? 2048u : 32u;
      const uint32_t b_lbo = b_mn ? kBoxBytes : 0u, b_kadv = b_mn ? 2048u : 32u;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++tile_count) {
        const Tile ti = decode_tile(p, t);
        // MT == 1: accumulators alternate per tile; MT == 2: accumulator = half, reused every tile
        const int acc = (MT == 1) ? (tile_count & 1) : half;
        const uint32_t acc_phase = (MT == 1) ? ((uint32_t)(tile_count >> 1) & 1u) : ((uint32_t)tile_count & 1u);
        ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)acc * kAccCols;
        const uint32_t idesc = ptx::umma_idesc_bf16(ti.umma_n, a_mn, b_mn);
        // a WGRAD half whose dW rows all lie past M has nothing to compute (it still takes part in the barriers)
        const bool live = (MODE != SDB_GEMM_WGRAD) || (ti.m_off + half * 128 < p.M);
        for (int it = 0; it < ti.num_k; ++it, ++gk) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          SDB_GTL(1 + half, gk, 0);
          const uint32_t sa = ptx::smem_u32(smem + (size_t)stage * p.stage_bytes) + (uint32_t)half * kAStageBytes;
          const uint32_t sb = ptx::smem_u32(smem + (size_t)stage * p.stage_bytes) + p.a_stage_bytes;
          if (ptx::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < kBlockK / 16; ++ks) {
              const uint64_t da = ptx::umma_desc_sw128(sa + ks * a_kadv, a_lbo, 1024u);
              const uint64_t db = ptx::umma_desc_sw128(sb + ks * b_kadv, b_lbo, 1024u);
              if (live && !SDB_DBG(p, 4)) ptx::umma_bf16(tmem_d, da, db, idesc, (it > 0 || ks > 0) ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
            if (it == ti.num_k - 1) ptx::umma_commit(&tmem_full[acc]);  // accumulator complete -> epilogue
          }
          __syncwarp();
          SDB_GTL(1 + half, gk, 1);
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int eset = (kEpiSets == 2 && warp >= 7) ? 1 : 0;
    const int c_first = (kEpiSets == 2 && MT == 1) ? eset * 32 : 0, c_step = (kEpiSets == 2 && MT == 1) ? 64 : 32;
    const int half_lo = (kEpiSets == 2 && MT == 2) ? eset : 0, half_hi = (kEpiSets == 2 && MT == 2) ? eset + 1 : MT;
    int tile_count = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++tile_count) {
      const Tile ti = decode_tile(p, t);
     for (int half = half_lo; half < half_hi; ++half) {
      const int row = half * 128 + q * 32 + lane;
      const int acc = (MT == 1) ? (tile_count & 1) : half;
      const uint32_t acc_phase = (MT == 1) ? ((uint32_t)(tile_count >> 1) & 1u) : ((uint32_t)tile_count & 1u);
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * kAccCols;

      if (MODE == SDB_GEMM_FWD) {
        const int wi = row % p.bw;
        const int r2 = row / p.bw;
        const int hi = r2 % p.bh;
        const int ni = r2 / p.bh;
        const int w = ti.w0 + wi, h = ti.h0 + hi, n = ti.n0 + ni;
        const bool valid = (w < p.W) && (h < p.H) && (n < p.NI) && !SDB_DBG(p, 2);
        const long long grow = ((long long)n * p.H + h) * p.W + w;
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.d) + (long long)n * p.dsN + (long long)h * p.dsH +
                             (long long)w * p.dsW + ti.n_off;
        const __nv_bfloat16* res = p.residual ? p.residual + grow * p.ldr + ti.n_off : nullptr;
        const float* bias = p.bias ? p.bias + ti.n_off : nullptr;
        const __nv_bfloat16* temb = p.temb ? p.temb + (long long)n * p.ldt + ti.n_off : nullptr;
        if (p.splits > 1) {
          // split-K: fp32 partial sums; sdb_splitk_finish adds the rest. Slab mode (default): every K range stores
          // into its own [pixels, N] slab and the finish pass folds the slabs in a fixed order — deterministic, no
          // memset, no atomics. red mode: red.global.add into one zeroed buffer (order depends on the schedule).
          float* acc_out = reinterpret_cast<float*>(p.d) + (long long)ti.split * p.slab + grow * p.ldd + ti.n_off;
          for (int c0 = c_first; c0 < ti.nw; c0 += c_step) {
            uint32_t v[32];
            ptx::tmem_ld_32x32(taddr + (uint32_t)c0, v);
            ptx::tmem_ld_wait();
            if (valid) {
#pragma unroll
              for (int j4 = 0; j4 < 32; j4 += 4) {
                if (c0 + j4 < ti.nw) {
                  if (p.slab > 0)
                    sdb::st_f4(acc_out + c0 + j4, make_float4(__uint_as_float(v[j4]), __uint_as_float(v[j4 + 1]),
                                                              __uint_as_float(v[j4 + 2]), __uint_as_float(v[j4 + 3])));
                  else
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(acc_out + c0 + j4),
                                 "f"(__uint_as_float(v[j4])), "f"(__uint_as_float(v[j4 + 1])),
                                 "f"(__uint_as_float(v[j4 + 2])), "f"(__uint_as_float(v[j4 + 3]))
                                 : "memory");
                }
              }
            }
          }
        } else
        for (int c0 = c_first; c0 < ti.nw; c0 += c_step) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(taddr + (uint32_t)c0, v);
          // every addend of the chunk is requested BEFORE waiting for the accumulator: the loads may not be hoisted over
          // the previous chunk's stores (possible aliasing), so issued where they are used each 8-column group exposed
          // a full load latency (40 per tile half: ~20 us of a 78 us convolution, SDB_GEMM_DBG experiments r2g)
          float4 bv[8];
          uint4 tv[4], rv[4];
          if (valid) {
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              if (c0 + g8 * 8 < ti.nw) {
                if (bias) {
                  bv[2 * g8] = __ldg(reinterpret_cast<const float4*>(bias + c0 + g8 * 8));
                  bv[2 * g8 + 1] = __ldg(reinterpret_cast<const float4*>(bias + c0 + g8 * 8 + 4));
                }
                if (temb) tv[g8] = sdb::ld_u4(temb + c0 + g8 * 8);
                if (res) rv[g8] = sdb::ld_u4(res + c0 + g8 * 8);
              }
            }
          }
          ptx::tmem_ld_wait();
          if (valid) {
            uint4 prev;
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              const int j8 = g8 * 8;
              if (c0 + j8 < ti.nw) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[j8 + j]);
                if (bias) {
                  const float4 b0 = bv[2 * g8], b1 = bv[2 * g8 + 1];
                  f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
                  f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
                }
                if (temb) {
                  const uint4 t4 = tv[g8];
                  f[0] += sdb::bf16_lo(t4.x); f[1] += sdb::bf16_hi(t4.x);
                  f[2] += sdb::bf16_lo(t4.y); f[3] += sdb::bf16_hi(t4.y);
                  f[4] += sdb::bf16_lo(t4.z); f[5] += sdb::bf16_hi(t4.z);
                  f[6] += sdb::bf16_lo(t4.w); f[7] += sdb::bf16_hi(t4.w);
                }
                if (res) {
                  const uint4 r4 = rv[g8];
                  f[0] += sdb::bf16_lo(r4.x); f[1] += sdb::bf16_hi(r4.x);
                  f[2] += sdb::bf16_lo(r4.y); f[3] += sdb::bf16_hi(r4.y);
                  f[4] += sdb::bf16_lo(r4.z); f[5] += sdb::bf16_hi(r4.z);
                  f[6] += sdb::bf16_lo(r4.w); f[7] += sdb::bf16_hi(r4.w);
                }
                uint4 o;
                o.x = sdb::pack_bf16x2(f[0], f[1]);
                o.y = sdb::pack_bf16x2(f[2], f[3]);
                o.z = sdb::pack_bf16x2(f[4], f[5]);
                o.w = sdb::pack_bf16x2(f[6], f[7]);
                // pairs of 8-column groups go out as one 32-byte store (a last unpaired group on its own)
                if (!p.st32 || ((g8 & 1) == 0 && c0 + j8 + 8 >= ti.nw)) sdb::st_u4(out + c0 + j8, o);
                else if (g8 & 1) sdb::st_u8(out + c0 + j8 - 8, prev, o);
                else prev = o;
              }
            }
          }
        }
      } else {
        const int m = ti.m_off + row;
        const bool valid = (m < p.M) && !SDB_DBG(p, 2);
        // split over pixel ranges: red.global.add into the zeroed dW (default), or — slab mode — plain stores into this
        // range's own dW-shaped slab, folded afterwards in a fixed order (sdb_splitk_finish_f32: run-to-run identical bits)
        float* out = reinterpret_cast<float*>(p.d) + (long long)ti.split * p.slab +
                     ((long long)p.tap_brow[ti.tap] + m) * p.ldd + ti.n_off;
        for (int c0 = c_first; c0 < ti.nw; c0 += c_step) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(taddr + (uint32_t)c0, v);
          ptx::tmem_ld_wait();
          if (valid) {
#pragma unroll
            for (int j4 = 0; j4 < 32; j4 += 4) {
              if (c0 + j4 < ti.nw) {
                float* dst = out + c0 + j4;
                if (p.splits > 1 && p.slab == 0) {
                  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst),
                               "f"(__uint_as_float(v[j4])), "f"(__uint_as_float(v[j4 + 1])),
                               "f"(__uint_as_float(v[j4 + 2])), "f"(__uint_as_float(v[j4 + 3]))
                               : "memory");
                } else if (!p.st32 || ((j4 & 4) == 0 && c0 + j4 + 4 >= ti.nw)) {
                  sdb::st_f4(dst, make_float4(__uint_as_float(v[j4]), __uint_as_float(v[j4 + 1]),
                                              __uint_as_float(v[j4 + 2]), __uint_as_float(v[j4 + 3])));
                } else if (j4 & 4) {
                  sdb::st_u8(dst - 4, make_uint4(v[j4 - 4], v[j4 - 3], v[j4 - 2], v[j4 - 1]),
                             make_uint4(v[j4], v[j4 + 1], v[j4 + 2], v[j4 + 3]));
                }
              }
            }
          }
        }
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(&tmem_empty[acc]);
     }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------- host side ----------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

int encode_view(CUtensorMap* map, const sdb_view4& v, int b0, int b1, int b2, int b3, const char* what) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_tc: cuTensorMapEncodeTiled entry point unavailable");
  if (v.ptr == nullptr || (((uintptr_t)v.ptr) & 15) != 0)
    return sdb_set_error(SDB_ERR_INVALID, "sdb_gemm_tc: %s pointer must be non-null and 16-byte aligned", what);
  if (v.C <= 0 || v.W <= 0 || v.H <= 0 || v.N <= 0)
    return sdb_set_error(SDB_ERR_INVALID, "sdb_gemm_tc: %s has a non-positive extent", what);
  if ((v.strideW % 8) || (v.strideH % 8) || (v.strideN % 8))
    return sdb_set_error(SDB_ERR_INVALID, "sdb_gemm_tc: %s strides must be multiples of 8 elements", what);
  cuuint64_t gdim[4] = {(cuuint64_t)v.C, (cuuint64_t)v.W, (cuuint64_t)v.H, (cuuint64_t)v.N};
  cuuint64_t gstr[3] = {(cuuint64_t)v.strideW * 2, (cuuint64_t)v.strideH * 2, (cuuint64_t)v.strideN * 2};
  cuuint32_t box[4] = {(cuuint32_t)b0, (cuuint32_t)b1, (cuuint32_t)b2, (cuuint32_t)b3};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(v.ptr), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return sdb_set_error(SDB_ERR_CUDA,
                         "sdb_gemm_tc: cuTensorMapEncodeTiled(%s) failed with %d (dims %d,%d,%d,%d strides %lld,%lld,%lld "
                         "box %d,%d,%d,%d)",
                         what, (int)r, v.C, v.W, v.H, v.N, v.strideW, v.strideH, v.strideN, b0, b1, b2, b3);
  return SDB_OK;
}

}  // namespace

extern "C" int sdb_gemm_tc(const sdb_gemm_desc* d, cudaStream_t stream) {
  SDB_CHECK_ARG(d != nullptr, "sdb_gemm_tc: null descriptor");
  SDB_CHECK_ARG(d->mode == SDB_GEMM_FWD || d->mode == SDB_GEMM_WGRAD, "sdb_gemm_tc: bad mode %d", d->mode);
  SDB_CHECK_ARG(d->num_src >= 1 && d->num_src <= SDB_MAX_SRC, "sdb_gemm_tc: num_src %d out of range", d->num_src);
  SDB_CHECK_ARG(d->ntaps >= 1 && d->ntaps <= SDB_MAX_TAPS, "sdb_gemm_tc: ntaps %d out of range", d->ntaps);
  SDB_CHECK_ARG(d->outW > 0 && d->outH > 0 && d->outN > 0, "sdb_gemm_tc: bad pixel space");
  const int box_pixels = d->bw * d->bh * d->bn;
  const bool wgrad = d->mode == SDB_GEMM_WGRAD;
  const int mt = d->m_halves == 2 ? 2 : 1;
  SDB_CHECK_ARG(d->m_halves >= 0 && d->m_halves <= 2, "sdb_gemm_tc: m_halves must be 0, 1 or 2");
  SDB_CHECK_ARG(d->bw > 0 && d->bh > 0 && d->bn > 0 && box_pixels == (wgrad ? 64 : 128 * mt) && d->bw <= 256 &&
                    d->bh <= 256 && d->bn <= 256,
                "sdb_gemm_tc: pixel box %dx%dx%d must hold exactly %d pixels", d->bw, d->bh, d->bn,
                wgrad ? 64 : 128 * mt);
  SDB_CHECK_ARG(d->N > 0 && d->block_n >= 16 && d->block_n <= 256 && d->block_n % 16 == 0,
                "sdb_gemm_tc: block_n %d must be a multiple of 16 in [16,256]", d->block_n);
  SDB_CHECK_ARG(d->d != nullptr, "sdb_gemm_tc: null output");
  for (int t = 0; t < d->ntaps; ++t)
    SDB_CHECK_ARG(d->tap_src[t] >= 0 && d->tap_src[t] < d->num_src, "sdb_gemm_tc: tap %d names source %d", t,
                  d->tap_src[t]);

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.mode = d->mode;
  p.mt = mt;
  p.a_stage_bytes = (uint32_t)mt * kAStageBytes;
  p.ntaps = d->ntaps;
  for (int t = 0; t < d->ntaps; ++t) {
    p.tap_src[t] = d->tap_src[t];
    p.tap_dh[t] = d->tap_dh[t];
    p.tap_dw[t] = d->tap_dw[t];
    p.tap_brow[t] = d->tap_brow[t];
  }
  p.W = d->outW; p.H = d->outH; p.NI = d->outN;
  p.bw = d->bw; p.bh = d->bh; p.bn = d->bn;
  p.tw = sdb::ceil_div(p.W, p.bw); p.th = sdb::ceil_div(p.H, p.bh); p.tn = sdb::ceil_div(p.NI, p.bn);
  p.N = d->N; p.BN = d->block_n;
  p.tiles_n = sdb::ceil_div(p.N, p.BN);
  p.d = d->d; p.ldd = d->ldd;
  if (d->d_strideW || d->d_strideH || d->d_strideN) {
    p.dsW = d->d_strideW; p.dsH = d->d_strideH; p.dsN = d->d_strideN;
  } else {
    p.dsW = d->ldd; p.dsH = d->ldd * d->outW; p.dsN = d->ldd * d->outW * d->outH;
  }
  {  // tile column offsets and chunk starts are multiples of 16 elements; the base and the strides must be too
    const int esz = (d->mode == SDB_GEMM_FWD) ? 2 : 4;
    const long long unit = 32 / esz;
    p.st32 = ((reinterpret_cast<uintptr_t>(d->d) & 31) == 0) && (p.dsW % unit == 0) && (p.dsH % unit == 0) &&
             (p.dsN % unit == 0) && (p.ldd % unit == 0) && (p.BN % 16 == 0);
  }
  p.bias = d->bias;
  p.temb = reinterpret_cast<const __nv_bfloat16*>(d->temb); p.ldt = d->ldt;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(d->residual); p.ldr = d->ldr;

  int rc;
  for (int i = 0; i < d->num_src; ++i) {
    if (wgrad) rc = encode_view(&p.mapA[i], d->a[i], 64, d->bw, d->bh, d->bn, "A (wgrad activations)");
    else rc = encode_view(&p.mapA[i], d->a[i], 64, d->bw, d->bh, d->bn, "A (activations)");
    if (rc) return rc;
  }
  for (int i = d->num_src; i < SDB_MAX_SRC; ++i) p.mapA[i] = p.mapA[0];

  if (!wgrad) {
    SDB_CHECK_ARG(d->K > 0, "sdb_gemm_tc: K must be positive");
    SDB_CHECK_ARG(p.dsW % 8 == 0 && p.dsH % 8 == 0 && p.dsN % 8 == 0, "sdb_gemm_tc: output strides must be multiples of 8");
    SDB_CHECK_ARG(d->N % 8 == 0 && d->ldd % 8 == 0 && (((uintptr_t)d->d) & 15) == 0,
                  "sdb_gemm_tc: N (%d) and ldd (%lld) must be multiples of 8 and D 16-byte aligned", d->N, d->ldd);
    SDB_CHECK_ARG(d->residual == nullptr || (d->ldr % 8 == 0 && (((uintptr_t)d->residual) & 15) == 0),
                  "sdb_gemm_tc: residual needs ldr %% 8 == 0 and 16-byte alignment");
    SDB_CHECK_ARG((((uintptr_t)d->bias) & 15) == 0 && (((uintptr_t)d->temb) & 15) == 0 && d->ldt % 8 == 0,
                  "sdb_gemm_tc: bias/temb need 16-byte alignment (ldt %% 8 == 0)");
    p.b_mn = d->b_mn_major ? 1 : 0;
    p.K = d->K;
    p.kblocks = sdb::ceil_div(p.K, kBlockK);
    p.tiles_m = p.tw * p.th * p.tn;
    p.total_pb = p.ntaps * p.kblocks;  // K blocks of one output tile
    int splits = d->splits < 1 ? 1 : d->splits;
    if (splits > p.total_pb) splits = p.total_pb;
    p.pb_per_split = sdb::ceil_div(p.total_pb, splits);
    p.splits = sdb::ceil_div(p.total_pb, p.pb_per_split);
    p.total_tiles = p.tiles_m * p.tiles_n * p.splits;
    if (p.splits > 1) {
      SDB_CHECK_ARG(d->bias == nullptr && d->temb == nullptr && d->residual == nullptr && d->d_strideW == 0 &&
                        d->d_strideH == 0 && d->d_strideN == 0 && d->N % 4 == 0 && d->ldd % 4 == 0,
                    "sdb_gemm_tc: split-K forward accumulates plain fp32 partial sums (no epilogue terms, dense output)");
      if (d->splitk_slab > 0) {
        SDB_CHECK_ARG(d->splitk_slab >= (long long)p.W * p.H * p.NI * d->ldd && d->splitk_slab % 4 == 0,
                      "sdb_gemm_tc: splitk_slab (%lld) must cover pixels x ldd and be a multiple of 4", d->splitk_slab);
        p.slab = d->splitk_slab;
      } else {
        const cudaError_t me = cudaMemsetAsync(d->d, 0, (size_t)p.W * p.H * p.NI * (size_t)d->ldd * sizeof(float), stream);
        if (me != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_tc: memset: %s", cudaGetErrorString(me));
      }
    }
    if (p.b_mn) {
      rc = encode_view(&p.mapB, d->b, 64, 64, 1, 1, "B (weights [K,N])");
      p.b_stage_bytes = (uint32_t)(((p.BN + 63) / 64) * kBoxBytes);
    } else {
      rc = encode_view(&p.mapB, d->b, 64, p.BN, 1, 1, "B (weights [N,K])");
      p.b_stage_bytes = (uint32_t)p.BN * 128u;
    }
    if (rc) return rc;
  } else {
    SDB_CHECK_ARG(d->M > 0, "sdb_gemm_tc: WGRAD needs M (rows of dW per tap)");
    SDB_CHECK_ARG(d->N % 4 == 0 && d->ldd % 4 == 0 && (((uintptr_t)d->d) & 15) == 0,
                  "sdb_gemm_tc: WGRAD needs N %% 4 == 0, ldd %% 4 == 0 and a 16-byte aligned output");
    p.b_mn = 1;
    p.M = d->M;
    p.tiles_m = sdb::ceil_div(p.M, 128 * mt);
    p.total_pb = p.tw * p.th * p.tn;
    int splits = d->splits < 1 ? 1 : d->splits;
    if (splits > p.total_pb) splits = p.total_pb;
    p.pb_per_split = sdb::ceil_div(p.total_pb, splits);
    p.splits = sdb::ceil_div(p.total_pb, p.pb_per_split);
    p.total_tiles = p.ntaps * p.tiles_m * p.tiles_n * p.splits;
    rc = encode_view(&p.mapB, d->b, 64, d->bw, d->bh, d->bn, "B (wgrad output gradients)");
    if (rc) return rc;
    int last_row = 0;
    for (int t = 0; t < d->ntaps; ++t) last_row = d->tap_brow[t] > last_row ? d->tap_brow[t] : last_row;
    if (p.splits > 1 && d->splitk_slab > 0) {  // deterministic mode: one dW-shaped slab per pixel range, plain stores
      SDB_CHECK_ARG(d->splitk_slab >= ((long long)last_row + d->M) * d->ldd && d->splitk_slab % 4 == 0,
                    "sdb_gemm_tc: splitk_slab (%lld) must cover the dW rows x ldd and be a multiple of 4", d->splitk_slab);
      p.slab = d->splitk_slab;
    } else if (p.splits > 1) {  // the pixel ranges are accumulated with red.global.add: start from zero (a memset node, no SMs)
      const cudaError_t me = cudaMemset2DAsync(d->d, (size_t)d->ldd * sizeof(float), 0, (size_t)d->N * sizeof(float),
                                               (size_t)last_row + (size_t)d->M, stream);
      if (me != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_tc: memset: %s", cudaGetErrorString(me));
    }
    p.b_stage_bytes = (uint32_t)(((p.BN + 63) / 64) * kBoxBytes);
  }
  p.b_stage_bytes = (p.b_stage_bytes + 1023u) & ~1023u;
  p.stage_bytes = p.a_stage_bytes + p.b_stage_bytes;
  const uint32_t budget = 225 * 1024;  // 227 KB per CTA minus the alignment slack and the barriers
  int stages = (int)(budget / p.stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_gemm_tc: stage too large");
  p.stages = stages;
  static const int dbg_env = getenv("SDB_GEMM_DBG") ? atoi(getenv("SDB_GEMM_DBG")) : 0;  // timing experiments only
  p.dbg = dbg_env;
  const size_t smem_bytes = (size_t)stages * p.stage_bytes + 1024 /*align slack*/ + 256 /*barriers*/;

  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    const int bytes = 227 * 1024;
    attr_err = cudaFuncSetAttribute(gemm_tc_kernel<SDB_GEMM_FWD, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(gemm_tc_kernel<SDB_GEMM_FWD, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(gemm_tc_kernel<SDB_GEMM_WGRAD, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(gemm_tc_kernel<SDB_GEMM_WGRAD, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  });
  if (attr_err != cudaSuccess)
    return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));

  const int grid = p.total_tiles < sdb::kNumSMs ? p.total_tiles : sdb::kNumSMs;
  cudaError_t lerr;
  const dim3 g3(grid), b3(kThreads);
  if (wgrad) {
    if (mt == 2) lerr = sdb::launch_pdl(gemm_tc_kernel<SDB_GEMM_WGRAD, 2>, g3, b3, smem_bytes, stream, p);
    else lerr = sdb::launch_pdl(gemm_tc_kernel<SDB_GEMM_WGRAD, 1>, g3, b3, smem_bytes, stream, p);
  } else {
    if (mt == 2) lerr = sdb::launch_pdl(gemm_tc_kernel<SDB_GEMM_FWD, 2>, g3, b3, smem_bytes, stream, p);
    else lerr = sdb::launch_pdl(gemm_tc_kernel<SDB_GEMM_FWD, 1>, g3, b3, smem_bytes, stream, p);
  }
  if (lerr != cudaSuccess) {
    cudaGetLastError();
    return sdb_set_error(SDB_ERR_CUDA, "sdb_gemm_tc: launch failed: %s", cudaGetErrorString(lerr));
  }
  SDB_CUDA_LAUNCH_CHECK("sdb_gemm_tc");
  return SDB_OK;
}
