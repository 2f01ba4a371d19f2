// Flash-style attention for sm_100a: online softmax with tcgen05.mma (S = QK^T and O += PV accumulate in TMEM),
// operands staged by TMA into 128B-swizzled shared memory. Forward + backward.
//
// Replaces diffusers-0.21.4 FlaxAttention with use_memory_efficient_attention=True — jax_memory_efficient_attention /
// _query_chunk_attention [upstream attention_flax.py, un-vendored] as patched by /root/reference/key_chunk_patch.patch:5-6
// (key_chunk_size = flatten_latent_dim); call site /root/reference/training_utils.py:678-684. The reference's
// query-chunk scan x key-chunk map is replaced by one pass over 128-key blocks. The patch's side effect (keys in the
// clamped last chunk are counted twice) is reproduced exactly by an additive per-key bias log2(m_k)
// (`kbias2`, log2 domain); the same array carries -inf for masked keys.
//
// Layout: q/k/v/o/dO are [B, S, H*d] row-major bf16 with independent row strides (heads interleaved in columns), so the
// projections' outputs are consumed in place: a 4-D tensor map (d, S, H, B) picks one head; columns d..63 of the last
// 64-wide swizzle atom are zero-filled by TMA, which pads d = 40 / 80 / 160 for the K = 16 UMMA steps.
//
// Forward (192 threads): warps 0-3 = softmax (one thread per query row: S row from TMEM with tcgen05.ld, no shuffles),
// warp 4 = tcgen05.mma issuer + TMEM owner, warp 5 = TMA producer. O stays in TMEM across key blocks and is rescaled
// lazily (only when the running max grows by > 2^8, FA4-style), P goes through shared memory as the A operand of PV.
//
// Backward: one kernel in two modes, no atomics, deterministic.
//   mode 0: a CTA owns 128 KEY rows and streams query blocks:  S^T = K Q^T, dP^T = V dO^T, P^T = exp2(...),
//           dS^T = P^T o (dP^T - D) * scale, dV += P^T dO, dK += dS^T Q.
//   mode 1: the same code with the roles swapped (a CTA owns 128 QUERY rows, streams key blocks): dQ += dS K.
#include "sdb_common.cuh"
#include "ptx_sm100.cuh"
#include "../../include/sdb200.h"
#include <cmath>
#include <mutex>

// -DSDB_ATTN_TIMELINE: CTA (0,0,0) of the backward kernel records clock64() at its hand-off points (tools/attn_timeline.py)
#ifdef SDB_ATTN_TIMELINE
__device__ long long g_attn_tl[8 * 64 * 4];
#define SDB_TL(role, i, slot)                                                                   \
  do {                                                                                          \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (i) >= 8 && (i) < 72 && (threadIdx.x & 31) == 0) \
      g_attn_tl[((role) * 64 + ((i) - 8)) * 4 + (slot)] = clock64();                            \
  } while (0)
extern "C" int sdb_attn_timeline(long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_attn_tl, sizeof(long long) * 8 * 64 * 4);
}
#else
#define SDB_TL(role, i, slot) do {} while (0)
#endif

// waits of the producer / issuer warps: -DSDB_ISSUER_SPIN makes them pure try_wait loops (no nanosleep back-off)
#ifdef SDB_ISSUER_SPIN
#define SDB_ISSUER_WAIT(bar, ph) ptx::mbar_wait(bar, ph)
#else
#define SDB_ISSUER_WAIT(bar, ph) ptx::mbar_wait_sleep(bar, ph)
#endif

namespace {

constexpr uint32_t kAtom128 = 128 * 128;  // bytes of a [128 rows x 64 bf16] swizzle atom column

__device__ __forceinline__ float u2f(uint32_t v) { return __uint_as_float(v); }

// 16-byte store of 8 bf16 into a K-major SWIZZLE_128B tile: row r, 16B chunk index c (0..7) of the 128B row
__device__ __forceinline__ void st_swz(uint8_t* atom_base, int r, int c, uint4 v) {
  *reinterpret_cast<uint4*>(atom_base + r * 128 + ((c ^ (r & 7)) << 4)) = v;
}

__device__ __forceinline__ void tma_tile(uint8_t* dst, const CUtensorMap* m, uint64_t* bar, int atoms,
                                         uint32_t atom_bytes, int row0, int h, int b) {
  for (int a = 0; a < atoms; ++a) ptx::tma_load_4d(dst + a * atom_bytes, m, bar, 64 * a, row0, h, b);
}

// Issue `ksteps` K = 16 MMAs from precomputed base descriptors. The smem address field is the low 14 bits of the
// descriptor (16-byte units), so stepping along K is an integer add: +2 per step inside a 64-column swizzle atom and a
// jump to the next atom every 4 steps (K-major), or +128 per 16 K-rows (MN-major). Keeping this loop free of divisions
// and descriptor re-encoding matters: ONE thread issues every MMA of the CTA, and its latency is on the critical path.
template <bool B_MN>
__device__ __forceinline__ void mma_ksteps(uint32_t tmem_d, uint64_t dA, uint32_t a_atom16, uint64_t dB, uint32_t b_atom16,
                                           uint32_t idesc, int ksteps, uint32_t acc_first) {
  uint32_t offA = 0, offB = 0;
  for (int k = 0; k < ksteps; ++k) {
    ptx::umma_bf16(tmem_d, dA + offA, dB + offB, idesc, k > 0 ? 1u : acc_first);
    const bool wrap = (k & 3) == 3;
    offA += wrap ? a_atom16 - 6u : 2u;
    offB += B_MN ? 128u : (wrap ? b_atom16 - 6u : 2u);
  }
}

// =============================================== forward ===============================================
// 384 threads: warps 0-7 = softmax (two warps per TMEM lane quarter: warp w and w+4 own the same 32 query rows and
// split the 128 keys of a block in halves), warp 8 = QK^T issuer + TMEM owner, warp 9 = TMA producer, warps 10-11 = PV
// issuers. S is double-buffered in TMEM and P in shared memory, so QK^T of block j+1 and PV of block j-1 run on the
// tensor pipe while the softmax warps (the MUFU-bound resource: 128x128 exp2 per block) work on block j.
//
// Several issuing warps: each waits on its own barriers (S needs the K tile and a free S buffer, PV needs P and the V
// tile), so QK^T gets its own warp and PV is split by key halves (k-steps 0-3 / 4-7) into two accumulators that the
// epilogue adds. The issue loops run on the whole warp with the MMAs under elect.sync (ptx::elect_one): written
// under `if (lane == 0)` every MMA cost ~120 clk of ELECT / R2UR serialisation (tools/mma_rate.cu).
constexpr int kFwdThreads = 384;
constexpr int kFwdMaxStages = 4;  // K and V rings: deep enough to cover the TMA latency under load
constexpr int kMaxBiasBlocks = 16;  // per-key bias (cross-attention quirk / masks given by the caller): Skv <= 2048

struct FwdParams {
  CUtensorMap mapQ, mapK, mapV;
  __nv_bfloat16* o;
  long long ldo;
  float* lse2;
  const float* kbias2;
  int B, H, Sq, Skv, d, dpad, atoms, kstages, vstages, nkb, npv;
  float scale2;
};

__global__ void __launch_bounds__(kFwdThreads, 1) attn_fwd_kernel(const __grid_constant__ FwdParams p) {
  sdb::pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const uint32_t tile_bytes = (uint32_t)p.atoms * kAtom128;
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + tile_bytes;
  uint8_t* sV = sK + (size_t)p.kstages * tile_bytes;
  uint8_t* sP = sV + (size_t)p.vstages * tile_bytes;             // [2][2 atoms]
  float* sX = reinterpret_cast<float*>(sP + 4 * kAtom128);       // [2 buf][2 half][128] row-max exchange
  float* sL = sX + 512;                                          // [2 half][128] row-sum exchange
  float* sKb = sL + 256;                                         // [nkb*128] per-key bias (only when kbias2 != null)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKb + (p.kbias2 ? p.nkb * 128 : 0));
  uint64_t* barQ = bars;
  uint64_t* k_full = bars + 1;                     // [kFwdMaxStages]
  uint64_t* k_empty = k_full + kFwdMaxStages;
  uint64_t* v_full = k_empty + kFwdMaxStages;
  uint64_t* v_empty = v_full + kFwdMaxStages;
  uint64_t* bar_S = v_empty + kFwdMaxStages;       // [2]
  uint64_t* bar_P = bar_S + 2;                     // [2]
  uint64_t* bar_PV = bar_P + 2;                    // [2]: PV(j) commits to bar_PV[j & 1]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_PV + 2);

  const int warp = ptx::warp_idx_uniform(), lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;

  if (warp == 9 && lane == 0) {
    ptx::tma_prefetch_desc(&p.mapQ);
    ptx::tma_prefetch_desc(&p.mapK);
    ptx::tma_prefetch_desc(&p.mapV);
    ptx::mbar_init(barQ, 1);
    for (int s = 0; s < kFwdMaxStages; ++s) {
      ptx::mbar_init(&k_full[s], 1);
      ptx::mbar_init(&k_empty[s], 1);
      ptx::mbar_init(&v_full[s], 1);
      ptx::mbar_init(&v_empty[s], (uint32_t)p.npv);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&bar_S[s], 1);
      ptx::mbar_init(&bar_P[s], 256);
    }
    ptx::mbar_init(&bar_PV[0], (uint32_t)p.npv);
    ptx::mbar_init(&bar_PV[1], (uint32_t)p.npv);
    ptx::fence_barrier_init();
  }
  if (warp == 8) {
    ptx::tmem_alloc(tmem_ptr_smem, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  sdb::pdl_wait();  // prologue above overlaps the previous kernel's tail
  const uint32_t tS = tmem_base;         // 2 x 128 columns
  const uint32_t tO = tmem_base + 256u;  // npv x dpad columns (one accumulator per PV issuer)

  if (warp == 9) {
    // ===================== TMA producer (whole warp; the TMA instructions under elect) =====================
    {
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(barQ, tile_bytes);
        tma_tile(sQ, &p.mapQ, barQ, p.atoms, kAtom128, q0, h, b);
      }
      __syncwarp();
      auto load_k = [&](int j) {
        const int s = j % p.kstages, n = j / p.kstages;
        SDB_ISSUER_WAIT(&k_empty[s], (uint32_t)((n & 1) ^ 1));
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&k_full[s], tile_bytes);
          tma_tile(sK + (size_t)s * tile_bytes, &p.mapK, &k_full[s], p.atoms, kAtom128, j * 128, h, b);
        }
        __syncwarp();
      };
      load_k(0);
      for (int j = 0; j < p.nkb; ++j) {
        if (j + 1 < p.nkb) load_k(j + 1);  // K runs one block ahead of V: S(j+1) is issued before PV(j)
        const int s = j % p.vstages, n = j / p.vstages;
        SDB_ISSUER_WAIT(&v_empty[s], (uint32_t)((n & 1) ^ 1));
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&v_full[s], tile_bytes);
          tma_tile(sV + (size_t)s * tile_bytes, &p.mapV, &v_full[s], p.atoms, kAtom128, j * 128, h, b);
        }
        __syncwarp();
      }
    }
  } else if (warp == 8) {
    // ===================== QK^T issuer (whole warp, MMAs under elect: see ptx::elect_one) =====================
    {
      const uint32_t idescS = ptx::umma_idesc_bf16(128, 0, 0);
      const uint64_t dQ = ptx::umma_desc_sw128(ptx::smem_u32(sQ), 0u, 1024u);
      const uint64_t dK0 = ptx::umma_desc_sw128(ptx::smem_u32(sK), 0u, 1024u);
      const uint32_t tile16 = tile_bytes >> 4, atom16 = kAtom128 >> 4;
      const int ksteps = p.dpad / 16;
      int ks = 0;
      uint32_t kph = 0;
      ptx::mbar_wait(barQ, 0);
      for (int j = 0; j < p.nkb; ++j) {
        // S buffer j&1 was last read by softmax(j-2): its arrival on bar_P comes after its TMEM loads
        if (j >= 2) SDB_ISSUER_WAIT(&bar_P[j & 1], (uint32_t)(((j - 2) >> 1) & 1));
        SDB_ISSUER_WAIT(&k_full[ks], kph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          mma_ksteps<false>(tS + (uint32_t)(j & 1) * 128u, dQ, atom16, dK0 + (uint32_t)ks * tile16, atom16, idescS,
                            ksteps, 0u);
          ptx::umma_commit(&k_empty[ks]);
          ptx::umma_commit(&bar_S[j & 1]);
        }
        __syncwarp();
        if (++ks == p.kstages) { ks = 0; kph ^= 1u; }
      }
    }
  } else if (warp >= 10) {
    // ===================== PV issuers: issuer hv owns k-steps [hv*8/npv, (hv+1)*8/npv) and accumulator hv ============
    const int hv = warp - 10;
    if (hv < p.npv) {
      const uint32_t idescO = ptx::umma_idesc_bf16(p.dpad, 0, 1);
      const uint64_t dV0 = ptx::umma_desc_sw128(ptx::smem_u32(sV), kAtom128, 1024u);  // MN-major: LBO = atom stride
      const uint64_t dP0 = ptx::umma_desc_sw128(ptx::smem_u32(sP), 0u, 1024u);
      const uint32_t tile16 = tile_bytes >> 4, atom16 = kAtom128 >> 4;
      const int nsteps = 8 / p.npv;
      const uint32_t tOh = tO + (uint32_t)(hv * p.dpad);
      const uint32_t a_off = (p.npv == 2) ? (uint32_t)hv * atom16 : 0u;   // P atom = key half
      const uint32_t b_off = (p.npv == 2) ? (uint32_t)hv * 4u * 128u : 0u;  // 16 key rows = 2048 B of an MN-major V atom
      int vs = 0;
      uint32_t vph = 0;
      for (int j = 0; j < p.nkb; ++j) {
        SDB_ISSUER_WAIT(&bar_P[j & 1], (uint32_t)((j >> 1) & 1));
        SDB_ISSUER_WAIT(&v_full[vs], vph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          mma_ksteps<true>(tOh, dP0 + (uint32_t)(j & 1) * 2u * atom16 + a_off, atom16, dV0 + (uint32_t)vs * tile16 + b_off,
                           0u, idescO, nsteps, j > 0 ? 1u : 0u);
          ptx::umma_commit(&v_empty[vs]);
          ptx::umma_commit(&bar_PV[j & 1]);
        }
        __syncwarp();
        if (++vs == p.vstages) { vs = 0; vph ^= 1u; }
      }
    }
  } else {
    // ===================== softmax / epilogue: thread = (query row, key half) =====================
    const int q4 = warp & 3, half = warp >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q4 * 32)) << 16;
    if (p.kbias2 != nullptr) {
      for (int i = threadIdx.x; i < p.nkb * 128; i += 256) sKb[i] = (i < p.Skv) ? p.kbias2[i] : -INFINITY;
      ptx::named_bar_sync(1, 256);
    }
    float m_used = -INFINITY, l = 0.f;
    for (int j = 0; j < p.nkb; ++j) {
      ptx::mbar_wait_sleep(&bar_S[j & 1], (uint32_t)((j >> 1) & 1), 20);
      ptx::tc_fence_after();
      const uint32_t tSj = tS + (uint32_t)(j & 1) * 128u + lane_base + (uint32_t)half * 64u;
      const int key0 = j * 128 + half * 64;
      // element modes: 0 plain, 1 tail mask computed inline, 2 per-key bias from shared memory
      const int emode = (p.kbias2 != nullptr) ? 2 : ((j * 128 + 128 > p.Skv) ? 1 : 0);
      const float* kb = sKb + key0;
      // pass 1: this thread's 64 scores stay in registers for pass 2; row max in the log2 domain with four
      // independent max chains (a single chain of 64 dependent FMNMX is pure latency at 2 warps per scheduler)
      uint32_t sv[64];
      ptx::tmem_ld_32x32(tSj, *reinterpret_cast<uint32_t(*)[32]>(&sv[0]));
      ptx::tmem_ld_32x32(tSj + 32, *reinterpret_cast<uint32_t(*)[32]>(&sv[32]));
      ptx::tmem_ld_wait();
      float mx;
      {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (emode == 0) {  // scale2 > 0: max of the raw scores, scaled once
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], u2f(sv[i]));
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale2;
        } else if (emode == 1) {
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], (key0 + i < p.Skv) ? u2f(sv[i]) : -INFINITY);
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale2;
        } else {
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], fmaf(u2f(sv[i]), p.scale2, kb[i]));
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        }
      }
      float* xbuf = sX + (j & 1) * 256;
      xbuf[half * 128 + row] = mx;
      ptx::named_bar_sync(1, 256);
      mx = fmaxf(mx, xbuf[(half ^ 1) * 128 + row]);
      if (j == 0) {
        m_used = mx;
      } else {
        // P buffer j&1 was last read by PV(j-2). PV(j-1) may still be running: it only has to be waited for in the
        // rare case that O must be rescaled (the lazy threshold), so the exp pass below overlaps it.
        if (j >= 2) ptx::mbar_wait_sleep(&bar_PV[j & 1], (uint32_t)(((j - 2) >> 1) & 1), 20);
        const bool need = mx > m_used + 8.0f;
        if (__any_sync(0xffffffffu, need)) {
          ptx::mbar_wait_sleep(&bar_PV[(j - 1) & 1], (uint32_t)(((j - 1) >> 1) & 1), 20);  // O holds blocks 0..j-1
          ptx::tc_fence_after();
          const float alpha = need ? ptx::ex2_approx(m_used - mx) : 1.0f;
          if (need) m_used = mx;
          l *= alpha;
          for (int c0 = half * 16; c0 < p.npv * p.dpad; c0 += 32) {  // the two warps of a row interleave 16-column chunks
            uint32_t v[16];
            ptx::tmem_ld_32x16(tO + lane_base + c0, v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(u2f(v[i]) * alpha);
            ptx::tmem_st_32x16(tO + lane_base + c0, v);
          }
          ptx::tmem_st_wait();
        }
      }
      // pass 2: P = exp2(s - m), bf16, into the swizzled A-operand tile (atom = key half); four partial row sums
      uint8_t* atom = sP + (size_t)(j & 1) * 2 * kAtom128 + (size_t)half * kAtom128;
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float pv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int i = c * 8 + e;
          if (emode == 0) pv[e] = ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, -m_used), i);
          else if (emode == 1) pv[e] = (key0 + i < p.Skv) ? ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, -m_used), i) : 0.f;
          else pv[e] = ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, kb[i] - m_used), i);
          l4[e & 3] += pv[e];
        }
        uint4 o;
        o.x = sdb::pack_bf16x2(pv[0], pv[1]);
        o.y = sdb::pack_bf16x2(pv[2], pv[3]);
        o.z = sdb::pack_bf16x2(pv[4], pv[5]);
        o.w = sdb::pack_bf16x2(pv[6], pv[7]);
        st_swz(atom, row, c, o);
      }
      l += (l4[0] + l4[1]) + (l4[2] + l4[3]);
      ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar_P[j & 1]);
    }
    // ---- epilogue: O / l -> global, lse ----
    sL[half * 128 + row] = l;
    ptx::named_bar_sync(1, 256);
    l += sL[(half ^ 1) * 128 + row];
    if (p.nkb >= 2) ptx::mbar_wait(&bar_PV[(p.nkb - 2) & 1], (uint32_t)(((p.nkb - 2) >> 1) & 1));
    ptx::mbar_wait(&bar_PV[(p.nkb - 1) & 1], (uint32_t)(((p.nkb - 1) >> 1) & 1));
    ptx::tc_fence_after();
    const int q = q0 + row;
    const bool valid = q < p.Sq;
    const float inv_l = 1.0f / l;
    __nv_bfloat16* orow = p.o + ((long long)b * p.Sq + q) * p.ldo + (long long)h * p.d;
    for (int c0 = half * 16; c0 < p.dpad; c0 += 32) {
      uint32_t v[16];
      ptx::tmem_ld_32x16(tO + lane_base + c0, v);
      if (p.npv == 2) {
        uint32_t v2[16];
        ptx::tmem_ld_32x16(tO + lane_base + (uint32_t)p.dpad + c0, v2);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(u2f(v[i]) + u2f(v2[i]));
      }
      ptx::tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c0 + c * 8 < p.d) {
            uint4 o;
            o.x = sdb::pack_bf16x2(u2f(v[c * 8 + 0]) * inv_l, u2f(v[c * 8 + 1]) * inv_l);
            o.y = sdb::pack_bf16x2(u2f(v[c * 8 + 2]) * inv_l, u2f(v[c * 8 + 3]) * inv_l);
            o.z = sdb::pack_bf16x2(u2f(v[c * 8 + 4]) * inv_l, u2f(v[c * 8 + 5]) * inv_l);
            o.w = sdb::pack_bf16x2(u2f(v[c * 8 + 6]) * inv_l, u2f(v[c * 8 + 7]) * inv_l);
            sdb::st_u4(orow + c0 + c * 8, o);
          }
        }
      }
    }
    if (valid && half == 0) p.lse2[((long long)b * p.H + h) * p.Sq + q] = m_used + log2f(l);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// =============================================== forward, two query tiles per CTA ===============================================
// For head dims <= 64 (the 64x64-latent layers of SD: d = 40 / 64). The one-tile kernel above keeps its 8 softmax warps in
// lockstep (a 256-thread barrier per key block for the row-max exchange), so with two warps per scheduler neither the MUFU
// pipe (40 %) nor the issue slots (39 %) are busy (profiles/r1n_ncu_full.md). Here a CTA owns 256 query rows as two
// independent 128-row tiles, each with its own 8 softmax warps, S / P / O columns in tensor memory and issuing warp; the
// two groups drift out of phase, so one group's exp2 pass overlaps the other's max pass, waits and MMAs, and every K / V
// tile is fetched once per 256 rows. P goes to the PV product through tensor memory (A operand from TMEM): no shared-
// memory staging, no proxy fence.
//   warps 0-7 softmax group 0, 8-15 group 1 (thread = (row, key half)), 16 TMA, 17 / 18 issuers (17 owns TMEM)
//   TMEM: [S0 128 | S1 128 | O0 dpad | O1 dpad | P0 64 | P1 64]  (P bf16: two keys per column)
constexpr int kFwd2Threads = 608;

__global__ void __launch_bounds__(kFwd2Threads, 1) attn_fwd2_kernel(const __grid_constant__ FwdParams p) {
  sdb::pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const uint32_t tile_bytes = (uint32_t)p.atoms * kAtom128;
  uint8_t* sQ = smem;                                            // [2 tiles]
  uint8_t* sK = sQ + 2 * tile_bytes;
  uint8_t* sV = sK + (size_t)p.kstages * tile_bytes;
  float* sX = reinterpret_cast<float*>(sV + (size_t)p.vstages * tile_bytes);  // [2 grp][2 buf][2 half][128] row max
  float* sL = sX + 1024;                                         // [2 grp][2 half][128] row-sum exchange
  float* sKb = sL + 512;                                         // [nkb*128] per-key bias (only when kbias2 != null)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKb + (p.kbias2 ? p.nkb * 128 : 0));
  uint64_t* barQ = bars;
  uint64_t* k_full = bars + 1;                     // [kFwdMaxStages]
  uint64_t* k_empty = k_full + kFwdMaxStages;      // count 2: both issuers
  uint64_t* v_full = k_empty + kFwdMaxStages;
  uint64_t* v_empty = v_full + kFwdMaxStages;
  uint64_t* bar_S = v_empty + kFwdMaxStages;       // [2 grp] QK^T(j) complete
  uint64_t* bar_Sfree = bar_S + 2;                 // [2 grp] the group holds S(j) in registers
  uint64_t* bar_P = bar_Sfree + 2;                 // [2 grp] P(j) written to tensor memory
  uint64_t* bar_PV = bar_P + 2;                    // [2 grp] PV(j) complete
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_PV + 2);

  const int warp = ptx::warp_idx_uniform(), lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 256, h = blockIdx.y, b = blockIdx.z;

  if (warp == 16 && lane == 0) {
    ptx::tma_prefetch_desc(&p.mapQ);
    ptx::tma_prefetch_desc(&p.mapK);
    ptx::tma_prefetch_desc(&p.mapV);
    ptx::mbar_init(barQ, 1);
    for (int s = 0; s < kFwdMaxStages; ++s) {
      ptx::mbar_init(&k_full[s], 1);
      ptx::mbar_init(&k_empty[s], 2);
      ptx::mbar_init(&v_full[s], 1);
      ptx::mbar_init(&v_empty[s], 2);
    }
    for (int g = 0; g < 2; ++g) {
      ptx::mbar_init(&bar_S[g], 1);
      ptx::mbar_init(&bar_Sfree[g], 256);
      ptx::mbar_init(&bar_P[g], 256);
      ptx::mbar_init(&bar_PV[g], 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 17) {
    ptx::tmem_alloc(tmem_ptr_smem, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  sdb::pdl_wait();
  const uint32_t tO0 = tmem_base + 256u;
  const uint32_t tP0 = tO0 + 2u * (uint32_t)p.dpad;

  if (warp == 16) {
    // ===================== TMA producer (whole warp; the TMA instructions under elect) =====================
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(barQ, 2 * tile_bytes);
      tma_tile(sQ, &p.mapQ, barQ, p.atoms, kAtom128, q0, h, b);
      tma_tile(sQ + tile_bytes, &p.mapQ, barQ, p.atoms, kAtom128, q0 + 128, h, b);
    }
    __syncwarp();
    auto load_k = [&](int j) {
      const int s = j % p.kstages, n = j / p.kstages;
      SDB_ISSUER_WAIT(&k_empty[s], (uint32_t)((n & 1) ^ 1));
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&k_full[s], tile_bytes);
        tma_tile(sK + (size_t)s * tile_bytes, &p.mapK, &k_full[s], p.atoms, kAtom128, j * 128, h, b);
      }
      __syncwarp();
    };
    load_k(0);
    for (int j = 0; j < p.nkb; ++j) {
      if (j + 1 < p.nkb) load_k(j + 1);
      const int s = j % p.vstages, n = j / p.vstages;
      SDB_ISSUER_WAIT(&v_empty[s], (uint32_t)((n & 1) ^ 1));
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&v_full[s], tile_bytes);
        tma_tile(sV + (size_t)s * tile_bytes, &p.mapV, &v_full[s], p.atoms, kAtom128, j * 128, h, b);
      }
      __syncwarp();
    }
  } else if (warp == 17 || warp == 18) {
    // ===================== issuer of group g: QK^T(0), then QK^T(j) followed by PV(j-1), finally PV(last) ==========
    const int g = warp - 17;
    const uint32_t idescS = ptx::umma_idesc_bf16(128, 0, 0);
    const uint32_t idescO = ptx::umma_idesc_bf16(p.dpad, 0, 1);
    const uint64_t dQ = ptx::umma_desc_sw128(ptx::smem_u32(sQ + (size_t)g * tile_bytes), 0u, 1024u);
    const uint64_t dK0 = ptx::umma_desc_sw128(ptx::smem_u32(sK), 0u, 1024u);
    const uint64_t dV0 = ptx::umma_desc_sw128(ptx::smem_u32(sV), kAtom128, 1024u);  // MN-major: LBO = atom stride
    const uint32_t tile16 = tile_bytes >> 4, atom16 = kAtom128 >> 4;
    const uint32_t tSg = tmem_base + (uint32_t)g * 128u;
    const uint32_t tOg = tO0 + (uint32_t)(g * p.dpad);
    const uint32_t tPg = tP0 + (uint32_t)g * 64u;
    const int ksteps = p.dpad / 16;
    int ks = 0, vs = 0;
    uint32_t kph = 0, vph = 0;
    ptx::mbar_wait(barQ, 0);
    for (int j = 0; j <= p.nkb; ++j) {
      if (j < p.nkb) {
        if (j >= 1) SDB_ISSUER_WAIT(&bar_Sfree[g], (uint32_t)((j - 1) & 1));
        SDB_ISSUER_WAIT(&k_full[ks], kph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          mma_ksteps<false>(tSg, dQ, atom16, dK0 + (uint32_t)ks * tile16, atom16, idescS, ksteps, 0u);
          ptx::umma_commit(&k_empty[ks]);
          ptx::umma_commit(&bar_S[g]);
        }
        __syncwarp();
        if (++ks == p.kstages) { ks = 0; kph ^= 1u; }
      }
      if (j >= 1) {
        SDB_ISSUER_WAIT(&bar_P[g], (uint32_t)((j - 1) & 1));
        SDB_ISSUER_WAIT(&v_full[vs], vph);
        ptx::tc_fence_after();
        const uint64_t dV = dV0 + (uint32_t)vs * tile16;
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            ptx::umma_bf16_ts(tOg, tPg + 8u * (uint32_t)k, dV + 128u * (uint32_t)k, idescO, (k > 0 || j > 1) ? 1u : 0u);
          ptx::umma_commit(&v_empty[vs]);
          ptx::umma_commit(&bar_PV[g]);
        }
        __syncwarp();
        if (++vs == p.vstages) { vs = 0; vph ^= 1u; }
      }
    }
  } else {
    // ===================== softmax / epilogue of group g: thread = (query row, key half) =====================
    const int g = warp >> 3, wg = warp & 7;
    const int q4 = wg & 3, half = wg >> 2;
    const int row = q4 * 32 + lane;
    const int gtid = (int)threadIdx.x & 255;
    const uint32_t lane_base = ((uint32_t)(q4 * 32)) << 16;
    const uint32_t tSg = tmem_base + (uint32_t)g * 128u + lane_base + (uint32_t)half * 64u;
    const uint32_t tOg = tO0 + (uint32_t)(g * p.dpad) + lane_base;
    const uint32_t tPg = tP0 + (uint32_t)g * 64u + lane_base + (uint32_t)half * 32u;
    float* sXg = sX + g * 512;
    if (p.kbias2 != nullptr) {
      for (int i = gtid; i < p.nkb * 128; i += 256) if (g == 0) sKb[i] = (i < p.Skv) ? p.kbias2[i] : -INFINITY;
      ptx::named_bar_sync(3, 512);
    }
    float m_used = -INFINITY, l = 0.f;
    for (int j = 0; j < p.nkb; ++j) {
      ptx::mbar_wait_sleep(&bar_S[g], (uint32_t)(j & 1), 20);
      ptx::tc_fence_after();
      const int key0 = j * 128 + half * 64;
      const int emode = (p.kbias2 != nullptr) ? 2 : ((j * 128 + 128 > p.Skv) ? 1 : 0);
      const float* kb = sKb + key0;
      uint32_t sv[64];
      ptx::tmem_ld_32x32(tSg, *reinterpret_cast<uint32_t(*)[32]>(&sv[0]));
      ptx::tmem_ld_32x32(tSg + 32, *reinterpret_cast<uint32_t(*)[32]>(&sv[32]));
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar_Sfree[g]);  // the scores are in registers: QK^T(j+1) may overwrite S
      float mx;
      {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (emode == 0) {
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], u2f(sv[i]));
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale2;
        } else if (emode == 1) {
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], (key0 + i < p.Skv) ? u2f(sv[i]) : -INFINITY);
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale2;
        } else {
#pragma unroll
          for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], fmaf(u2f(sv[i]), p.scale2, kb[i]));
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        }
      }
      float* xbuf = sXg + (j & 1) * 256;
      xbuf[half * 128 + row] = mx;
      ptx::named_bar_sync(1 + g, 256);
      mx = fmaxf(mx, xbuf[(half ^ 1) * 128 + row]);
      if (j == 0) {
        m_used = mx;
      } else {
        // PV(j-1) has read P(j-1) and updated O: needed before P is overwritten below and before O is rescaled
        ptx::mbar_wait_sleep(&bar_PV[g], (uint32_t)((j - 1) & 1), 20);
        ptx::tc_fence_after();
        const bool need = mx > m_used + 8.0f;
        if (__any_sync(0xffffffffu, need)) {
          const float alpha = need ? ptx::ex2_approx(m_used - mx) : 1.0f;
          if (need) m_used = mx;
          l *= alpha;
          for (int c0 = half * 16; c0 < p.dpad; c0 += 32) {  // the two warps of a row interleave 16-column chunks
            uint32_t v[16];
            ptx::tmem_ld_32x16(tOg + c0, v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(u2f(v[i]) * alpha);
            ptx::tmem_st_32x16(tOg + c0, v);
          }
        }
      }
      // P = exp2(s - m) as bf16 pairs into the group's P columns (16 keys = 8 columns per store)
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[8];
#pragma unroll
        for (int e2 = 0; e2 < 8; ++e2) {
          float pv[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int i = c * 16 + e2 * 2 + e;
            if (emode == 0) pv[e] = ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, -m_used), i);
            else if (emode == 1) pv[e] = (key0 + i < p.Skv) ? ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, -m_used), i) : 0.f;
            else pv[e] = ptx::ex2_mixed(fmaf(u2f(sv[i]), p.scale2, kb[i] - m_used), i);
            l4[(e2 * 2 + e) & 3] += pv[e];
          }
          o[e2] = sdb::pack_bf16x2(pv[0], pv[1]);
        }
        ptx::tmem_st_32x8(tPg + 8u * (uint32_t)c, o);
      }
      l += (l4[0] + l4[1]) + (l4[2] + l4[3]);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar_P[g]);
    }
    // ---- epilogue: O / l -> global, lse ----
    float* sLg = sL + g * 256;
    sLg[half * 128 + row] = l;
    ptx::named_bar_sync(1 + g, 256);
    l += sLg[(half ^ 1) * 128 + row];
    ptx::mbar_wait(&bar_PV[g], (uint32_t)((p.nkb - 1) & 1));
    ptx::tc_fence_after();
    const int q = q0 + g * 128 + row;
    const bool valid = q < p.Sq;
    const float inv_l = 1.0f / l;
    __nv_bfloat16* orow = p.o + ((long long)b * p.Sq + q) * p.ldo + (long long)h * p.d;
    for (int c0 = half * 16; c0 < p.dpad; c0 += 32) {
      uint32_t v[16];
      ptx::tmem_ld_32x16(tOg + c0, v);
      ptx::tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c0 + c * 8 < p.d) {
            uint4 o;
            o.x = sdb::pack_bf16x2(u2f(v[c * 8 + 0]) * inv_l, u2f(v[c * 8 + 1]) * inv_l);
            o.y = sdb::pack_bf16x2(u2f(v[c * 8 + 2]) * inv_l, u2f(v[c * 8 + 3]) * inv_l);
            o.z = sdb::pack_bf16x2(u2f(v[c * 8 + 4]) * inv_l, u2f(v[c * 8 + 5]) * inv_l);
            o.w = sdb::pack_bf16x2(u2f(v[c * 8 + 6]) * inv_l, u2f(v[c * 8 + 7]) * inv_l);
            sdb::st_u4(orow + c0 + c * 8, o);
          }
        }
      }
    }
    if (valid && half == 0) p.lse2[((long long)b * p.H + h) * p.Sq + q] = m_used + log2f(l);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 17) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// =============================================== backward ===============================================
// 672 threads: warps 0-15 = compute (two out-of-phase groups of 8 warps: group g owns the column blocks i = g, g+2, ..
// and S / dP buffer g; thread = (row, column half) inside a group; no row reductions are needed in backward because
// lse and D are inputs), warp 17 = TMA producer, and FOUR issuing warps — 16 scores (+ TMEM owner), 18 dP, 19 first
// accumulator (dK / dQ), 20 second accumulator (dV, mode 0) — each waiting on its own barriers; the issue loops run on
// the whole warp with the MMAs under elect.sync (ptx::elect_one). Sixteen compute warps (four per scheduler) rather
// than eight: the kernel is bound by the compute warps' exp2 / conversion instruction stream, whose TMEM-load and MUFU
// latencies two warps per scheduler could not hide (46 % issue utilisation, profiles/r1n_ncu_full.md). With nbuf = 2
// (head dim <= 128) S / dP are double-buffered in TMEM; the bf16 P^T / dS^T operands of the accumulating MMAs are
// written back to tensor memory (A operand from TMEM), so shared memory carries only the TMA tiles.
constexpr int kBwdThreads = 672;
constexpr int kBwdComputeThreads = 512;
constexpr int kBwdMaxStages = 8;  // TMA latency under load is ~1.5 us and a block takes ~0.5 us: the ring must run well ahead

struct BwdParams {
  CUtensorMap mapR1, mapR2, mapC1, mapC2;
  const float* lse2;
  const float* dvec;
  const float* kbias2;
  __nv_bfloat16* out1;
  long long ld1;
  __nv_bfloat16* out2;
  long long ld2;
  int mode;  // 0: rows = keys (out1 = dK, out2 = dV); 1: rows = queries (out1 = dQ)
  int B, H, Sq, Skv, Srow, Scol, d, dpad, atoms, bc, stages, ncb, nbuf;
  int alias;  // 1: the bf16 P^T / dS^T operands overwrite the S / dP columns they came from (no TMEM left for a copy)
  float scale2, scale;
};

template <int MODE, int NBUF>
__global__ void __launch_bounds__(kBwdThreads, 1) attn_bwd_kernel(const __grid_constant__ BwdParams p) {
  sdb::pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const uint32_t rtile = (uint32_t)p.atoms * kAtom128;          // row tile: 128 rows
  const uint32_t catom = (uint32_t)p.bc * 128u;                 // column tile atom: bc rows x 128 B
  const uint32_t ctile = (uint32_t)p.atoms * catom;
  uint8_t* sR1 = smem;
  uint8_t* sR2 = sR1 + rtile;
  uint8_t* sC1 = sR2 + rtile;
  uint8_t* sC2 = sC1 + (size_t)p.stages * ctile;
  float* sColA = reinterpret_cast<float*>(sC2 + (size_t)p.stages * ctile);  // [2][128]
  float* sColB = sColA + 256;                                             // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sColB + 256);
  uint64_t* barR = bars;
  uint64_t* c_full = bars + 1;                     // [kBwdMaxStages]
  uint64_t* c_empty = c_full + kBwdMaxStages;      // [kBwdMaxStages]
  uint64_t* bar_SdP = c_empty + kBwdMaxStages;     // [2]
  uint64_t* bar_Sfree = bar_SdP + 2;               // [2]  compute has the scores of the buffer in registers
  uint64_t* bar_PdS = bar_Sfree + 2;               // [2]  bf16 P^T / dS^T of the buffer written to tensor memory
  uint64_t* bar_acc = bar_PdS + 2;                 // [2]  accumulation that read them complete
  uint64_t* bar_done = bar_acc + 2;                // every accumulation of the CTA has completed
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_done + 1);

  const int warp = ptx::warp_idx_uniform(), lane = threadIdx.x & 31;
  const int r0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;

  if (warp == 17 && lane == 0) {
    ptx::tma_prefetch_desc(&p.mapR1);
    ptx::tma_prefetch_desc(&p.mapR2);
    ptx::tma_prefetch_desc(&p.mapC1);
    ptx::tma_prefetch_desc(&p.mapC2);
    ptx::mbar_init(barR, 1);
    constexpr uint32_t kAccIssuers = (MODE == 0) ? 2u : 1u;
    for (int s = 0; s < kBwdMaxStages; ++s) {
      ptx::mbar_init(&c_full[s], 1);
      ptx::mbar_init(&c_empty[s], kAccIssuers);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&bar_SdP[s], 2);
      ptx::mbar_init(&bar_Sfree[s], NBUF == 2 ? 256 : kBwdComputeThreads);
      ptx::mbar_init(&bar_PdS[s], NBUF == 2 ? 256 : kBwdComputeThreads);
      ptx::mbar_init(&bar_acc[s], kAccIssuers);
    }
    ptx::mbar_init(bar_done, kAccIssuers);
    ptx::fence_barrier_init();
  }
  if (warp == 16) {
    ptx::tmem_alloc(tmem_ptr_smem, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  sdb::pdl_wait();  // prologue above overlaps the previous kernel's tail
  // [S0 | dP0 | (S1 | dP1) | A1 | A2 | bf16 operands: (dS^T0 | P^T0) (dS^T1 | P^T1)]. The compute warps hand P^T / dS^T to
  // the accumulating MMAs THROUGH TENSOR MEMORY (A operand from TMEM): shared memory carries only the TMA tiles, and it
  // was the saturated resource when the operands were staged there (profiles/r1w_attn_bwd_timeline.md). A thread owns
  // row = lane and writes the 16 columns of a slice it has just processed as 8 packed columns; k-step k of a block
  // starts at column (k / kps) * slice_stride + (k % kps) * 8 with kps = k-steps per thread slice. When 512 columns
  // cannot hold a separate copy (p.alias) the operands overwrite the front of each thread slice of S / dP in place.
  const uint32_t tS = tmem_base;
  const uint32_t tA1 = tmem_base + 2u * (uint32_t)p.bc * (uint32_t)NBUF;
  const uint32_t tA2 = tA1 + (uint32_t)p.dpad;
  const uint32_t tOp = tA1 + (MODE == 0 ? 2u : 1u) * (uint32_t)p.dpad;  // separate operand copies (not aliased)
  const uint32_t op_stride = (MODE == 0 ? 2u : 1u) * (uint32_t)(p.bc / 2);
  const int wcols_k = p.bc / (NBUF == 2 ? 2 : 4);                           // columns per thread slice
  const uint32_t slice_stride = p.alias ? (uint32_t)wcols_k : (uint32_t)(wcols_k / 2);
  auto op_base = [&](int buf, bool pt) -> uint32_t {  // pt: P^T (second accumulator), else dS^T
    if (p.alias) return tS + (uint32_t)buf * 2u * (uint32_t)p.bc + (pt ? 0u : (uint32_t)p.bc);
    return tOp + (uint32_t)buf * op_stride + (pt ? (uint32_t)(p.bc / 2) : 0u);
  };

  if (warp == 17) {
    {  // whole warp; the TMA instructions under elect
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(barR, 2 * rtile);
        tma_tile(sR1, &p.mapR1, barR, p.atoms, kAtom128, r0, h, b);
        tma_tile(sR2, &p.mapR2, barR, p.atoms, kAtom128, r0, h, b);
      }
      __syncwarp();
      for (int i = 0; i < p.ncb; ++i) {
        const int s = i % p.stages, n = i / p.stages;
        SDB_ISSUER_WAIT(&c_empty[s], (uint32_t)((n & 1) ^ 1));
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&c_full[s], 2 * ctile);
          tma_tile(sC1 + (size_t)s * ctile, &p.mapC1, &c_full[s], p.atoms, catom, i * p.bc, h, b);
          tma_tile(sC2 + (size_t)s * ctile, &p.mapC2, &c_full[s], p.atoms, catom, i * p.bc, h, b);
        }
        __syncwarp();
      }
    }
  } else if (warp == 16 || warp == 18) {
    // ===================== score issuers: warp 16  S = R1 C1^T,  warp 18  dP = R2 C2^T =====================
    // (whole warp, warp-uniform values, MMAs under elect: see ptx::elect_one)
    {
      const bool second = warp == 18;
      const uint32_t idescS = ptx::umma_idesc_bf16(p.bc, 0, 0);
      const uint64_t dR = ptx::umma_desc_sw128(ptx::smem_u32(second ? sR2 : sR1), 0u, 1024u);
      const uint64_t dCk = ptx::umma_desc_sw128(ptx::smem_u32(second ? sC2 : sC1), 0u, 1024u);  // column tile as K-major B
      const uint32_t atom16 = kAtom128 >> 4, catom16 = catom >> 4, ctile16 = ctile >> 4;
      const int ksteps = p.dpad / 16;
      int ss = 0;
      uint32_t sph = 0;
      ptx::mbar_wait(barR, 0);
      for (int i = 0; i < p.ncb; ++i) {
        const int buf = i & (NBUF - 1);
        // S/dP buffer `buf` was last read by compute(i - NBUF), which arrives on bar_Sfree as soon as the scores are
        // in its registers (before the exp2 / pack / store work on them)
        if (i >= NBUF) {
          SDB_ISSUER_WAIT(&bar_Sfree[buf], (uint32_t)(((i - NBUF) >> (NBUF - 1)) & 1));
          // aliased operands live in the S / dP columns until the accumulations of that block have read them
          if (p.alias) SDB_ISSUER_WAIT(&bar_acc[buf], (uint32_t)(((i - NBUF) >> (NBUF - 1)) & 1));
        }
        SDB_TL(second ? 1 : 0, i, 0);
        SDB_ISSUER_WAIT(&c_full[ss], sph);
        ptx::tc_fence_after();
        SDB_TL(second ? 1 : 0, i, 1);
        const uint32_t tSi = tS + (uint32_t)buf * 2u * (uint32_t)p.bc + (second ? (uint32_t)p.bc : 0u);
        if (ptx::elect_one()) {
          mma_ksteps<false>(tSi, dR, atom16, dCk + (uint32_t)ss * ctile16, catom16, idescS, ksteps, 0u);
          ptx::umma_commit(&bar_SdP[buf]);
        }
        __syncwarp();
        SDB_TL(second ? 1 : 0, i, 2);
        if (++ss == p.stages) { ss = 0; sph ^= 1u; }
      }
    }
  } else if (warp == 19 || warp == 20) {
    // ===================== accumulate issuers: warp 19  A1 += dS C1,  warp 20 (mode 0)  A2 += P^T C2 =====================
    if (warp == 19 || MODE == 0) {
      const bool second = warp == 20;
      const uint32_t idescA = ptx::umma_idesc_bf16(p.dpad, 0, 1);
      const uint64_t dCm = ptx::umma_desc_sw128(ptx::smem_u32(second ? sC2 : sC1), catom, 1024u);  // MN-major B
      const uint32_t ctile16 = ctile >> 4;
      const int k2 = p.bc / 16, kps = wcols_k / 16;
      const uint32_t tA = second ? tA2 : tA1;
      int as = 0;
      uint32_t aph = 0;
      for (int i = 0; i < p.ncb; ++i) {
        const int buf = i & (NBUF - 1);
        SDB_TL(second ? 3 : 2, i, 0);
        SDB_ISSUER_WAIT(&bar_PdS[buf], (uint32_t)((i >> (NBUF - 1)) & 1));
        SDB_ISSUER_WAIT(&c_full[as], aph);  // already complete (the scores of block i used it); orders the async proxy
        ptx::tc_fence_after();
        SDB_TL(second ? 3 : 2, i, 1);
        const uint32_t aop = op_base(buf, second);
        const uint64_t dB = dCm + (uint32_t)as * ctile16;
        if (ptx::elect_one()) {
          for (int k = 0, ks = 0, sl = 0; k < k2; ++k) {
            ptx::umma_bf16_ts(tA, aop + (uint32_t)sl * slice_stride + (uint32_t)ks * 8u, dB + 128u * (uint32_t)k, idescA,
                              (k > 0 || i > 0) ? 1u : 0u);
            if (++ks == kps) { ks = 0; ++sl; }
          }
          ptx::umma_commit(&c_empty[as]);
          ptx::umma_commit(&bar_acc[buf]);
          if (i == p.ncb - 1) ptx::umma_commit(bar_done);
        }
        __syncwarp();
        SDB_TL(second ? 3 : 2, i, 2);
        if (++as == p.stages) { as = 0; aph ^= 1u; }
      }
    }
  } else {
    // ===================== compute threads =====================
    // Two independent groups of 8 warps: group g works on the column blocks i = g, g + 2, ... (TMEM / smem buffer g),
    // thread = (row, column half) inside a group. The groups run out of phase, so on every scheduler one pair of warps
    // computes while the other waits for its block's MMAs (with all 16 warps on the same block everybody waited at
    // the same time).
    const int q4 = warp & 3, part = warp >> 2;       // part: column quarter for the final drain
    const int grp = (NBUF == 2) ? (warp >> 3) : 0;   // single-buffered (head dim > 128): one group of 16 warps
    const int nsub = (NBUF == 2) ? 2 : 4;            // column slices per block inside a group
    const int sub = (NBUF == 2) ? ((warp >> 2) & 1) : part;
    const int gthreads = (NBUF == 2) ? 256 : 512;
    const int gtid = (NBUF == 2) ? ((int)threadIdx.x & 255) : (int)threadIdx.x;
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q4 * 32)) << 16;
    const int r = r0 + row;
    const long long bh = (long long)b * p.H + h;
    float rowA, rowD = 0.f;  // MODE 0: rowA = key bias; MODE 1: rowA = -lse2[q], rowD = D[q]
    if (MODE == 0) {
      rowA = (r < p.Skv) ? (p.kbias2 ? p.kbias2[r] : 0.f) : -INFINITY;
    } else {
      rowA = (r < p.Sq) ? -p.lse2[bh * p.Sq + r] : -INFINITY;
      rowD = (r < p.Sq) ? p.dvec[bh * p.Sq + r] : 0.f;
    }
    // per-column metadata of block i, fetched one block ahead into registers by the first bc threads of the group.
    // MODE 1 needs it only when a key bias is given or the block is cut by Skv (plain blocks skip the LDS entirely).
    float nA = 0.f, nB = 0.f;
    auto fetch_cols = [&](int i) {
      if (gtid < p.bc && i < p.ncb) {
        const int c = i * p.bc + gtid;
        if (MODE == 0) {
          nA = (c < p.Sq) ? -p.lse2[bh * p.Sq + c] : -INFINITY;
          nB = (c < p.Sq) ? p.dvec[bh * p.Sq + c] : 0.f;
        } else {
          nA = (c < p.Skv) ? (p.kbias2 ? p.kbias2[c] : 0.f) : -INFINITY;
        }
      }
    };
    const int istep = (NBUF == 2) ? 2 : 1;
    fetch_cols(grp);
    const int wcols = p.bc / nsub;        // 32 (two slices) or 16 (four slices) columns per thread and block
    for (int i = grp; i < p.ncb; i += istep) {
      const int buf = i & (NBUF - 1);     // == grp when double-buffered
      float* cA = sColA + buf * 128;
      float* cB = sColB + buf * 128;
      const bool col_meta = (MODE == 0) || (p.kbias2 != nullptr) || ((i + 1) * p.bc > p.Skv);
      // the group's previous block (i - istep) used the same cA/cB: every thread of the group has finished reading
      // them before it arrived on bar_PdS, and this barrier orders those reads before the overwrite below
      ptx::named_bar_sync(1 + grp, gthreads);
      if (gtid < p.bc) {
        cA[gtid] = nA;
        if (MODE == 0) cB[gtid] = nB;
      }
      ptx::named_bar_sync(1 + grp, gthreads);
      fetch_cols(i + istep);
      if (gtid == 0) SDB_TL(4 + grp, i, 0);
      ptx::mbar_wait_sleep(&bar_SdP[buf], (uint32_t)((i >> (NBUF - 1)) & 1), 20);
      ptx::tc_fence_after();
      if (gtid == 0) SDB_TL(4 + grp, i, 1);
      const uint32_t tSi = tS + (uint32_t)buf * 2u * (uint32_t)p.bc + lane_base;
      const uint32_t tdPi = tSi + (uint32_t)p.bc;
      const uint32_t opD = op_base(buf, false) + lane_base + (uint32_t)sub * slice_stride;
      const uint32_t opP = op_base(buf, true) + lane_base + (uint32_t)sub * slice_stride;
#pragma unroll 1
      for (int cc = 0; cc < wcols; cc += 16) {
        const int col0 = sub * wcols + cc;
        uint32_t sv[16], dv[16];
        ptx::tmem_ld_32x16(tSi + col0, sv);
        ptx::tmem_ld_32x16(tdPi + col0, dv);
        ptx::tmem_ld_wait();
        if (cc + 16 >= wcols) {  // last slice loaded: the tensor pipe may overwrite S / dP[buf] with block i + NBUF
          ptx::tc_fence_before();
          ptx::mbar_arrive(&bar_Sfree[buf]);
        }
        float pv[16], ds[16];  // dS without the softmax scale: it is applied once when the accumulator is drained
        if (col_meta) {
          const float4* cA4 = reinterpret_cast<const float4*>(cA + col0);
          const float4* cB4 = reinterpret_cast<const float4*>(cB + col0);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            const float4 a4 = cA4[k4];
            const float av[4] = {a4.x, a4.y, a4.z, a4.w};
            float bv[4] = {rowD, rowD, rowD, rowD};
            if (MODE == 0) {
              const float4 b4 = cB4[k4];
              bv[0] = b4.x; bv[1] = b4.y; bv[2] = b4.z; bv[3] = b4.w;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int k = k4 * 4 + e;
              const float pk = ptx::ex2_mixed(fmaf(u2f(sv[k]), p.scale2, av[e] + rowA), k);  // bias - lse; -inf masks
              pv[k] = pk;
              ds[k] = pk * (u2f(dv[k]) - bv[e]);
            }
          }
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const float pk = ptx::ex2_mixed(fmaf(u2f(sv[k]), p.scale2, rowA), k);
            pv[k] = pk;
            ds[k] = pk * (u2f(dv[k]) - rowD);
          }
        }
        // the operand columns of this buffer were last read by the accumulations of block i - NBUF
        if (cc == 0 && i >= NBUF) {
          ptx::mbar_wait_sleep(&bar_acc[buf], (uint32_t)(((i - NBUF) >> (NBUF - 1)) & 1), 20);
          ptx::tc_fence_after();
        }
        if (gtid == 0 && cc == 0) SDB_TL(4 + grp, i, 2);
        uint32_t o[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) o[c] = sdb::pack_bf16x2(ds[2 * c], ds[2 * c + 1]);
        ptx::tmem_st_32x8(opD + (uint32_t)(cc >> 1), o);
        if (MODE == 0) {
#pragma unroll
          for (int c = 0; c < 8; ++c) o[c] = sdb::pack_bf16x2(pv[2 * c], pv[2 * c + 1]);
          ptx::tmem_st_32x8(opP + (uint32_t)(cc >> 1), o);
        }
      }
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&bar_PdS[buf]);
      if (gtid == 0) SDB_TL(4 + grp, i, 3);
    }
    // ---- drain the accumulators: bar_done completes when every issuer's last accumulation has (a thread must not
    // wait on bar_acc of the OTHER group's buffer: it has not followed that barrier's phases) ----
    ptx::mbar_wait(bar_done, 0);
    ptx::tc_fence_after();
    const bool valid = r < p.Srow;
    for (int which = 0; which < (MODE == 0 ? 2 : 1); ++which) {
      const uint32_t tA = which == 0 ? tA1 : tA2;
      const float osc = which == 0 ? p.scale : 1.0f;  // dK / dQ carry the softmax scale; dV does not
      __nv_bfloat16* base = which == 0 ? p.out1 : p.out2;
      const long long ld = which == 0 ? p.ld1 : p.ld2;
      __nv_bfloat16* orow = base + ((long long)b * p.Srow + r) * ld + (long long)h * p.d;
      for (int c0 = part * 16; c0 < p.dpad; c0 += 64) {
        uint32_t v[16];
        ptx::tmem_ld_32x16(tA + lane_base + c0, v);
        ptx::tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            if (c0 + c * 8 < p.d) {
              uint4 o;
              o.x = sdb::pack_bf16x2(u2f(v[c * 8 + 0]) * osc, u2f(v[c * 8 + 1]) * osc);
              o.y = sdb::pack_bf16x2(u2f(v[c * 8 + 2]) * osc, u2f(v[c * 8 + 3]) * osc);
              o.z = sdb::pack_bf16x2(u2f(v[c * 8 + 4]) * osc, u2f(v[c * 8 + 5]) * osc);
              o.w = sdb::pack_bf16x2(u2f(v[c * 8 + 6]) * osc, u2f(v[c * 8 + 7]) * osc);
              sdb::st_u4(orow + c0 + c * 8, o);
            }
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 16) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// D[b,h,q] = sum_c dO[b,q,h,c] * O[b,q,h,c]
__global__ void __launch_bounds__(256)
attn_dvec_kernel(const __nv_bfloat16* __restrict__ o, long long ldo, const __nv_bfloat16* __restrict__ dO,
                 long long lddo, float* __restrict__ dvec, int B, int H, int Sq, int d) {
  sdb::pdl_trigger();
  const long long total = (long long)B * Sq * H;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int h = (int)(idx % H);
    const long long bq = idx / H;
    const int q = (int)(bq % Sq);
    const int b = (int)(bq / Sq);
    const __nv_bfloat16* po = o + bq * ldo + (long long)h * d;
    const __nv_bfloat16* pd = dO + bq * lddo + (long long)h * d;
    float acc = 0.f;
    for (int c = 0; c < d; c += 8) {
      const uint4 a = sdb::ld_u4(po + c), g = sdb::ld_u4(pd + c);
      acc = fmaf(sdb::bf16_lo(a.x), sdb::bf16_lo(g.x), acc); acc = fmaf(sdb::bf16_hi(a.x), sdb::bf16_hi(g.x), acc);
      acc = fmaf(sdb::bf16_lo(a.y), sdb::bf16_lo(g.y), acc); acc = fmaf(sdb::bf16_hi(a.y), sdb::bf16_hi(g.y), acc);
      acc = fmaf(sdb::bf16_lo(a.z), sdb::bf16_lo(g.z), acc); acc = fmaf(sdb::bf16_hi(a.z), sdb::bf16_hi(g.z), acc);
      acc = fmaf(sdb::bf16_lo(a.w), sdb::bf16_lo(g.w), acc); acc = fmaf(sdb::bf16_hi(a.w), sdb::bf16_hi(g.w), acc);
    }
    dvec[((long long)b * H + h) * Sq + q] = acc;
  }
}

// ---------------- host side ----------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn attn_encode_fn() {
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

// [B, S, H*d] tensor with row stride ld -> map over (d, S, H, B), box (64, box_rows, 1, 1)
int encode_heads(CUtensorMap* map, const void* ptr, long long ld, int B, int S, int H, int d, int box_rows,
                 const char* what) {
  EncodeTiledFn fn = attn_encode_fn();
  if (!fn) return sdb_set_error(SDB_ERR_CUDA, "sdb_attn: cuTensorMapEncodeTiled entry point unavailable");
  if (ptr == nullptr || (((uintptr_t)ptr) & 15) != 0 || ld % 8 != 0 || d % 8 != 0)
    return sdb_set_error(SDB_ERR_INVALID, "sdb_attn: %s must be 16-byte aligned with ld %% 8 == 0 and d %% 8 == 0", what);
  cuuint64_t gdim[4] = {(cuuint64_t)d, (cuuint64_t)S, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstr[3] = {(cuuint64_t)ld * 2, (cuuint64_t)d * 2, (cuuint64_t)S * (cuuint64_t)ld * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return sdb_set_error(SDB_ERR_CUDA, "sdb_attn: cuTensorMapEncodeTiled(%s) failed with %d (B %d S %d H %d d %d ld %lld)",
                         what, (int)r, B, S, H, d, ld);
  return SDB_OK;
}

int check_shape(const char* name, int B, int H, int Sq, int Skv, int d) {
  SDB_CHECK_ARG(B > 0 && H > 0 && Sq > 0 && Skv > 0, "%s: bad shape", name);
  SDB_CHECK_ARG(d % 8 == 0 && d >= 16 && d <= 192, "%s: head dim %d must be a multiple of 8 in [16,192]", name, d);
  SDB_CHECK_ARG(H <= 65535 && B <= 65535, "%s: H/B too large for the grid", name);
  return SDB_OK;
}

}  // namespace

extern "C" int sdb_attn_fwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                            void* o, long long ldo, float* lse2, const float* kbias2, int B, int H, int Sq, int Skv,
                            int d, float scale, cudaStream_t stream) {
  int rc = check_shape("sdb_attn_fwd", B, H, Sq, Skv, d);
  if (rc) return rc;
  SDB_CHECK_ARG(o != nullptr && lse2 != nullptr && (((uintptr_t)o) & 15) == 0 && ldo % 8 == 0,
                "sdb_attn_fwd: output must be 16-byte aligned with ldo %% 8 == 0");
  FwdParams p;
  memset(&p, 0, sizeof(p));
  if ((rc = encode_heads(&p.mapQ, q, ldq, B, Sq, H, d, 128, "q"))) return rc;
  if ((rc = encode_heads(&p.mapK, k, ldk, B, Skv, H, d, 128, "k"))) return rc;
  if ((rc = encode_heads(&p.mapV, v, ldv, B, Skv, H, d, 128, "v"))) return rc;
  p.o = reinterpret_cast<__nv_bfloat16*>(o);
  p.ldo = ldo;
  p.lse2 = lse2;
  p.kbias2 = kbias2;
  p.B = B; p.H = H; p.Sq = Sq; p.Skv = Skv; p.d = d;
  p.dpad = (d + 15) & ~15;
  p.atoms = (d + 63) / 64;
  p.nkb = (Skv + 127) / 128;
  p.npv = (256 + 2 * p.dpad <= 512) ? 2 : 1;  // two PV accumulators when they fit beside the double-buffered S
  p.scale2 = scale * 1.4426950408889634f;
  if (kbias2 != nullptr && p.nkb > kMaxBiasBlocks)
    return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_attn_fwd: a per-key bias supports Skv <= %d (got %d)",
                         kMaxBiasBlocks * 128, Skv);
  const size_t tile = (size_t)p.atoms * kAtom128;
  static const int fwd_env = [] { const char* e = getenv("SDB_ATTN_FWD"); return e ? atoi(e) : 0; }();
  // two query tiles per CTA when S, O and a bf16 copy of P for both fit the 512 tensor-memory columns (head dim <= 64)
  if (p.dpad <= 64 && Sq > 128 && fwd_env != 1) {
    const size_t fixed2 = 2 * tile + (1024 + 512) * sizeof(float) +
                          (kbias2 ? (size_t)p.nkb * 128 * sizeof(float) : 0) + 256 /*barriers*/ + 1024 /*align*/;
    p.kstages = p.vstages = (p.nkb < kFwdMaxStages) ? p.nkb : kFwdMaxStages;
    const size_t smem2 = fixed2 + (size_t)(p.kstages + p.vstages) * tile;
    static std::once_flag once2;
    static cudaError_t attr_err2 = cudaSuccess;
    std::call_once(once2, [] {
      attr_err2 = cudaFuncSetAttribute(attn_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    });
    if (attr_err2 != cudaSuccess)
      return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err2));
    dim3 grid2((Sq + 255) / 256, H, B);
    if (sdb::launch_pdl(attn_fwd2_kernel, grid2, dim3(kFwd2Threads), smem2, stream, p) != cudaSuccess) {
      const cudaError_t e = cudaGetLastError();
      return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_fwd: launch failed: %s", cudaGetErrorString(e));
    }
    SDB_CUDA_LAUNCH_CHECK("sdb_attn_fwd");
    return SDB_OK;
  }
  const size_t fixed = tile + 4 * kAtom128 + (512 + 256) * sizeof(float) +
                       (kbias2 ? (size_t)p.nkb * 128 * sizeof(float) : 0) + 256 /*barriers*/ + 1024 /*align*/;
  const size_t budget = 227 * 1024;
  p.kstages = p.vstages = (p.nkb < kFwdMaxStages) ? p.nkb : kFwdMaxStages;
  while (fixed + (size_t)(p.kstages + p.vstages) * tile > budget && p.kstages + p.vstages > 2) {
    if (p.vstages >= p.kstages) --p.vstages; else --p.kstages;
  }
  const size_t smem_bytes = fixed + (size_t)(p.kstages + p.vstages) * tile;
  if (smem_bytes > budget) return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_attn_fwd: head dim %d needs too much smem", d);
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  if (attr_err != cudaSuccess)
    return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
  dim3 grid((Sq + 127) / 128, H, B);
  if (sdb::launch_pdl(attn_fwd_kernel, grid, dim3(kFwdThreads), smem_bytes, stream, p) != cudaSuccess) {
    const cudaError_t e = cudaGetLastError();
    return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_fwd: launch failed: %s", cudaGetErrorString(e));
  }
  SDB_CUDA_LAUNCH_CHECK("sdb_attn_fwd");
  return SDB_OK;
}

extern "C" size_t sdb_attn_bwd_workspace_bytes(int B, int H, int Sq) { return (size_t)B * H * Sq * sizeof(float); }

extern "C" int sdb_attn_bwd(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                            const void* o, long long ldo, const void* dO, long long lddo, const float* lse2,
                            const float* kbias2, void* dq, long long lddq, void* dk, long long lddk, void* dv,
                            long long lddv, int B, int H, int Sq, int Skv, int d, float scale, void* workspace,
                            size_t ws_bytes, cudaStream_t stream) {
  int rc = check_shape("sdb_attn_bwd", B, H, Sq, Skv, d);
  if (rc) return rc;
  if (ws_bytes < sdb_attn_bwd_workspace_bytes(B, H, Sq))
    return sdb_set_error(SDB_ERR_WORKSPACE, "sdb_attn_bwd: workspace %zu < %zu", ws_bytes,
                         sdb_attn_bwd_workspace_bytes(B, H, Sq));
  SDB_CHECK_ARG(dq && dk && dv && ((((uintptr_t)dq) | ((uintptr_t)dk) | ((uintptr_t)dv)) & 15) == 0 &&
                    lddq % 8 == 0 && lddk % 8 == 0 && lddv % 8 == 0 && ldo % 8 == 0 && lddo % 8 == 0,
                "sdb_attn_bwd: gradient outputs must be 16-byte aligned with row strides %% 8 == 0");
  float* dvec = reinterpret_cast<float*>(workspace);
  {
    const long long total = (long long)B * Sq * H;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    attn_dvec_kernel<<<(int)blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(o), ldo,
                                                      reinterpret_cast<const __nv_bfloat16*>(dO), lddo, dvec, B, H, Sq, d);
    SDB_CUDA_LAUNCH_CHECK("sdb_attn_bwd(dvec)");
  }
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(attn_bwd_kernel<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(attn_bwd_kernel<0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(attn_bwd_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(attn_bwd_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  if (attr_err != cudaSuccess)
    return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));

  for (int mode = 0; mode < 2; ++mode) {
    BwdParams p;
    memset(&p, 0, sizeof(p));
    p.mode = mode;
    p.B = B; p.H = H; p.Sq = Sq; p.Skv = Skv; p.d = d;
    p.dpad = (d + 15) & ~15;
    p.atoms = (d + 63) / 64;
    p.bc = 64;
    p.nbuf = (p.dpad <= 128) ? 2 : 1;
    p.Srow = mode == 0 ? Skv : Sq;
    p.Scol = mode == 0 ? Sq : Skv;
    p.ncb = (p.Scol + p.bc - 1) / p.bc;
    p.scale = scale;
    p.scale2 = scale * 1.4426950408889634f;
    p.lse2 = lse2; p.dvec = dvec; p.kbias2 = kbias2;
    if (mode == 0) {
      if ((rc = encode_heads(&p.mapR1, k, ldk, B, Skv, H, d, 128, "k"))) return rc;
      if ((rc = encode_heads(&p.mapR2, v, ldv, B, Skv, H, d, 128, "v"))) return rc;
      if ((rc = encode_heads(&p.mapC1, q, ldq, B, Sq, H, d, p.bc, "q"))) return rc;
      if ((rc = encode_heads(&p.mapC2, dO, lddo, B, Sq, H, d, p.bc, "dO"))) return rc;
      p.out1 = reinterpret_cast<__nv_bfloat16*>(dk); p.ld1 = lddk;
      p.out2 = reinterpret_cast<__nv_bfloat16*>(dv); p.ld2 = lddv;
    } else {
      if ((rc = encode_heads(&p.mapR1, q, ldq, B, Sq, H, d, 128, "q"))) return rc;
      if ((rc = encode_heads(&p.mapR2, dO, lddo, B, Sq, H, d, 128, "dO"))) return rc;
      if ((rc = encode_heads(&p.mapC1, k, ldk, B, Skv, H, d, p.bc, "k"))) return rc;
      if ((rc = encode_heads(&p.mapC2, v, ldv, B, Skv, H, d, p.bc, "v"))) return rc;
      p.out1 = reinterpret_cast<__nv_bfloat16*>(dq); p.ld1 = lddq;
      p.out2 = nullptr; p.ld2 = 0;
    }
    const size_t rtile = (size_t)p.atoms * kAtom128;
    const size_t ctile = (size_t)p.atoms * p.bc * 128;
    int stages = 1;
    size_t fixed = 0;
    for (;;) {
      fixed = 2 * rtile + 512 * sizeof(float) + 256 /*barriers*/ + 1024 /*align*/;
      stages = (int)((227 * 1024 - (long long)fixed) / (long long)(2 * ctile));
      if (stages > kBwdMaxStages) stages = kBwdMaxStages;
      if (stages > p.ncb) stages = p.ncb;
      // the double-buffered schedule issues S/dP of block i+1 before block i's operands are released: needs 2 stages
      if (p.nbuf == 2 && p.ncb > 1 && stages < 2) { p.nbuf = 1; continue; }
      break;
    }
    {  // tensor memory: S / dP buffers, accumulators, and (when they fit) separate bf16 operand columns
      const int nacc = mode == 0 ? 2 : 1;
      const int need = 2 * p.bc * p.nbuf + nacc * p.dpad + p.nbuf * nacc * (p.bc / 2);
      static const int alias_env = [] { const char* e = getenv("SDB_ATTN_ALIAS"); return e ? atoi(e) : 0; }();
      p.alias = (need > 512 || alias_env) ? 1 : 0;
    }
    if (stages < 1) return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_attn_bwd: head dim %d needs too much smem", d);
    p.stages = stages;
    const size_t smem_bytes = fixed + (size_t)p.stages * 2 * ctile;
    dim3 grid((p.Srow + 127) / 128, H, B);
    const dim3 b3(kBwdThreads);
    cudaError_t lerr;
    if (mode == 0 && p.nbuf == 2) lerr = sdb::launch_pdl(attn_bwd_kernel<0, 2>, grid, b3, smem_bytes, stream, p);
    else if (mode == 0) lerr = sdb::launch_pdl(attn_bwd_kernel<0, 1>, grid, b3, smem_bytes, stream, p);
    else if (p.nbuf == 2) lerr = sdb::launch_pdl(attn_bwd_kernel<1, 2>, grid, b3, smem_bytes, stream, p);
    else lerr = sdb::launch_pdl(attn_bwd_kernel<1, 1>, grid, b3, smem_bytes, stream, p);
    if (lerr != cudaSuccess) {
      cudaGetLastError();
      return sdb_set_error(SDB_ERR_CUDA, "sdb_attn_bwd: launch failed: %s", cudaGetErrorString(lerr));
    }
    SDB_CUDA_LAUNCH_CHECK("sdb_attn_bwd");
  }
  return SDB_OK;
}
