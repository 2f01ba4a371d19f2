// In-step sampling with jax.random's counter-based generator (Threefry-2x32-20), on the device.
//
// Replaces (reference): training_utils.py:590-624 — `noise = jax.random.normal(noise_rng, latents.shape)`
// (+ offset noise 594-606, + perturbation noise 608-615) and `timesteps = jax.random.randint(timestep_rng, (B,), 0,
// num_train_timesteps)`. The key splits (530, 590-592) are 2-6 blocks of integer arithmetic and stay on the host
// (jax_random.py); the bulk draws run here. jax itself is an absent third-party dependency: the algorithm is the
// published one of jax 0.4.x with the default threefry2x32 implementation (oracle/threefry.py states it and is pinned
// by the Random123 known answers and the key / sample values of jax's documentation).
//
// Counter layout (jax._src.prng.threefry_2x32 on iota(n)): the n counters are zero-padded to an even length and split
// in two halves; block i = (count[i], count[i + half]) yields output elements i and i + half. A data-parallel rank
// draws only ITS rows of the global array: element indices are global (offset + j), so N ranks together produce
// exactly the array one process would.
//
// Bits, uniforms and randint are bit-exact against the oracle. normal = sqrt(2) * erf_inv(u) follows XLA's ErfInv for
// F32 (Giles' two degree-8 polynomials after w = -log1p(-u^2)); log1pf / sqrtf round like CUDA's libdevice, which may
// differ from another backend's in the last bit — compared with a few-ulp tolerance (tests/test_rng_gpu.py).
#include "sdb_common.cuh"
#include "../../include/sdb200.h"

namespace {

struct Key2 {
  uint32_t k0, k1;
};

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return __funnelshift_l(x, x, r); }

__device__ __forceinline__ void threefry2x32(Key2 k, uint32_t& x0, uint32_t& x1) {
  const uint32_t ks[3] = {k.k0, k.k1, k.k0 ^ k.k1 ^ 0x1BD11BDAu};
  x0 += ks[0];
  x1 += ks[1];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    if ((i & 1) == 0) {
      x0 += x1; x1 = rotl32(x1, 13); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 15); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 26); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 6);  x1 ^= x0;
    } else {
      x0 += x1; x1 = rotl32(x1, 17); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 29); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 16); x1 ^= x0;
      x0 += x1; x1 = rotl32(x1, 24); x1 ^= x0;
    }
    x0 += ks[(i + 1) % 3];
    x1 += ks[(i + 2) % 3] + (uint32_t)(i + 1);
  }
}

// element i of threefry_2x32(key, iota(n))
__device__ __forceinline__ uint32_t bits_at(Key2 k, long long i, long long n) {
  const long long half = (n + 1) >> 1;
  const bool first = i < half;
  const long long b = first ? i : i - half;
  uint32_t x0 = (uint32_t)b;
  uint32_t x1 = (b + half < n) ? (uint32_t)(b + half) : 0u;  // the zero pad of an odd-length count array
  threefry2x32(k, x0, x1);
  return first ? x0 : x1;
}

// jax.random.uniform(minval = nextafter(-1, 0), maxval = 1): mantissa trick, (maxval - minval) == 2.0f exactly
__device__ __forceinline__ float uniform_pm1(uint32_t bits) {
  const float lo = -0.99999994f;
  const float f = __fsub_rn(__uint_as_float((bits >> 9) | 0x3F800000u), 1.0f);
  return fmaxf(lo, __fadd_rn(__fmul_rn(f, 2.0f), lo));
}

// XLA ErfInv (F32), individually rounded multiply / add like the unfused HLO
__device__ __forceinline__ float erf_inv_f32(float x) {
  float w = -log1pf(-__fmul_rn(x, x));
  const bool lt = w < 5.0f;
  w = lt ? __fsub_rn(w, 2.5f) : __fsub_rn(__fsqrt_rn(w), 3.0f);
  float p = lt ? 2.81022636e-08f : -0.000200214257f;
  p = __fadd_rn(lt ? 3.43273939e-07f : 0.000100950558f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -3.5233877e-06f : 0.00134934322f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -4.39150654e-06f : -0.00367342844f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 0.00021858087f : 0.00573950773f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -0.00125372503f : -0.0076224613f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -0.00417768164f : 0.00943887047f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 0.246640727f : 1.00167406f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 1.50140941f : 2.83297682f, __fmul_rn(p, w));
  if (fabsf(x) == 1.0f) return x * __int_as_float(0x7f800000);
  return __fmul_rn(p, x);
}

__device__ __forceinline__ float normal_at(Key2 k, long long i, long long n) {
  return __fmul_rn(1.41421354f, erf_inv_f32(uniform_pm1(bits_at(k, i, n))));
}

__global__ void __launch_bounds__(256)
bits_kernel(Key2 k, uint32_t* __restrict__ out, long long n_total, long long offset, long long n_local) {
  sdb::pdl_trigger();
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n_local; j += (long long)gridDim.x * blockDim.x)
    out[j] = bits_at(k, offset + j, n_total);
}

// noise[b, c, p] = N(noise_key)[b, c, p]  (+ N(offset_key)[b, c] * offset_mag)  (+ perturb_mag * N(perturb_key)[b, c, p])
// rows b are GLOBAL batch rows b0 .. b0 + B_local - 1 of a [B_global, C, HW] draw.
__global__ void __launch_bounds__(256)
sample_noise_kernel(Key2 kn, Key2 ko, Key2 kp, float offset_mag, float perturb_mag, float* __restrict__ out,
                    int B_global, int b0, int B_local, int C, int HW) {
  sdb::pdl_trigger();
  const long long per = (long long)C * HW;
  const long long n_total = (long long)B_global * per;
  const long long n_local = (long long)B_local * per;
  const long long base = (long long)b0 * per;
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n_local; j += (long long)gridDim.x * blockDim.x) {
    const long long i = base + j;
    float v = normal_at(kn, i, n_total);
    if (offset_mag != 0.0f)
      v = __fadd_rn(v, __fmul_rn(normal_at(ko, i / HW, (long long)B_global * C), offset_mag));
    if (perturb_mag != 0.0f) v = __fadd_rn(v, __fmul_rn(perturb_mag, normal_at(kp, i, n_total)));
    out[j] = v;
  }
}

// jax.random.randint(key, (n_total,), minval, maxval) for int32: k1, k2 = split(key); hi = bits(k1), lo = bits(k2);
// ((hi % span) * ((2^16 % span)^2 % span) + lo % span) % span + minval
__global__ void __launch_bounds__(128)
randint_kernel(Key2 k, int* __restrict__ out, long long n_total, long long offset, long long n_local, int minval,
               uint32_t span) {
  sdb::pdl_trigger();
  // split(key, 2) = threefry_2x32(key, iota(4)).reshape(2, 2): blocks (0, 2) and (1, 3)
  uint32_t a0 = 0, a1 = 2, b0 = 1, b1 = 3;
  threefry2x32(k, a0, a1);
  threefry2x32(k, b0, b1);
  const Key2 k1{a0, b0}, k2{a1, b1};
  uint32_t mult = 65536u % span;
  mult = (mult * mult) % span;
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n_local; j += (long long)gridDim.x * blockDim.x) {
    const uint32_t hi = bits_at(k1, offset + j, n_total), lo = bits_at(k2, offset + j, n_total);
    const uint32_t off = ((hi % span) * mult + (lo % span)) % span;
    out[j] = minval + (int)off;
  }
}

int grid_for(long long n, int threads) {
  long long blocks = (n + threads - 1) / threads;
  long long cap = (long long)sdb::kNumSMs * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

extern "C" int sdb_threefry_bits(uint32_t key0, uint32_t key1, uint32_t* out, long long n_total, long long offset,
                                 long long n_local, cudaStream_t stream) {
  SDB_CHECK_ARG(n_total > 0 && n_total <= 0xFFFFFFFFll && offset >= 0 && n_local >= 0 && offset + n_local <= n_total,
                "sdb_threefry_bits: bad range (n_total %lld, offset %lld, n_local %lld)", n_total, offset, n_local);
  if (n_local == 0) return SDB_OK;
  bits_kernel<<<grid_for(n_local, 256), 256, 0, stream>>>(Key2{key0, key1}, out, n_total, offset, n_local);
  SDB_CUDA_LAUNCH_CHECK("sdb_threefry_bits");
  return SDB_OK;
}

extern "C" int sdb_sample_noise(const uint32_t* noise_key_host, const uint32_t* offset_key_host,
                                const uint32_t* perturb_key_host, float offset_noise_magnitude,
                                float perturbation_noise_magnitude, float* noise_out, int batch_global, int batch_begin,
                                int batch_local, int channels, int hw, cudaStream_t stream) {
  SDB_CHECK_ARG(noise_key_host != nullptr, "sdb_sample_noise: the noise key is required");
  SDB_CHECK_ARG(batch_global > 0 && batch_begin >= 0 && batch_local > 0 && batch_begin + batch_local <= batch_global &&
                    channels > 0 && hw > 0,
                "sdb_sample_noise: bad shape");
  SDB_CHECK_ARG((long long)batch_global * channels * hw <= 0xFFFFFFFFll, "sdb_sample_noise: more than 2^32 elements");
  SDB_CHECK_ARG(offset_noise_magnitude == 0.0f || offset_key_host != nullptr, "sdb_sample_noise: offset key missing");
  SDB_CHECK_ARG(perturbation_noise_magnitude == 0.0f || perturb_key_host != nullptr,
                "sdb_sample_noise: perturbation key missing");
  const Key2 kn{noise_key_host[0], noise_key_host[1]};
  const Key2 ko = offset_key_host ? Key2{offset_key_host[0], offset_key_host[1]} : Key2{0, 0};
  const Key2 kp = perturb_key_host ? Key2{perturb_key_host[0], perturb_key_host[1]} : Key2{0, 0};
  const long long n_local = (long long)batch_local * channels * hw;
  sample_noise_kernel<<<grid_for(n_local, 256), 256, 0, stream>>>(kn, ko, kp, offset_noise_magnitude,
                                                                 perturbation_noise_magnitude, noise_out, batch_global,
                                                                 batch_begin, batch_local, channels, hw);
  SDB_CUDA_LAUNCH_CHECK("sdb_sample_noise");
  return SDB_OK;
}

extern "C" int sdb_threefry_randint(uint32_t key0, uint32_t key1, int* out, long long n_total, long long offset,
                                    long long n_local, int minval, int maxval, cudaStream_t stream) {
  SDB_CHECK_ARG(n_total > 0 && n_total <= 0xFFFFFFFFll && offset >= 0 && n_local >= 0 && offset + n_local <= n_total,
                "sdb_threefry_randint: bad range");
  if (n_local == 0) return SDB_OK;
  const uint32_t span = (maxval > minval) ? (uint32_t)((long long)maxval - (long long)minval) : 1u;
  randint_kernel<<<grid_for(n_local, 128), 128, 0, stream>>>(Key2{key0, key1}, out, n_total, offset, n_local, minval,
                                                            span);
  SDB_CUDA_LAUNCH_CHECK("sdb_threefry_randint");
  return SDB_OK;
}
