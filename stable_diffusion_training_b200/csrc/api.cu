// Error string plumbing and library-level queries of the C ABI.
#include "sdb_common.cuh"
#include "../../include/sdb200.h"
#include <cstring>
#include <atomic>

static thread_local char g_err[512] = "";

int sdb_set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

static std::atomic<long long> g_launches{0};
void sdb_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long long sdb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" const char* sdb_last_error_string(void) { return g_err; }

extern "C" int sdb_abi_version(void) { return SDB200_ABI_VERSION; }

// Compiled-for architecture check: the kernels are sm_100a only; refuse to run anywhere else (no fallback).
extern "C" int sdb_check_device(int device) {
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return sdb_set_error(SDB_ERR_CUDA, "sdb_check_device: %s", cudaGetErrorString(e));
  if (prop.major != 10)
    return sdb_set_error(SDB_ERR_UNSUPPORTED, "sdb_check_device: device %d is sm_%d%d; this library is sm_100a only",
                         device, prop.major, prop.minor);
  return SDB_OK;
}
