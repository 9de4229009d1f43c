// Shared helpers for the nerf_b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "../../include/nerf_b200.h"

namespace nrf {

void set_error(const char* fmt, ...);

#define NRF_REQUIRE(cond, ...)                         \
  do {                                                 \
    if (!(cond)) {                                     \
      ::nrf::set_error(__VA_ARGS__);                   \
      return NRF_ERR_INVALID;                          \
    }                                                  \
  } while (0)

#define NRF_CUDA(call)                                                                  \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess) {                                                           \
      ::nrf::set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return NRF_ERR_CUDA;                                                              \
    }                                                                                   \
  } while (0)

extern unsigned long long g_launches;  // kernels launched by this library (statistics only)
#define NRF_LAUNCH_CHECK()        \
  do {                            \
    ++::nrf::g_launches;          \
    NRF_CUDA(cudaGetLastError()); \
  } while (0)

// Optional per-kernel device timing (nrf_prof_enable / nrf_prof_read, error.cu): an event pair is recorded around
// the launch on the launching stream; a no-op unless enabled.
extern bool g_prof_on;
void prof_record(int kernel_id, cudaStream_t st, bool begin, int64_t rows);
struct ProfScope {
  int id;
  cudaStream_t st;
  ProfScope(int kernel_id, cudaStream_t s, int64_t rows) : id(kernel_id), st(s) { if (g_prof_on) prof_record(id, st, true, rows); }
  ~ProfScope() { if (g_prof_on) prof_record(id, st, false, 0); }
};

// Opt-in to > 48 KB of dynamic shared memory.  The attribute is per DEVICE: `done` is a per-kernel bit mask indexed by
// the device ordinal, so a process that launches on a second GPU configures the kernel there too.
#define NRF_OPT_IN_SMEM(kernel, bytes)                                                              \
  do {                                                                                              \
    static unsigned long long done__ = 0;                                                           \
    int dev__ = 0;                                                                                  \
    NRF_CUDA(cudaGetDevice(&dev__));                                                                \
    if (dev__ >= 64 || !((done__ >> dev__) & 1ull)) {                                               \
      NRF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (bytes))); \
      if (dev__ < 64) done__ |= 1ull << dev__;                                                      \
    }                                                                                               \
  } while (0)

static inline cudaStream_t as_stream(nrf_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;  // B200

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace nrf
