#include <cstdarg>
#include <cstring>
#include "common.cuh"

namespace nrf {
static thread_local char g_err[512] = "";
unsigned long long g_launches = 0;
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace nrf

extern "C" int nrf_abi_version(void) { return 1; }
extern "C" uint64_t nrf_launch_count(void) { return nrf::g_launches; }
extern "C" const char* nrf_last_error(void) { return nrf::g_err; }
extern "C" int nrf_check_device(void) {
  int dev = 0;
  NRF_CUDA(cudaGetDevice(&dev));
  int major = 0;
  NRF_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    nrf::set_error("nerf_b200 is built for sm_100a only; device %d has compute capability %d.x", dev, major);
    return NRF_ERR_UNSUPPORTED;
  }
  return NRF_OK;
}
