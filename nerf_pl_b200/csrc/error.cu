#include <cstdarg>
#include <cstring>
#include <vector>
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

// ---- per-kernel device timing -------------------------------------------------------------------
bool g_prof_on = false;
namespace {
struct ProfRec { cudaEvent_t e0, e1; int id; bool closed; int64_t rows; };
std::vector<ProfRec> g_prof;       // records of the current session
std::vector<cudaEvent_t> g_pool;   // events to reuse
cudaEvent_t get_event() {
  cudaEvent_t e = nullptr;
  if (!g_pool.empty()) { e = g_pool.back(); g_pool.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
}  // namespace
void prof_record(int kernel_id, cudaStream_t st, bool begin, int64_t rows) {
  if (begin) {
    ProfRec r{get_event(), get_event(), kernel_id, false, rows};
    cudaEventRecord(r.e0, st);
    g_prof.push_back(r);
  } else {
    for (size_t i = g_prof.size(); i-- > 0;) {
      if (g_prof[i].id == kernel_id && !g_prof[i].closed) {
        cudaEventRecord(g_prof[i].e1, st);
        g_prof[i].closed = true;
        break;
      }
    }
  }
}
}  // namespace nrf

extern "C" int nrf_prof_enable(int on) {
  if (on) {   // a new session: recycle the previous one's events
    for (auto& r : nrf::g_prof) { nrf::g_pool.push_back(r.e0); nrf::g_pool.push_back(r.e1); }
    nrf::g_prof.clear();
  }
  nrf::g_prof_on = on != 0;
  return NRF_OK;
}
extern "C" int nrf_prof_read(int kernel_id, double* total_ms, int64_t* launches, int64_t* rows) {
  NRF_REQUIRE(total_ms && launches && rows, "nrf_prof_read: null pointer");
  double ms = 0.0;
  int64_t n = 0, nrows = 0;
  for (auto& r : nrf::g_prof) {
    if (r.id != kernel_id || !r.closed) continue;
    NRF_CUDA(cudaEventSynchronize(r.e1));
    float t = 0.f;
    NRF_CUDA(cudaEventElapsedTime(&t, r.e0, r.e1));
    ms += t;
    ++n;
    nrows += r.rows;
  }
  *total_ms = ms;
  *launches = n;
  *rows = nrows;
  return NRF_OK;
}

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
