// Arguments and debug hooks shared by the forward kernels (mlp_tc.cu: one CTA per tile; mlp_tc_pair.cu: CTA pairs).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace nrf {
namespace tc {

struct FwdArgs {
  const uint8_t* packed;
  const float* rays;
  const float* z;
  int64_t n_rows;
  int S;
  int num_tiles;
  float* sigma_out;
  float* rgb_out;
  uint8_t* stash;      // training only
  uint32_t* masks;     // training only
  unsigned long long* dbg;  // optional per-CTA cycle counters (nrf_dbg_set_counters), null normally
  int dbg_flags;            // debug build only (NRF_DBG_FLAGS): 1 = issue no MMAs, 2 = epilogue does no arithmetic
};

#ifndef NRF_FWD_DEBUG
#define NRF_FWD_DEBUG 0   // build with NVCC_EXTRA=-DNRF_FWD_DEBUG=1 for the cycle counters / timeline (tools/fwd_*.py)
#endif
// Debug timeline (CTA 0 only, when dbg != null): dbg[kTraceBase + role * kTraceCap + n] = clock64() of event n.
constexpr int kTraceBase = 148 * 8, kTraceCap = 4096;
// dbg_flags bit 4 (16): LIGHT mode - only the NRF_WTRACE points are recorded (a clock read + one store each), so that the
// timeline is not stretched by the counters / event traces below
// NRF_FWD_DEBUG=2 compiles ONLY the light trace points in (the heavy counters alone slow the kernel by ~30 %).
#define NRF_HEAVY (NRF_FWD_DEBUG == 1 && a.dbg && !(a.dbg_flags & 16))
#define NRF_TRACE(role, n)                                                                          \
  do {                                                                                              \
    if (NRF_HEAVY && blockIdx.x == 0 && (n) < kTraceCap) a.dbg[kTraceBase + (role) * kTraceCap + (n)++] = clock64(); \
  } while (0)

// Progress markers (debug build, dbg = PINNED HOST memory so that they survive a trapped launch): the last wait every
// role of every CTA entered, dbg[kMarkBase + cta * 32 + slot] = code (tools/fwd_deadlock.py decodes them).
constexpr int kMarkBase = kTraceBase + 4 * kTraceCap;
#define NRF_MARK(slot, code)                                                                        \
  do {                                                                                              \
    if (NRF_HEAVY) ((volatile unsigned long long*)a.dbg)[kMarkBase + blockIdx.x * 32 + (slot)] = (unsigned long long)(code); \
  } while (0)


// Per-warp epilogue timeline (debug build, CTA 0, its third tile): dbg[kWarpBase + epilogue_warp * 64 + layer * 4 + e],
// e = 0 accumulator seen complete, 1 / 2 first / last hand-over done.  Row 16 = the UMMA issuer: e = 0 first A operand
// of the layer seen, 1 first group's MMAs issued, 2 last A operand seen, 3 accumulator commit issued  (tools/fwd_warp_skew.py)
constexpr int kWarpBase = kMarkBase + 148 * 32;
#define NRF_WTRACE(on, ew, layer, e)                                                                \
  do {                                                                                              \
    if (NRF_FWD_DEBUG && a.dbg && (on)) a.dbg[kWarpBase + (ew) * 64 + (layer) * 4 + (e)] = clock64(); \
  } while (0)

// forward with the hidden activations as tensor-memory operands, N = 256 (mlp_tc_tmem.cu): the training forward
int launch_fwd_tmem(bool full, bool train, const FwdArgs& a, cudaStream_t st);
// forward with the activations in tensor memory and N-split layers (debug/mlp_tc_ts.cu)
int launch_fwd_ts(bool full, bool train, const FwdArgs& a, cudaStream_t st);
// CTA-pair forward (mlp_tc_pair.cu)
int launch_fwd_pair(bool full, bool train, const FwdArgs& a, cudaStream_t st);

}  // namespace tc
}  // namespace nrf
