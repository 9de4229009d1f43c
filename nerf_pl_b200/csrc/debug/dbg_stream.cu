// Measurement aid (not on the product path): how fast can persistent CTAs pull data into shared memory
// with lane-parallel cp.async.bulk copies, as a function of piece size and address pattern?  Used to
// choose the stash layout (HBM streams of the wgrad kernel) and the weight-ring geometry (L2 streams of
// the forward / dgrad kernels).  DESIGN.md §4 quotes the numbers; tools/stream_bench.py drives it.
#include "common.cuh"
#include "tc_common.cuh"

namespace nrf {
namespace tc {

struct StreamArgs {
  const uint8_t* base;
  unsigned long long cta_stride, stream_stride, tile_stride;
  uint32_t piece;      // bytes per copy (multiple of 16)
  int n_streams;       // copies per stage (<= 32)
  int ppc;             // pieces per chunk: consecutive steps walk `ppc` adjacent pieces, then jump by tile_stride
  int wrap_tiles;      // tiles wrap around after this many (0 = never): L2-resident working sets
  int steps, stages;
};

__global__ void __launch_bounds__(64, 1) stream_kernel(const __grid_constant__ StreamArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t stage_bytes = a.piece * a.n_streams;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + (size_t)a.stages * stage_bytes);
  uint64_t* bar_empty = bar_full + a.stages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    fence_barrier_init();
  }
  __syncthreads();
  if (warp == 0) {
    const uint8_t* src = a.base + (size_t)blockIdx.x * a.cta_stride + (size_t)lane * a.stream_stride;
    int stage = 0, ph = 0, tile = 0, sub = 0;
    for (int s = 0; s < a.steps; ++s) {
      if (lane == 0) {
        mbar_wait(bar_empty + stage, ph ^ 1);
        mbar_arrive_expect_tx(bar_full + stage, stage_bytes);
      }
      __syncwarp();
      if (lane < a.n_streams)
        bulk_g2s(smem + (size_t)stage * stage_bytes + lane * a.piece,
                 src + (size_t)tile * a.tile_stride + (size_t)sub * a.piece, a.piece, bar_full + stage);
      if (++stage == a.stages) { stage = 0; ph ^= 1; }
      if (++sub == a.ppc) { sub = 0; if (++tile == a.wrap_tiles) tile = 0; }
    }
  } else if (lane == 0) {
    int stage = 0, ph = 0;
    for (int s = 0; s < a.steps; ++s) {
      mbar_wait(bar_full + stage, ph);
      mbar_arrive(bar_empty + stage);
      if (++stage == a.stages) { stage = 0; ph ^= 1; }
    }
  }
}

// Measurement aid: bulk STORE rate shared memory -> global (the stash path of the forward / dgrad kernels).  One thread
// per CTA issues `steps` cp.async.bulk stores of `piece` bytes from a constant smem buffer to consecutive addresses,
// keeping `depth` groups in flight; out[cta] = cycles.
__global__ void __launch_bounds__(32, 1) store_rate_kernel(uint8_t* dst, unsigned long long cta_stride, uint32_t piece,
                                                           int steps, int depth, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  for (uint32_t i = threadIdx.x; i < piece / 4; i += 32) reinterpret_cast<uint32_t*>(smem)[i] = i;
  fence_async_smem();
  __syncwarp();
  if (threadIdx.x == 0) {
    uint8_t* p = dst + (size_t)blockIdx.x * cta_stride;
    const long long t0 = clock64();
    for (int s = 0; s < steps; ++s) {
      bulk_s2g(p + (size_t)s * piece, smem, piece);
      bulk_commit();
      if (depth == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      else if (depth == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      else asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
    }
    bulk_wait_all0();
    out[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

// Measurement aid: cycles per tcgen05.mma (kind::f16, M=128, N=256, K=16, operands from shared memory) when ONE
// thread issues `reps` of them back to back into two alternating accumulators; out[cta] = total cycles.
__global__ void __launch_bounds__(64, 1) mma_rate_kernel(int reps, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 2 * 32768);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 1);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 2 * 32768 / 4; i += 64) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, 256, 0, 0);
    const uint64_t a_desc = make_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t b_desc = make_desc_sw128(smem_u32(smem + 32768), 16, 1024);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r)
      umma_ss(tmem_base + (r & 1) * 256, a_desc + 2 * (r & 3), b_desc + 2 * (r & 3), idesc, r >= 2);
    umma_commit(bar);
    mbar_wait(bar, 0);
    out[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;

// Launches `grid` CTAs; returns NRF_OK.  Total bytes moved = grid * steps * n_streams * piece.
extern "C" int nrf_dbg_stream(const void* base, unsigned long long cta_stride, unsigned long long stream_stride,
                              unsigned long long tile_stride, uint32_t piece, int n_streams, int ppc, int wrap_tiles,
                              int steps, int stages, int grid, nrf_stream_t stream) {
  NRF_REQUIRE(base && piece % 16 == 0 && n_streams >= 1 && n_streams <= 32 && stages >= 1 && ppc >= 1, "nrf_dbg_stream: bad args");
  const size_t smem_bytes = (size_t)stages * piece * n_streams + 16 * stages + 1024;
  NRF_REQUIRE(smem_bytes <= 227 * 1024, "nrf_dbg_stream: ring too large");
  NRF_CUDA(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  StreamArgs a{static_cast<const uint8_t*>(base), cta_stride, stream_stride, tile_stride, piece, n_streams, ppc,
               wrap_tiles, steps, stages};
  stream_kernel<<<grid, 64, smem_bytes, as_stream(stream)>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// Launches `grid` CTAs issuing `reps` MMAs each; out (grid x u64, device) receives the cycles each CTA needed.
extern "C" int nrf_dbg_mma_rate(int reps, int grid, void* out, nrf_stream_t stream) {
  NRF_REQUIRE(out && reps >= 2 && grid >= 1, "nrf_dbg_mma_rate: bad args");
  const size_t smem_bytes = 2 * 32768 + 64;
  NRF_CUDA(cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  mma_rate_kernel<<<grid, 64, smem_bytes, as_stream(stream)>>>(reps, static_cast<unsigned long long*>(out));
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// `grid` CTAs each store steps x piece bytes (piece <= 64 KB, multiple of 16) starting at dst + cta * cta_stride.
extern "C" int nrf_dbg_store_rate(void* dst, unsigned long long cta_stride, uint32_t piece, int steps, int depth, int grid,
                                  void* out, nrf_stream_t stream) {
  NRF_REQUIRE(dst && out && piece % 16 == 0 && piece >= 16 && piece <= 65536 && steps >= 1 && grid >= 1, "nrf_dbg_store_rate: bad args");
  NRF_CUDA(cudaFuncSetAttribute(store_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  store_rate_kernel<<<grid, 32, 65536, as_stream(stream)>>>(static_cast<uint8_t*>(dst), cta_stride, piece, steps, depth,
                                                           static_cast<unsigned long long*>(out));
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
