// Measurement aid (not on the product path): how fast can persistent CTAs pull data into shared memory
// with lane-parallel cp.async.bulk copies, as a function of piece size and address pattern?  Used to
// choose the stash layout (HBM streams of the wgrad kernel) and the weight-ring geometry (L2 streams of
// the forward / dgrad kernels).  DESIGN.md §4 quotes the numbers; tools/stream_bench.py drives it.
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"

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
  int replicas;        // > 0: CTA b reads replica (b % replicas) at base + replica * cta_stride (L2-resident copies of one blob)
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
    const uint8_t* src = a.base + (size_t)(a.replicas > 0 ? blockIdx.x % a.replicas : blockIdx.x) * a.cta_stride + (size_t)lane * a.stream_stride;
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
__global__ void __launch_bounds__(64, 1) mma_rate_kernel(int reps, int n_cols, unsigned long long* out) {
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
    const uint32_t idesc = make_idesc_bf16(128, n_cols, 0, 0);
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


// ---- MMA rate under shared-memory traffic ---------------------------------------------------------------
// One thread issues `reps` M128 x N256 x K16 MMAs back to back (mode bit 0: A operand from tensor memory instead of
// shared memory) while, with mode bit 1, another thread keeps 32 KB bulk copies from an L2-resident buffer landing in
// a 4 x 32 KB shared-memory ring (the forward kernel's weight stream) and, with mode bit 2, 16 warps write a 64 KB
// operand image with 16-byte shared stores over and over (the epilogue's stores).  out[cta * 3] = cycles of the MMAs,
// out[cta * 3 + 1] = 32 KB copies landed meanwhile, out[cta * 3 + 2] = 32 KB bulk stores issued meanwhile (mode bit 7).
__global__ void __launch_bounds__(64 + 16 * 32, 1) mma_contend_kernel(int reps, int mode, const uint8_t* src, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];   // [0,64K) A image / store target, [64K,96K) B chunk, [96K,224K) ring
  __shared__ uint64_t bar, bar_ring[4];
  __shared__ uint32_t s_tmem;
  __shared__ volatile int s_stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 4; ++i) mbar_init(bar_ring + i, (mode & 256) ? 32 : 1);
    fence_barrier_init();
    s_stop = 0;
  }
  for (int i = threadIdx.x; i < 3 * 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_async_smem();
  if (warp == 1) tmem_alloc(&s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = make_idesc_bf16(128, (mode & 8) ? 128 : ((mode & 16) ? 64 : 256), 0, 0);   // mode bit 3: N = 128, bit 4: N = 64
    const uint64_t a_desc = make_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t b_desc = make_desc_sw128(smem_u32(smem + 65536), 16, 1024);
    const long long t0 = clock64();
    if (mode & 1) {
      for (int r = 0; r < reps; ++r) umma_ts(tmem_base, tmem_base + 256 + 8 * (r & 15), b_desc + 2 * (r & 3), idesc, r >= 1);
    } else {
      for (int r = 0; r < reps; ++r) umma_ss(tmem_base, a_desc + 2 * (r & 3) + (uint64_t)((r >> 2) & 3) * (16384 >> 4), b_desc + 2 * (r & 3), idesc, r >= 1);
    }
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    out[blockIdx.x * 3] = (unsigned long long)(clock64() - t0);
    s_stop = 1;
  } else if (warp == 0 && (mode & 256)) {
    // the weight stream through cp.async (LDGSTS) instead of the TMA engine: every lane copies 64 x 16 bytes of a stage
    unsigned long long landed = 0;
    uint32_t ph = 0;
    const uint32_t ring_s = smem_u32(smem + 98304);
    auto issue = [&](int i, unsigned long long blk) {
      const uint8_t* g = src + (size_t)(blk & 31) * 32768 + lane * 16;
      const uint32_t d = ring_s + i * 32768 + lane * 16;
#pragma unroll 8
      for (int k = 0; k < 64; ++k) cp_async16(d + k * 512, g + k * 512);
      cp_async_mbar_arrive(bar_ring + i);
    };
    for (int i = 0; i < 4; ++i) issue(i, i);
    while (!s_stop) {
      for (int i = 0; i < 4; ++i) {
        mbar_wait(bar_ring + i, ph);
        ++landed;
        issue(i, landed + i);
      }
      ph ^= 1;
    }
    for (int i = 0; i < 4; ++i) mbar_wait(bar_ring + i, ph);
    if (lane == 0) out[blockIdx.x * 3 + 1] = landed;
  } else if (warp == 0 && lane == 0 && (mode & 2)) {
    unsigned long long landed = 0;
    uint32_t ph = 0;
    for (int i = 0; i < 4; ++i) {
      mbar_arrive_expect_tx(bar_ring + i, 32768);
      bulk_g2s(smem + 98304 + i * 32768, src + (size_t)i * 32768, 32768, bar_ring + i);
    }
    while (!s_stop) {
      for (int i = 0; i < 4; ++i) {
        mbar_wait(bar_ring + i, ph);
        ++landed;
        mbar_arrive_expect_tx(bar_ring + i, 32768);
        bulk_g2s(smem + 98304 + i * 32768, src + (size_t)((landed + i) & 31) * 32768, 32768, bar_ring + i);
      }
      ph ^= 1;
    }
    for (int i = 0; i < 4; ++i) mbar_wait(bar_ring + i, ph);
    out[blockIdx.x * 3 + 1] = landed;
  } else if (warp == 2 && lane == 0 && (mode & 128)) {
    // the stash store: 32 KB bulk stores that READ the A image (the tile the MMAs' A operand comes from in SS mode), two
    // in flight, into a 64 KB window of `src` per CTA (L2 resident)
    uint8_t* dst = const_cast<uint8_t*>(src) + (1 << 20) + (size_t)blockIdx.x * 65536;
    unsigned long long n = 0;
    while (!s_stop) {
      bulk_s2g(dst + (n & 1) * 32768, smem + (n & 1) * 32768, 32768);
      bulk_commit();
      bulk_wait_read1();
      ++n;
    }
    bulk_wait_all0();
    out[blockIdx.x * 3 + 2] = n;
  } else if (warp >= 3 && (mode & 64)) {
    // arithmetic-bound warps on every SM sub-partition (the epilogue's FADD / F2FP / LDS mix): does the issuing thread
    // still get its issue slots?
    float x0 = (float)threadIdx.x, x1 = 1.f, x2 = 2.f, x3 = 3.f;
    while (!s_stop) {
#pragma unroll
      for (int i = 0; i < 64; ++i) { x0 = fmaf(x0, 1.0001f, x1); x1 = fmaf(x1, 0.9999f, x2); x2 = fmaf(x2, 1.0002f, x3); x3 = fmaf(x3, 0.9998f, x0); }
    }
    if (x0 + x1 + x2 + x3 == 12345.f) out[0] = 1;
  } else if (warp >= 2 && (mode & 32)) {
    // the epilogue's tensor-memory traffic: every warp reads 2 x 16 accumulator columns and writes 2 x 8 operand columns
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const int sub = (warp - 2) >> 2;
    uint32_t va[16], vb[16], pk[8];
    while (!s_stop) {
      tmem_ld16(lane_base + 128 + sub * 16, va);
      tmem_ld16(lane_base + 192 + sub * 16, vb);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 8; ++i) pk[i] = va[2 * i] ^ vb[2 * i + 1];
      tmem_st8(lane_base + 384 + sub * 8, pk);
      tmem_st8(lane_base + 416 + sub * 8, pk);
      tmem_st_wait();
    }
  } else if (warp >= 2 && (mode & 4)) {
    const int t = threadIdx.x - 64;
    const uint32_t base = smem_u32(smem);
    uint32_t k = 0;
    while (!s_stop) {
#pragma unroll
      for (int c = 0; c < 8; ++c) st_shared_v4(base + ((uint32_t)(c * 512 + t) * 16u), k, k, k, k);   // 64 KB per pass
      ++k;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

// ---- TMEM read rate ------------------------------------------------------------------------------------
// `warps` (4, 8 or 16) warps each read their TMEM lane quarter with tcgen05.ld.32x32b.x16 (or .x32), `reps` passes over
// the 512 columns split among the warps of a quarter; optionally one more thread issues back-to-back MMAs into the same
// TMEM meanwhile (mode bit 1).  out[cta * 2] = cycles of the slowest reader warp, out[cta * 2 + 1] = MMAs retired.
__global__ void __launch_bounds__(64 + 16 * 32, 1) tmem_rate_kernel(int warps, int reps, int mode, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t s_tmem;
  __shared__ unsigned long long s_max;
  __shared__ volatile int s_stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); s_max = 0; s_stop = 0; }
  for (int i = threadIdx.x; i < 2 * 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_async_smem();
  if (warp == 1) tmem_alloc(&s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem;
  unsigned long long mmas = 0;
  if (warp == 1 && lane == 0 && (mode & 2)) {
    const uint32_t idesc = make_idesc_bf16(128, 256, 0, 0);
    const uint64_t a_desc = make_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t b_desc = make_desc_sw128(smem_u32(smem + 32768), 16, 1024);
    uint32_t ph = 0;
    while (!s_stop) {
      for (int r = 0; r < 16; ++r) umma_ss(tmem_base + (r & 1) * 256, a_desc + 2 * (r & 3), b_desc + 2 * (r & 3), idesc, 1);
      umma_commit(&bar);
      mbar_wait(&bar, ph);
      ph ^= 1;
      mmas += 16;
    }
    out[blockIdx.x * 2 + 1] = mmas;
  } else if (warp >= 2 && warp < 2 + warps) {
    const int q = warp & 3, sub = (warp - 2) >> 2, per_q = warps / 4;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t acc = 0;
    const long long t0 = clock64();
    if (mode & 12) {
      // shared-memory traffic instead of TMEM reads: mode bit 2 = STS.128 (512 B per warp instruction, conflict free),
      // bit 3 = LDS.128; 64 KB per pass and CTA, into the upper half of the operand buffers (the MMAs read garbage)
      uint32_t* base = reinterpret_cast<uint32_t*>(smem + 32768 + 16384) + (warp - 2) * 128 + lane * 4;   // 16 KB window
      for (int r = 0; r < reps; ++r) {
        for (int i = 0; i < 128 / warps * 1; ++i) {     // warps * (128 / warps) * 512 B = 64 KB per pass
          uint32_t* p = base + (i & 1) * (warps * 128);
          if (mode & 4) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(smem_u32(p)), "r"(acc) : "memory");
          } else {
            uint32_t a0, a1, a2, a3;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "r"(smem_u32(p)) : "memory");
            acc ^= a0 ^ a1 ^ a2 ^ a3;
          }
        }
      }
    } else
    for (int r = 0; r < reps; ++r) {
      if (mode & 1) {
        for (int c = sub * 32; c < 512; c += per_q * 32) {
          uint32_t v[32];
          tmem_ld32(t_lane + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc ^= v[j];
        }
      } else {
        for (int c = sub * 16; c < 512; c += per_q * 16) {
          uint32_t v[16];
          tmem_ld16(t_lane + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) acc ^= v[j];
        }
      }
    }
    const unsigned long long dt = (unsigned long long)(clock64() - t0);
    if (acc == 0x12345678u) out[blockIdx.x * 2 + 1] = acc;
    atomicMax(&s_max, dt);
  }
  // readers done -> stop the MMA thread
  if (warp >= 2 && warp < 2 + warps) {
    named_bar_sync(3, warps * 32);      // reader warps only
    if (threadIdx.x == 64) { s_stop = 1; out[blockIdx.x * 2] = s_max; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;
extern "C" int nrf_dbg_mma_rate_n(int reps, int n_cols, int grid, void* out, nrf_stream_t stream);

static int g_stream_replicas = 0;
// the next nrf_dbg_stream launches read replica (cta % replicas) at base + replica * cta_stride (0: cta * cta_stride)
extern "C" int nrf_dbg_stream_replicas(int replicas) { g_stream_replicas = replicas; return NRF_OK; }

// Launches `grid` CTAs; returns NRF_OK.  Total bytes moved = grid * steps * n_streams * piece.
extern "C" int nrf_dbg_stream(const void* base, unsigned long long cta_stride, unsigned long long stream_stride,
                              unsigned long long tile_stride, uint32_t piece, int n_streams, int ppc, int wrap_tiles,
                              int steps, int stages, int grid, nrf_stream_t stream) {
  NRF_REQUIRE(base && piece % 16 == 0 && n_streams >= 1 && n_streams <= 32 && stages >= 1 && ppc >= 1, "nrf_dbg_stream: bad args");
  const size_t smem_bytes = (size_t)stages * piece * n_streams + 16 * stages + 1024;
  NRF_REQUIRE(smem_bytes <= 227 * 1024, "nrf_dbg_stream: ring too large");
  NRF_CUDA(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  StreamArgs a{static_cast<const uint8_t*>(base), cta_stride, stream_stride, tile_stride, piece, n_streams, ppc,
               wrap_tiles, steps, stages, g_stream_replicas};
  stream_kernel<<<grid, 64, smem_bytes, as_stream(stream)>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// Launches `grid` CTAs issuing `reps` MMAs each; out (grid x u64, device) receives the cycles each CTA needed.
extern "C" int nrf_dbg_mma_rate(int reps, int grid, void* out, nrf_stream_t stream) { return nrf_dbg_mma_rate_n(reps, 256, grid, out, stream); }

// the same with N = n_cols (16..256, multiple of 16)
extern "C" int nrf_dbg_mma_rate_n(int reps, int n_cols, int grid, void* out, nrf_stream_t stream) {
  NRF_REQUIRE(out && reps >= 2 && grid >= 1 && n_cols >= 16 && n_cols <= 256 && n_cols % 16 == 0, "nrf_dbg_mma_rate: bad args");
  const size_t smem_bytes = 2 * 32768 + 64;
  NRF_CUDA(cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  mma_rate_kernel<<<grid, 64, smem_bytes, as_stream(stream)>>>(reps, n_cols, static_cast<unsigned long long*>(out));
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

// MMA rate under shared-memory traffic: see mma_contend_kernel.  src: >= 1 MB + grid x 64 KB device buffer; out: grid x 3 u64.
extern "C" int nrf_dbg_mma_contend(int reps, int mode, int grid, const void* src, void* out, nrf_stream_t stream) {
  NRF_REQUIRE(src && out && reps >= 2 && grid >= 1, "nrf_dbg_mma_contend: bad args");
  const size_t smem_bytes = 98304 + 4 * 32768;
  NRF_CUDA(cudaFuncSetAttribute(mma_contend_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  mma_contend_kernel<<<grid, 64 + 16 * 32, smem_bytes, as_stream(stream)>>>(reps, mode, static_cast<const uint8_t*>(src),
                                                                            static_cast<unsigned long long*>(out));
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// TMEM read-rate microbenchmark: see tmem_rate_kernel.  mode bit 0: .x32 loads (else .x16); bit 1: concurrent MMAs.
extern "C" int nrf_dbg_tmem_rate(int warps, int reps, int mode, int grid, void* out, nrf_stream_t stream) {
  NRF_REQUIRE(out && (warps == 4 || warps == 8 || warps == 16) && reps >= 1 && grid >= 1, "nrf_dbg_tmem_rate: bad args");
  const size_t smem_bytes = 2 * 32768;
  NRF_CUDA(cudaFuncSetAttribute(tmem_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  tmem_rate_kernel<<<grid, 64 + 16 * 32, smem_bytes, as_stream(stream)>>>(warps, reps, mode, static_cast<unsigned long long*>(out));
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
