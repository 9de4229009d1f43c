// bf16 tensor-core NeRF MLP, weight gradients ("wgrad"): dW_l = sum over sample rows of G_l^T X_l.
//
// Both operands are the [rows][64 x bf16] SWIZZLE_128B images written by the forward (X: layer
// inputs) and dgrad (G: pre-activation gradients) kernels.  Read as MN-major UMMA operands the
// reduction dimension K is the ROW index, so no transpose is ever materialised:
//   D[n_out (M=128 half), k_in (N<=256)] += A(MN-major, M = 128 cols of G) * B(MN-major, N = cols of X),
//   K = 16 rows per tcgen05.mma.
// A persistent CTA owns one (weight matrix, slab of tiles) unit at a time: the full fp32 dW block
// (2 halves x 256 columns = all 512 TMEM columns) stays in TMEM across the slab and is flushed
// once with fp32 atomics.  Stages are 32-row slices (4 KB per chunk) streamed by bulk copies
// through a 4-deep ring.  Bias gradients (column sums of G) and the tiny sigma / rgb head
// gradients run on the otherwise idle warps from the same shared-memory stages.
// HBM-bound: ~10 KB/sample of images are read exactly once.
#include <algorithm>
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kWgStages = 4;
constexpr int kWgSliceRows = 32;
constexpr uint32_t kWgSlice = kWgSliceRows * 128;         // 4 KB: one chunk's 32-row slice
constexpr int kWgASlots = 4, kWgBSlots = 8;
constexpr uint32_t kWgStageBytes = (kWgASlots + kWgBSlots) * kWgSlice;  // 48 KB
constexpr int kWgSideThreads = 256;
constexpr int kWgThreads = 64 + kWgSideThreads;
constexpr uint32_t kWgSmBars = kWgStages * kWgStageBytes;
constexpr uint32_t kWgSmTotal = kWgSmBars + 128;
constexpr int kWgMaxJobs = 12;

struct WgradJob {
  int a_slot, a_chunks;        // G image: first chunk slot in the G stash, 4 (N_out=256) or 2 (N_out=128) chunks
  int b_n, b_mma;              // X chunks loaded / of which take part in the MMA (rest: side work only)
  int b_slot[kWgBSlots];       // forward-stash chunk slot of each X chunk
  int b_valid[5], b_dst[5];    // per MMA X chunk: valid columns, destination column offset
  float* dW; int ld;           // destination (row = n_out)
  float* db;                   // bias gradient = column sums of G, or null
  int side, side_off;          // 0: none; 1: sigma head, h7 = X chunks side_off..+3; 2: rgb head, dmid = X chunks side_off..+1
  int first_unit, n_units, tiles_per_unit;
};

struct WgradArgs {
  WgradJob jobs[kWgMaxJobs];
  int n_jobs, total_units, num_tiles;
  int64_t n_rows;
  const uint8_t* xstash; size_t x_tile_bytes;
  const uint8_t* gstash; size_t g_tile_bytes;
  const float* d_sigma;        // (rows)
  const float* dpre;           // (tiles*128, 4)
  float* dW_sigma; float* db_sigma; float* dW_rgb; float* db_rgb;
};

__device__ __forceinline__ float ld_img(const uint8_t* slice, int r, int c) {  // element (row r, col c) of a 64-col slice
  return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(slice + sw128_off(r, c >> 3) + (c & 7) * 2));
}

__global__ void __launch_bounds__(kWgThreads, 1) mlp_wgrad_kernel(const WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgSmBars);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kWgStages;
  uint64_t* bar_acc = bars + 2 * kWgStages;        // accumulators complete (MMA -> epilogue)
  uint64_t* bar_drained = bars + 2 * kWgStages + 1;  // accumulators read out (epilogue -> MMA)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * kWgStages + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kWgStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1 + kWgSideThreads); }
    mbar_init(bar_acc, 1);
    mbar_init(bar_drained, kWgSideThreads);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  uint32_t it = 0;       // stage counter (same sequence in all roles)
  uint32_t unit_it = 0;  // units processed by this CTA
  for (int unit = blockIdx.x; unit < a.total_units; unit += gridDim.x, ++unit_it) {
    int j = 0;
    while (j + 1 < a.n_jobs && unit >= a.jobs[j + 1].first_unit) ++j;
    const WgradJob& job = a.jobs[j];
    const int t0 = (unit - job.first_unit) * job.tiles_per_unit;
    const int t1 = min(a.num_tiles, t0 + job.tiles_per_unit);
    const int n_stage = (t1 - t0) * 4;

    if (warp == 0) {
      if (lane == 0) {
        const uint32_t bytes = (uint32_t)(job.a_chunks + job.b_n) * kWgSlice;
        for (int s = 0; s < n_stage; ++s, ++it) {
          const int tile = t0 + (s >> 2), rs = s & 3;
          const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
          mbar_wait(bar_empty + stage, ph ^ 1);
          mbar_arrive_expect_tx(bar_full + stage, bytes);
          uint8_t* dst = smem + stage * kWgStageBytes;
          const uint8_t* g = a.gstash + (size_t)tile * a.g_tile_bytes + (size_t)job.a_slot * kActChunkBytes + rs * kWgSlice;
          for (int c = 0; c < job.a_chunks; ++c) bulk_g2s(dst + c * kWgSlice, g + (size_t)c * kActChunkBytes, kWgSlice, bar_full + stage);
          const uint8_t* x = a.xstash + (size_t)tile * a.x_tile_bytes + rs * kWgSlice;
          for (int c = 0; c < job.b_n; ++c)
            bulk_g2s(dst + (kWgASlots + c) * kWgSlice, x + (size_t)job.b_slot[c] * kActChunkBytes, kWgSlice, bar_full + stage);
        }
      } else {
        it += n_stage;
      }
    } else if (warp == 1) {
      if (lane == 0) {
        if (unit_it > 0) {  // previous unit's accumulators must be drained before they are overwritten
          mbar_wait(bar_drained, (unit_it - 1) & 1);
          tc_fence_after();
        }
        const int halves = job.a_chunks >> 1;
        const int n_main = job.b_mma > 4 ? 4 : job.b_mma;
        const uint32_t idesc_main = make_idesc_bf16(128, 64 * n_main, 1, 1);
        const uint32_t idesc_tail = make_idesc_bf16(128, 64, 1, 1);
        for (int s = 0; s < n_stage; ++s, ++it) {
          const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
          mbar_wait(bar_full + stage, ph);
          tc_fence_after();
          const uint32_t base = smem_u32(smem + stage * kWgStageBytes);
          for (int hh = 0; hh < halves; ++hh) {
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              const uint64_t ad = make_desc_sw128(base + hh * 2 * kWgSlice + ks * 2048, kWgSlice, 1024);
              const uint32_t acc = (s | ks) != 0;
              umma_ss(tmem_base + hh * 256, ad, make_desc_sw128(base + kWgASlots * kWgSlice + ks * 2048, kWgSlice, 1024),
                      idesc_main, acc);
              if (job.b_mma > 4)
                umma_ss(tmem_base + hh * 256 + 256, ad,
                        make_desc_sw128(base + (kWgASlots + 4) * kWgSlice + ks * 2048, kWgSlice, 1024), idesc_tail, acc);
            }
          }
          umma_commit(bar_empty + stage);
        }
        umma_commit(bar_acc);
      } else {
        it += n_stage;
      }
    } else {
      // ---------------- side warps: column sums + head gradients, then the accumulator flush ----------------
      const int t = threadIdx.x - 64;
      float bsum = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f, sb = 0.f;
      const bool do_bias = job.db != nullptr && t < 64 * job.a_chunks;
      for (int s = 0; s < n_stage; ++s, ++it) {
        const int tile = t0 + (s >> 2), rs = s & 3;
        const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
        mbar_wait(bar_full + stage, ph);
        const uint8_t* st = smem + stage * kWgStageBytes;
        if (do_bias) {
          const uint8_t* sl = st + (t >> 6) * kWgSlice;
#pragma unroll 8
          for (int r = 0; r < kWgSliceRows; ++r) bsum += ld_img(sl, r, t & 63);
        }
        if (job.side == 1) {
          const uint8_t* sl = st + (kWgASlots + job.side_off + (t >> 6)) * kWgSlice;
          const int64_t r0 = (int64_t)tile * kTileM + rs * kWgSliceRows;
#pragma unroll 8
          for (int r = 0; r < kWgSliceRows; ++r) {
            const float ds = (r0 + r < a.n_rows) ? __ldg(a.d_sigma + r0 + r) : 0.f;
            s0 += ds * ld_img(sl, r, t & 63);
            if (t == 0) sb += ds;
          }
        } else if (job.side == 2 && t < 128) {
          const uint8_t* sl = st + (kWgASlots + job.side_off + (t >> 6)) * kWgSlice;
          const float4* dp = reinterpret_cast<const float4*>(a.dpre) + (size_t)tile * kTileM + rs * kWgSliceRows;
#pragma unroll 8
          for (int r = 0; r < kWgSliceRows; ++r) {
            const float4 d = __ldg(dp + r);
            const float x = ld_img(sl, r, t & 63);
            s0 += d.x * x; s1 += d.y * x; s2 += d.z * x;
            if (t < 3) sb += (t == 0 ? d.x : (t == 1 ? d.y : d.z));
          }
        }
        mbar_arrive(bar_empty + stage);
      }
      if (do_bias) atomicAdd(job.db + t, bsum);
      if (job.side == 1) {
        atomicAdd(a.dW_sigma + t, s0);
        if (t == 0) atomicAdd(a.db_sigma, sb);
      } else if (job.side == 2 && t < 128) {
        atomicAdd(a.dW_rgb + t, s0); atomicAdd(a.dW_rgb + 128 + t, s1); atomicAdd(a.dW_rgb + 256 + t, s2);
        if (t < 3) atomicAdd(a.db_rgb + t, sb);
      }
      // accumulator flush: thread = TMEM lane (n_out within the half); the two warp groups split the columns
      mbar_wait(bar_acc, unit_it & 1);
      tc_fence_after();
      const int q = warp & 3, grp = (warp - 2) >> 2;
      const int halves = job.a_chunks >> 1;
      const int nblk = job.b_mma * 2;  // 32-column blocks
      for (int hh = 0; hh < halves; ++hh) {
        float* drow = job.dW + (size_t)(hh * 128 + q * 32 + lane) * job.ld;
        for (int blk = grp; blk < nblk; blk += 2) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + hh * 256 + blk * 32, v);
          tmem_ld_wait();
          const int ch = blk >> 1, cbase = (blk & 1) * 32;
          const int valid = job.b_valid[ch] - cbase;
          float* d = drow + job.b_dst[ch] + cbase;
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c < valid) atomicAdd(d + c, __uint_as_float(v[c]));
        }
      }
      tc_fence_before();
      mbar_arrive(bar_drained);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}


int launch_dgrad(bool full, const uint8_t* packed, const uint32_t* masks, const float* d_sigma, const float* d_rgb,
                 const float* rgb_out, int64_t n_rows, int num_tiles, uint8_t* gstash, float* dpre, cudaStream_t st);

static void add_job(WgradArgs& a, int a_slot, int a_chunks, std::initializer_list<int> b_slots, int b_mma,
                    std::initializer_list<int> valid, float* dW, int ld, float* db, int side, int side_off) {
  WgradJob& j = a.jobs[a.n_jobs++];
  j = WgradJob{};
  j.a_slot = a_slot; j.a_chunks = a_chunks; j.b_n = (int)b_slots.size(); j.b_mma = b_mma;
  int i = 0;
  for (int s : b_slots) j.b_slot[i++] = s;
  i = 0;
  for (int v : valid) { j.b_valid[i] = v; j.b_dst[i] = 64 * i; ++i; }
  j.dW = dW; j.ld = ld; j.db = db; j.side = side; j.side_off = side_off;
}

static int launch_wgrad(bool full, const nrf_mlp_grads* g, const uint8_t* xstash, const uint8_t* gstash,
                        const float* d_sigma, const float* dpre, int64_t n_rows, int num_tiles, cudaStream_t st) {
  WgradArgs a{};
  a.num_tiles = num_tiles; a.n_rows = n_rows;
  a.xstash = xstash; a.x_tile_bytes = stash_tile_bytes(full);
  a.gstash = gstash; a.g_tile_bytes = gstash_tile_bytes(full);
  a.d_sigma = d_sigma; a.dpre = dpre;
  a.dW_sigma = g->sigma_w; a.db_sigma = g->sigma_b; a.dW_rgb = g->rgb_w; a.db_rgb = g->rgb_b;
  auto H = [](int l) { return kStashH0 + 4 * l; };
  add_job(a, 0, 4, {kStashEnc}, 1, {63}, g->xyz_w[0], 63, g->xyz_b[0], 0, 0);
  for (int l = 1; l < 8; ++l) {
    const int x = H(l - 1);
    if (l == 4) {
      add_job(a, 16, 4, {x, x + 1, x + 2, x + 3}, 4, {64, 64, 64, 64}, g->xyz_w[4] + 63, 319, g->xyz_b[4], 0, 0);
      add_job(a, 16, 4, {kStashEnc}, 1, {63}, g->xyz_w[4], 319, nullptr, 0, 0);
    } else if (l == 7 && !full) {
      const int h7 = H(7);
      add_job(a, 28, 4, {x, x + 1, x + 2, x + 3, h7, h7 + 1, h7 + 2, h7 + 3}, 4, {64, 64, 64, 64}, g->xyz_w[7], 256,
              g->xyz_b[7], 1, 4);
    } else {
      add_job(a, 4 * l, 4, {x, x + 1, x + 2, x + 3}, 4, {64, 64, 64, 64}, g->xyz_w[l], 256, g->xyz_b[l], 0, 0);
    }
  }
  if (full) {
    const int h7 = H(7), f = kStashFeat, m = kStashMid;
    add_job(a, kGstashG8, 4, {h7, h7 + 1, h7 + 2, h7 + 3}, 4, {64, 64, 64, 64}, g->final_w, 256, g->final_b, 1, 0);
    add_job(a, kGstashG9, 2, {f, f + 1, f + 2, f + 3, kStashDir, m, m + 1}, 5, {64, 64, 64, 64, 27}, g->dir_w, 283,
            g->dir_b, 2, 5);
  }
  // slabs: about two units per SM in total, split between jobs in proportion to the bytes they stream
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  double total_cost = 0;
  for (int i = 0; i < a.n_jobs; ++i) total_cost += a.jobs[i].a_chunks + a.jobs[i].b_n;
  int unit = 0;
  for (int i = 0; i < a.n_jobs; ++i) {
    WgradJob& j = a.jobs[i];
    int want = (int)(2.0 * sms * (j.a_chunks + j.b_n) / total_cost + 0.5);
    want = std::max(1, std::min(want, num_tiles));
    j.tiles_per_unit = (num_tiles + want - 1) / want;
    j.n_units = (num_tiles + j.tiles_per_unit - 1) / j.tiles_per_unit;
    j.first_unit = unit;
    unit += j.n_units;
  }
  a.total_units = unit;
  static bool configured = false;
  const int smem_bytes = (int)kWgSmTotal + 1024;
  if (!configured) {
    NRF_CUDA(cudaFuncSetAttribute(mlp_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  mlp_wgrad_kernel<<<std::min(a.total_units, sms), kWgThreads, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;

extern "C" size_t nrf_mlp_tc_bwd_workspace_bytes(int64_t rows, int sigma_only) {
  if (rows <= 0) return 0;
  const size_t tiles = (size_t)((rows + kTileM - 1) / kTileM);
  return tiles * (gstash_tile_bytes(!sigma_only) + (size_t)kTileM * 16) + 1024;
}

extern "C" int nrf_mlp_tc_bwd(const void* packed, const nrf_mlp_grads* g, const float* rays, const float* z,
                              int64_t n_rays, int n_samples, int sigma_only, const float* d_sigma, const float* d_rgb,
                              const float* rgb_out, const void* stash, size_t stash_bytes, void* workspace,
                              size_t workspace_bytes, nrf_stream_t stream) {
  (void)rays; (void)z;
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_mlp_tc_bwd: bad sizes");
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(packed && g && d_sigma && stash && workspace, "nrf_mlp_tc_bwd: null pointer");
  NRF_REQUIRE(sigma_only || (d_rgb && rgb_out), "nrf_mlp_tc_bwd: full mode needs d_rgb and rgb_out");
  const bool full = !sigma_only;
  const int64_t rows = n_rays * n_samples;
  const int tiles = (int)((rows + kTileM - 1) / kTileM);
  if (stash_bytes < nrf_mlp_tc_stash_bytes(rows, sigma_only) ||
      workspace_bytes < nrf_mlp_tc_bwd_workspace_bytes(rows, sigma_only)) {
    set_error("nrf_mlp_tc_bwd: stash or workspace too small");
    return NRF_ERR_WORKSPACE;
  }
  const uint8_t* xstash = static_cast<const uint8_t*>(stash);
  const uint32_t* masks = reinterpret_cast<const uint32_t*>(xstash + (size_t)tiles * stash_tile_bytes(full));
  uint8_t* gst = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~uintptr_t(1023));
  float* dpre = reinterpret_cast<float*>(gst + (size_t)tiles * gstash_tile_bytes(full));
  cudaStream_t st = as_stream(stream);
  int rc = launch_dgrad(full, static_cast<const uint8_t*>(packed), masks, d_sigma, d_rgb, rgb_out, rows, tiles, gst,
                        dpre, st);
  if (rc) return rc;
  return launch_wgrad(full, g, xstash, gst, d_sigma, dpre, rows, tiles, st);
}
