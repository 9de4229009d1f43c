// bf16 tensor-core NeRF MLP, weight gradients ("wgrad"): dW_l = sum over sample rows of G_l^T X_l.
//
// Both operands are the [rows][64 x bf16] SWIZZLE_128B images written by the forward (X: layer
// inputs) and dgrad (G: pre-activation gradients) kernels.  Read as MN-major UMMA operands the
// reduction dimension K is the ROW index, so no transpose is ever materialised:
//   D[n_out (M=128 half), k_in (N<=256)] += A(MN-major, M = 128 cols of G) * B(MN-major, N = cols of X),
//   K = 16 rows per tcgen05.mma.
// Work is cut into units = (weight matrix, slab of tiles); persistent CTAs pull units from a global
// atomic counter (units of different matrices stream different byte counts, so static assignment
// leaves SMs idle).  The full fp32 dW block (2 halves x 256 columns = all 512 TMEM columns) stays in
// TMEM across a unit and is flushed once with fp32 reductions (red.global.add, vectorised when the
// destination row pitch allows).  Stages are 32-row slices (4 KB per chunk) streamed through a
// 6-deep ring; the copies of one stage are issued by different lanes of the producer warp (a single
// thread issuing them serially was the bottleneck of the first version: profiles/r01_*baseline*).
// Bias gradients (column sums of G) and the tiny sigma / rgb head gradients run on the otherwise
// idle warps from the same shared-memory stages.
// HBM-bound: ~10 KB/sample of images are read exactly once.
// fp32 atomics make the summation order (not the values beyond rounding) run-dependent.
#include <algorithm>
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kWgStages = 6;
constexpr int kWgSliceRows = 32;
constexpr uint32_t kWgSlice = kWgSliceRows * 128;         // 4 KB: one chunk's 32-row slice
constexpr int kWgASlots = 4, kWgBSlots = 5;               // G slices, X slices per stage (max)
constexpr uint32_t kWgStageBytes = (kWgASlots + kWgBSlots) * kWgSlice;  // 36 KB
constexpr int kWgSideWarps = 8;
constexpr int kWgSideThreads = kWgSideWarps * 32;
constexpr int kWgThreads = 64 + kWgSideThreads;
constexpr uint32_t kWgSmBars = kWgStages * kWgStageBytes;
constexpr uint32_t kWgSmTotal = kWgSmBars + 256;
constexpr int kWgMaxJobs = 14;

struct WgradJob {
  int a_slot, a_chunks;        // G image: first chunk slot in the G stash; 4 (N_out=256), 2 (N_out=128) or 0 (side work only)
  int b_n, b_mma;              // X chunks loaded / of which take part in the MMA (rest: side work only)
  int b_slot[kWgBSlots];       // forward-stash chunk slot of each X chunk
  int b_valid[kWgBSlots];      // per MMA X chunk: valid columns (destination column offset = 64 * index)
  float* dW; int ld;           // destination (row = n_out)
  float* db;                   // bias gradient = column sums of G, or null
  int side, side_off;          // 0: none; 1: sigma head, h7 = X slices side_off..+3; 2: rgb head, dmid = X slices side_off..+1
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
  int* unit_counter;           // zeroed before launch
};

// two adjacent bf16 columns (2*cp, 2*cp+1) of row r of a 64-column SWIZZLE_128B slice
__device__ __forceinline__ float2 ld_img2(const uint8_t* slice, int r, int cp) {
  const uint32_t w = *reinterpret_cast<const uint32_t*>(slice + sw128_off(r, cp >> 2) + (cp & 3) * 4);
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(kWgThreads, 1) mlp_wgrad_kernel(const __grid_constant__ WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgSmBars);
  uint64_t* bar_full = bars;                          // [kWgStages] stage landed (producer -> MMA, side)
  uint64_t* bar_empty = bars + kWgStages;             // [kWgStages] stage consumed (MMA commit + side warps -> producer)
  uint64_t* bar_acc = bars + 2 * kWgStages;           // accumulators complete (MMA -> flush)
  uint64_t* bar_drained = bars + 2 * kWgStages + 1;   // accumulators read out (flush -> MMA)
  uint64_t* bar_uq_full = bars + 2 * kWgStages + 2;   // [2] unit id published (producer -> MMA, side)
  uint64_t* bar_uq_empty = bars + 2 * kWgStages + 4;  // [2] unit id consumed
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * kWgStages + 6);
  volatile int* s_unit = reinterpret_cast<volatile int*>(s_tmem + 1);  // [2]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kWgStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1 + kWgSideWarps); }
    mbar_init(bar_acc, 1);
    mbar_init(bar_drained, kWgSideWarps);
    for (int s = 0; s < 2; ++s) { mbar_init(bar_uq_full + s, 1); mbar_init(bar_uq_empty + s, 1 + kWgSideWarps); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  uint32_t it = 0;  // stage counter (same sequence in all roles)
  for (uint32_t uq = 0;; ++uq) {
    const uint32_t us = uq & 1, uph = (uq >> 1) & 1;
    int unit;
    if (warp == 0) {
      // ---------------- scheduler: claim the next unit and publish it ----------------
      unit = 0;
      if (lane == 0) {
        mbar_wait(bar_uq_empty + us, uph ^ 1);
        unit = atomicAdd(a.unit_counter, 1);
        if (unit >= a.total_units) unit = -1;
        s_unit[us] = unit;
        mbar_arrive(bar_uq_full + us);
      }
      unit = __shfl_sync(0xffffffffu, unit, 0);
    } else {
      mbar_wait(bar_uq_full + us, uph);
      unit = s_unit[us];
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_uq_empty + us);
    }
    if (unit < 0) break;
    int j = 0;
    while (j + 1 < a.n_jobs && unit >= a.jobs[j + 1].first_unit) ++j;
    const WgradJob& job = a.jobs[j];
    const int t0 = (unit - job.first_unit) * job.tiles_per_unit;
    const int t1 = min(a.num_tiles, t0 + job.tiles_per_unit);
    const int n_stage = (t1 - t0) * 4;
    const int a_chunks = job.a_chunks, b_n = job.b_n;

    if (warp == 0) {
      // ---------------- producer: lane c copies slice c of every stage ----------------
      const int n_copies = a_chunks + b_n;
      const uint8_t* src = nullptr;
      size_t tile_stride = 0;
      uint32_t dst_off = 0;
      if (lane < a_chunks) {
        src = a.gstash + (size_t)(job.a_slot + lane) * kActChunkBytes;
        tile_stride = a.g_tile_bytes;
        dst_off = lane * kWgSlice;
      } else if (lane < n_copies) {
        src = a.xstash + (size_t)job.b_slot[lane - a_chunks] * kActChunkBytes;
        tile_stride = a.x_tile_bytes;
        dst_off = (kWgASlots + lane - a_chunks) * kWgSlice;
      }
      src += (size_t)t0 * tile_stride;
      const uint32_t bytes = (uint32_t)n_copies * kWgSlice;
      for (int s = 0; s < n_stage; ++s, ++it) {
        const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
        if (lane == 0) {
          mbar_wait(bar_empty + stage, ph ^ 1);
          mbar_arrive_expect_tx(bar_full + stage, bytes);
        }
        __syncwarp();
        if (lane < n_copies)
          bulk_g2s(smem + stage * kWgStageBytes + dst_off, src + (size_t)(s >> 2) * tile_stride + (s & 3) * kWgSlice,
                   kWgSlice, bar_full + stage);
      }
    } else if (warp == 1) {
      // ---------------- UMMA issuer ----------------
      if (lane == 0) {
        if (uq > 0) {  // previous unit's accumulators must be drained before they are overwritten
          mbar_wait(bar_drained, (uq - 1) & 1);
          tc_fence_after();
        }
        const int halves = a_chunks >> 1;
        const int n_main = job.b_mma > 4 ? 4 : job.b_mma;
        const uint32_t idesc_main = make_idesc_bf16(128, 64 * n_main, 1, 1);
        const uint32_t idesc_tail = make_idesc_bf16(128, 64, 1, 1);
        const bool tail = job.b_mma > 4;
        for (int s = 0; s < n_stage; ++s, ++it) {
          const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
          mbar_wait(bar_full + stage, ph);
          if (halves == 0) {
            mbar_arrive(bar_empty + stage);
            continue;
          }
          tc_fence_after();
          const uint32_t base = smem_u32(smem + stage * kWgStageBytes);
          for (int hh = 0; hh < halves; ++hh) {
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              const uint64_t ad = make_desc_sw128(base + hh * 2 * kWgSlice + ks * 2048, kWgSlice, 1024);
              const uint32_t acc = (s | ks) != 0;
              umma_ss(tmem_base + hh * 256, ad, make_desc_sw128(base + kWgASlots * kWgSlice + ks * 2048, kWgSlice, 1024),
                      idesc_main, acc);
              if (tail)
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
      const int cp = t & 31;                 // column pair within a 64-column slice
      float b0 = 0.f, b1 = 0.f;              // bias partial sums (columns 2cp, 2cp+1 of G slice `bch`)
      float h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f, h4 = 0.f, h5 = 0.f, hb0 = 0.f, hb1 = 0.f, hb2 = 0.f;
      const int bch = (t >> 5) & 3, brh = t >> 7;            // bias / sigma head: slice, 16-row half
      const int rch = (t >> 5) & 1, rrq = t >> 6;            // rgb head: slice, 8-row quarter
      const bool do_bias = job.db != nullptr && bch < a_chunks;
      const int side = job.side;
      for (int s = 0; s < n_stage; ++s, ++it) {
        const int tile = t0 + (s >> 2), rs = s & 3;
        const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
        mbar_wait(bar_full + stage, ph);
        const uint8_t* st = smem + stage * kWgStageBytes;
        if (do_bias) {
          const uint8_t* sl = st + bch * kWgSlice;
#pragma unroll
          for (int r = 0; r < 16; ++r) {
            const float2 g = ld_img2(sl, brh * 16 + r, cp);
            b0 += g.x; b1 += g.y;
          }
        }
        if (side == 1) {
          const uint8_t* sl = st + (kWgASlots + job.side_off + bch) * kWgSlice;
          const int64_t r0 = (int64_t)tile * kTileM + rs * kWgSliceRows + brh * 16;
#pragma unroll
          for (int r = 0; r < 16; ++r) {
            const float ds = (r0 + r < a.n_rows) ? __ldg(a.d_sigma + r0 + r) : 0.f;
            const float2 x = ld_img2(sl, brh * 16 + r, cp);
            h0 += ds * x.x; h1 += ds * x.y;
            hb0 += ds;
          }
        } else if (side == 2) {
          const uint8_t* sl = st + (kWgASlots + job.side_off + rch) * kWgSlice;
          const float4* dp = reinterpret_cast<const float4*>(a.dpre) + (size_t)tile * kTileM + rs * kWgSliceRows + rrq * 8;
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            const float4 d = __ldg(dp + r);
            const float2 x = ld_img2(sl, rrq * 8 + r, cp);
            h0 += d.x * x.x; h1 += d.x * x.y;
            h2 += d.y * x.x; h3 += d.y * x.y;
            h4 += d.z * x.x; h5 += d.z * x.y;
            hb0 += d.x; hb1 += d.y; hb2 += d.z;
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + stage);
      }
      if (do_bias) {
        atomicAdd(job.db + bch * 64 + 2 * cp, b0);
        atomicAdd(job.db + bch * 64 + 2 * cp + 1, b1);
      }
      if (side == 1) {
        atomicAdd(a.dW_sigma + bch * 64 + 2 * cp, h0);
        atomicAdd(a.dW_sigma + bch * 64 + 2 * cp + 1, h1);
        if ((t & 127) == 0) atomicAdd(a.db_sigma, hb0);
      } else if (side == 2) {
        const int k = rch * 64 + 2 * cp;
        atomicAdd(a.dW_rgb + k, h0); atomicAdd(a.dW_rgb + k + 1, h1);
        atomicAdd(a.dW_rgb + 128 + k, h2); atomicAdd(a.dW_rgb + 128 + k + 1, h3);
        atomicAdd(a.dW_rgb + 256 + k, h4); atomicAdd(a.dW_rgb + 256 + k + 1, h5);
        if ((t & 63) == 0) { atomicAdd(a.db_rgb, hb0); atomicAdd(a.db_rgb + 1, hb1); atomicAdd(a.db_rgb + 2, hb2); }
      }
      // accumulator flush: thread = TMEM lane (n_out within the half); the two warp groups split the column blocks
      mbar_wait(bar_acc, uq & 1);
      tc_fence_after();
      const int q = warp & 3, grp = (warp - 2) >> 2;
      const int halves = a_chunks >> 1;
      const int nblk = job.b_mma * 2;  // 32-column blocks
      const bool vec_ok = (job.ld & 3) == 0 && (reinterpret_cast<uintptr_t>(job.dW) & 15) == 0;
      for (int hh = 0; hh < halves; ++hh) {
        float* drow = job.dW + (size_t)(hh * 128 + q * 32 + lane) * job.ld;
        for (int blk = grp; blk < nblk; blk += 2) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + hh * 256 + blk * 32, v);
          tmem_ld_wait();
          const int ch = blk >> 1, cbase = (blk & 1) * 32;
          const int valid = job.b_valid[ch] - cbase;
          float* d = drow + ch * 64 + cbase;
          if (vec_ok && valid >= 32) {
#pragma unroll
            for (int c = 0; c < 32; c += 4)
              red_add_v4(d + c, __uint_as_float(v[c]), __uint_as_float(v[c + 1]), __uint_as_float(v[c + 2]),
                         __uint_as_float(v[c + 3]));
          } else {
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (c < valid) atomicAdd(d + c, __uint_as_float(v[c]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_drained);
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
  for (int v : valid) j.b_valid[i++] = v;
  j.dW = dW; j.ld = ld; j.db = db; j.side = side; j.side_off = side_off;
}

static int launch_wgrad(bool full, const nrf_mlp_grads* g, const uint8_t* xstash, const uint8_t* gstash,
                        const float* d_sigma, const float* dpre, int64_t n_rows, int num_tiles, int* unit_counter,
                        cudaStream_t st) {
  WgradArgs a{};
  a.num_tiles = num_tiles; a.n_rows = n_rows;
  a.xstash = xstash; a.x_tile_bytes = stash_tile_bytes(full);
  a.gstash = gstash; a.g_tile_bytes = gstash_tile_bytes(full);
  a.d_sigma = d_sigma; a.dpre = dpre;
  a.dW_sigma = g->sigma_w; a.db_sigma = g->sigma_b; a.dW_rgb = g->rgb_w; a.db_rgb = g->rgb_b;
  a.unit_counter = unit_counter;
  auto H = [](int l) { return kStashH0 + 4 * l; };
  add_job(a, 0, 4, {kStashEnc}, 1, {63}, g->xyz_w[0], 63, g->xyz_b[0], 0, 0);
  for (int l = 1; l < 8; ++l) {
    const int x = H(l - 1);
    if (l == 4) {
      add_job(a, 16, 4, {x, x + 1, x + 2, x + 3}, 4, {64, 64, 64, 64}, g->xyz_w[4] + 63, 319, g->xyz_b[4], 0, 0);
      add_job(a, 16, 4, {kStashEnc}, 1, {63}, g->xyz_w[4], 319, nullptr, 0, 0);
    } else {
      add_job(a, 4 * l, 4, {x, x + 1, x + 2, x + 3}, 4, {64, 64, 64, 64}, g->xyz_w[l], 256, g->xyz_b[l], 0, 0);
    }
  }
  const int h7 = H(7);
  if (full) {
    const int f = kStashFeat, m = kStashMid;
    add_job(a, kGstashG8, 4, {h7, h7 + 1, h7 + 2, h7 + 3}, 4, {64, 64, 64, 64}, g->final_w, 256, g->final_b, 1, 0);
    add_job(a, kGstashG9, 2, {f, f + 1, f + 2, f + 3, kStashDir}, 5, {64, 64, 64, 64, 27}, g->dir_w, 283, g->dir_b, 0, 0);
    add_job(a, 0, 0, {m, m + 1}, 0, {}, nullptr, 0, nullptr, 2, 0);          // rgb head (side work only)
  } else {
    add_job(a, 0, 0, {h7, h7 + 1, h7 + 2, h7 + 3}, 0, {}, nullptr, 0, nullptr, 1, 0);  // sigma head (side work only)
  }
  // units of about equal byte count, ~6 per SM in total; CTAs claim them dynamically
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  double total_cost = 0;
  for (int i = 0; i < a.n_jobs; ++i) total_cost += a.jobs[i].a_chunks + a.jobs[i].b_n;
  const double unit_cost = std::max(1.0, (double)num_tiles * total_cost / (6.0 * sms));
  int unit = 0;
  for (int i = 0; i < a.n_jobs; ++i) {
    WgradJob& j = a.jobs[i];
    int tpu = (int)(unit_cost / (j.a_chunks + j.b_n) + 0.5);
    tpu = std::max(1, std::min(tpu, num_tiles));
    j.tiles_per_unit = tpu;
    j.n_units = (num_tiles + tpu - 1) / tpu;
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
  NRF_CUDA(cudaMemsetAsync(unit_counter, 0, sizeof(int), st));
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
  return tiles * (gstash_tile_bytes(!sigma_only) + (size_t)kTileM * 16) + 2048;
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
  int* unit_counter = reinterpret_cast<int*>(dpre + (size_t)tiles * kTileM * 4);
  cudaStream_t st = as_stream(stream);
  int rc = launch_dgrad(full, static_cast<const uint8_t*>(packed), masks, d_sigma, d_rgb, rgb_out, rows, tiles, gst,
                        dpre, st);
  if (rc) return rc;
  return launch_wgrad(full, g, xstash, gst, d_sigma, dpre, rows, tiles, unit_counter, st);
}
