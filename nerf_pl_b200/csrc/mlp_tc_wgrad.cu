// bf16 tensor-core NeRF MLP, weight gradients ("wgrad"): dW_l = sum over sample rows of G_l^T X_l.
//
// Both operands are the [rows][64 x bf16] SWIZZLE_128B images written by the forward (X: layer
// inputs) and dgrad (G: pre-activation gradients) kernels.  Read as MN-major UMMA operands the
// reduction dimension K is the ROW index, so no transpose is ever materialised:
//   D[n_out (M=128 half), k_in (N<=256)] += A(MN-major, M = 128 cols of G) * B(MN-major, N = cols of X),
//   K = 16 rows per tcgen05.mma.
// Work = jobs (one weight matrix each) cut into granules of a few tiles that persistent CTAs claim through
// per-job atomic counters.  A CTA stays on one job while that job has granules left: the full fp32 dW block
// (2 halves x 256 columns = all 512 TMEM columns) keeps accumulating in TMEM and is flushed with fp32
// reductions (red.global.add, vectorised when the destination row pitch allows) only when the CTA moves to
// another job -- the flush of 64 K atomics was measured at ~0.3 ms per step when it ran 8 times per SM.
// CTAs start spread over the jobs in proportion to the jobs' byte counts and help out elsewhere once their job
// is exhausted, so the tail is one granule.  Stages are 32-row slices (4 KB per chunk) streamed through a
// 6-deep ring; the copies of one stage are issued by different lanes of the producer warp (a single
// thread issuing them serially was the bottleneck of the first version: profiles/r01_*baseline*).
// Bias gradients (column sums of G) and the tiny sigma / rgb head gradients run on the otherwise
// idle warps from the same shared-memory stages.
// HBM-bound: ~10 KB/sample of images are read exactly once.
// fp32 atomics make the summation order (not the values beyond rounding) run-dependent.
#include <algorithm>
#include <cstdlib>
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
constexpr int kWgSideArrivals = kWgSideWarps / 2;   // side warps that hold a stage: those of SM sub-partitions 2 and 3 (see below)

#ifndef NRF_FWD_DEBUG
#define NRF_FWD_DEBUG 0
#endif
// experiment flags (tools/wgrad_probe.py) exist in debug builds only; the product kernel carries none of the branches
#define NRF_WG_DBG(bits) (NRF_FWD_DEBUG && (a.dbg_flags & (bits)))

struct WgradJob {
  int a_slot, a_chunks;        // G image: first chunk slot in the G stash; 4 (N_out=256), 2 (N_out=128) or 0 (side work only)
  int b_n, b_mma;              // X chunks loaded / of which take part in the MMA (rest: side work only)
  int b_slot[kWgBSlots];       // forward-stash chunk slot of each X chunk
  int b_img[kWgBSlots];        // first slot of the image that chunk belongs to
  int b_imgn[kWgBSlots];       // chunks in that image
  int b_valid[kWgBSlots];      // per MMA X chunk: valid columns (destination column offset = 64 * index)
  float* dW; int ld;           // destination (row = n_out)
  float* db;                   // bias gradient = column sums of G, or null
  int side, side_off;          // 0: none; 1: sigma head, h7 = X slices side_off..+3; 2: rgb head, dmid = X slices side_off..+1
  int n_gran, tiles_per_gran;  // work granules of this job (claimed dynamically through counters[job])
  int first_cta;               // CTAs [first_cta, next job's first_cta) start on this job
};

struct WgradArgs {
  WgradJob jobs[kWgMaxJobs];
  int n_jobs, num_tiles;
  int64_t n_rows;
  const uint8_t* xstash;
  const uint8_t* gstash;
  const float* d_sigma;        // (rows)
  const float* dpre;           // (tiles*128, 4)
  float* dW_sigma; float* db_sigma; float* dW_rgb; float* db_rgb;
  int* counters;               // [n_jobs] next granule of each job, zeroed before launch
  int dbg_flags;               // experiments only (NRF_DBG_FLAGS): 1 = skip side work, 2 = skip MMAs, 4 = side loads without
                               // arithmetic, 64 = skip the flush, 8192 = bias sums only, 16384 = head gradients only
};

// Head jobs (side 1: sigma head, side 2: rgb head) need per-row upstream gradients: d_sigma (rows) floats, dpre (rows, 4)
// floats.  They travel in X slot 4 of the stage, which no head job uses for an image (b_n <= 4 there).
constexpr uint32_t kWgSideVecOff = (kWgASlots + 4) * kWgSlice;
__host__ __device__ inline uint32_t side_vec_bytes(int side) {
  return side == 1 ? kWgSliceRows * 4u : (side == 2 ? kWgSliceRows * 16u : 0u);
}
// d_sigma holds n_rows floats: the slice that straddles the end is read element-wise instead (dpre is padded to whole tiles)
__host__ __device__ inline bool side_vec_in_range(int side, int64_t row0, int64_t n_rows) {
  return side == 2 || row0 + kWgSliceRows <= n_rows;
}

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
  volatile int* s_unit = reinterpret_cast<volatile int*>(s_tmem + 1);  // [2] unit queue, [2] = scratch
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kWgStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1 + kWgSideArrivals); }
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

  // Work distribution: a job (one layer's weight gradient) is cut into granules of a few tiles, claimed through a
  // per-job counter.  A CTA stays on ONE job for as long as that job has granules left, so its TMEM accumulators keep
  // summing and are flushed (64 K atomic adds) only when it moves on to another job; CTAs start spread over the jobs
  // in proportion to the jobs' cost and help out elsewhere (cyclic scan) once their own job is exhausted.
  uint32_t it = 0;     // stage counter (same sequence in all roles)
  uint32_t grp = 0;    // accumulator groups flushed so far (UMMA issuer and side warps count alike)
  int prev_job = -1;   // job of the previous unit
  int sched_job = 0;   // scheduler: job this CTA currently draws from
  if (warp == 0) {
    while (sched_job + 1 < a.n_jobs && (int)blockIdx.x >= a.jobs[sched_job + 1].first_cta) ++sched_job;
  }
  // side-warp register accumulators (column sums of G; head gradients); live across the units of one job
  float bs[8], hs[8], hg[8], hbv[8], hb0 = 0.f, hb1 = 0.f, hb2 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) { bs[e] = 0.f; hs[e] = 0.f; hg[e] = 0.f; hbv[e] = 0.f; }
  bool fresh = true;   // UMMA issuer: the next stage starts a new accumulator group

  // side warps: end-of-group reduction of the register accumulators + accumulator flush of job `pj`
  auto flush_group = [&](const WgradJob& pj) {
    const int t = threadIdx.x - 64;
    const int cb = t & 7, rql = (t >> 3) & 3, sw = t >> 5;
    const bool active = (warp & 2) != 0;                // side work runs on the warps of SM sub-partitions 2 and 3 only
    const int ai = (sw & 1) | ((sw >> 2) << 1);         // 0..3 over the active warps
    const int k4 = ai, k2 = ai & 1;
    const int a_chunks = pj.a_chunks;
    const bool do_bias = active && pj.db != nullptr && !NRF_WG_DBG(1 | 16384);
    const int side = (!active || NRF_WG_DBG(1 | 8192)) ? 0 : pj.side;
    // reduce over the four row groups that live in one warp (lane bits 3, 4), then one atomic per column and warp
    auto warp_rows_sum = [&](float v) {
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      return v;
    };
    if (do_bias) {   // warp-uniform (k4 depends on the warp only)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float v = warp_rows_sum(bs[e]);
        if (rql == 0 && k4 < a_chunks) atomicAdd(pj.db + k4 * 64 + cb * 8 + e, v);
      }
    }
    if (side == 1) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float v = warp_rows_sum(hs[e]);
        if (rql == 0) atomicAdd(a.dW_sigma + k4 * 64 + cb * 8 + e, v);
      }
      const float vb = warp_rows_sum(hb0);
      if (k4 == 0 && lane == 0) atomicAdd(a.db_sigma, vb);   // cb = 0 lanes of slice-0 warps cover every row once
    } else if (side == 2) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float v0 = warp_rows_sum(hs[e]), v1 = warp_rows_sum(hg[e]), v2 = warp_rows_sum(hbv[e]);
        if (rql == 0) {
          atomicAdd(a.dW_rgb + k2 * 64 + cb * 8 + e, v0);
          atomicAdd(a.dW_rgb + 128 + k2 * 64 + cb * 8 + e, v1);
          atomicAdd(a.dW_rgb + 256 + k2 * 64 + cb * 8 + e, v2);
        }
      }
      const float w0 = warp_rows_sum(hb0), w1 = warp_rows_sum(hb1), w2 = warp_rows_sum(hb2);
      if (k2 == 0 && lane == 0) { atomicAdd(a.db_rgb, w0); atomicAdd(a.db_rgb + 1, w1); atomicAdd(a.db_rgb + 2, w2); }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) { bs[e] = 0.f; hs[e] = 0.f; hg[e] = 0.f; hbv[e] = 0.f; }
    hb0 = hb1 = hb2 = 0.f;
    // accumulator flush: thread = TMEM lane (n_out within the half); the two warp groups split the column blocks
    mbar_wait(bar_acc, grp & 1);
    tc_fence_after();
    const int q = warp & 3, wg = (warp - 2) >> 2;
    const int halves = NRF_WG_DBG(64) ? 0 : a_chunks >> 1;
    const int nblk = pj.b_mma * 2;  // 32-column blocks
    const bool vec_ok = (pj.ld & 3) == 0 && (reinterpret_cast<uintptr_t>(pj.dW) & 15) == 0;
    for (int hh = 0; hh < halves; ++hh) {
      float* drow = pj.dW + (size_t)(hh * 128 + q * 32 + lane) * pj.ld;
      for (int blk = wg; blk < nblk; blk += 2) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + hh * 256 + blk * 32, v);
        tmem_ld_wait();
        const int ch = blk >> 1, cbase = (blk & 1) * 32;
        const int valid = pj.b_valid[ch] - cbase;
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
  };

  for (uint32_t uq = 0;; ++uq) {
    const uint32_t us = uq & 1, uph = (uq >> 1) & 1;
    int unit;   // (job << 20) | granule, or -1: no work left
    if (warp == 0) {
      // ---------------- scheduler: claim the next granule and publish it ----------------
      unit = 0;
      if (lane == 0) {
        mbar_wait(bar_uq_empty + us, uph ^ 1);
        unit = -1;
        for (int tries = 0; tries < a.n_jobs; ++tries) {
          const int g = atomicAdd(a.counters + sched_job, 1);
          if (g < a.jobs[sched_job].n_gran) { unit = (sched_job << 20) | g; break; }
          if (++sched_job == a.n_jobs) sched_job = 0;
        }
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
    const int j = unit >> 20;
    const WgradJob& job = a.jobs[j];
    const int t0 = (unit & 0xfffff) * job.tiles_per_gran;
    const int t1 = min(a.num_tiles, t0 + job.tiles_per_gran);
    const int n_stage = (t1 - t0) * 4;
    const int a_chunks = job.a_chunks, b_n = job.b_n;
    const bool new_job = j != prev_job;

    if (warp == 0) {
      // ---------------- producer: lane c copies slice c of every stage ----------------
      const int n_copies = a_chunks + b_n;
      const uint8_t* src = nullptr;
      size_t tile_stride = 0;
      uint32_t dst_off = 0;
      if (lane < a_chunks) {
        src = a.gstash + img_off(job.a_slot, a_chunks, t0, a.num_tiles) + (size_t)lane * kActChunkBytes;
        tile_stride = (size_t)a_chunks * kActChunkBytes;
        dst_off = lane * kWgSlice;
      } else if (lane < n_copies) {
        const int bi = lane - a_chunks;
        src = a.xstash + img_off(job.b_img[bi], job.b_imgn[bi], t0, a.num_tiles) +
              (size_t)(job.b_slot[bi] - job.b_img[bi]) * kActChunkBytes;
        tile_stride = (size_t)job.b_imgn[bi] * kActChunkBytes;
        dst_off = (kWgASlots + bi) * kWgSlice;
      }
      const uint32_t bytes = (uint32_t)n_copies * kWgSlice;
      // head jobs: the stage's 32 rows of d_sigma (128 B) / dpre (512 B) land with it in the free X slot 4, so the side
      // warps never wait for a global load (one DRAM latency per stage made the two head jobs latency-bound: 0.15 ms)
      const uint32_t vec_bytes = side_vec_bytes(job.side);
      const uint64_t stream_pol = l2_policy_evict_first();   // every image byte is read once
      for (int s = 0; s < n_stage; ++s, ++it) {
        const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
        const int64_t row0 = (int64_t)(t0 + (s >> 2)) * kTileM + (s & 3) * kWgSliceRows;
        const bool vec = vec_bytes != 0 && side_vec_in_range(job.side, row0, a.n_rows);
        if (lane == 0) {
          mbar_wait(bar_empty + stage, ph ^ 1);
          mbar_arrive_expect_tx(bar_full + stage, bytes + (vec ? vec_bytes : 0u));
        }
        __syncwarp();
        if (lane < n_copies)
          bulk_g2s(smem + stage * kWgStageBytes + dst_off, src + (size_t)(s >> 2) * tile_stride + (s & 3) * kWgSlice,
                   kWgSlice, bar_full + stage, stream_pol);
        else if (lane == n_copies && vec)
          bulk_g2s(smem + stage * kWgStageBytes + kWgSideVecOff,
                   job.side == 1 ? reinterpret_cast<const uint8_t*>(a.d_sigma + row0) : reinterpret_cast<const uint8_t*>(a.dpre + row0 * 4),
                   vec_bytes, bar_full + stage);
      }
    } else if (warp == 1) {
      // ---------------- UMMA issuer ----------------
      if (lane == 0) {
        if (new_job) {
          if (prev_job >= 0) {   // the previous job's accumulators are complete: hand them over and wait for the flush
            umma_commit(bar_acc);
            mbar_wait(bar_drained, grp & 1);
            tc_fence_after();
          }
          fresh = true;
        }
        const int halves = NRF_WG_DBG(2) ? 0 : a_chunks >> 1;
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
              const uint32_t acc = !(fresh && ks == 0);
              umma_ss(tmem_base + hh * 256, ad, make_desc_sw128(base + kWgASlots * kWgSlice + ks * 2048, kWgSlice, 1024),
                      idesc_main, acc);
              if (tail)
                umma_ss(tmem_base + hh * 256 + 256, ad,
                        make_desc_sw128(base + (kWgASlots + 4) * kWgSlice + ks * 2048, kWgSlice, 1024), idesc_tail, acc);
            }
          }
          umma_commit(bar_empty + stage);
          fresh = false;
        }
      } else {
        it += n_stage;
      }
    } else {
      // ---------------- side warps: column sums + head gradients, then the accumulator flush ----------------
      // A thread owns ONE logical 8-column block cb of ONE slice k and a few rows of every stage: LDS.128 per row
      // (physical piece = cb ^ (row & 7)), register accumulators per column, so the end-of-unit reduction is two
      // shuffles and a handful of atomics.  All loads of a stage are issued first and the stage is released
      // before the arithmetic.
      if (new_job && prev_job >= 0) flush_group(a.jobs[prev_job]);
      // The arithmetic of the side work competes for issue slots with whatever shares its SM sub-partition; the
      // producer and the UMMA issuer (warps 0, 1) are latency critical, so only the side warps of sub-partitions 2
      // and 3 (warp & 2) do it (measured: -0.1 ms per step); the others only help with the accumulator flush.  (r02:
      // letting the warps of sub-partitions 0 and 1 merely wait for and release every stage costs +0.1 ms; the four
      // that do take part cost nothing beyond their arithmetic: with the side warps out of the ring altogether the
      // kernel runs exactly as fast as with their work skipped, tools/wgrad_probe.py.)
      const bool active = (warp & 2) != 0;
      if (!active) {
        it += n_stage;
      } else {
      const int t = threadIdx.x - 64;
      const int cb = t & 7, rql = (t >> 3) & 3, sw = t >> 5;
      const int ai = (sw & 1) | ((sw >> 2) << 1);       // 0..3 over the active warps
      const bool do_bias = job.db != nullptr && !NRF_WG_DBG(1 | 16384);   // 16384: experiment, heads only
      const int side = NRF_WG_DBG(1 | 8192) ? 0 : job.side;               // 8192: experiment, bias sums only
      // bias / sigma head: slice k4 = ai, rows rql*4 .. +3 and the same +16;  rgb head: slice k2 = ai & 1, rows
      // ((ai >> 1) * 8 + rql * 2) .. +1 and the same +16
      const int k4 = ai;
      const int k2 = ai & 1;
      uint32_t off4[8], off2[4];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = rql * 4 + (i & 3) + (i >> 2) * 16;
        off4[i] = (uint32_t)(r * 128 + ((cb ^ (r & 7)) << 4));
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = (ai >> 1) * 8 + rql * 2 + (i & 1) + (i >> 1) * 16;
        off2[i] = (uint32_t)(r * 128 + ((cb ^ (r & 7)) << 4));
      }
      const bool bias_on = do_bias && k4 < a_chunks;
      auto unpack = [](const uint4& w, float (&f)[8]) {
        f[0] = __uint_as_float(w.x << 16); f[1] = __uint_as_float(w.x & 0xffff0000u);
        f[2] = __uint_as_float(w.y << 16); f[3] = __uint_as_float(w.y & 0xffff0000u);
        f[4] = __uint_as_float(w.z << 16); f[5] = __uint_as_float(w.z & 0xffff0000u);
        f[6] = __uint_as_float(w.w << 16); f[7] = __uint_as_float(w.w & 0xffff0000u);
      };
      for (int s = 0; s < n_stage; ++s, ++it) {
        const int tile = t0 + (s >> 2), rs = s & 3;
        const uint32_t stage = it % kWgStages, ph = (it / kWgStages) & 1;
        const int64_t row0 = (int64_t)tile * kTileM + rs * kWgSliceRows;
        float ds[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float4 dp[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) dp[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool vec = side_vec_in_range(side, row0, a.n_rows);   // the producer applies the same test
        if (side == 1 && !vec) {           // the slice that straddles the end of d_sigma: element-wise, guarded
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int64_t r = row0 + rql * 4 + (i & 3) + (i >> 2) * 16;
            ds[i] = (r < a.n_rows) ? __ldg(a.d_sigma + r) : 0.f;
          }
        }
        mbar_wait(bar_full + stage, ph);
        const uint8_t* st = smem + stage * kWgStageBytes;
        uint4 gw[8], xw[8];
        uint32_t dep = 0;
        if (bias_on) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { gw[i] = *reinterpret_cast<const uint4*>(st + k4 * kWgSlice + off4[i]); dep ^= gw[i].w; }
        }
        if (side == 1) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            xw[i] = *reinterpret_cast<const uint4*>(st + (kWgASlots + job.side_off + k4) * kWgSlice + off4[i]);
            dep ^= xw[i].w;
          }
          if (vec) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              ds[i] = *reinterpret_cast<const float*>(st + kWgSideVecOff + (rql * 4 + (i & 3) + (i >> 2) * 16) * 4);
              dep ^= __float_as_uint(ds[i]);
            }
          }
        } else if (side == 2) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            xw[i] = *reinterpret_cast<const uint4*>(st + (kWgASlots + job.side_off + k2) * kWgSlice + off2[i]);
            dp[i] = *reinterpret_cast<const float4*>(st + kWgSideVecOff + ((ai >> 1) * 8 + rql * 2 + (i & 1) + (i >> 1) * 16) * 16);
            dep ^= xw[i].w ^ __float_as_uint(dp[i].w);
          }
        }
        // every load above must have RETURNED before the stage goes back to the async-proxy producer: a (never
        // taken) store that depends on all loaded words keeps the scoreboard wait ahead of the arrive
        if (dep == 0x5bd1e995u) s_unit[2] = (int)dep;   // s_unit[2]: scratch word nobody reads
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + stage);
        if (NRF_WG_DBG(4)) continue;   // experiment: loads only, no arithmetic
        float f[8];
        if (bias_on) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            unpack(gw[i], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) bs[e] += f[e];
          }
        }
        if (side == 1) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            unpack(xw[i], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) hs[e] = fmaf(ds[i], f[e], hs[e]);
            hb0 += ds[i];
          }
        } else if (side == 2) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            unpack(xw[i], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              hs[e] = fmaf(dp[i].x, f[e], hs[e]);
              hg[e] = fmaf(dp[i].y, f[e], hg[e]);
              hbv[e] = fmaf(dp[i].z, f[e], hbv[e]);
            }
            hb0 += dp[i].x; hb1 += dp[i].y; hb2 += dp[i].z;
          }
        }
      }
      }   // active
    }
    if (new_job && prev_job >= 0) ++grp;
    prev_job = j;
  }
  // the last job's accumulators
  if (prev_job >= 0) {
    if (warp == 1) {
      if (lane == 0) umma_commit(bar_acc);
    } else if (warp >= 2) {
      flush_group(a.jobs[prev_job]);
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
  for (int s : b_slots) {
    j.b_slot[i] = s;
    // image the chunk belongs to (tc_layout.cuh): enc, dir: 1 chunk; h_l, feat: 4; dmid: 2
    if (s < kStashH0) { j.b_img[i] = s; j.b_imgn[i] = 1; }
    else if (s < kStashMid) { j.b_img[i] = kStashH0 + ((s - kStashH0) / 4) * 4; j.b_imgn[i] = 4; }
    else { j.b_img[i] = kStashMid; j.b_imgn[i] = 2; }
    ++i;
  }
  i = 0;
  for (int v : valid) j.b_valid[i++] = v;
  j.dW = dW; j.ld = ld; j.db = db; j.side = side; j.side_off = side_off;
}

static int launch_wgrad(bool full, const nrf_mlp_grads* g, const uint8_t* xstash, const uint8_t* gstash,
                        const float* d_sigma, const float* dpre, int64_t n_rows, int num_tiles, int* unit_counter,
                        cudaStream_t st) {
  WgradArgs a{};
#if NRF_FWD_DEBUG   // debug build only: results are WRONG under these flags (tools/wgrad_probe.py)
  { const char* e = getenv("NRF_DBG_FLAGS"); a.dbg_flags = e ? atoi(e) : 0; }
#endif
  a.num_tiles = num_tiles; a.n_rows = n_rows;
  a.xstash = xstash;
  a.gstash = gstash;
  a.d_sigma = d_sigma; a.dpre = dpre;
  a.dW_sigma = g->sigma_w; a.db_sigma = g->sigma_b; a.dW_rgb = g->rgb_w; a.db_rgb = g->rgb_b;
  a.counters = unit_counter;
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
  // granules of a few tiles, claimed per job; CTAs start spread over the jobs in proportion to the jobs' cost
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // per-stage cost ~ bytes, with a floor: thin stages are latency- rather than bandwidth-paced
  auto stage_cost = [](const WgradJob& j) { return std::max(5, j.a_chunks + j.b_n); };
  double total_cost = 0;
  for (int i = 0; i < a.n_jobs; ++i) total_cost += stage_cost(a.jobs[i]);
  // granule = 16 tiles (measured best of 2..64 at 4096 x 64/128 rows), less when a CTA's share of a job is small
  const long per_cta = (long)num_tiles * a.n_jobs / sms;
  int tpg = (int)std::max<long>(1, std::min<long>(16, per_cta / 8));
  if (NRF_WG_DBG(8)) tpg = 4; else if (NRF_WG_DBG(16)) tpg = 8; else if (NRF_WG_DBG(32)) tpg = 32;
  long total_gran = 0;
  for (int i = 0; i < a.n_jobs; ++i) {
    WgradJob& j = a.jobs[i];
    j.tiles_per_gran = std::max(1, std::min(tpg, num_tiles));
    j.n_gran = (num_tiles + j.tiles_per_gran - 1) / j.tiles_per_gran;
    total_gran += j.n_gran;
  }
  const int grid = (int)std::min<long>(total_gran, sms);
  double cum = 0;
  for (int i = 0; i < a.n_jobs; ++i) {
    a.jobs[i].first_cta = (int)(grid * cum / total_cost);
    cum += stage_cost(a.jobs[i]);
  }
  const int smem_bytes = (int)kWgSmTotal + 1024;
  NRF_OPT_IN_SMEM(mlp_wgrad_kernel, smem_bytes);
  NRF_CUDA(cudaMemsetAsync(unit_counter, 0, kWgMaxJobs * sizeof(int), st));
  ProfScope prof(full ? NRF_K_WGRAD_FULL : NRF_K_WGRAD_SIGMA, st, n_rows);
  mlp_wgrad_kernel<<<grid, kWgThreads, smem_bytes, st>>>(a);
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
  NRF_REQUIRE((reinterpret_cast<uintptr_t>(d_sigma) & 15) == 0, "nrf_mlp_tc_bwd: d_sigma must be 16-byte aligned (bulk-copied into the wgrad stages)");
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
