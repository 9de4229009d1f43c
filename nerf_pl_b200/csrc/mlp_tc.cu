// bf16 tensor-core NeRF MLP, forward (models/nerf.py:83-124 fused with the positional encoding
// nerf.py:21-38 and the point generation rendering.py:234).
//
// Persistent kernel, one CTA per SM, 128-row tiles of sample points; the layer chain is software
// pipelined at 64-column (= one K chunk) granularity so the tensor pipe rarely waits for the epilogue:
//   warp 0      weight producer: linear bulk copies (TMA engine) of pre-packed 32 KB weight chunks
//               L2 -> 4-stage shared-memory ring, mbarrier complete_tx
//   warp 1      UMMA issuer: tcgen05.mma kind::f16, M=128, N=256 (128 for the direction layer),
//               K=16 x 4 per chunk, A and B from shared memory (SWIZZLE_128B K-major).  Two fp32
//               accumulators in TMEM (2 x 256 columns) alternate between layers; the MMAs of layer
//               L+1 over K chunk c start as soon as the epilogue of layer L has produced chunk c.
//   warps 2-17  epilogue (4 warps per TMEM lane quarter, 16 columns of every chunk each): tcgen05.ld ->
//               +bias -> ReLU -> bf16 -> in place into the A-operand buffer -> per-chunk mbarrier;
//               sigma / rgb heads as fp32 dot products on the accumulators.  One 16 KB buffer holds the
//               xyz encoding (layers 0..4) and then the direction encoding (layer 9) of a tile; the NEXT
//               tile's xyz encoding is written as soon as the last layer's MMAs are done, before that
//               layer's own epilogue, so the next chain starts while the heads are still evaluated.
//               Training mode also streams the operand images + ReLU sign masks to the HBM stash.
// Roofline: tensor pipe (593 408 MAC/sample).  HBM traffic: 32 B/ray in, 16 B/sample
// out (+5.3 KB/sample stash when training); weights (1.17 MB) stay L2 resident (measured L2 -> SMEM
// bulk-copy rate with all 148 CTAs on the same blob: > 19 TB/s, tools/stream_bench.py).
#include <cstdlib>
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"
#include "tc_fwd.cuh"

namespace nrf {
namespace tc {

constexpr int kStages = 4;
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreadsInfer = 64 + kEpiThreads;
constexpr int kThreadsTrain = kThreadsInfer + 32;   // + the stash store warp
constexpr uint32_t kTmemCols = 512;
// Hand-over groups: the epilogue announces its output (fence.proxy.async + mbarrier arrive) per GROUP of 64-column chunks,
// and the MMAs of the next layer start per group.  NRF_FWD_GRP: 1 = {0}{1}{2}{3}, 2 = {0,1}{2,3}, 4 = {0,1,2,3},
// 112 = {0}{1}{2,3}.  Every fence stalls all epilogue warps of an SM sub-partition together; measured on B200 (r02,
// tools/step_probe.py): schemes 1, 2 and 112 are within 2 % of each other, 2 is kept.
#ifndef NRF_FWD_GRP
#define NRF_FWD_GRP 2
#endif
constexpr int kGrpScheme = NRF_FWD_GRP;
static_assert(kGrpScheme == 1 || kGrpScheme == 2 || kGrpScheme == 4 || kGrpScheme == 112, "NRF_FWD_GRP");
// first / last chunk of the group chunk c belongs to; barriers and buffers are indexed by the group's LAST chunk
__host__ __device__ constexpr int grp_first(int c) {
  return kGrpScheme == 1 ? c : kGrpScheme == 2 ? (c & ~1) : kGrpScheme == 4 ? 0 : (c < 2 ? c : 2);
}
__host__ __device__ constexpr int grp_last(int c) {
  return kGrpScheme == 1 ? c : kGrpScheme == 2 ? (c | 1) : kGrpScheme == 4 ? 3 : (c < 2 ? c : 3);
}
constexpr bool kSplit01 = grp_last(0) == 0;   // chunks 0 and 1 are handed over separately

// shared-memory map (dynamic shared memory is declared 1024-byte aligned; no slack is left for re-alignment)
constexpr uint32_t kSmEnc = 0;                                    // xyz encoding, later direction encoding
constexpr uint32_t kSmAct = kSmEnc + kActChunkBytes;
constexpr uint32_t kSmRing = kSmAct + 4 * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStages * kChunkBytes256;
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;  // [4][128][3] floats
constexpr uint32_t kSmBars = kSmScratch + 4 * 128 * 3 * 4;
constexpr uint32_t kSmTotal = kSmBars + 256;
static_assert(kSmTotal <= 232448, "shared memory budget");

// A-operand readiness barriers: act chunks 0..3, xyz encoding, direction encoding
constexpr int kBarEnc = 4, kBarDir = 5, kNumBarA = 6;
// "stash store has read the buffer" barriers (training): act chunks 0..3, encoding buffer
constexpr int kFreeEnc = 4, kNumBarFree = 5;
// images nobody but the store warp reads (a tile's last layer; dmid) are announced on barriers of their own: a parity
// wait only tells "one completion behind" from "up to date", so every waiter must consume every completion of a barrier
constexpr int kNumBarS = 4;

template <bool kFull, bool kTrain>
__global__ void __launch_bounds__(kTrain ? kThreadsTrain : kThreadsInfer, 1) mlp_fwd_kernel(const __grid_constant__ FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_act = smem + kSmAct;
  uint8_t* s_ring = smem + kSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;                 // [kStages] weight chunk landed
  uint64_t* bar_empty = bars + kStages;      // [kStages] weight chunk consumed
  uint64_t* bar_a = bars + 2 * kStages;      // [kNumBarA] A-operand chunk written (epilogue -> MMA)
  uint64_t* bar_acc = bar_a + kNumBarA;      // [2] accumulator complete (MMA -> epilogue)
  uint64_t* bar_free = bar_acc + 2;          // [kNumBarFree] stash store finished reading a buffer (store warp -> epilogue)
  uint64_t* bar_s = bar_free + kNumBarFree;  // [kNumBarS] store-only image chunk written (epilogue -> store warp)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_s + kNumBarS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kLayers = kFull ? 10 : 8;
  constexpr int kChunksPerTile = kFull ? kFwdChunksFull : kFwdChunksSigma;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();    // SWIZZLE_128B operands need the 1024-byte alignment
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < kNumBarA; ++s) mbar_init(bar_a + s, kEpiWarps);
    mbar_init(bar_acc, 1);
    mbar_init(bar_acc + 1, 1);
    for (int s = 0; s < kNumBarFree; ++s) mbar_init(bar_free + s, 1);
    for (int s = 0; s < kNumBarS; ++s) mbar_init(bar_s + s, kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, kTmemCols);
  if (warp >= 2 && warp < 2 + kEpiWarps) {
    const int et = threadIdx.x - 64;
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = et; i < AuxOffsets::count; i += kEpiThreads) s_aux[i] = gaux[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===================== weight producer =====================
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      int tn0 = 0;
      const uint64_t keep = l2_policy_evict_last();   // the weights stay in L2 under the stash write stream
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunksPerTile; ++ci) {
          NRF_MARK(0, 0x1000000 | (tile << 8) | ci);
          mbar_wait(bar_empty + stage, ph ^ 1);
          NRF_TRACE(0, tn0);
          const bool half = ci >= kFwdChunksSigma + 4;
          const uint32_t bytes = half ? kChunkBytes128 : kChunkBytes256;
          const size_t off = half ? (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 +
                                        (size_t)(ci - kFwdChunksSigma - 4) * kChunkBytes128
                                  : (size_t)ci * kChunkBytes256;
          mbar_arrive_expect_tx(bar_full + stage, bytes);
          bulk_g2s(s_ring + stage * kChunkBytes256, a.packed + kPackedFwdOff + off, bytes, bar_full + stage, keep);
          if (++stage == kStages) { stage = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    // The WHOLE warp runs this loop with warp-uniform values and one elected lane issues the tcgen05 instructions: the
    // compiler then keeps descriptors and barrier addresses in uniform registers.  (Under `if (lane == 0)` every UTCHMMA
    // was wrapped in a per-thread elect loop with R2UR moves, ~165 cycles per MMA: the issuing thread, not the tensor
    // pipe (128 cycles per MMA), paced the kernel - tools/fwd_warp_skew.py.)
    {
      uint32_t stage = 0, ph = 0;
      uint32_t a_par = 0;   // bit i: parity of the next completion of bar_a[i] to wait for
      long long w_a = 0, w_full = 0, t_begin = clock64();
      int tn1 = 0;
      const uint32_t idesc256 = make_idesc_bf16(128, 256, 0, 0), idesc128 = make_idesc_bf16(128, 128, 0, 0);
      const uint64_t enc_desc = make_desc_sw128(smem_u32(s_enc), 16, 1024);
      const uint64_t act_desc = make_desc_sw128(smem_u32(s_act), 16, 1024);
      const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
      int mk = 0;   // (tile, layer) for the progress markers
      auto wait_a = [&](int src) {
        const long long t0 = NRF_HEAVY ? clock64() : 0;
        NRF_MARK(1, 0x2000000 | (mk << 4) | src);
#ifdef NRF_SPIN_MMA
        mbar_wait_spin(bar_a + src, (a_par >> src) & 1u);
#else
        mbar_wait(bar_a + src, (a_par >> src) & 1u);
#endif
        a_par ^= 1u << src;
        if (NRF_HEAVY) w_a += clock64() - t0;
      };
      // Weight stages are awaited AHEAD of use, while the thread would otherwise idle waiting for the epilogue's
      // first chunk: the per-chunk critical path is then one barrier wait + the issue of 4 MMAs (K = 16 each).
      int ready = 0;   // upcoming ring stages already known to be full
      auto prewait = [&](int n) {
        uint32_t st = stage, p = ph;
        for (int i = 0; i < ready; ++i) { if (++st == kStages) { st = 0; p ^= 1; } }
        for (; ready < n; ++ready) {
          mbar_wait(bar_full + st, p);
          if (++st == kStages) { st = 0; p ^= 1; }
        }
      };
      auto chunk = [&](uint64_t a_desc, uint32_t d_tmem, uint32_t idesc, bool first) {
        const long long t0 = NRF_HEAVY ? clock64() : 0;
        NRF_TRACE(1, tn1);
        NRF_MARK(1, 0x3000000 | (mk << 4));
        if (ready == 0) mbar_wait(bar_full + stage, ph); else --ready;
        NRF_TRACE(1, tn1);
        if (NRF_HEAVY) w_full += clock64() - t0;
        tc_fence_after();
        const uint64_t b_desc = ring_desc + (uint64_t)(stage * (kChunkBytes256 >> 4));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (!first || k) ? 1u : 0u);
          umma_commit(bar_empty + stage);
        }
        __syncwarp();
        NRF_TRACE(1, tn1);
        if (++stage == kStages) { stage = 0; ph ^= 1; }
      };
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
#pragma unroll 1
        for (int L = 0; L < kLayers; ++L) {
          mk = (tile << 4) | L;
          const bool mtr = NRF_FWD_DEBUG && blockIdx.x == 0 && lane == 0 && tile == 2 * (int)gridDim.x;
          const uint32_t d_tmem = tmem_base + (uint32_t)(L & 1) * 256;
          const uint32_t idesc = (L == 9) ? idesc128 : idesc256;
          bool first = true;
          if (L == 0 || L == 4) {          // the xyz encoding leads (skip connection: models/nerf.py:96-97)
            if (L == 0) wait_a(kBarEnc);   // at layer 4 the same completion is still current
            chunk(enc_desc, d_tmem, idesc, true);
            first = false;
          }
          if (L != 0) {
            prewait(3);
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              if (c == grp_first(c)) wait_a(grp_last(c));
              if (c == 0) NRF_WTRACE(mtr, 16, L, 0);
              if (c == grp_first(3)) NRF_WTRACE(mtr, 16, L, 2);
              chunk(act_desc + (uint64_t)(c * (kActChunkBytes >> 4)), d_tmem, idesc, first);
              if (c == grp_last(0)) NRF_WTRACE(mtr, 16, L, 1);
              first = false;
            }
          }
          if (L == 9) {                    // the direction encoding has replaced the xyz encoding in s_enc
            wait_a(kBarDir);
            chunk(enc_desc, d_tmem, idesc, false);
          }
          if (elect_one()) umma_commit(bar_acc + (L & 1));
          __syncwarp();
          NRF_WTRACE(mtr, 16, L, 3);
        }
      }
      if (NRF_FWD_DEBUG == 1 && a.dbg && lane == 0) {
        a.dbg[blockIdx.x * 8 + 0] = (unsigned long long)(clock64() - t_begin);
        a.dbg[blockIdx.x * 8 + 1] = (unsigned long long)w_a;
        a.dbg[blockIdx.x * 8 + 2] = (unsigned long long)w_full;
      }
    }
  } else if (warp < 2 + kEpiWarps) {
    // ===================== epilogue warps =====================
    const int et = threadIdx.x - 64;
    const int q = warp & 3, sub = (warp - 2) >> 2;   // TMEM lane quarter; 16-column quarter of every 64-column chunk
    const int row = q * 32 + lane;
    const RowSwz rsw(row);
    const uint32_t act_s = smem_u32(s_act), enc_s = smem_u32(s_enc);
    const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma);
    uint32_t acc_par = 0;
    long long w_acc = 0, w_free = 0, t_begin = clock64();
    int tn2 = 0;
    // training: every image written to s_act / s_enc is also copied to the stash by the store warp; a buffer may be
    // overwritten once that copy has read it (bar_free).  bit i of `stored`: a copy of buffer i has been requested.
    uint32_t free_par = 0, stored = 0;
    int mk = 0;
    auto acquire = [&](int buf) {
      if (!kTrain) return;
      if ((stored >> buf) & 1u) {
        if (lane == 0) NRF_MARK(warp, 0x4000000 | (mk << 4) | buf);
        const long long tf0 = NRF_HEAVY ? clock64() : 0;
        if (!(NRF_FWD_DEBUG && (a.dbg_flags & 64)))   // 64: timing experiment, buffers overwritten without waiting for the stash store
          mbar_wait(bar_free + buf, (free_par >> buf) & 1u);
        if (NRF_HEAVY) w_free += clock64() - tf0;
        free_par ^= 1u << buf;
      }
      stored |= 1u << buf;
    };

    // this warp's part of an A-operand chunk is written: make it visible to the async proxy and tell the MMA thread
    auto signal = [&](int idx) {
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a + idx);
    };
    auto signal_store = [&](int c) {   // an image chunk only the store warp reads
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_s + c);
    };
    // positional encoding of a tile's sample points -> s_enc (thread (er, part) fills 16 columns of row er)
    auto encode_xyz = [&](int tile) {
      const int er = et & 127, part = et >> 7;
      int64_t gr = (int64_t)tile * kTileM + er;
      if (gr >= a.n_rows) gr = a.n_rows - 1;
      const int64_t ray = gr / a.S;
      const float zv = __ldg(a.z + gr);
      float p[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) p[c] = sample_point(a.rays, ray, zv, c);
      acquire(kFreeEnc);
      encode_store16<10>(part, p, enc_s, er);
      signal(kBarEnc);
    };
    // direction encoding of the same rows, into the same buffer (the xyz encoding is dead after layer 4)
    auto encode_dir = [&](int tile) {
      const int er = et & 127, part = et >> 7;
      int64_t gr = (int64_t)tile * kTileM + er;
      if (gr >= a.n_rows) gr = a.n_rows - 1;
      const int64_t ray = gr / a.S;
      float d[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) d[c] = __ldg(a.rays + ray * 8 + 3 + c);
      acquire(kFreeEnc);
      encode_store16<4>(part, d, enc_s, er);
      signal(kBarDir);
    };

    int tile = blockIdx.x;
    if (tile < a.num_tiles) encode_xyz(tile);
    for (; tile < a.num_tiles; tile += gridDim.x) {
      const int next = tile + gridDim.x;
      uint32_t* mask_tile = kTrain ? a.masks + (size_t)tile * mask_tile_words(kFull) : nullptr;
      const int64_t grow = (int64_t)tile * kTileM + row;
      const bool valid = grow < a.n_rows;
#pragma unroll 1
      for (int L = 0; L < kLayers; ++L) {
        const uint32_t b = L & 1;
        const long long tw0 = NRF_HEAVY ? clock64() : 0;
        mk = (tile << 4) | L;
        if (lane == 0) NRF_MARK(warp, 0x5000000 | (mk << 4));
#ifdef NRF_SPIN_EPI
        mbar_wait_spin(bar_acc + b, (acc_par >> b) & 1u);
#else
        mbar_wait(bar_acc + b, (acc_par >> b) & 1u);
#endif
        if (NRF_HEAVY) w_acc += clock64() - tw0;
        if (et == 0) NRF_TRACE(2, tn2);
        const bool wtr = NRF_FWD_DEBUG && blockIdx.x == 0 && lane == 0 && tile == 2 * (int)gridDim.x;
        NRF_WTRACE(wtr, warp - 2, L, 0);
        acc_par ^= 1u << b;
        tc_fence_after();
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + b * 256;
        if (L < 9) {
          const bool feed = L + 1 < kLayers;   // the next layer's MMAs consume these chunks
          const uint32_t bias_s = smem_u32(s_aux + (L < 8 ? AuxOffsets::bias + L * 256 : AuxOffsets::bias_final));
          uint32_t mreg = 0, mw0 = 0, mw1 = 0;
          float dot = 0.f;
          uint32_t va[16], vb[16];
          tmem_ld16(t_acc + sub * 16, va);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t(&v)[16] = (c & 1) ? vb : va;
            uint32_t(&vn)[16] = (c & 1) ? va : vb;
            tmem_ld_wait_dep16(v);
            if (c + 1 < 4) tmem_ld16(t_acc + (c + 1) * 64 + sub * 16, vn);
            const int c0 = c * 64 + sub * 16;
            const uint32_t chunk_s = act_s + (uint32_t)c * kActChunkBytes;
            if (c == grp_first(c)) acquire(grp_last(c));
            if (L == 7) epi_fwd16<true, kTrain, true>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 2, mreg, dot);
            else if (L < 8) epi_fwd16<true, kTrain, false>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 2, mreg, dot);
            else epi_fwd16<false, false, false>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 2, mreg, dot);
            if (kTrain && (c & 1)) {
              if (c == 1) mw0 = ~mreg; else mw1 = ~mreg;
              mreg = 0;
            }
            if (c == grp_last(c)) {
              if (feed) signal(c);             // the MMA thread (and, in training, the store warp) take it from here
              else if (kTrain) signal_store(c);
              if (c == grp_last(0)) NRF_WTRACE(wtr, warp - 2, L, 1);
              if (c == 3) NRF_WTRACE(wtr, warp - 2, L, 2);
            }
            if (et == 0) NRF_TRACE(2, tn2);
          }
          if (kTrain && L < 8)
            *reinterpret_cast<uint2*>(mask_tile + ((size_t)L * 128 + row) * 8 + sub * 2) = make_uint2(mw0, mw1);
          if (L == 7) {
            s_scr[sub * 128 + row] = dot;
            named_bar_sync(2, kEpiThreads);
            if (sub == 0 && valid)
              a.sigma_out[grow] = (s_scr[row] + s_scr[128 + row]) + (s_scr[256 + row] + s_scr[384 + row]) + s_aux[AuxOffsets::b_sigma];
          }
          if (L == 4) {
            if (kFull) encode_dir(tile);                        // s_enc: xyz encoding -> direction encoding
            else if (next < a.num_tiles) encode_xyz(next);     // sigma-only: the buffer is free from here on
          }
        } else {
          // direction layer: N = 128; rgb head on the fp32 accumulators (nerf.py:119-121).
          // Its MMAs are done, so s_enc is free: start the next tile's chain before evaluating the heads.
          if (next < a.num_tiles) encode_xyz(next);
          float r0 = 0.f, r1 = 0.f, r2 = 0.f;
          uint32_t mreg = 0;
          const uint32_t bdir_s = smem_u32(s_aux + AuxOffsets::bias_dir), wrgb_s = smem_u32(s_aux + AuxOffsets::w_rgb);
          uint32_t va[16], vb[16];
          tmem_ld16(t_acc + sub * 16, va);
          tmem_ld16(t_acc + 64 + sub * 16, vb);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t(&v)[16] = c ? vb : va;
            tmem_ld_wait_dep16(v);
            const int c0 = c * 64 + sub * 16;
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              const float4 bb = ld_shared_f4(bdir_s + (c0 + j) * 4);
              const float x0 = __uint_as_float(v[j]) + bb.x, x1 = __uint_as_float(v[j + 1]) + bb.y;
              const float x2 = __uint_as_float(v[j + 2]) + bb.z, x3 = __uint_as_float(v[j + 3]) + bb.w;
              if (kTrain) {
                mreg = push_sign(mreg, x0); mreg = push_sign(mreg, x1);
                mreg = push_sign(mreg, x2); mreg = push_sign(mreg, x3);
              }
              const float f0 = fmaxf(x0, 0.f), f1 = fmaxf(x1, 0.f), f2 = fmaxf(x2, 0.f), f3 = fmaxf(x3, 0.f);
              const float4 w0 = ld_shared_f4(wrgb_s + (c0 + j) * 4), w1 = ld_shared_f4(wrgb_s + (128 + c0 + j) * 4),
                           w2 = ld_shared_f4(wrgb_s + (256 + c0 + j) * 4);
              r0 += f0 * w0.x + f1 * w0.y + f2 * w0.z + f3 * w0.w;
              r1 += f0 * w1.x + f1 * w1.y + f2 * w1.z + f3 * w1.w;
              r2 += f0 * w2.x + f1 * w2.y + f2 * w2.z + f3 * w2.w;
              if (kTrain) {
                pk[j >> 1] = cvt_bf16x2(f0, f1);
                pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
              }
            }
            if (kTrain) {
              const uint32_t chunk_s = act_s + (uint32_t)c * kActChunkBytes;
              if (kSplit01 || c == 0) acquire(kSplit01 ? c : grp_last(0));
              st_shared_v4(chunk_s + rsw.piece(sub * 2), pk[0], pk[1], pk[2], pk[3]);
              st_shared_v4(chunk_s + rsw.piece(sub * 2 + 1), pk[4], pk[5], pk[6], pk[7]);
              if (kSplit01 || c == 1) signal_store(kSplit01 ? c : grp_last(0));
            }
          }
          if (kTrain) mask_tile[8 * 128 * 8 + (size_t)row * 4 + sub] = ~mreg;
          float* sc = s_scr + (sub * 128 + row) * 3;
          sc[0] = r0; sc[1] = r1; sc[2] = r2;
          named_bar_sync(2, kEpiThreads);
          if (sub == 0 && valid) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float pre = (s_scr[row * 3 + c] + s_scr[(128 + row) * 3 + c]) +
                                (s_scr[(256 + row) * 3 + c] + s_scr[(384 + row) * 3 + c]) + s_aux[AuxOffsets::b_rgb + c];
              a.rgb_out[grow * 3 + c] = 1.0f / (1.0f + expf(-pre));
            }
          }
        }
      }
    }
    if (NRF_FWD_DEBUG == 1 && a.dbg && et == 0) {
      a.dbg[blockIdx.x * 8 + 4] = (unsigned long long)(clock64() - t_begin);
      a.dbg[blockIdx.x * 8 + 5] = (unsigned long long)w_acc;
      a.dbg[blockIdx.x * 8 + 6] = (unsigned long long)w_free;   // training: waiting for the stash store to have read a buffer
    }
  } else if (kTrain && lane == 0) {
    // ===================== stash store warp (training) =====================
    // One thread follows the epilogue's production order: as soon as a 16 KB operand image is complete (bar_a) it is
    // bulk-copied to the stash; once the copy has read the buffer (one copy behind, so two are in flight) the
    // epilogue may overwrite it (bar_free).
    uint32_t s_par = 0;
    int pending = -1;
    int mk = 0;
    const uint64_t stream_pol = l2_policy_evict_first();   // written once, read next by wgrad from DRAM anyway
    auto store = [&](int bar_idx, int buf, const uint8_t* src, uint8_t* dst, uint32_t bytes = kActChunkBytes) {   // bar_idx >= 8: bar_s[bar_idx - 8]
      NRF_MARK(18, 0x6000000 | (mk << 4) | bar_idx);
      uint64_t* bar = bar_idx < 8 ? bar_a + bar_idx : bar_s + (bar_idx - 8);
      const uint32_t par = (s_par >> bar_idx) & 1u;
      // about to block: hand the pending buffer back first (the epilogue may need it to produce this very image)
      if (pending >= 0 && !mbar_try_wait(bar, par)) {
        bulk_wait_read0();
        mbar_arrive(bar_free + pending);
        pending = -1;
      }
      mbar_wait(bar, par);
      s_par ^= 1u << bar_idx;
      // timing experiments: 128 = no stash writes; 256 = every write lands in one L2-resident 64 KB window per CTA
      if (NRF_FWD_DEBUG && (a.dbg_flags & 256)) dst = a.stash + (size_t)blockIdx.x * 65536 + ((dst - a.stash) & 0xffff);
      if (!(NRF_FWD_DEBUG && (a.dbg_flags & 128))) bulk_s2g(dst, src, bytes, stream_pol);
      bulk_commit();
      if (pending >= 0) {
        bulk_wait_read1();
        mbar_arrive(bar_free + pending);
      }
      pending = buf;
    };
    int tile = blockIdx.x;
    if (tile < a.num_tiles) store(kBarEnc, kFreeEnc, s_enc, a.stash + img_off(kStashEnc, 1, tile, a.num_tiles));
    for (; tile < a.num_tiles; tile += gridDim.x) {
      const int next = tile + gridDim.x;
#pragma unroll 1
      for (int L = 0; L < kLayers; ++L) {
        mk = (tile << 4) | L;
        if (L < 9) {
          uint8_t* dst = a.stash + img_off(L < 8 ? kStashH0 + 4 * L : kStashFeat, 4, tile, a.num_tiles);
          const int bar0 = (L + 1 < kLayers) ? 0 : 8;
          for (int c = 0; c < 4; c = grp_last(c) + 1)
            store(bar0 + grp_last(c), grp_last(c), s_act + c * kActChunkBytes, dst + (size_t)c * kActChunkBytes,
                  (uint32_t)(grp_last(c) - c + 1) * kActChunkBytes);
          if (L == 4) {
            if (kFull) store(kBarDir, kFreeEnc, s_enc, a.stash + img_off(kStashDir, 1, tile, a.num_tiles));
            else if (next < a.num_tiles) store(kBarEnc, kFreeEnc, s_enc, a.stash + img_off(kStashEnc, 1, next, a.num_tiles));
          }
        } else {
          if (next < a.num_tiles) store(kBarEnc, kFreeEnc, s_enc, a.stash + img_off(kStashEnc, 1, next, a.num_tiles));
          uint8_t* dst = a.stash + img_off(kStashMid, 2, tile, a.num_tiles);
          if (kSplit01) {
            for (int c = 0; c < 2; ++c) store(8 + c, c, s_act + c * kActChunkBytes, dst + (size_t)c * kActChunkBytes);
          } else {
            store(8 + grp_last(0), grp_last(0), s_act, dst, 2 * kActChunkBytes);
          }
        }
      }
    }
    bulk_wait_all0();
    NRF_MARK(18, 0x7000000);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <bool kFull, bool kTrain>
static int launch_fwd(const FwdArgs& a, cudaStream_t st) {
  const int smem_bytes = (int)kSmTotal;
  NRF_OPT_IN_SMEM((mlp_fwd_kernel<kFull, kTrain>), smem_bytes);
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = a.num_tiles < sms ? a.num_tiles : sms;
  ProfScope prof((kTrain ? NRF_K_FWD_SIGMA_TRAIN : NRF_K_FWD_SIGMA_INFER) + (kFull ? 1 : 0), st, a.n_rows);
  mlp_fwd_kernel<kFull, kTrain><<<grid, kTrain ? kThreadsTrain : kThreadsInfer, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;

#if NRF_FWD_DEBUG
// Debug build only (NRF_DEBUG_BUILD=1, include/nerf_b200_debug.h); the product library reads no environment variable.
static unsigned long long* g_dbg_counters = nullptr;
// device buffer of >= 8 * gridDim 64-bit counters filled by the next forward launches (null = off)
extern "C" void nrf_dbg_set_counters(void* dev_ptr) { g_dbg_counters = static_cast<unsigned long long*>(dev_ptr); }
// NRF_FWD_IMPL=pair selects the experimental CTA-pair kernel (debug/mlp_tc_pair.cu)
static bool use_pair_kernel() {
  const char* e = getenv("NRF_FWD_IMPL");   // read per call: tools/pair_check.py switches it at run time
  return e && e[0] == 'p';
}
// NRF_FWD_IMPL=ts selects the experimental tensor-memory / N-split kernel (debug/mlp_tc_ts.cu, tools/ts_check.py)
static bool use_ts_kernel() {
  const char* e = getenv("NRF_FWD_IMPL");
  return e && e[0] == 't';
}
// NRF_FWD_IMPL=m selects debug/mlp_tc_tmem.cu (hidden activations as tensor-memory operands, N = 256) for every mode
static int force_impl() {
  const char* e = getenv("NRF_FWD_IMPL");
  return (e && e[0] == 'm') ? 2 : 0;
}
#endif

extern "C" size_t nrf_mlp_tc_stash_bytes(int64_t rows, int sigma_only) {
  if (rows <= 0) return 0;
  const size_t tiles = (size_t)((rows + kTileM - 1) / kTileM);
  return tiles * (stash_tile_bytes(!sigma_only) + mask_tile_words(!sigma_only) * 4);
}

extern "C" int nrf_mlp_tc_fwd(const void* packed, const float* rays, const float* z, int64_t n_rays, int n_samples,
                              int sigma_only, float* sigma_out, float* rgb_out, void* stash, size_t stash_bytes,
                              nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_mlp_tc_fwd: bad sizes");
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(packed && rays && z && sigma_out, "nrf_mlp_tc_fwd: null pointer");
  NRF_REQUIRE(sigma_only || rgb_out, "nrf_mlp_tc_fwd: rgb_out is null in full mode");
  const int64_t rows = n_rays * n_samples;
  NRF_REQUIRE(rows < ((int64_t)1 << 31) * kTileM, "nrf_mlp_tc_fwd: too many rows");
  FwdArgs a{};
  a.packed = static_cast<const uint8_t*>(packed);
  a.rays = rays; a.z = z; a.n_rows = rows; a.S = n_samples;
  a.num_tiles = (int)((rows + kTileM - 1) / kTileM);
  a.sigma_out = sigma_out; a.rgb_out = rgb_out;
#if NRF_FWD_DEBUG
  a.dbg = g_dbg_counters;
  { const char* e = getenv("NRF_DBG_FLAGS"); a.dbg_flags = e ? atoi(e) : 0; }
#endif
  cudaStream_t st = as_stream(stream);
  if (stash != nullptr) {
    const size_t need = nrf_mlp_tc_stash_bytes(rows, sigma_only);
    if (stash_bytes < need) {
      set_error("nrf_mlp_tc_fwd: stash %zu < %zu bytes", stash_bytes, need);
      return NRF_ERR_WORKSPACE;
    }
    a.stash = static_cast<uint8_t*>(stash);
    a.masks = reinterpret_cast<uint32_t*>(a.stash + (size_t)a.num_tiles * stash_tile_bytes(!sigma_only));
#if NRF_FWD_DEBUG
    if (use_pair_kernel()) return launch_fwd_pair(!sigma_only, true, a, st);
#endif
#if NRF_FWD_DEBUG
    if (use_ts_kernel()) return launch_fwd_ts(!sigma_only, true, a, st);
#endif
#if NRF_FWD_DEBUG
    if (force_impl() == 2) return launch_fwd_tmem(!sigma_only, true, a, st);
#endif
    return sigma_only ? launch_fwd<false, true>(a, st) : launch_fwd<true, true>(a, st);
  }
#if NRF_FWD_DEBUG
  if (use_pair_kernel()) return launch_fwd_pair(!sigma_only, false, a, st);
#endif
#if NRF_FWD_DEBUG
  if (use_ts_kernel()) return launch_fwd_ts(!sigma_only, false, a, st);
#endif
#if NRF_FWD_DEBUG
  if (force_impl() == 2) return launch_fwd_tmem(!sigma_only, false, a, st);
#endif
  return sigma_only ? launch_fwd<false, false>(a, st) : launch_fwd<true, false>(a, st);
}

