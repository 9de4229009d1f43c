// bf16 tensor-core NeRF MLP, backward data path ("dgrad"): gradients w.r.t. every layer's
// pre-activation, G_l = dL/da_l, for 128-row tiles of sample points.
//
//   G9 = (d_rgb * rgb(1-rgb)) . W_rgb  (x) relu'(dir layer)         CUDA cores (K = 3)
//   G8 = G9 . W_dir[:, :256]                                        UMMA  M128 N256 K128
//   G7 = (G8 . W_final + d_sigma w_sigma) (x) relu'(h7)             UMMA  M128 N256 K256 + rank-1 on CUDA cores
//   G_{l-1} = (G_l . W_l[:, h part]) (x) relu'(h_{l-1}),  l = 7..1  UMMA  M128 N256 K256
//
// Same warp-specialised, chunk-pipelined structure as the forward kernel (mlp_tc.cu): bulk-copy weight
// ring (transposed, pre-packed images), one UMMA issuer thread, two alternating TMEM accumulators,
// eight epilogue warps that turn the accumulator into the next A operand (bf16, SWIZZLE_128B) in shared
// memory 64 columns at a time while the next layer's MMAs already consume the finished chunks.  ReLU
// derivatives come from the 1-bit sign masks the forward kernel stashed (32 B per row and layer).
// Every G_l image is also streamed to HBM with a bulk store; the wgrad kernel
// (mlp_tc_wgrad.cu) consumes those images as MN-major UMMA operands.
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kDgStages = 4;
constexpr int kDgEpiWarps = 16;
constexpr int kDgEpiThreads = kDgEpiWarps * 32;
constexpr int kDgThreads = 64 + kDgEpiThreads + 32;   // producer, UMMA issuer, epilogue warps, G-image store warp
constexpr uint32_t kDgSmAct = 0;
constexpr uint32_t kDgSmPro = kDgSmAct + 4 * kActChunkBytes;      // full mode: staging of the next tile's G9 (2 chunks)
constexpr uint32_t kDgSmRing = kDgSmPro + 2 * kActChunkBytes;
constexpr uint32_t kDgSmAux = kDgSmRing + kDgStages * kChunkBytes256;
constexpr int kDgAuxFirst = AuxOffsets::w_sigma;                   // only the head weights are needed here
constexpr int kDgAuxCount = AuxOffsets::count - kDgAuxFirst;
constexpr uint32_t kDgSmBars = kDgSmAux + kDgAuxCount * 4;
constexpr uint32_t kDgSmTotal = kDgSmBars + 256;
static_assert(kDgSmTotal <= 232448, "shared memory budget");

struct DgradArgs {
  const uint8_t* packed;
  const uint32_t* masks;   // forward stash masks
  const float* d_sigma;    // (rows)
  const float* d_rgb;      // (rows,3) or null
  const float* rgb_out;    // (rows,3) forward sigmoid outputs, or null
  int64_t n_rows;
  int num_tiles;
  uint8_t* gstash;         // per tile: G images (tc_layout.cuh: gstash_*)
  float* dpre;             // (tiles*128, 4): d_rgb * rgb (1-rgb), written for the wgrad kernel (full mode)
};

constexpr int kDgBarPro = 4, kDgNumBarA = 5;   // A-operand readiness: act chunks 0..3, G9 staging
constexpr int kDgNumBarFree = 5;               // same buffers: the G-image store has read them
// G0 (nobody but the store warp reads it) is announced on barriers of its own: a parity wait only tells "one completion
// behind" from "up to date", so every waiter must consume every completion of a barrier
constexpr int kDgNumBarS = 4;
// hand-over groups of a layer's four chunks (see NRF_FWD_GRP in mlp_tc.cu): 1 = {0}{1}{2}{3}, 2 = {0,1}{2,3}, 4 = all,
// 112 = {0}{1}{2,3}
#ifndef NRF_DG_GRP
#define NRF_DG_GRP 2
#endif
constexpr int kDgScheme = NRF_DG_GRP;
static_assert(kDgScheme == 1 || kDgScheme == 2 || kDgScheme == 4 || kDgScheme == 112, "NRF_DG_GRP");
__host__ __device__ constexpr int dg_first(int c) {
  return kDgScheme == 1 ? c : kDgScheme == 2 ? (c & ~1) : kDgScheme == 4 ? 0 : (c < 2 ? c : 2);
}
__host__ __device__ constexpr int dg_last(int c) {
  return kDgScheme == 1 ? c : kDgScheme == 2 ? (c | 1) : kDgScheme == 4 ? 3 : (c < 2 ? c : 3);
}

template <bool kFull>
__global__ void __launch_bounds__(kDgThreads, 1) mlp_dgrad_kernel(const __grid_constant__ DgradArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_act = smem + kDgSmAct;
  uint8_t* s_pro = smem + kDgSmPro;
  uint8_t* s_ring = smem + kDgSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kDgSmAux) - kDgAuxFirst;   // indexed with AuxOffsets (>= w_sigma)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kDgSmBars);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kDgStages;
  uint64_t* bar_a = bars + 2 * kDgStages;     // [kDgNumBarA]
  uint64_t* bar_acc = bar_a + kDgNumBarA;     // [2]
  uint64_t* bar_free = bar_acc + 2;           // [kDgNumBarFree]
  uint64_t* bar_s = bar_free + kDgNumBarFree;  // [kDgNumBarS]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_s + kDgNumBarS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kMmaLayers = kFull ? 9 : 7;
  constexpr int kFirstChunk = kFull ? 0 : 6;  // sigma-only starts at W7^T
  constexpr int kChunks = kBwdChunks - kFirstChunk;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    for (int s = 0; s < kDgStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < kDgNumBarA; ++s) mbar_init(bar_a + s, kDgEpiWarps);
    mbar_init(bar_acc, 1);
    mbar_init(bar_acc + 1, 1);
    for (int s = 0; s < kDgNumBarFree; ++s) mbar_init(bar_free + s, 1);
    for (int s = 0; s < kDgNumBarS; ++s) mbar_init(bar_s + s, kDgEpiWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  if (warp >= 2 && warp < 2 + kDgEpiWarps) {
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = kDgAuxFirst + threadIdx.x - 64; i < AuxOffsets::count; i += kDgEpiThreads) s_aux[i] = gaux[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      const uint64_t keep = l2_policy_evict_last();   // the weights stay in L2 under the G-image write stream
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunks; ++ci) {
          mbar_wait(bar_empty + stage, ph ^ 1);
          mbar_arrive_expect_tx(bar_full + stage, kChunkBytes256);
          bulk_g2s(s_ring + stage * kChunkBytes256,
                   a.packed + kPackedBwdOff + (size_t)(kFirstChunk + ci) * kChunkBytes256, kChunkBytes256,
                   bar_full + stage, keep);
          if (++stage == kDgStages) { stage = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // the whole warp runs the loop with warp-uniform values; one elected lane issues the tcgen05 instructions (see mlp_tc.cu)
    {
      uint32_t stage = 0, ph = 0, a_par = 0, g = 0;   // g: running layer counter (accumulators alternate across tiles too)
      const uint32_t idesc = make_idesc_bf16(128, 256, 0, 0);
      const uint64_t act_desc = make_desc_sw128(smem_u32(s_act), 16, 1024);
      const uint64_t pro_desc = make_desc_sw128(smem_u32(s_pro), 16, 1024);
      const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
      // weight stages are awaited ahead of use, while the thread would otherwise idle waiting for the epilogue
      int ready = 0;
      auto prewait = [&](int n) {
        uint32_t st = stage, p = ph;
        for (int i = 0; i < ready; ++i) { if (++st == kDgStages) { st = 0; p ^= 1; } }
        for (; ready < n; ++ready) {
          mbar_wait(bar_full + st, p);
          if (++st == kDgStages) { st = 0; p ^= 1; }
        }
      };
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
#pragma unroll 1
        for (int i = 0; i < kMmaLayers; ++i, ++g) {
          const bool from_pro = kFull && i == 0;
          const int nchunks = from_pro ? 2 : 4;
          const uint32_t d_tmem = tmem_base + (g & 1u) * 256;
          prewait(from_pro ? 2 : 3);
#pragma unroll 1
          for (int c = 0; c < nchunks; ++c) {
            const int src = from_pro ? kDgBarPro : c;
            if (from_pro ? c == 0 : (c == dg_first(c))) {
              const int b = from_pro ? src : dg_last(c);
              mbar_wait(bar_a + b, (a_par >> b) & 1u);
              a_par ^= 1u << b;
            }
            if (ready == 0) mbar_wait(bar_full + stage, ph); else --ready;
            tc_fence_after();
            const uint64_t a_desc = (from_pro ? pro_desc : act_desc) + (uint64_t)(c * (kActChunkBytes >> 4));
            const uint64_t b_desc = ring_desc + (uint64_t)(stage * (kChunkBytes256 >> 4));
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (c | k) != 0);
              umma_commit(bar_empty + stage);
            }
            __syncwarp();
            if (++stage == kDgStages) { stage = 0; ph ^= 1; }
          }
          if (elect_one()) umma_commit(bar_acc + (g & 1u));
          __syncwarp();
        }
      }
    }
  } else if (warp < 2 + kDgEpiWarps) {
    const int et = threadIdx.x - 64;
    const int q = warp & 3, sub = (warp - 2) >> 2;   // TMEM lane quarter; 16-column quarter of every 64-column chunk
    const int row = q * 32 + lane;
    const RowSwz rsw(row);
    const uint32_t act_s = smem_u32(s_act), pro_s = smem_u32(s_pro);
    const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma), wrgb_s = smem_u32(s_aux + AuxOffsets::w_rgb);
    uint32_t g = 0;   // running layer counter, mirrors the UMMA issuer
    // Every image written to s_act / s_pro is also copied to the G stash by the store warp; a buffer may be overwritten
    // once that copy has read it (bar_free).  bit i of `stored`: a copy of buffer i has been requested.
    uint32_t free_par = 0, stored = 0;
    auto acquire = [&](int buf) {
      if ((stored >> buf) & 1u) {
        mbar_wait(bar_free + buf, (free_par >> buf) & 1u);
        free_par ^= 1u << buf;
      }
      stored |= 1u << buf;
    };

    auto signal = [&](int idx) {
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a + idx);
    };
    // first A operand of a tile's chain, on CUDA cores.  Returns d_sigma of this thread's row.
    //   full mode : G9 -> s_pro (2 chunks), also writes dpre and the G9 image
    //   sigma-only: G7 = d_sigma * w_sigma (x) relu'(h7) -> s_act (4 chunks), also writes the G7 image
    auto prologue = [&](int tile) -> float {
      const uint32_t* mask_tile = a.masks + (size_t)tile * mask_tile_words(kFull);
      const int64_t grow = (int64_t)tile * kTileM + row;
      const bool valid = grow < a.n_rows;
      const float dsig = valid ? __ldg(a.d_sigma + grow) : 0.f;
      if (kFull) {
        float dp[3] = {0.f, 0.f, 0.f};
        if (valid) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const float y = __ldg(a.rgb_out + grow * 3 + c);
            dp[c] = __ldg(a.d_rgb + grow * 3 + c) * y * (1.0f - y);  // sigmoid' (nerf.py:79-81)
          }
        }
        if (sub == 0) *reinterpret_cast<float4*>(a.dpre + ((size_t)tile * kTileM + row) * 4) = make_float4(dp[0], dp[1], dp[2], 0.f);
        const uint32_t m = __ldcs(mask_tile + 8 * 128 * 8 + (size_t)row * 4 + sub);   // read once: streaming
        acquire(kDgBarPro);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const int col = c * 64 + sub * 16 + j;
            const float4 w0 = ld_shared_f4(wrgb_s + col * 4), w1 = ld_shared_f4(wrgb_s + (128 + col) * 4),
                         w2 = ld_shared_f4(wrgb_s + (256 + col) * 4);
            float f0 = dp[0] * w0.x + dp[1] * w1.x + dp[2] * w2.x;
            float f1 = dp[0] * w0.y + dp[1] * w1.y + dp[2] * w2.y;
            float f2 = dp[0] * w0.z + dp[1] * w1.z + dp[2] * w2.z;
            float f3 = dp[0] * w0.w + dp[1] * w1.w + dp[2] * w2.w;
            f0 = mask_bit16(m, c, j) ? f0 : 0.f;
            f1 = mask_bit16(m, c, j + 1) ? f1 : 0.f;
            f2 = mask_bit16(m, c, j + 2) ? f2 : 0.f;
            f3 = mask_bit16(m, c, j + 3) ? f3 : 0.f;
            pk[j >> 1] = cvt_bf16x2(f0, f1);
            pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
          }
          const uint32_t chunk_s = pro_s + (uint32_t)c * kActChunkBytes;
          st_shared_v4(chunk_s + rsw.piece(sub * 2), pk[0], pk[1], pk[2], pk[3]);
          st_shared_v4(chunk_s + rsw.piece(sub * 2 + 1), pk[4], pk[5], pk[6], pk[7]);
        }
        signal(kDgBarPro);
      } else {
        const uint2 mw2 = __ldcs(reinterpret_cast<const uint2*>(mask_tile + ((size_t)7 * 128 + row) * 8 + sub * 2));
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t m = (c < 2) ? mw2.x : mw2.y;
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 w = ld_shared_f4(wsig_s + (c * 64 + sub * 16 + j) * 4);
            const float f0 = mask_bit16(m, c, j) ? dsig * w.x : 0.f, f1 = mask_bit16(m, c, j + 1) ? dsig * w.y : 0.f;
            const float f2 = mask_bit16(m, c, j + 2) ? dsig * w.z : 0.f, f3 = mask_bit16(m, c, j + 3) ? dsig * w.w : 0.f;
            pk[j >> 1] = cvt_bf16x2(f0, f1);
            pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
          }
          const uint32_t chunk_s = act_s + (uint32_t)c * kActChunkBytes;
          if (c == dg_first(c)) acquire(dg_last(c));
          st_shared_v4(chunk_s + rsw.piece(sub * 2), pk[0], pk[1], pk[2], pk[3]);
          st_shared_v4(chunk_s + rsw.piece(sub * 2 + 1), pk[4], pk[5], pk[6], pk[7]);
          if (c == dg_last(c)) signal(c);
        }
      }
      return dsig;
    };

    int tile = blockIdx.x;
    float dsig = 0.f, dsig_next = 0.f;
    if (kFull && tile < a.num_tiles) dsig_next = prologue(tile);
    for (; tile < a.num_tiles; tile += gridDim.x) {
      const int next = tile + gridDim.x;
      const uint32_t* mask_tile = a.masks + (size_t)tile * mask_tile_words(kFull);
      if (kFull) {
        dsig = dsig_next;
      } else {
        dsig = prologue(tile);
      }
#pragma unroll 1
      for (int i = 0; i < kMmaLayers; ++i, ++g) {
        const int t = (kFull ? 8 : 6) - i;  // this epilogue produces G_t
        uint2 mw = make_uint2(0xffffffffu, 0xffffffffu);
        if (t < 8)   // issued before the wait so that the load latency hides behind the MMAs
          mw = __ldcs(reinterpret_cast<const uint2*>(mask_tile + ((size_t)t * 128 + row) * 8 + sub * 2));
        const uint32_t b = g & 1u;
        mbar_wait(bar_acc + b, (g >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + b * 256;
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
          const uint32_t m = (c < 2) ? mw.x : mw.y;
          if (c == dg_first(c)) acquire(dg_last(c));
          if (kFull && t == 7) epi_dgrad16<true>(v, c, c0, m, dsig, wsig_s, rsw, chunk_s, sub * 2);
          else epi_dgrad16<false>(v, c, c0, m, dsig, wsig_s, rsw, chunk_s, sub * 2);
          if (c != dg_last(c)) {
            // hand-over per group of chunks
          } else if (t > 0) {
            signal(c);   // the UMMA issuer and the store warp take it from here
          } else {       // G0: the store warp only
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_s + c);
          }
        }
        // full mode: s_pro is free once this tile's first GEMM is done -> stage the next tile's G9 now
        if (kFull && i == 0 && next < a.num_tiles) dsig_next = prologue(next);
      }
    }
  } else if (lane == 0) {
    // ===================== G-image store warp =====================
    // One thread follows the epilogue's production order: as soon as an operand image is complete (bar_a) it is
    // bulk-copied to the G stash; once the copy has read the buffer (one copy behind, so two are in flight) the
    // epilogue may overwrite it (bar_free).
    uint32_t s_par = 0;
    int pending = -1;
    const uint64_t stream_pol = l2_policy_evict_first();
    auto store = [&](int idx, bool own_bar, const uint8_t* src, uint8_t* dst, uint32_t bytes) {
      const int pb = own_bar ? 8 + idx : idx;
      uint64_t* bar = own_bar ? bar_s + idx : bar_a + idx;
      const uint32_t par = (s_par >> pb) & 1u;
      // about to block: hand the pending buffer back first (the epilogue may need it to produce this very image)
      if (pending >= 0 && !mbar_try_wait(bar, par)) {
        bulk_wait_read0();
        mbar_arrive(bar_free + pending);
        pending = -1;
      }
      mbar_wait(bar, par);
      s_par ^= 1u << pb;
      bulk_s2g(dst, src, bytes, stream_pol);
      bulk_commit();
      if (pending >= 0) {
        bulk_wait_read1();
        mbar_arrive(bar_free + pending);
      }
      pending = idx;
    };
    auto store_act = [&](int slot, int tile) {
      uint8_t* dst = a.gstash + img_off(slot, 4, tile, a.num_tiles);
      for (int c = 0; c < 4; c = dg_last(c) + 1)
        store(dg_last(c), slot == 0, s_act + c * kActChunkBytes, dst + (size_t)c * kActChunkBytes,
              (uint32_t)(dg_last(c) - c + 1) * kActChunkBytes);
    };
    int tile = blockIdx.x;
    if (kFull && tile < a.num_tiles)
      store(kDgBarPro, false, s_pro, a.gstash + img_off(kGstashG9, 2, tile, a.num_tiles), 2 * kActChunkBytes);
    for (; tile < a.num_tiles; tile += gridDim.x) {
      const int next = tile + gridDim.x;
      if (!kFull) store_act(4 * 7, tile);
#pragma unroll 1
      for (int i = 0; i < kMmaLayers; ++i) {
        const int t = (kFull ? 8 : 6) - i;
        store_act(t == 8 ? kGstashG8 : 4 * t, tile);
        if (kFull && i == 0 && next < a.num_tiles)
          store(kDgBarPro, false, s_pro, a.gstash + img_off(kGstashG9, 2, next, a.num_tiles), 2 * kActChunkBytes);
      }
    }
    bulk_wait_all0();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <bool kFull>
static int launch_dgrad_t(const DgradArgs& a, cudaStream_t st) {
  const int smem_bytes = (int)kDgSmTotal;
  NRF_OPT_IN_SMEM(mlp_dgrad_kernel<kFull>, smem_bytes);
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  ProfScope prof(kFull ? NRF_K_DGRAD_FULL : NRF_K_DGRAD_SIGMA, st, a.n_rows);
  mlp_dgrad_kernel<kFull><<<a.num_tiles < sms ? a.num_tiles : sms, kDgThreads, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

int launch_dgrad(bool full, const uint8_t* packed, const uint32_t* masks, const float* d_sigma, const float* d_rgb,
                 const float* rgb_out, int64_t n_rows, int num_tiles, uint8_t* gstash, float* dpre, cudaStream_t st) {
  DgradArgs a{packed, masks, d_sigma, d_rgb, rgb_out, n_rows, num_tiles, gstash, dpre};
  return full ? launch_dgrad_t<true>(a, st) : launch_dgrad_t<false>(a, st);
}

}  // namespace tc
}  // namespace nrf
