// bf16 tensor-core NeRF MLP, forward, activations in TENSOR MEMORY (models/nerf.py:83-124 fused with the positional
// encoding nerf.py:21-38 and the point generation rendering.py:234).  Same maths, stash and mask layout, and the same
// per-element summation order as mlp_tc.cu (bit-identical outputs); what changes is where the layer chain lives:
//
//   * The hidden activations never go through shared memory on their way to the next GEMM: the epilogue writes them
//     with tcgen05.st into one of two 128-column A-operand buffers in TMEM and the next layer's MMAs read them from
//     there (tcgen05.mma with the A operand in tensor memory).  Only the encodings (layers 0, 4, 9) are shared-memory
//     operands.  TMEM: accumulator 256 columns + 2 x 128 columns of bf16 A operand = 512.
//   * Every layer is issued as two N = 128 halves.  While the tensor pipe works on the second half, the epilogue
//     warps already turn the first half's accumulator into chunks 0, 1 of the next operand - written into the OTHER
//     A buffer, so nothing the running MMAs read is overwritten (in shared memory that needed a second 64 KB tile).
//     One tile per SM then keeps the tensor pipe busy through most of what used to be the accumulator hand-over.
//   * A weight stage is {K-chunk pair, N half}: two 16 KB [128 x 64] sub-chunks, each used by exactly one group of
//     MMAs, so the ring streams (no stage has to stay resident for the other half).
//
// EXPERIMENTAL (debug builds, NRF_FWD_IMPL=ts): bit-identical to mlp_tc.cu in all four modes (tools/ts_check.py) but
// SLOWER: 0.69 ms against 0.57 ms (inference, 524 288 rows), 0.79 against 0.75 ms (training).  Measured cause: a thread
// that issues N = 128 MMAs (64 cycles of tensor pipe each) needs an issue slot every few cycles, and the four busy
// epilogue warps of its SM sub-partition take them - tools/mma_contend.py: 64 cycles per MMA alone, 90-100 with
// arithmetic-bound warps on the sub-partition, 130-160 with this kernel's issue loop (N = 256 MMAs keep 128 either way).
// Overlapping the epilogue with the MMAs is exactly what makes those warps busy while the issuer needs the slots.
//
//   warp 0      weight producer: bulk copies of the 16 KB half chunks, L2 -> 4-stage ring (32 KB stages)
//   warp 1      UMMA issuer: kind::f16, M = 128, N = 128, K = 16 x 4 per sub-chunk
//   warps 2-17  epilogue (4 warps per TMEM lane quarter, 16 columns of every 64-column chunk each)
//   warp 18     training: stash store warp (operand images: shared memory -> HBM)
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"
#include "tc_fwd.cuh"

namespace nrf {
namespace tc {
namespace ts {

// waits on the layer chain's critical path poll (mbarrier.test_wait) instead of suspending: a suspended try_wait wakes
// up hundreds of cycles after the phase completes
#ifndef NRF_TS_SPIN
#define NRF_TS_SPIN 0   // measured: 16 polling epilogue warps slow the issuer down (0.74 -> 0.89 ms)
#endif
#if NRF_TS_SPIN
#define NRF_TS_WAIT(bar, par) mbar_wait_spin(bar, par)
#else
#define NRF_TS_WAIT(bar, par) mbar_wait(bar, par)
#endif

// Weight ring: one stage is in flight (L2 latency + transfer, ~2100 cycles for 32 KB) or held (landed, waiting for its
// MMAs, being read) at any time, so the ring's BYTES set the weight supply: 128 KB per ~3900-cycle round trip = 33 B/clk
// paces one tile per SM at ~3900 cycles per 128 KB layer, whatever the tensor pipe could do (tools/ts_trace.py).
// Inference does not need the 64 KB activation tile in shared memory (the operand lives in tensor memory), so its ring
// takes that space: 6 stages.  Training keeps the tile (the stash store's source) and 4 stages.
constexpr int kStagesTrain = 4, kStagesInfer = 6, kStagesMax = 6;
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32;
#ifndef NRF_TS_ISSUER_WARP
#define NRF_TS_ISSUER_WARP 1   // 18 (experiment): the UMMA issuer is the YOUNGEST warp of its SM sub-partition instead of the oldest
#endif
constexpr int kIssuerWarp = NRF_TS_ISSUER_WARP;
constexpr int kStoreWarp = kIssuerWarp == 18 ? 19 : 18;
constexpr int kThreadsInfer = (kIssuerWarp == 18 ? 19 : 18) * 32;
constexpr int kThreadsTrain = (kStoreWarp + 1) * 32;
constexpr uint32_t kSub = kChunkBytes128;   // 16 KB: [128 weight rows][64 K] = one N half of a chunk (a whole chunk of layer 9)
constexpr uint32_t kStageBytes = 2 * kSub;

// shared-memory map (1024-byte aligned).  s_act is written in training only (the stash store's source).
constexpr uint32_t kSmEnc = 0;
constexpr uint32_t kSmAct = kSmEnc + kActChunkBytes;
constexpr uint32_t kSmRing = kSmAct + 4 * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStagesTrain * kStageBytes;   // inference: the ring starts at kSmAct and ends here too
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;   // [4][128][3] floats
constexpr uint32_t kSmBars = kSmScratch + 4 * 128 * 3 * 4;
constexpr uint32_t kSmTotal = kSmBars + 256;
static_assert(kSmTotal <= 232448, "shared memory budget");

// A-operand readiness: activation chunk groups {0,1} -> 1, {2,3} -> 3; xyz encoding; direction encoding
constexpr int kBarEnc = 4, kBarDir = 5, kNumBarA = 6;
constexpr int kFreeEnc = 4, kNumBarFree = 5;   // training: "the stash store has read the buffer"
constexpr int kNumBarS = 4;                    // training: images only the store warp reads

// TMEM columns
constexpr uint32_t kTmAcc = 0;      // [2 halves][128]
constexpr uint32_t kTmA = 256;      // [2 buffers][4 chunks][32 columns]: layer L reads buffer L & 1

// K chunks of layer L in the packed forward blob (consumption order, tc_layout.cuh)
__device__ __forceinline__ int layer_first_chunk(int L) { return L == 0 ? 0 : (L <= 4 ? 1 + 4 * (L - 1) : 2 + 4 * (L - 1)); }
__device__ __forceinline__ int layer_chunks(int L) { return L == 0 ? 1 : ((L == 4 || L == 9) ? 5 : 4); }

template <bool kFull, bool kTrain>
__global__ void __launch_bounds__(kTrain ? kThreadsTrain : kThreadsInfer, 1) mlp_fwd_ts_kernel(const __grid_constant__ FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_act = smem + kSmAct;
  constexpr int kStages = kTrain ? kStagesTrain : kStagesInfer;
  uint8_t* s_ring = smem + (kTrain ? kSmRing : kSmAct);
  static_assert(kSmAct + kStagesInfer * kStageBytes <= kSmAux, "inference ring");
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;                   // [kStages] weight stage landed
  uint64_t* bar_empty = bars + kStagesMax;     // [kStages] weight stage consumed
  uint64_t* bar_a = bars + 2 * kStagesMax;     // [kNumBarA] A operand written (epilogue -> MMA, store warp)
  uint64_t* bar_acc = bar_a + kNumBarA;        // [2] accumulator half complete (MMA -> epilogue)
  uint64_t* bar_drained = bar_acc + 2;         // [2] a tile's LAST use of an accumulator half has been read out
  uint64_t* bar_free = bar_drained + 2;        // [kNumBarFree] stash store finished reading a buffer
  uint64_t* bar_s = bar_free + kNumBarFree;    // [kNumBarS] store-only image written
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_s + kNumBarS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kLayers = kFull ? 10 : 8;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < kNumBarA; ++s) mbar_init(bar_a + s, kEpiWarps);
    for (int s = 0; s < 2; ++s) { mbar_init(bar_acc + s, 1); mbar_init(bar_drained + s, kEpiWarps); }
    for (int s = 0; s < kNumBarFree; ++s) mbar_init(bar_free + s, 1);
    for (int s = 0; s < kNumBarS; ++s) mbar_init(bar_s + s, kEpiWarps);
    fence_barrier_init();
  }
  if (warp == kIssuerWarp) tmem_alloc(s_tmem, 512);
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
      const uint64_t keep = l2_policy_evict_last();
      const uint8_t* blob = a.packed + kPackedFwdOff;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int L = 0; L < kLayers; ++L) {
          const int c0 = layer_first_chunk(L), n = layer_chunks(L), nh = (L == 9) ? 1 : 2;
          for (int h = 0; h < nh; ++h) {
            for (int j = 0; j < n; j += 2) {
              const int ns = (j + 1 < n) ? 2 : 1;
              mbar_wait(bar_empty + stage, ph ^ 1);
              mbar_arrive_expect_tx(bar_full + stage, ns * kSub);
              for (int i = 0; i < ns; ++i) {
                const uint8_t* src = (L == 9) ? blob + (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 + (size_t)(j + i) * kSub
                                              : blob + (size_t)(c0 + j + i) * kChunkBytes256 + (size_t)h * kSub;
                bulk_g2s(s_ring + stage * kStageBytes + i * kSub, src, kSub, bar_full + stage, keep);
              }
              if (++stage == kStages) { stage = 0; ph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == kIssuerWarp) {
    // ===================== UMMA issuer (whole warp, warp-uniform; one elected lane issues) =====================
    uint32_t stage = 0, ph = 0;
    uint32_t a_par = 0, d_par = 0;
    const uint32_t idesc = make_idesc_bf16(128, 128, 0, 0);
    const uint64_t enc_desc = make_desc_sw128(smem_u32(s_enc), 16, 1024);
    const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
    auto wait_a = [&](int src) {
      if (!(NRF_FWD_DEBUG && (a.dbg_flags & 64)))     // 64: timing experiment, operands used without waiting for them
        NRF_TS_WAIT(bar_a + src, (a_par >> src) & 1u);
      a_par ^= 1u << src;
    };
    bool first_tile = true;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, first_tile = false) {
#pragma unroll 1
      for (int L = 0; L < kLayers; ++L) {
        const int n = layer_chunks(L), nh = (L == 9) ? 1 : 2;
        const bool mtr = NRF_FWD_DEBUG && blockIdx.x == 0 && lane == 0 && tile == 2 * (int)gridDim.x;
        const uint32_t a_tm = tmem_base + kTmA + (uint32_t)(L & 1) * 128;
#pragma unroll 1
        for (int h = 0; h < nh; ++h) {
          // layer 9 (N = 128) accumulates in half 1: the next tile's layer 0 starts in half 0 while the heads still read it
          const int hacc = (L == 9) ? 1 : h;
          const uint32_t d_tmem = tmem_base + kTmAcc + (uint32_t)hacc * 128;
          // the previous tile's last reader of this accumulator half (sigma-only: layer 7, both halves; full: layer 9, half 1)
          if (L == 0 && !first_tile && (!kFull || h == 1)) {
            NRF_TS_WAIT(bar_drained + h, (d_par >> h) & 1u);
            d_par ^= 1u << h;
          }
#pragma unroll 1
          for (int j = 0; j < n; j += 2) {
            const int ns = (j + 1 < n) ? 2 : 1;
            // A operands of this stage's K chunks: known complete before the stage's MMAs are issued
            if (h == 0) {
              for (int i = 0; i < ns; ++i) {
                const int kc = j + i;
                const bool is_enc = (L == 0) || (L == 4 && kc == 0);
                const bool is_dir = (L == 9 && kc == 4);
                if (is_enc) { if (L == 0) wait_a(kBarEnc); }         // at layer 4 the same completion is still current
                else if (is_dir) wait_a(kBarDir);
                else {
                  const int c = (L == 4) ? kc - 1 : kc;
                  if (!(c & 1)) wait_a(c | 1);
                }
              }
            }
            if (j == 2) NRF_WTRACE(mtr, 17, L, 2 * h);       // operands of the half's SECOND stage known complete (previous stage issued)
            if (!(NRF_FWD_DEBUG && (a.dbg_flags & 32)))   // 32: timing experiment, weights used without waiting for them
              NRF_TS_WAIT(bar_full + stage, ph);
            if (j == 2) NRF_WTRACE(mtr, 17, L, 2 * h + 1);   // ... and its weights landed
            tc_fence_after();
            if (h == 0) NRF_WTRACE(mtr, 16, L, j ? 1 : 0);   // layer's first / second stage of half 0: operands and weights there
            // Issue: the common stage (two K chunks, both from tensor memory) goes out as ONE group of eight MMAs - the
            // chunks are adjacent in the A buffer (32 columns apart) and in the stage (16 KB apart).  The issuer warp
            // shares its SM sub-partition with four busy epilogue warps: every instruction it does not execute counts.
            const int kc0 = j, c_first = (L == 4) ? kc0 - 1 : kc0;
            const bool enc0 = (L == 0) || (L == 4 && kc0 == 0), dir_last = (L == 9 && j + ns - 1 == 4);
            const uint64_t b0 = ring_desc + (uint64_t)((stage * kStageBytes) >> 4);
            if (!(NRF_FWD_DEBUG && (a.dbg_flags & 1))) {
              if (ns == 2 && !enc0 && !dir_last) {
                const uint32_t a0 = a_tm + (uint32_t)(c_first * 32);
                const uint32_t acc0 = kc0 ? 1u : 0u;
                if (elect_one()) {
#pragma unroll
                  for (int k = 0; k < 8; ++k)
                    umma_ts(d_tmem, a0 + (uint32_t)(k * 8), b0 + (uint64_t)((k >> 2) * (kSub >> 4) + 2 * (k & 3)), idesc, k ? 1u : acc0);
                }
              } else {
                for (int i = 0; i < ns; ++i) {
                  const int kc = j + i;
                  const bool smem_a = (L == 0) || (L == 4 && kc == 0) || (L == 9 && kc == 4);
                  const uint32_t a_addr = a_tm + (uint32_t)(((L == 4) ? kc - 1 : kc) * 32);
                  const uint64_t b_desc = b0 + (uint64_t)((i * kSub) >> 4);
                  const uint32_t acc0 = kc ? 1u : 0u;
                  if (smem_a) {
                    if (elect_one()) {
#pragma unroll
                      for (int k = 0; k < 4; ++k) umma_ss(d_tmem, enc_desc + 2 * k, b_desc + 2 * k, idesc, k ? 1u : acc0);
                    }
                  } else {
                    if (elect_one()) {
#pragma unroll
                      for (int k = 0; k < 4; ++k) umma_ts(d_tmem, a_addr + (uint32_t)(k * 8), b_desc + 2 * k, idesc, k ? 1u : acc0);
                    }
                  }
                }
              }
            }
            if (elect_one()) umma_commit(bar_empty + stage);
            __syncwarp();
            if (++stage == kStages) { stage = 0; ph ^= 1; }
          }
          if (elect_one()) umma_commit(bar_acc + hacc);
          __syncwarp();
          NRF_WTRACE(mtr, 16, L, 2 + h);                    // half h's accumulator commit issued
        }
      }
    }
  } else if (warp >= 2 && warp < 2 + kEpiWarps) {
#include "mlp_tc_ts_epilogue.inc"
  } else if (kTrain && warp == kStoreWarp && lane == 0) {
#include "mlp_tc_ts_store.inc"
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace ts

template <bool kFull, bool kTrain>
static int launch_fwd_ts_t(const FwdArgs& a, cudaStream_t st) {
  const int smem_bytes = (int)ts::kSmTotal;
  NRF_OPT_IN_SMEM((ts::mlp_fwd_ts_kernel<kFull, kTrain>), smem_bytes);
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = a.num_tiles < sms ? a.num_tiles : sms;
  ProfScope prof((kTrain ? NRF_K_FWD_SIGMA_TRAIN : NRF_K_FWD_SIGMA_INFER) + (kFull ? 1 : 0), st, a.n_rows);
  ts::mlp_fwd_ts_kernel<kFull, kTrain><<<grid, kTrain ? ts::kThreadsTrain : ts::kThreadsInfer, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

int launch_fwd_ts(bool full, bool train, const FwdArgs& a, cudaStream_t st) {
  if (full) return train ? launch_fwd_ts_t<true, true>(a, st) : launch_fwd_ts_t<true, false>(a, st);
  return train ? launch_fwd_ts_t<false, true>(a, st) : launch_fwd_ts_t<false, false>(a, st);
}

}  // namespace tc
}  // namespace nrf
