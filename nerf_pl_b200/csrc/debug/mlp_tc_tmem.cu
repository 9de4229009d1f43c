// bf16 tensor-core NeRF MLP, forward, hidden activations as TENSOR-MEMORY operands (models/nerf.py:83-124 fused with the
// positional encoding nerf.py:21-38 and the point generation rendering.py:234).  Same maths, stash and mask layout and
// per-element summation order as mlp_tc.cu (bit-identical outputs in all four modes, tools/ts_check.py with TS_IMPL=m).
//
// EXPERIMENTAL (debug builds, NRF_FWD_IMPL=m), and a negative result.  Hypothesis: the training forward is bound by how
// fast its stash stores drain - 64 KB of operand image per layer leave through bulk stores that read the activation
// tile in shared memory, and while the tensor pipe reads the SAME tile as its A operand those stores get 8 B/clk/SM, 16
// once the pipe idles (tools/mma_contend.py, modes 128 / 129); 2338 cycles of MMAs at 8 B/clk + the other 45 KB at
// 16 B/clk = 5170 cycles per layer would explain the measured training period (5260 against 3744 in inference).  With
// the A operand in tensor memory the MMAs do not touch the tile.  Measured: training forward 0.767 ms against 0.764 ms,
// inference 0.58 ms either way - the store drain is not what paces the training forward.
//
//   The epilogue writes every hidden activation with tcgen05.st into one of two 128-column A buffers in TMEM (the next
//   layer's operand; tcgen05.mma with the A operand in tensor memory) and, in training, with st.shared into the tile
//   (the stash store's source, and nothing else).  TMEM = ONE 256-column fp32 accumulator + 2 x 128 operand columns, so
//   the accumulator cannot alternate between layers as in mlp_tc.cu: the epilogue first loads all of it into registers
//   and says so (bar_drained) before the next layer's MMAs may overwrite it - they could not start before the first
//   hand-over anyway.  Weights, ring, encodings (shared-memory operands of layers 0, 4, 9) and the store warp are those
//   of mlp_tc.cu.
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"
#include "tc_fwd.cuh"

namespace nrf {
namespace tc {
namespace tm {

constexpr int kStages = 4;
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreadsInfer = 64 + kEpiThreads;
constexpr int kThreadsTrain = kThreadsInfer + 32;

constexpr uint32_t kSmEnc = 0;
constexpr uint32_t kSmAct = kSmEnc + kActChunkBytes;
constexpr uint32_t kSmRing = kSmAct + 4 * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStages * kChunkBytes256;
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;   // [4][128][3] floats
constexpr uint32_t kSmBars = kSmScratch + 4 * 128 * 3 * 4;
constexpr uint32_t kSmTotal = kSmBars + 256;
static_assert(kSmTotal <= 232448, "shared memory budget");

// A-operand readiness: activation chunk groups {0,1} -> 1, {2,3} -> 3; xyz encoding; direction encoding
constexpr int kBarEnc = 4, kBarDir = 5, kNumBarA = 6;
constexpr int kFreeEnc = 4, kNumBarFree = 5;
constexpr int kNumBarS = 4;
constexpr uint32_t kTmAcc = 0, kTmA = 256;   // TMEM columns: accumulator; [2 buffers][4 chunks][32 columns], layer L reads L & 1

template <bool kFull, bool kTrain>
__global__ void __launch_bounds__(kTrain ? kThreadsTrain : kThreadsInfer, 1) mlp_fwd_tmem_kernel(const __grid_constant__ FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_act = smem + kSmAct;
  uint8_t* s_ring = smem + kSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;                   // [kStages] weight chunk landed
  uint64_t* bar_empty = bars + kStages;        // [kStages] weight chunk consumed
  uint64_t* bar_a = bars + 2 * kStages;        // [kNumBarA] A operand written (epilogue -> MMA, store warp)
  uint64_t* bar_acc = bar_a + kNumBarA;        // accumulator complete (MMA -> epilogue)
  uint64_t* bar_drained = bar_acc + 1;         // accumulator read out into registers (epilogue -> MMA)
  uint64_t* bar_free = bar_drained + 1;        // [kNumBarFree] stash store finished reading a buffer
  uint64_t* bar_s = bar_free + kNumBarFree;    // [kNumBarS] store-only image written
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_s + kNumBarS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kLayers = kFull ? 10 : 8;
  constexpr int kChunksPerTile = kFull ? kFwdChunksFull : kFwdChunksSigma;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < kNumBarA; ++s) mbar_init(bar_a + s, kEpiWarps);
    mbar_init(bar_acc, 1);
    mbar_init(bar_drained, kEpiWarps);
    for (int s = 0; s < kNumBarFree; ++s) mbar_init(bar_free + s, 1);
    for (int s = 0; s < kNumBarS; ++s) mbar_init(bar_s + s, kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
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
    // ===================== weight producer (as mlp_tc.cu) =====================
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      const uint64_t keep = l2_policy_evict_last();
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunksPerTile; ++ci) {
          mbar_wait(bar_empty + stage, ph ^ 1);
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
    // ===================== UMMA issuer (whole warp, warp-uniform; one elected lane issues) =====================
    uint32_t stage = 0, ph = 0;
    uint32_t a_par = 0, d_par = 0;
    const uint32_t idesc256 = make_idesc_bf16(128, 256, 0, 0), idesc128 = make_idesc_bf16(128, 128, 0, 0);
    const uint64_t enc_desc = make_desc_sw128(smem_u32(s_enc), 16, 1024);
    const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
    const uint32_t d_tmem = tmem_base + kTmAcc;
    auto wait_a = [&](int src) {
      mbar_wait(bar_a + src, (a_par >> src) & 1u);
      a_par ^= 1u << src;
    };
    int ready = 0;   // upcoming ring stages already known to be full (awaited while the operands are not there yet)
    auto prewait = [&](int n) {
      uint32_t st = stage, p = ph;
      for (int i = 0; i < ready; ++i) { if (++st == kStages) { st = 0; p ^= 1; } }
      for (; ready < n; ++ready) {
        mbar_wait(bar_full + st, p);
        if (++st == kStages) { st = 0; p ^= 1; }
      }
    };
    // one K chunk: four MMAs, A from the encoding buffer in shared memory or from tensor-memory columns a_tm .. a_tm + 31
    auto chunk = [&](bool smem_a, uint32_t a_tm, uint32_t idesc, bool first) {
      if (ready == 0) mbar_wait(bar_full + stage, ph); else --ready;
      tc_fence_after();
      const uint64_t b_desc = ring_desc + (uint64_t)(stage * (kChunkBytes256 >> 4));
      if (smem_a) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss(d_tmem, enc_desc + 2 * k, b_desc + 2 * k, idesc, (!first || k) ? 1u : 0u);
          umma_commit(bar_empty + stage);
        }
      } else {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ts(d_tmem, a_tm + (uint32_t)(k * 8), b_desc + 2 * k, idesc, (!first || k) ? 1u : 0u);
          umma_commit(bar_empty + stage);
        }
      }
      __syncwarp();
      if (++stage == kStages) { stage = 0; ph ^= 1; }
    };
    bool have_acc = false;   // an earlier layer's accumulator may still be in the registers-to-be: wait for its drain
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
#pragma unroll 1
      for (int L = 0; L < kLayers; ++L) {
        const uint32_t idesc = (L == 9) ? idesc128 : idesc256;
        const uint32_t a_tm = tmem_base + kTmA + (uint32_t)(L & 1) * 128;
        bool first = true;
        if (L == 0 || L == 4) {          // the xyz encoding leads (skip connection: models/nerf.py:96-97)
          if (L == 0) wait_a(kBarEnc);   // at layer 4 the same completion is still current
        }
        if (L != 0) { prewait(3); wait_a(1); }   // the first activation group of this layer's operand
        // the single accumulator: the previous layer's values must be in the epilogue's registers before they are overwritten
        if (have_acc) {
          mbar_wait(bar_drained, d_par);
          d_par ^= 1u;
        }
        if (L == 0 || L == 4) {
          chunk(true, 0, idesc, true);
          first = false;
        }
        if (L != 0) {
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            if (c == 2) wait_a(3);
            chunk(false, a_tm + (uint32_t)(c * 32), idesc, first);
            first = false;
          }
        }
        if (L == 9) {                    // the direction encoding has replaced the xyz encoding in s_enc
          wait_a(kBarDir);
          chunk(true, 0, idesc, false);
        }
        if (elect_one()) umma_commit(bar_acc);
        __syncwarp();
        have_acc = true;
      }
    }
  } else if (warp < 2 + kEpiWarps) {
#include "mlp_tc_tmem_epilogue.inc"
  } else if (kTrain && lane == 0) {
#include "mlp_tc_tmem_store.inc"
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace tm

template <bool kFull, bool kTrain>
static int launch_fwd_tmem_t(const FwdArgs& a, cudaStream_t st) {
  const int smem_bytes = (int)tm::kSmTotal;
  NRF_OPT_IN_SMEM((tm::mlp_fwd_tmem_kernel<kFull, kTrain>), smem_bytes);
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = a.num_tiles < sms ? a.num_tiles : sms;
  ProfScope prof((kTrain ? NRF_K_FWD_SIGMA_TRAIN : NRF_K_FWD_SIGMA_INFER) + (kFull ? 1 : 0), st, a.n_rows);
  tm::mlp_fwd_tmem_kernel<kFull, kTrain><<<grid, kTrain ? tm::kThreadsTrain : tm::kThreadsInfer, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

int launch_fwd_tmem(bool full, bool train, const FwdArgs& a, cudaStream_t st) {
  if (full) return train ? launch_fwd_tmem_t<true, true>(a, st) : launch_fwd_tmem_t<true, false>(a, st);
  return train ? launch_fwd_tmem_t<false, true>(a, st) : launch_fwd_tmem_t<false, false>(a, st);
}

}  // namespace tc
}  // namespace nrf
