// bf16 tensor-core NeRF MLP forward on CTA PAIRS (tcgen05 cta_group::2), two tiles in flight per CTA.
// Same maths, stash and mask layout as mlp_tc.cu (models/nerf.py:83-124 fused with nerf.py:21-38 and
// rendering.py:234); what changes is how the work is laid over the SMs:
//
//   * A cluster of two CTAs (two SMs of one TPC) runs every UMMA as M = 256: 128 sample rows from each CTA's
//     shared memory, the weight chunk split in two N halves, one half per CTA (16 KB instead of 32 KB per chunk
//     and SM).  One thread of the leader CTA issues the MMAs for both; completions are multicast to both CTAs.
//   * The halved weight ring frees room for the operands of TWO 128-row tiles per CTA ("slots").  The layer chain
//     of slot B runs half a step behind slot A: while the epilogue warps turn A's accumulator into A's next
//     operand, the tensor pipe works on B, and vice versa.  Each slot owns one 256-column TMEM accumulator.
//     (One tile per CTA leaves the tensor pipe idle during every accumulator hand-over: DESIGN.md §4.3.)
//
//   warp 0      weight producer: this CTA's half of every chunk, bulk copy L2 -> 3-stage ring
//   warp 1      leader: UMMA issuer.  peer: relay, forwards "my half has landed" to the leader's barrier
//   warps 2-17  epilogue, alternating between the slots (same thread mapping as mlp_tc.cu)
//   warp 18     training: stash store warp (whole images, one bulk store each)
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"
#include "tc_fwd.cuh"

namespace nrf {
namespace tc {

namespace pair {

#ifndef NRF_PAIR_HACK
#define NRF_PAIR_HACK 0   // 1: timing experiment only (both slots share ONE enc buffer -> wrong results, frees a 4th stage)
#endif
#ifndef NRF_PAIR_SPIN
#define NRF_PAIR_SPIN 0   // 1: the waits on the weight round trip (producer, relay, issuer) poll instead of suspending
#endif
constexpr int kStages = NRF_PAIR_HACK ? 4 : 3;
constexpr uint32_t kEncStride = NRF_PAIR_HACK ? 0 : kActChunkBytes;
constexpr uint32_t kHalfChunk = kChunkBytes256 / 2;   // 16 KB: this CTA's N half of a weight chunk
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreadsInfer = 64 + kEpiThreads;
constexpr int kThreadsTrain = kThreadsInfer + 32;

// shared-memory map (per CTA)
constexpr uint32_t kSmAct = 0;                                     // [2 slots][4 chunks]
constexpr uint32_t kSmEnc = kSmAct + 2 * 4 * kActChunkBytes;       // [2 slots]: xyz encoding, later direction encoding
constexpr uint32_t kSmRing = kSmEnc + (NRF_PAIR_HACK ? 1 : 2) * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStages * kHalfChunk;
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;    // [4][128][3] floats
constexpr uint32_t kSmBars = kSmScratch + 4 * 128 * 3 * 4;
constexpr uint32_t kSmTotal = kSmBars + 384;
static_assert(kSmTotal <= 232448, "shared memory budget");

// weight chunks of layer L in the packed forward blob (consumption order, tc_layout.cuh)
__device__ __forceinline__ int layer_first_chunk(int L) { return L == 0 ? 0 : (L <= 4 ? 1 + 4 * (L - 1) : 2 + 4 * (L - 1)); }
__device__ __forceinline__ int layer_chunks(int L) { return L == 0 ? 1 : ((L == 4 || L == 9) ? 5 : 4); }
// Ring stage loads of layer L (16 KB each; the direction layer's 8 KB half chunks go two to a stage), and whether
// ONE copy of the layer's weights serves both slots (it must fit the ring together: not the 5-chunk skip layer)
__device__ __forceinline__ int layer_loads(int L) { return L == 0 ? 1 : (L == 4 ? 5 : (L == 9 ? 3 : 4)); }
__device__ __forceinline__ bool layer_shared(int L) { return kStages >= 4 && L != 4; }

// buffers the stash store warp copies from (training): act of slot 0/1, enc of slot 0/1
constexpr int kBufAct = 0, kBufEnc = 2, kNumBuf = 4;

template <bool kFull, bool kTrain>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTrain ? kThreadsTrain : kThreadsInfer, 1)
    mlp_fwd_pair_kernel(const __grid_constant__ FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_act = smem + kSmAct;
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_ring = smem + kSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;                    // [kStages] this CTA's half chunk landed (leader: AND the peer's, via its relay)
  uint64_t* bar_empty = bar_full + kStages;     // [kStages] chunk consumed (multicast commit)
  uint64_t* bar_a = bar_empty + kStages;        // [2] leader only: a slot's next A operand is complete in BOTH CTAs
  uint64_t* bar_acc = bar_a + 2;                // [2] a slot's accumulator is complete (multicast commit)
  uint64_t* bar_st = bar_acc + 2;               // [kNumBuf] an image is complete in a buffer (epilogue -> store warp)
  uint64_t* bar_free = bar_st + kNumBuf;        // [kNumBuf] the stash store has read the buffer (store warp -> epilogue)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_free + kNumBuf);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  constexpr int kLayers = kFull ? 10 : 8;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, (rank == 0 && !(a.dbg_flags & 8)) ? 2 : 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(bar_a + s, 2 * kEpiWarps); mbar_init(bar_acc + s, 1); }
    for (int s = 0; s < kNumBuf; ++s) { mbar_init(bar_st + s, kEpiWarps); mbar_init(bar_free + s, 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(s_tmem, 512);
  if (warp >= 2 && warp < 2 + kEpiWarps) {
    const int et = threadIdx.x - 64;
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = et; i < AuxOffsets::count; i += kEpiThreads) s_aux[i] = gaux[i];
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  // tiles of this CTA in pair-iteration `it`: slot s -> ((it * n_pairs + pair_id) * 4 + 2 * rank + s
  auto iter_base = [&](int it) { return (it * n_pairs + pair_id) * 4; };

  if (warp == 0) {
    // ===================== weight producer (both CTAs: own N half of every chunk) =====================
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      const uint8_t* blob = a.packed + kPackedFwdOff;
      for (int it = 0; iter_base(it) < a.num_tiles; ++it) {
        for (int L = 0; L < kLayers; ++L) {
          const int c0 = layer_first_chunk(L), nl = layer_loads(L);
          const int reps = layer_shared(L) ? 1 : 2;
          for (int r = 0; r < reps; ++r) {
            for (int j = 0; j < nl; ++j) {
              NRF_MARK(0, 0x1000000 | (it << 12) | (L << 8) | (r << 4) | j);
              if (NRF_PAIR_SPIN) mbar_wait_spin(bar_empty + stage, ph ^ 1); else mbar_wait(bar_empty + stage, ph ^ 1);
              uint8_t* dst = s_ring + stage * kHalfChunk;
              if (L == 9) {   // N = 128: this CTA's half of a chunk is 8 KB; chunks 2j and 2j+1 (the last load: dir chunk only)
                const uint8_t* src = blob + (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 + (size_t)rank * (kChunkBytes128 / 2);
                const int n = (j == 2) ? 1 : 2;
                mbar_arrive_expect_tx(bar_full + stage, n * (kChunkBytes128 / 2));
                for (int h = 0; h < n; ++h)
                  bulk_g2s(dst + h * (kChunkBytes128 / 2), src + (size_t)(2 * j + h) * kChunkBytes128, kChunkBytes128 / 2, bar_full + stage);
              } else {
                mbar_arrive_expect_tx(bar_full + stage, kHalfChunk);
                bulk_g2s(dst, blob + (size_t)(c0 + j) * kChunkBytes256 + (size_t)rank * kHalfChunk, kHalfChunk, bar_full + stage);
              }
              if (++stage == kStages) { stage = 0; ph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      // ===================== UMMA issuer (leader CTA) =====================
      uint32_t stage = 0, ph = 0, a_par = 0;
      long long w_a = 0, w_full = 0, w_pfull = 0, w_mma = 0, n_mma = 0, t_begin = clock64();
      uint32_t acc_seen = 0;
      const uint32_t idesc256 = make_idesc_bf16(256, 256, 0, 0), idesc128 = make_idesc_bf16(256, 128, 0, 0);
      const uint64_t enc_desc = make_desc_sw128(smem_u32(s_enc), 16, 1024);
      const uint64_t act_desc = make_desc_sw128(smem_u32(s_act), 16, 1024);
      const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
      for (int it = 0; iter_base(it) < a.num_tiles; ++it) {
#pragma unroll 1
        for (int L = 0; L < kLayers; ++L) {
          const int nc = layer_chunks(L), nl = layer_loads(L);
          const bool shared = layer_shared(L);
          const uint32_t idesc = (L == 9) ? idesc128 : idesc256;
#pragma unroll 1
          for (int slot = 0; slot < 2; ++slot) {
            NRF_MARK(1, 0x2000000 | (it << 12) | (L << 8) | (slot << 4));
            const long long t0 = (NRF_FWD_DEBUG && a.dbg) ? clock64() : 0;
            if (NRF_PAIR_SPIN) mbar_wait_cluster_spin(bar_a + slot, (a_par >> slot) & 1u); else mbar_wait_cluster(bar_a + slot, (a_par >> slot) & 1u);
            if (NRF_FWD_DEBUG && a.dbg) w_a += clock64() - t0;
            a_par ^= 1u << slot;
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)slot * 256;
            const uint64_t s_enc_desc = enc_desc + (uint64_t)(slot * (kEncStride >> 4));
            const uint64_t s_act_desc = act_desc + (uint64_t)(slot * 4 * (kActChunkBytes >> 4));
            // a shared layer's stages are walked by slot 0 without release and again, releasing, by slot 1
            const bool release = !shared || slot == 1;
            uint32_t st = stage, p = ph;
            const long long t_issue = (NRF_FWD_DEBUG && a.dbg) ? clock64() : 0;
#pragma unroll 1
            for (int j = 0; j < nc; ++j) {
              // A operand of K chunk j: L0: enc; L4: enc then act 0..3 (skip, nerf.py:96-97); L9: act 0..3 then dir
              uint64_t a_desc;
              if (L == 0 || (L == 4 && j == 0) || (L == 9 && j == 4)) a_desc = s_enc_desc;
              else a_desc = s_act_desc + (uint64_t)((L == 4 ? j - 1 : j) * (kActChunkBytes >> 4));
              const bool new_stage = (L != 9) || !(j & 1);          // direction layer: two K chunks per stage
              const bool last_of_stage = (L != 9) || (j & 1) || j == 4;
              if (new_stage && (!shared || slot == 0)) {   // slot 1 of a shared layer re-walks stages slot 0 has seen
                NRF_MARK(1, 0x3000000 | (it << 12) | (L << 8) | (slot << 4) | j);
                const long long t1 = (NRF_FWD_DEBUG && a.dbg) ? clock64() : 0;
                if (NRF_PAIR_SPIN) mbar_wait_cluster_spin(bar_full + st, p); else mbar_wait_cluster(bar_full + st, p);
                if (NRF_FWD_DEBUG && a.dbg) w_pfull += clock64() - t1;
                tc_fence_after();
              }
              const uint64_t b_desc = ring_desc + (uint64_t)((st * kHalfChunk + ((L == 9 && (j & 1)) ? kChunkBytes128 / 2 : 0)) >> 4);
              const int reps = (NRF_FWD_DEBUG && (a.dbg_flags & 32)) ? 4 : 1;   // timing experiment: every MMA issued 4 times
              for (int rep = 0; rep < reps; ++rep)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (!(NRF_FWD_DEBUG && (a.dbg_flags & 1))) umma_ss_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (j | k | rep) ? 1u : 0u);
              if (last_of_stage) {
                if (release) umma_commit_pair(bar_empty + st, 0x3);
                if (++st == kStages) { st = 0; p ^= 1; }
              }
            }
            umma_commit_pair(bar_acc + slot, 0x3);
            if (NRF_FWD_DEBUG && (a.dbg_flags & 4)) {   // timing experiment: how long do one slot-layer's MMAs take?
              mbar_wait(bar_acc + slot, (acc_seen >> slot) & 1u);
              acc_seen ^= 1u << slot;
              if (slot == 1 && shared && L >= 1 && L <= 8) { w_mma += clock64() - t_issue; ++n_mma; }
            }
            if (release) { stage = st; ph = p; }
          }
        }
      }
      if (NRF_FWD_DEBUG && a.dbg) {
        a.dbg[blockIdx.x * 8 + 0] = (unsigned long long)(clock64() - t_begin);
        a.dbg[blockIdx.x * 8 + 1] = (unsigned long long)w_a;
        a.dbg[blockIdx.x * 8 + 2] = (unsigned long long)w_full;
        a.dbg[blockIdx.x * 8 + 3] = (unsigned long long)w_pfull;
        a.dbg[blockIdx.x * 8 + 6] = (unsigned long long)w_mma;
        a.dbg[blockIdx.x * 8 + 7] = (unsigned long long)n_mma;
      }
    } else if (lane == 0 && !(a.dbg_flags & 8)) {
      // ===================== relay (peer CTA): my half of the stage has landed -> leader =====================
      uint32_t stage = 0, ph = 0;
      for (int it = 0; iter_base(it) < a.num_tiles; ++it) {
        for (int L = 0; L < kLayers; ++L) {
          const int n = (layer_shared(L) ? 1 : 2) * layer_loads(L);
          for (int j = 0; j < n; ++j) {
            NRF_MARK(1, 0x8000000 | (it << 12) | (L << 8) | j);
            if (NRF_PAIR_SPIN) mbar_wait_spin(bar_full + stage, ph); else mbar_wait(bar_full + stage, ph);
            mbar_arrive_cluster(mapa_u32(smem_u32(bar_full + stage), 0));
            if (++stage == kStages) { stage = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp < 2 + kEpiWarps) {
#include "mlp_tc_pair_epilogue.inc"
  } else if (kTrain && lane == 0) {
#include "mlp_tc_pair_store.inc"
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // neither CTA leaves (or frees TMEM) while the other may still read its memory or signal it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

}  // namespace pair

template <bool kFull, bool kTrain>
static int launch_pair_t(const FwdArgs& a, cudaStream_t st) {
  static bool configured = false;   // debug build, single device
  const int smem_bytes = (int)pair::kSmTotal;
  if (!configured) {
    NRF_CUDA(cudaFuncSetAttribute(pair::mlp_fwd_pair_kernel<kFull, kTrain>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int quads = (a.num_tiles + 3) / 4;
  const int pairs = quads < sms / 2 ? quads : sms / 2;
  ProfScope prof((kTrain ? NRF_K_FWD_SIGMA_TRAIN : NRF_K_FWD_SIGMA_INFER) + (kFull ? 1 : 0), st, a.n_rows);
  pair::mlp_fwd_pair_kernel<kFull, kTrain><<<2 * pairs, kTrain ? pair::kThreadsTrain : pair::kThreadsInfer, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

int launch_fwd_pair(bool full, bool train, const FwdArgs& a, cudaStream_t st) {
  if (full) return train ? launch_pair_t<true, true>(a, st) : launch_pair_t<true, false>(a, st);
  return train ? launch_pair_t<false, true>(a, st) : launch_pair_t<false, false>(a, st);
}

}  // namespace tc
}  // namespace nrf
