// bf16 tensor-core NeRF MLP, backward data path ("dgrad"): gradients w.r.t. every layer's
// pre-activation, G_l = dL/da_l, for 128-row tiles of sample points.
//
//   G9 = (d_rgb * rgb(1-rgb)) . W_rgb  (x) relu'(dir layer)         CUDA cores (K = 3)
//   G8 = G9 . W_dir[:, :256]                                        UMMA  M128 N256 K128
//   G7 = (G8 . W_final + d_sigma w_sigma) (x) relu'(h7)             UMMA  M128 N256 K256 + rank-1 on CUDA cores
//   G_{l-1} = (G_l . W_l[:, h part]) (x) relu'(h_{l-1}),  l = 7..1  UMMA  M128 N256 K256
//
// Same warp-specialised structure as the forward kernel (mlp_tc.cu): bulk-copy weight ring
// (transposed, pre-packed images), one UMMA issuer thread, eight epilogue warps that turn the
// TMEM accumulator into the next A operand (bf16, SWIZZLE_128B) in shared memory.  ReLU
// derivatives come from the 1-bit sign masks the forward kernel stashed (32 B per row and layer).
// Every G_l image is also streamed to HBM with a bulk store; the wgrad kernel
// (mlp_tc_wgrad.cu) consumes those images as MN-major UMMA operands.
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kDgStages = 3;
constexpr int kDgEpiThreads = 256;
constexpr int kDgThreads = 64 + kDgEpiThreads;
constexpr uint32_t kDgSmAct = 0;
constexpr uint32_t kDgSmRing = kDgSmAct + 4 * kActChunkBytes;
constexpr uint32_t kDgSmAux = kDgSmRing + kDgStages * kChunkBytes256;
constexpr uint32_t kDgSmBars = kDgSmAux + AuxOffsets::count * 4;
constexpr uint32_t kDgSmTotal = kDgSmBars + 128;

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

template <bool kFull>
__global__ void __launch_bounds__(kDgThreads, 1) mlp_dgrad_kernel(const DgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_act = smem + kDgSmAct;
  uint8_t* s_ring = smem + kDgSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kDgSmAux);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kDgSmBars);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kDgStages;
  uint64_t* bar_act = bars + 2 * kDgStages;
  uint64_t* bar_acc = bars + 2 * kDgStages + 1;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * kDgStages + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kMmaLayers = kFull ? 9 : 7;
  constexpr int kFirstChunk = kFull ? 0 : 6;  // sigma-only starts at W7^T
  constexpr int kChunks = kBwdChunks - kFirstChunk;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kDgStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    mbar_init(bar_act, kDgEpiThreads);
    mbar_init(bar_acc, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 256);
  if (warp >= 2) {
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = threadIdx.x - 64; i < AuxOffsets::count; i += kDgEpiThreads) s_aux[i] = gaux[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunks; ++ci, ++it) {
          const uint32_t stage = it % kDgStages, ph = (it / kDgStages) & 1;
          mbar_wait(bar_empty + stage, ph ^ 1);
          mbar_arrive_expect_tx(bar_full + stage, kChunkBytes256);
          bulk_g2s(s_ring + stage * kChunkBytes256,
                   a.packed + kPackedBwdOff + (size_t)(kFirstChunk + ci) * kChunkBytes256, kChunkBytes256,
                   bar_full + stage);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t it = 0, act_it = 0;
      const uint32_t idesc = make_idesc_bf16(128, 256, 0, 0);
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int i = 0; i < kMmaLayers; ++i) {
          mbar_wait(bar_act, act_it & 1);
          ++act_it;
          tc_fence_after();
          const int nchunks = (kFull && i == 0) ? 2 : 4;
          for (int c = 0; c < nchunks; ++c, ++it) {
            const uint32_t stage = it % kDgStages, ph = (it / kDgStages) & 1;
            mbar_wait(bar_full + stage, ph);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(s_act + c * kActChunkBytes), b_addr = smem_u32(s_ring + stage * kChunkBytes256);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_ss(tmem_base, make_desc_sw128(a_addr + k * 32, 16, 1024), make_desc_sw128(b_addr + k * 32, 16, 1024),
                      idesc, (c | k) != 0);
            umma_commit(bar_empty + stage);
          }
          umma_commit(bar_acc);
        }
      }
    }
  } else {
    const int et = threadIdx.x - 64;
    const int q = warp & 3, h = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t acc_it = 0;
    bool store_pending = false;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
      uint8_t* g_tile = a.gstash + (size_t)tile * gstash_tile_bytes(kFull);
      const uint32_t* mask_tile = a.masks + (size_t)tile * mask_tile_words(kFull);
      const int64_t grow = (int64_t)tile * kTileM + row;
      const bool valid = grow < a.n_rows;
      const float dsig = valid ? __ldg(a.d_sigma + grow) : 0.f;
      // ---- prologue: the first A operand of the chain, on CUDA cores ----
      if (store_pending) {  // the previous tile's last image store still reads s_act
        if (et == 0) bulk_wait_read0();
        named_bar_sync(1, kDgEpiThreads);
      }
      if (kFull) {
        float dp[3] = {0.f, 0.f, 0.f};
        if (valid) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const float y = __ldg(a.rgb_out + grow * 3 + c);
            dp[c] = __ldg(a.d_rgb + grow * 3 + c) * y * (1.0f - y);  // sigmoid' (nerf.py:79-81)
          }
        }
        if (h == 0) *reinterpret_cast<float4*>(a.dpre + ((size_t)tile * kTileM + row) * 4) = make_float4(dp[0], dp[1], dp[2], 0.f);
        const uint2 mw = *reinterpret_cast<const uint2*>(mask_tile + 8 * 128 * 8 + (size_t)row * 4 + h * 2);
        const uint32_t chunk = smem_u32(s_act) + (uint32_t)h * kActChunkBytes;
        const uint32_t wrgb_s = smem_u32(s_aux + AuxOffsets::w_rgb);
        const RowSwz rsw(row);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const uint32_t m = g ? mw.y : mw.x;
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const int c = h * 64 + g * 32 + j;
            const float4 w0 = ld_shared_f4(wrgb_s + c * 4), w1 = ld_shared_f4(wrgb_s + (128 + c) * 4),
                         w2 = ld_shared_f4(wrgb_s + (256 + c) * 4);
            float f0 = dp[0] * w0.x + dp[1] * w1.x + dp[2] * w2.x;
            float f1 = dp[0] * w0.y + dp[1] * w1.y + dp[2] * w2.y;
            float f2 = dp[0] * w0.z + dp[1] * w1.z + dp[2] * w2.z;
            float f3 = dp[0] * w0.w + dp[1] * w1.w + dp[2] * w2.w;
            f0 = mask_bit(m, j) ? f0 : 0.f;
            f1 = mask_bit(m, j + 1) ? f1 : 0.f;
            f2 = mask_bit(m, j + 2) ? f2 : 0.f;
            f3 = mask_bit(m, j + 3) ? f3 : 0.f;
            pk[j >> 1] = cvt_bf16x2(f0, f1);
            pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
          }
#pragma unroll
          for (int p = 0; p < 4; ++p)
            st_shared_v4(chunk + rsw.piece(g * 4 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
        }
      } else {
        const uint4 mw4 = *reinterpret_cast<const uint4*>(mask_tile + ((size_t)7 * 128 + row) * 8 + h * 4);
        const uint32_t mw[4] = {mw4.x, mw4.y, mw4.z, mw4.w};
        const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma), act_s = smem_u32(s_act);
        const RowSwz rsw(row);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int c0 = h * 128 + g * 32;
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 w = ld_shared_f4(wsig_s + (c0 + j) * 4);
            const float f0 = mask_bit(mw[g], j) ? dsig * w.x : 0.f, f1 = mask_bit(mw[g], j + 1) ? dsig * w.y : 0.f;
            const float f2 = mask_bit(mw[g], j + 2) ? dsig * w.z : 0.f, f3 = mask_bit(mw[g], j + 3) ? dsig * w.w : 0.f;
            pk[j >> 1] = cvt_bf16x2(f0, f1);
            pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
          }
          const uint32_t chunk = act_s + (uint32_t)(c0 >> 6) * kActChunkBytes;
          const int p0 = (c0 & 63) >> 3;
#pragma unroll
          for (int p = 0; p < 4; ++p)
            st_shared_v4(chunk + rsw.piece(p0 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
        }
      }
      fence_async_smem();
      named_bar_sync(1, kDgEpiThreads);
      if (et == 0) {
        if (kFull) bulk_s2g(g_tile + (size_t)kGstashG9 * kActChunkBytes, s_act, 2 * kActChunkBytes);
        else bulk_s2g(g_tile + (size_t)(4 * 7) * kActChunkBytes, s_act, 4 * kActChunkBytes);
        bulk_commit();
      }
      store_pending = true;
      tc_fence_before();
      mbar_arrive(bar_act);

      for (int i = 0; i < kMmaLayers; ++i) {
        const int t = (kFull ? 8 : 6) - i;  // this epilogue produces G_t
        mbar_wait(bar_acc, acc_it & 1);
        ++acc_it;
        tc_fence_after();
        if (et == 0) bulk_wait_read0();
        named_bar_sync(1, kDgEpiThreads);
        uint32_t mw[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        if (t < 8) {
          const uint4 m4 = *reinterpret_cast<const uint4*>(mask_tile + ((size_t)t * 128 + row) * 8 + h * 4);
          mw[0] = m4.x; mw[1] = m4.y; mw[2] = m4.z; mw[3] = m4.w;
        }
        {
          const RowSwz rsw(row);
          const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma), act_s = smem_u32(s_act);
          if (kFull && t == 7) epi_dgrad128<true>(t_row, h * 128, rsw, act_s, mw, dsig, wsig_s);
          else epi_dgrad128<false>(t_row, h * 128, rsw, act_s, mw, dsig, wsig_s);
        }
        fence_async_smem();
        named_bar_sync(1, kDgEpiThreads);
        if (et == 0) {
          const int slot = (t == 8) ? kGstashG8 : 4 * t;
          bulk_s2g(g_tile + (size_t)slot * kActChunkBytes, s_act, 4 * kActChunkBytes);
          bulk_commit();
        }
        if (t > 0) {
          tc_fence_before();
          mbar_arrive(bar_act);
        }
      }
    }
    if (et == 0) bulk_wait_all0();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

template <bool kFull>
static int launch_dgrad_t(const DgradArgs& a, cudaStream_t st) {
  static bool configured = false;
  const int smem_bytes = (int)kDgSmTotal + 1024;
  if (!configured) {
    NRF_CUDA(cudaFuncSetAttribute(mlp_dgrad_kernel<kFull>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
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
