// bf16 tensor-core NeRF MLP, forward (models/nerf.py:83-124 fused with the positional encoding
// nerf.py:21-38 and the point generation rendering.py:234).
//
// Persistent kernel, one CTA per SM, one 128-row tile of sample points in flight per CTA:
//   warp 0      weight producer: linear bulk copies (TMA engine) of pre-packed 32 KB weight chunks
//               L2 -> 3-stage shared-memory ring, mbarrier complete_tx
//   warp 1      UMMA issuer: tcgen05.mma kind::f16, M=128, N=256 (128 for the direction layer),
//               K=16 x 4 per chunk, A and B from shared memory (SWIZZLE_128B K-major),
//               fp32 accumulator in TMEM (256 columns); tcgen05.commit frees ring stages / signals
//               the epilogue
//   warps 2-9   epilogue: encoding of the tile's points straight into the layer-0 operand buffer,
//               then per layer tcgen05.ld -> +bias -> ReLU -> bf16 -> next layer's A operand in
//               shared memory; sigma / rgb heads as fp32 dot products on the accumulators;
//               training mode also streams the operand images + ReLU sign masks to the HBM stash
//               with bulk stores.
// Roofline: tensor pipe (593 408 MAC/sample).  HBM traffic: 32 B/ray in, 16 B/sample
// out (+5.3 KB/sample stash when training); weights (1.17 MB) stay L2 resident.
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kStages = 3;
constexpr int kEpiThreads = 256;
constexpr int kThreads = 64 + kEpiThreads;
constexpr uint32_t kTmemCols = 256;

// shared-memory map (offsets from a 1024-byte aligned base)
constexpr uint32_t kSmEnc = 0;
constexpr uint32_t kSmDir = kSmEnc + kActChunkBytes;
constexpr uint32_t kSmAct = kSmDir + kActChunkBytes;
constexpr uint32_t kSmRing = kSmAct + 4 * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStages * kChunkBytes256;
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;  // [2][128][4] floats
constexpr uint32_t kSmBars = kSmScratch + 2 * 128 * 4 * 4;
constexpr uint32_t kSmTotal = kSmBars + 128;

struct FwdArgs {
  const uint8_t* packed;
  const float* rays;
  const float* z;
  int64_t n_rows;
  int S;
  int num_tiles;
  float* sigma_out;
  float* rgb_out;
  uint8_t* stash;      // training only
  uint32_t* masks;     // training only
};

__device__ __forceinline__ void st_bf16(uint8_t* tile, int row, int col, float v) {
  *reinterpret_cast<__nv_bfloat16*>(tile + sw128_off(row, col >> 3) + (col & 7) * 2) = __float2bfloat16_rn(v);
}

template <bool kFull, bool kTrain>
__global__ void __launch_bounds__(kThreads, 1) mlp_fwd_kernel(const FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_dir = smem + kSmDir;
  uint8_t* s_act = smem + kSmAct;
  uint8_t* s_ring = smem + kSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;             // [kStages]
  uint64_t* bar_empty = bars + kStages;  // [kStages]
  uint64_t* bar_act = bars + 2 * kStages;
  uint64_t* bar_acc = bars + 2 * kStages + 1;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kLayers = kFull ? 10 : 8;
  constexpr int kChunksPerTile = kFull ? kFwdChunksFull : kFwdChunksSigma;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    mbar_init(bar_act, kEpiThreads);
    mbar_init(bar_acc, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, kTmemCols);
  if (warp >= 2) {
    const int et = threadIdx.x - 64;
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = et; i < AuxOffsets::count; i += kEpiThreads) s_aux[i] = gaux[i];
    uint4* zp = reinterpret_cast<uint4*>(s_enc);  // enc + dir are contiguous: zero the padding columns once
    for (int i = et; i < (int)(2 * kActChunkBytes / 16); i += kEpiThreads) zp[i] = make_uint4(0, 0, 0, 0);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===================== weight producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunksPerTile; ++ci, ++it) {
          const uint32_t stage = it % kStages, ph = (it / kStages) & 1;
          mbar_wait(bar_empty + stage, ph ^ 1);
          const bool half = ci >= kFwdChunksSigma + 4;
          const uint32_t bytes = half ? kChunkBytes128 : kChunkBytes256;
          const size_t off = half ? (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 +
                                        (size_t)(ci - kFwdChunksSigma - 4) * kChunkBytes128
                                  : (size_t)ci * kChunkBytes256;
          mbar_arrive_expect_tx(bar_full + stage, bytes);
          bulk_g2s(s_ring + stage * kChunkBytes256, a.packed + kPackedFwdOff + off, bytes, bar_full + stage);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    if (lane == 0) {
      uint32_t it = 0, act_it = 0;
      const uint32_t idesc256 = make_idesc_bf16(128, 256, 0, 0), idesc128 = make_idesc_bf16(128, 128, 0, 0);
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int L = 0; L < kLayers; ++L) {
          mbar_wait(bar_act, act_it & 1);
          ++act_it;
          tc_fence_after();
          const int nchunks = (L == 0) ? 1 : ((L == 4 || L == 9) ? 5 : 4);
          const uint32_t idesc = (L == 9) ? idesc128 : idesc256;
          for (int c = 0; c < nchunks; ++c, ++it) {
            const uint8_t* A;
            if (L == 0) A = s_enc;
            else if (L == 4) A = (c == 0) ? s_enc : s_act + (c - 1) * kActChunkBytes;
            else if (L == 9) A = (c < 4) ? s_act + c * kActChunkBytes : s_dir;
            else A = s_act + c * kActChunkBytes;
            const uint32_t stage = it % kStages, ph = (it / kStages) & 1;
            mbar_wait(bar_full + stage, ph);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(A), b_addr = smem_u32(s_ring + stage * kChunkBytes256);
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
    // ===================== epilogue warps =====================
    const int et = threadIdx.x - 64;
    const int q = warp & 3, h = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t acc_it = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
      uint8_t* stash_tile = kTrain ? a.stash + (size_t)tile * stash_tile_bytes(kFull) : nullptr;
      uint32_t* mask_tile = kTrain ? a.masks + (size_t)tile * mask_tile_words(kFull) : nullptr;
      const int64_t grow = (int64_t)tile * kTileM + row;
      const bool valid = grow < a.n_rows;
      if (kTrain) {  // previous tile's last bulk store must have finished reading act / enc / dir
        if (et == 0) bulk_wait_read0();
        named_bar_sync(1, kEpiThreads);
      }
      // ---- encode this tile's points: thread (er, part) fills 32 columns of row er ----
      {
        const int er = et & 127, part = et >> 7;
        int64_t gr = (int64_t)tile * kTileM + er;
        if (gr >= a.n_rows) gr = a.n_rows - 1;
        const int64_t ray = gr / a.S;
        const float zv = __ldg(a.z + gr);
        float p[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) p[c] = sample_point(a.rays, ray, zv, c);
        encode_store<10>(part, p, smem_u32(s_enc), er);
        if (kFull) {
          float d[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) d[c] = __ldg(a.rays + ray * 8 + 3 + c);
          if (part == 0) encode_store<4>(0, d, smem_u32(s_dir), er);
        }
      }
      fence_async_smem();
      if (kTrain) {
        named_bar_sync(1, kEpiThreads);
        if (et == 0) {
          bulk_s2g(stash_tile + (size_t)kStashEnc * kActChunkBytes, s_enc, kActChunkBytes);
          if (kFull) bulk_s2g(stash_tile + (size_t)kStashDir * kActChunkBytes, s_dir, kActChunkBytes);
          bulk_commit();
        }
      }
      tc_fence_before();
      mbar_arrive(bar_act);

      for (int L = 0; L < kLayers; ++L) {
        mbar_wait(bar_acc, acc_it & 1);
        ++acc_it;
        tc_fence_after();
        if (kTrain && L > 0) {  // the stash store of the previous layer's image reads s_act
          if (et == 0) bulk_wait_read0();
          named_bar_sync(1, kEpiThreads);
        }
        if (L < 9) {
          const bool relu = L < 8;
          const uint32_t bias_s = smem_u32(s_aux + (L < 8 ? AuxOffsets::bias + L * 256 : AuxOffsets::bias_final));
          const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma);
          const uint32_t act_s = smem_u32(s_act);
          const RowSwz rsw(row);
          float sig_part = 0.f;
          uint32_t mw[4];
          if (L == 7) epi_hidden128<true, kTrain, true>(t_row, h * 128, rsw, act_s, bias_s, wsig_s, mw, sig_part);
          else if (relu) epi_hidden128<true, kTrain, false>(t_row, h * 128, rsw, act_s, bias_s, wsig_s, mw, sig_part);
          else epi_hidden128<false, false, false>(t_row, h * 128, rsw, act_s, bias_s, wsig_s, mw, sig_part);
          if (kTrain && relu)
            *reinterpret_cast<uint4*>(mask_tile + ((size_t)L * 128 + row) * 8 + h * 4) = make_uint4(mw[0], mw[1], mw[2], mw[3]);
          if (L == 7) {
            s_scr[h * 128 + row] = sig_part;
            named_bar_sync(2, kEpiThreads);
            if (h == 0 && valid) a.sigma_out[grow] = s_scr[row] + s_scr[128 + row] + s_aux[AuxOffsets::b_sigma];
          }
          fence_async_smem();
          if (kTrain) {
            named_bar_sync(1, kEpiThreads);
            if (et == 0) {
              const int slot = (L < 8) ? kStashH0 + 4 * L : kStashFeat;
              bulk_s2g(stash_tile + (size_t)slot * kActChunkBytes, s_act, 4 * kActChunkBytes);
              bulk_commit();
            }
          }
          if (L + 1 < kLayers) {
            tc_fence_before();
            mbar_arrive(bar_act);
          }
        } else {
          // direction layer: N = 128; rgb head on the fp32 accumulators (nerf.py:119-121)
          float r0 = 0.f, r1 = 0.f, r2 = 0.f;
          uint32_t mw[2];
          const uint32_t bdir_s = smem_u32(s_aux + AuxOffsets::bias_dir), wrgb_s = smem_u32(s_aux + AuxOffsets::w_rgb);
          const RowSwz rsw(row);
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const int c0 = h * 64 + g * 32;
            uint32_t v[32];
            tmem_ld32(t_row + c0, v);
            tmem_ld_wait_dep(v);
            uint32_t m = 0, pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = ld_shared_f4(bdir_s + (c0 + j) * 4);
              const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
              const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
              if (kTrain) { m = push_sign(m, x0); m = push_sign(m, x1); m = push_sign(m, x2); m = push_sign(m, x3); }
              const float f0 = fmaxf(x0, 0.f), f1 = fmaxf(x1, 0.f), f2 = fmaxf(x2, 0.f), f3 = fmaxf(x3, 0.f);
              const float4 w0 = ld_shared_f4(wrgb_s + (c0 + j) * 4), w1 = ld_shared_f4(wrgb_s + (128 + c0 + j) * 4),
                           w2 = ld_shared_f4(wrgb_s + (256 + c0 + j) * 4);
              r0 += f0 * w0.x + f1 * w0.y + f2 * w0.z + f3 * w0.w;
              r1 += f0 * w1.x + f1 * w1.y + f2 * w1.z + f3 * w1.w;
              r2 += f0 * w2.x + f1 * w2.y + f2 * w2.z + f3 * w2.w;
              pk[j >> 1] = cvt_bf16x2(f0, f1);
              pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
            }
            mw[g] = ~m;
            if (kTrain) {
              const uint32_t chunk = smem_u32(s_act) + (uint32_t)h * kActChunkBytes;  // 64 columns per half = one chunk
#pragma unroll
              for (int p = 0; p < 4; ++p)
                st_shared_v4(chunk + rsw.piece(g * 4 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
            }
          }
          if (kTrain) *reinterpret_cast<uint2*>(mask_tile + 8 * 128 * 8 + (size_t)row * 4 + h * 2) = make_uint2(mw[0], mw[1]);
          float* sc = s_scr + (h * 128 + row) * 4;
          sc[0] = r0; sc[1] = r1; sc[2] = r2;
          named_bar_sync(2, kEpiThreads);
          if (h == 0 && valid) {
            const float* o = s_scr + (128 + row) * 4;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float pre = sc[c] + o[c] + s_aux[AuxOffsets::b_rgb + c];
              a.rgb_out[grow * 3 + c] = 1.0f / (1.0f + expf(-pre));
            }
          }
          if (kTrain) {
            fence_async_smem();
            named_bar_sync(1, kEpiThreads);
            if (et == 0) {
              bulk_s2g(stash_tile + (size_t)kStashMid * kActChunkBytes, s_act, 2 * kActChunkBytes);
              bulk_commit();
            }
          }
        }
      }
    }
    if (kTrain && et == 0) bulk_wait_all0();
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
  static bool configured = false;
  const int smem_bytes = (int)kSmTotal + 1024;
  if (!configured) {
    NRF_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<kFull, kTrain>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  int dev = 0, sms = kNumSMs;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = a.num_tiles < sms ? a.num_tiles : sms;
  mlp_fwd_kernel<kFull, kTrain><<<grid, kThreads, smem_bytes, st>>>(a);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;

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
  cudaStream_t st = as_stream(stream);
  if (stash != nullptr) {
    const size_t need = nrf_mlp_tc_stash_bytes(rows, sigma_only);
    if (stash_bytes < need) {
      set_error("nrf_mlp_tc_fwd: stash %zu < %zu bytes", stash_bytes, need);
      return NRF_ERR_WORKSPACE;
    }
    a.stash = static_cast<uint8_t*>(stash);
    a.masks = reinterpret_cast<uint32_t*>(a.stash + (size_t)a.num_tiles * stash_tile_bytes(!sigma_only));
    return sigma_only ? launch_fwd<false, true>(a, st) : launch_fwd<true, true>(a, st);
  }
  return sigma_only ? launch_fwd<false, false>(a, st) : launch_fwd<true, false>(a, st);
}

