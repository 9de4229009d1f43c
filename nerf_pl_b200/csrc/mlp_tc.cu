// bf16 tensor-core NeRF MLP, forward (models/nerf.py:83-124 fused with the positional encoding
// nerf.py:21-38 and the point generation rendering.py:234).
//
// Persistent kernel, one CTA per SM, 128-row tiles of sample points; the layer chain is software
// pipelined at 64-column (= one K chunk) granularity so the tensor pipe rarely waits for the epilogue:
//   warp 0      weight producer: linear bulk copies (TMA engine) of pre-packed 32 KB weight chunks
//               L2 -> 3-stage shared-memory ring, mbarrier complete_tx
//   warp 1      UMMA issuer: tcgen05.mma kind::f16, M=128, N=256 (128 for the direction layer),
//               K=16 x 4 per chunk, A and B from shared memory (SWIZZLE_128B K-major).  Two fp32
//               accumulators in TMEM (2 x 256 columns) alternate between layers; the MMAs of layer
//               L+1 over K chunk c start as soon as the epilogue of layer L has produced chunk c.
//   warps 2-9   epilogue, chunk by chunk: tcgen05.ld (32 lanes x 32 columns per thread) -> +bias ->
//               ReLU -> bf16 -> in place into the A-operand buffer -> per-chunk mbarrier; sigma / rgb
//               heads as fp32 dot products on the accumulators; the encoding of the NEXT tile is written
//               while layers 5..9 of the current one run (the xyz encoding buffer is free after layer
//               4, the direction buffer after layer 9).  Training mode also streams the operand images
//               + ReLU sign masks to the HBM stash with bulk stores.
// Roofline: tensor pipe (593 408 MAC/sample).  HBM traffic: 32 B/ray in, 16 B/sample
// out (+5.3 KB/sample stash when training); weights (1.17 MB) stay L2 resident (measured L2 -> SMEM
// bulk-copy rate with all 148 CTAs on the same blob: > 19 TB/s, tools/stream_bench.py).
#include "common.cuh"
#include "embed.cuh"
#include "tc_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

constexpr int kStages = 3;
constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads;
constexpr uint32_t kTmemCols = 512;

// shared-memory map (offsets from a 1024-byte aligned base)
constexpr uint32_t kSmEnc = 0;
constexpr uint32_t kSmDir = kSmEnc + kActChunkBytes;
constexpr uint32_t kSmAct = kSmDir + kActChunkBytes;
constexpr uint32_t kSmRing = kSmAct + 4 * kActChunkBytes;
constexpr uint32_t kSmAux = kSmRing + kStages * kChunkBytes256;
constexpr uint32_t kSmScratch = kSmAux + AuxOffsets::count * 4;  // [2][128][4] floats
constexpr uint32_t kSmBars = kSmScratch + 2 * 128 * 4 * 4;
constexpr uint32_t kSmTotal = kSmBars + 256;

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
  unsigned long long* dbg;  // optional per-CTA cycle counters (nrf_dbg_set_counters), null normally
};

// A-operand readiness barriers: act chunks 0..3, xyz encoding, direction encoding
constexpr int kBarEnc = 4, kBarDir = 5, kNumBarA = 6;

template <bool kFull, bool kTrain>
__global__ void __launch_bounds__(kThreads, 1) mlp_fwd_kernel(const __grid_constant__ FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_enc = smem + kSmEnc;
  uint8_t* s_dir = smem + kSmDir;
  uint8_t* s_act = smem + kSmAct;
  uint8_t* s_ring = smem + kSmRing;
  float* s_aux = reinterpret_cast<float*>(smem + kSmAux);
  float* s_scr = reinterpret_cast<float*>(smem + kSmScratch);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBars);
  uint64_t* bar_full = bars;                 // [kStages] weight chunk landed
  uint64_t* bar_empty = bars + kStages;      // [kStages] weight chunk consumed
  uint64_t* bar_a = bars + 2 * kStages;      // [kNumBarA] A-operand chunk written (epilogue -> MMA)
  uint64_t* bar_acc = bar_a + kNumBarA;      // [2] accumulator complete (MMA -> epilogue)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_acc + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kLayers = kFull ? 10 : 8;
  constexpr int kChunksPerTile = kFull ? kFwdChunksFull : kFwdChunksSigma;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
    for (int s = 0; s < kNumBarA; ++s) mbar_init(bar_a + s, kEpiWarps);
    mbar_init(bar_acc, 1);
    mbar_init(bar_acc + 1, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, kTmemCols);
  if (warp >= 2) {
    const int et = threadIdx.x - 64;
    const float* gaux = reinterpret_cast<const float*>(a.packed + kPackedAuxOff);
    for (int i = et; i < AuxOffsets::count; i += kEpiThreads) s_aux[i] = gaux[i];
    uint4* zp = reinterpret_cast<uint4*>(s_dir);  // direction chunk: columns 27..63 stay zero for the whole kernel
    for (int i = et; i < (int)(kActChunkBytes / 16); i += kEpiThreads) zp[i] = make_uint4(0, 0, 0, 0);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===================== weight producer =====================
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int ci = 0; ci < kChunksPerTile; ++ci) {
          mbar_wait(bar_empty + stage, ph ^ 1);
          const bool half = ci >= kFwdChunksSigma + 4;
          const uint32_t bytes = half ? kChunkBytes128 : kChunkBytes256;
          const size_t off = half ? (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 +
                                        (size_t)(ci - kFwdChunksSigma - 4) * kChunkBytes128
                                  : (size_t)ci * kChunkBytes256;
          mbar_arrive_expect_tx(bar_full + stage, bytes);
          bulk_g2s(s_ring + stage * kChunkBytes256, a.packed + kPackedFwdOff + off, bytes, bar_full + stage);
          if (++stage == kStages) { stage = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    // One thread; descriptors are precomputed and only their 14-bit start-address field is advanced, so the
    // per-chunk issue cost stays well below the 512 cycles the four MMAs of a chunk execute for.
    if (lane == 0) {
      uint32_t stage = 0, ph = 0;
      uint32_t a_par = 0;   // bit i: parity of the next completion of bar_a[i] to wait for
      long long w_a = 0, w_full = 0, t_begin = clock64();
      const uint32_t idesc256 = make_idesc_bf16(128, 256, 0, 0), idesc128 = make_idesc_bf16(128, 128, 0, 0);
      const uint64_t enc_desc = make_desc_sw128(smem_u32(s_enc), 16, 1024);
      const uint64_t dir_desc = make_desc_sw128(smem_u32(s_dir), 16, 1024);
      const uint64_t act_desc = make_desc_sw128(smem_u32(s_act), 16, 1024);
      const uint64_t ring_desc = make_desc_sw128(smem_u32(s_ring), 16, 1024);
      auto wait_a = [&](int src) {
        const long long t0 = a.dbg ? clock64() : 0;
        mbar_wait(bar_a + src, (a_par >> src) & 1u);
        a_par ^= 1u << src;
        if (a.dbg) w_a += clock64() - t0;
      };
      // one K chunk (64 columns of A against one ring stage): 4 MMAs of K = 16
      auto chunk = [&](uint64_t a_desc, uint32_t d_tmem, uint32_t idesc, bool first) {
        const long long t0 = a.dbg ? clock64() : 0;
        mbar_wait(bar_full + stage, ph);
        if (a.dbg) w_full += clock64() - t0;
        tc_fence_after();
        const uint64_t b_desc = ring_desc + (uint64_t)(stage * (kChunkBytes256 >> 4));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (!first || k) ? 1u : 0u);
        umma_commit(bar_empty + stage);
        if (++stage == kStages) { stage = 0; ph ^= 1; }
      };
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
#pragma unroll 1
        for (int L = 0; L < kLayers; ++L) {
          const uint32_t d_tmem = tmem_base + (uint32_t)(L & 1) * 256;
          const uint32_t idesc = (L == 9) ? idesc128 : idesc256;
          bool first = true;
          if (L == 0 || L == 4) {          // the xyz encoding leads (skip connection: models/nerf.py:96-97)
            if (L == 0) wait_a(kBarEnc);   // at layer 4 the same completion is still current
            chunk(enc_desc, d_tmem, idesc, true);
            first = false;
          }
          if (L != 0) {
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              wait_a(c);
              chunk(act_desc + (uint64_t)(c * (kActChunkBytes >> 4)), d_tmem, idesc, first);
              first = false;
            }
          }
          if (L == 9) {
            wait_a(kBarDir);
            chunk(dir_desc, d_tmem, idesc, false);
          }
          umma_commit(bar_acc + (L & 1));
        }
      }
      if (a.dbg) {
        a.dbg[blockIdx.x * 8 + 0] = (unsigned long long)(clock64() - t_begin);
        a.dbg[blockIdx.x * 8 + 1] = (unsigned long long)w_a;
        a.dbg[blockIdx.x * 8 + 2] = (unsigned long long)w_full;
      }
    }
  } else {
    // ===================== epilogue warps =====================
    const int et = threadIdx.x - 64;
    const int q = warp & 3, sub = (warp - 2) >> 2;   // TMEM lane quarter; 32-column half of every 64-column chunk
    const int row = q * 32 + lane;
    const RowSwz rsw(row);
    const uint32_t act_s = smem_u32(s_act), enc_s = smem_u32(s_enc), dir_s = smem_u32(s_dir);
    const uint32_t wsig_s = smem_u32(s_aux + AuxOffsets::w_sigma);
    uint32_t acc_par = 0;
    long long w_acc = 0, t_begin = clock64();

    // this warp's part of an A-operand chunk is written: make it visible to the async proxy and tell the MMA thread
    auto signal = [&](int idx) {
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a + idx);
    };
    // positional encoding of a tile's sample points -> s_enc (thread (er, part) fills 32 columns of row er)
    auto encode_xyz = [&](int tile) {
      const int er = et & 127, part = et >> 7;
      int64_t gr = (int64_t)tile * kTileM + er;
      if (gr >= a.n_rows) gr = a.n_rows - 1;
      const int64_t ray = gr / a.S;
      const float zv = __ldg(a.z + gr);
      float p[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) p[c] = sample_point(a.rays, ray, zv, c);
      encode_store<10>(part, p, enc_s, er);
      signal(kBarEnc);
      if (kTrain) {
        named_bar_sync(1, kEpiThreads);
        if (et == 0) {
          bulk_s2g(a.stash + img_off(kStashEnc, 1, tile, a.num_tiles), s_enc, kActChunkBytes);
          bulk_commit();
        }
      }
    };
    auto encode_dir = [&](int tile) {
      const int er = et & 127, part = et >> 7;
      if (part == 0) {
        int64_t gr = (int64_t)tile * kTileM + er;
        if (gr >= a.n_rows) gr = a.n_rows - 1;
        const int64_t ray = gr / a.S;
        float d[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) d[c] = __ldg(a.rays + ray * 8 + 3 + c);
        encode_store<4>(0, d, dir_s, er);
      }
      signal(kBarDir);
      if (kTrain) {
        named_bar_sync(1, kEpiThreads);
        if (et == 0) {
          bulk_s2g(a.stash + img_off(kStashDir, 1, tile, a.num_tiles), s_dir, kActChunkBytes);
          bulk_commit();
        }
      }
    };

    int tile = blockIdx.x;
    if (tile < a.num_tiles) {
      encode_xyz(tile);
      if (kFull) encode_dir(tile);
    }
    for (; tile < a.num_tiles; tile += gridDim.x) {
      const int next = tile + gridDim.x;
      uint32_t* mask_tile = kTrain ? a.masks + (size_t)tile * mask_tile_words(kFull) : nullptr;
      const int64_t grow = (int64_t)tile * kTileM + row;
      const bool valid = grow < a.n_rows;
#pragma unroll 1
      for (int L = 0; L < kLayers; ++L) {
        const uint32_t b = L & 1;
        const long long tw0 = a.dbg ? clock64() : 0;
        mbar_wait(bar_acc + b, (acc_par >> b) & 1u);
        if (a.dbg) w_acc += clock64() - tw0;
        acc_par ^= 1u << b;
        tc_fence_after();
        if (kTrain) {  // the stash store of the previous image may still be reading s_act
          if (et == 0) bulk_wait_read0();
          named_bar_sync(1, kEpiThreads);
        }
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + b * 256;
        if (L < 9) {
          const bool feed = L + 1 < kLayers;   // the next layer's MMAs consume these chunks
          const uint32_t bias_s = smem_u32(s_aux + (L < 8 ? AuxOffsets::bias + L * 256 : AuxOffsets::bias_final));
          uint32_t mw[4] = {0, 0, 0, 0};
          float dot = 0.f;
          uint32_t va[32], vb[32];
          tmem_ld32(t_acc + sub * 32, va);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t(&v)[32] = (c & 1) ? vb : va;
            uint32_t(&vn)[32] = (c & 1) ? va : vb;
            tmem_ld_wait_dep(v);
            if (c + 1 < 4) tmem_ld32(t_acc + (c + 1) * 64 + sub * 32, vn);
            const int c0 = c * 64 + sub * 32;
            const uint32_t chunk_s = act_s + (uint32_t)c * kActChunkBytes;
            if (L == 7) epi_fwd32<true, kTrain, true>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 4, mw[c], dot);
            else if (L < 8) epi_fwd32<true, kTrain, false>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 4, mw[c], dot);
            else epi_fwd32<false, false, false>(v, c0, bias_s, wsig_s, rsw, chunk_s, sub * 4, mw[c], dot);
            if (feed) signal(c);
          }
          if (kTrain && L < 8)
            *reinterpret_cast<uint4*>(mask_tile + ((size_t)L * 128 + row) * 8 + sub * 4) = make_uint4(mw[0], mw[1], mw[2], mw[3]);
          if (L == 7) {
            s_scr[sub * 128 + row] = dot;
            named_bar_sync(2, kEpiThreads);
            if (sub == 0 && valid) a.sigma_out[grow] = s_scr[row] + s_scr[128 + row] + s_aux[AuxOffsets::b_sigma];
          }
          if (kTrain) {
            if (!feed) fence_async_smem();
            named_bar_sync(1, kEpiThreads);
            if (et == 0) {
              const int slot = (L < 8) ? kStashH0 + 4 * L : kStashFeat;
              bulk_s2g(a.stash + img_off(slot, 4, tile, a.num_tiles), s_act, 4 * kActChunkBytes);
              bulk_commit();
            }
          }
        } else {
          // direction layer: N = 128; rgb head on the fp32 accumulators (nerf.py:119-121)
          float r0 = 0.f, r1 = 0.f, r2 = 0.f;
          uint32_t mwd[2] = {0, 0};
          const uint32_t bdir_s = smem_u32(s_aux + AuxOffsets::bias_dir), wrgb_s = smem_u32(s_aux + AuxOffsets::w_rgb);
          uint32_t va[32], vb[32];
          tmem_ld32(t_acc + sub * 32, va);
          tmem_ld32(t_acc + 64 + sub * 32, vb);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t(&v)[32] = c ? vb : va;
            tmem_ld_wait_dep(v);
            const int c0 = c * 64 + sub * 32;
            uint32_t m = 0, pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 bb = ld_shared_f4(bdir_s + (c0 + j) * 4);
              const float x0 = __uint_as_float(v[j]) + bb.x, x1 = __uint_as_float(v[j + 1]) + bb.y;
              const float x2 = __uint_as_float(v[j + 2]) + bb.z, x3 = __uint_as_float(v[j + 3]) + bb.w;
              if (kTrain) { m = push_sign(m, x0); m = push_sign(m, x1); m = push_sign(m, x2); m = push_sign(m, x3); }
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
            mwd[c] = ~m;
            if (kTrain) {
              const uint32_t chunk_s = act_s + (uint32_t)c * kActChunkBytes;
#pragma unroll
              for (int p = 0; p < 4; ++p)
                st_shared_v4(chunk_s + rsw.piece(sub * 4 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
            }
          }
          if (kTrain) *reinterpret_cast<uint2*>(mask_tile + 8 * 128 * 8 + (size_t)row * 4 + sub * 2) = make_uint2(mwd[0], mwd[1]);
          float* sc = s_scr + (sub * 128 + row) * 4;
          sc[0] = r0; sc[1] = r1; sc[2] = r2;
          named_bar_sync(2, kEpiThreads);
          if (sub == 0 && valid) {
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
              bulk_s2g(a.stash + img_off(kStashMid, 2, tile, a.num_tiles), s_act, 2 * kActChunkBytes);
              bulk_commit();
            }
          }
        }
        // the encodings of the next tile: s_enc is free once layer 4's MMAs are done, s_dir after layer 9's
        if (L == 4 && next < a.num_tiles) encode_xyz(next);
        if (kFull && L == 9 && next < a.num_tiles) encode_dir(next);
      }
    }
    if (kTrain && et == 0) bulk_wait_all0();
    if (a.dbg && et == 0) {
      a.dbg[blockIdx.x * 8 + 4] = (unsigned long long)(clock64() - t_begin);
      a.dbg[blockIdx.x * 8 + 5] = (unsigned long long)w_acc;
    }
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

static unsigned long long* g_dbg_counters = nullptr;
// Diagnostics: device buffer of >= 8 * gridDim 64-bit counters filled by the next forward launches (null = off).
extern "C" void nrf_dbg_set_counters(void* dev_ptr) { g_dbg_counters = static_cast<unsigned long long*>(dev_ptr); }

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
  a.dbg = g_dbg_counters;
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

