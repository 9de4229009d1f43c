// nrf_mlp_tc_pack: fp32 checkpoint-layout parameters -> bf16 UMMA operand images (tc_layout.cuh).
// Runs once per optimizer step (host side caches by parameter version); ~2.3 MB written.
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_layout.cuh"

namespace nrf {
namespace tc {

struct PackChunk {
  const float* base;   // element (n, kk) = base[n*stride_n + kk*stride_k]
  int32_t stride_n, stride_k;
  int32_t n_rows;      // 256 or 128
  int32_t k_valid;     // columns >= k_valid are zero padding
  uint32_t dst_off;    // byte offset in the packed buffer
};

constexpr int kMaxPackChunks = kFwdChunksFull + kBwdChunks;
struct PackTable {
  PackChunk c[kMaxPackChunks];
  int n;
};

struct AuxSrc {
  const float* bias[8];
  const float* bias_final; const float* bias_dir; const float* w_sigma; const float* w_rgb;
  const float* b_sigma; const float* b_rgb;
};

__global__ void pack_chunks_kernel(const PackTable tab, uint8_t* __restrict__ packed) {
  const PackChunk ch = tab.c[blockIdx.x];
  uint8_t* dst = packed + ch.dst_off;
  const int total16 = ch.n_rows * 8;  // 16-byte pieces
  for (int p = threadIdx.x; p < total16; p += blockDim.x) {
    const int n = p >> 3, c16 = p & 7;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k0 = c16 * 8 + 2 * j;
      const float lo = (k0 < ch.k_valid) ? ch.base[(int64_t)n * ch.stride_n + (int64_t)k0 * ch.stride_k] : 0.f;
      const float hi = (k0 + 1 < ch.k_valid) ? ch.base[(int64_t)n * ch.stride_n + (int64_t)(k0 + 1) * ch.stride_k] : 0.f;
      w[j] = pack_bf16(lo, hi);
    }
    *reinterpret_cast<uint4*>(dst + sw128_off(n, c16)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void pack_aux_kernel(const AuxSrc s, float* __restrict__ aux) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 8 * 256) aux[AuxOffsets::bias + t] = s.bias[t >> 8][t & 255];
  if (t < 256) {
    aux[AuxOffsets::bias_final + t] = s.bias_final[t];
    aux[AuxOffsets::w_sigma + t] = s.w_sigma[t];
  }
  if (t < 128) aux[AuxOffsets::bias_dir + t] = s.bias_dir[t];
  if (t < 384) aux[AuxOffsets::w_rgb + t] = s.w_rgb[t];
  if (t < 3) aux[AuxOffsets::b_rgb + t] = s.b_rgb[t];
  if (t == 0) aux[AuxOffsets::b_sigma] = s.b_sigma[0];
}

}  // namespace tc
}  // namespace nrf

using namespace nrf;
using namespace nrf::tc;

extern "C" size_t nrf_mlp_tc_packed_bytes(void) { return kPackedBytes; }

extern "C" int nrf_mlp_tc_pack(const nrf_mlp_params* p, void* packed, nrf_stream_t stream) {
  NRF_REQUIRE(p && packed, "nrf_mlp_tc_pack: null pointer");
  if (!(p->D == 8 && p->W == 256 && p->in_xyz == 63 && p->in_dir == 27 && p->skip_mask == (1u << 4))) {
    set_error("nrf_mlp_tc_pack: the tensor-core path supports D=8 W=256 in_xyz=63 in_dir=27 skips=[4] only");
    return NRF_ERR_UNSUPPORTED;
  }
  for (int i = 0; i < 8; ++i) NRF_REQUIRE(p->xyz_w[i] && p->xyz_b[i], "nrf_mlp_tc_pack: null layer %d", i);
  NRF_REQUIRE(p->final_w && p->final_b && p->dir_w && p->dir_b && p->sigma_w && p->sigma_b && p->rgb_w && p->rgb_b,
              "nrf_mlp_tc_pack: null head");
  PackTable tab{};
  int n = 0;
  uint32_t off = (uint32_t)kPackedFwdOff;
  auto add = [&](const float* base, int sn, int sk, int rows, int kv) {
    tab.c[n++] = PackChunk{base, sn, sk, rows, kv, off};
    off += (uint32_t)rows * kChunkK * 2;
  };
  // ---- forward: B[n][k] = W[n][k0+k] -------------------------------------------------------------
  add(p->xyz_w[0], 63, 1, 256, 63);                                        // L0 enc
  for (int l = 1; l < 8; ++l) {
    if (l == 4) {
      add(p->xyz_w[4], 319, 1, 256, 63);                                   // skip: encoding first (nerf.py:108-109)
      for (int c = 0; c < 4; ++c) add(p->xyz_w[4] + 63 + 64 * c, 319, 1, 256, 64);
    } else {
      for (int c = 0; c < 4; ++c) add(p->xyz_w[l] + 64 * c, 256, 1, 256, 64);
    }
  }
  for (int c = 0; c < 4; ++c) add(p->final_w + 64 * c, 256, 1, 256, 64);   // L8
  for (int c = 0; c < 4; ++c) add(p->dir_w + 64 * c, 283, 1, 128, 64);     // L9 feature part
  add(p->dir_w + 256, 283, 1, 128, 27);                                    // L9 direction part
  NRF_REQUIRE(n == kFwdChunksFull && off == kPackedBwdOff, "internal: forward blob layout");
  // ---- dgrad: B'[n'][k'] = W[k0+k'][c0+n'] --------------------------------------------------------
  for (int c = 0; c < 2; ++c) add(p->dir_w + (int64_t)64 * c * 283, 1, 283, 256, 64);       // d feat <- d dmid
  for (int c = 0; c < 4; ++c) add(p->final_w + (int64_t)64 * c * 256, 1, 256, 256, 64);     // d h7  <- d feat
  for (int l = 7; l >= 1; --l) {
    const int ld = (l == 4) ? 319 : 256, c0 = (l == 4) ? 63 : 0;
    for (int c = 0; c < 4; ++c) add(p->xyz_w[l] + (int64_t)64 * c * ld + c0, 1, ld, 256, 64);
  }
  NRF_REQUIRE(n == kMaxPackChunks && off == kPackedAuxOff, "internal: dgrad blob layout");
  tab.n = n;
  cudaStream_t st = as_stream(stream);
  pack_chunks_kernel<<<n, 256, 0, st>>>(tab, static_cast<uint8_t*>(packed));
  NRF_LAUNCH_CHECK();
  AuxSrc a{};
  for (int i = 0; i < 8; ++i) a.bias[i] = p->xyz_b[i];
  a.bias_final = p->final_b; a.bias_dir = p->dir_b; a.w_sigma = p->sigma_w; a.w_rgb = p->rgb_w;
  a.b_sigma = p->sigma_b; a.b_rgb = p->rgb_b;
  pack_aux_kernel<<<8, 256, 0, st>>>(a, reinterpret_cast<float*>(static_cast<uint8_t*>(packed) + kPackedAuxOff));
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
