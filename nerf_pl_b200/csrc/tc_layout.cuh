// Memory layouts shared by the tensor-core MLP kernels (pack, forward, dgrad, wgrad).
//
// Network (models/nerf.py:41-124, D=8 W=256 L=(10,4) skips=[4]) as a chain of UMMA GEMMs on
// 128-row tiles of sample points; K is consumed in 64-element chunks (= one 128-byte
// SWIZZLE_128B row), N = 256 (128 for the direction layer):
//
//   layer  A operand (K chunks)              N    epilogue
//   L0     enc(1)                            256  +b, ReLU -> act
//   L1-3   act(4)                            256  +b, ReLU -> act
//   L4     enc(1) , act(4)   [skip, enc 1st] 256  +b, ReLU -> act
//   L5-7   act(4)                            256  +b, ReLU -> act ; L7 also sigma = w_s.h + b_s
//   L8     act(4)            [xyz_final]     256  +b        -> act
//   L9     act(4) , dir(1)   [dir_encoding]  128  +b, ReLU ; rgb = sigmoid(W_rgb.d + b_rgb)
//
// Every operand tile, in shared memory AND in the HBM stash, is the same image:
//   [rows][64 x bf16], SWIZZLE_128B, 8-row groups of 1024 B  (16 KB per 128-row chunk).
// Weights are pre-packed into that image once per optimizer step (nrf_mlp_tc_pack), so a
// weight chunk is one linear bulk copy; the stash is written by bulk copies of the activation
// buffer, so wgrad reads it back as MN-major UMMA operands without any transpose.
#pragma once
#include <cstddef>
#include <cstdint>

namespace nrf {
namespace tc {

constexpr int kTileM = 128;
constexpr int kW = 256;
constexpr int kChunkK = 64;
constexpr uint32_t kChunkBytes256 = 256 * kChunkK * 2;  // 32 KB: weight chunk with N = 256
constexpr uint32_t kChunkBytes128 = 128 * kChunkK * 2;  // 16 KB: N = 128, and one activation chunk
constexpr uint32_t kActChunkBytes = kTileM * kChunkK * 2;  // 16 KB

// ---- forward weight blob: chunks in consumption order -------------------------------------------
constexpr int kFwdChunksSigma = 1 + 4 * 3 + 5 + 4 * 3;       // 30: L0..L7
constexpr int kFwdChunksFull = kFwdChunksSigma + 4 + 5;       // 39: + L8, L9
constexpr size_t kFwdBlobBytes = (size_t)(kFwdChunksSigma + 4) * kChunkBytes256 + 5 * (size_t)kChunkBytes128;

// ---- dgrad weight blob (transposed weights): L9feat(2) L8(4) L7..L1 (4 each; L4 = h part) --------
constexpr int kBwdChunks = 2 + 4 + 7 * 4;                     // 34, all N' = 256
constexpr size_t kBwdBlobBytes = (size_t)kBwdChunks * kChunkBytes256;

// ---- fp32 side vectors ----------------------------------------------------------------------------
struct AuxOffsets {  // in floats
  static constexpr int bias = 0;                  // 8 x 256 hidden-layer biases
  static constexpr int bias_final = 8 * 256;      // 256
  static constexpr int bias_dir = bias_final + 256;  // 128
  static constexpr int w_sigma = bias_dir + 128;  // 256
  static constexpr int w_rgb = w_sigma + 256;     // 3 x 128
  static constexpr int b_sigma = w_rgb + 384;     // 1
  static constexpr int b_rgb = b_sigma + 1;       // 3
  static constexpr int count = ((b_rgb + 3 + 31) / 32) * 32;
};

constexpr size_t kPackedFwdOff = 0;
constexpr size_t kPackedBwdOff = kFwdBlobBytes;
constexpr size_t kPackedAuxOff = kPackedBwdOff + kBwdBlobBytes;
constexpr size_t kPackedBytes = kPackedAuxOff + (size_t)AuxOffsets::count * 4;

// ---- per-tile stash written by the training forward ----------------------------------------------
// images: enc(1 chunk) dir(1) h0..h7 (4 each) feat(4) dmid(2)  = 40 chunks of 16 KB
constexpr int kStashEnc = 0, kStashDir = 1, kStashH0 = 2, kStashFeat = 2 + 32, kStashMid = kStashFeat + 4;
constexpr int kStashChunksFull = kStashMid + 2;   // 40
constexpr int kStashChunksSigma = kStashFeat;     // 34 (no feat / dmid; dir slot unused)
// ReLU sign masks: per tile, 8 hidden layers x 128 rows x 8 words, then dir layer 128 rows x 4 words.
// A thread of the pipelined kernels owns one row and the 32-column group `sub` (0/1) of every 64-column chunk c:
// its words are stored contiguously, hidden layers word index = sub * 4 + c, dir layer = sub * 2 + c;
// bit (31 - j) of a word <-> column c * 64 + sub * 32 + j positive (tc_epilogue.cuh: push_sign / mask_bit).
constexpr int kMaskWordsFull = 8 * 128 * 8 + 128 * 4;
constexpr int kMaskWordsSigma = 8 * 128 * 8;

// ---- per-tile G stash written by the dgrad kernel (dL/d pre-activation images) ---------------------
// G0..G7 at chunk 4*l, G8 (xyz_final) at 32..35, G9 (dir layer, 128 wide) at 36..37
constexpr int kGstashG8 = 32, kGstashG9 = 36;
__host__ __device__ inline size_t gstash_tile_bytes(bool full) { return (size_t)(full ? 38 : 32) * kActChunkBytes; }

__host__ __device__ inline size_t stash_tile_bytes(bool full) {
  return (size_t)(full ? kStashChunksFull : kStashChunksSigma) * kActChunkBytes;
}
// Stash addressing is IMAGE-major: an image = the chunks one bulk store writes for one tile (enc 1, dir 1, h_l 4,
// feat 4, dmid 2; G_l 4, G9 2), and all tiles' copies of an image are contiguous.  The wgrad kernel streams one
// image (or two) over many tiles, so each of its streams walks a dense region of HBM (tools/stream_bench.py:
// 6.6 TB/s versus 4.9 TB/s for the tile-major order).  slot_start = index of the image's first chunk.
__host__ __device__ inline size_t img_off(int slot_start, int img_chunks, int tile, int num_tiles) {
  return ((size_t)slot_start * (size_t)num_tiles + (size_t)tile * (size_t)img_chunks) * kActChunkBytes;
}

__host__ __device__ inline size_t mask_tile_words(bool full) { return full ? kMaskWordsFull : kMaskWordsSigma; }

}  // namespace tc
}  // namespace nrf
