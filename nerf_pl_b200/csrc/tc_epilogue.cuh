// Epilogue building blocks shared by the tensor-core MLP kernels: packed conversions, ReLU sign
// masks, the power-of-two positional encoder, and pipelined TMEM loads.
//
// Instruction budget (the first version spent ~10.7 SASS instructions per accumulator element in
// the epilogue and the tensor pipe waited for it, profiles/r01_ncu_full_baseline_summary.csv):
//   bias add 1 (FADD) + 1/4 (LDS.128)   relu+bf16 pack 1/2 (F2FP.RELU)   mask 1 (SHF funnel)   store 1/8
#pragma once
#include "tc_common.cuh"

namespace nrf {
namespace tc {

// {hi, lo} -> bf16x2 with round-to-nearest-even; lo lands in the low half (lower column)
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// same with ReLU fused into the conversion (max(x, 0); NaN -> canonical NaN)
__device__ __forceinline__ uint32_t cvt_relu_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// Sign-bit shift register: after pushing elements j = 0..31 in order, bit (31 - j) holds the SIGN of
// element j.  relu'(x) = !sign(x) (x = +0 counts as positive: its output is 0 either way and the
// event has measure zero; the reference's relu'(0) = 0).
__device__ __forceinline__ uint32_t push_sign(uint32_t m, float x) { return __funnelshift_l(__float_as_uint(x), m, 1); }
// true iff element j of a 32-element group was positive, given the INVERTED register (~m) stored by the forward
__device__ __forceinline__ bool mask_bit(uint32_t pos_mask, int j) { return (pos_mask >> (31 - j)) & 1u; }

// tcgen05.wait::ld that also names the destination registers, so that no use of them can be
// scheduled above the wait when several loads are kept in flight.
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.wait::ld.sync.aligned;"
      : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
        "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
        "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
        "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
      :
      : "memory");
}

// Row addressing of a [rows][64 x bf16] SWIZZLE_128B chunk: byte offset of 16-byte piece p of the row
struct RowSwz {
  uint32_t base;  // (r >> 3) * 1024 + (r & 7) * 128
  uint32_t r7;    // r & 7
  __device__ __forceinline__ explicit RowSwz(int r) : base((uint32_t)((r >> 3) * 1024 + (r & 7) * 128)), r7((uint32_t)(r & 7)) {}
  __device__ __forceinline__ uint32_t piece(int p) const { return base + (((uint32_t)p ^ r7) << 4); }
};

__device__ __forceinline__ void st_shared_v4(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ---- positional encoding with power-of-two frequencies (models/nerf.py:16-17,21-38) ----------------
// sin/cos(2^k p): 2^k p is exact in fp32, so the reference evaluates sin of exactly that number.  We
// reduce once: r = p / (2 pi) as an unevaluated sum hi + lo (|error| ~ 2^-48 |p|), then per octave
// frac(2^k r) is exact (power-of-two scaling and a - rint(a) are both exact), and the remaining
// argument in [-pi, pi] goes to the SFU (abs error ~ 1e-6, far below the bf16 rounding that follows).
struct TurnArg {
  float hi, lo;
  __device__ __forceinline__ explicit TurnArg(float p) {
    const float c_hi = 0.15915494f;           // fl(1 / (2 pi))
    const float c_lo = 6.4206382e-09f;        // 1/(2 pi) - c_hi
    hi = p * c_hi;
    lo = fmaf(p, c_hi, -hi) + p * c_lo;
  }
  // sin and cos of 2^k * p
  __device__ __forceinline__ void sincos_pow2(int k, float* s, float* c) const {
    const float sc = __int_as_float((127 + k) << 23);   // 2^k
    const float a = hi * sc, b = lo * sc;
    const float t = (a - rintf(a)) + b;                 // turns, |t| <= 0.5 + eps
    const float x = t * 6.28318530717958647692f;
    *s = __sinf(x);
    *c = __cosf(x);
  }
};


// 32 consecutive columns [kPart*32, kPart*32+32) of the encoded row [p, sin(2^0 p), cos(2^0 p), ...]
// (channel index 3 + (2k+f)*3 + c, models/nerf.py:33-38) as 16 bf16x2 words; columns >= 3+6*kL are zero.
template <int kPart, int kL>
__device__ __forceinline__ void encode_half(const float (&p)[3], uint32_t (&w)[16]) {
  constexpr int kFirst = kPart * 32, kLast = kFirst + 31;
  constexpr int kK0 = kFirst < 3 ? 0 : (kFirst - 3) / 6;
  constexpr int kK1 = (kLast < 3 ? 0 : (kLast - 3) / 6) < kL - 1 ? (kLast < 3 ? 0 : (kLast - 3) / 6) : kL - 1;
  constexpr int kNK = kK1 >= kK0 ? kK1 - kK0 + 1 : 0;
  float sn[kNK > 0 ? kNK : 1][3], cs[kNK > 0 ? kNK : 1][3];
  if (kNK > 0) {
    const TurnArg ta[3] = {TurnArg(p[0]), TurnArg(p[1]), TurnArg(p[2])};
#pragma unroll
    for (int k = 0; k < kNK; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) ta[c].sincos_pow2(kK0 + k, &sn[k][c], &cs[k][c]);
  }
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const int col = kFirst + i;
    float x = 0.f;
    if (col < 3) {
      x = p[col];
    } else if (col < 3 + 6 * kL) {
      const int e = col - 3, k = e / 6, f = (e % 6) / 3, c = e % 3;
      x = f ? cs[k - kK0][c] : sn[k - kK0][c];
    }
    v[i] = x;
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) w[i] = cvt_bf16x2(v[2 * i], v[2 * i + 1]);
}

template <int kL>
__device__ __forceinline__ void encode_store(int part, const float (&p)[3], uint32_t tile_saddr, int row) {
  uint32_t w[16];
  if (part == 0) encode_half<0, kL>(p, w);
  else encode_half<1, kL>(p, w);
  const RowSwz rs(row);
#pragma unroll
  for (int q = 0; q < 4; ++q)
    st_shared_v4(tile_saddr + rs.piece(part * 4 + q), w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

__device__ __forceinline__ float4 ld_shared_f4(uint32_t saddr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(saddr));
  return r;
}

// One hidden-layer epilogue for a thread that owns TMEM lane `row` and 128 accumulator columns
// [col0, col0+128): +bias, (ReLU), bf16, store as the next layer's A operand (SWIZZLE_128B chunks of 64
// columns at act_saddr + chunk * 16 KB).  TMEM loads are double-buffered.
//   kMask : also return the "positive" masks of the four 32-column groups (bit 31-j <-> column j)
//   kDot  : also accumulate dot(relu(x), wdot[col0..]) in fp32 (the sigma head, models/nerf.py:104)
template <bool kRelu, bool kMask, bool kDot>
__device__ __forceinline__ void epi_hidden128(uint32_t t_addr, int col0, const RowSwz& rs, uint32_t act_saddr,
                                              uint32_t bias_saddr, uint32_t wdot_saddr, uint32_t (&mw)[4], float& dot) {
  uint32_t va[32], vb[32];
  tmem_ld32(t_addr + col0, va);
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t(&v)[32] = (g & 1) ? vb : va;
    uint32_t(&vn)[32] = (g & 1) ? va : vb;
    tmem_ld_wait_dep(v);
    if (g + 1 < 4) tmem_ld32(t_addr + col0 + (g + 1) * 32, vn);
    const int c0 = col0 + g * 32;
    uint32_t m = 0, pk[16];
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = ld_shared_f4(bias_saddr + (c0 + j) * 4);
      const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
      const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
      if (kMask) { m = push_sign(m, x0); m = push_sign(m, x1); m = push_sign(m, x2); m = push_sign(m, x3); }
      if (kDot) {
        const float4 w = ld_shared_f4(wdot_saddr + (c0 + j) * 4);
        dot += fmaxf(x0, 0.f) * w.x + fmaxf(x1, 0.f) * w.y + fmaxf(x2, 0.f) * w.z + fmaxf(x3, 0.f) * w.w;
      }
      pk[j >> 1] = kRelu ? cvt_relu_bf16x2(x0, x1) : cvt_bf16x2(x0, x1);
      pk[(j >> 1) + 1] = kRelu ? cvt_relu_bf16x2(x2, x3) : cvt_bf16x2(x2, x3);
    }
    if (kMask) mw[g] = ~m;
    const uint32_t chunk = act_saddr + (uint32_t)(c0 >> 6) * 16384u;
    const int p0 = (c0 & 63) >> 3;
#pragma unroll
    for (int p = 0; p < 4; ++p) st_shared_v4(chunk + rs.piece(p0 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
  }
}

// dgrad epilogue for a thread that owns TMEM lane `row` and 128 accumulator columns [col0, col0+128):
// G = (acc [+ dsig * w_sigma]) masked by relu'(h) (mw: positive masks of the four 32-column groups,
// all ones when the layer has no ReLU), stored as bf16 A operand / G image.
template <bool kAddSig>
__device__ __forceinline__ void epi_dgrad128(uint32_t t_addr, int col0, const RowSwz& rs, uint32_t act_saddr,
                                             const uint32_t (&mw)[4], float dsig, uint32_t wsig_saddr) {
  uint32_t va[32], vb[32];
  tmem_ld32(t_addr + col0, va);
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t(&v)[32] = (g & 1) ? vb : va;
    uint32_t(&vn)[32] = (g & 1) ? va : vb;
    tmem_ld_wait_dep(v);
    if (g + 1 < 4) tmem_ld32(t_addr + col0 + (g + 1) * 32, vn);
    const int c0 = col0 + g * 32;
    const uint32_t m = mw[g];
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      float f0 = __uint_as_float(v[j]), f1 = __uint_as_float(v[j + 1]), f2 = __uint_as_float(v[j + 2]),
            f3 = __uint_as_float(v[j + 3]);
      if (kAddSig) {
        const float4 w = ld_shared_f4(wsig_saddr + (c0 + j) * 4);
        f0 = fmaf(dsig, w.x, f0); f1 = fmaf(dsig, w.y, f1); f2 = fmaf(dsig, w.z, f2); f3 = fmaf(dsig, w.w, f3);
      }
      f0 = mask_bit(m, j) ? f0 : 0.f;
      f1 = mask_bit(m, j + 1) ? f1 : 0.f;
      f2 = mask_bit(m, j + 2) ? f2 : 0.f;
      f3 = mask_bit(m, j + 3) ? f3 : 0.f;
      pk[j >> 1] = cvt_bf16x2(f0, f1);
      pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
    }
    const uint32_t chunk = act_saddr + (uint32_t)(c0 >> 6) * 16384u;
    const int p0 = (c0 & 63) >> 3;
#pragma unroll
    for (int p = 0; p < 4; ++p) st_shared_v4(chunk + rs.piece(p0 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
  }
}

// ---- chunk-granular epilogues (pipelined kernels): 32 accumulator columns already in registers ------
// forward: x = acc + bias[c0..c0+31]; optional sign mask / fp32 dot with relu(x); (ReLU) + bf16; stored as
// pieces p0..p0+3 of the thread's row in the 64-column SWIZZLE_128B chunk at chunk_saddr.
template <bool kRelu, bool kMask, bool kDot>
__device__ __forceinline__ void epi_fwd32(const uint32_t (&v)[32], int c0, uint32_t bias_saddr, uint32_t wdot_saddr,
                                          const RowSwz& rs, uint32_t chunk_saddr, int p0, uint32_t& pos_mask, float& dot) {
  uint32_t m = 0, pk[16];
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 b = ld_shared_f4(bias_saddr + (c0 + j) * 4);
    const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
    const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
    if (kMask) { m = push_sign(m, x0); m = push_sign(m, x1); m = push_sign(m, x2); m = push_sign(m, x3); }
    if (kDot) {
      const float4 w = ld_shared_f4(wdot_saddr + (c0 + j) * 4);
      dot += fmaxf(x0, 0.f) * w.x + fmaxf(x1, 0.f) * w.y + fmaxf(x2, 0.f) * w.z + fmaxf(x3, 0.f) * w.w;
    }
    pk[j >> 1] = kRelu ? cvt_relu_bf16x2(x0, x1) : cvt_bf16x2(x0, x1);
    pk[(j >> 1) + 1] = kRelu ? cvt_relu_bf16x2(x2, x3) : cvt_bf16x2(x2, x3);
  }
  if (kMask) pos_mask = ~m;
#pragma unroll
  for (int p = 0; p < 4; ++p) st_shared_v4(chunk_saddr + rs.piece(p0 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
}

// dgrad: G = (acc [+ dsig * w_sigma[c0..]]) masked by the forward's positive mask word.
template <bool kAddSig>
__device__ __forceinline__ void epi_dgrad32(const uint32_t (&v)[32], int c0, uint32_t pos_mask, float dsig,
                                            uint32_t wsig_saddr, const RowSwz& rs, uint32_t chunk_saddr, int p0) {
  uint32_t pk[16];
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    float f0 = __uint_as_float(v[j]), f1 = __uint_as_float(v[j + 1]), f2 = __uint_as_float(v[j + 2]),
          f3 = __uint_as_float(v[j + 3]);
    if (kAddSig) {
      const float4 w = ld_shared_f4(wsig_saddr + (c0 + j) * 4);
      f0 = fmaf(dsig, w.x, f0); f1 = fmaf(dsig, w.y, f1); f2 = fmaf(dsig, w.z, f2); f3 = fmaf(dsig, w.w, f3);
    }
    f0 = mask_bit(pos_mask, j) ? f0 : 0.f;
    f1 = mask_bit(pos_mask, j + 1) ? f1 : 0.f;
    f2 = mask_bit(pos_mask, j + 2) ? f2 : 0.f;
    f3 = mask_bit(pos_mask, j + 3) ? f3 : 0.f;
    pk[j >> 1] = cvt_bf16x2(f0, f1);
    pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) st_shared_v4(chunk_saddr + rs.piece(p0 + p), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
}

// ---- 16-column granularity (16 epilogue warps: 4 per TMEM lane quarter, each owning columns
// [c*64 + sub*16, +16) of every 64-column chunk c) -------------------------------------------------------
// Sign masks: a thread pushes its 16 columns of chunks 0,1 into one word and of chunks 2,3 into a second one;
// column j of chunk c is bit 31 - ((c & 1) * 16 + j) of word c >> 1 (after inversion: 1 = positive).
__device__ __forceinline__ bool mask_bit16(uint32_t pos_word, int c, int j) { return (pos_word >> (31 - ((c & 1) * 16 + j))) & 1u; }

template <bool kRelu, bool kMask, bool kDot>
__device__ __forceinline__ void epi_fwd16(const uint32_t (&v)[16], int c0, uint32_t bias_saddr, uint32_t wdot_saddr,
                                          const RowSwz& rs, uint32_t chunk_saddr, int p0, uint32_t& sign_reg, float& dot) {
  uint32_t pk[8];
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 b = ld_shared_f4(bias_saddr + (c0 + j) * 4);
    const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
    const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
    if (kMask) {
      sign_reg = push_sign(sign_reg, x0); sign_reg = push_sign(sign_reg, x1);
      sign_reg = push_sign(sign_reg, x2); sign_reg = push_sign(sign_reg, x3);
    }
    if (kDot) {
      const float4 w = ld_shared_f4(wdot_saddr + (c0 + j) * 4);
      dot += fmaxf(x0, 0.f) * w.x + fmaxf(x1, 0.f) * w.y + fmaxf(x2, 0.f) * w.z + fmaxf(x3, 0.f) * w.w;
    }
    pk[j >> 1] = kRelu ? cvt_relu_bf16x2(x0, x1) : cvt_bf16x2(x0, x1);
    pk[(j >> 1) + 1] = kRelu ? cvt_relu_bf16x2(x2, x3) : cvt_bf16x2(x2, x3);
  }
  st_shared_v4(chunk_saddr + rs.piece(p0), pk[0], pk[1], pk[2], pk[3]);
  st_shared_v4(chunk_saddr + rs.piece(p0 + 1), pk[4], pk[5], pk[6], pk[7]);
}

// the same arithmetic, the 16 bf16 results returned as 8 packed words (lower column in the low half) instead of stored
template <bool kRelu, bool kMask, bool kDot>
__device__ __forceinline__ void epi_fwd16_pk(const uint32_t (&v)[16], int c0, uint32_t bias_saddr, uint32_t wdot_saddr,
                                             uint32_t& sign_reg, float& dot, uint32_t (&pk)[8]) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 b = ld_shared_f4(bias_saddr + (c0 + j) * 4);
    const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
    const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
    if (kMask) {
      sign_reg = push_sign(sign_reg, x0); sign_reg = push_sign(sign_reg, x1);
      sign_reg = push_sign(sign_reg, x2); sign_reg = push_sign(sign_reg, x3);
    }
    if (kDot) {
      const float4 w = ld_shared_f4(wdot_saddr + (c0 + j) * 4);
      dot += fmaxf(x0, 0.f) * w.x + fmaxf(x1, 0.f) * w.y + fmaxf(x2, 0.f) * w.z + fmaxf(x3, 0.f) * w.w;
    }
    pk[j >> 1] = kRelu ? cvt_relu_bf16x2(x0, x1) : cvt_bf16x2(x0, x1);
    pk[(j >> 1) + 1] = kRelu ? cvt_relu_bf16x2(x2, x3) : cvt_bf16x2(x2, x3);
  }
}

template <bool kAddSig>
__device__ __forceinline__ void epi_dgrad16(const uint32_t (&v)[16], int c, int c0, uint32_t pos_word, float dsig,
                                            uint32_t wsig_saddr, const RowSwz& rs, uint32_t chunk_saddr, int p0) {
  uint32_t pk[8];
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    float f0 = __uint_as_float(v[j]), f1 = __uint_as_float(v[j + 1]), f2 = __uint_as_float(v[j + 2]),
          f3 = __uint_as_float(v[j + 3]);
    if (kAddSig) {
      const float4 w = ld_shared_f4(wsig_saddr + (c0 + j) * 4);
      f0 = fmaf(dsig, w.x, f0); f1 = fmaf(dsig, w.y, f1); f2 = fmaf(dsig, w.z, f2); f3 = fmaf(dsig, w.w, f3);
    }
    f0 = mask_bit16(pos_word, c, j) ? f0 : 0.f;
    f1 = mask_bit16(pos_word, c, j + 1) ? f1 : 0.f;
    f2 = mask_bit16(pos_word, c, j + 2) ? f2 : 0.f;
    f3 = mask_bit16(pos_word, c, j + 3) ? f3 : 0.f;
    pk[j >> 1] = cvt_bf16x2(f0, f1);
    pk[(j >> 1) + 1] = cvt_bf16x2(f2, f3);
  }
  st_shared_v4(chunk_saddr + rs.piece(p0), pk[0], pk[1], pk[2], pk[3]);
  st_shared_v4(chunk_saddr + rs.piece(p0 + 1), pk[4], pk[5], pk[6], pk[7]);
}

// 16 consecutive columns [kPart*16, +16) of the encoded row as 8 bf16x2 words (see encode_half)
template <int kPart, int kL>
__device__ __forceinline__ void encode_cols16(const float (&p)[3], uint32_t (&w)[8]) {
  constexpr int kFirst = kPart * 16, kLast = kFirst + 15;
  constexpr int kK0 = kFirst < 3 ? 0 : (kFirst - 3) / 6;
  constexpr int kKl = kLast < 3 ? 0 : (kLast - 3) / 6;
  constexpr int kK1 = kKl < kL - 1 ? kKl : kL - 1;
  constexpr bool kAny = kLast >= 3 && kK1 >= kK0 && kFirst < 3 + 6 * kL;
  constexpr int kNK = kAny ? kK1 - kK0 + 1 : 0;
  float sn[kNK > 0 ? kNK : 1][3], cs[kNK > 0 ? kNK : 1][3];
  if (kNK > 0) {
    const TurnArg ta[3] = {TurnArg(p[0]), TurnArg(p[1]), TurnArg(p[2])};
#pragma unroll
    for (int k = 0; k < kNK; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) ta[c].sincos_pow2(kK0 + k, &sn[k][c], &cs[k][c]);
  }
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int col = kFirst + i;
    float x = 0.f;
    if (col < 3) {
      x = p[col];
    } else if (col < 3 + 6 * kL) {
      const int e = col - 3, k = e / 6, f = (e % 6) / 3, c = e % 3;
      x = f ? cs[k - kK0][c] : sn[k - kK0][c];
    }
    v[i] = x;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = cvt_bf16x2(v[2 * i], v[2 * i + 1]);
}

// thread (row, part 0..3) writes 16 columns of the encoded row into the chunk at tile_saddr
template <int kL>
__device__ __forceinline__ void encode_store16(int part, const float (&p)[3], uint32_t tile_saddr, int row) {
  uint32_t w[8];
  if (part == 0) encode_cols16<0, kL>(p, w);
  else if (part == 1) encode_cols16<1, kL>(p, w);
  else if (part == 2) encode_cols16<2, kL>(p, w);
  else encode_cols16<3, kL>(p, w);
  const RowSwz rs(row);
  st_shared_v4(tile_saddr + rs.piece(part * 2), w[0], w[1], w[2], w[3]);
  st_shared_v4(tile_saddr + rs.piece(part * 2 + 1), w[4], w[5], w[6], w[7]);
}

}  // namespace tc
}  // namespace nrf
