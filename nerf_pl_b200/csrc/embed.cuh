// Internal declarations shared between embed.cu, the fp32 MLP and the tensor-core MLP.
#pragma once
#include "common.cuh"

namespace nrf {

// Component c of the sample point o + d*z (models/rendering.py:234-235): multiply then add,
// each rounded to fp32 — no FMA — so that the encoding input matches the reference bit for bit.
__device__ __forceinline__ float sample_point(const float* __restrict__ rays, int64_t ray, float zv, int c) {
  return __fadd_rn(__ldg(rays + ray * 8 + c), __fmul_rn(__ldg(rays + ray * 8 + 3 + c), zv));
}

int launch_encode_samples(const float* rays, const float* z, int64_t n_rays, int S, int Lx, float* enc, int ld_enc,
                          cudaStream_t st);
int launch_encode_dirs(const float* rays, int64_t n_rays, int Ld, float* out, cudaStream_t st);

}  // namespace nrf
