// Positional encoding (models/nerf.py:21-38) as standalone kernels, plus the fused
// "ray -> sample point -> encoding" producer used by the fp32 MLP path
// (models/rendering.py:213,234-235 + nerf.py:33-38).  sinf/cosf are the accurate versions:
// with the fork's near=1/far=200 bounds the arguments reach ~1e5 rad, where the
// fast intrinsics are off by more than the 1e-5 parity budget.
#include "common.cuh"
#include "embed.cuh"

namespace nrf {

// out[r, :] = [x, sin(f0 x), cos(f0 x), sin(f1 x), ...]; channel index 3+(2k+f)*C+c for C channels
__global__ void embed_fwd_kernel(const float* __restrict__ x, int64_t rows, int C, const float* __restrict__ freqs,
                                 int L, float* __restrict__ out) {
  const int OC = C * (2 * L + 1);
  const int64_t total = rows * OC;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / OC;
    const int col = (int)(idx - r * OC);
    float v;
    if (col < C) {
      v = x[r * C + col];
    } else {
      const int q = (col - C) / C, c = (col - C) - q * C;
      const float arg = __fmul_rn(__ldg(freqs + (q >> 1)), x[r * C + c]);
      v = (q & 1) ? cosf(arg) : sinf(arg);
    }
    out[idx] = v;
  }
}

__global__ void embed_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g, int64_t rows, int C,
                                 const float* __restrict__ freqs, int L, float* __restrict__ gx) {
  const int OC = C * (2 * L + 1);
  const int64_t total = rows * C;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / C;
    const int c = (int)(idx - r * C);
    const float xv = x[idx];
    const float* gr = g + r * OC;
    float acc = gr[c];
    for (int k = 0; k < L; ++k) {
      const float f = __ldg(freqs + k);
      float s, co;
      sincosf(__fmul_rn(f, xv), &s, &co);
      acc += f * (co * gr[C + (2 * k) * C + c] - s * gr[C + (2 * k + 1) * C + c]);
    }
    gx[idx] = acc;
  }
}

// rows = n_rays*S; enc (rows, 3*(2*Lx+1)) with leading dimension ld_enc; dir_enc (n_rays, 3*(2*Ld+1))
__global__ void encode_samples_kernel(const float* __restrict__ rays, const float* __restrict__ z, int64_t n_rays,
                                      int S, int Lx, float* __restrict__ enc, int ld_enc) {
  const int OC = 3 * (2 * Lx + 1);
  const int64_t total = n_rays * S * OC;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / OC;
    const int col = (int)(idx - row * OC);
    const int64_t ray = row / S;
    const int c = col < 3 ? col : (col - 3) % 3;
    const float p = sample_point(rays, ray, z[row], c);
    float v = p;
    if (col >= 3) {
      const int q = (col - 3) / 3;
      const float arg = __fmul_rn(exp2f((float)(q >> 1)), p);
      v = (q & 1) ? cosf(arg) : sinf(arg);
    }
    enc[row * ld_enc + col] = v;
  }
}

__global__ void encode_dirs_kernel(const float* __restrict__ rays, int64_t n_rays, int Ld, float* __restrict__ out) {
  const int OC = 3 * (2 * Ld + 1);
  const int64_t total = n_rays * OC;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t ray = idx / OC;
    const int col = (int)(idx - ray * OC);
    const int c = col < 3 ? col : (col - 3) % 3;
    const float p = __ldg(rays + ray * 8 + 3 + c);
    float v = p;
    if (col >= 3) {
      const int q = (col - 3) / 3;
      const float arg = __fmul_rn(exp2f((float)(q >> 1)), p);
      v = (q & 1) ? cosf(arg) : sinf(arg);
    }
    out[idx] = v;
  }
}

static inline int ew_grid(int64_t total) {
  int64_t b = (total + 255) / 256;
  const int64_t cap = (int64_t)kNumSMs * 32;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

int launch_encode_samples(const float* rays, const float* z, int64_t n_rays, int S, int Lx, float* enc, int ld_enc,
                          cudaStream_t st) {
  if (n_rays == 0) return NRF_OK;
  encode_samples_kernel<<<ew_grid(n_rays * S * 3 * (2 * Lx + 1)), 256, 0, st>>>(rays, z, n_rays, S, Lx, enc, ld_enc);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

int launch_encode_dirs(const float* rays, int64_t n_rays, int Ld, float* out, cudaStream_t st) {
  if (n_rays == 0) return NRF_OK;
  encode_dirs_kernel<<<ew_grid(n_rays * 3 * (2 * Ld + 1)), 256, 0, st>>>(rays, n_rays, Ld, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

}  // namespace nrf

using namespace nrf;

extern "C" int nrf_embed_fwd(const float* x, int64_t rows, int channels, const float* freqs, int n_freqs,
                             float* out, nrf_stream_t stream) {
  NRF_REQUIRE(rows >= 0 && channels >= 1 && n_freqs >= 0, "nrf_embed_fwd: bad sizes");
  if (rows == 0) return NRF_OK;
  NRF_REQUIRE(x && out && (freqs || n_freqs == 0), "nrf_embed_fwd: null pointer");
  embed_fwd_kernel<<<ew_grid(rows * channels * (2 * n_freqs + 1)), 256, 0, as_stream(stream)>>>(
      x, rows, channels, freqs, n_freqs, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_embed_bwd(const float* x, const float* grad_out, int64_t rows, int channels, const float* freqs,
                             int n_freqs, float* grad_x, nrf_stream_t stream) {
  NRF_REQUIRE(rows >= 0 && channels >= 1 && n_freqs >= 0, "nrf_embed_bwd: bad sizes");
  if (rows == 0) return NRF_OK;
  NRF_REQUIRE(x && grad_out && grad_x && (freqs || n_freqs == 0), "nrf_embed_bwd: null pointer");
  embed_bwd_kernel<<<ew_grid(rows * channels), 256, 0, as_stream(stream)>>>(x, grad_out, rows, channels, freqs,
                                                                            n_freqs, grad_x);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
