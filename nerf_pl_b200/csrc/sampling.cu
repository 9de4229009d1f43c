// Ray-depth sampling kernels: stratified coarse sampling, the fork's sample_pdf, per-ray sort,
// and the row-wise searchsorted that replaces the torchsearchsorted extension.
//
// All of these are HBM-bound streaming kernels (a few bytes of arithmetic per byte moved):
// one warp per ray, coalesced lane-strided global access, per-warp shared-memory staging of
// the ray's cdf / depth list, warp-scan primitives.  Arithmetic that feeds the bin search is
// written with explicit round-to-nearest intrinsics (no FMA contraction) so that, given the
// same inputs, results are bit-identical to the reference's fp32 torch ops.
#include <cfloat>
#include <cmath>
#include "common.cuh"

namespace nrf {

constexpr int kWarpsPerBlock = 4;

// ---------------------------------------------------------------------------------------------
// a4  models/rendering.py:216-232
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float lerp_depth(float near, float far, float t, bool use_disp) {
  const float omt = __fsub_rn(1.0f, t);
  if (use_disp) {  // 1/(1/near*(1-t) + 1/far*t), rendering.py:221
    const float a = __fmul_rn(__fdiv_rn(1.0f, near), omt);
    const float b = __fmul_rn(__fdiv_rn(1.0f, far), t);
    return __fdiv_rn(1.0f, __fadd_rn(a, b));
  }
  return __fadd_rn(__fmul_rn(near, omt), __fmul_rn(far, t));  // rendering.py:219
}

__global__ void sample_coarse_kernel(const float* __restrict__ rays, const float* __restrict__ t_steps,
                                     int64_t n_rays, int S, int use_disp, float perturb,
                                     const float* __restrict__ u, float* __restrict__ z_out) {
  const int64_t total = n_rays * S;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / S;
    const int i = (int)(idx - r * S);
    const float near = __ldg(rays + r * 8 + 6), far = __ldg(rays + r * 8 + 7);
    const float zi = lerp_depth(near, far, __ldg(t_steps + i), use_disp);
    float z = zi;
    if (perturb > 0.f) {  // rendering.py:225-232
      float lower = zi, upper = zi;
      if (i > 0) {
        const float zm = lerp_depth(near, far, __ldg(t_steps + i - 1), use_disp);
        lower = __fmul_rn(0.5f, __fadd_rn(zm, zi));
      }
      if (i < S - 1) {
        const float zp = lerp_depth(near, far, __ldg(t_steps + i + 1), use_disp);
        upper = __fmul_rn(0.5f, __fadd_rn(zi, zp));
      }
      const float pr = __fmul_rn(perturb, u[idx]);
      z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), pr));
    }
    z_out[idx] = z;
  }
}

// ---------------------------------------------------------------------------------------------
// a11  models/rendering.py:14-48.  One warp per ray; the cdf lives in shared memory.
// torch.cumsum on CPU accumulates fp32 inputs in double and rounds each prefix to fp32
// (verified in the build container); we do the same, so the cdf is bit-identical whenever the
// normaliser sum(w) is.
// ---------------------------------------------------------------------------------------------
__global__ void sample_pdf_kernel(const float* __restrict__ rays, const float* __restrict__ weights,
                                  int64_t n_rays, int n_coarse, int n_imp, const float* __restrict__ u,
                                  const float* __restrict__ jitter, float eps, float* __restrict__ z_out,
                                  int64_t* __restrict__ idx_out, float* __restrict__ cdf_out) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_bins = n_coarse - 2;
  float* cdf = smem + warp * (n_bins + 1);
  const int chunk = (n_bins + 31) / 32;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    const float* w = weights + ray * n_coarse + 1;  // weights[:, 1:-1]
    float part = 0.f;
    for (int k = lane; k < n_bins; k += 32) {
      const float wk = __fadd_rn(__ldg(w + k), eps);
      cdf[k + 1] = wk;
      part += wk;
    }
    const float total = warp_sum(part);
    __syncwarp();
    // lane-local sequential prefix in double over a contiguous chunk, then warp scan of chunk sums
    const int k0 = lane * chunk, k1 = min(k0 + chunk, n_bins);
    double acc = 0.0;
    for (int k = k0; k < k1; ++k) acc += (double)__fdiv_rn(cdf[k + 1], total);
    double incl = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    double run = incl - acc;  // exclusive offset of this lane's chunk
    __syncwarp();
    for (int k = k0; k < k1; ++k) {
      run += (double)__fdiv_rn(cdf[k + 1], total);
      cdf[k + 1] = (float)run;  // each lane only rewrites its own chunk after reading it
    }
    if (lane == 0) cdf[0] = 0.f;
    __syncwarp();
    if (cdf_out != nullptr)
      for (int k = lane; k <= n_bins; k += 32) cdf_out[ray * (n_bins + 1) + k] = cdf[k];
    const float near = __ldg(rays + ray * 8 + 6), far = __ldg(rays + ray * 8 + 7);
    const float nb = (float)n_bins;
    for (int j = lane; j < n_imp; j += 32) {
      const float uj = __ldg(u + ray * n_imp + j);
      int lo = 0, hi = n_bins + 1;  // count of cdf entries <= uj (searchsorted side='right')
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] <= uj) lo = mid + 1; else hi = mid;
      }
      const float ind = fmaxf((float)lo - 1.0f, 0.0f);  // rendering.py:37-39
      const float t = __fdiv_rn(__fadd_rn(ind, __ldg(jitter + ray * n_imp + j)), nb);  // :41
      z_out[ray * n_imp + j] = __fadd_rn(__fmul_rn(near, __fsub_rn(1.0f, t)), __fmul_rn(far, t));  // :45
      if (idx_out != nullptr) idx_out[ray * n_imp + j] = (int64_t)ind;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// a13  models/rendering.py:257 — per-ray ascending sort of the concatenated depths (bitonic
// network in shared memory, one warp per ray, padded with +inf to a power of two).
// ---------------------------------------------------------------------------------------------
__global__ void merge_sort_kernel(const float* __restrict__ zc, const float* __restrict__ zf, int64_t n_rays,
                                  int nc, int nf, int P, float* __restrict__ z_out) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* s = smem + warp * P;
  const int n = nc + nf;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    for (int i = lane; i < P; i += 32) {
      float v = INFINITY;
      if (i < nc) v = __ldg(zc + ray * nc + i);
      else if (i < n) v = __ldg(zf + ray * nf + (i - nc));
      s[i] = v;
    }
    __syncwarp();
    for (int k = 2; k <= P; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < (P >> 1); t += 32) {
          const int i = 2 * t - (t & (j - 1));
          const int l = i + j;
          const float a = s[i], b = s[l];
          const bool asc = (i & k) == 0;
          if ((a > b) == asc) { s[i] = b; s[l] = a; }
        }
        __syncwarp();
      }
    }
    for (int i = lane; i < n; i += 32) z_out[ray * n + i] = s[i];
    __syncwarp();
  }
}

// Same result for the shape render_rays produces (rendering.py:257: the coarse depths are already ascending): bitonic
// sort of the nf fine depths only, then a rank merge - a coarse element goes to i + #(fine < it), a fine element to
// j + #(coarse <= it), two binary searches in shared memory.  The output is the same sorted multiset bit for bit.  Rays
// whose coarse part is NOT ascending (callers outside render_rays) take the full network of merge_sort_kernel's body.
__global__ void merge_sorted_kernel(const float* __restrict__ zc, const float* __restrict__ zf, int64_t n_rays,
                                    int nc, int nf, int Pf, int P, float* __restrict__ z_out) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = nc + nf;
  const int per_warp = (nc + Pf > P ? nc + Pf : P) + n;
  float* sc = smem + warp * per_warp;   // [nc] coarse, then [Pf] fine (also the [P] scratch of the fallback)
  float* sf = sc + nc;
  float* so = sc + (nc + Pf > P ? nc + Pf : P);   // [n] merged
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    bool ok = true;
    for (int i = lane; i < nc; i += 32) sc[i] = __ldg(zc + ray * nc + i);
    for (int i = lane; i < Pf; i += 32) sf[i] = i < nf ? __ldg(zf + ray * nf + i) : INFINITY;
    __syncwarp();
    for (int i = lane; i + 1 < nc; i += 32) ok = ok && (sc[i] <= sc[i + 1]);
    ok = __all_sync(0xffffffffu, ok);
    if (ok) {
      for (int k = 2; k <= Pf; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
          for (int t = lane; t < (Pf >> 1); t += 32) {
            const int i = 2 * t - (t & (j - 1));
            const int l = i + j;
            const float a = sf[i], b = sf[l];
            const bool asc = (i & k) == 0;
            if ((a > b) == asc) { sf[i] = b; sf[l] = a; }
          }
          __syncwarp();
        }
      }
      for (int i = lane; i < nc; i += 32) {
        const float v = sc[i];
        int lo = 0, hi = nf;                       // #(fine < v)
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (sf[mid] < v) lo = mid + 1; else hi = mid; }
        so[i + lo] = v;
      }
      for (int j = lane; j < nf; j += 32) {
        const float v = sf[j];
        int lo = 0, hi = nc;                       // #(coarse <= v)
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (sc[mid] <= v) lo = mid + 1; else hi = mid; }
        so[j + lo] = v;
      }
      __syncwarp();
      for (int i = lane; i < n; i += 32) z_out[ray * n + i] = so[i];
    } else {
      float* s = sc;                               // full bitonic network over the concatenation, as merge_sort_kernel
      __syncwarp();
      for (int i = lane; i < P; i += 32) {
        float v = INFINITY;
        if (i < nc) v = __ldg(zc + ray * nc + i);
        else if (i < n) v = __ldg(zf + ray * nf + (i - nc));
        s[i] = v;
      }
      __syncwarp();
      for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
          for (int t = lane; t < (P >> 1); t += 32) {
            const int i = 2 * t - (t & (j - 1));
            const int l = i + j;
            const float a = s[i], b = s[l];
            const bool asc = (i & k) == 0;
            if ((a > b) == asc) { s[i] = b; s[l] = a; }
          }
          __syncwarp();
        }
      }
      for (int i = lane; i < n; i += 32) z_out[ray * n + i] = s[i];
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// a12  torchsearchsorted (searchsorted_cuda_kernel.cu:83-107): one thread per query.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void searchsorted_kernel(const T* __restrict__ a, const T* __restrict__ v,
                                    int64_t* __restrict__ out, int64_t nrow_a, int64_t nrow_v,
                                    int64_t ncol_a, int64_t ncol_v, int64_t nrow_out, int side_left) {
  const int64_t total = nrow_out * ncol_v;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / ncol_v, col = idx - row * ncol_v;
    const T* ar = a + (nrow_a == 1 ? 0 : row) * ncol_a;
    const T val = v[(nrow_v == 1 ? 0 : row) * ncol_v + col];
    int64_t lo = 0, hi = ncol_a;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      const T am = __ldg(ar + mid);
      const bool go_right = side_left ? (am < val) : (am <= val);
      if (go_right) lo = mid + 1; else hi = mid;
    }
    out[idx] = lo;
  }
}

static inline int grid_for(int64_t work_items, int per_block, int max_blocks = kNumSMs * 16) {
  int64_t b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

}  // namespace nrf

using namespace nrf;

extern "C" int nrf_sample_coarse(const float* rays, const float* t_steps, int64_t n_rays, int n_samples,
                                 int use_disp, float perturb, const float* u, float* z_out, nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_sample_coarse: bad sizes n_rays=%lld n_samples=%d",
              (long long)n_rays, n_samples);
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(rays && t_steps && z_out, "nrf_sample_coarse: null pointer");
  NRF_REQUIRE(!(perturb > 0.f) || u != nullptr, "nrf_sample_coarse: perturb > 0 needs the uniform tensor u");
  sample_coarse_kernel<<<grid_for(n_rays * n_samples, 256), 256, 0, as_stream(stream)>>>(
      rays, t_steps, n_rays, n_samples, use_disp, perturb, u, z_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_sample_pdf(const float* rays, const float* weights_coarse, int64_t n_rays, int n_coarse,
                              int n_importance, const float* u, const float* jitter, float eps,
                              float* z_fine_out, int64_t* idx_out, float* cdf_out, nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && n_coarse >= 3 && n_importance >= 1,
              "nrf_sample_pdf: need n_coarse >= 3 and n_importance >= 1 (got %d, %d)", n_coarse, n_importance);
  NRF_REQUIRE(n_coarse <= 4096, "nrf_sample_pdf: n_coarse=%d > 4096 unsupported", n_coarse);
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(rays && weights_coarse && u && jitter && z_fine_out, "nrf_sample_pdf: null pointer");
  const size_t smem = (size_t)kWarpsPerBlock * (n_coarse - 1) * sizeof(float);
  if (smem > 48 * 1024)
    NRF_CUDA(cudaFuncSetAttribute(sample_pdf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sample_pdf_kernel<<<grid_for(n_rays, kWarpsPerBlock), kWarpsPerBlock * 32, smem, as_stream(stream)>>>(
      rays, weights_coarse, n_rays, n_coarse, n_importance, u, jitter, eps, z_fine_out, idx_out, cdf_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_merge_sort(const float* z_coarse, const float* z_fine, int64_t n_rays, int nc, int nf,
                              float* z_out, nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && nc >= 0 && nf >= 0 && nc + nf >= 1, "nrf_merge_sort: bad sizes");
  NRF_REQUIRE(nc + nf <= 8192, "nrf_merge_sort: %d samples per ray > 8192 unsupported", nc + nf);
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE((z_coarse || nc == 0) && (z_fine || nf == 0) && z_out, "nrf_merge_sort: null pointer");
  int P = 2;
  while (P < nc + nf) P <<= 1;
  if (nc >= 2 && nf >= 1) {
    int Pf = 2;
    while (Pf < nf) Pf <<= 1;
    const size_t per_warp = (size_t)((nc + Pf > P ? nc + Pf : P) + nc + nf);
    const size_t smem2 = (size_t)kWarpsPerBlock * per_warp * sizeof(float);
    if (smem2 <= 200 * 1024) {
      if (smem2 > 48 * 1024)
        NRF_CUDA(cudaFuncSetAttribute(merge_sorted_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
      merge_sorted_kernel<<<grid_for(n_rays, kWarpsPerBlock), kWarpsPerBlock * 32, smem2, as_stream(stream)>>>(
          z_coarse, z_fine, n_rays, nc, nf, Pf, P, z_out);
      NRF_LAUNCH_CHECK();
      return NRF_OK;
    }
  }
  const size_t smem = (size_t)kWarpsPerBlock * P * sizeof(float);
  if (smem > 48 * 1024)
    NRF_CUDA(cudaFuncSetAttribute(merge_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  merge_sort_kernel<<<grid_for(n_rays, kWarpsPerBlock), kWarpsPerBlock * 32, smem, as_stream(stream)>>>(
      z_coarse, z_fine, n_rays, nc, nf, P, z_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

template <typename T>
static int searchsorted_launch(const void* a, const void* v, int64_t* out, int64_t nrow_a, int64_t nrow_v, int64_t ncol_a,
                               int64_t ncol_v, int side_left, cudaStream_t st) {
  const int64_t nrow_out = nrow_a > nrow_v ? nrow_a : nrow_v;
  searchsorted_kernel<T><<<grid_for(nrow_out * ncol_v, 256), 256, 0, st>>>(
      static_cast<const T*>(a), static_cast<const T*>(v), out, nrow_a, nrow_v, ncol_a, ncol_v, nrow_out, side_left);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_searchsorted_typed(const void* a, const void* v, int64_t* out, int64_t nrow_a, int64_t nrow_v,
                                      int64_t ncol_a, int64_t ncol_v, int side_left, int dtype, nrf_stream_t stream) {
  // same contract as torchsearchsorted/src/torchsearchsorted/searchsorted.py:23-31
  NRF_REQUIRE(nrow_a >= 1 && nrow_v >= 1 && ncol_a >= 1 && ncol_v >= 0, "nrf_searchsorted: bad sizes");
  NRF_REQUIRE(nrow_a == nrow_v || nrow_a == 1 || nrow_v == 1,
              "`a` and `v` must have the same number of rows or one of them must have only one");
  if (ncol_v == 0) return NRF_OK;
  NRF_REQUIRE(a && v && out, "nrf_searchsorted: null pointer");
  cudaStream_t st = as_stream(stream);
  switch (dtype) {   // AT_DISPATCH_ALL_TYPES (searchsorted_cuda_kernel.cu:134)
    case NRF_DT_F32: return searchsorted_launch<float>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_F64: return searchsorted_launch<double>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_I64: return searchsorted_launch<long long>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_I32: return searchsorted_launch<int>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_I16: return searchsorted_launch<short>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_I8: return searchsorted_launch<signed char>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
    case NRF_DT_U8: return searchsorted_launch<unsigned char>(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, st);
  }
  set_error("nrf_searchsorted: \"searchsorted cuda\" not implemented for dtype code %d", dtype);
  return NRF_ERR_INVALID;
}

extern "C" int nrf_searchsorted(const float* a, const float* v, int64_t* out, int64_t nrow_a, int64_t nrow_v,
                                int64_t ncol_a, int64_t ncol_v, int side_left, nrf_stream_t stream) {
  return nrf_searchsorted_typed(a, v, out, nrow_a, nrow_v, ncol_a, ncol_v, side_left, NRF_DT_F32, stream);
}
