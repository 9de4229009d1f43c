// Alpha-compositing integral, forward and backward (models/rendering.py:170-198; disparity
// from rendering_shadows.py:198 / rendering_rgb_sm.py:197).
//
// One warp per ray.  Sample i of a ray is handled by lane i%32 in chunk i/32, so every global
// access is a coalesced 128-byte row segment; transmittance is an exclusive product scan
// (__shfl_up) inside a chunk with a running carry between chunks.  HBM-bound: forward moves
// 24(+4 noise) B/sample, backward 40 B/sample; nothing is re-read.
//
// Backward uses the same formula autograd derives for cumprod (reverse cumulative sum of
// grad*output divided by the input) — it is evaluated as a true reverse scan, never as
// "total - prefix", because the factors can be as small as 1e-10 (opaque samples).
#include <cmath>
#include "common.cuh"

namespace nrf {

constexpr int kCompWarps = 4;

struct SampleTerms {
  float alpha, factor, delta_e;  // alpha_i, (1-alpha_i+1e-10), delta_i*exp(-delta_i*relu(raw_i))
  bool active;                   // raw_i > 0 (ReLU gate)
};

__device__ __forceinline__ SampleTerms sample_terms(float sigma, float nz, float zi, float znext, bool last,
                                                    float dnorm) {
  SampleTerms t;
  float delta = last ? 1e10f : __fsub_rn(znext, zi);  // rendering.py:170-172
  delta = __fmul_rn(delta, dnorm);                    // :176
  const float raw = __fadd_rn(sigma, nz);
  const float r = fmaxf(raw, 0.f);
  const float e = expf(-__fmul_rn(delta, r));          // :181
  t.alpha = __fsub_rn(1.0f, e);
  t.factor = __fadd_rn(__fsub_rn(1.0f, t.alpha), 1e-10f);  // :183
  t.delta_e = delta * e;
  t.active = raw > 0.f;
  return t;
}

__device__ __forceinline__ float warp_excl_prod(float f, int lane, float* total) {
  float incl = f;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl *= up;
  }
  *total = __shfl_sync(0xffffffffu, incl, 31);
  const float ex = __shfl_up_sync(0xffffffffu, incl, 1);
  return lane == 0 ? 1.0f : ex;
}

__device__ __forceinline__ float disp_from(float depth, float opacity) {
  const float q = depth / opacity;  // NaN when opacity == 0, as in the reference
  const float m = (q > 1e-10f || isnan(q)) ? q : 1e-10f;
  return 1.0f / m;
}

struct ChunkIn {
  float zi, zn, sg, nz, r, g, b;
};

// Lane-strided reads of the 12-byte rgb records: measured FASTER than three coalesced pieces transposed through shared
// memory (fwd 34.9 vs 37.8 us, bwd 54 vs 62 us at 32768 x 128): L1 serves the three passes over a line.
__device__ __forceinline__ ChunkIn load_chunk(const float* __restrict__ sg, const float* __restrict__ zz,
                                              const float* __restrict__ rgb, const float* __restrict__ noise,
                                              int64_t ray, int S, int lane, int c) {
  ChunkIn in;
  in.zi = in.zn = in.sg = in.nz = in.r = in.g = in.b = 0.f;
  const int i = c * 32 + lane;
  if (i < S) {
    in.zi = zz[i];
    in.zn = (i + 1 < S) ? zz[i + 1] : 0.f;
    in.sg = sg[i];
    if (noise) in.nz = noise[ray * S + i];
    if (rgb) {
      const float* c3 = rgb + (ray * S + i) * 3;
      in.r = c3[0]; in.g = c3[1]; in.b = c3[2];
    }
  }
  return in;
}

__global__ void composite_fwd_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb,
                                     const float* __restrict__ z, const float* __restrict__ rays,
                                     const float* __restrict__ noise, float noise_std, int white_back,
                                     int64_t n_rays, int S, float* __restrict__ weights_out,
                                     float* __restrict__ rgb_out, float* __restrict__ depth_out,
                                     float* __restrict__ opacity_out, float* __restrict__ disp_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunks = (S + 31) >> 5;
  for (int64_t ray = blockIdx.x * (int64_t)kCompWarps + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kCompWarps) {
    const float dx = __ldg(rays + ray * 8 + 3), dy = __ldg(rays + ray * 8 + 4), dz = __ldg(rays + ray * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float* sg = sigma + ray * S;
    const float* zz = z + ray * S;
    float carry = 1.0f, a_r = 0.f, a_g = 0.f, a_b = 0.f, a_d = 0.f, a_w = 0.f;
    // software pipeline: the loads of chunk c+1 are in flight while chunk c is scanned (the scan is a dependent chain of
    // five shuffles; without the prefetch a warp has one chunk's loads outstanding at a time)
    ChunkIn cur = load_chunk(sg, zz, rgb, noise, ray, S, lane, 0);
    for (int c = 0; c < chunks; ++c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      ChunkIn nxt = cur;
      if (c + 1 < chunks) nxt = load_chunk(sg, zz, rgb, noise, ray, S, lane, c + 1);
      float w = 0.f, zi = cur.zi, f = 1.0f, alpha = 0.f;
      if (valid) {
        const SampleTerms t = sample_terms(cur.sg, cur.nz * noise_std, cur.zi, cur.zn, i == S - 1, dnorm);
        f = t.factor;
        alpha = t.alpha;
      }
      float chunk_total;
      const float ex = warp_excl_prod(f, lane, &chunk_total);
      if (valid) {
        w = alpha * (carry * ex);  // :185
        weights_out[ray * S + i] = w;
        a_w += w;
        a_d += w * zi;
        if (rgb) {
          a_r += w * cur.r;
          a_g += w * cur.g;
          a_b += w * cur.b;
        }
      }
      carry *= chunk_total;
      cur = nxt;
    }
    a_w = warp_sum(a_w);
    a_d = warp_sum(a_d);
    if (rgb) { a_r = warp_sum(a_r); a_g = warp_sum(a_g); a_b = warp_sum(a_b); }
    if (lane == 0) {
      opacity_out[ray] = a_w;
      depth_out[ray] = a_d;
      if (disp_out) disp_out[ray] = disp_from(a_d, a_w);
      if (rgb && rgb_out) {
        const float bg = white_back ? (1.0f - a_w) : 0.f;  // :195-196
        rgb_out[ray * 3 + 0] = a_r + bg;
        rgb_out[ray * 3 + 1] = a_g + bg;
        rgb_out[ray * 3 + 2] = a_b + bg;
      }
    }
  }
}

__global__ void composite_bwd_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb,
                                     const float* __restrict__ z, const float* __restrict__ rays,
                                     const float* __restrict__ noise, float noise_std, int white_back,
                                     int64_t n_rays, int S, const float* __restrict__ g_rgb,
                                     const float* __restrict__ g_depth, const float* __restrict__ g_opacity,
                                     const float* __restrict__ g_disp, const float* __restrict__ g_weights,
                                     float* __restrict__ d_sigma, float* __restrict__ d_rgb) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunks = (S + 31) >> 5;  // <= 32 (checked by the launcher)
  for (int64_t ray = blockIdx.x * (int64_t)kCompWarps + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kCompWarps) {
    const float dx = __ldg(rays + ray * 8 + 3), dy = __ldg(rays + ray * 8 + 4), dz = __ldg(rays + ray * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float* sg = sigma + ray * S;
    const float* zz = z + ray * S;
    // sweep 1: chunk-start transmittances (lane c keeps chunk c's) and the ray totals
    float carry = 1.0f, my_carry = 1.0f, a_d = 0.f, a_w = 0.f;
    for (int c = 0; c < chunks; ++c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      float f = 1.0f, alpha = 0.f, zi = 0.f;
      if (valid) {
        zi = zz[i];
        const float zn = (i + 1 < S) ? zz[i + 1] : 0.f;
        const float nzv = noise ? noise[ray * S + i] * noise_std : 0.f;
        const SampleTerms t = sample_terms(sg[i], nzv, zi, zn, i == S - 1, dnorm);
        f = t.factor;
        alpha = t.alpha;
      }
      if (lane == c) my_carry = carry;
      float chunk_total;
      const float ex = warp_excl_prod(f, lane, &chunk_total);
      if (valid) {
        const float w = alpha * (carry * ex);
        a_w += w;
        a_d += w * zi;
      }
      carry *= chunk_total;
    }
    a_w = warp_sum(a_w);
    a_d = warp_sum(a_d);
    // gradients of the per-ray outputs w.r.t. the three sums (rgb handled per channel)
    float gr = 0.f, gg = 0.f, gb = 0.f;
    if (rgb && g_rgb) { gr = g_rgb[ray * 3]; gg = g_rgb[ray * 3 + 1]; gb = g_rgb[ray * 3 + 2]; }
    float gd = g_depth ? g_depth[ray] : 0.f;
    float go = g_opacity ? g_opacity[ray] : 0.f;
    if (white_back && rgb) go -= (gr + gg + gb);
    if (g_disp) {
      const float q = a_d / a_w;
      if (q > 1e-10f || isnan(q)) {  // disp = 1/q ; q = depth/opacity
        const float gq = -g_disp[ray] / (q * q);
        gd += gq / a_w;
        go += -gq * a_d / (a_w * a_w);
      }
    }
    // sweep 2 (reverse): suffix sums of w_j*G_j, then d alpha -> d sigma
    float suffix = 0.f;  // sum over samples in later chunks
    for (int c = chunks - 1; c >= 0; --c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      SampleTerms t;
      t.alpha = 0.f; t.factor = 1.0f; t.delta_e = 0.f; t.active = false;
      float zi = 0.f;
      if (valid) {
        zi = zz[i];
        const float zn = (i + 1 < S) ? zz[i + 1] : 0.f;
        const float nzv = noise ? noise[ray * S + i] * noise_std : 0.f;
        t = sample_terms(sg[i], nzv, zi, zn, i == S - 1, dnorm);
      }
      const float start = __shfl_sync(0xffffffffu, my_carry, c);
      float chunk_total;
      const float ex = warp_excl_prod(t.factor, lane, &chunk_total);
      const float T = start * ex;
      const float w = t.alpha * T;
      float G = 0.f;
      if (valid) {
        G = gd * zi + go;
        if (g_weights) G += g_weights[ray * S + i];
        if (rgb) {
          const float* c3 = rgb + (ray * S + i) * 3;
          G += gr * c3[0] + gg * c3[1] + gb * c3[2];
          if (d_rgb) {
            float* o3 = d_rgb + (ray * S + i) * 3;
            o3[0] = w * gr; o3[1] = w * gg; o3[2] = w * gb;
          }
        }
      }
      const float wG = valid ? w * G : 0.f;
      float incl = wG;  // reverse inclusive scan within the chunk
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float dn = __shfl_down_sync(0xffffffffu, incl, o);
        if (lane + o < 32) incl += dn;
      }
      // exclusive reverse scan by a lane shift (never "incl - wG": the difference is divided by a factor
      // that can be as small as 1e-10, which would amplify the cancellation error)
      const float nxt = __shfl_down_sync(0xffffffffu, incl, 1);
      const float after = (lane < 31 ? nxt : 0.f) + suffix;  // sum_{j>i} w_j G_j
      if (valid) {
        const float dalpha = T * G - after / t.factor;
        d_sigma[ray * S + i] = t.active ? dalpha * t.delta_e : 0.f;
      }
      suffix += __shfl_sync(0xffffffffu, incl, 0);
    }
  }
}

// Backward for S <= 256 (NC = chunks of 32, compile time): every input of a ray is loaded ONCE, all loads are issued
// before the first scan, and the per-sample terms of the forward sweep stay in registers for the reverse sweep (the
// generic kernel above reads z / sigma / noise twice and evaluates expf twice).  Same arithmetic, same order.
template <int NC>
__global__ void __launch_bounds__(kCompWarps * 32)
composite_bwd_small_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb, const float* __restrict__ z,
                           const float* __restrict__ rays, const float* __restrict__ noise, float noise_std,
                           int white_back, int64_t n_rays, int S, const float* __restrict__ g_rgb,
                           const float* __restrict__ g_depth, const float* __restrict__ g_opacity,
                           const float* __restrict__ g_disp, const float* __restrict__ g_weights,
                           float* __restrict__ d_sigma, float* __restrict__ d_rgb) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t ray = blockIdx.x * (int64_t)kCompWarps + warp; ray < n_rays;
       ray += (int64_t)gridDim.x * kCompWarps) {
    const float dx = __ldg(rays + ray * 8 + 3), dy = __ldg(rays + ray * 8 + 4), dz = __ldg(rays + ray * 8 + 5);
    ChunkIn in[NC];
    float gw[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      in[c] = load_chunk(sigma + ray * S, z + ray * S, rgb, noise, ray, S, lane, c);
      const int i = c * 32 + lane;
      gw[c] = (g_weights && i < S) ? g_weights[ray * S + i] : 0.f;
    }
    float gr = 0.f, gg = 0.f, gb = 0.f;
    if (rgb && g_rgb) { gr = g_rgb[ray * 3]; gg = g_rgb[ray * 3 + 1]; gb = g_rgb[ray * 3 + 2]; }
    float gd = g_depth ? g_depth[ray] : 0.f;
    float go = g_opacity ? g_opacity[ray] : 0.f;
    const float gdisp = g_disp ? g_disp[ray] : 0.f;
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    // sweep 1
    SampleTerms tm[NC];
    float T[NC];
    float carry = 1.0f, a_d = 0.f, a_w = 0.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      tm[c].alpha = 0.f; tm[c].factor = 1.0f; tm[c].delta_e = 0.f; tm[c].active = false;
      if (valid) tm[c] = sample_terms(in[c].sg, in[c].nz * noise_std, in[c].zi, in[c].zn, i == S - 1, dnorm);
      float chunk_total;
      const float ex = warp_excl_prod(tm[c].factor, lane, &chunk_total);
      T[c] = carry * ex;
      if (valid) {
        const float w = tm[c].alpha * T[c];
        a_w += w;
        a_d += w * in[c].zi;
      }
      carry *= chunk_total;
    }
    a_w = warp_sum(a_w);
    a_d = warp_sum(a_d);
    if (white_back && rgb) go -= (gr + gg + gb);
    if (g_disp) {
      const float q = a_d / a_w;
      if (q > 1e-10f || isnan(q)) {
        const float gq = -gdisp / (q * q);
        gd += gq / a_w;
        go += -gq * a_d / (a_w * a_w);
      }
    }
    // sweep 2 (reverse)
    float suffix = 0.f;
#pragma unroll
    for (int c = NC - 1; c >= 0; --c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      const float w = tm[c].alpha * T[c];
      float G = 0.f;
      if (valid) {
        G = gd * in[c].zi + go;
        if (g_weights) G += gw[c];
        if (rgb) {
          G += gr * in[c].r + gg * in[c].g + gb * in[c].b;
          if (d_rgb) {
            float* o3 = d_rgb + (ray * S + i) * 3;
            o3[0] = w * gr; o3[1] = w * gg; o3[2] = w * gb;
          }
        }
      }
      const float wG = valid ? w * G : 0.f;
      float incl = wG;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float dn = __shfl_down_sync(0xffffffffu, incl, o);
        if (lane + o < 32) incl += dn;
      }
      const float nxt = __shfl_down_sync(0xffffffffu, incl, 1);
      const float after = (lane < 31 ? nxt : 0.f) + suffix;
      if (valid) {
        const float dalpha = T[c] * G - after / tm[c].factor;
        d_sigma[ray * S + i] = tm[c].active ? dalpha * tm[c].delta_e : 0.f;
      }
      suffix += __shfl_sync(0xffffffffu, incl, 0);
    }
  }
}

}  // namespace nrf

using namespace nrf;

static int comp_grid(int64_t n_rays) {
  int64_t b = (n_rays + kCompWarps - 1) / kCompWarps;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

extern "C" int nrf_composite_fwd(const float* sigma, const float* rgb, const float* z, const float* rays,
                                 const float* noise, float noise_std, int white_back, int64_t n_rays,
                                 int n_samples, float* weights_out, float* rgb_out, float* depth_out,
                                 float* opacity_out, float* disp_out, nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_composite_fwd: bad sizes");
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(sigma && z && rays && weights_out && depth_out && opacity_out, "nrf_composite_fwd: null pointer");
  NRF_REQUIRE(!rgb || rgb_out, "nrf_composite_fwd: rgb given but rgb_out is null");
  composite_fwd_kernel<<<comp_grid(n_rays), kCompWarps * 32, 0, as_stream(stream)>>>(
      sigma, rgb, z, rays, noise, noise_std, white_back, n_rays, n_samples, weights_out, rgb_out, depth_out,
      opacity_out, disp_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_composite_bwd(const float* sigma, const float* rgb, const float* z, const float* rays,
                                 const float* noise, float noise_std, int white_back, int64_t n_rays,
                                 int n_samples, const float* g_rgb, const float* g_depth, const float* g_opacity,
                                 const float* g_disp, const float* g_weights, float* d_sigma, float* d_rgb,
                                 nrf_stream_t stream) {
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_composite_bwd: bad sizes");
  NRF_REQUIRE(n_samples <= 1024, "nrf_composite_bwd: %d samples per ray > 1024 unsupported", n_samples);
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(sigma && z && rays && d_sigma, "nrf_composite_bwd: null pointer");
  NRF_REQUIRE(!rgb || d_rgb, "nrf_composite_bwd: rgb given but d_rgb is null");
  const int nc = (n_samples + 31) / 32;
#define NRF_COMP_BWD(NC)                                                                                          \
  composite_bwd_small_kernel<NC><<<comp_grid(n_rays), kCompWarps * 32, 0, as_stream(stream)>>>(                  \
      sigma, rgb, z, rays, noise, noise_std, white_back, n_rays, n_samples, g_rgb, g_depth, g_opacity, g_disp,    \
      g_weights, d_sigma, d_rgb)
  switch (nc) {
    case 1: NRF_COMP_BWD(1); break;
    case 2: NRF_COMP_BWD(2); break;
    case 3: NRF_COMP_BWD(3); break;
    case 4: NRF_COMP_BWD(4); break;
    case 5: NRF_COMP_BWD(5); break;
    case 6: NRF_COMP_BWD(6); break;
    case 7: NRF_COMP_BWD(7); break;
    case 8: NRF_COMP_BWD(8); break;
    default:
      composite_bwd_kernel<<<comp_grid(n_rays), kCompWarps * 32, 0, as_stream(stream)>>>(
          sigma, rgb, z, rays, noise, noise_std, white_back, n_rays, n_samples, g_rgb, g_depth, g_opacity, g_disp,
          g_weights, d_sigma, d_rgb);
  }
#undef NRF_COMP_BWD
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
