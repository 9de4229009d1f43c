// Differentiable shadow test (models/efficient_shadow_mapping.py:19-130, models/camera.py:121-132)
// for a batch of camera rays against one light-view depth map, including the same-camera
// segmentation that the reference does with a Python loop and one host sync per ray
// (models/rendering_shadows.py:385-406).  Everything stays on the device:
//   k0  light map normalisation     get_normed_w on the light image          (:47-62)
//   k1  segment starts              run-length split on eye_pos               (rendering_shadows.py:385-406)
//   k2  project + gather            get_normed_w, get_diff_projections, get_projected_depths (:47-100)
//   k3  shadow value                generate_shadow_map                       (:103-130)
// Backward recomputes k0-k2 and adds the reductions autograd performs for min()/max().
// HBM-bound, tiny: ~100 B per camera ray.
#include <cfloat>
#include <mutex>
#include "common.cuh"

namespace nrf {

constexpr float kShadowEps = 1e-5f;  // efficient_shadow_mapping.py:8

__device__ __forceinline__ int float_key(float f) {  // order-preserving float -> int
  const int b = __float_as_int(f);
  return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

struct ShadowWs {
  float* w_light;   // [HW] normalised light depth
  float* nrm_light; // [HW]
  int* seg_start;   // [n]
  float* diff;      // [n]
  int* gidx;        // [n]
  float* nrm_cam;   // [n]
  float* p2;        // [n]  third component of R*pix
  int* seg_min;     // [n] keys
  int* seg_max;     // [n] keys
  float* s1;        // [n]
  float* s2;        // [n]
  int* cnt_min;     // [n]
  int* cnt_max;     // [n]
  float* g_wlight;  // [HW]
  size_t total;
};

static inline size_t up256(size_t v) { return (v + 255) & ~size_t(255); }

static ShadowWs carve(void* base, int64_t n, int64_t hw) {
  ShadowWs w{};
  char* p = static_cast<char*>(base);
  size_t o = 0;
  auto take = [&](size_t bytes) { char* q = p ? p + o : nullptr; o += up256(bytes); return q; };
  w.w_light = (float*)take(hw * 4); w.nrm_light = (float*)take(hw * 4);
  w.seg_start = (int*)take(n * 4); w.diff = (float*)take(n * 4); w.gidx = (int*)take(n * 4);
  w.nrm_cam = (float*)take(n * 4); w.p2 = (float*)take(n * 4);
  w.seg_min = (int*)take(n * 4); w.seg_max = (int*)take(n * 4);
  w.s1 = (float*)take(n * 4); w.s2 = (float*)take(n * 4);
  w.cnt_min = (int*)take(n * 4); w.cnt_max = (int*)take(n * 4);
  w.g_wlight = (float*)take(hw * 4);
  w.total = o;
  return w;
}

__device__ __forceinline__ float mat_row_dot(const float* __restrict__ M, int r, float a, float b, float c) {
  // sum over the last axis of pix[None,:] * M  (efficient_shadow_mapping.py:55-56)
  return (a * M[r * 3 + 0] + b * M[r * 3 + 1]) + c * M[r * 3 + 2];
}

__global__ void light_norm_kernel(const float* __restrict__ light_pixels, const float* __restrict__ light_depth,
                                  const float* __restrict__ light_M, int64_t hw, float* __restrict__ w_light,
                                  float* __restrict__ nrm_light) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < hw; k += (int64_t)gridDim.x * blockDim.x) {
    if (light_pixels == nullptr) {  // light_depth is already normalised (run_shadow_mapping's contract, :29-30)
      nrm_light[k] = 1.0f;
      w_light[k] = light_depth[k];
      continue;
    }
    const float a = light_pixels[k * 3], b = light_pixels[k * 3 + 1], c = light_pixels[k * 3 + 2];
    const float x = mat_row_dot(light_M, 0, a, b, c), y = mat_row_dot(light_M, 1, a, b, c),
                zc = mat_row_dot(light_M, 2, a, b, c);
    const float nrm = sqrtf(x * x + y * y + zc * zc) + kShadowEps;
    nrm_light[k] = nrm;
    w_light[k] = light_depth[k] / nrm;
  }
}

// Single block: seg_start[i] = index of the first ray of the same-eye_pos run containing i.
__global__ void segment_kernel(const float* __restrict__ eye, int64_t n, int* __restrict__ seg_start,
                               int* __restrict__ seg_min, int* __restrict__ seg_max, float* __restrict__ s1,
                               float* __restrict__ s2, int* __restrict__ cnt_min, int* __restrict__ cnt_max) {
  __shared__ int warp_max[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < n; base += blockDim.x) {
    const int64_t i = base + tid;
    int v = 0;
    if (i < n) {
      bool start = (i == 0);
      if (!start) {
        start = eye[i * 3] != eye[(i - 1) * 3] || eye[i * 3 + 1] != eye[(i - 1) * 3 + 1] ||
                eye[i * 3 + 2] != eye[(i - 1) * 3 + 2];
      }
      v = start ? (int)i : 0;
      if (seg_min) {
        seg_min[i] = float_key(FLT_MAX);
        seg_max[i] = float_key(-FLT_MAX);
        s1[i] = 0.f; s2[i] = 0.f; cnt_min[i] = 0; cnt_max[i] = 0;
      }
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v = max(v, up);
    }
    if (lane == 31) warp_max[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int m = warp_max[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, m, o);
        if (lane >= o) m = max(m, up);
      }
      warp_max[lane] = m;
    }
    __syncthreads();
    int pre = carry_s;
    if (warp > 0) pre = max(pre, warp_max[warp - 1]);
    v = max(v, pre);
    if (i < n) seg_start[i] = v;
    __syncthreads();
    if (tid == blockDim.x - 1) carry_s = v;
    __syncthreads();
  }
}

__global__ void project_kernel(const float* __restrict__ cam_pixels, const float* __restrict__ cam_depth,
                               const float* __restrict__ cam_M, const float* __restrict__ cam_eye,
                               const float* __restrict__ light_Minv, const float* __restrict__ light_eye,
                               const float* __restrict__ w_light, int64_t n, int res_w, int res_h, int track_minmax,
                               const int* __restrict__ seg_start, float* __restrict__ diff, int* __restrict__ gidx,
                               float* __restrict__ nrm_cam, float* __restrict__ p2_out, int* __restrict__ seg_min,
                               int* __restrict__ seg_max) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int s = seg_start[i];
    const float* M = cam_M + (int64_t)s * 9;
    const float a = cam_pixels[i * 3], b = cam_pixels[i * 3 + 1], c = cam_pixels[i * 3 + 2];
    const float cx = mat_row_dot(M, 0, a, b, c), cy = mat_row_dot(M, 1, a, b, c), cz = mat_row_dot(M, 2, a, b, c);
    const float nrm = sqrtf(cx * cx + cy * cy + cz * cz) + kShadowEps;
    const float wc = cam_depth[i] / nrm;  // :58-59
    // R = ML^-1 * MC ; Q = ML^-1 * (OC - OL)   (camera.py:121-132)
    float R[9], Q[3];
    const float ex = cam_eye[(int64_t)s * 3] - light_eye[0], ey = cam_eye[(int64_t)s * 3 + 1] - light_eye[1],
                ez = cam_eye[(int64_t)s * 3 + 2] - light_eye[2];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float l0 = light_Minv[r * 3], l1 = light_Minv[r * 3 + 1], l2 = light_Minv[r * 3 + 2];
#pragma unroll
      for (int q = 0; q < 3; ++q) R[r * 3 + q] = (l0 * M[q] + l1 * M[3 + q]) + l2 * M[6 + q];
      Q[r] = (l0 * ex + l1 * ey) + l2 * ez;
    }
    const float p0 = mat_row_dot(R, 0, a, b, c), p1 = mat_row_dot(R, 1, a, b, c), p2 = mat_row_dot(R, 2, a, b, c);
    const float u = wc * p0 + Q[0], v = wc * p1 + Q[1], wl = wc * p2 + Q[2];  // :76-78
    float ul = u / wl, vl = v / wl;
    ul = fminf(fmaxf(ul, 0.f), (float)res_w - 1.0f);  // :93-96
    vl = fminf(fmaxf(vl, 0.f), (float)res_h - 1.0f);
    const int gi = (int)vl * res_h + (int)ul;          // w_light.view(w,h)[vl, ul]  (:98)
    const float d = wl - w_light[gi];
    diff[i] = d;
    gidx[i] = gi;
    nrm_cam[i] = nrm;
    p2_out[i] = p2;
    if (track_minmax) {
      atomicMin(seg_min + s, float_key(d));
      atomicMax(seg_max + s, float_key(d));
    }
  }
}

struct ShadowCfg {
  int mode, sigmoid;
  float delta, epsilon, new_min, new_max, out_bias;
  float norm_eps = kShadowEps;  // 0 for models/shadow_mapping_utils.py:8-9 (whole-image variant, no epsilon)
};

__device__ __forceinline__ float shadow_value(float d, float mn, float mx, const ShadowCfg& cfg) {
  if (cfg.mode == 1) return fmaxf(d / cfg.delta, cfg.epsilon);  // :119-121
  float y = (d - mn) / (mx - mn + cfg.norm_eps) * (cfg.new_max - cfg.new_min) + cfg.new_min;  // :10-11
  if (cfg.sigmoid) y = 1.0f / (1.0f + expf(-y));
  return y;
}

__global__ void shade_kernel(const float* __restrict__ diff, const int* __restrict__ seg_start,
                             const int* __restrict__ seg_min, const int* __restrict__ seg_max, int64_t n,
                             ShadowCfg cfg, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float mn = 0.f, mx = 0.f;
    if (cfg.mode == 2) { const int s = seg_start[i]; mn = key_float(seg_min[s]); mx = key_float(seg_max[s]); }
    float v = shadow_value(diff[i], mn, mx, cfg);
    v = fminf(fmaxf(v, 0.f), 1.f) + cfg.out_bias;  // :129 + rendering_shadows.py:459
    out[i * 3] = v; out[i * 3 + 1] = v; out[i * 3 + 2] = v;
  }
}

// gradient w.r.t. the (pre-clip) normalised value y for method 2, or w.r.t. diff for method 1
__device__ __forceinline__ float upstream(const float* __restrict__ g_out, int64_t i, float d, float mn, float mx,
                                          const ShadowCfg& cfg) {
  const float g = (g_out[i * 3] + g_out[i * 3 + 1]) + g_out[i * 3 + 2];
  const float v = shadow_value(d, mn, mx, cfg);
  if (!(v >= 0.f && v <= 1.f)) return 0.f;  // clamp backward
  if (cfg.mode == 1) {
    const float a = d / cfg.delta;
    const float pass = a > cfg.epsilon ? 1.f : (a == cfg.epsilon ? 0.5f : 0.f);  // torch.max(a, b) backward
    return g * pass / cfg.delta;
  }
  float gy = g;
  if (cfg.sigmoid) gy *= v * (1.f - v);
  return gy * (cfg.new_max - cfg.new_min);
}

__global__ void bwd_reduce_kernel(const float* __restrict__ g_out, const float* __restrict__ diff,
                                  const int* __restrict__ seg_start, const int* __restrict__ seg_min,
                                  const int* __restrict__ seg_max, int64_t n, ShadowCfg cfg, float* __restrict__ s1,
                                  float* __restrict__ s2, int* __restrict__ cnt_min, int* __restrict__ cnt_max) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int s = seg_start[i];
    const float mn = key_float(seg_min[s]), mx = key_float(seg_max[s]);
    const float d = diff[i];
    const float gy = upstream(g_out, i, d, mn, mx, cfg);
    atomicAdd(s1 + s, gy);
    atomicAdd(s2 + s, gy * (d - mn));
    if (d == mn) atomicAdd(cnt_min + s, 1);
    if (d == mx) atomicAdd(cnt_max + s, 1);
  }
}

__global__ void bwd_apply_kernel(const float* __restrict__ g_out, const float* __restrict__ diff,
                                 const int* __restrict__ gidx, const float* __restrict__ nrm_cam,
                                 const float* __restrict__ p2, const int* __restrict__ seg_start,
                                 const int* __restrict__ seg_min, const int* __restrict__ seg_max,
                                 const float* __restrict__ s1, const float* __restrict__ s2,
                                 const int* __restrict__ cnt_min, const int* __restrict__ cnt_max, int64_t n,
                                 ShadowCfg cfg, float* __restrict__ g_cam_depth, float* __restrict__ g_wlight) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = diff[i];
    float gd;
    if (cfg.mode == 1) {
      gd = upstream(g_out, i, d, 0.f, 0.f, cfg);
    } else {
      const int s = seg_start[i];
      const float mn = key_float(seg_min[s]), mx = key_float(seg_max[s]);
      const float D = mx - mn + kShadowEps;
      gd = upstream(g_out, i, d, mn, mx, cfg) / D;
      const float g_mn = -s1[s] / D + s2[s] / (D * D);  // through tensor.min()
      const float g_mx = -s2[s] / (D * D);              // through tensor.max()
      if (d == mn) gd += g_mn / (float)cnt_min[s];
      if (d == mx) gd += g_mx / (float)cnt_max[s];
    }
    g_cam_depth[i] = gd * p2[i] / nrm_cam[i];  // wl = (depth/nrm)*p2 + Q2
    if (g_wlight) atomicAdd(g_wlight + gidx[i], -gd);
  }
}

__global__ void bwd_light_kernel(const float* __restrict__ g_wlight, const float* __restrict__ nrm_light, int64_t hw,
                                 float* __restrict__ g_light_depth) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < hw; k += (int64_t)gridDim.x * blockDim.x)
    g_light_depth[k] += g_wlight[k] / nrm_light[k];
}

static inline int sgrid(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < 1 ? 1 : (b > kNumSMs * 8 ? kNumSMs * 8 : b));
}

static int shadow_common(const float* cam_pixels, const float* cam_depth, const float* cam_M, const float* cam_eye,
                         const float* light_Minv, const float* light_M, const float* light_eye,
                         const float* light_pixels, const float* light_depth, int64_t n, int res_w, int res_h,
                         int mode, const ShadowWs& w, cudaStream_t st) {
  const int64_t hw = (int64_t)res_w * res_h;
  light_norm_kernel<<<sgrid(hw), 256, 0, st>>>(light_pixels, light_depth, light_M, hw, w.w_light, w.nrm_light);
  NRF_LAUNCH_CHECK();
  segment_kernel<<<1, 1024, 0, st>>>(cam_eye, n, w.seg_start, w.seg_min, w.seg_max, w.s1, w.s2, w.cnt_min, w.cnt_max);
  NRF_LAUNCH_CHECK();
  project_kernel<<<sgrid(n), 256, 0, st>>>(cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_eye, w.w_light, n,
                                           res_w, res_h, mode == 2, w.seg_start, w.diff, w.gidx, w.nrm_cam, w.p2,
                                           w.seg_min, w.seg_max);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

static int shadow_check(const char* who, const void* a, const void* b, const void* c, const void* d, const void* e,
                        const void* f, const void* g, const void* h, const void* i, int64_t n, int res_w, int res_h,
                        int mode, void* ws, size_t ws_bytes) {
  NRF_REQUIRE(n >= 0 && res_w >= 1 && res_h >= 1, "%s: bad sizes", who);
  NRF_REQUIRE(res_w == res_h, "%s: only square light maps are supported, like the reference "
              "(efficient_shadow_mapping.py:3-4), got %dx%d", who, res_w, res_h);
  NRF_REQUIRE(mode == 1 || mode == 2, "%s: shadow mode %d not found", who, mode);
  NRF_REQUIRE(n < (int64_t)1 << 31, "%s: too many rays", who);
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(a && b && c && d && e && g && i && ws, "%s: null pointer", who);
  NRF_REQUIRE((f && h) || (!f && !h), "%s: light_M and light_pixels must be both given or both null", who);
  const size_t need = carve(nullptr, n, (int64_t)res_w * res_h).total;
  if (ws_bytes < need) {
    set_error("%s: workspace %zu < %zu bytes", who, ws_bytes, need);
    return NRF_ERR_WORKSPACE;
  }
  return NRF_OK;
}

}  // namespace nrf

using namespace nrf;

extern "C" size_t nrf_shadow_workspace_bytes(int64_t n, int64_t n_light) {
  if (n < 0 || n_light < 0) return 0;
  return carve(nullptr, n, n_light).total + 256;
}

extern "C" int nrf_shadow_fwd(const float* cam_pixels, const float* cam_depth, const float* cam_M,
                              const float* cam_eye, const float* light_Minv, const float* light_M,
                              const float* light_eye, const float* light_pixels, const float* light_depth, int64_t n,
                              int res_w, int res_h, int mode, float delta, float epsilon, float new_min, float new_max,
                              int sigmoid, float out_bias, float* out, void* workspace, size_t workspace_bytes,
                              nrf_stream_t stream) {
  int rc = shadow_check("nrf_shadow_fwd", cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye,
                        light_pixels, light_depth, n, res_w, res_h, mode, workspace, workspace_bytes);
  if (rc || n == 0) return rc;
  NRF_REQUIRE(out != nullptr, "nrf_shadow_fwd: null output");
  cudaStream_t st = as_stream(stream);
  const ShadowWs w = carve(workspace, n, (int64_t)res_w * res_h);
  rc = shadow_common(cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye, light_pixels, light_depth,
                     n, res_w, res_h, mode, w, st);
  if (rc) return rc;
  const ShadowCfg cfg{mode, sigmoid, delta, epsilon, new_min, new_max, out_bias};
  shade_kernel<<<sgrid(n), 256, 0, st>>>(w.diff, w.seg_start, w.seg_min, w.seg_max, n, cfg, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// nrf_shadow_fwd_static: the same four launches with the scratch in a file-scope device array instead of a caller
// workspace, for callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// Sized for one config-5 batch: NRF_SHADOW_STATIC_MAX_RAYS camera rays against a light map of
// NRF_SHADOW_STATIC_MAX_TEXELS texels.  Every GPU that loads the library holds the array until exit.
// ---------------------------------------------------------------------------------------------------------------------
namespace nrf {
constexpr size_t cup256(size_t v) { return (v + 255) & ~size_t(255); }
// mirrors carve(): 3 light-map arrays + 11 per-ray arrays, each rounded up to 256 bytes
constexpr size_t kShadowStaticBytes = 3 * cup256((size_t)NRF_SHADOW_STATIC_MAX_TEXELS * 4) +
                                      11 * cup256((size_t)NRF_SHADOW_STATIC_MAX_RAYS * 4);
static __device__ __align__(256) unsigned char g_shadow_static_scratch[kShadowStaticBytes];

// In host code `g_shadow_static_scratch` is only a stand-in for the symbol: its device address is looked up once per
// device, with that device current, and cached.
static int shadow_static_ptr(void** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static void* ptrs[kMaxDev] = {};
  int dev = -1;
  NRF_CUDA(cudaGetDevice(&dev));
  NRF_REQUIRE(dev >= 0 && dev < kMaxDev, "nrf_shadow_fwd_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    NRF_CUDA(cudaGetSymbolAddress(&p, g_shadow_static_scratch));
    NRF_REQUIRE(p != nullptr, "nrf_shadow_fwd_static: no device address for the static scratch");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return NRF_OK;
}
}  // namespace nrf

extern "C" int nrf_shadow_fwd_static(const float* cam_pixels, const float* cam_depth, const float* cam_M,
                                     const float* cam_eye, const float* light_Minv, const float* light_M,
                                     const float* light_eye, const float* light_pixels, const float* light_depth,
                                     int64_t n, int res_w, int res_h, int mode, float delta, float epsilon,
                                     float new_min, float new_max, int sigmoid, float out_bias, float* out,
                                     nrf_stream_t stream) {
  if (n > (int64_t)NRF_SHADOW_STATIC_MAX_RAYS || (int64_t)res_w * res_h > (int64_t)NRF_SHADOW_STATIC_MAX_TEXELS) {
    set_error("nrf_shadow_fwd_static: %lld rays / %lld light texels exceed the static scratch (at most %d rays, %d texels); "
              "use nrf_shadow_fwd with a caller workspace", (long long)n, (long long)res_w * res_h,
              NRF_SHADOW_STATIC_MAX_RAYS, NRF_SHADOW_STATIC_MAX_TEXELS);
    return NRF_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  int rc = shadow_static_ptr(&scratch);
  if (rc) return rc;
  rc = shadow_check("nrf_shadow_fwd_static", cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye,
                    light_pixels, light_depth, n, res_w, res_h, mode, scratch, kShadowStaticBytes);
  if (rc || n == 0) return rc;
  NRF_REQUIRE(out != nullptr, "nrf_shadow_fwd_static: null output");
  cudaStream_t st = as_stream(stream);
  // nothing in the scratch has to start at zero: light_norm / segment / project overwrite every word the shade
  // kernel reads (segment_kernel re-initialises the per-segment min / max keys), so the previous call's contents are dead
  const ShadowWs w = carve(scratch, n, (int64_t)res_w * res_h);
  rc = shadow_common(cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye, light_pixels, light_depth,
                     n, res_w, res_h, mode, w, st);
  if (rc) return rc;
  const ShadowCfg cfg{mode, sigmoid, delta, epsilon, new_min, new_max, out_bias};
  shade_kernel<<<sgrid(n), 256, 0, st>>>(w.diff, w.seg_start, w.seg_min, w.seg_max, n, cfg, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_shadow_bwd(const float* cam_pixels, const float* cam_depth, const float* cam_M,
                              const float* cam_eye, const float* light_Minv, const float* light_M,
                              const float* light_eye, const float* light_pixels, const float* light_depth, int64_t n,
                              int res_w, int res_h, int mode, float delta, float epsilon, float new_min, float new_max,
                              int sigmoid, const float* g_out, float* g_cam_depth, float* g_light_depth,
                              void* workspace, size_t workspace_bytes, nrf_stream_t stream) {
  int rc = shadow_check("nrf_shadow_bwd", cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye,
                        light_pixels, light_depth, n, res_w, res_h, mode, workspace, workspace_bytes);
  if (rc || n == 0) return rc;
  NRF_REQUIRE(g_out && g_cam_depth, "nrf_shadow_bwd: null pointer");
  cudaStream_t st = as_stream(stream);
  const int64_t hw = (int64_t)res_w * res_h;
  const ShadowWs w = carve(workspace, n, hw);
  rc = shadow_common(cam_pixels, cam_depth, cam_M, cam_eye, light_Minv, light_M, light_eye, light_pixels, light_depth,
                     n, res_w, res_h, mode, w, st);
  if (rc) return rc;
  const ShadowCfg cfg{mode, sigmoid, delta, epsilon, new_min, new_max, 0.f};
  if (g_light_depth) NRF_CUDA(cudaMemsetAsync(w.g_wlight, 0, hw * sizeof(float), st));
  if (mode == 2) {
    bwd_reduce_kernel<<<sgrid(n), 256, 0, st>>>(g_out, w.diff, w.seg_start, w.seg_min, w.seg_max, n, cfg, w.s1, w.s2,
                                                w.cnt_min, w.cnt_max);
    NRF_LAUNCH_CHECK();
  }
  bwd_apply_kernel<<<sgrid(n), 256, 0, st>>>(g_out, w.diff, w.gidx, w.nrm_cam, w.p2, w.seg_start, w.seg_min, w.seg_max,
                                             w.s1, w.s2, w.cnt_min, w.cnt_max, n, cfg, g_cam_depth,
                                             g_light_depth ? w.g_wlight : nullptr);
  NRF_LAUNCH_CHECK();
  if (g_light_depth) {
    bwd_light_kernel<<<sgrid(hw), 256, 0, st>>>(w.g_wlight, w.nrm_light, hw, g_light_depth);
    NRF_LAUNCH_CHECK();
  }
  return NRF_OK;
}

// a14 stand-alone get_normed_w (efficient_shadow_mapping.py:47-62): out = [pass[:, :3], pd[:,3]/(|M pd[:, :3]|+eps)].
// With pass == pd this is the forward; with pass == grad_out and pd[:,3] replaced by grad_out[:,3]
// it is the backward w.r.t. the depth column (the map is linear in depth).
namespace nrf {
__global__ void normed_w_kernel(const float* __restrict__ pd, const float* __restrict__ val,
                                const float* __restrict__ pass, const float* __restrict__ M, int64_t n,
                                float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float a = pd[i * 4], b = pd[i * 4 + 1], c = pd[i * 4 + 2];
    const float x = mat_row_dot(M, 0, a, b, c), y = mat_row_dot(M, 1, a, b, c), zc = mat_row_dot(M, 2, a, b, c);
    const float nrm = sqrtf(x * x + y * y + zc * zc) + kShadowEps;
    out[i * 4] = pass[i * 4]; out[i * 4 + 1] = pass[i * 4 + 1]; out[i * 4 + 2] = pass[i * 4 + 2];
    out[i * 4 + 3] = val[i * 4 + 3] / nrm;
  }
}
}  // namespace nrf

extern "C" int nrf_normed_w(const float* pixel_depth, const float* value, const float* passthrough, const float* M,
                            int64_t n, float* out, nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0, "nrf_normed_w: bad size");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(pixel_depth && value && passthrough && M && out, "nrf_normed_w: null pointer");
  normed_w_kernel<<<sgrid(n), 256, 0, as_stream(stream)>>>(pixel_depth, value, passthrough, M, n, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}


// ---- §8f rank 2: the pieces the light-sampler trainer calls one by one (train_light_sampler.py:141-196) -----------------
// a24 get_K / get_projections / get_diff_projections (rendering_shadows.py:484-546, efficient_shadow_mapping.py:13-17,
//     :65-83): K = [ul, vl, wl] per camera ray, cameras per ray (segmented by eye_pos like the reference's loop) or one
//     camera for the whole batch; `normalised` = the depth column already is w (get_diff_projections).
// a25 generate_shadow_map / normalize_min_max (:10-11, :103-130) on explicit (wl, w_light_bounded) vectors: whole-vector
//     min/max, warp-reduced before the atomics.
// a26 get_projected_depths (:85-100): clamp, truncate, gather from the light map; backward scatters.
namespace nrf {

struct ProjCam {
  float R[9], Q[3], nrm;
};

__device__ __forceinline__ ProjCam proj_setup(const float* __restrict__ M, const float* __restrict__ eye,
                                              const float* __restrict__ light_Minv, const float* __restrict__ light_eye,
                                              float a, float b, float c, int normalised) {
  ProjCam pc;
  const float cx = mat_row_dot(M, 0, a, b, c), cy = mat_row_dot(M, 1, a, b, c), cz = mat_row_dot(M, 2, a, b, c);
  pc.nrm = normalised == 1 ? 1.0f : sqrtf(cx * cx + cy * cy + cz * cz) + (normalised == 2 ? 0.0f : kShadowEps);
  const float ex = eye[0] - light_eye[0], ey = eye[1] - light_eye[1], ez = eye[2] - light_eye[2];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float l0 = light_Minv[r * 3], l1 = light_Minv[r * 3 + 1], l2 = light_Minv[r * 3 + 2];
#pragma unroll
    for (int q = 0; q < 3; ++q) pc.R[r * 3 + q] = (l0 * M[q] + l1 * M[3 + q]) + l2 * M[6 + q];
    pc.Q[r] = (l0 * ex + l1 * ey) + l2 * ez;
  }
  return pc;
}

// g_K == nullptr: forward, out = K (n,3).  Otherwise backward: out = g_depth (n).
__global__ void project_to_light_kernel(const float* __restrict__ cam_pixels, const float* __restrict__ cam_depth,
                                        const float* __restrict__ cam_M, const float* __restrict__ cam_eye,
                                        const int* __restrict__ seg_start, const float* __restrict__ light_Minv,
                                        const float* __restrict__ light_eye, int64_t n, int normalised,
                                        const float* __restrict__ g_K, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = seg_start ? seg_start[i] : 0;
    const float a = cam_pixels[i * 3], b = cam_pixels[i * 3 + 1], c = cam_pixels[i * 3 + 2];
    const ProjCam pc = proj_setup(cam_M + s * 9, cam_eye + s * 3, light_Minv, light_eye, a, b, c, normalised);
    const float wc = cam_depth[i] / pc.nrm;
    const float p0 = mat_row_dot(pc.R, 0, a, b, c), p1 = mat_row_dot(pc.R, 1, a, b, c), p2 = mat_row_dot(pc.R, 2, a, b, c);
    const float u = wc * p0 + pc.Q[0], v = wc * p1 + pc.Q[1], wl = wc * p2 + pc.Q[2];
    if (!g_K) {
      out[i * 3] = u / wl; out[i * 3 + 1] = v / wl; out[i * 3 + 2] = wl;
    } else {
      // d(u/wl)/dwc = p0/wl - u p2/wl^2, same for v; dwl/dwc = p2; dwc/ddepth = 1/nrm
      const float inv = 1.0f / wl;
      const float g_wc = g_K[i * 3] * (p0 - u * inv * p2) * inv + g_K[i * 3 + 1] * (p1 - v * inv * p2) * inv +
                         g_K[i * 3 + 2] * p2;
      out[i] = g_wc / pc.nrm;
    }
  }
}

// ---- a25: [0] min key, [1] max key, then float s1, s2, int cnt_min, cnt_max
struct MapWs {
  int* keys;
  float* sums;
  int* cnt;
};

__global__ void map_init_kernel(MapWs w) {
  w.keys[0] = float_key(FLT_MAX); w.keys[1] = float_key(-FLT_MAX);
  w.sums[0] = 0.f; w.sums[1] = 0.f; w.cnt[0] = 0; w.cnt[1] = 0;
}

__device__ __forceinline__ float map_diff(const float* __restrict__ wl, const float* __restrict__ wlb, int64_t i) {
  return wlb ? wl[i] - wlb[i] : wl[i];
}

__global__ void map_minmax_kernel(const float* __restrict__ wl, const float* __restrict__ wlb, int64_t n, MapWs w) {
  int mn = float_key(FLT_MAX), mx = float_key(-FLT_MAX);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = float_key(map_diff(wl, wlb, i));
    mn = min(mn, k); mx = max(mx, k);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(w.keys, mn); atomicMax(w.keys + 1, mx); }
}

__global__ void map_shade_kernel(const float* __restrict__ wl, const float* __restrict__ wlb, int64_t n, MapWs w,
                                 ShadowCfg cfg, int stack_clip, float* __restrict__ out) {
  float mn = 0.f, mx = 0.f;
  if (cfg.mode == 2) { mn = key_float(w.keys[0]); mx = key_float(w.keys[1]); }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = shadow_value(map_diff(wl, wlb, i), mn, mx, cfg);
    if (stack_clip) {
      v = fminf(fmaxf(v, 0.f), 1.f);
      out[i * 3] = v; out[i * 3 + 1] = v; out[i * 3 + 2] = v;
    } else {
      out[i] = v;
    }
  }
}

__device__ __forceinline__ float map_upstream(const float* __restrict__ g_out, int64_t i, float d, float mn, float mx,
                                              const ShadowCfg& cfg, int stack_clip) {
  if (stack_clip) return upstream(g_out, i, d, mn, mx, cfg);
  float gy = g_out[i];                                            // normalize_min_max alone: no clip, no stack
  if (cfg.mode == 1) {
    const float a = d / cfg.delta;
    return gy * (a > cfg.epsilon ? 1.f : (a == cfg.epsilon ? 0.5f : 0.f)) / cfg.delta;
  }
  if (cfg.sigmoid) { const float v = shadow_value(d, mn, mx, cfg); gy *= v * (1.f - v); }
  return gy * (cfg.new_max - cfg.new_min);
}

__global__ void map_bwd_reduce_kernel(const float* __restrict__ g_out, const float* __restrict__ wl,
                                      const float* __restrict__ wlb, int64_t n, MapWs w, ShadowCfg cfg, int stack_clip) {
  const float mn = key_float(w.keys[0]), mx = key_float(w.keys[1]);
  float s1 = 0.f, s2 = 0.f;
  int c0 = 0, c1 = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = map_diff(wl, wlb, i);
    const float gy = map_upstream(g_out, i, d, mn, mx, cfg, stack_clip);
    s1 += gy; s2 += gy * (d - mn);
    c0 += d == mn; c1 += d == mx;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    c0 += __shfl_xor_sync(0xffffffffu, c0, o); c1 += __shfl_xor_sync(0xffffffffu, c1, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(w.sums, s1); atomicAdd(w.sums + 1, s2);
    if (c0) atomicAdd(w.cnt, c0);
    if (c1) atomicAdd(w.cnt + 1, c1);
  }
}

__global__ void map_bwd_apply_kernel(const float* __restrict__ g_out, const float* __restrict__ wl,
                                     const float* __restrict__ wlb, int64_t n, MapWs w, ShadowCfg cfg, int stack_clip,
                                     float* __restrict__ g_diff) {
  float mn = 0.f, mx = 0.f, D = 1.f, g_mn = 0.f, g_mx = 0.f;
  if (cfg.mode == 2) {
    mn = key_float(w.keys[0]); mx = key_float(w.keys[1]);
    D = mx - mn + cfg.norm_eps;
    g_mn = (-w.sums[0] / D + w.sums[1] / (D * D)) / (float)max(w.cnt[0], 1);   // through tensor.min()
    g_mx = (-w.sums[1] / (D * D)) / (float)max(w.cnt[1], 1);                    // through tensor.max()
  }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = map_diff(wl, wlb, i);
    float gd = map_upstream(g_out, i, d, mn, mx, cfg, stack_clip);
    if (cfg.mode == 2) {
      gd /= D;
      if (d == mn) gd += g_mn;
      if (d == mx) gd += g_mx;
    }
    g_diff[i] = gd;
  }
}

// ---- a26
__global__ void light_gather_kernel(const float* __restrict__ K, const float* __restrict__ w_light, int64_t n, int res_w,
                                    int res_h, float* __restrict__ wlb, int* __restrict__ gidx) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float ul = fminf(fmaxf(K[i * 3], 0.f), (float)res_w - 1.0f);
    const float vl = fminf(fmaxf(K[i * 3 + 1], 0.f), (float)res_h - 1.0f);
    const int gi = (int)vl * res_h + (int)ul;          // w_light.view(w,h)[vl, ul]  (:98)
    gidx[i] = gi;
    wlb[i] = w_light[gi];
  }
}

__global__ void light_scatter_kernel(const float* __restrict__ g_wlb, const int* __restrict__ gidx, int64_t n,
                                     float* __restrict__ g_wlight) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(g_wlight + gidx[i], g_wlb[i]);
}

static MapWs map_carve(void* ws) {
  MapWs w;
  w.keys = static_cast<int*>(ws);
  w.sums = reinterpret_cast<float*>(w.keys + 2);
  w.cnt = w.keys + 4;
  return w;
}
}  // namespace nrf

extern "C" int nrf_project_to_light(const float* cam_pixels, const float* cam_depth, const float* cam_M,
                                    const float* cam_eye, int64_t n_cam, const float* light_Minv, const float* light_eye,
                                    int64_t n, int normalised, const float* g_K, float* out, int* seg_ws,
                                    nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0 && n < (int64_t)1 << 31, "nrf_project_to_light: bad size");
  NRF_REQUIRE(n_cam == 1 || n_cam == n, "nrf_project_to_light: %lld cameras for %lld rays (1 or one per ray)",
              (long long)n_cam, (long long)n);
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(cam_pixels && cam_depth && cam_M && cam_eye && light_Minv && light_eye && out,
              "nrf_project_to_light: null pointer");
  cudaStream_t st = as_stream(stream);
  const int* seg = nullptr;
  if (n_cam != 1) {
    NRF_REQUIRE(seg_ws != nullptr, "nrf_project_to_light: per-ray cameras need a workspace of n ints");
    segment_kernel<<<1, 1024, 0, st>>>(cam_eye, n, seg_ws, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    NRF_LAUNCH_CHECK();
    seg = seg_ws;
  }
  project_to_light_kernel<<<sgrid(n), 256, 0, st>>>(cam_pixels, cam_depth, cam_M, cam_eye, seg, light_Minv, light_eye, n,
                                                    normalised, g_K, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" size_t nrf_shadow_map_workspace_bytes(void) { return 32; }

static int map_check(const char* who, const float* wl, int64_t n, int mode, void* ws) {
  NRF_REQUIRE(n >= 0, "%s: bad size", who);
  NRF_REQUIRE(mode == 1 || mode == 2, "%s: shadow mode %d not found", who, mode);
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(wl && ws, "%s: null pointer", who);
  return NRF_OK;
}

extern "C" int nrf_shadow_map_fwd(const float* wl, const float* w_light_bounded, int64_t n, int mode, float delta,
                                  float epsilon, float new_min, float new_max, int sigmoid, int flags, float* out,
                                  void* workspace, nrf_stream_t stream) {
  int rc = map_check("nrf_shadow_map_fwd", wl, n, mode, workspace);
  if (rc || n == 0) return rc;
  NRF_REQUIRE(out != nullptr, "nrf_shadow_map_fwd: null output");
  cudaStream_t st = as_stream(stream);
  const MapWs w = map_carve(workspace);
  ShadowCfg cfg{mode, sigmoid, delta, epsilon, new_min, new_max, 0.f};
  if (flags & 2) cfg.norm_eps = 0.f;
  const int stack_clip = flags & 1;
  if (mode == 2) {
    map_init_kernel<<<1, 1, 0, st>>>(w);
    NRF_LAUNCH_CHECK();
    map_minmax_kernel<<<sgrid(n), 256, 0, st>>>(wl, w_light_bounded, n, w);
    NRF_LAUNCH_CHECK();
  }
  map_shade_kernel<<<sgrid(n), 256, 0, st>>>(wl, w_light_bounded, n, w, cfg, stack_clip, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_shadow_map_bwd(const float* wl, const float* w_light_bounded, int64_t n, int mode, float delta,
                                  float epsilon, float new_min, float new_max, int sigmoid, int flags,
                                  const float* g_out, float* g_diff, void* workspace, nrf_stream_t stream) {
  int rc = map_check("nrf_shadow_map_bwd", wl, n, mode, workspace);
  if (rc || n == 0) return rc;
  NRF_REQUIRE(g_out && g_diff, "nrf_shadow_map_bwd: null pointer");
  cudaStream_t st = as_stream(stream);
  const MapWs w = map_carve(workspace);
  ShadowCfg cfg{mode, sigmoid, delta, epsilon, new_min, new_max, 0.f};
  if (flags & 2) cfg.norm_eps = 0.f;
  const int stack_clip = flags & 1;
  if (mode == 2) {
    map_init_kernel<<<1, 1, 0, st>>>(w);
    NRF_LAUNCH_CHECK();
    map_minmax_kernel<<<sgrid(n), 256, 0, st>>>(wl, w_light_bounded, n, w);
    NRF_LAUNCH_CHECK();
    map_bwd_reduce_kernel<<<sgrid(n), 256, 0, st>>>(g_out, wl, w_light_bounded, n, w, cfg, stack_clip);
    NRF_LAUNCH_CHECK();
  }
  map_bwd_apply_kernel<<<sgrid(n), 256, 0, st>>>(g_out, wl, w_light_bounded, n, w, cfg, stack_clip, g_diff);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_light_gather(const float* K, const float* w_light, int64_t n, int res_w, int res_h,
                                float* w_light_bounded, int* gidx, nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0 && res_w >= 1 && res_h >= 1, "nrf_light_gather: bad sizes");
  NRF_REQUIRE(res_w == res_h, "nrf_light_gather: only square light maps are supported, like the reference "
              "(efficient_shadow_mapping.py:3-4), got %dx%d", res_w, res_h);
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(K && w_light && w_light_bounded && gidx, "nrf_light_gather: null pointer");
  light_gather_kernel<<<sgrid(n), 256, 0, as_stream(stream)>>>(K, w_light, n, res_w, res_h, w_light_bounded, gidx);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_light_scatter(const float* g_wlb, const int* gidx, int64_t n, int64_t n_light, float* g_wlight,
                                 nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0 && n_light >= 0, "nrf_light_scatter: bad sizes");
  if (n_light == 0) return NRF_OK;
  NRF_REQUIRE(g_wlight != nullptr, "nrf_light_scatter: null output");
  NRF_CUDA(cudaMemsetAsync(g_wlight, 0, (size_t)n_light * sizeof(float), as_stream(stream)));
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(g_wlb && gidx, "nrf_light_scatter: null pointer");
  light_scatter_kernel<<<sgrid(n), 256, 0, as_stream(stream)>>>(g_wlb, gidx, n, g_wlight);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}


// a27 whole-image range -> w (models/shadow_mapping_utils.py:39-66): w = range / |M [i+.5, j+.5, 1]^T| for explicit pixel
// rows, no epsilon unless asked for.  Linear in `value`, so the backward is the same launch with value = grad_out.
namespace nrf {
__global__ void w_mapping_kernel(const float* __restrict__ pixels, const float* __restrict__ value,
                                 const float* __restrict__ M, int64_t n, float eps, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float a = pixels[i * 3], b = pixels[i * 3 + 1], c = pixels[i * 3 + 2];
    const float x = mat_row_dot(M, 0, a, b, c), y = mat_row_dot(M, 1, a, b, c), zc = mat_row_dot(M, 2, a, b, c);
    out[i] = value[i] / (sqrtf(x * x + y * y + zc * zc) + eps);
  }
}
}  // namespace nrf

extern "C" int nrf_w_mapping(const float* pixels, const float* value, const float* M, int64_t n, float eps, float* out,
                             nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0, "nrf_w_mapping: bad size");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(pixels && value && M && out, "nrf_w_mapping: null pointer");
  nrf::w_mapping_kernel<<<nrf::sgrid(n), 256, 0, nrf::as_stream(stream)>>>(pixels, value, M, n, eps, out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
