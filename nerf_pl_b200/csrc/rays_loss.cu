// The steps either side of render_rays: camera-ray generation on the GPU and the fused MSE loss.
//
//   a21 nrf_gen_rays     datasets/ray_utils.py:5-93 (get_ray_directions + get_rays + get_ndc_rays) and the row
//                        assembly of the datasets (datasets/blender.py:80-84: [o, d, near, far]) in one pass:
//                        no (H,W,3) direction grid, no rays_o/rays_d temporaries, nothing crosses PCIe but 12 floats.
//   a22 nrf_mse_pair     losses.py:4-14 (MSELoss / SMMSELoss: mean((a-t)^2) + mean((b-t)^2)) and metrics.py:4-10
//                        (mse with an optional mask) as ONE deterministic reduction launch;
//       nrf_mse_pair_bwd the gradients 2 (x - t) / n * upstream for both predictions in one launch.
// Both are HBM-bound streaming kernels: 32 B written per ray; 12 (24) B read per element of the loss.
#include <cmath>
#include "common.cuh"

namespace nrf {

struct C2W { float m[12]; };   // row-major 3x4, by value in the kernel parameters

__global__ void gen_rays_kernel(const __grid_constant__ C2W c2w, int H, int W, float focal, const int64_t* __restrict__ pix,
                                int64_t n, float near, float far, int ndc, float* __restrict__ rays) {
  const float half_w = (float)W / 2.0f, half_h = (float)H / 2.0f;   // W/2, H/2 (python true division)
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < n; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = pix ? __ldg(pix + idx) : idx;
    const float i = (float)(p % W), j = (float)(p / W);              // grid x (column), y (row): ray_utils.py:17-18
    // ray_utils.py:21-22 (no +0.5 pixel centre)
    const float dx = __fdiv_rn(__fsub_rn(i, half_w), focal);
    const float dy = -__fdiv_rn(__fsub_rn(j, half_h), focal);
    const float dz = -1.0f;
    // ray_utils.py:43: directions @ c2w[:, :3].T, accumulated left to right
    float d[3], o[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      d[k] = __fadd_rn(__fadd_rn(__fmul_rn(dx, c2w.m[k * 4 + 0]), __fmul_rn(dy, c2w.m[k * 4 + 1])), __fmul_rn(dz, c2w.m[k * 4 + 2]));
      o[k] = c2w.m[k * 4 + 3];                                      // ray_utils.py:46
    }
    const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1])), __fmul_rn(d[2], d[2])));
#pragma unroll
    for (int k = 0; k < 3; ++k) d[k] = __fdiv_rn(d[k], nrm);        // ray_utils.py:44
    if (ndc) {                                                      // ray_utils.py:75-91 with the scalar near plane
      const float t = __fdiv_rn(-__fadd_rn(near, o[2]), d[2]);
#pragma unroll
      for (int k = 0; k < 3; ++k) o[k] = __fadd_rn(o[k], __fmul_rn(t, d[k]));
      const float ox_oz = __fdiv_rn(o[0], o[2]), oy_oz = __fdiv_rn(o[1], o[2]);
      const float sx = __fdiv_rn(-1.0f, __fdiv_rn((float)W, __fmul_rn(2.0f, focal)));
      const float sy = __fdiv_rn(-1.0f, __fdiv_rn((float)H, __fmul_rn(2.0f, focal)));
      const float o0 = __fmul_rn(sx, ox_oz), o1 = __fmul_rn(sy, oy_oz);
      const float o2 = __fadd_rn(1.0f, __fdiv_rn(__fmul_rn(2.0f, near), o[2]));
      const float d0 = __fmul_rn(sx, __fsub_rn(__fdiv_rn(d[0], d[2]), ox_oz));
      const float d1 = __fmul_rn(sy, __fsub_rn(__fdiv_rn(d[1], d[2]), oy_oz));
      const float d2 = __fsub_rn(1.0f, o2);
      o[0] = o0; o[1] = o1; o[2] = o2; d[0] = d0; d[1] = d1; d[2] = d2;
    }
    float4* out = reinterpret_cast<float4*>(rays + idx * 8);
    out[0] = make_float4(o[0], o[1], o[2], d[0]);
    out[1] = make_float4(d[1], d[2], ndc ? 0.0f : near, ndc ? 1.0f : far);   // datasets/llff.py:236-239: NDC rays span [0, 1]
  }
}

// ---- fused MSE ---------------------------------------------------------------------------------
constexpr int kMseThreads = 512;

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kMseThreads / 32; ++w) r += sh[w];
  __syncthreads();
  return r;   // valid in thread 0
}

// partial[cta] = {sum (a-t)^2, sum (b-t)^2, count}; the CTA that takes the last ticket adds the partials in index
// order (deterministic) and writes loss = sum_a / count (+ sum_b / count).  `ticket` is left at zero.
__global__ void __launch_bounds__(kMseThreads) mse_pair_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                               const float* __restrict__ t, const uint8_t* __restrict__ mask,
                                                               int64_t n, double* __restrict__ partial,
                                                               unsigned int* __restrict__ ticket, float* __restrict__ loss) {
  __shared__ double sh[kMseThreads / 32];
  __shared__ bool last;
  double sa = 0.0, sb = 0.0, cnt = 0.0;
  int64_t first = 0;
  const bool vec = !mask && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(t) |
                              (b ? reinterpret_cast<uintptr_t>(b) : 0)) & 15) == 0;
  if (vec) {   // 16-byte loads over the aligned body; fp32 partial sums of 4 squares, fp64 across iterations
    const int64_t n4 = n >> 2;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
      const float4 tv = __ldg(reinterpret_cast<const float4*>(t) + i), av = __ldg(reinterpret_cast<const float4*>(a) + i);
      const float d0 = av.x - tv.x, d1 = av.y - tv.y, d2 = av.z - tv.z, d3 = av.w - tv.w;
      sa += (double)d0 * d0 + (double)d1 * d1 + (double)d2 * d2 + (double)d3 * d3;
      if (b) {
        const float4 bv = __ldg(reinterpret_cast<const float4*>(b) + i);
        const float e0 = bv.x - tv.x, e1 = bv.y - tv.y, e2 = bv.z - tv.z, e3 = bv.w - tv.w;
        sb += (double)e0 * e0 + (double)e1 * e1 + (double)e2 * e2 + (double)e3 * e3;
      }
      cnt += 4.0;
    }
    first = n4 << 2;
  }
  for (int64_t i = first + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    const float tv = __ldg(t + i);
    const float da = __fsub_rn(__ldg(a + i), tv);
    sa += (double)__fmul_rn(da, da);
    if (b) {
      const float db = __fsub_rn(__ldg(b + i), tv);
      sb += (double)__fmul_rn(db, db);
    }
    cnt += 1.0;
  }
  const double ra = block_sum(sa, sh), rb = block_sum(sb, sh), rc = block_sum(cnt, sh);
  if (threadIdx.x == 0) {
    partial[blockIdx.x * 3 + 0] = ra;
    partial[blockIdx.x * 3 + 1] = rb;
    partial[blockIdx.x * 3 + 2] = rc;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double ta = 0.0, tb = 0.0, tc = 0.0;
    for (unsigned int c = 0; c < gridDim.x; ++c) {
      ta += __ldcg(partial + c * 3 + 0);   // written by other SMs: read through L2
      tb += __ldcg(partial + c * 3 + 1);
      tc += __ldcg(partial + c * 3 + 2);
    }
    float l = (float)(ta / tc);          // nan for an empty selection, like torch.mean of an empty tensor
    if (b) l = __fadd_rn(l, (float)(tb / tc));
    *loss = l;
    *ticket = 0u;
  }
}

__global__ void mse_pair_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ t,
                                    int64_t n, const float* __restrict__ upstream, float* __restrict__ ga,
                                    float* __restrict__ gb) {
  const float s = __fdiv_rn(__fmul_rn(2.0f, __ldg(upstream)), (float)n);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float tv = __ldg(t + i);
    if (ga) ga[i] = __fmul_rn(s, __fsub_rn(__ldg(a + i), tv));
    if (gb) gb[i] = __fmul_rn(s, __fsub_rn(__ldg(b + i), tv));
  }
}

}  // namespace nrf

using namespace nrf;

extern "C" int nrf_gen_rays(const float* c2w, int H, int W, float focal, const int64_t* pix, int64_t n, float near_plane,
                            float far_plane, int ndc, float* rays_out, nrf_stream_t stream) {
  NRF_REQUIRE(H >= 1 && W >= 1 && focal > 0.f && n >= 0, "nrf_gen_rays: bad sizes");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(c2w && rays_out, "nrf_gen_rays: null pointer");
  NRF_REQUIRE(pix || n <= (int64_t)H * W, "nrf_gen_rays: more rays than pixels");
  C2W m;
  for (int i = 0; i < 12; ++i) m.m[i] = c2w[i];
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8);
  gen_rays_kernel<<<blocks, threads, 0, as_stream(stream)>>>(m, H, W, focal, pix, n, near_plane, far_plane, ndc, rays_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" size_t nrf_mse_workspace_bytes(void) { return (size_t)kNumSMs * 3 * sizeof(double) + 16; }

extern "C" int nrf_mse_pair(const float* pred_a, const float* pred_b, const float* target, const uint8_t* mask,
                            int64_t n_elems, float* loss_out, void* workspace, nrf_stream_t stream) {
  NRF_REQUIRE(n_elems >= 0, "nrf_mse_pair: bad size");
  NRF_REQUIRE(pred_a && target && loss_out && workspace, "nrf_mse_pair: null pointer");
  double* partial = static_cast<double*>(workspace);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(partial + (size_t)kNumSMs * 3);
  const int64_t want = (n_elems + kMseThreads * 8 - 1) / (kMseThreads * 8);
  const int blocks = (int)(want < 1 ? 1 : (want < kNumSMs ? want : kNumSMs));
  mse_pair_kernel<<<blocks, kMseThreads, 0, as_stream(stream)>>>(pred_a, pred_b, target, mask, n_elems, partial, ticket, loss_out);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_mse_pair_bwd(const float* pred_a, const float* pred_b, const float* target, int64_t n_elems,
                                const float* upstream, float* grad_a, float* grad_b, nrf_stream_t stream) {
  NRF_REQUIRE(n_elems >= 0, "nrf_mse_pair_bwd: bad size");
  if (n_elems == 0) return NRF_OK;
  NRF_REQUIRE(target && upstream && (grad_a || grad_b), "nrf_mse_pair_bwd: null pointer");
  NRF_REQUIRE((!grad_a || pred_a) && (!grad_b || pred_b), "nrf_mse_pair_bwd: gradient without its prediction");
  const int threads = 256;
  const int64_t want = (n_elems + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8);
  mse_pair_bwd_kernel<<<blocks, threads, 0, as_stream(stream)>>>(pred_a, pred_b, target, n_elems, upstream, grad_a, grad_b);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// ---- a23 fused Adam over the flat parameter / gradient buffers ----------------------------
// torch.optim.Adam as the reference constructs it (utils/__init__.py:17-19: lr, eps = 1e-8, weight_decay as an L2 term
// added to the gradient; betas default (0.9, 0.999); no amsgrad).  One launch for all parameters of all models: they
// live in ONE flat fp32 buffer (optim.FusedAdam), next to the flat gradient buffer the backward kernels accumulate into.
// Hyper-parameters arrive as doubles and are rounded the way torch rounds its Python scalars: beta and (1 - beta)
// separately, so (1 - 0.999) is 0.001f and not 1.0f - 0.999f.
namespace nrf {
struct AdamK {
  float step_size, beta2, eps, weight_decay, w1, w2, bc2_sqrt;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamK& k) {
  if (k.weight_decay != 0.f) g = fmaf(k.weight_decay, p, g);           // grad.add(param, alpha=weight_decay)
  m = fmaf(k.w1, g - m, m);                                              // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(k.w2, g * g, k.beta2 * v);                                    // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = sqrtf(v) / k.bc2_sqrt + k.eps;                    // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
  p = p - k.step_size * (m / denom);                                     // param.addcdiv_(exp_avg, denom, value=-step_size)
}

// Device-resident optimizer state for CUDA-graph replays (the host's step count would be baked into the graph):
// state = { int64 step; double lr; float step_size; float bc2_sqrt }.  One thread advances it ahead of the update.
struct AdamState {
  long long step;
  double lr;
  float step_size, bc2_sqrt;
};

__global__ void adam_tick_kernel(AdamState* st, double beta1, double beta2) {
  const long long t = ++st->step;
  st->step_size = (float)(st->lr / (1.0 - pow(beta1, (double)t)));
  st->bc2_sqrt = (float)sqrt(1.0 - pow(beta2, (double)t));
}

__global__ void __launch_bounds__(256) adam_step_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v, int64_t n, AdamK k,
                                                        const AdamState* __restrict__ st) {
  if (st) { k.step_size = st->step_size; k.bc2_sqrt = st->bc2_sqrt; }
  const int64_t n4 = n >> 2;
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n4; i += nth) {                              // 16-byte lanes: 4 loads in flight per thread
    float4 pv = reinterpret_cast<float4*>(p)[i], mv = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
    adam_one(pv.x, gv.x, mv.x, vv.x, k);
    adam_one(pv.y, gv.y, mv.y, vv.y, k);
    adam_one(pv.z, gv.z, mv.z, vv.z, k);
    adam_one(pv.w, gv.w, mv.w, vv.w, k);
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (int64_t i = (n4 << 2) + tid; i < n; i += nth) adam_one(p[i], g[i], m[i], v[i], k);
}

__global__ void __launch_bounds__(256) adam_step_scalar_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                               float* __restrict__ m, float* __restrict__ v, int64_t n, AdamK k,
                                                               const AdamState* __restrict__ st) {
  if (st) { k.step_size = st->step_size; k.bc2_sqrt = st->bc2_sqrt; }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    adam_one(p[i], g[i], m[i], v[i], k);
}
}  // namespace nrf

static int adam_launch(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, nrf::AdamK k,
                       const nrf::AdamState* st, cudaStream_t stream) {
  const int threads = 256;
  const bool vec = ((reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) |
                     reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0;
  const int64_t want = ((vec ? (n + 3) / 4 : n) + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8);
  if (vec) nrf::adam_step_kernel<<<blocks, threads, 0, stream>>>(param, grad, exp_avg, exp_avg_sq, n, k, st);
  else nrf::adam_step_scalar_kernel<<<blocks, threads, 0, stream>>>(param, grad, exp_avg, exp_avg_sq, n, k, st);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr,
                             double beta1, double beta2, double eps, double weight_decay, int64_t step,
                             nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0 && step >= 1, "nrf_adam_step: bad size or step");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(param && grad && exp_avg && exp_avg_sq, "nrf_adam_step: null pointer");
  nrf::AdamK k;
  k.step_size = (float)(lr / (1.0 - std::pow(beta1, (double)step)));
  k.beta2 = (float)beta2;
  k.eps = (float)eps;
  k.weight_decay = (float)weight_decay;
  k.w1 = (float)(1.0 - beta1);
  k.w2 = (float)(1.0 - beta2);
  k.bc2_sqrt = (float)std::sqrt(1.0 - std::pow(beta2, (double)step));
  return adam_launch(param, grad, exp_avg, exp_avg_sq, n, k, nullptr, as_stream(stream));
}

extern "C" size_t nrf_adam_state_bytes(void) { return sizeof(nrf::AdamState); }

extern "C" int nrf_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                                 double beta1, double beta2, double eps, double weight_decay, void* state,
                                 nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0, "nrf_adam_step_dev: bad size");
  NRF_REQUIRE(state && (reinterpret_cast<uintptr_t>(state) & 7) == 0, "nrf_adam_step_dev: state must be 8-byte aligned");
  NRF_REQUIRE(n == 0 || (param && grad && exp_avg && exp_avg_sq), "nrf_adam_step_dev: null pointer");
  nrf::AdamState* st = static_cast<nrf::AdamState*>(state);
  nrf::adam_tick_kernel<<<1, 1, 0, as_stream(stream)>>>(st, beta1, beta2);
  NRF_LAUNCH_CHECK();
  if (n == 0) return NRF_OK;
  nrf::AdamK k;
  k.step_size = 0.f;
  k.bc2_sqrt = 1.f;
  k.beta2 = (float)beta2;
  k.eps = (float)eps;
  k.weight_decay = (float)weight_decay;
  k.w1 = (float)(1.0 - beta1);
  k.w2 = (float)(1.0 - beta2);
  return adam_launch(param, grad, exp_avg, exp_avg_sq, n, k, st, as_stream(stream));
}

// ---- a28 the reference's other optimizers over the same flat buffers (utils/__init__.py:14-27) -------
// sgd    torch.optim.SGD(lr, momentum, weight_decay): L2 term added to the gradient, momentum buffer seeded with the first
//        gradient, dampening 0, no nesterov.
// radam  utils/optimizers.py:6-95: raw-gradient moments, rectified step (N_sma >= 5) or the SGD-like fallback, weight
//        decay applied to the parameter (p += -wd*lr*p) ahead of the update.
// ranger utils/optimizers.py:266-404: RAdam (threshold N_sma > N_sma_threshhold, decay always applied) + lookahead every
//        k steps (slow += alpha (p - slow); p = slow), slow buffer seeded with the initial parameters by the caller.
// The step-dependent scalars (N_sma, step_size) are computed on the host in double like the reference's Python floats.
namespace nrf {
struct RAdamK {
  float beta1, beta2, w1, w2, eps, decay, step, alpha;   // decay = -wd*lr, step = -step_size*lr
  int rectified, update, lookahead;
};

__device__ __forceinline__ void radam_one(float& p, float g, float& m, float& v, float* slow, const RAdamK& k) {
  v = v * k.beta2 + k.w2 * g * g;                    // exp_avg_sq.mul_(beta2).addcmul_(1 - beta2, grad, grad)
  m = m * k.beta1 + k.w1 * g;                        // exp_avg.mul_(beta1).add_(1 - beta1, grad)
  if (k.update) {
    if (k.decay != 0.f) p = p + k.decay * p;
    if (k.rectified) p = p + k.step * (m / (sqrtf(v) + k.eps));
    else p = p + k.step * m;
  }
  if (k.lookahead) {
    const float s = *slow + k.alpha * (p - *slow);
    *slow = s;
    p = s;
  }
}

__global__ void __launch_bounds__(256) radam_step_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                         float* __restrict__ m, float* __restrict__ v,
                                                         float* __restrict__ slow, int64_t n, RAdamK k) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    radam_one(p[i], g[i], m[i], v[i], slow ? slow + i : nullptr, k);
}

__global__ void __launch_bounds__(256) sgd_step_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ buf, int64_t n, float lr, float momentum,
                                                       float weight_decay, int first) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float d = g[i];
    if (weight_decay != 0.f) d = fmaf(weight_decay, p[i], d);          // grad.add(param, alpha=weight_decay)
    if (buf) {
      d = first ? d : buf[i] * momentum + d;                           // buf.mul_(momentum).add_(grad)
      buf[i] = d;
    }
    p[i] = p[i] - lr * d;                                              // param.add_(grad, alpha=-lr)
  }
}

static inline int eltwise_grid(int64_t n) {
  const int64_t want = (n + 255) / 256;
  return (int)(want < 1 ? 1 : (want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8));
}
}  // namespace nrf

extern "C" int nrf_sgd_step(float* param, const float* grad, float* momentum_buf, int64_t n, double lr, double momentum,
                            double weight_decay, int first_step, nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0, "nrf_sgd_step: bad size");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(param && grad, "nrf_sgd_step: null pointer");
  NRF_REQUIRE(momentum == 0.0 || momentum_buf, "nrf_sgd_step: momentum %g needs a buffer", momentum);
  nrf::sgd_step_kernel<<<nrf::eltwise_grid(n), 256, 0, as_stream(stream)>>>(
      param, grad, momentum != 0.0 ? momentum_buf : nullptr, n, (float)lr, (float)momentum, (float)weight_decay, first_step);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

extern "C" int nrf_radam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* slow, int64_t n,
                              double lr, double beta1, double beta2, double eps, double weight_decay, int64_t step,
                              int ranger, double n_sma_threshold, int degenerated_to_sgd, double alpha, int k,
                              nrf_stream_t stream) {
  NRF_REQUIRE(n >= 0 && step >= 1, "nrf_radam_step: bad size or step");
  NRF_REQUIRE(!ranger || k >= 1, "nrf_radam_step: lookahead period must be >= 1");
  if (n == 0) return NRF_OK;
  NRF_REQUIRE(param && grad && exp_avg && exp_avg_sq && (!ranger || slow), "nrf_radam_step: null pointer");
  // utils/optimizers.py:66-80 (RAdam), :363-373 (Ranger), in double like the Python floats
  const double beta2_t = std::pow(beta2, (double)step);
  const double n_sma_max = 2.0 / (1.0 - beta2) - 1.0;
  const double n_sma = n_sma_max - 2.0 * (double)step * beta2_t / (1.0 - beta2_t);
  const double bc1 = 1.0 - std::pow(beta1, (double)step);
  const bool rect = ranger ? n_sma > n_sma_threshold : n_sma >= 5.0;
  double step_size;
  if (rect)
    step_size = std::sqrt((1.0 - beta2_t) * (n_sma - 4.0) / (n_sma_max - 4.0) * (n_sma - 2.0) / n_sma * n_sma_max /
                          (n_sma_max - 2.0)) / bc1;
  else
    step_size = (ranger || degenerated_to_sgd) ? 1.0 / bc1 : -1.0;
  nrf::RAdamK kk;
  kk.beta1 = (float)beta1; kk.beta2 = (float)beta2;
  kk.w1 = (float)(1.0 - beta1); kk.w2 = (float)(1.0 - beta2);
  kk.eps = (float)eps;
  kk.decay = weight_decay != 0.0 ? (float)(-weight_decay * lr) : 0.f;
  kk.step = (float)(-step_size * lr);
  kk.alpha = (float)alpha;
  kk.rectified = rect;
  kk.update = rect || step_size > 0.0;
  kk.lookahead = ranger && (step % k == 0);
  nrf::radam_step_kernel<<<nrf::eltwise_grid(n), 256, 0, as_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq,
                                                                               ranger ? slow : nullptr, n, kk);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
