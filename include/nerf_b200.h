/* nerf_b200.h — C ABI of the B200-native nerf_pl ray-rendering hot path.
 *
 * One shared library (nerf_pl_b200/lib/libnerf_b200.so), sm_100a only, no CPU fallback.
 * Every entry point:
 *   - takes DEVICE pointers to contiguous row-major buffers (fp32 unless stated), plain sizes,
 *     and the CUDA stream to launch on (a cudaStream_t passed as void*; NULL = legacy default);
 *   - returns 0 on success, otherwise an NRF_ERR_* code; nrf_last_error() gives the text;
 *   - is asynchronous w.r.t. the host (no device synchronisation inside).
 *
 * The reference (ktiwary2/nerf_pl @ 99d40cb) has no FFI of its own for this path except the
 * torchsearchsorted extension; its "operator API" is the Python functions cited below.  Each
 * entry point names the reference code it replaces (paths relative to the reference root).
 */
#ifndef NERF_B200_H_
#define NERF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRF_OK 0
#define NRF_ERR_INVALID 1     /* bad argument (shape, null pointer, unsupported size) */
#define NRF_ERR_CUDA 2        /* a CUDA runtime call or launch failed */
#define NRF_ERR_UNSUPPORTED 3 /* valid request this build cannot serve (e.g. W != 256 on the tensor-core path) */
#define NRF_ERR_WORKSPACE 4   /* caller-provided workspace too small */

#define NRF_MAX_D 16

typedef void* nrf_stream_t;

/* Library / device ------------------------------------------------------------------------- */
int nrf_abi_version(void);
const char* nrf_last_error(void);
/* 0 iff the current device is compute capability 10.x (the only supported target). */
int nrf_check_device(void);
/* Number of CUDA kernels this library has launched in the calling process so far. */
uint64_t nrf_launch_count(void);

/* Per-kernel device timing of the tensor-core MLP kernels (measurement aid; bench.py's roofline).  While enabled,
 * every launch is bracketed by a CUDA event pair on its own stream.  nrf_prof_enable(1) discards earlier records and
 * starts a session, nrf_prof_enable(0) stops recording (records stay readable); nrf_prof_read waits for the
 * recorded launches of one kernel and returns their summed duration, their count and the sample rows they processed. */
#define NRF_K_FWD_SIGMA_INFER 0
#define NRF_K_FWD_FULL_INFER 1
#define NRF_K_FWD_SIGMA_TRAIN 2
#define NRF_K_FWD_FULL_TRAIN 3
#define NRF_K_DGRAD_SIGMA 4
#define NRF_K_DGRAD_FULL 5
#define NRF_K_WGRAD_SIGMA 6
#define NRF_K_WGRAD_FULL 7
#define NRF_K_COUNT 8
int nrf_prof_enable(int on);
int nrf_prof_read(int kernel_id, double* total_ms, int64_t* launches, int64_t* rows);

/* a21 camera-ray generation — datasets/ray_utils.py:5-93 (get_ray_directions, get_rays, get_ndc_rays) fused with the
 * datasets' row assembly (datasets/blender.py:80-84).  c2w: HOST pointer to the 3x4 camera-to-world matrix (row-major);
 * pix (n) device int64 flat pixel indices row*W+col, or NULL for the first n pixels in row-major order;
 * rays_out (n,8) = [o, d/|d|, near, far].  ndc != 0: rays are mapped to NDC with the near plane `near_plane`
 * (ray_utils.py:52-93) and the depth bounds written are [0, 1] (datasets/llff.py:236-239). */
int nrf_gen_rays(const float* c2w, int H, int W, float focal, const int64_t* pix, int64_t n, float near_plane,
                 float far_plane, int ndc, float* rays_out, nrf_stream_t stream);

/* a22 fused MSE — losses.py:4-26 (MSELoss / SMMSELoss: mean((a-t)^2) [+ mean((b-t)^2)]) and metrics.py:4-10 (mse with an
 * optional element mask).  pred_b, mask may be NULL; mask is one byte per element (non-zero = counted).  One launch,
 * deterministic (fixed summation order); loss_out (1) device float.  workspace: nrf_mse_workspace_bytes() bytes, zeroed
 * once by the caller (the kernel leaves it reusable).  nrf_mse_pair_bwd writes 2 (x - t) / n_elems * upstream[0]
 * for each non-NULL gradient (no mask). */
size_t nrf_mse_workspace_bytes(void);
int nrf_mse_pair(const float* pred_a, const float* pred_b, const float* target, const uint8_t* mask, int64_t n_elems,
                 float* loss_out, void* workspace, nrf_stream_t stream);
int nrf_mse_pair_bwd(const float* pred_a, const float* pred_b, const float* target, int64_t n_elems, const float* upstream,
                     float* grad_a, float* grad_b, nrf_stream_t stream);

/* a23 fused Adam — torch.optim.Adam as constructed by utils/__init__.py:17-19 (L2 weight decay added to
 * the gradient, bias-corrected, no amsgrad) over flat fp32 buffers of n elements; `step` counts from 1.  Hyper-parameters
 * are doubles, rounded to fp32 the way torch rounds its Python scalars (beta and 1 - beta separately). */
int nrf_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr, double beta1,
                  double beta2, double eps, double weight_decay, int64_t step, nrf_stream_t stream);
/* The same step with the optimizer state on the device, for CUDA-graph replays (a host step count would be baked into
 * the graph).  `state` = nrf_adam_state_bytes() bytes, 8-byte aligned: { int64 step (starts at 0); double lr; 2 floats
 * of scratch }.  Each call advances `step` by one on the device and reads `lr` from there; the host writes `lr` (offset 8)
 * whenever a scheduler changes it. */
size_t nrf_adam_state_bytes(void);
int nrf_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double beta1,
                      double beta2, double eps, double weight_decay, void* state, nrf_stream_t stream);
/* a28 the reference's other optimizers over the same flat buffers (utils/__init__.py:14-27).
 * nrf_sgd_step: torch.optim.SGD(lr, momentum, weight_decay) — L2 term added to the gradient, momentum buffer seeded
 * with the first gradient (first_step != 0), dampening 0, no nesterov; momentum_buf may be NULL when momentum == 0.
 * nrf_radam_step: ranger == 0: RAdam (utils/optimizers.py:6-95; rectified when N_sma >= 5, else the SGD-like step when
 * degenerated_to_sgd, else moments only); ranger != 0: Ranger (utils/optimizers.py:266-404; threshold N_sma >
 * n_sma_threshold, lookahead slow += alpha (p - slow), p = slow every k steps; `slow` holds the initial parameters before
 * the first step).  `step` counts from 1.  Weight decay multiplies the parameter (p += -wd*lr*p) ahead of the update. */
int nrf_sgd_step(float* param, const float* grad, float* momentum_buf, int64_t n, double lr, double momentum,
                 double weight_decay, int first_step, nrf_stream_t stream);
int nrf_radam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* slow, int64_t n, double lr,
                   double beta1, double beta2, double eps, double weight_decay, int64_t step, int ranger,
                   double n_sma_threshold, int degenerated_to_sgd, double alpha, int k, nrf_stream_t stream);

/* a4  stratified depth sampling — models/rendering.py:216-232
 * rays (N,8) = [o(3) d(3) near far]; t_steps (n_samples) = linspace(0,1,n_samples);
 * u (N,n_samples) uniform [0,1) or NULL (required iff perturb > 0); z_out (N,n_samples). */
int nrf_sample_coarse(const float* rays, const float* t_steps, int64_t n_rays, int n_samples,
                      int use_disp, float perturb, const float* u, float* z_out, nrf_stream_t stream);

/* a6  positional encoding — models/nerf.py:21-38 (Embedding.forward)
 * x (rows,channels) -> out (rows, channels*(2*n_freqs+1)); freqs (n_freqs) device array. */
int nrf_embed_fwd(const float* x, int64_t rows, int channels, const float* freqs, int n_freqs,
                  float* out, nrf_stream_t stream);
int nrf_embed_bwd(const float* x, const float* grad_out, int64_t rows, int channels, const float* freqs,
                  int n_freqs, float* grad_x, nrf_stream_t stream);

/* a10 alpha compositing — models/rendering.py:170-198 (+disp rendering_shadows.py:198)
 * sigma (N,S); rgb (N,S,3) or NULL (depth-only / weights-only); z (N,S); rays (N,8);
 * noise (N,S) or NULL.  Outputs: weights (N,S); rgb_out (N,3) or NULL; depth/opacity (N);
 * disp_out (N) or NULL. */
int nrf_composite_fwd(const float* sigma, const float* rgb, const float* z, const float* rays,
                      const float* noise, float noise_std, int white_back, int64_t n_rays, int n_samples,
                      float* weights_out, float* rgb_out, float* depth_out, float* opacity_out,
                      float* disp_out, nrf_stream_t stream);
/* Gradients of a scalar loss w.r.t. sigma (N,S) and rgb (N,S,3), given the gradients of the
 * outputs (any of g_* may be NULL = zero).  g_weights (N,S) or NULL. */
int nrf_composite_bwd(const float* sigma, const float* rgb, const float* z, const float* rays,
                      const float* noise, float noise_std, int white_back, int64_t n_rays, int n_samples,
                      const float* g_rgb, const float* g_depth, const float* g_opacity, const float* g_disp,
                      const float* g_weights, float* d_sigma, float* d_rgb, nrf_stream_t stream);

/* a11 hierarchical resampling — models/rendering.py:14-48 (sample_pdf as called at :253)
 * weights_coarse (N,n_coarse): the FULL coarse weights; bins are weights[:,1:-1].
 * u, jitter (N,n_importance).  z_fine_out (N,n_importance); idx_out (N,n_importance) int64 or
 * NULL (clamped bin index); cdf_out (N,n_coarse-1) or NULL. */
int nrf_sample_pdf(const float* rays, const float* weights_coarse, int64_t n_rays, int n_coarse,
                   int n_importance, const float* u, const float* jitter, float eps, float* z_fine_out,
                   int64_t* idx_out, float* cdf_out, nrf_stream_t stream);

/* a13 sort(cat([z_coarse, z_fine])) per ray — models/rendering.py:257.  z_out (N,nc+nf). */
int nrf_merge_sort(const float* z_coarse, const float* z_fine, int64_t n_rays, int nc, int nf,
                   float* z_out, nrf_stream_t stream);

/* a12 row-wise searchsorted — torchsearchsorted/src/cuda/searchsorted_cuda_kernel.cu:83-142,
 * src/torchsearchsorted/searchsorted.py:20-53.  a (nrow_a,ncol_a) sorted rows, v (nrow_v,ncol_v),
 * out (max(nrow_a,nrow_v), ncol_v) int64; one of nrow_a/nrow_v may be 1 (broadcast). */
int nrf_searchsorted(const float* a, const float* v, int64_t* out, int64_t nrow_a, int64_t nrow_v,
                     int64_t ncol_a, int64_t ncol_v, int side_left, nrf_stream_t stream);
/* The same for every dtype the reference dispatches (AT_DISPATCH_ALL_TYPES, searchsorted_cuda_kernel.cu:134): `dtype` is
 * one of NRF_DT_*; a and v share it.  Other codes return NRF_ERR_INVALID with the reference's "not implemented for" text. */
enum { NRF_DT_F32 = 0, NRF_DT_F64 = 1, NRF_DT_I64 = 2, NRF_DT_I32 = 3, NRF_DT_I16 = 4, NRF_DT_I8 = 5, NRF_DT_U8 = 6 };
int nrf_searchsorted_typed(const void* a, const void* v, int64_t* out, int64_t nrow_a, int64_t nrow_v,
                           int64_t ncol_a, int64_t ncol_v, int side_left, int dtype, nrf_stream_t stream);

/* a8  NeRF MLP — models/nerf.py:41-124.  Parameters in the checkpoint layout
 * (utils/__init__.py:55-76): weights (out,in) row-major fp32, biases (out). */
typedef struct {
  int32_t D, W, in_xyz, in_dir;
  uint32_t skip_mask;                 /* bit i set <=> layer i concatenates the encoding first */
  const float* xyz_w[NRF_MAX_D];
  const float* xyz_b[NRF_MAX_D];
  const float* final_w; const float* final_b;
  const float* dir_w;   const float* dir_b;
  const float* sigma_w; const float* sigma_b;
  const float* rgb_w;   const float* rgb_b;
} nrf_mlp_params;

typedef struct {                      /* same shapes; gradients are ACCUMULATED (+=) */
  float* xyz_w[NRF_MAX_D];
  float* xyz_b[NRF_MAX_D];
  float* final_w; float* final_b;
  float* dir_w;   float* dir_b;
  float* sigma_w; float* sigma_b;
  float* rgb_w;   float* rgb_b;
} nrf_mlp_grads;

/* fp32 (parity) MLP.  Two input forms:
 *   rays form     : points o+d*z are encoded on the fly (rendering.py:234, nerf.py:21-38):
 *                   rays (N,8), z (N,S), rows = N*S, row r belongs to ray r/S.
 *   embedded form : x (rows, in_xyz[+in_dir]) already encoded (NeRF.forward, nerf.py:83).
 * sigma_out (rows); rgb_out (rows,3) or NULL when sigma_only.
 * stash: workspace of >= nrf_mlp_fp32_workspace_bytes(...) bytes; with need_grad it holds the
 * activations consumed by nrf_mlp_fp32_bwd. */
size_t nrf_mlp_fp32_workspace_bytes(const nrf_mlp_params* p, int64_t rows, int64_t n_rays, int sigma_only,
                                    int need_grad);
int nrf_mlp_fp32_fwd_rays(const nrf_mlp_params* p, const float* rays, const float* z, int64_t n_rays,
                          int n_samples, int xyz_freqs, int dir_freqs, int sigma_only, int need_grad,
                          float* sigma_out, float* rgb_out, void* workspace, size_t workspace_bytes,
                          nrf_stream_t stream);
int nrf_mlp_fp32_fwd_embedded(const nrf_mlp_params* p, const float* x, int64_t rows, int sigma_only,
                              int need_grad, float* sigma_out, float* rgb_out, void* workspace,
                              size_t workspace_bytes, nrf_stream_t stream);
/* d_sigma (rows), d_rgb (rows,3) or NULL; `rgb_out` is the forward output (sigmoid backward).
 * x/n_rays/n_samples describe the same input form as the forward that filled `workspace`
 * (x == NULL for the rays form). */
int nrf_mlp_fp32_bwd(const nrf_mlp_params* p, const nrf_mlp_grads* g, const float* x, int64_t rows,
                     int64_t n_rays, int n_samples, int sigma_only, const float* d_sigma, const float* d_rgb,
                     const float* rgb_out, void* workspace, size_t workspace_bytes, nrf_stream_t stream);

/* bf16 tensor-core (tcgen05/TMEM/TMA) MLP, D=8 W=256 L=(10,4) skips=[4] only.
 * packed: device buffer of nrf_mlp_tc_packed_bytes() bytes filled by nrf_mlp_tc_pack(). */
size_t nrf_mlp_tc_packed_bytes(void);
int nrf_mlp_tc_pack(const nrf_mlp_params* p, void* packed, nrf_stream_t stream);
size_t nrf_mlp_tc_stash_bytes(int64_t rows, int sigma_only);
int nrf_mlp_tc_fwd(const void* packed, const float* rays, const float* z, int64_t n_rays, int n_samples,
                   int sigma_only, float* sigma_out, float* rgb_out, void* stash, size_t stash_bytes,
                   nrf_stream_t stream);
/* d_sigma (rows) must be 16-byte aligned: slices of it are bulk-copied into the weight-gradient kernel's stages. */
size_t nrf_mlp_tc_bwd_workspace_bytes(int64_t rows, int sigma_only);
int nrf_mlp_tc_bwd(const void* packed, const nrf_mlp_grads* g, const float* rays, const float* z,
                   int64_t n_rays, int n_samples, int sigma_only, const float* d_sigma, const float* d_rgb,
                   const float* rgb_out, const void* stash, size_t stash_bytes, void* workspace,
                   size_t workspace_bytes, nrf_stream_t stream);

/* a14-a19 shadow test — models/efficient_shadow_mapping.py:19-130, models/camera.py:121-132,
 * models/rendering_shadows.py:359-482 (segmentation by eye_pos :385-406).
 * cam_pixels (n,3) [i+.5, j+.5, 1]; cam_depth (n); cam_M (n,3,3); cam_eye (n,3);
 * light_Minv (3,3) = inverse(light.camera); light_M (3,3); light_eye (3);
 * light_pixels (W*H,3); light_depth (W*H).  If light_M and light_pixels are both NULL,
 * light_depth is taken as already normalised (the contract of run_shadow_mapping :29-30).
 * mode: 1 = shadow_method_1, 2 = shadow_method_2.
 * out (n,3) = clip(shadow,0,1) + out_bias.  workspace >= nrf_shadow_workspace_bytes(n, W*H). */
size_t nrf_shadow_workspace_bytes(int64_t n, int64_t n_light);
int nrf_shadow_fwd(const float* cam_pixels, const float* cam_depth, const float* cam_M, const float* cam_eye,
                   const float* light_Minv, const float* light_M, const float* light_eye,
                   const float* light_pixels, const float* light_depth, int64_t n, int res_w, int res_h,
                   int mode, float delta, float epsilon, float new_min, float new_max, int sigmoid,
                   float out_bias, float* out, void* workspace, size_t workspace_bytes, nrf_stream_t stream);
/* nrf_shadow_fwd without a workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, arithmetic and output as nrf_shadow_fwd; the scratch is a file-scope
 * `static __device__` array in shadow.cu sized for one config-5 batch - at most NRF_SHADOW_STATIC_MAX_RAYS camera rays
 * against a light map of at most NRF_SHADOW_STATIC_MAX_TEXELS texels (368 KB, held by every GPU that loads the library
 * until exit); larger sizes return NRF_ERR_WORKSPACE with a message and launch nothing.  The call performs no cudaMalloc /
 * cudaFree (its device address is looked up once per device with cudaGetSymbolAddress).  All calls on a device share the
 * one array, which keeps the previous call's contents (every word read is rewritten first, nothing relies on zeros):
 * calls that overlap on the GPU (two streams, two host threads) race - serialise them on one stream. */
#define NRF_SHADOW_STATIC_MAX_RAYS 4096
#define NRF_SHADOW_STATIC_MAX_TEXELS 16384
int nrf_shadow_fwd_static(const float* cam_pixels, const float* cam_depth, const float* cam_M, const float* cam_eye,
                          const float* light_Minv, const float* light_M, const float* light_eye,
                          const float* light_pixels, const float* light_depth, int64_t n, int res_w, int res_h,
                          int mode, float delta, float epsilon, float new_min, float new_max, int sigmoid,
                          float out_bias, float* out, nrf_stream_t stream);
/* g_out (n,3) -> g_cam_depth (n) (overwritten), g_light_depth (W*H) (ACCUMULATED, may be NULL). */
int nrf_shadow_bwd(const float* cam_pixels, const float* cam_depth, const float* cam_M, const float* cam_eye,
                   const float* light_Minv, const float* light_M, const float* light_eye,
                   const float* light_pixels, const float* light_depth, int64_t n, int res_w, int res_h,
                   int mode, float delta, float epsilon, float new_min, float new_max, int sigmoid,
                   const float* g_out, float* g_cam_depth, float* g_light_depth, void* workspace,
                   size_t workspace_bytes, nrf_stream_t stream);

/* a14 get_normed_w — models/efficient_shadow_mapping.py:47-62.  pixel_depth (n,4) [i,j,1,depth], M (3,3).
 * out (n,4) = [passthrough[:, :3], value[:,3] / (|M pixel_depth[:, :3]| + 1e-5)].  Forward: value =
 * passthrough = pixel_depth.  Backward w.r.t. depth: value = passthrough = grad_out (linear map). */
int nrf_normed_w(const float* pixel_depth, const float* value, const float* passthrough, const float* M,
                 int64_t n, float* out, nrf_stream_t stream);

/* SURVEY 8f rank 2 — the shadow-mapping pieces the light-sampler trainer calls one by one
 * (train_light_sampler.py:141-196).
 * a24 get_K / get_projections / get_diff_projections — models/rendering_shadows.py:484-546,
 * models/efficient_shadow_mapping.py:13-17, :65-83.  K (n,3) = [ul, vl, wl] of every camera ray in the light's image.
 * n_cam = n: cam_M (n,3,3), cam_eye (n,3) per ray, segmented by eye_pos like the reference's loop (seg_ws: n ints);
 * n_cam = 1: one camera for the batch (seg_ws may be NULL).  normalised: 0 = cam_depth is a range,
 * w = depth / (|M pix| + 1e-5) (get_normed_w); 1 = cam_depth already is w (get_diff_projections); 2 = range without
 * the epsilon (models/shadow_mapping_utils.py:39-66, whole-image variant).  g_K == NULL: forward, out = K (n,3); else backward w.r.t. cam_depth, out = g_depth (n). */
int nrf_project_to_light(const float* cam_pixels, const float* cam_depth, const float* cam_M, const float* cam_eye,
                         int64_t n_cam, const float* light_Minv, const float* light_eye, int64_t n, int normalised,
                         const float* g_K, float* out, int* seg_ws, nrf_stream_t stream);
/* a25 generate_shadow_map / normalize_min_max — models/efficient_shadow_mapping.py:10-11, :103-130 on explicit
 * vectors: diff = wl - w_light_bounded (w_light_bounded NULL: diff = wl).  flags bit 0 (stack_clip): out (n,3) = stack and
 * clip(0,1) (generate_shadow_map), else out (n) (mode 2 without sigmoid = normalize_min_max); flags bit 1: min-max
 * normalisation without the 1e-5 in the denominator (models/shadow_mapping_utils.py:8-9, :149-174).
 * Backward: g_out -> g_diff (n) (= d/dwl; d/dw_light_bounded is its negative).  workspace: 32 bytes. */
size_t nrf_shadow_map_workspace_bytes(void);
int nrf_shadow_map_fwd(const float* wl, const float* w_light_bounded, int64_t n, int mode, float delta, float epsilon,
                       float new_min, float new_max, int sigmoid, int flags, float* out, void* workspace,
                       nrf_stream_t stream);
int nrf_shadow_map_bwd(const float* wl, const float* w_light_bounded, int64_t n, int mode, float delta, float epsilon,
                       float new_min, float new_max, int sigmoid, int flags, const float* g_out, float* g_diff,
                       void* workspace, nrf_stream_t stream);
/* a26 get_projected_depths — models/efficient_shadow_mapping.py:85-100: w_light_bounded[i] =
 * w_light.view(w,h)[clamp(vl), clamp(ul)] (truncated), gidx[i] = the flat index read.  nrf_light_scatter is its
 * backward: g_wlight (n_light) = 0, then += g_wlb[i] at gidx[i]. */
int nrf_light_gather(const float* K, const float* w_light, int64_t n, int res_w, int res_h, float* w_light_bounded,
                     int* gidx, nrf_stream_t stream);
int nrf_light_scatter(const float* g_wlb, const int* gidx, int64_t n, int64_t n_light, float* g_wlight,
                      nrf_stream_t stream);
/* a27 differentiable_w_mapping — models/shadow_mapping_utils.py:39-66 (whole-image shadow mapping of
 * train_shadow_mapping.py): out (n) = value / (|M pixels_i| + eps), pixels (n,3) = [i+.5, j+.5, 1]; eps = 0 there.
 * Linear in value: the backward is the same call with value = grad_out. */
int nrf_w_mapping(const float* pixels, const float* value, const float* M, int64_t n, float eps, float* out,
                  nrf_stream_t stream);

/* (e) Gradient exchange of the ray-sharded training step over NVLink peer memory - replaces the NCCL all-reduce that
 * DistributedDataParallel issues for the reference (train.py:174-175).  The flat fp32 gradient buffer of every
 * rank is a cudaMalloc allocation shared with the other ranks' processes through CUDA IPC:
 *   nrf_peer_alloc / nrf_peer_free      zero-filled device allocation owned by this process
 *   nrf_peer_export                     64-byte IPC handle of an allocation (cudaIpcGetMemHandle)
 *   nrf_peer_import / nrf_peer_close    map a peer's allocation into this process (peer access enabled lazily)
 *   nrf_peer_allreduce_mean             bufs[q][0..n) <- mean over q of bufs[q][0..n), in place, in every rank's buffer:
 *                                       ONE launch per rank; rank r reduces slice r in rank order 0..world-1 (all ranks end
 *                                       up bit-identical) and stores it into all buffers; two flag barriers in the peers'
 *                                       memory (flags[q]: world words of rank q, written by the peers; seq / counter: this
 *                                       rank's round number and CTA counter, zero-initialised).  Every rank must call it the
 *                                       same number of times; capturable in a CUDA graph. */
#define NRF_PEER_MAX 8
typedef struct {
  void* bufs[NRF_PEER_MAX];    /* gradient buffer of every rank, as mapped in THIS process */
  void* flags[NRF_PEER_MAX];   /* flag words of every rank (>= NRF_PEER_MAX x 4 bytes each), as mapped in this process */
  int32_t rank, world;
  int64_t n;                   /* fp32 elements */
  void* seq;                   /* this rank: 4 bytes, round number */
  void* counter;               /* this rank: 4 bytes, CTA counter */
} nrf_peer_comm;
int nrf_peer_alloc(size_t bytes, void** ptr);
int nrf_peer_free(void* ptr);
int nrf_peer_export(const void* ptr, void* handle64);
int nrf_peer_import(const void* handle64, void** ptr);
int nrf_peer_close(void* ptr);
int nrf_peer_allreduce_mean(const nrf_peer_comm* comm, nrf_stream_t stream);

/* Diagnostics and microbenchmarks are NOT part of this library: they are compiled only with NRF_DEBUG_BUILD=1 and
 * declared in include/nerf_b200_debug.h. */

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H_ */
