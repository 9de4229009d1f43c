// fp32 NeRF MLP (models/nerf.py:41-124): forward, and backward w.r.t. all parameters.
//
// This is the *parity* mode of the MLP (end-to-end 1e-5 criterion): plain fp32 FFMA, no
// tensor cores, one strided SIMT GEMM kernel used for forward (X·Wᵀ), dgrad (dY·W) and wgrad
// (dYᵀ·X, split over the sample dimension with fp32 atomics).  The concatenations of the
// reference (skip connection nerf.py:108-109, direction input :118) are never materialised:
// a concatenated operand is consumed as two GEMM passes over the two sources (beta=1 on the
// second), and the per-ray direction encoding is broadcast by a row divisor instead of
// repeat_interleave (rendering.py:145).  The performance path is mlp_tc.cu.
#include "common.cuh"
#include "embed.cuh"

namespace nrf {

constexpr int BM = 128, BN = 128, BK = 16, PAD = 4;

struct GemmOp {
  // C[i,j] (+)= sum_r Aop(i,r) * Bop(r,j)
  const float* A; int64_t sa_i, sa_r; int a_div_i, a_div_r;
  const float* B; int64_t sb_r, sb_j; int b_div_r;
  float* C; int64_t ldc;
  int64_t I, J, R;
  const float* bias;      // per j, or null
  const float* mask;      // same layout as C: multiply by (mask > 0), or null
  float* a_colsum;        // wgrad only: a_colsum[i] += sum_r Aop(i,r) (bias gradient), or null
  int beta;               // 1: add existing C before the epilogue
  int act;                // 0 none, 1 relu, 2 sigmoid
  int atomic;             // 1: atomicAdd acc into C (split-R accumulation), epilogue flags ignored
  int64_t r_per_split;
};

// Register-prefetched, double-buffered SIMT GEMM: the next K tile's global loads are issued before the FFMAs of the
// current one, index arithmetic (row offsets, the broadcast divisors) is hoisted out of the K loop, one barrier per tile.
template <bool A_RCONTIG, bool B_JCONTIG>
__global__ void __launch_bounds__(256, 2) gemm_fp32_kernel(const GemmOp op) {
  __shared__ float As[2][BK][BM + PAD];
  __shared__ float Bs[2][BK][BN + PAD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t i0 = (int64_t)blockIdx.y * BM, j0 = (int64_t)blockIdx.x * BN;
  const int64_t r_begin = (int64_t)blockIdx.z * op.r_per_split;
  const int64_t r_end = min(op.R, r_begin + op.r_per_split);
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
  float asum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const bool do_colsum = op.a_colsum != nullptr && blockIdx.x == 0 && tx == 0;

  // element e of this thread's 8 A loads sits at (i, r) = A_RCONTIG ? ((tid>>4) + 16 e, tid & 15) : (tid & 127, (tid>>7) + 2 e)
  // and of its 8 B loads at (r, j) = B_JCONTIG ? ((tid>>7) + 2 e, tid & 127) : (tid & 15, (tid>>4) + 16 e)
  const int a_i0 = A_RCONTIG ? (tid >> 4) : (tid & 127), a_r0 = A_RCONTIG ? (tid & 15) : (tid >> 7);
  const int b_j0 = B_JCONTIG ? (tid & 127) : (tid >> 4), b_r0 = B_JCONTIG ? (tid >> 7) : (tid & 15);
  const float* a_row[A_RCONTIG ? 8 : 1];          // row bases (null: out of range)
  const float* b_col[B_JCONTIG ? 1 : 8];
#pragma unroll
  for (int e = 0; e < (A_RCONTIG ? 8 : 1); ++e) {
    const int64_t gi = i0 + a_i0 + (A_RCONTIG ? 16 * e : 0);
    a_row[e] = gi < op.I ? op.A + (gi / op.a_div_i) * op.sa_i : nullptr;
  }
#pragma unroll
  for (int e = 0; e < (B_JCONTIG ? 1 : 8); ++e) {
    const int64_t gj = j0 + b_j0 + (B_JCONTIG ? 0 : 16 * e);
    b_col[e] = gj < op.J ? op.B + gj * op.sb_j : nullptr;
  }
  const unsigned b_div = (unsigned)op.b_div_r;
  float pa[8], pb[8];

  auto fetch = [&](int64_t r0) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int64_t gr = r0 + a_r0 + (A_RCONTIG ? 0 : 2 * e);
      const float* row = a_row[A_RCONTIG ? e : 0];
      pa[e] = (row != nullptr && gr < r_end) ? __ldg(row + gr * op.sa_r) : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int64_t gr = r0 + b_r0 + (B_JCONTIG ? 2 * e : 0);
      const float* col = b_col[B_JCONTIG ? 0 : e];
      const int64_t br = b_div == 1u ? gr : (int64_t)((unsigned)gr / b_div);     // rows < 2^31 (grid.y limit)
      pb[e] = (col != nullptr && gr < r_end) ? __ldg(col + br * op.sb_r) : 0.f;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      if (A_RCONTIG) As[buf][a_r0][a_i0 + 16 * e] = pa[e]; else As[buf][a_r0 + 2 * e][a_i0] = pa[e];
      if (B_JCONTIG) Bs[buf][b_r0 + 2 * e][b_j0] = pb[e]; else Bs[buf][b_r0][b_j0 + 16 * e] = pb[e];
    }
  };

  int cur = 0;
  if (r_begin < r_end) {
    fetch(r_begin);
    stash(0);
  }
  __syncthreads();
  for (int64_t r0 = r_begin; r0 < r_end; r0 += BK) {
    const bool more = r0 + BK < r_end;
    if (more) fetch(r0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[8], b[8];
      *reinterpret_cast<float4*>(&a[0]) = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8]);
      *reinterpret_cast<float4*>(&a[4]) = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8 + 4]);
      // columns tx*4 .. +3 and 64 + tx*4 .. +3: a quarter warp reads 128 contiguous bytes (no bank conflicts)
      *reinterpret_cast<float4*>(&b[0]) = *reinterpret_cast<const float4*>(&Bs[cur][kk][tx * 4]);
      *reinterpret_cast<float4*>(&b[4]) = *reinterpret_cast<const float4*>(&Bs[cur][kk][64 + tx * 4]);
#pragma unroll
      for (int x = 0; x < 8; ++x)
#pragma unroll
        for (int y = 0; y < 8; ++y) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
      if (do_colsum) {
#pragma unroll
        for (int x = 0; x < 8; ++x) asum[x] += a[x];
      }
    }
    if (more) stash(cur ^ 1);
    __syncthreads();
    cur ^= 1;
  }

#pragma unroll
  for (int x = 0; x < 8; ++x) {
    const int64_t gi = i0 + ty * 8 + x;
    if (gi >= op.I) continue;
    if (do_colsum) atomicAdd(op.a_colsum + gi, asum[x]);
#pragma unroll
    for (int y = 0; y < 8; ++y) {
      const int64_t gj = j0 + (y < 4 ? tx * 4 + y : 64 + tx * 4 + (y - 4));
      if (gj >= op.J) continue;
      float* c = op.C + gi * op.ldc + gj;
      if (op.atomic) {
        atomicAdd(c, acc[x][y]);
        continue;
      }
      float v = acc[x][y];
      if (op.beta) v += *c;
      if (op.bias) v += __ldg(op.bias + gj);
      if (op.act == 1) v = fmaxf(v, 0.f);
      else if (op.act == 2) v = 1.0f / (1.0f + expf(-v));
      if (op.mask && !(op.mask[gi * op.ldc + gj] > 0.f)) v = 0.f;
      *c = v;
    }
  }
}

static int run_gemm(GemmOp op, bool a_rcontig, bool b_jcontig, int splits, cudaStream_t st) {
  if (op.I == 0 || op.J == 0) return NRF_OK;
  if (splits < 1) splits = 1;
  op.r_per_split = ((op.R + splits - 1) / splits + BK - 1) / BK * BK;
  if (op.r_per_split < BK) op.r_per_split = BK;
  splits = (int)((op.R + op.r_per_split - 1) / op.r_per_split);
  if (splits < 1) splits = 1;
  if (splits > 1 && !op.atomic) {
    set_error("internal: split-R gemm without atomic accumulation");
    return NRF_ERR_INVALID;
  }
  dim3 grid((unsigned)((op.J + BN - 1) / BN), (unsigned)((op.I + BM - 1) / BM), (unsigned)splits);
  if (grid.y > 65535) {
    set_error("fp32 MLP: more than %d rows per call unsupported", 65535 * BM);
    return NRF_ERR_INVALID;
  }
  if (a_rcontig && !b_jcontig) gemm_fp32_kernel<true, false><<<grid, 256, 0, st>>>(op);
  else if (a_rcontig && b_jcontig) gemm_fp32_kernel<true, true><<<grid, 256, 0, st>>>(op);
  else if (!a_rcontig && b_jcontig) gemm_fp32_kernel<false, true><<<grid, 256, 0, st>>>(op);
  else gemm_fp32_kernel<false, false><<<grid, 256, 0, st>>>(op);
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}

// Y[rows, n_out] = act(beta*Y + X[rows, k] · W[n_out, k]ᵀ + b)
static int linear_fwd(const float* X, int64_t ldx, int x_div, const float* W, int64_t ldw, const float* b, float* Y,
                      int64_t ldy, int64_t rows, int n_out, int k, int beta, int act, cudaStream_t st) {
  GemmOp op{};
  op.A = X; op.sa_i = ldx; op.sa_r = 1; op.a_div_i = x_div; op.a_div_r = 1;
  op.B = W; op.sb_r = 1; op.sb_j = ldw; op.b_div_r = 1;
  op.C = Y; op.ldc = ldy; op.I = rows; op.J = n_out; op.R = k;
  op.bias = b; op.beta = beta; op.act = act;
  return run_gemm(op, true, false, 1, st);
}

// dX[rows, k] = (beta*dX + dY[rows, n_out] · W[n_out, k]) ⊙ (mask > 0)
static int linear_dgrad(const float* dY, int64_t ldy, const float* W, int64_t ldw, float* dX, int64_t ldx,
                        int64_t rows, int n_out, int k, int beta, const float* mask, cudaStream_t st) {
  GemmOp op{};
  op.A = dY; op.sa_i = ldy; op.sa_r = 1; op.a_div_i = 1; op.a_div_r = 1;
  op.B = W; op.sb_r = ldw; op.sb_j = 1; op.b_div_r = 1;
  op.C = dX; op.ldc = ldx; op.I = rows; op.J = k; op.R = n_out;
  op.beta = beta; op.mask = mask;
  return run_gemm(op, true, true, 1, st);
}

// dW[n_out, k] += dY[rows, n_out]ᵀ · X[rows, k] ; db[n_out] += colsum(dY) (if db)
static int linear_wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, int x_div, float* dW, int64_t ldw,
                        float* db, int64_t rows, int n_out, int k, cudaStream_t st) {
  GemmOp op{};
  op.A = dY; op.sa_i = 1; op.sa_r = ldy; op.a_div_i = 1; op.a_div_r = 1;
  op.B = X; op.sb_r = ldx; op.sb_j = 1; op.b_div_r = x_div;
  op.C = dW; op.ldc = ldw; op.I = n_out; op.J = k; op.R = rows;
  op.a_colsum = db; op.atomic = 1;
  int splits = (int)((rows + 2047) / 2048);
  if (splits > 512) splits = 512;
  return run_gemm(op, false, true, splits, st);
}

__global__ void sigmoid_bwd_kernel(const float* __restrict__ g, const float* __restrict__ y, int64_t n,
                                   float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float yv = y[i];
    out[i] = g[i] * yv * (1.0f - yv);
  }
}

struct Fp32Plan {
  int64_t rows, n_rays, chunk_rows;
  int D, W, in_xyz, in_dir, full, need_grad;
  size_t off_enc, off_dir, off_h, off_feat, off_mid, off_da, off_db, off_dmid, off_drgb, total;
};

static inline size_t align_up(size_t v) { return (v + 255) & ~size_t(255); }

static Fp32Plan make_plan(const nrf_mlp_params* p, int64_t rows, int64_t n_rays, int sigma_only, int need_grad) {
  Fp32Plan pl{};
  pl.rows = rows; pl.n_rays = n_rays; pl.D = p->D; pl.W = p->W; pl.in_xyz = p->in_xyz; pl.in_dir = p->in_dir;
  pl.full = !sigma_only; pl.need_grad = need_grad;
  pl.chunk_rows = need_grad ? rows : (rows < (1 << 18) ? rows : (1 << 18));
  const size_t R = (size_t)pl.chunk_rows, f = sizeof(float);
  size_t o = 0;
  pl.off_enc = o; o += align_up(R * p->in_xyz * f);
  pl.off_dir = o; o += align_up((size_t)(n_rays > 0 ? n_rays : 0) * p->in_dir * f);
  const int n_h = need_grad ? p->D : 2;
  pl.off_h = o; o += (size_t)n_h * align_up(R * p->W * f);
  pl.off_feat = o; o += pl.full ? align_up(R * p->W * f) : 0;
  pl.off_mid = o; o += pl.full ? align_up(R * (p->W / 2) * f) : 0;
  if (need_grad) {
    pl.off_da = o; o += align_up(R * p->W * f);
    pl.off_db = o; o += align_up(R * p->W * f);
    pl.off_dmid = o; o += pl.full ? align_up(R * (p->W / 2) * f) : 0;
    pl.off_drgb = o; o += pl.full ? align_up(R * 3 * f) : 0;
  }
  pl.total = o;
  return pl;
}

static int check_params(const nrf_mlp_params* p, const char* who) {
  NRF_REQUIRE(p != nullptr, "%s: null params", who);
  NRF_REQUIRE(p->D >= 1 && p->D <= NRF_MAX_D && p->W >= 2 && p->in_xyz >= 1 && p->in_dir >= 0,
              "%s: bad MLP shape D=%d W=%d in_xyz=%d in_dir=%d", who, p->D, p->W, p->in_xyz, p->in_dir);
  NRF_REQUIRE((p->skip_mask & 1u) == 0, "%s: a skip connection at layer 0 is meaningless", who);
  for (int i = 0; i < p->D; ++i) NRF_REQUIRE(p->xyz_w[i] && p->xyz_b[i], "%s: null xyz layer %d", who, i);
  NRF_REQUIRE(p->sigma_w && p->sigma_b, "%s: null sigma head", who);
  return NRF_OK;
}

// One forward pass over `rows` rows whose encodings are enc (ld ld_enc) and whose direction
// encodings are dirs (row r uses dirs[(r/dir_div)*ld_dir ...]).
static int forward_rows(const nrf_mlp_params* p, const Fp32Plan& pl, char* ws, const float* enc, int64_t ld_enc,
                        const float* dirs, int64_t ld_dir, int dir_div, int64_t rows, float* sigma_out, float* rgb_out,
                        cudaStream_t st) {
  const int D = p->D, W = p->W;
  const size_t hstride = align_up((size_t)pl.chunk_rows * W * sizeof(float));
  auto H = [&](int i) { return reinterpret_cast<float*>(ws + pl.off_h + (size_t)(pl.need_grad ? i : (i & 1)) * hstride); };
  int rc;
  for (int i = 0; i < D; ++i) {
    float* out = H(i);
    if (i == 0) {
      rc = linear_fwd(enc, ld_enc, 1, p->xyz_w[0], p->in_xyz, p->xyz_b[0], out, W, rows, W, p->in_xyz, 0, 1, st);
    } else if (p->skip_mask & (1u << i)) {
      const int64_t ldw = p->in_xyz + W;
      rc = linear_fwd(enc, ld_enc, 1, p->xyz_w[i], ldw, nullptr, out, W, rows, W, p->in_xyz, 0, 0, st);
      if (rc) return rc;
      rc = linear_fwd(H(i - 1), W, 1, p->xyz_w[i] + p->in_xyz, ldw, p->xyz_b[i], out, W, rows, W, W, 1, 1, st);
    } else {
      rc = linear_fwd(H(i - 1), W, 1, p->xyz_w[i], W, p->xyz_b[i], out, W, rows, W, W, 0, 1, st);
    }
    if (rc) return rc;
  }
  const float* hl = H(D - 1);
  rc = linear_fwd(hl, W, 1, p->sigma_w, W, p->sigma_b, sigma_out, 1, rows, 1, W, 0, 0, st);
  if (rc || !pl.full) return rc;
  float* feat = reinterpret_cast<float*>(ws + pl.off_feat);
  float* mid = reinterpret_cast<float*>(ws + pl.off_mid);
  rc = linear_fwd(hl, W, 1, p->final_w, W, p->final_b, feat, W, rows, W, W, 0, 0, st);
  if (rc) return rc;
  const int64_t ldw = W + p->in_dir;
  rc = linear_fwd(feat, W, 1, p->dir_w, ldw, nullptr, mid, W / 2, rows, W / 2, W, 0, 0, st);
  if (rc) return rc;
  rc = linear_fwd(dirs, ld_dir, dir_div, p->dir_w + W, ldw, p->dir_b, mid, W / 2, rows, W / 2, p->in_dir, 1, 1, st);
  if (rc) return rc;
  return linear_fwd(mid, W / 2, 1, p->rgb_w, W / 2, p->rgb_b, rgb_out, 3, rows, 3, W / 2, 0, 2, st);
}

}  // namespace nrf

using namespace nrf;

extern "C" size_t nrf_mlp_fp32_workspace_bytes(const nrf_mlp_params* p, int64_t rows, int64_t n_rays, int sigma_only,
                                               int need_grad) {
  if (!p || rows < 0) return 0;
  return make_plan(p, rows, n_rays, sigma_only, need_grad).total + 256;
}

extern "C" int nrf_mlp_fp32_fwd_rays(const nrf_mlp_params* p, const float* rays, const float* z, int64_t n_rays,
                                     int n_samples, int xyz_freqs, int dir_freqs, int sigma_only, int need_grad,
                                     float* sigma_out, float* rgb_out, void* workspace, size_t workspace_bytes,
                                     nrf_stream_t stream) {
  int rc = check_params(p, "nrf_mlp_fp32_fwd_rays");
  if (rc) return rc;
  NRF_REQUIRE(n_rays >= 0 && n_samples >= 1, "nrf_mlp_fp32_fwd_rays: bad sizes");
  NRF_REQUIRE(p->in_xyz == 3 * (2 * xyz_freqs + 1), "nrf_mlp_fp32_fwd_rays: in_xyz=%d does not match %d frequencies",
              p->in_xyz, xyz_freqs);
  NRF_REQUIRE(sigma_only || p->in_dir == 3 * (2 * dir_freqs + 1),
              "nrf_mlp_fp32_fwd_rays: in_dir=%d does not match %d frequencies", p->in_dir, dir_freqs);
  NRF_REQUIRE(sigma_only || (p->final_w && p->dir_w && p->rgb_w && rgb_out), "nrf_mlp_fp32_fwd_rays: null rgb branch");
  if (n_rays == 0) return NRF_OK;
  NRF_REQUIRE(rays && z && sigma_out && workspace, "nrf_mlp_fp32_fwd_rays: null pointer");
  const int64_t rows = n_rays * n_samples;
  const Fp32Plan pl = make_plan(p, rows, n_rays, sigma_only, need_grad);
  if (workspace_bytes < pl.total) {
    set_error("nrf_mlp_fp32_fwd_rays: workspace %zu < %zu bytes", workspace_bytes, pl.total);
    return NRF_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  char* ws = static_cast<char*>(workspace);
  float* enc = reinterpret_cast<float*>(ws + pl.off_enc);
  float* dirs = reinterpret_cast<float*>(ws + pl.off_dir);
  if (!sigma_only) {
    rc = launch_encode_dirs(rays, n_rays, dir_freqs, dirs, st);
    if (rc) return rc;
  }
  int64_t chunk_rays = pl.chunk_rows / n_samples;
  if (chunk_rays < 1) chunk_rays = 1;
  NRF_REQUIRE(!need_grad || chunk_rays >= n_rays, "internal: need_grad forward must be a single chunk");
  for (int64_t r0 = 0; r0 < n_rays; r0 += chunk_rays) {
    const int64_t nr = (n_rays - r0 < chunk_rays) ? (n_rays - r0) : chunk_rays;
    NRF_REQUIRE(nr * n_samples <= pl.chunk_rows || pl.chunk_rows >= n_samples, "internal: chunking");
    if (nr * n_samples > pl.chunk_rows) {
      set_error("nrf_mlp_fp32_fwd_rays: n_samples=%d exceeds the internal chunk", n_samples);
      return NRF_ERR_UNSUPPORTED;
    }
    rc = launch_encode_samples(rays + r0 * 8, z + r0 * n_samples, nr, n_samples, xyz_freqs, enc, p->in_xyz, st);
    if (rc) return rc;
    rc = forward_rows(p, pl, ws, enc, p->in_xyz, dirs + r0 * p->in_dir, p->in_dir, n_samples, nr * n_samples,
                      sigma_out + r0 * n_samples, rgb_out ? rgb_out + r0 * n_samples * 3 : nullptr, st);
    if (rc) return rc;
  }
  return NRF_OK;
}

extern "C" int nrf_mlp_fp32_fwd_embedded(const nrf_mlp_params* p, const float* x, int64_t rows, int sigma_only,
                                         int need_grad, float* sigma_out, float* rgb_out, void* workspace,
                                         size_t workspace_bytes, nrf_stream_t stream) {
  int rc = check_params(p, "nrf_mlp_fp32_fwd_embedded");
  if (rc) return rc;
  NRF_REQUIRE(rows >= 0, "nrf_mlp_fp32_fwd_embedded: bad sizes");
  NRF_REQUIRE(sigma_only || (p->final_w && p->dir_w && p->rgb_w && rgb_out), "nrf_mlp_fp32_fwd_embedded: null rgb branch");
  if (rows == 0) return NRF_OK;
  NRF_REQUIRE(x && sigma_out && workspace, "nrf_mlp_fp32_fwd_embedded: null pointer");
  const Fp32Plan pl = make_plan(p, rows, 0, sigma_only, need_grad);
  if (workspace_bytes < pl.total) {
    set_error("nrf_mlp_fp32_fwd_embedded: workspace %zu < %zu bytes", workspace_bytes, pl.total);
    return NRF_ERR_WORKSPACE;
  }
  const int64_t ldx = p->in_xyz + (sigma_only ? 0 : p->in_dir);
  char* ws = static_cast<char*>(workspace);
  for (int64_t r0 = 0; r0 < rows; r0 += pl.chunk_rows) {
    const int64_t nr = (rows - r0 < pl.chunk_rows) ? (rows - r0) : pl.chunk_rows;
    const float* xc = x + r0 * ldx;
    rc = forward_rows(p, pl, ws, xc, ldx, xc + p->in_xyz, ldx, 1, nr, sigma_out + r0,
                      rgb_out ? rgb_out + r0 * 3 : nullptr, as_stream(stream));
    if (rc) return rc;
  }
  return NRF_OK;
}

extern "C" int nrf_mlp_fp32_bwd(const nrf_mlp_params* p, const nrf_mlp_grads* g, const float* x, int64_t rows,
                                int64_t n_rays, int n_samples, int sigma_only, const float* d_sigma,
                                const float* d_rgb, const float* rgb_out, void* workspace, size_t workspace_bytes,
                                nrf_stream_t stream) {
  int rc = check_params(p, "nrf_mlp_fp32_bwd");
  if (rc) return rc;
  NRF_REQUIRE(g != nullptr && rows >= 0, "nrf_mlp_fp32_bwd: bad arguments");
  if (rows == 0) return NRF_OK;
  NRF_REQUIRE(d_sigma && workspace, "nrf_mlp_fp32_bwd: null pointer");
  NRF_REQUIRE(sigma_only || (d_rgb && rgb_out), "nrf_mlp_fp32_bwd: full mode needs d_rgb and rgb_out");
  const bool rays_form = (x == nullptr);
  NRF_REQUIRE(!rays_form || (n_rays > 0 && n_rays * (int64_t)n_samples == rows), "nrf_mlp_fp32_bwd: rows != n_rays*n_samples");
  const Fp32Plan pl = make_plan(p, rows, rays_form ? n_rays : 0, sigma_only, 1);
  if (workspace_bytes < pl.total) {
    set_error("nrf_mlp_fp32_bwd: workspace %zu < %zu bytes", workspace_bytes, pl.total);
    return NRF_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  char* ws = static_cast<char*>(workspace);
  const int D = p->D, W = p->W;
  const size_t hstride = align_up((size_t)pl.chunk_rows * W * sizeof(float));
  auto H = [&](int i) { return reinterpret_cast<float*>(ws + pl.off_h + (size_t)i * hstride); };
  const float* enc; int64_t ld_enc; const float* dirs; int64_t ld_dir; int dir_div;
  if (rays_form) {
    enc = reinterpret_cast<float*>(ws + pl.off_enc); ld_enc = p->in_xyz;
    dirs = reinterpret_cast<float*>(ws + pl.off_dir); ld_dir = p->in_dir; dir_div = n_samples;
  } else {
    ld_enc = p->in_xyz + (sigma_only ? 0 : p->in_dir);
    enc = x; dirs = x + p->in_xyz; ld_dir = ld_enc; dir_div = 1;
  }
  float* da = reinterpret_cast<float*>(ws + pl.off_da);
  float* db = reinterpret_cast<float*>(ws + pl.off_db);
  const float* hl = H(D - 1);
  float* cur = da;   // gradient w.r.t. the pre-activation of layer D-1 ends up here
  if (!sigma_only) {
    float* feat = reinterpret_cast<float*>(ws + pl.off_feat);
    float* mid = reinterpret_cast<float*>(ws + pl.off_mid);
    float* dmid = reinterpret_cast<float*>(ws + pl.off_dmid);
    float* drgb = reinterpret_cast<float*>(ws + pl.off_drgb);
    sigmoid_bwd_kernel<<<kNumSMs * 4, 256, 0, st>>>(d_rgb, rgb_out, rows * 3, drgb);
    NRF_LAUNCH_CHECK();
    if ((rc = linear_wgrad(drgb, 3, mid, W / 2, 1, g->rgb_w, W / 2, g->rgb_b, rows, 3, W / 2, st))) return rc;
    if ((rc = linear_dgrad(drgb, 3, p->rgb_w, W / 2, dmid, W / 2, rows, 3, W / 2, 0, mid, st))) return rc;
    const int64_t ldw = W + p->in_dir;
    if ((rc = linear_wgrad(dmid, W / 2, feat, W, 1, g->dir_w, ldw, g->dir_b, rows, W / 2, W, st))) return rc;
    if ((rc = linear_wgrad(dmid, W / 2, dirs, ld_dir, dir_div, g->dir_w + W, ldw, nullptr, rows, W / 2, p->in_dir, st))) return rc;
    if ((rc = linear_dgrad(dmid, W / 2, p->dir_w, ldw, db, W, rows, W / 2, W, 0, nullptr, st))) return rc;  // d feat
    if ((rc = linear_wgrad(db, W, hl, W, 1, g->final_w, W, g->final_b, rows, W, W, st))) return rc;
    if ((rc = linear_dgrad(db, W, p->final_w, W, da, W, rows, W, W, 0, nullptr, st))) return rc;           // d h_last (partial)
  }
  if ((rc = linear_wgrad(d_sigma, 1, hl, W, 1, g->sigma_w, W, g->sigma_b, rows, 1, W, st))) return rc;
  if ((rc = linear_dgrad(d_sigma, 1, p->sigma_w, W, da, W, rows, 1, W, sigma_only ? 0 : 1, hl, st))) return rc;
  float* other = db;
  for (int i = D - 1; i >= 0; --i) {
    const bool skip = (p->skip_mask >> i) & 1u;
    if (i == 0) {
      rc = linear_wgrad(cur, W, enc, ld_enc, 1, g->xyz_w[0], p->in_xyz, g->xyz_b[0], rows, W, p->in_xyz, st);
      if (rc) return rc;
      break;
    }
    const int64_t ldw = skip ? (p->in_xyz + W) : W;
    const int off = skip ? p->in_xyz : 0;
    if (skip) {
      if ((rc = linear_wgrad(cur, W, enc, ld_enc, 1, g->xyz_w[i], ldw, nullptr, rows, W, p->in_xyz, st))) return rc;
    }
    if ((rc = linear_wgrad(cur, W, H(i - 1), W, 1, g->xyz_w[i] + off, ldw, g->xyz_b[i], rows, W, W, st))) return rc;
    if ((rc = linear_dgrad(cur, W, p->xyz_w[i] + off, ldw, other, W, rows, W, W, 0, H(i - 1), st))) return rc;
    float* t = cur; cur = other; other = t;
  }
  return NRF_OK;
}
