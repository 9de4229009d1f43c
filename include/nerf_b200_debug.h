/* nerf_b200_debug.h - entry points that exist only in a debug build of libnerf_b200.so
 * (NRF_DEBUG_BUILD=1 python -m nerf_pl_b200._build --force): microbenchmarks that chose the stash layout and the
 * weight-ring geometry, and the cycle counters of the forward kernels.  The product library exports none of them and
 * reads no environment variable.  Used by tools/{stream_bench,mma_rate,store_rate,fwd_counters,fwd_trace,fwd_deadlock,
 * pair_check,pair_counters,wgrad_probe}.py. */
#ifndef NERF_B200_DEBUG_H_
#define NERF_B200_DEBUG_H_
#include "nerf_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Diagnostics (not on the product path): lane-parallel cp.async.bulk streaming benchmark used to choose
 * the stash layout and the weight-ring geometry (tools/stream_bench.py).  Moves
 * grid * steps * n_streams * piece bytes from `base` (+ strides) into shared memory. */
/* Per-CTA cycle counters of the tensor-core forward kernel (8 x uint64 per CTA: [0] MMA-thread total, [1] waiting
 * for A operands, [2] waiting for weights, [4] epilogue total, [5] epilogue waiting for accumulators); NULL = off. */
void nrf_dbg_set_counters(void* dev_ptr);
int nrf_dbg_stream_replicas(int replicas);   /* > 0: the next nrf_dbg_stream launches read replica cta % replicas at base + replica * cta_stride */
int nrf_dbg_stream(const void* base, unsigned long long cta_stride, unsigned long long stream_stride,
                   unsigned long long tile_stride, uint32_t piece, int n_streams, int ppc, int wrap_tiles,
                   int steps, int stages, int grid, nrf_stream_t stream);
/* Measurement aid: `grid` CTAs each issue `reps` tcgen05.mma (M=128, N=256, K=16, bf16) back to back; out[cta] (u64,
 * device) = cycles from the first issue to the completion of the last. */
int nrf_dbg_mma_rate(int reps, int grid, void* out, nrf_stream_t stream);
int nrf_dbg_mma_rate_n(int reps, int n_cols, int grid, void* out, nrf_stream_t stream);   /* the same with N = n_cols */
/* Measurement aid: `grid` CTAs each bulk-store steps x piece bytes from shared memory to dst + cta * cta_stride with
 * `depth` (1, 2 or 4) stores in flight; out[cta] (u64, device) = cycles. */
int nrf_dbg_store_rate(void* dst, unsigned long long cta_stride, uint32_t piece, int steps, int depth, int grid, void* out,
                       nrf_stream_t stream);

/* Measurement aid: `warps` (4 / 8 / 16) warps per CTA read the 512 TMEM columns of their lane quarter `reps` times with
 * tcgen05.ld 32x32b .x16 (mode bit 0: .x32); mode bit 1: one thread issues back-to-back MMAs meanwhile.
 * out[2 cta] (u64, device) = cycles of the slowest warp, out[2 cta + 1] = MMAs retired meanwhile. */
int nrf_dbg_tmem_rate(int warps, int reps, int mode, int grid, void* out, nrf_stream_t stream);
/* MMA rate (M128 N256 K16, `reps` back to back) under shared-memory traffic.  mode bit 0: A operand from tensor memory;
 * bit 1: a 4 x 32 KB ring of bulk copies from `src` (>= 1 MB, L2 resident) keeps landing; bit 2: 16 warps keep writing a
 * 64 KB image with 16-byte shared stores; bit 7: one thread keeps 32 KB bulk stores READING the A tile in flight; bits 3 / 4:
 * N = 128 / 64; bit 5: 16 warps of tcgen05.ld / st; bit 6: 16 arithmetic-bound warps.  src >= 1 MB + grid x 64 KB.
 * out: grid x 3 u64 = {cycles of the MMAs, 32 KB copies landed, 32 KB stores issued meanwhile}. */
int nrf_dbg_mma_contend(int reps, int mode, int grid, const void* src, void* out, nrf_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_DEBUG_H_ */
