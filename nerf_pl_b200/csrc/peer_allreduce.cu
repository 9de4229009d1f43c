// Gradient exchange of the ray-sharded training step over NVLink peer memory (the reference gets it from
// DistributedDataParallel's NCCL all-reduce, train.py:174-175): a two-shot all-reduce (reduce-scatter + all-gather) of the
// flat fp32 gradient buffer written directly against the peers' buffers - every rank loads its slice from all W buffers
// (ld over NVLink from the peers' mapped memory), averages it in a FIXED rank order (so every rank ends up with
// bit-identical gradients, which the flat Adam step then turns into bit-identical weights), and stores the mean back
// into all W buffers.  One launch, no NCCL channel set-up or proxy thread on the path, 2 x 4.2 MB per rank through the
// switch at W = 8: the 4.77 MB exchange is latency-bound, so the point is the two flag barriers (release / acquire at
// system scope on words in the peers' memory) instead of a collective's launch + protocol latency.
//
// Buffers are plain cudaMalloc allocations shared between the ranks' processes with CUDA IPC (nrf_peer_export / import);
// the host side (nerf_pl_b200/dist.py: GradBucket) exchanges the handles through torch.distributed and falls back to
// NCCL when the ranks are not on one node.  Capturable in a CUDA graph: the round number lives in device memory.
#include <cstring>
#include "common.cuh"

namespace nrf {

constexpr int kPeerMax = NRF_PEER_MAX;

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peers' data: never through this SM's L1
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// wait until every rank's word in MY flag array has reached `ticket` (threads 0..world-1 poll one word each); bounded
__device__ __forceinline__ void wait_all(const unsigned* my_flags, int world, unsigned ticket) {
  if ((int)threadIdx.x < world) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(my_flags + threadIdx.x) - ticket) < 0) {
      if (clock64() - t0 > 20000000000ll) __trap();     // ~10 s: a rank never arrived; fail loudly instead of hanging
    }
  }
  __syncthreads();
}

template <int kWorld>
__global__ void __launch_bounds__(256) peer_allreduce_mean_kernel(const nrf_peer_comm c) {
  __shared__ unsigned s_round;
  __shared__ int s_last;
  const int tid = threadIdx.x;
  const unsigned* my_flags = static_cast<const unsigned*>(c.flags[c.rank]);
  if (tid == 0) s_round = *static_cast<volatile unsigned*>(c.seq);
  __syncthreads();
  const unsigned round = s_round, t_in = 2 * round + 1, t_out = 2 * round + 2;

  // ---- entry barrier: every rank's backward has finished (stream order) and nobody reads stale gradients ----
  if (blockIdx.x == 0 && tid < kWorld) st_release_sys(static_cast<unsigned*>(c.flags[tid]) + c.rank, t_in);
  wait_all(my_flags, kWorld, t_in);

  // ---- my slice: sum over the ranks in rank order, mean, broadcast ----
  const int64_t per = (((c.n + kWorld - 1) / kWorld) + 3) & ~int64_t(3);
  const int64_t lo = (int64_t)c.rank * per, hi = lo + per < c.n ? lo + per : c.n;
  const float inv = 1.0f / (float)kWorld;
  float* bufs[kWorld];
#pragma unroll
  for (int q = 0; q < kWorld; ++q) bufs[q] = static_cast<float*>(c.bufs[q]);
  // two float4 per thread and pass: 2 x kWorld independent 16-byte loads in flight per thread hide the NVLink round trip
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
  for (int64_t i = lo + ((int64_t)blockIdx.x * blockDim.x + tid) * 4; i < hi; i += 2 * stride) {
    const int64_t i2 = i + stride;
    const bool full1 = i + 4 <= hi, full2 = i2 + 4 <= hi;
    if (full1) {
      float4 v[kWorld], w[kWorld];
#pragma unroll
      for (int q = 0; q < kWorld; ++q) v[q] = ld_peer_f4(bufs[q] + i);
      if (full2) {
#pragma unroll
        for (int q = 0; q < kWorld; ++q) w[q] = ld_peer_f4(bufs[q] + i2);
      }
      float4 s = v[0];
#pragma unroll
      for (int q = 1; q < kWorld; ++q) { s.x += v[q].x; s.y += v[q].y; s.z += v[q].z; s.w += v[q].w; }
      s.x *= inv; s.y *= inv; s.z *= inv; s.w *= inv;
#pragma unroll
      for (int q = 0; q < kWorld; ++q) st_peer_f4(bufs[q] + i, s);
      if (full2) {
        float4 t = w[0];
#pragma unroll
        for (int q = 1; q < kWorld; ++q) { t.x += w[q].x; t.y += w[q].y; t.z += w[q].z; t.w += w[q].w; }
        t.x *= inv; t.y *= inv; t.z *= inv; t.w *= inv;
#pragma unroll
        for (int q = 0; q < kWorld; ++q) st_peer_f4(bufs[q] + i2, t);
      }
    }
    // ragged tails (n not a multiple of 4): element by element
    for (int pass = 0; pass < 2; ++pass) {
      const int64_t b = pass ? i2 : i;
      if ((pass ? full2 : full1) || b >= hi) continue;
      for (int64_t j = b; j < hi; ++j) {
        float s = 0.f;
        for (int q = 0; q < kWorld; ++q) s += *reinterpret_cast<volatile float*>(bufs[q] + j);
        s *= inv;
        for (int q = 0; q < kWorld; ++q) *reinterpret_cast<volatile float*>(bufs[q] + j) = s;
      }
    }
  }

  // ---- exit barrier: my slice has landed in every buffer; leave once every rank's slice has landed in mine ----
  __threadfence_system();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(static_cast<int*>(c.counter), 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (s_last) {                                   // the last CTA of this rank: all local stores are fenced
    if (tid == 0) {
      *static_cast<volatile int*>(c.counter) = 0;
      *static_cast<volatile unsigned*>(c.seq) = round + 1;
      __threadfence();
    }
    __syncthreads();
    if (tid < kWorld) st_release_sys(static_cast<unsigned*>(c.flags[tid]) + c.rank, t_out);
  }
  wait_all(my_flags, kWorld, t_out);
}

}  // namespace nrf

using namespace nrf;

extern "C" int nrf_peer_alloc(size_t bytes, void** ptr) {
  NRF_REQUIRE(ptr && bytes > 0, "nrf_peer_alloc: bad arguments");
  NRF_CUDA(cudaMalloc(ptr, bytes));
  NRF_CUDA(cudaMemset(*ptr, 0, bytes));
  return NRF_OK;
}
extern "C" int nrf_peer_free(void* ptr) {
  if (ptr) NRF_CUDA(cudaFree(ptr));
  return NRF_OK;
}
extern "C" int nrf_peer_export(const void* ptr, void* handle64) {
  NRF_REQUIRE(ptr && handle64, "nrf_peer_export: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  NRF_CUDA(cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle64), const_cast<void*>(ptr)));
  return NRF_OK;
}
extern "C" int nrf_peer_import(const void* handle64, void** ptr) {
  NRF_REQUIRE(ptr && handle64, "nrf_peer_import: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  NRF_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return NRF_OK;
}
extern "C" int nrf_peer_close(void* ptr) {
  if (ptr) NRF_CUDA(cudaIpcCloseMemHandle(ptr));
  return NRF_OK;
}

extern "C" int nrf_peer_allreduce_mean(const nrf_peer_comm* c, nrf_stream_t stream) {
  NRF_REQUIRE(c && c->world >= 2 && c->world <= kPeerMax && c->rank >= 0 && c->rank < c->world && c->n > 0 && c->seq && c->counter,
              "nrf_peer_allreduce_mean: bad communicator");
  for (int q = 0; q < c->world; ++q) NRF_REQUIRE(c->bufs[q] && c->flags[q], "nrf_peer_allreduce_mean: rank %d is not mapped", q);
  cudaStream_t st = as_stream(stream);
  const int64_t per = (c->n + c->world - 1) / c->world;
  int grid = (int)((per / 8 + 255) / 256);              // two float4 per thread
  grid = grid < 1 ? 1 : (grid > 128 ? 128 : grid);     // every CTA must be resident: the exit signal is sent by the last one
  switch (c->world) {
#define NRF_PEER_CASE(W) case W: peer_allreduce_mean_kernel<W><<<grid, 256, 0, st>>>(*c); break;
    NRF_PEER_CASE(2) NRF_PEER_CASE(3) NRF_PEER_CASE(4) NRF_PEER_CASE(5) NRF_PEER_CASE(6) NRF_PEER_CASE(7) NRF_PEER_CASE(8)
#undef NRF_PEER_CASE
  }
  NRF_LAUNCH_CHECK();
  return NRF_OK;
}
