// Placeholder until the tcgen05 path lands (replaced in a later commit).
#include "common.cuh"
using namespace nrf;
extern "C" size_t nrf_mlp_tc_packed_bytes(void) { return 0; }
extern "C" int nrf_mlp_tc_pack(const nrf_mlp_params*, void*, nrf_stream_t) { set_error("tensor-core MLP not built"); return NRF_ERR_UNSUPPORTED; }
extern "C" size_t nrf_mlp_tc_stash_bytes(int64_t, int) { return 0; }
extern "C" int nrf_mlp_tc_fwd(const void*, const float*, const float*, int64_t, int, int, float*, float*, void*, size_t, nrf_stream_t) { set_error("tensor-core MLP not built"); return NRF_ERR_UNSUPPORTED; }
extern "C" size_t nrf_mlp_tc_bwd_workspace_bytes(int64_t, int) { return 0; }
extern "C" int nrf_mlp_tc_bwd(const void*, const nrf_mlp_grads*, const float*, const float*, int64_t, int, int, const float*, const float*, const float*, const void*, size_t, void*, size_t, nrf_stream_t) { set_error("tensor-core MLP not built"); return NRF_ERR_UNSUPPORTED; }
