"""ctypes binding of libnerf_b200.so (include/nerf_b200.h).  No fallback: if the library is
missing or a call fails, a RuntimeError is raised."""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.environ.get("NRF_LIBDIR") or os.path.join(_HERE, "lib"), "libnerf_b200.so")
NRF_MAX_D = 16

ERR_NAMES = {1: "invalid argument", 2: "CUDA error", 3: "unsupported", 4: "workspace too small"}


class MlpParams(Structure):
    _fields_ = [("D", c_int32), ("W", c_int32), ("in_xyz", c_int32), ("in_dir", c_int32), ("skip_mask", c_uint32),
                ("xyz_w", c_void_p * NRF_MAX_D), ("xyz_b", c_void_p * NRF_MAX_D),
                ("final_w", c_void_p), ("final_b", c_void_p), ("dir_w", c_void_p), ("dir_b", c_void_p),
                ("sigma_w", c_void_p), ("sigma_b", c_void_p), ("rgb_w", c_void_p), ("rgb_b", c_void_p)]


class MlpGrads(Structure):
    _fields_ = [("xyz_w", c_void_p * NRF_MAX_D), ("xyz_b", c_void_p * NRF_MAX_D),
                ("final_w", c_void_p), ("final_b", c_void_p), ("dir_w", c_void_p), ("dir_b", c_void_p),
                ("sigma_w", c_void_p), ("sigma_b", c_void_p), ("rgb_w", c_void_p), ("rgb_b", c_void_p)]


P, I, L, F, Z, D = c_void_p, c_int, c_int64, c_float, c_size_t, c_double
PP, PG = POINTER(MlpParams), POINTER(MlpGrads)

# name -> (restype, argtypes); mirrors include/nerf_b200.h declaration by declaration
SIGNATURES = {
    "nrf_abi_version": (I, []),
    "nrf_last_error": (c_char_p, []),
    "nrf_check_device": (I, []),
    "nrf_launch_count": (ctypes.c_uint64, []),
    "nrf_prof_enable": (I, [I]),
    "nrf_prof_read": (I, [I, P, P, P]),
    "nrf_gen_rays": (I, [P, I, I, F, P, L, F, F, I, P, P]),
    "nrf_mse_workspace_bytes": (Z, []),
    "nrf_mse_pair": (I, [P, P, P, P, L, P, P, P]),
    "nrf_mse_pair_bwd": (I, [P, P, P, L, P, P, P, P]),
    "nrf_adam_step": (I, [P, P, P, P, L, D, D, D, D, D, L, P]),
    "nrf_adam_state_bytes": (Z, []),
    "nrf_adam_step_dev": (I, [P, P, P, P, L, D, D, D, D, P, P]),
    "nrf_sgd_step": (I, [P, P, P, L, D, D, D, I, P]),
    "nrf_radam_step": (I, [P, P, P, P, P, L, D, D, D, D, D, L, I, D, I, D, I, P]),
    "nrf_sample_coarse": (I, [P, P, L, I, I, F, P, P, P]),
    "nrf_embed_fwd": (I, [P, L, I, P, I, P, P]),
    "nrf_embed_bwd": (I, [P, P, L, I, P, I, P, P]),
    "nrf_composite_fwd": (I, [P, P, P, P, P, F, I, L, I, P, P, P, P, P, P]),
    "nrf_composite_bwd": (I, [P, P, P, P, P, F, I, L, I, P, P, P, P, P, P, P, P]),
    "nrf_sample_pdf": (I, [P, P, L, I, I, P, P, F, P, P, P, P]),
    "nrf_merge_sort": (I, [P, P, L, I, I, P, P]),
    "nrf_searchsorted": (I, [P, P, P, L, L, L, L, I, P]),
    "nrf_searchsorted_typed": (I, [P, P, P, L, L, L, L, I, I, P]),
    "nrf_mlp_fp32_workspace_bytes": (Z, [PP, L, L, I, I]),
    "nrf_mlp_fp32_fwd_rays": (I, [PP, P, P, L, I, I, I, I, I, P, P, P, Z, P]),
    "nrf_mlp_fp32_fwd_embedded": (I, [PP, P, L, I, I, P, P, P, Z, P]),
    "nrf_mlp_fp32_bwd": (I, [PP, PG, P, L, L, I, I, P, P, P, P, Z, P]),
    "nrf_mlp_tc_packed_bytes": (Z, []),
    "nrf_mlp_tc_pack": (I, [PP, P, P]),
    "nrf_mlp_tc_stash_bytes": (Z, [L, I]),
    "nrf_mlp_tc_fwd": (I, [P, P, P, L, I, I, P, P, P, Z, P]),
    "nrf_mlp_tc_bwd_workspace_bytes": (Z, [L, I]),
    "nrf_mlp_tc_bwd": (I, [P, PG, P, P, L, I, I, P, P, P, P, Z, P, Z, P]),
    "nrf_shadow_workspace_bytes": (Z, [L, L]),
    "nrf_shadow_fwd": (I, [P, P, P, P, P, P, P, P, P, L, I, I, I, F, F, F, F, I, F, P, P, Z, P]),
    "nrf_shadow_fwd_static": (I, [P, P, P, P, P, P, P, P, P, L, I, I, I, F, F, F, F, I, F, P, P]),
    "nrf_normed_w": (I, [P, P, P, P, L, P, P]),
    "nrf_project_to_light": (I, [P, P, P, P, L, P, P, L, I, P, P, P, P]),
    "nrf_shadow_map_workspace_bytes": (Z, []),
    "nrf_shadow_map_fwd": (I, [P, P, L, I, F, F, F, F, I, I, P, P, P]),
    "nrf_shadow_map_bwd": (I, [P, P, L, I, F, F, F, F, I, I, P, P, P, P]),
    "nrf_light_gather": (I, [P, P, L, I, I, P, P, P]),
    "nrf_light_scatter": (I, [P, P, L, L, P, P]),
    "nrf_w_mapping": (I, [P, P, P, L, F, P, P]),
    "nrf_shadow_bwd": (I, [P, P, P, P, P, P, P, P, P, L, I, I, I, F, F, F, F, I, P, P, P, P, Z, P]),
}

class PeerComm(ctypes.Structure):
    """nrf_peer_comm (include/nerf_b200.h)."""
    _fields_ = [("bufs", ctypes.c_void_p * 8), ("flags", ctypes.c_void_p * 8), ("rank", ctypes.c_int32),
                ("world", ctypes.c_int32), ("n", ctypes.c_int64), ("seq", ctypes.c_void_p), ("counter", ctypes.c_void_p)]


SIGNATURES.update({
    "nrf_peer_alloc": (I, [Z, ctypes.POINTER(ctypes.c_void_p)]),
    "nrf_peer_free": (I, [P]),
    "nrf_peer_export": (I, [P, P]),
    "nrf_peer_import": (I, [P, ctypes.POINTER(ctypes.c_void_p)]),
    "nrf_peer_close": (I, [P]),
    "nrf_peer_allreduce_mean": (I, [ctypes.POINTER(PeerComm), P]),
})

# include/nerf_b200_debug.h: present only in a library built with NRF_DEBUG_BUILD=1 (tools/ microbenchmarks)
DEBUG_SIGNATURES = {
    "nrf_dbg_set_counters": (None, [P]),
    "nrf_dbg_mma_rate": (I, [I, I, P, P]),
    "nrf_dbg_mma_rate_n": (I, [I, I, I, P, P]),
    "nrf_dbg_tmem_rate": (I, [I, I, I, I, P, P]),
    "nrf_dbg_mma_contend": (I, [I, I, I, P, P, P]),
    "nrf_dbg_store_rate": (I, [P, ctypes.c_uint64, c_uint32, I, I, I, P, P]),
    "nrf_dbg_stream_replicas": (I, [I]),
    "nrf_dbg_stream": (I, [P, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, c_uint32, I, I, I, I, I, I, P]),
}

_lib = None


def load():
    """Load the shared library (once) and declare every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m nerf_pl_b200._build` "
            "(or __graft_entry__.build()).  nerf_pl_b200 has no CPU / PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    for name, (res, args) in DEBUG_SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype = res
            fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().nrf_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"nerf_b200 {what}: {ERR_NAMES.get(rc, rc)}: {msg}")


def call(name, *args):
    fn = getattr(load(), name, None)
    if fn is None:
        raise RuntimeError(f"{name} is not exported by {LIB_PATH}"
                           + (": debug entry points need a library built with NRF_DEBUG_BUILD=1" if name in DEBUG_SIGNATURES else ""))
    check(fn(*args), name)
