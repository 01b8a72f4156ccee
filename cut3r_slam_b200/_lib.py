"""ctypes binding of libdba.so (include/dba.h).  No compute happens on this side.

The library must have been built (`python -m cut3r_slam_b200.build`); a missing library is a hard
error -- there is no CPU or PyTorch fallback for any op.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DBA_LIB_PATH") or os.path.join(HERE, "libdba.so")     # override: developer builds (tracing)

_lib = None

c_f = C.c_void_p      # device/host float*   (raw addresses from tensor.data_ptr())
c_i = C.c_void_p      # int64*

SIGNATURES = {
    "dba_create": (C.c_void_p, [C.c_int]),
    "dba_destroy": (None, [C.c_void_p]),
    "dba_last_error": (C.c_char_p, [C.c_void_p]),
    "dba_version": (C.c_char_p, []),
    "dba_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_double]),
    "dba_ba": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int,
                         C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, c_f, c_f,
                         C.c_void_p]),
    "dba_ba_cached": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int,
                                C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, c_f, c_f,
                                C.c_void_p, C.c_uint64, c_i, c_i]),
    "dba_ba_host": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int,
                              C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, c_f, c_f,
                              C.c_void_p]),
    "dba_frame_distance": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                     C.c_float, c_f, C.c_void_p]),
    "dba_covis_distance": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int, c_f,
                                     C.c_void_p]),
    "dba_iproj": (C.c_int, [C.c_void_p, c_f, c_f, c_f, C.c_int64, C.c_int, C.c_int, c_f, C.c_void_p]),
    "dba_depth_filter": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_i, c_f, C.c_int64, C.c_int64, C.c_int, C.c_int, c_f,
                                   C.c_void_p]),
    "dba_proj_trans": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                 C.c_int64, c_f, c_f, C.c_void_p]),
    "dba_bi_inter": (C.c_int, [C.c_void_p, c_f, c_f, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, c_f, c_f,
                               C.c_void_p]),
    "dba_jdsa": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int,
                           C.c_int, C.c_int, C.c_int, C.c_int64, C.c_float, C.c_float, C.c_float, c_f, C.c_void_p]),
    "dba_pgba": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, c_i, c_i, c_f, c_f, C.c_int64, C.c_int64,
                           C.c_int64, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float,
                           c_f, c_f, C.c_void_p]),
    "dba_depth_cov": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int,
                                C.c_int, C.c_int64, C.c_int, C.c_int, C.c_float, C.c_float, c_f, C.c_void_p]),
    "dba_rel_pose_cov": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                   c_f, c_f, c_f, C.c_void_p]),
    "dba_reproject": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_i, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int, c_f, c_f,
                                C.c_void_p]),
    "dba_motion_features": (C.c_int, [C.c_void_p, c_f, c_f, c_f, c_i, c_i, c_f, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                      c_f, c_f, c_f, C.c_void_p]),
    "dba_damping_gather": (C.c_int, [C.c_void_p, c_f, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int64, c_f,
                                     C.c_void_p]),
    "dba_damping_gather_noalloc": (C.c_int, [C.c_void_p, c_f, c_i, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int64, c_f,
                                             C.c_void_p]),
    "dba_cvx_upsample": (C.c_int, [C.c_void_p, c_f, c_f, C.c_int64, C.c_int, C.c_int, C.c_int, c_f, C.c_void_p]),
    "dba_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "dba_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "dba_comm_destroy": (C.c_int, [C.c_void_p]),
    "dba_comm_p2p_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "dba_comm_p2p_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "dba_debug_read": (C.c_int64, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]),
    "dba_last_launch_count": (C.c_int64, [C.c_void_p]),
    "dba_plan_build": (C.c_void_p, [c_i, c_i, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int,
                                    C.c_char_p, C.c_int]),
    "dba_plan_free": (None, [C.c_void_p]),
    "dba_graph_neighbourhood": (C.c_int64, [C.c_int64, C.c_int64, C.c_int64, c_i, c_i, C.c_int64]),
    "dba_graph_filter_repeated": (C.c_int64, [c_i, c_i, C.c_int64, c_i, c_i, C.c_int64, C.c_void_p]),
    "dba_graph_evict_mask": (C.c_int64, [c_i, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "dba_graph_rm_keyframe": (C.c_int64, [c_i, c_i, c_i, C.c_int64, C.c_int64]),
    "dba_graph_proximity": (C.c_int64, [c_f, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_float,
                                        C.c_int64, c_i, c_i, C.c_int64, c_i, c_i, C.c_int64]),
    "dba_plan_get": (C.c_int64, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]),
}


def load():
    """Load libdba.so and declare every entry point of include/dba.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m cut3r_slam_b200.build` "
            "(the DBA ops are CUDA-only; there is no fallback)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
