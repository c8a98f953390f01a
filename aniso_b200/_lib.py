"""ctypes binding of libaniso_b200.so (include/aniso_b200.h).  The library is built in-tree by
aniso_b200/csrc/Makefile (see __graft_entry__.build); there is no Python or CPU fallback: if the
shared library is missing, importing a compute entry point raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaniso_b200.so")

_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_H = C.c_void_p
_V = C.c_void_p  # raw device pointer

# host preconditioner callback of aniso_gmres_prec / aniso_block_gmres_prec: int (*)(void* ctx, int n, const double* r, double* z)
PREC_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double))

# host kernel callback of aniso_fmm_cache: double (*)(void* ctx, double sx, double sy, double tx, double ty)
EVAL_FN = C.CFUNCTYPE(C.c_double, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double)

# name -> (restype, argtypes): exactly the symbols include/aniso_b200.h declares
SIGNATURES = {
    "aniso_last_error": (C.c_char_p, []),
    "aniso_version": (C.c_char_p, []),
    "aniso_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(_H)]),
    "aniso_create_sharded": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_void_p, C.c_int, C.POINTER(_H)]),
    "aniso_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "aniso_create_multi": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, _ip, C.POINTER(_H)]),
    "aniso_destroy": (C.c_int, [_H]),
    "aniso_num_nodes": (C.c_int, [_H]),
    "aniso_kernel_size": (C.c_int, [_H]),
    "aniso_get_nodes": (C.c_int, [_H, _dp, _dp, _dp]),
    "aniso_gl_table": (C.c_int, [C.c_int, _dp, _dp]),
    "aniso_geometry_tables": (C.c_int, [_H, _dp, _dp, _dp, _dp, _dp, _dp]),
    "aniso_sing_tables": (C.c_int, [_H, _dp, _dp, _dp]),
    "aniso_tree_stats": (C.c_int, [_H, _lp]),
    "aniso_tree_nodes": (C.c_int, [_H, C.c_int, _dp, _ip]),
    "aniso_node_lists": (C.c_int, [_H, C.c_int, C.c_int, _ip, _ip]),
    "aniso_slot_permutation": (C.c_int, [_H, _ip]),
    "aniso_partition": (C.c_int, [_H, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "aniso_owned_range": (C.c_int, [_H, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "aniso_set_coeff": (C.c_int, [_H, _dp, _dp]),
    "aniso_sigma_coeff": (C.c_int, [_H, _dp, _dp]),
    "aniso_cache": (C.c_int, [_H, C.c_int]),
    "aniso_cache_part": (C.c_int, [_H, C.c_int, C.c_int]),
    "aniso_set_option": (C.c_int, [_H, C.c_char_p, C.c_char_p]),
    "aniso_cache_bytes": (C.c_size_t, [_H, C.c_int]),
    "aniso_cache_counts": (C.c_int, [_H, _lp]),
    "aniso_cache_times": (C.c_int, [_H, _dp]),
    "aniso_mapping": (C.c_int, [_H, _dp, C.c_int, _dp]),
    "aniso_mapping_dev": (C.c_int, [_H, _V, C.c_int, _V]),
    "aniso_forward_dev": (C.c_int, [_H, _V, _V]),
    "aniso_interpolation": (C.c_int, [_H, _dp, _dp]),
    "aniso_run_kernels_fast": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_run_kernels_fast_sing": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_near_removal": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_refine_add_on_fast": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_singular_add_fast": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_gmres": (C.c_int, [_H, _dp, _dp, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]),
    "aniso_gmres_dev": (C.c_int, [_H, _V, _V, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]),
    "aniso_gmres_prec": (C.c_int, [_H, _dp, _dp, C.c_int, C.c_int, C.c_double, PREC_FN, C.c_void_p, C.POINTER(C.c_int), _dp, C.c_int]),
    "aniso_block_gmres_prec": (C.c_int, [_H, _dp, _dp, C.c_int, C.c_int, C.c_double, PREC_FN, C.c_void_p, C.POINTER(C.c_int), _dp,
                                          C.c_int]),
    "aniso_block_forward": (C.c_int, [_H, _dp, C.c_int, _dp]),
    "aniso_block_forward_dev": (C.c_int, [_H, _V, C.c_int, _V]),
    "aniso_block_gmres": (C.c_int, [_H, _dp, _dp, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]),
    "aniso_block_gmres_dev": (C.c_int, [_H, _V, _V, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]),
    "aniso_fmm_create": (C.c_int, [C.c_int, C.c_int, _dp, _dp, C.c_int, C.c_int, C.POINTER(_H)]),
    "aniso_fmm_cache": (C.c_int, [_H, EVAL_FN, C.c_void_p, C.c_int]),
    "aniso_fmm_apply": (C.c_int, [_H, _dp, _dp]),
    "aniso_line_integral": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_line_integral_noalloc": (C.c_int, [_H, C.c_int, _dp, _dp]),
    "aniso_kernel_eval": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, _dp, _dp]),
    "aniso_stream": (C.c_void_p, [_H]),
    "aniso_device": (C.c_int, [_H]),
    "aniso_synchronize": (C.c_int, [_H]),
    "aniso_dev_alloc": (C.c_int, [_H, C.c_size_t, C.POINTER(C.c_void_p)]),
    "aniso_dev_free": (C.c_int, [_H, C.c_void_p]),
    "aniso_host_alloc": (C.c_int, [_H, C.c_size_t, C.POINTER(C.c_void_p)]),
    "aniso_host_free": (C.c_int, [_H, C.c_void_p]),
    "aniso_to_slots": (C.c_int, [_H, _dp, _V]),
    "aniso_from_slots": (C.c_int, [_H, _V, _dp]),
    "aniso_launch_count": (C.c_int64, [_H]),
    "aniso_apply_bytes": (C.c_int, [_H, _lp]),
    "aniso_time_kernel": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "aniso_time_forward": (C.c_int, [_H, C.c_int, C.POINTER(C.c_double)]),
    "aniso_get_config": (C.c_char_p, [_H]),
}

NOALLOC_MAX_SEGMENTS = 32768  # ANISO_NOALLOC_MAX_SEGMENTS of include/aniso_b200.h

_lib = None


class AnisoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("aniso_b200 error %d: %s" % (code, msg))
        self.code = code


def load():
    """Loads the shared library (raises if it has not been built: no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(or make -C aniso_b200/csrc); aniso_b200 has no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc < 0:
        raise AnisoError(rc, load().aniso_last_error().decode())
    return rc
