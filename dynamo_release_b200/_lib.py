"""ctypes binding of libdvf_b200.so (the C ABI declared in include/dvf_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DVF_LIB_PATH") or os.path.join(_HERE, "libdvf_b200.so")  # env: developer experiments

SOLVER_CHOLESKY, SOLVER_NORMAL_LU = 0, 1   # DVF_SOLVER_* of include/dvf_b200.h
DVF_OK, DVF_ERR_ARG, DVF_ERR_CUDA, DVF_ERR_NUMERIC, DVF_ERR_COMM, DVF_ERR_WORKSPACE = 0, -1, -2, -3, -4, -5


class DvfError(RuntimeError):
    """CUDA / NCCL / argument failure reported by libdvf_b200."""

    def __init__(self, code, msg):
        super().__init__("libdvf_b200 error %d: %s" % (code, msg))
        self.code = code


class EmParams(C.Structure):  # dvf_em_params
    _fields_ = [("a", C.c_double), ("gamma", C.c_double), ("lambda_", C.c_double), ("minP", C.c_double),
                ("ecr", C.c_double), ("theta", C.c_double), ("max_iter", C.c_int), ("check_every", C.c_int),
                ("solver", C.c_int), ("reserved", C.c_int)]


class EmResult(C.Structure):  # dvf_em_result
    _fields_ = [("n_iter", C.c_int), ("status", C.c_int), ("sigma2", C.c_double), ("gamma", C.c_double),
                ("n_inlier", C.c_double)]


_p, _i, _i64, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t

# name -> (restype, argtypes); mirrors include/dvf_b200.h one to one (tests/test_abi.py checks the header)
SIGNATURES = {
    "dvf_abi_version": (_i, []),
    "dvf_last_error": (C.c_char_p, []),
    "dvf_profile_enable": (_i, [_i]),
    "dvf_profile_get": (_i, [C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_longlong)]),
    "dvf_kernel_matrix": (_i, [_p, _i64, _i, _p, _i, _d, _p, _i64, _p]),
    "dvf_eval": (_i, [_p, _i64, _i, _p, _i, _d, _p, _p, _p]),
    "dvf_jacobian": (_i, [_p, _i64, _i, _p, _i, _d, _p, _p, _p]),
    "dvf_jacobian_workspace_bytes": (_i, [_i64, _i, _i, C.POINTER(_sz)]),
    "dvf_jacobian_ws": (_i, [_p, _i64, _i, _p, _i, _d, _p, _p, _p, _sz, _p]),
    "dvf_lap_terms": (_i, [_p, _p, _i64, _i, _p, _i, _d, _p, _p, _p, _p]),
    "dvf_div_acc_curv": (_i, [_p, _i64, _i, _p, _i, _d, _p, _i, _p, _p, _p, _p, _p, _p, _p]),
    "dvf_gene_transform": (_i, [_p, _i64, _i, _p, _i, _p, _i, _i, _p, _i, _p, _p, _p]),
    "dvf_gene_gemm_workspace_bytes": (_i, [_i64, _i, _i, _i, _i, _i, C.POINTER(_sz)]),
    "dvf_gene_transform_gemm": (_i, [_p, _i64, _i, _p, _i, _p, _i, _i, _p, _i, _p, _p, _p, _sz, _p]),
    "dvf_gene_transform_elem": (_i, [_p, _i64, _i, _p, _p, _p, _p]),
    "dvf_gene_transform_reduce": (_i, [_p, _i64, _i, _p, _i, _p, _i, _i, _i, _p, _p, _i, _p, _p, _sz, _p]),
    "dvf_vector_transform": (_i, [_p, _i64, _i, _p, _i, _p, _p]),
    "dvf_jacobian_det": (_i, [_p, _i64, _i, _p, _p]),
    "dvf_cell_angles": (_i, [_p, _p, _i64, _i, _p, _p]),
    "dvf_row_norms": (_i, [_p, _i64, _i, _p, _p]),
    "dvf_knn_dists": (_i, [_p, _i, _i, _i, _p, _p]),
    "dvf_f32_row_means_pairwise": (_i, [_p, _i64, _i64, _p, _i, _p, _p]),
    "dvf_f32_group_means_seq": (_i, [_p, _i64, _i64, _p, _i, _p, _p, _p]),
    "dvf_hessian": (_i, [_p, _i64, _i, _p, _i, _d, _p, _p, _p]),
    "dvf_laplacian": (_i, [_p, _i64, _i, _p, _i, _d, _p, _p, _p]),
    "dvf_ivp_workspace_bytes": (_i, [_i, _i, C.POINTER(_sz)]),
    "dvf_integrate_rk45": (_i, [_p, _i64, _i, _p, _i, _d, _p, _d, _d, _d, _d, _d, _d, _d, _i, _p, _p, _p, _p, _p,
                                _p, _sz, _p]),
    "dvf_ivp_dense_eval": (_i, [_p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _i64, _i, _p, _p]),
    "dvf_fit_row_width": (_i64, [_i, _i]),
    "dvf_fit_workspace_bytes": (_i, [_i64, _i, _i, C.POINTER(_sz)]),
    "dvf_fit_em": (_i, [_p, _p, _i64, _i64, _i, _p, _i, _d, C.POINTER(EmParams), _p, _p, _p, _p, _p,
                        C.POINTER(EmResult), _p, _sz, _p, _p]),
    "dvf_build_uy": (_i, [_p, _p, _i64, _i, _p, _i, _d, _p, _p]),
    "dvf_estep": (_i, [_p, _i64, _i, _d, _d, _d, _p, _p, _p]),
    "dvf_gram_ld": (_i, [_i, _i]),
    "dvf_gram_scratch_bytes": (_i, [_i64, _i, _i, C.POINTER(_sz)]),
    "dvf_gram": (_i, [_p, _p, _d, _i64, _i, _i, _p, _p, _sz, _p]),
    "dvf_solve": (_i, [_p, _i64, _p, _i, _i, _d, _p, _p, _p, _p]),
    "dvf_solve_normal_workspace_bytes": (_i, [_i, _i, C.POINTER(_sz)]),
    "dvf_solve_normal": (_i, [_p, _i64, _p, _i, _i, _d, _p, _sz, _p, _p, _p]),
    "dvf_update_v": (_i, [_p, _p, _p, _d, _i64, _i, _i, _p, _p, _p, _p]),
    "dvf_update_v_noalloc": (_i, [_p, _p, _p, _d, _i64, _i, _i, _p, _p, _p, _p]),
    "dvf_comm_unique_id": (_i, [_p]),
    "dvf_comm_init": (_i, [_p, _i, _i, C.POINTER(_p)]),
    "dvf_comm_destroy": (_i, [_p]),
    "dvf_comm_allreduce_sum": (_i, [_p, _p, _i64, _p]),
    "dvf_host_cdf": (_i, [_p, _i64, _p]),
    "dvf_host_scale": (_i, [_p, _i64, C.c_double, _p]),
    "dvf_host_pairwise_sum": (_i, [_p, _i64, _p]),
    "dvf_host_choice_cumsum": (_i, [_p, _i64, C.c_double, _p, _i64, _p]),
    "dvf_host_searchsorted_div": (_i, [_p, _i64, _p, _i64, _p]),
    "dvf_memcpy2d_async": (_i, [_p, _sz, _p, _sz, _sz, _sz, _i, _p]),
}

_lib = None


def load():
    """Load libdvf_b200.so (once).  Raises if it has not been built: python -c 'import __graft_entry__ as g; g.build()'."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libdvf_b200.so not found at %s -- build it with `make -C dynamo_release_b200/csrc` "
            "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().dvf_last_error()
        raise DvfError(rc, msg.decode("utf-8", "replace") if msg else "")
