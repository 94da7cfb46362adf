"""ctypes mirror of include/dimreduce4gpu.h.

``params`` and ``_load_tsvd_lib`` / ``_load_pca_lib`` keep the reference's names and meaning
(dimreduce4gpu/lib_dimreduce4gpu.py:6-18, :21-61, :64-105); the dr4g_* bindings are additive.
"""
from __future__ import annotations

import ctypes

c_float_p = ctypes.POINTER(ctypes.c_float)
c_double_p = ctypes.POINTER(ctypes.c_double)


class params(ctypes.Structure):
    # include/dimreduce4gpu.h `params` == reference include/tsvd.h:24-35 (48 bytes, by value)
    _fields_ = [
        ("X_n", ctypes.c_int),
        ("X_m", ctypes.c_int),
        ("k", ctypes.c_int),
        ("algorithm", ctypes.c_char_p),
        ("n_iter", ctypes.c_int),
        ("random_state", ctypes.c_int),
        ("tol", ctypes.c_float),
        ("verbose", ctypes.c_int),
        ("gpu_id", ctypes.c_int),
        ("whiten", ctypes.c_bool),
    ]


def _bind(lib: ctypes.CDLL) -> None:
    vp, ll, ci = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
    for name, fp in (("float", c_float_p), ("double", c_double_p)):
        f = getattr(lib, f"truncated_svd_{name}")
        f.argtypes, f.restype = [fp] * 7 + [params], None
        f = getattr(lib, f"pca_{name}")
        f.argtypes, f.restype = [fp] * 8 + [params], None
        # device-pointer API: pointers are passed as integers (tensor.data_ptr())
        f = getattr(lib, f"dr4g_fit_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, ci, ci, params, vp, vp, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_fit_transform_host_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, ci, ci, params, vp, vp, vp, vp, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_partial_fit_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, ci], ci
        f = getattr(lib, f"dr4g_partial_fit_finalize_{name}")
        f.argtypes, f.restype = [vp, ci, params, vp, vp, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_transform_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, ci, vp, vp, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_inverse_transform_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, ci, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_csr_to_dense_{name}")
        f.argtypes, f.restype = [vp, vp, vp, vp, ll, ll, ci, vp], ci
        f = getattr(lib, f"dr4g_colstats_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, vp, vp, vp], ci
        f = getattr(lib, f"dr4g_gram_{name}")
        f.argtypes, f.restype = [vp, vp, ll, ci, vp, vp], ci
    lib.dr4g_last_error.argtypes, lib.dr4g_last_error.restype = [], ctypes.c_char_p
    lib.dr4g_abi_version.argtypes, lib.dr4g_abi_version.restype = [], ci
    lib.dr4g_ctx_create.argtypes, lib.dr4g_ctx_create.restype = [ci], vp
    lib.dr4g_ctx_destroy.argtypes, lib.dr4g_ctx_destroy.restype = [vp], None
    lib.dr4g_ctx_set_stream.argtypes, lib.dr4g_ctx_set_stream.restype = [vp, vp], ci
    lib.dr4g_sync.argtypes, lib.dr4g_sync.restype = [vp], ci
    lib.dr4g_set_option.argtypes, lib.dr4g_set_option.restype = [vp, ctypes.c_char_p, ctypes.c_double], ci
    lib.dr4g_nccl_unique_id.argtypes, lib.dr4g_nccl_unique_id.restype = [vp], ci
    lib.dr4g_comm_init.argtypes, lib.dr4g_comm_init.restype = [vp, vp, ci, ci], ci
    lib.dr4g_comm_world.argtypes, lib.dr4g_comm_world.restype = [vp], ci
    lib.dr4g_eigh_double.argtypes, lib.dr4g_eigh_double.restype = [vp, vp, ci, vp, vp], ci
    lib.dr4g_eigh_topk_double.argtypes, lib.dr4g_eigh_topk_double.restype = [vp, vp, ci, ci, vp, vp], ci
    lib.dr4g_eigh_topk_static_double.argtypes, lib.dr4g_eigh_topk_static_double.restype = [vp, vp, ci, ci, vp, vp], ci
    lib.dr4g_eigh_topk_static_max_d.argtypes, lib.dr4g_eigh_topk_static_max_d.restype = [], ci
    lib.dr4g_stage_times.argtypes, lib.dr4g_stage_times.restype = [vp, c_double_p, ci], ci
    lib.dr4g_last_gram_path.argtypes, lib.dr4g_last_gram_path.restype = [vp], ci
    lib.dr4g_last_proj_path.argtypes, lib.dr4g_last_proj_path.restype = [vp], ci
    lib.dr4g_last_fold.argtypes, lib.dr4g_last_fold.restype = [vp], ci
    lib.dr4g_p2p_prepare.argtypes, lib.dr4g_p2p_prepare.restype = [vp, ctypes.c_longlong, vp], ci
    lib.dr4g_p2p_connect.argtypes, lib.dr4g_p2p_connect.restype = [vp, vp], ci
    lib.dr4g_last_exchange_path.argtypes, lib.dr4g_last_exchange_path.restype = [vp], ci
    lib.dr4g_launch_count.argtypes, lib.dr4g_launch_count.restype = [vp], ll
    lib.dr4g_partial_fit_reset.argtypes, lib.dr4g_partial_fit_reset.restype = [vp], ci
    lib.dr4g_partial_fit_rows.argtypes, lib.dr4g_partial_fit_rows.restype = [vp], ll


# every symbol include/dimreduce4gpu.h declares (checked by tests/test_abi.py with `nm -D` semantics)
EXPORTED_SYMBOLS = (
    "truncated_svd_float", "truncated_svd_double", "pca_float", "pca_double",
    "dr4g_last_error", "dr4g_abi_version", "dr4g_ctx_create", "dr4g_ctx_destroy", "dr4g_ctx_set_stream",
    "dr4g_sync", "dr4g_set_option", "dr4g_nccl_unique_id", "dr4g_comm_init", "dr4g_comm_world",
    "dr4g_fit_float", "dr4g_fit_double", "dr4g_transform_float", "dr4g_transform_double",
    "dr4g_inverse_transform_float", "dr4g_inverse_transform_double", "dr4g_colstats_float",
    "dr4g_colstats_double", "dr4g_gram_float", "dr4g_gram_double", "dr4g_eigh_double", "dr4g_eigh_topk_double", "dr4g_stage_times",
    "dr4g_last_gram_path", "dr4g_last_proj_path", "dr4g_last_fold", "dr4g_launch_count",
    "dr4g_eigh_topk_static_double", "dr4g_eigh_topk_static_max_d",
    "dr4g_p2p_prepare", "dr4g_p2p_connect", "dr4g_last_exchange_path",
    "dr4g_fit_transform_host_float", "dr4g_fit_transform_host_double",
    "dr4g_csr_to_dense_float", "dr4g_csr_to_dense_double",
    "dr4g_partial_fit_float", "dr4g_partial_fit_double", "dr4g_partial_fit_finalize_float",
    "dr4g_partial_fit_finalize_double", "dr4g_partial_fit_reset", "dr4g_partial_fit_rows",
)


def _load_tsvd_lib(dtype: str = "float"):
    """Function pointer of truncated_svd_float (reference lib_dimreduce4gpu.py:21-61)."""
    from ._native import load_library

    return getattr(load_library(), f"truncated_svd_{dtype}")


def _load_pca_lib(dtype: str = "float"):
    """Function pointer of pca_float (reference lib_dimreduce4gpu.py:64-105)."""
    from ._native import load_library

    return getattr(load_library(), f"pca_{dtype}")
