"""ctypes binding of libhmb200.so (include/hmb200.h).  There is no fallback: a missing library is an error."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_char, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_void_p
from pathlib import Path

from .flat import hmb200_flat

HOST, DEVICE = 0, 1
import os as _os

# HMB200_LIB: another build of the same ABI (A/B runs of two kernel versions in one GPU session)
LIB_PATH = Path(_os.environ["HMB200_LIB"]).resolve() if _os.environ.get("HMB200_LIB") else Path(__file__).resolve().parent / "libhmb200.so"


class hmb200_options(ctypes.Structure):
    _fields_ = [(n, c_int32) for n in ("stage_bytes", "stage_z_bytes", "n_stages", "consumer_warps", "piece1_bytes", "piece2_bytes",
                                       "bin_rows", "unit_stages", "ctas", "build_transpose", "no_matmat", "mm_task_cols", "mm1_kg", "mm2_kg", "reserved0")] + [("reserved", c_int32 * 1)]


class hmb200_info(ctypes.Structure):
    _fields_ = [("n_rows", c_int32), ("n_cols", c_int32), ("is_complex", c_int32), ("n_bins", c_int32),
                ("n_leaves", c_int64), ("n_pieces1", c_int64), ("n_pieces2", c_int64), ("n_stages1", c_int64),
                ("n_stages2", c_int64), ("n_reduce_segments", c_int64), ("stream1_bytes", c_int64), ("stream2_bytes", c_int64),
                ("table_bytes", c_int64), ("workspace_bytes", c_int64), ("stored_coefficients", c_double),
                ("pack_seconds", c_double), ("upload_seconds", c_double), ("mm_stream_bytes", c_int64), ("mm_tiles", c_int64),
                ("mm_stages", c_int64), ("mm_rows", c_int64)]


EXPORTS = {
    "hmb200_create_d": (c_void_p, [POINTER(hmb200_flat), c_int, c_void_p, POINTER(hmb200_options)]),
    "hmb200_create_z": (c_void_p, [POINTER(hmb200_flat), c_int, c_void_p, POINTER(hmb200_options)]),
    "hmb200_destroy": (None, [c_void_p]),
    "hmb200_get_info": (c_int, [c_void_p, POINTER(hmb200_info)]),
    "hmb200_set_stream": (c_int, [c_void_p, c_void_p]),
    "hmb200_set_graphs": (c_int, [c_void_p, c_int]),
    "hmb200_add_vector_product_d": (c_int, [c_void_p, c_char, c_void_p, c_void_p, c_void_p, c_void_p, c_int]),
    "hmb200_add_vector_product_z": (c_int, [c_void_p, c_char, c_void_p, c_void_p, c_void_p, c_void_p, c_int]),
    "hmb200_add_matrix_product_row_major_d": (c_int, [c_void_p, c_char, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "hmb200_add_matrix_product_row_major_z": (c_int, [c_void_p, c_char, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "hmb200_set_cluster_offsets": (c_int, [c_void_p, c_int, c_int]),
    "hmb200_add_sub_matrix_product_to_local_d": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int]),
    "hmb200_add_sub_matrix_product_to_local_z": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int]),
    "hmb200_peer_create": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_void_p]),
    "hmb200_peer_attach": (c_int, [c_void_p, c_void_p]),
    "hmb200_peer_buffer": (c_int, [c_void_p, c_int, POINTER(c_void_p)]),
    "hmb200_peer_status": (c_int, [c_void_p, POINTER(c_int)]),
    "hmb200_set_global_permutations": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64]),
    "hmb200_matvec_user_chunked_d": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64]),
    "hmb200_matvec_user_chunked_z": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64]),
    "hmb200_matmat_user_chunked_d": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int]),
    "hmb200_matmat_user_chunked_z": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int]),
    "hmb200_peer_mm_create": (c_int, [c_void_p, c_int, c_void_p]),
    "hmb200_peer_mm_attach": (c_int, [c_void_p, c_void_p]),
    "hmb200_peer_mm_detach": (c_int, [c_void_p]),
    "hmb200_peer_mm_buffer": (c_int, [c_void_p, c_int, POINTER(c_void_p)]),
    "hmb200_matrix_product_allgather_d": (c_int, [c_void_p, c_void_p, c_void_p, c_int, POINTER(c_void_p)]),
    "hmb200_matrix_product_allgather_z": (c_int, [c_void_p, c_void_p, c_void_p, c_int, POINTER(c_void_p)]),
    "hmb200_peer_allgather": (c_int, [c_void_p, c_void_p, POINTER(c_void_p)]),
    "hmb200_peer_allgather_range": (c_int, [c_void_p, c_void_p, c_int64, c_int64, POINTER(c_void_p)]),
    "hmb200_peer_allreduce_f64": (c_int, [c_void_p, c_void_p, c_int]),
    "hmb200_peer_exchange_sum": (c_int, [c_void_p, c_char, c_void_p, c_void_p, c_void_p, c_void_p]),
    "hmb200_vector_product_allgather_d": (c_int, [c_void_p, c_void_p, c_void_p, POINTER(c_void_p)]),
    "hmb200_vector_product_allgather_z": (c_int, [c_void_p, c_void_p, c_void_p, POINTER(c_void_p)]),
    "hmb200_permute_gather": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "hmb200_permute_scatter": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "hmb200_block_to_rows": (c_int, [c_void_p, c_int, c_int, c_void_p, ctypes.c_longlong, c_void_p, c_int, c_void_p]),
    "hmb200_rows_to_block": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, ctypes.c_longlong, c_int, c_void_p]),
    "hmb200_set_permutations": (c_int, [c_void_p, c_void_p, c_void_p]),
    "hmb200_matvec_user_d": (c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    "hmb200_matvec_user_z": (c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    "hmb200_matmat_user_d": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "hmb200_matmat_user_z": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "hmb200_krylov_scratch_elems": (c_int64, [c_int64, c_int]),
    "hmb200_krylov_dots": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "hmb200_krylov_combine": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_double, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "hmb200_krylov_axpby": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "hmb200_krylov_scale": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_void_p]),
    "hmb200_krylov_dots_static": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int, c_void_p]),
    "hmb200_last_kernel_ms": (c_int, [c_void_p, POINTER(c_float * 4)]),
    "hmb200_kernel_ms_avg": (c_int, [c_void_p, POINTER(c_float * 4), POINTER(c_int)]),
    "hmb200_enable_timing": (c_int, [c_void_p, c_int]),
    "hmb200_launch_count": (c_int64, [c_void_p]),
    "hmb200_fp64_rate": (c_int, [c_int, POINTER(c_double), POINTER(c_double)]),
    "hmb200_host_flag_publish": (c_int, [c_void_p, c_int64]),
    "hmb200_host_flags_wait": (c_int, [c_void_p, c_int64, c_int, c_int64, c_double, POINTER(c_int)]),
    "hmb200_last_error": (c_char_p, []),
    "hmb200_version": (c_char_p, []),
}

_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is not built (python -m htool_python_b200.build cuda); the engine has no fallback path")
        _lib = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in EXPORTS.items():
            fn = getattr(_lib, name)  # AttributeError if the symbol is missing
            fn.restype, fn.argtypes = res, args
    return _lib


def last_error() -> str:
    return lib().hmb200_last_error().decode()


def check(status: int) -> None:
    if status != 0:
        raise RuntimeError("hmb200: " + last_error())
