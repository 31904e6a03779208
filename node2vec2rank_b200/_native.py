"""ctypes binding of libn2v2r_b200.so (include/n2v2r_b200.h).

There is NO CPU fallback: every compute entry point raises if the library is missing or
no sm_100 GPU is visible.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libn2v2r_b200.so")

OK, EINVAL, ENOMEM, ECUDA, ECOMM, ENOTCONV, EUNSUPPORTED = range(7)
F32, F64 = 0, 1
EUCLIDEAN, COSINE, CORRELATION = 0, 1, 2
SEQUENTIAL, ONE_VS_BEFORE, ONE_VS_REST = 0, 1, 2
ALGO_TCGEN05, ALGO_SIMT = 0, 1

METRICS = {"euclidean": EUCLIDEAN, "cosine": COSINE, "correlation": CORRELATION}
STRATEGIES = {"sequential": SEQUENTIAL, "one_vs_before": ONE_VS_BEFORE, "one_vs_rest": ONE_VS_REST}


class NativeError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"n2v2r_b200 status {status}: {message}")
        self.status = status


STORE_PLANES, STORE_TILES, STORE_E4M3, STORE_ROWS = 0, 1, 2, 3
MAX_CHECK = 32


class Layer(C.Structure):
    _fields_ = [("d_planes", C.c_void_p), ("d_planes_t", C.c_void_p), ("n_planes", C.c_int32),
                ("storage", C.c_int32), ("ld", C.c_int64),
                ("d_tile_ptr", C.c_void_p), ("d_entries", C.c_void_p), ("d_tile_ptr_t", C.c_void_p), ("d_entries_t", C.c_void_p),
                ("n_entries", C.c_int64), ("n_entries_t", C.c_int64), ("tile_rows", C.c_int32), ("panel_cols", C.c_int32),
                ("row_offset", C.c_int64), ("rows", C.c_int64), ("ld_t", C.c_int64), ("list_cap", C.c_int32), ("list_cap_t", C.c_int32)]


class UaseParams(C.Structure):
    _fields_ = [("n", C.c_int64), ("T_local", C.c_int32), ("T_global", C.c_int32), ("layer_offset", C.c_int32),
                ("k", C.c_int32), ("l", C.c_int32), ("max_iters", C.c_int32), ("min_iters", C.c_int32),
                ("cheb_degree", C.c_int32), ("algo", C.c_int32), ("k_chunk", C.c_int32), ("n_check", C.c_int32),
                ("tol", C.c_double), ("seed", C.c_uint64), ("h_check_dims", C.POINTER(C.c_int32)),
                ("floor_mult", C.c_double), ("rel_floor", C.c_double), ("first_degree", C.c_int32), ("no_early_lock", C.c_int32),
                ("trace", C.c_int32), ("comm_mode", C.c_int32), ("group_comm", C.c_void_p), ("batch", C.c_void_p), ("batch_slot", C.c_int32),
                ("reserved2", C.c_int32)]


class UaseStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("layer_passes", C.c_int32), ("converged", C.c_int32),
                ("kernel_launches", C.c_int32), ("outer_iterations", C.c_int32), ("locked", C.c_int32),
                ("stream_ms", C.c_double), ("mma_ms", C.c_double), ("mma_launches", C.c_int32), ("bound_met", C.c_int32),
                ("total_ms", C.c_double),
                ("stream_bytes", C.c_double), ("max_angle_bound", C.c_double),
                ("resid", C.c_double * 96), ("ritz", C.c_double * 96), ("angle_bounds", C.c_double * MAX_CHECK)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p)

# name -> (restype, argtypes); every symbol include/n2v2r_b200.h declares
_vp, _i, _i64, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t
SIGNATURES = {
    "n2v2r_abi_version": (_i, []),
    "n2v2r_last_error": (C.c_char_p, []),
    "n2v2r_device_info": (_i, [_i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz)]),
    "n2v2r_sbm_generate": (_i, [_vp, _vp, _i64, _i64, _i, C.POINTER(C.c_float), C.c_uint64, _i, _i]),
    "n2v2r_select_workspace_bytes": (_sz, []),
    "n2v2r_select_nonzero": (_i, [_vp, _vp, _i, _i64, _i64, _i64, _i, _i, _d, _d, _i, _vp, _sz,
                                  C.POINTER(_i64), C.POINTER(_d), C.POINTER(_d), C.POINTER(_d)]),
    "n2v2r_transform": (_i, [_vp, _vp, _i, _i64, _i64, _i64, _i, _i, _d, _i, _d, _i, _vp, _i64, _vp, _i, _i64, _vp]),
    "n2v2r_planes_needed": (_i, [_vp, _vp, _i64, _i64, _i64, C.POINTER(_i)]),
    "n2v2r_plane_to_e4m3": (_i, [_vp, _vp, _i64, _i64, _i64, _vp, _i64, C.POINTER(_i)]),
    "n2v2r_is_symmetric": (_i, [_vp, _vp, _i, _i64, _i64, C.POINTER(_i)]),
    "n2v2r_transpose_planes": (_i, [_vp, _vp, _i, _i64, _i64, _vp]),
    "n2v2r_tiles_count": (_i, [_vp, _vp, _i64, _i64, _vp, _vp, C.POINTER(_i), C.POINTER(_i)]),
    "n2v2r_tiles_scan": (_i, [_vp, _vp, _i64, _i, _vp, C.POINTER(_i64)]),
    "n2v2r_tiles_fill": (_i, [_vp, _vp, _i64, _i64, _i, _i, _vp, _vp]),
    "n2v2r_tiles_to_dense": (_i, [_vp, _vp, _vp, _i64, _i, _vp, _i64]),
    "n2v2r_rows_counts_bytes": (_sz, [_i64, _i, _i]),
    "n2v2r_rows_count": (_i, [_vp, _vp, _i64, _i64, _i, _i, _vp, _vp, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "n2v2r_rows_fill": (_i, [_vp, _vp, _i64, _i64, _i, _i, _vp, _vp, _vp]),
    "n2v2r_rows_to_dense": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp, _i64]),
    "n2v2r_project_workspace_bytes": (_sz, [_i64, _i64, _i]),
    "n2v2r_project": (_i, [_vp, _vp, _i, _i64, _i64, _i64, _i, _vp, _i64, _vp, _sz]),
    "n2v2r_uase_workspace_bytes": (_sz, [C.POINTER(UaseParams)]),
    "n2v2r_uase": (_i, [_vp, C.POINTER(UaseParams), C.POINTER(Layer), _vp, _sz, ALLREDUCE_FN, _vp, _vp, _vp,
                        C.POINTER(_d), C.POINTER(UaseStats)]),
    "n2v2r_batch_workspace_bytes": (_sz, [C.POINTER(UaseParams), _i]),
    "n2v2r_batch_create": (_i, [C.POINTER(UaseParams), C.POINTER(Layer), _i, _vp, _sz, C.POINTER(_vp)]),
    "n2v2r_batch_destroy": (_i, [_vp, C.POINTER(_d)]),
    "n2v2r_comm_unique_id": (_i, [_vp]),
    "n2v2r_comm_init": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "n2v2r_comm_destroy": (_i, [_vp]),
    "n2v2r_comm_allreduce_f32": (_i, [_vp, _vp, _vp, _i64]),
    "n2v2r_apply_workspace_bytes": (_sz, [_i64, _i, _i]),
    "n2v2r_layer_apply": (_i, [_vp, C.POINTER(Layer), _i64, _i, _i, _vp, _vp, _i, _i, _vp, _sz]),
    "n2v2r_distances": (_i, [_vp, _vp, _i, _i, _i64, _i, C.POINTER(C.c_int32), C.POINTER(C.c_int32), _i, _i, _vp]),
    "n2v2r_rank_workspace_bytes": (_sz, [_i64, _i]),
    "n2v2r_rank_borda": (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp, _vp, _sz]),
    "n2v2r_rank_borda_noalloc": (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp]),
    "n2v2r_signed_scores": (_i, [_vp, _vp, _i, _i64, _i, _vp, _vp]),
    "n2v2r_ndcg": (_i, [_vp, _vp, _i64, _i64, _vp, _i64, _vp, _vp]),
    "n2v2r_scatter_edges": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _i, _i64, _i64]),
    "n2v2r_gather_submatrix": (_i, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i, _vp, _i64]),
    "n2v2r_host_pack_nonzeros": (_i, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _vp, _i, _vp]),
    "n2v2r_unpack_nonzeros": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _vp, _i64]),
}

_lib = None


def load():
    """Load the shared library (no GPU needed for loading / symbol checks)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build the CUDA extension with "
                "`python -m node2vec2rank_b200.csrc.build` (or __graft_entry__.build()). "
                "node2vec2rank_b200 has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)            # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        if lib.n2v2r_abi_version() != 3:
            raise ImportError("libn2v2r_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(status, allow=()):
    if status != OK and status not in allow:
        raise NativeError(status, load().n2v2r_last_error().decode("utf-8", "replace"))
    return status


def require_gpu():
    """Fail loudly unless a CUDA device with compute capability 10.x is current."""
    import torch
    load()
    if not torch.cuda.is_available():
        raise RuntimeError("node2vec2rank_b200 needs a CUDA device (B200, sm_100a); none is visible and there is no CPU fallback")
    major, _minor = torch.cuda.get_device_capability()
    if major != 10:
        raise RuntimeError(f"node2vec2rank_b200 kernels are built for sm_100a only; found compute capability {major}.x")
    return torch.device("cuda", torch.cuda.current_device())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())
