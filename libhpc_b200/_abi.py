"""ctypes binding of libhpcx.so — one prototype per entry point of include/hpcx.h.

The library is the product; this module only declares it.  There is no fallback: if the
shared object is missing, `load()` raises, and every compute entry point of the library
itself fails with HPCX_CUDA_ERROR when no sm_100 device is present.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libhpcx.so")

OK, INVALID_ARG, OOM, CUDA_ERROR, NCCL_ERROR, UNSUPPORTED, OUT_OF_RANGE = range(7)
F32, F64 = 0, 1
MEM_HOST, MEM_DEVICE, MEM_MANAGED = 0, 1, 2
STENCIL_BOX_X, STENCIL_BOX_Y, STENCIL_JACOBI7 = 1, 2, 3
VARIANT_AUTO, VARIANT_TMA, VARIANT_DIRECT, VARIANT_TEMPORAL = 0, 1, 2, 3
SPMV_AUTO, SPMV_STREAM, SPMV_VECTOR, SPMV_SELL, SPMV_HUB = 0, 1, 2, 3, 4

STATUS_NAMES = {0: "HPCX_OK", 1: "HPCX_INVALID_ARG", 2: "HPCX_OOM", 3: "HPCX_CUDA_ERROR",
                4: "HPCX_NCCL_ERROR", 5: "HPCX_UNSUPPORTED", 6: "HPCX_OUT_OF_RANGE"}


class GridDesc(C.Structure):
    _fields_ = [("ndim", C.c_int32), ("dtype", C.c_int32), ("lo", C.c_int32), ("hi", C.c_int32),
                ("dim", C.c_int64 * 3)]


class StencilDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("radius", C.c_int32), ("variant", C.c_int32),
                ("zsplit", C.c_int32)]


class CsrDesc(C.Structure):
    _fields_ = [("nrows", C.c_int64), ("ncols", C.c_int64), ("nnz", C.c_int64),
                ("row_ptr", C.c_void_p), ("idx_bytes", C.c_int32), ("dtype", C.c_int32),
                ("col_idx", C.c_void_p), ("values", C.c_void_p), ("mem", C.c_int32),
                ("reserved", C.c_int32)]


class CsrOpts(C.Structure):
    _fields_ = [("kernel", C.c_int32), ("tile_nnz", C.c_int32), ("sell_c", C.c_int32),
                ("sell_sigma", C.c_int32), ("lanes_per_row", C.c_int32),
                ("short_row_max", C.c_int32), ("reserved", C.c_int32 * 2)]


class CsrArrays(C.Structure):
    _fields_ = [("nnz", C.c_int64), ("row_ptr", C.c_void_p), ("col_idx", C.c_void_p),
                ("values", C.c_void_p)]


class CgResult(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("converged", C.c_int32), ("rr", C.c_double),
                ("bb", C.c_double)]


class CellMap(C.Structure):
    _fields_ = [("nrows", C.c_int64), ("cell_row", C.c_void_p), ("row_x", C.c_void_p), ("row_y", C.c_void_p)]


class HaloSizes(C.Structure):
    _fields_ = [("nloc", C.c_int64), ("nnz_loc", C.c_int64), ("nhalo", C.c_int64),
                ("nsend", C.c_int64)]


_vp, _i32, _i64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
_P = C.POINTER

# name -> (restype, argtypes); must list every symbol include/hpcx.h declares.
PROTOTYPES = {
    "hpcx_version": (C.c_char_p, []),
    "hpcx_init": (_i32, [_i32, _P(_vp)]),
    "hpcx_finalize": (_i32, [_vp]),
    "hpcx_last_error_string": (C.c_char_p, [_vp]),
    "hpcx_synchronize": (_i32, [_vp]),
    "hpcx_launch_count": (_i64, [_vp]),
    "hpcx_device_count": (_i32, []),
    "hpcx_host_alloc": (_i32, [_vp, _sz, _P(_vp)]),
    "hpcx_host_free": (_i32, [_vp, _vp]),
    "hpcx_device_alloc": (_i32, [_vp, _sz, _P(_vp)]),
    "hpcx_device_free": (_i32, [_vp, _vp]),
    "hpcx_mem_alloc": (_i32, [_i32, _sz, _P(_vp)]),
    "hpcx_mem_free": (_i32, [_i32, _vp]),
    "hpcx_memcpy": (_i32, [_vp, _vp, _vp, _sz]),
    "hpcx_grid_stride_total": (_i32, [_P(GridDesc), _P(_i64), _P(_i64)]),
    "hpcx_grid_linearize": (_i32, [_P(GridDesc), _P(_i64), _i32, _P(_i64)]),
    "hpcx_grid_create": (_i32, [_vp, _P(GridDesc), _P(_vp)]),
    "hpcx_grid_destroy": (_i32, [_vp]),
    "hpcx_grid_upload": (_i32, [_vp, _vp, _vp]),
    "hpcx_grid_download": (_i32, [_vp, _vp, _vp]),
    "hpcx_grid_device_layout": (_i32, [_vp, _P(_i64), _P(_i64), _P(_i64), _P(_vp)]),
    "hpcx_grid_export_halo_maps": (_i32, [_P(GridDesc), _i32, _i32, _i32, _P(_i64), _P(_i64), _vp, _vp]),
    "hpcx_stencil_sweep": (_i32, [_vp, _vp, _P(StencilDesc), _i32, _vp]),
    "hpcx_conv2d": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "hpcx_conv2d_static_scratch": (_i32, [_vp, _vp, _vp, _i32, _i32, _vp]),
    "hpcx_grid_transpose": (_i32, [_vp, _vp, _vp]),
    "hpcx_grid_face_elems": (_i32, [_vp, _i32, _i32, _P(_i64)]),
    "hpcx_grid_pack_face": (_i32, [_vp, _i32, _i32, _i32, _vp, _i32, _vp]),
    "hpcx_grid_unpack_face": (_i32, [_vp, _i32, _i32, _i32, _vp, _i32, _vp]),
    "hpcx_csr_create": (_i32, [_vp, _P(CsrDesc), _P(CsrOpts), _P(_vp)]),
    "hpcx_csr_destroy": (_i32, [_vp]),
    "hpcx_spmv": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp]),
    "hpcx_csr_download": (_i32, [_vp, _vp, _vp, _vp]),
    "hpcx_csr_rowblocks": (_i32, [_vp, _P(_i64), _vp, _i64]),
    "hpcx_csr_row_bins": (_i32, [_vp, _vp]),
    "hpcx_csr_sell_info": (_i32, [_vp, _P(_i64), _P(_i64), _P(_i32), _P(_i32)]),
    "hpcx_csr_sell_download": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "hpcx_csr_from_coo": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _i32, _i32, _i32, _P(CsrArrays)]),
    "hpcx_csr_from_cells": (_i32, [_vp, _i64, _vp, _vp, _i32, _i32, _i32, _P(CsrArrays), _P(CellMap)]),
    "hpcx_synth_rmat_coo": (_i32, [_vp, _i32, _i64, C.c_uint64, _vp, _vp, _vp]),
    "hpcx_synth_field": (_i32, [_vp, C.c_uint64, _i64, _i64, _i32, _vp]),
    "hpcx_partition_rows": (_i32, [_i64, _vp, _i32, _i32, _i32, _vp]),
    "hpcx_partition_slabs": (_i32, [_i64, _i32, _vp]),
    "hpcx_csr_localize": (_i32, [_i64, _vp, _i32, _vp, _vp, _i32, _i32, _P(HaloSizes), _vp, _vp,
                                 _vp, _vp, _vp, _vp, _vp]),
    "hpcx_csr_localize_rows": (_i32, [_i64, _i64, _vp, _i32, _vp, _vp, _i32, _i32, _P(HaloSizes), _vp, _vp, _vp, _vp]),
    "hpcx_halo_send_lists": (_i32, [_vp, _i32, _i32, _vp, _vp, _vp]),
    "hpcx_comm_unique_id": (_i32, [_vp, _vp]),
    "hpcx_comm_init": (_i32, [_vp, _i32, _i32, _vp]),
    "hpcx_comm_finalize": (_i32, [_vp]),
    "hpcx_comm_barrier": (_i32, [_vp]),
    "hpcx_comm_ipc_export": (_i32, [_vp, _vp, _vp]),
    "hpcx_comm_ipc_open": (_i32, [_vp, _vp, _P(_vp)]),
    "hpcx_comm_ipc_close": (_i32, [_vp, _vp]),
    "hpcx_comm_set_timeout": (_i32, [_vp, _i64]),
    "hpcx_comm_check": (_i32, [_vp]),
    "hpcx_stencil_dist_plan": (_i32, [_vp, _vp, _vp, _P(StencilDesc), _i32, _P(_vp)]),
    "hpcx_stencil_dist_plan_blocks": (_i32, [_vp, _vp, _vp, _P(StencilDesc), _i32, _i32, _P(_vp)]),
    "hpcx_stencil_dist_sweep": (_i32, [_vp, _i32, _vp]),
    "hpcx_stencil_dist_destroy": (_i32, [_vp]),
    "hpcx_csr_set_halo": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "hpcx_csr_halo_transport": (_i32, [_vp, _i32]),
    "hpcx_csr_x_buffer": (_i32, [_vp, _P(_vp)]),
    "hpcx_spmv_dist": (_i32, [_vp, _vp, _vp, _vp]),
    "hpcx_spmv_dist_host": (_i32, [_vp, _vp, _vp, _vp]),
    "hpcx_init_multi": (_i32, [_i32, _vp, _P(_vp)]),
    "hpcx_multi_finalize": (_i32, [_vp]),
    "hpcx_multi_size": (_i32, [_vp]),
    "hpcx_multi_ctx": (_vp, [_vp, _i32]),
    "hpcx_multi_last_error_string": (C.c_char_p, [_vp]),
    "hpcx_multi_csr_create": (_i32, [_vp, _P(CsrDesc), _P(CsrOpts), _i32, _P(_vp)]),
    "hpcx_multi_csr_destroy": (_i32, [_vp]),
    "hpcx_multi_csr_bounds": (_i32, [_vp, _vp]),
    "hpcx_multi_spmv": (_i32, [_vp, _vp, _vp]),
    "hpcx_multi_grid_create": (_i32, [_vp, _P(GridDesc), _P(_vp)]),
    "hpcx_multi_grid_destroy": (_i32, [_vp]),
    "hpcx_multi_grid_upload": (_i32, [_vp, _vp]),
    "hpcx_multi_grid_download": (_i32, [_vp, _vp]),
    "hpcx_multi_stencil_sweep": (_i32, [_vp, _vp, _P(StencilDesc), _i32]),
    "hpcx_spmv_allgather": (_i32, [_vp, _P(_i64), _vp, _vp, _vp, _vp]),
    "hpcx_cg_create": (_i32, [_vp, _P(_vp)]),
    "hpcx_cg_destroy": (_i32, [_vp]),
    "hpcx_cg_solve": (_i32, [_vp, _vp, _i32, _vp, _i32, _i32, C.c_double, _i32, _P(CgResult), _vp]),
}

_LIB = None


def load() -> C.CDLL:
    """dlopen libhpcx.so (built in-tree by `python -m libhpc_b200.build`) and type it."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `python -m libhpc_b200.build` (there is no fallback path)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib
