"""libhpc_b200 — B200 (sm_100a) implementation of libHPC's data-parallel sparse hot path:
CSR SpMV y = A*x and ghost-cell stencil sweeps, behind the C ABI of include/hpcx.h.

The product is libhpcx.so (csrc/, built by `python -m libhpc_b200.build`); this package is
the host-side mirror of the reference's container interface over that ABI.
"""
from . import _abi  # noqa: F401
from .api import (CgSolver, Context, ManagedArray, MultiContext, MultiCsr, MultiGrid, multi_stencil_sweep, CsrMatrix, DeviceCsr, DeviceCsrArrays, DeviceGrid, HPCHighDimensionFlatArray,  # noqa: F401
                  HpcxError, conv2d, grid_halo_maps, csr_from_cells, csr_from_coo, csr_localize, grid_transpose, partition_rows, partition_slabs, spmv, stencil_sweep,
                  synth_field, synth_rmat_csr,
                  stencil_sweep_host)
from ._abi import (F32, F64, MEM_DEVICE, MEM_HOST, MEM_MANAGED, SPMV_AUTO, SPMV_HUB, SPMV_SELL, SPMV_STREAM, SPMV_VECTOR,  # noqa: F401
                   STENCIL_BOX_X, STENCIL_BOX_Y, STENCIL_JACOBI7, VARIANT_AUTO, VARIANT_DIRECT, VARIANT_TEMPORAL,
                   VARIANT_TMA)
