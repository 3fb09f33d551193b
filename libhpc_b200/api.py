"""Host-side mirror of the reference's container surface over the libhpcx C ABI.

Names and argument meaning follow the reference (Liupeter01/libHPC):

  HPCHighDimensionFlatArray  lib/hpc/include/HPCHighDimensionFlatArray.hpp:54-224
      ctor(dims...), data(), operator()(idx...) [unchecked], at({idx...}) [std::out_of_range]
  CsrMatrix / spmv           new next to lib/sparse (the reference has no CSR)
  stencil_sweep              replaces the BM_*_blur loop nests
                             (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368,444-457)

Error convention: status codes become the exception the reference throws:
HPCX_OUT_OF_RANGE -> IndexError (std::out_of_range), HPCX_OOM -> MemoryError (std::bad_alloc),
everything else -> HpcxError (std::system_error) carrying hpcx_last_error_string().

Every compute call goes to the CUDA library; nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from ._abi import (F32, F64, MEM_DEVICE, MEM_HOST, SPMV_AUTO, SPMV_HUB, SPMV_SELL, SPMV_STREAM, SPMV_VECTOR,
                   STENCIL_BOX_X, STENCIL_BOX_Y, STENCIL_JACOBI7, VARIANT_AUTO, VARIANT_DIRECT,
                   VARIANT_TMA)


class HpcxError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{_abi.STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


def _raise(status: int, message: str):
    if status == _abi.OUT_OF_RANGE:
        raise IndexError(message or "invalid index, out of boundary")
    if status == _abi.OOM:
        raise MemoryError(message)
    raise HpcxError(status, message)


def _dtype_code(dt) -> int:
    dt = np.dtype(dt)
    if dt == np.float32:
        return F32
    if dt == np.float64:
        return F64
    raise TypeError(f"unsupported dtype {dt}: libhpcx computes in float32 / float64")


def _ptr(buf):
    """(address, HPCX_MEM_*) of a numpy array, a torch tensor or a (ptr, mem) pair."""
    if buf is None:
        return None, MEM_HOST
    if isinstance(buf, np.ndarray):
        if not buf.flags["C_CONTIGUOUS"]:
            raise ValueError("host buffers must be C-contiguous")
        return buf.ctypes.data, MEM_HOST
    if isinstance(buf, tuple):
        return int(buf[0]), int(buf[1])
    if hasattr(buf, "data_ptr"):  # torch.Tensor
        if not buf.is_contiguous():
            raise ValueError("tensors must be contiguous")
        return buf.data_ptr(), (MEM_DEVICE if buf.is_cuda else MEM_HOST)
    raise TypeError(f"cannot take the address of {type(buf)}")


def _stream(stream):
    if stream is None:
        return None
    return int(getattr(stream, "cuda_stream", stream))


# ------------------------------------------------------------------ context
class Context:
    """One per process and device (hpcx_init)."""

    def __init__(self, device: int = 0):
        self.lib = _abi.load()
        h = C.c_void_p()
        st = self.lib.hpcx_init(device, C.byref(h))
        self.h = h
        if st != _abi.OK:
            msg = self.lib.hpcx_last_error_string(h).decode() if h else "hpcx_init failed"
            if h:
                self.lib.hpcx_finalize(h)
            self.h = None
            _raise(st, msg)
        self.device = device
        self.nranks, self.rank = 1, 0

    def check(self, st: int):
        if st != _abi.OK:
            _raise(st, self.lib.hpcx_last_error_string(self.h).decode())

    def synchronize(self):
        self.check(self.lib.hpcx_synchronize(self.h))

    def launch_count(self) -> int:
        return int(self.lib.hpcx_launch_count(self.h))

    def pinned(self, shape, dtype) -> np.ndarray:
        """numpy view of a cudaHostAlloc buffer (kept alive by the context)."""
        dt = np.dtype(dtype)
        n = int(np.prod(shape))
        p = C.c_void_p()
        self.check(self.lib.hpcx_host_alloc(self.h, max(n * dt.itemsize, 1), C.byref(p)))
        buf = (C.c_char * max(n * dt.itemsize, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dt, count=n).reshape(shape)
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        return arr

    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self.check(self.lib.hpcx_device_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def device_free(self, ptr: int):
        self.check(self.lib.hpcx_device_free(self.h, ptr))

    def memcpy(self, dst, src, nbytes: int):
        self.check(self.lib.hpcx_memcpy(self.h, _ptr(dst)[0], _ptr(src)[0], nbytes))

    def close(self):
        if self.h:
            for p in getattr(self, "_pinned", []):
                self.lib.hpcx_host_free(self.h, p)
            self._pinned = []
            self.lib.hpcx_comm_finalize(self.h)
            self.lib.hpcx_finalize(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class ManagedArray:
    """numpy view of cudaMallocManaged memory (hpcx_mem_alloc): the Python spelling of the
    reference's `std::vector<T, CudaAllocator<T, CudaMemManaged>>` (lib/gpu/util/include/
    cudaAllocator.hpp:39-83, radix_sort_gpu.h:22-23).  `.a` is host-addressable; `.arg` is the
    (pointer, HPCX_MEM_MANAGED) pair the library's entry points take."""

    def __init__(self, shape, dtype, mem: int = _abi.MEM_MANAGED):
        dt = np.dtype(dtype)
        n = int(np.prod(shape))
        self.mem, self.nbytes = mem, max(n * dt.itemsize, 1)
        p = C.c_void_p()
        st = _abi.load().hpcx_mem_alloc(mem, self.nbytes, C.byref(p))
        if st != _abi.OK:
            _raise(st, "hpcx_mem_alloc")
        self.ptr = p.value
        self.a = np.frombuffer((C.c_char * self.nbytes).from_address(self.ptr), dtype=dt, count=n).reshape(shape)

    @property
    def arg(self):
        return (self.ptr, self.mem)

    def free(self):
        if self.ptr:
            self.a = None
            _abi.load().hpcx_mem_free(self.mem, self.ptr)
            self.ptr = None


# ------------------------------------------------------------------ dense ghost-padded array
class HPCHighDimensionFlatArray:
    """hpc::HPCHighDimensionFlatArray<Dimension, T, Low_Bound, High_Bound> on the host.

    `HPCHighDimensionFlatArray(ny, nx, dtype=np.float32, lo=8)` is the reference's
    `HPCHighDimensionFlatArray<2, float, 8> a(ny, nx)` (High_Bound defaults to Low_Bound,
    HPCHighDimensionFlatArray.hpp:54-55).  Storage is the same row-major flat buffer,
    value-initialised (ghosts zero); index math runs through the C ABI so that it is the
    code the tests pin against the reference header.
    """

    def __init__(self, *dims, dtype=np.float32, lo: int = 0, hi: int | None = None, buffer=None):
        if not 1 <= len(dims) <= 3:
            raise ValueError("Dimension must be 1..3")
        self.dims = tuple(int(d) for d in dims)
        self.dtype = np.dtype(dtype)
        self.lo = int(lo)
        self.hi = self.lo if hi is None else int(hi)
        self.desc = _abi.GridDesc(len(dims), _dtype_code(dtype), self.lo, self.hi,
                                  (C.c_int64 * 3)(*self.dims))
        st = (C.c_int64 * 3)()
        tot = C.c_int64()
        rc = _abi.load().hpcx_grid_stride_total(C.byref(self.desc), st, C.byref(tot))
        if rc != _abi.OK:
            _raise(rc, "bad array shape")
        self.stride = tuple(st[i] for i in range(len(dims)))
        self.total = int(tot.value)
        if self.total <= 0:  # resize() asserts total > 0 (…:138)
            raise ValueError("resize() attempted to create zero-sized flat array!")
        self._flat = np.zeros(self.total, dtype=self.dtype) if buffer is None else buffer
        assert self._flat.size == self.total

    def data(self) -> np.ndarray:
        return self._flat

    def _lin(self, idx, checked: int) -> int:
        if len(idx) != len(self.dims):
            raise TypeError("index rank mismatch")
        off = C.c_int64()
        rc = _abi.load().hpcx_grid_linearize(C.byref(self.desc), (C.c_int64 * 3)(*idx), checked,
                                             C.byref(off))
        if rc != _abi.OK:
            _raise(rc, "invalid index, out of boundary")
        return int(off.value)

    def __call__(self, *idx):  # operator(): unchecked
        return self._flat[self._lin(idx, 0)]

    def set(self, idx, value):
        self._flat[self._lin(tuple(idx), 0)] = value

    def at(self, idx):  # bounds-checked, raises IndexError where the reference throws
        return self._flat[self._lin(tuple(idx), 1)]

    def padded_view(self) -> np.ndarray:
        """N-d view over the padded extents (ghosts included)."""
        return self._flat.reshape(tuple(d + self.lo + self.hi for d in self.dims))

    def interior(self) -> np.ndarray:
        v = self.padded_view()
        return v[tuple(slice(self.lo, self.lo + d) for d in self.dims)]


class DeviceGrid:
    """Device mirror of an HPCHighDimensionFlatArray (hpcx_grid_*)."""

    def __init__(self, ctx: Context, dims, dtype=np.float32, lo: int = 0, hi: int | None = None):
        self.ctx = ctx
        self.dims = tuple(int(d) for d in dims)
        self.dtype = np.dtype(dtype)
        self.lo = int(lo)
        self.hi = self.lo if hi is None else int(hi)
        self.desc = _abi.GridDesc(len(dims), _dtype_code(dtype), self.lo, self.hi,
                                  (C.c_int64 * 3)(*self.dims))
        h = C.c_void_p()
        ctx.check(ctx.lib.hpcx_grid_create(ctx.h, C.byref(self.desc), C.byref(h)))
        self.h = h
        self.total_host = int(np.prod([d + self.lo + self.hi for d in self.dims]))

    @classmethod
    def from_host(cls, ctx: Context, a: HPCHighDimensionFlatArray) -> "DeviceGrid":
        g = cls(ctx, a.dims, a.dtype, a.lo, a.hi)
        g.upload(a.data())
        return g

    def upload(self, host_flat, stream=None):
        p, mem = _ptr(host_flat)
        assert mem == MEM_HOST and host_flat.size == self.total_host and host_flat.dtype == self.dtype
        self.ctx.check(self.ctx.lib.hpcx_grid_upload(self.h, p, _stream(stream)))

    def download(self, out: np.ndarray | None = None, stream=None) -> np.ndarray:
        if out is None:
            out = np.empty(self.total_host, dtype=self.dtype)
        assert out.size == self.total_host and out.dtype == self.dtype
        self.ctx.check(self.ctx.lib.hpcx_grid_download(self.h, _ptr(out)[0], _stream(stream)))
        return out

    def layout(self):
        xoff, pitch, total = C.c_int64(), C.c_int64(), C.c_int64()
        p = C.c_void_p()
        self.ctx.check(self.ctx.lib.hpcx_grid_device_layout(self.h, C.byref(xoff), C.byref(pitch),
                                                            C.byref(total), C.byref(p)))
        return dict(xoff=xoff.value, pitch=pitch.value, total=total.value, ptr=p.value)

    def face_elems(self, axis: int, width: int) -> int:
        n = C.c_int64()
        self.ctx.check(self.ctx.lib.hpcx_grid_face_elems(self.h, axis, width, C.byref(n)))
        return int(n.value)

    def pack_face(self, axis: int, side: int, width: int, buf=None, stream=None):
        if buf is None:
            buf = np.empty(self.face_elems(axis, width), dtype=self.dtype)
        p, mem = _ptr(buf)
        self.ctx.check(self.ctx.lib.hpcx_grid_pack_face(self.h, axis, side, width, p, mem,
                                                        _stream(stream)))
        return buf

    def unpack_face(self, axis: int, side: int, width: int, buf, stream=None):
        p, mem = _ptr(buf)
        self.ctx.check(self.ctx.lib.hpcx_grid_unpack_face(self.h, axis, side, width, p, mem,
                                                          _stream(stream)))

    def destroy(self):
        if self.h:
            self.ctx.lib.hpcx_grid_destroy(self.h)
            self.h = None


def grid_halo_maps(dims, lo: int, hi: int, axis: int, side: int, width: int, dtype=np.float32):
    """hpcx_grid_export_halo_maps: (run_len, send_start[], recv_start[]) of one face, flat offsets in
    the reference layout."""
    desc = _abi.GridDesc(len(dims), _dtype_code(dtype), lo, hi, (C.c_int64 * 3)(*[int(d) for d in dims]))
    nruns, run_len = C.c_int64(), C.c_int64()
    lib = _abi.load()
    rc = lib.hpcx_grid_export_halo_maps(C.byref(desc), axis, side, width, C.byref(nruns), C.byref(run_len), None, None)
    if rc != _abi.OK:
        _raise(rc, "hpcx_grid_export_halo_maps")
    send, recv = np.empty(nruns.value, dtype=np.int64), np.empty(nruns.value, dtype=np.int64)
    rc = lib.hpcx_grid_export_halo_maps(C.byref(desc), axis, side, width, C.byref(nruns), C.byref(run_len),
                                        send.ctypes.data, recv.ctypes.data)
    if rc != _abi.OK:
        _raise(rc, "hpcx_grid_export_halo_maps")
    return int(run_len.value), send, recv


def stencil_desc(kind: int, radius: int = 0, variant: int = VARIANT_AUTO, zsplit: int = 0):
    return _abi.StencilDesc(kind, radius, variant, zsplit)


def stencil_sweep(gin: DeviceGrid, gout: DeviceGrid, kind: int, radius: int = 0, steps: int = 1,
                  variant: int = VARIANT_AUTO, zsplit: int = 0, stream=None):
    """hpcx_stencil_sweep on device-resident grids (see include/hpcx.h for the contract)."""
    d = stencil_desc(kind, radius, variant, zsplit)
    gin.ctx.check(gin.ctx.lib.hpcx_stencil_sweep(gin.h, gout.h, C.byref(d), steps, _stream(stream)))


def conv2d(gin: DeviceGrid, gout: DeviceGrid, weights: np.ndarray, variant: int = VARIANT_AUTO, stream=None):
    """hpcx_conv2d: gout(y,x) += sum_{l,k} gin(y+l, x+k) * weights[l,k] (BM_conv,
    test_hpc_benchmark.cpp:793-806); weights is a host (kh, kw) array of the grids' dtype."""
    w = np.ascontiguousarray(weights, dtype=gin.dtype)
    if w.ndim != 2:
        raise ValueError("weights must be a (kh, kw) array")
    gin.ctx.check(gin.ctx.lib.hpcx_conv2d(gin.h, gout.h, w.ctypes.data, w.shape[0], w.shape[1], variant,
                                          _stream(stream)))


def grid_transpose(gin: DeviceGrid, gout: DeviceGrid, stream=None):
    """hpcx_grid_transpose: gout(x, y) = gin(y, x) (BM_transpose, test_hpc_benchmark.cpp:672-682)."""
    gin.ctx.check(gin.ctx.lib.hpcx_grid_transpose(gin.h, gout.h, _stream(stream)))


def stencil_sweep_host(ctx: Context, a: HPCHighDimensionFlatArray, b: HPCHighDimensionFlatArray,
                       kind: int, radius: int = 0, steps: int = 1, variant: int = VARIANT_AUTO):
    """The reference-facing call: host containers in, host containers out.

    Box sweeps: `b` receives the sweep of `a` (BM_x_blur / BM_y_blur).  JACOBI7: both hold the
    same ghosts; after the call `b` holds the newest iterate and `a` the one before it.
    """
    ga, gb = DeviceGrid.from_host(ctx, a), DeviceGrid.from_host(ctx, b)
    try:
        stencil_sweep(ga, gb, kind, radius, steps, variant)
        gb.download(b.data())
        if kind == STENCIL_JACOBI7 and steps > 1:
            ga.download(a.data())
    finally:
        ga.destroy()
        gb.destroy()


# ------------------------------------------------------------------ CSR matrix + SpMV
class CsrMatrix:
    """sparse::CsrMatrix<T, Index> on the host: row_ptr[N+1] (int32 or int64), col_idx[nnz]
    (int32, ascending within a row), values[nnz]."""

    def __init__(self, row_ptr, col_idx, values, ncols: int | None = None):
        self.row_ptr = np.ascontiguousarray(row_ptr)
        if self.row_ptr.dtype not in (np.int32, np.int64):
            raise TypeError("row_ptr must be int32 or int64")
        self.col_idx = np.ascontiguousarray(col_idx, dtype=np.int32)
        self.values = np.ascontiguousarray(values)
        _dtype_code(self.values.dtype)
        self.nrows = len(self.row_ptr) - 1
        self.ncols = self.nrows if ncols is None else int(ncols)
        self.nnz = len(self.col_idx)
        if len(self.values) != self.nnz:
            raise ValueError("col_idx and values differ in length")

    @property
    def dtype(self):
        return self.values.dtype

    def algorithmic_bytes(self) -> int:
        """nnz*(sizeof T + 4) + N*(4 + 2*sizeof T) + 4."""
        es = self.values.dtype.itemsize
        return self.nnz * (es + 4) + self.nrows * (4 + 2 * es) + 4


class DeviceCsr:
    """Device-resident matrix behind hpcx_csr_create."""

    def __init__(self, ctx: Context, A: CsrMatrix, kernel: int = SPMV_AUTO, tile_nnz: int = 0,
                 sell_c: int = 0, sell_sigma: int = 0, lanes_per_row: int = 0, short_row_max: int = 0):
        self.ctx, self.host = ctx, A
        d = _abi.CsrDesc(A.nrows, A.ncols, A.nnz, A.row_ptr.ctypes.data,
                         A.row_ptr.dtype.itemsize, _dtype_code(A.dtype),
                         A.col_idx.ctypes.data if A.nnz else None,
                         A.values.ctypes.data if A.nnz else None, getattr(A, "mem", MEM_HOST), 0)
        o = _abi.CsrOpts(kernel, tile_nnz, sell_c, sell_sigma, lanes_per_row, short_row_max)
        h = C.c_void_p()
        ctx.check(ctx.lib.hpcx_csr_create(ctx.h, C.byref(d), C.byref(o), C.byref(h)))
        self.h = h
        self.nrows, self.ncols, self.nnz, self.dtype = A.nrows, A.ncols, A.nnz, A.dtype
        self.kernel = kernel
        self._rp_dtype = A.row_ptr.dtype

    @classmethod
    def from_device_arrays(cls, ctx: Context, arrays: "DeviceCsrArrays", ncols: int | None = None,
                           kernel: int = SPMV_AUTO, tile_nnz: int = 0, sell_c: int = 0,
                           sell_sigma: int = 0, lanes_per_row: int = 0, short_row_max: int = 0) -> "DeviceCsr":
        """Matrix whose CSR arrays already live on the device (csr_from_coo)."""
        self = cls.__new__(cls)
        self.ctx, self.host = ctx, None
        ncols = arrays.nrows if ncols is None else int(ncols)
        d = _abi.CsrDesc(arrays.nrows, ncols, arrays.nnz, arrays.row_ptr, arrays.idx_bytes,
                         _dtype_code(arrays.dtype), arrays.col_idx, arrays.values, MEM_DEVICE, 0)
        o = _abi.CsrOpts(kernel, tile_nnz, sell_c, sell_sigma, lanes_per_row, short_row_max)
        h = C.c_void_p()
        ctx.check(ctx.lib.hpcx_csr_create(ctx.h, C.byref(d), C.byref(o), C.byref(h)))
        self.h = h
        self.nrows, self.ncols, self.nnz, self.dtype = arrays.nrows, ncols, arrays.nnz, arrays.dtype
        self.kernel = kernel
        self._rp_dtype = np.dtype(np.int32 if arrays.idx_bytes == 4 else np.int64)
        return self

    def algorithmic_bytes(self) -> int:
        es = np.dtype(self.dtype).itemsize
        return self.nnz * (es + 4) + self.nrows * (4 + 2 * es) + 4

    def spmv(self, x, y, stream=None):
        """y = A*x; x and y may be numpy arrays (host) or CUDA tensors / (ptr, mem) pairs."""
        px, mx = _ptr(x)
        py, my = _ptr(y)
        self.ctx.check(self.ctx.lib.hpcx_spmv(self.h, px, mx, py, my, _stream(stream)))
        return y

    def download(self):
        rp = np.empty(self.nrows + 1, dtype=self._rp_dtype)
        col = np.empty(self.nnz, dtype=np.int32)
        val = np.empty(self.nnz, dtype=self.dtype)
        self.ctx.check(self.ctx.lib.hpcx_csr_download(self.h, rp.ctypes.data, col.ctypes.data,
                                                      val.ctypes.data))
        return rp, col, val

    def rowblocks(self) -> np.ndarray:
        n = C.c_int64()
        self.ctx.check(self.ctx.lib.hpcx_csr_rowblocks(self.h, C.byref(n), None, 0))
        out = np.empty((n.value, 6), dtype=np.int64)
        self.ctx.check(self.ctx.lib.hpcx_csr_rowblocks(self.h, C.byref(n), out.ctypes.data, n.value))
        return out

    def row_bins(self) -> np.ndarray:
        b = np.zeros(34, dtype=np.int64)
        self.ctx.check(self.ctx.lib.hpcx_csr_row_bins(self.h, b.ctypes.data))
        return b

    def sell(self):
        ns, pad = C.c_int64(), C.c_int64()
        c, sg = C.c_int32(), C.c_int32()
        self.ctx.check(self.ctx.lib.hpcx_csr_sell_info(self.h, C.byref(ns), C.byref(pad),
                                                       C.byref(c), C.byref(sg)))
        perm = np.empty(self.nrows, dtype=np.int32)
        sptr = np.empty(ns.value + 1, dtype=np.int64)
        slen = np.empty(ns.value, dtype=np.int32)
        col = np.empty(pad.value, dtype=np.int32)
        val = np.empty(pad.value, dtype=self.dtype)
        self.ctx.check(self.ctx.lib.hpcx_csr_sell_download(
            self.h, perm.ctypes.data, sptr.ctypes.data, slen.ctypes.data, col.ctypes.data,
            val.ctypes.data))
        return dict(c=c.value, sigma=sg.value, perm=perm, slice_ptr=sptr, slice_len=slen, col=col,
                    val=val)

    # distributed
    def set_halo(self, loc: dict):
        self.ctx.check(self.ctx.lib.hpcx_csr_set_halo(
            self.h, loc["nloc"], loc["nhalo"], loc["recv_counts"].ctypes.data,
            loc["send_counts"].ctypes.data, loc["send_idx"].ctypes.data if loc["nsend"] else None,
            loc["row_kind"].ctypes.data))

    def halo_transport(self, transport: int):
        """0 = NCCL send/recv, 1 = fused peer stores + flags (collective; after set_halo)."""
        self.ctx.check(self.ctx.lib.hpcx_csr_halo_transport(self.h, transport))

    def x_buffer(self) -> int:
        p = C.c_void_p()
        self.ctx.check(self.ctx.lib.hpcx_csr_x_buffer(self.h, C.byref(p)))
        return p.value

    def spmv_dist(self, x_local, y_local, stream=None):
        self.ctx.check(self.ctx.lib.hpcx_spmv_dist(self.h, _ptr(x_local)[0], _ptr(y_local)[0],
                                                   _stream(stream)))

    def spmv_allgather(self, bounds, x_local, x_full, y_local, stream=None):
        """hpcx_spmv_allgather: this handle is a row block with GLOBAL column ids; every rank's slice
        of x (bounds, world + 1 entries) is gathered into x_full, then the block's kernel runs."""
        b = np.ascontiguousarray(bounds, dtype=np.int64)
        self.ctx.check(self.ctx.lib.hpcx_spmv_allgather(self.h, b.ctypes.data_as(C.POINTER(C.c_int64)),
                                                        _ptr(x_local)[0], _ptr(x_full)[0], _ptr(y_local)[0],
                                                        _stream(stream)))

    def spmv_dist_host(self, x_host, y_host, stream=None):
        """hpcx_spmv_dist_host: host x / y (this rank's owned entries), pipelined against PCIe."""
        px, mx = _ptr(x_host)
        py, my = _ptr(y_host)
        assert mx == MEM_HOST and my == MEM_HOST
        self.ctx.check(self.ctx.lib.hpcx_spmv_dist_host(self.h, px, py, _stream(stream)))

    def destroy(self):
        if self.h:
            self.ctx.lib.hpcx_csr_destroy(self.h)
            self.h = None


class DeviceCsrArrays:
    """CSR arrays in device memory produced by hpcx_csr_from_coo (owned here until destroy())."""

    def __init__(self, ctx: Context, nrows: int, idx_bytes: int, dtype, raw: "_abi.CsrArrays"):
        self.ctx, self.nrows, self.idx_bytes, self.dtype = ctx, int(nrows), idx_bytes, np.dtype(dtype)
        self.nnz, self.row_ptr, self.col_idx, self.values = raw.nnz, raw.row_ptr, raw.col_idx, raw.values

    def download(self):
        rp = np.empty(self.nrows + 1, dtype=np.int32 if self.idx_bytes == 4 else np.int64)
        col = np.empty(self.nnz, dtype=np.int32)
        val = np.empty(self.nnz, dtype=self.dtype)
        self.ctx.memcpy(rp, (self.row_ptr, MEM_DEVICE), rp.nbytes)
        if self.nnz:
            self.ctx.memcpy(col, (self.col_idx, MEM_DEVICE), col.nbytes)
            self.ctx.memcpy(val, (self.values, MEM_DEVICE), val.nbytes)
        return rp, col, val

    def destroy(self):
        for name in ("row_ptr", "col_idx", "values"):
            p = getattr(self, name)
            if p:
                self.ctx.device_free(p)
                setattr(self, name, None)


def csr_from_coo(ctx: Context, nrows: int, rows, cols, vals, idx_bytes: int = 4,
                 nedges: int | None = None, dtype=None) -> DeviceCsrArrays:
    """Device assembly of a CSR matrix from an edge list (hpcx_csr_from_coo): sort by
    (row, col, list position), merge duplicates by adding in list order.  rows/cols/vals are
    numpy arrays (host), CUDA tensors or (ptr, mem) pairs (then give nedges and dtype)."""
    pr, mr = _ptr(rows)
    pc, mc = _ptr(cols)
    pv, mv = _ptr(vals)
    if not (mr == mc == mv):
        raise ValueError("rows, cols and vals must live in the same memory space")
    if isinstance(rows, np.ndarray) and (rows.dtype != np.int32 or cols.dtype != np.int32):
        raise TypeError("rows and cols must be int32")
    if nedges is None:
        nedges = len(rows)
    if dtype is None:
        dtype = vals.dtype if isinstance(vals, np.ndarray) else {"torch.float32": np.float32,
                                                                  "torch.float64": np.float64}[str(vals.dtype)]
    raw = _abi.CsrArrays()
    ctx.check(ctx.lib.hpcx_csr_from_coo(ctx.h, nrows, nedges, pr, pc, pv, _dtype_code(dtype), mr,
                                        idx_bytes, C.byref(raw)))
    return DeviceCsrArrays(ctx, nrows, idx_bytes, dtype, raw)


def csr_from_cells(ctx: Context, xs, ys, dtype=np.float64, idx_bytes: int = 4, ncells: int | None = None):
    """Sparse-grid bridge (hpcx_csr_from_cells): the 5-point operator over the active cells
    (x, y) of a sparse 2-D grid — what sparse::RootGrid::foreach enumerates.  Returns
    (DeviceCsrArrays, map) with map = dict(nrows, cell_row, row_x, row_y) as host int32 arrays."""
    px, mx = _ptr(xs)
    py, my = _ptr(ys)
    if mx != my:
        raise ValueError("xs and ys must live in the same memory space")
    if isinstance(xs, np.ndarray) and (xs.dtype != np.int32 or ys.dtype != np.int32):
        raise TypeError("xs and ys must be int32")
    n = len(xs) if ncells is None else int(ncells)
    raw, cm = _abi.CsrArrays(), _abi.CellMap()
    ctx.check(ctx.lib.hpcx_csr_from_cells(ctx.h, n, px, py, mx, _dtype_code(dtype), idx_bytes,
                                          C.byref(raw), C.byref(cm)))
    arrays = DeviceCsrArrays(ctx, cm.nrows, idx_bytes, dtype, raw)
    out = dict(nrows=int(cm.nrows))
    for name, count in (("cell_row", n), ("row_x", cm.nrows), ("row_y", cm.nrows)):
        a = np.empty(count, dtype=np.int32)
        p = getattr(cm, name)
        if count:
            ctx.memcpy(a, (p, MEM_DEVICE), a.nbytes)
        if p:
            ctx.device_free(p)
        out[name] = a
    return arrays, out


def synth_rmat_csr(ctx: Context, scale: int, edge_factor: int = 16, seed: int = 24,
                   idx_bytes: int = 4) -> DeviceCsrArrays:
    """C5 on the device: R-MAT edges (hpcx_synth_rmat_coo) assembled by hpcx_csr_from_coo.
    Bit-identical to synth.rmat / the oracle's orc_rmat_csr."""
    ne = (1 << scale) * edge_factor
    r, c, v = ctx.device_alloc(4 * ne), ctx.device_alloc(4 * ne), ctx.device_alloc(4 * ne)
    try:
        ctx.check(ctx.lib.hpcx_synth_rmat_coo(ctx.h, scale, ne, seed, r, c, v))
        return csr_from_coo(ctx, 1 << scale, (r, MEM_DEVICE), (c, MEM_DEVICE), (v, MEM_DEVICE),
                            idx_bytes, nedges=ne, dtype=np.float32)
    finally:
        for p in (r, c, v):
            ctx.device_free(p)


def synth_field(ctx: Context, seed: int, n: int, out, dtype, first: int = 0):
    """out[i] = 2u(seed, first+i) - 1 generated on the device (out: CUDA tensor or (ptr, mem))."""
    ctx.check(ctx.lib.hpcx_synth_field(ctx.h, seed, first, n, _dtype_code(dtype), _ptr(out)[0]))
    return out


class CgSolver:
    """Conjugate gradients around a DeviceCsr (hpcx_cg_*): fused SpMV + dot, device scalars."""

    def __init__(self, A: DeviceCsr):
        self.A, self.ctx = A, A.ctx
        h = C.c_void_p()
        self.ctx.check(self.ctx.lib.hpcx_cg_create(A.h, C.byref(h)))
        self.h = h

    def solve(self, b, x, max_iters: int, rel_tol: float = 0.0, check_every: int = 0, stream=None,
              want_result: bool = True):
        """x is the start vector and receives the iterate; returns a dict (or None)."""
        pb, mb = _ptr(b)
        px, mx = _ptr(x)
        res = _abi.CgResult()
        self.ctx.check(self.ctx.lib.hpcx_cg_solve(self.h, pb, mb, px, mx, max_iters, rel_tol, check_every,
                                                  C.byref(res) if want_result else None, _stream(stream)))
        if not want_result:
            return None
        return dict(iterations=res.iterations, converged=res.converged == 1, breakdown=res.converged < 0,
                    rr=res.rr, bb=res.bb)

    def destroy(self):
        if self.h:
            self.ctx.lib.hpcx_cg_destroy(self.h)
            self.h = None


def spmv(ctx: Context, A: CsrMatrix, x: np.ndarray, y: np.ndarray | None = None, **opts):
    """The reference-facing free function sparse::spmv(A, x, y) with host containers.
    The device mirror of A is cached on the host matrix (keyed by context and options)."""
    key = (id(ctx), tuple(sorted(opts.items())))
    cache = A.__dict__.setdefault("_device", {})
    if key not in cache:
        cache[key] = DeviceCsr(ctx, A, **opts)
    if y is None:
        y = np.empty(A.nrows, dtype=A.dtype)
    return cache[key].spmv(np.ascontiguousarray(x, dtype=A.dtype), y)


# ------------------------------------------------------------------ partitioner (host only)
def partition_rows(nrows: int, row_ptr, nparts: int, mode: int = 0) -> np.ndarray:
    b = np.empty(nparts + 1, dtype=np.int64)
    rp = None if row_ptr is None else np.ascontiguousarray(row_ptr)
    rc = _abi.load().hpcx_partition_rows(nrows, None if rp is None else rp.ctypes.data,
                                         4 if rp is None else rp.dtype.itemsize, nparts, mode,
                                         b.ctypes.data)
    if rc != _abi.OK:
        _raise(rc, "hpcx_partition_rows")
    return b


def partition_slabs(n: int, nparts: int) -> np.ndarray:
    b = np.empty(nparts + 1, dtype=np.int64)
    rc = _abi.load().hpcx_partition_slabs(n, nparts, b.ctypes.data)
    if rc != _abi.OK:
        _raise(rc, "hpcx_partition_slabs")
    return b


def csr_localize(A: CsrMatrix, bounds, rank: int) -> dict:
    """hpcx_csr_localize with the two-call protocol; returns the tables as numpy arrays."""
    lib = _abi.load()
    bounds = np.ascontiguousarray(bounds, dtype=np.int64)
    P = len(bounds) - 1
    sz = _abi.HaloSizes()
    head = (A.nrows, A.row_ptr.ctypes.data, A.row_ptr.dtype.itemsize,
            A.col_idx.ctypes.data if A.nnz else None, bounds.ctypes.data, P, rank, C.byref(sz))
    rc = lib.hpcx_csr_localize(*head, None, None, None, None, None, None, None)
    if rc != _abi.OK:
        _raise(rc, "hpcx_csr_localize")
    out = dict(
        local_row_ptr=np.empty(sz.nloc + 1, dtype=np.int64),
        local_col=np.empty(sz.nnz_loc, dtype=np.int32),
        halo_gcol=np.empty(sz.nhalo, dtype=np.int64),
        recv_counts=np.empty(P, dtype=np.int64), send_counts=np.empty(P, dtype=np.int64),
        send_idx=np.empty(sz.nsend, dtype=np.int32), row_kind=np.empty(sz.nloc, dtype=np.uint8))
    rc = lib.hpcx_csr_localize(*head, *[out[k].ctypes.data for k in (
        "local_row_ptr", "local_col", "halo_gcol", "recv_counts", "send_counts", "send_idx",
        "row_kind")])
    if rc != _abi.OK:
        _raise(rc, "hpcx_csr_localize")
    out.update(nloc=int(sz.nloc), nnz_loc=int(sz.nnz_loc), nhalo=int(sz.nhalo), nsend=int(sz.nsend))
    p0 = int(A.row_ptr[bounds[rank]])
    out["values"] = A.values[p0:p0 + out["nnz_loc"]]
    return out


# ------------------------------------------------------------------ one process, several GPUs
class MultiContext:
    """hpcx_init_multi: N GPUs driven from this one process / thread (peer access, no NCCL, no IPC).
    devices=None takes every visible GPU."""

    def __init__(self, devices=None):
        self.lib = _abi.load()
        if devices is None:
            devices = list(range(max(int(self.lib.hpcx_device_count()), 0)))
        self.devices = [int(d) for d in devices]
        ids = (C.c_int32 * len(self.devices))(*self.devices)
        h = C.c_void_p()
        st = self.lib.hpcx_init_multi(len(self.devices), ids, C.byref(h))
        self.h = h
        if st != _abi.OK:
            msg = self.lib.hpcx_multi_last_error_string(h).decode() if h else "hpcx_init_multi failed"
            if h:
                self.lib.hpcx_multi_finalize(h)
            self.h = None
            _raise(st, msg)

    @property
    def size(self) -> int:
        return int(self.lib.hpcx_multi_size(self.h))

    def check(self, st: int):
        if st != _abi.OK:
            _raise(st, self.lib.hpcx_multi_last_error_string(self.h).decode())

    def close(self):
        if self.h:
            self.lib.hpcx_multi_finalize(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class MultiCsr:
    """A host CsrMatrix partitioned over the devices of a MultiContext (hpcx_multi_csr_create)."""

    def __init__(self, mctx: MultiContext, A: CsrMatrix, mode: int = 0, kernel: int = SPMV_AUTO, tile_nnz: int = 0):
        self.m, self.host = mctx, A
        d = _abi.CsrDesc(A.nrows, A.ncols, A.nnz, A.row_ptr.ctypes.data, A.row_ptr.dtype.itemsize,
                         _dtype_code(A.dtype), A.col_idx.ctypes.data if A.nnz else None,
                         A.values.ctypes.data if A.nnz else None, MEM_HOST, 0)
        o = _abi.CsrOpts(kernel, tile_nnz, 0, 0, 0, 0)
        h = C.c_void_p()
        mctx.check(mctx.lib.hpcx_multi_csr_create(mctx.h, C.byref(d), C.byref(o), mode, C.byref(h)))
        self.h = h

    def bounds(self) -> np.ndarray:
        b = np.empty(self.m.size + 1, dtype=np.int64)
        self.m.check(self.m.lib.hpcx_multi_csr_bounds(self.h, b.ctypes.data))
        return b

    def spmv(self, x: np.ndarray, y: np.ndarray | None = None) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=self.host.dtype)
        if y is None:
            y = np.empty(self.host.nrows, dtype=self.host.dtype)
        self.m.check(self.m.lib.hpcx_multi_spmv(self.h, x.ctypes.data, y.ctypes.data))
        return y

    def destroy(self):
        if self.h:
            self.m.lib.hpcx_multi_csr_destroy(self.h)
            self.h = None


class MultiGrid:
    """A ghost-padded grid split along its slowest axis over the devices (hpcx_multi_grid_*)."""

    def __init__(self, mctx: MultiContext, dims, dtype=np.float32, lo: int = 0, hi: int | None = None):
        self.m = mctx
        self.dims = tuple(int(d) for d in dims)
        self.dtype = np.dtype(dtype)
        self.lo = int(lo)
        self.hi = self.lo if hi is None else int(hi)
        self.desc = _abi.GridDesc(len(dims), _dtype_code(dtype), self.lo, self.hi, (C.c_int64 * 3)(*self.dims))
        h = C.c_void_p()
        mctx.check(mctx.lib.hpcx_multi_grid_create(mctx.h, C.byref(self.desc), C.byref(h)))
        self.h = h
        self.total_host = int(np.prod([d + self.lo + self.hi for d in self.dims]))

    def upload(self, host_flat: np.ndarray):
        assert host_flat.size == self.total_host and host_flat.dtype == self.dtype and host_flat.flags["C_CONTIGUOUS"]
        self.m.check(self.m.lib.hpcx_multi_grid_upload(self.h, host_flat.ctypes.data))

    def download(self, out: np.ndarray) -> np.ndarray:
        """Writes the owned slices into `out` (the global flat array; its ghosts are left as they are)."""
        assert out.size == self.total_host and out.dtype == self.dtype and out.flags["C_CONTIGUOUS"]
        self.m.check(self.m.lib.hpcx_multi_grid_download(self.h, out.ctypes.data))
        return out

    def destroy(self):
        if self.h:
            self.m.lib.hpcx_multi_grid_destroy(self.h)
            self.h = None


def multi_stencil_sweep(gin: MultiGrid, gout: MultiGrid, kind: int, radius: int = 0, steps: int = 1,
                        variant: int = VARIANT_AUTO, zsplit: int = 0):
    d = stencil_desc(kind, radius, variant, zsplit)
    gin.m.check(gin.m.lib.hpcx_multi_stencil_sweep(gin.h, gout.h, C.byref(d), steps))
