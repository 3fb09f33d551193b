// grid.cu — ghost-padded grid: host index math, device layout, upload/download,
// face pack/unpack.  Replaces (device side) hpc::HPCHighDimensionFlatArray storage,
// reference lib/hpc/include/HPCHighDimensionFlatArray.hpp:135-208.
#include "grid.cuh"

using namespace hpcx;

static bool desc_ok(const hpcx_grid_desc *d) {
  if (!d || d->ndim < 1 || d->ndim > 3) return false;
  if (d->dtype != HPCX_F32 && d->dtype != HPCX_F64) return false;
  if (d->lo < 0 || d->hi < 0) return false;
  for (int i = 0; i < d->ndim; ++i)
    if (d->dim[i] < 0) return false;
  return true;
}

extern "C" {

// compute_stride_and_total, HPCHighDimensionFlatArray.hpp:161-171: walk dimensions from
// the fastest to the slowest, stride = running product of padded extents.
hpcx_status hpcx_grid_stride_total(const hpcx_grid_desc *d, int64_t *stride, int64_t *total) {
  if (!desc_ok(d)) return HPCX_INVALID_ARG;
  int64_t run = 1;
  for (int i = d->ndim; i-- > 0;) {
    if (stride) stride[i] = run;
    run *= d->dim[i] + d->lo + d->hi;
  }
  if (total) *total = run;
  return HPCX_OK;
}

// unsafe_linearize (…:180-187) / safe_linearize (…:197-208).
hpcx_status hpcx_grid_linearize(const hpcx_grid_desc *d, const int64_t *idx, int32_t checked,
                                int64_t *offset) {
  if (!desc_ok(d) || !idx || !offset) return HPCX_INVALID_ARG;
  int64_t stride[3], total;
  hpcx_grid_stride_total(d, stride, &total);
  int64_t off = 0;
  for (int i = d->ndim; i-- > 0;) {
    if (checked && (idx[i] < -(int64_t)d->lo || idx[i] >= d->dim[i] + d->hi))
      return HPCX_OUT_OF_RANGE;
    off += stride[i] * (idx[i] + d->lo);
  }
  *offset = off;
  return HPCX_OK;
}

// Halo index map of one face in the reference layout.  With the padded extents e[i] and strides
// st[i] of compute_stride_and_total, the `width` slices of axis a starting at padded coordinate c0
// are, for every combination of the slower coordinates, one run of width * st[a] consecutive
// elements starting at sum_{i<a} idx[i]*st[i] + c0*st[a]; runs are listed with the slower
// coordinates in row-major order.  Send = interior slices next to the side, recv = its ghost slices.
hpcx_status hpcx_grid_export_halo_maps(const hpcx_grid_desc *d, int32_t axis, int32_t side, int32_t width,
                                       int64_t *nruns, int64_t *run_len, int64_t *send_start,
                                       int64_t *recv_start) {
  if (!desc_ok(d) || axis < 0 || axis >= d->ndim || (side != 0 && side != 1) || !nruns || !run_len)
    return HPCX_INVALID_ARG;
  if (width < 1 || width > d->dim[axis] || width > (side ? d->hi : d->lo)) return HPCX_INVALID_ARG;
  int64_t st[3], total;
  hpcx_grid_stride_total(d, st, &total);
  int64_t runs = 1;
  for (int i = 0; i < axis; ++i) runs *= d->dim[i] + d->lo + d->hi;
  *nruns = runs;
  *run_len = (int64_t)width * st[axis];
  const int64_t send_c0 = side ? d->lo + d->dim[axis] - width : d->lo;  // padded coordinate of the first slice
  const int64_t recv_c0 = side ? d->lo + d->dim[axis] : d->lo - width;
  const int64_t outer = axis > 0 ? st[axis - 1] : 0;  // stride of one step of the combined slower coordinates
  for (int64_t k = 0; k < runs; ++k) {
    if (send_start) send_start[k] = k * outer + send_c0 * st[axis];
    if (recv_start) recv_start[k] = k * outer + recv_c0 * st[axis];
  }
  return HPCX_OK;
}

hpcx_status hpcx_grid_create(hpcx_ctx ctx, const hpcx_grid_desc *d, hpcx_grid *out) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  *out = nullptr;
  HPCX_REQUIRE(ctx, desc_ok(d), "bad grid descriptor");
  hpcx_grid g = new (std::nothrow) hpcx_grid_s();
  if (!g) return fail(ctx, HPCX_OOM, "host allocation failed");
  g->ctx = ctx;
  g->desc = *d;
  g->es = (int)dtype_size(d->dtype);
  const int shift = 3 - d->ndim;
  for (int i = 0; i < d->ndim; ++i) {
    g->n[i + shift] = d->dim[i];
    g->glo[i + shift] = d->lo;
    g->ghi[i + shift] = d->hi;
  }
  for (int i = 0; i < 3; ++i) g->ext[i] = g->n[i] + g->glo[i] + g->ghi[i];
  // resize() asserts total > 0 (HPCHighDimensionFlatArray.hpp:138).
  if (g->ext[0] * g->ext[1] * g->ext[2] <= 0) {
    delete g;
    return fail(ctx, HPCX_INVALID_ARG, "zero-sized flat array");
  }
  // Dense rows: interior x = 0 starts on a 32-byte sector boundary and the pitch is a multiple
  // of 32 bytes (TMA needs 16-byte strides; float4 / double2 stores stay sector aligned).
  // Padding rows out to 128-byte lines cost 13 % extra DRAM reads on 512^3 (ncu, profiles/r01):
  // every touched line is fetched whole, including the pad.
  const int64_t per32 = 32 / g->es;
  g->xoff = round_up(g->glo[2], per32);
  g->pitch = round_up(g->xoff + g->n[2] + g->ghi[2], per32);
  g->rows = g->ext[0] * g->ext[1];
  g->total = g->rows * g->pitch;
  cudaError_t e = cudaMalloc(&g->dev, (size_t)g->total * g->es);
  if (e != cudaSuccess) {
    delete g;
    return fail(ctx, HPCX_OOM, std::string("cudaMalloc(grid): ") + cudaGetErrorString(e));
  }
  e = cudaMemsetAsync(g->dev, 0, (size_t)g->total * g->es, ctx->compute);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->compute);
  if (e != cudaSuccess) {
    cudaFree(g->dev);
    delete g;
    return fail(ctx, HPCX_CUDA_ERROR, std::string("memset(grid): ") + cudaGetErrorString(e));
  }
  *out = g;
  return HPCX_OK;
}

hpcx_status hpcx_grid_destroy(hpcx_grid g) {
  if (!g) return HPCX_INVALID_ARG;
  cudaFree(g->dev);
  delete g;
  return HPCX_OK;
}

static hpcx_status grid_copy(hpcx_grid g, void *host, bool to_device, void *stream) {
  if (!g || !host) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = g->ctx;
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  char *dev = (char *)g->dev + (g->xoff - g->glo[2]) * g->es;
  const size_t width = (size_t)g->ext[2] * g->es;
  if (to_device)
    HPCX_CUDA(ctx, cudaMemcpy2DAsync(dev, (size_t)g->pitch * g->es, host, width, width,
                                     (size_t)g->rows, cudaMemcpyHostToDevice, s));
  else
    HPCX_CUDA(ctx, cudaMemcpy2DAsync(host, width, dev, (size_t)g->pitch * g->es, width,
                                     (size_t)g->rows, cudaMemcpyDeviceToHost, s));
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}

hpcx_status hpcx_grid_upload(hpcx_grid g, const void *host_flat, void *stream) {
  return grid_copy(g, const_cast<void *>(host_flat), true, stream);
}
hpcx_status hpcx_grid_download(hpcx_grid g, void *host_flat, void *stream) {
  return grid_copy(g, host_flat, false, stream);
}

hpcx_status hpcx_grid_device_layout(hpcx_grid g, int64_t *xoff, int64_t *pitch, int64_t *total,
                                    void **dev_ptr) {
  if (!g) return HPCX_INVALID_ARG;
  if (xoff) *xoff = g->xoff;
  if (pitch) *pitch = g->pitch;
  if (total) *total = g->total;
  if (dev_ptr) *dev_ptr = g->dev;
  return HPCX_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ pack / unpack
// Buffer shape = padded extents with the extent of `axis` replaced by `width`
// (slowest first).  One thread per buffer element, x fastest => coalesced both sides.
template <typename T, bool PACK>
__global__ void __launch_bounds__(256)
face_copy_kernel(T *__restrict__ grid, T *__restrict__ buf, int64_t b0, int64_t b1, int64_t b2,
                 int64_t s0, int64_t s1, int64_t s2,  // start (padded coordinates) per axis
                 int64_t ext1, int64_t pitch, int64_t xshift) {
  const int64_t total = b0 * b1 * b2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t x = i % b2;
    const int64_t r = i / b2;
    const int64_t y = r % b1;
    const int64_t z = r / b1;
    const int64_t gi = ((z + s0) * ext1 + (y + s1)) * pitch + xshift + x + s2;
    if (PACK)
      buf[i] = grid[gi];
    else
      grid[gi] = buf[i];
  }
}

static hpcx_status face_geometry(hpcx_grid g, int axis, int side, int width, bool ghost,
                                 int64_t b[3], int64_t s[3]) {
  hpcx_ctx ctx = g->ctx;
  HPCX_REQUIRE(ctx, axis >= 0 && axis < g->desc.ndim, "axis out of range");
  HPCX_REQUIRE(ctx, side == 0 || side == 1, "side must be 0 or 1");
  const int a = axis + (3 - g->desc.ndim);
  HPCX_REQUIRE(ctx, width >= 1 && width <= g->n[a], "face width out of range");
  if (ghost) HPCX_REQUIRE(ctx, width <= (side ? g->ghi[a] : g->glo[a]), "width exceeds ghost cells");
  for (int i = 0; i < 3; ++i) {
    b[i] = g->ext[i];
    s[i] = 0;
  }
  b[a] = width;
  if (ghost)
    s[a] = side ? g->glo[a] + g->n[a] : g->glo[a] - width;
  else
    s[a] = side ? g->glo[a] + g->n[a] - width : g->glo[a];
  return HPCX_OK;
}

template <bool PACK>
static hpcx_status face_copy(hpcx_grid g, int axis, int side, int width, void *buf, int buf_mem,
                             void *stream) {
  if (!g || !buf) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = g->ctx;
  int64_t b[3], s[3];
  hpcx_status st = face_geometry(g, axis, side, width, !PACK, b, s);
  if (st != HPCX_OK) return st;
  cudaStream_t cs = stream ? (cudaStream_t)stream : ctx->compute;
  const int64_t elems = b[0] * b[1] * b[2];
  void *dbuf = buf;
  if (buf_mem == HPCX_MEM_HOST) {
    HPCX_CUDA(ctx, cudaMallocAsync(&dbuf, (size_t)elems * g->es, cs));
    if (!PACK)
      HPCX_CUDA(ctx, cudaMemcpyAsync(dbuf, buf, (size_t)elems * g->es, cudaMemcpyHostToDevice, cs));
  }
  const int blocks = (int)std::min<int64_t>((elems + 255) / 256, (int64_t)ctx->sm_count * 16);
  const int64_t xshift = g->xoff - g->glo[2];
  if (g->es == 4)
    face_copy_kernel<float, PACK><<<blocks, 256, 0, cs>>>((float *)g->dev, (float *)dbuf, b[0],
                                                         b[1], b[2], s[0], s[1], s[2], g->ext[1],
                                                         g->pitch, xshift);
  else
    face_copy_kernel<double, PACK><<<blocks, 256, 0, cs>>>((double *)g->dev, (double *)dbuf, b[0],
                                                          b[1], b[2], s[0], s[1], s[2], g->ext[1],
                                                          g->pitch, xshift);
  st = after_launch(ctx, PACK ? "pack_face" : "unpack_face");
  if (st != HPCX_OK) return st;
  if (buf_mem == HPCX_MEM_HOST) {
    if (PACK)
      HPCX_CUDA(ctx, cudaMemcpyAsync(buf, dbuf, (size_t)elems * g->es, cudaMemcpyDeviceToHost, cs));
    HPCX_CUDA(ctx, cudaFreeAsync(dbuf, cs));
    HPCX_CUDA(ctx, cudaStreamSynchronize(cs));
  } else if (!stream) {
    HPCX_CUDA(ctx, cudaStreamSynchronize(cs));
  }
  return HPCX_OK;
}

extern "C" {

hpcx_status hpcx_grid_face_elems(hpcx_grid g, int32_t axis, int32_t width, int64_t *elems) {
  if (!g || !elems) return HPCX_INVALID_ARG;
  int64_t b[3], s[3];
  hpcx_status st = face_geometry(g, axis, 0, width, false, b, s);
  if (st != HPCX_OK) return st;
  *elems = b[0] * b[1] * b[2];
  return HPCX_OK;
}

hpcx_status hpcx_grid_pack_face(hpcx_grid g, int32_t axis, int32_t side, int32_t width, void *buf,
                                int32_t buf_mem, void *stream) {
  return face_copy<true>(g, axis, side, width, buf, buf_mem, stream);
}

hpcx_status hpcx_grid_unpack_face(hpcx_grid g, int32_t axis, int32_t side, int32_t width,
                                  const void *buf, int32_t buf_mem, void *stream) {
  return face_copy<false>(g, axis, side, width, const_cast<void *>(buf), buf_mem, stream);
}

}  // extern "C"
