// dist.cuh — internals shared by the multi-GPU transports: dist.cu (one process per GPU: NCCL
// and CUDA IPC windows) and multi.cu (one process, several GPUs: plain peer pointers).
#pragma once
#include "grid.cuh"
#include "spmv.cuh"

struct hpcx_halo_plan_s {
  hpcx_ctx ctx = nullptr;
  hpcx_grid in = nullptr, out = nullptr;
  hpcx_stencil_desc st{};
  int transport = 0;
  // 2-D block split of the box sweeps (hpcx_stencil_dist_plan_blocks): the ranks form a py x px
  // grid, rank = ry * px + rx; px == 1 is the y-slab split.  face_send / face_recv: packed x faces
  // (side 0 = towards rx - 1), ext_y x radius elements each.
  int py = 0, px = 1;
  void *face_send[2] = {nullptr, nullptr}, *face_recv[2] = {nullptr, nullptr};
  // ev_bnd[pass & 1]: boundary strips of a pass done; ev_int: interior of a pass done
  cudaEvent_t ev_bnd[2] = {nullptr, nullptr}, ev_recv = nullptr, ev_start = nullptr, ev_int = nullptr;
  cudaError_t create_events() {
    cudaError_t e = cudaSuccess;
    for (cudaEvent_t *ev : {&ev_bnd[0], &ev_bnd[1], &ev_recv, &ev_start, &ev_int})
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
    return e;
  }
  void destroy_events() {
    for (cudaEvent_t ev : {ev_bnd[0], ev_bnd[1], ev_recv, ev_start, ev_int})
      if (ev) cudaEventDestroy(ev);
  }
  // transport 1 (peer stores): neighbour side 0 = rank-1, 1 = rank+1
  uint32_t *flags = nullptr;          // local block: [0],[1] flags, [4],[5] counters, [8] push counter
  void *buf[2] = {nullptr, nullptr};  // this rank's two grid buffers at plan time (A = in, B = out)
  void *peer_open[2][3] = {};         // bases returned by cudaIpcOpenMemHandle (A, B, flags)
  float *peer_buf[2][2] = {};         // the neighbour's A and B
  uint32_t *peer_flags[2] = {nullptr, nullptr};
  int64_t peer_n0[2] = {0, 0};
  uint32_t epoch = 0;                 // sweeps completed through this plan
  // NCCL transport: one captured CUDA graph per (source buffer, destination buffer, steps)
  struct GraphEntry {
    void *in, *out;
    int steps;
    cudaGraphExec_t exec;
    int64_t launches;  // kernels of this library one replay launches (for hpcx_launch_count)
  };
  std::vector<GraphEntry> graphs;
  bool graphs_ok = true;  // cleared when a capture fails: the call then runs eagerly
};


namespace hpcx {
// kernels of the peer transports loaded ahead of their first launch (see stencil.cu)
hpcx_status preload_stencil_kernels(hpcx_ctx ctx);
hpcx_status preload_dist_kernels(hpcx_ctx ctx);
hpcx_status preload_spmv_kernels(hpcx_ctx ctx);
hpcx_status plan_prepare(hpcx_halo_plan p);  // local flag block + the plan's two buffers
// `steps` sweeps through the fused peer-store transport on stream cs (the neighbours' windows in
// p->peer_* must be wired); sync: wait for the stream and report a flag timeout.
hpcx_status stencil_peer_sweep(hpcx_halo_plan p, int32_t steps, cudaStream_t cs, bool sync);
}  // namespace hpcx
