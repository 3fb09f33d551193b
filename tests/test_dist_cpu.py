"""N>1 host logic on CPU: world_size-2 (and 3) gloo runs of the exchange schedule that
hpcx_spmv_dist / hpcx_stencil_dist_sweep issue with NCCL — same partition, same halo tables,
same message order — with the oracle as the local kernel.  Checks that the partitioned
result equals the 1-way result bit-for-bit (SURVEY.md §8c vi)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _exchange(xfull, L, P, rank):
    """Grouped send/recv in the order of hpcx_spmv_dist: for q ascending, send segment q of the
    gathered boundary x, receive halo segment q into [nloc + roff ...)."""
    sendbuf = xfull[:L["nloc"]][L["send_idx"]]
    reqs, soff, roff = [], 0, 0
    recv = []
    for q in range(P):
        ns, nr = int(L["send_counts"][q]), int(L["recv_counts"][q])
        if ns:
            reqs.append(dist.isend(torch.from_numpy(sendbuf[soff:soff + ns].copy()), q))
        if nr:
            t = torch.empty(nr, dtype=torch.from_numpy(xfull).dtype)
            recv.append((roff, t))
            reqs.append(dist.irecv(t, q))
        soff += ns
        roff += nr
    for r in reqs:
        r.wait()
    for roff, t in recv:
        xfull[L["nloc"] + roff:L["nloc"] + roff + len(t)] = t.numpy()


def _worker(rank, world, port, case):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import libhpc_b200 as hp
    import oracle
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        if case == "spmv_op27":
            rp, col, val = oracle.op27(8)
            mode = 0
        else:
            rp, col, val = oracle.rmat(10, 8, 31)
            val = val.astype(np.float64)
            mode = 1
        n = len(rp) - 1
        if case != "jacobi":
            x = oracle.fill(n, 42, np.float64)
            y_global = oracle.spmv(rp, col, val, x)
            A = hp.CsrMatrix(rp, col, val)
            bounds = hp.partition_rows(n, rp, world, mode)
            L = hp.csr_localize(A, bounds, rank)
            xfull = np.zeros(L["nloc"] + L["nhalo"])
            xfull[:L["nloc"]] = x[bounds[rank]:bounds[rank + 1]]
            _exchange(xfull, L, world, rank)
            assert np.array_equal(xfull[L["nloc"]:], x[L["halo_gcol"]])
            y = oracle.spmv(L["local_row_ptr"], L["local_col"], np.ascontiguousarray(L["values"]), xfull)
            assert np.array_equal(y, y_global[bounds[rank]:bounds[rank + 1]])
            # interior rows never touch the halo: they can run before the exchange completes
            xonly = xfull.copy()
            xonly[L["nloc"]:] = np.nan
            y2 = oracle.spmv(L["local_row_ptr"], L["local_col"], np.ascontiguousarray(L["values"]), xonly)
            interior = L["row_kind"] == 0
            assert np.array_equal(y2[interior], y[interior]) and np.isnan(y2[~interior]).all()
            # the same tables from this rank's OWN rows only (hpcx_csr_localize_rows + one exchange of
            # column lists + hpcx_halo_send_lists): bit-identical to the global-matrix localiser and
            # to the oracle's
            from libhpc_b200 import dist as hd
            r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
            p0, p1 = int(rp[r0]), int(rp[r1])
            D = hd.localize_rows_distributed(rp[r0:r1 + 1] - p0, col[p0:p1], val[p0:p1], bounds, rank, world, n)
            O = oracle.localize(rp, col, bounds, rank)
            for k in ("local_row_ptr", "local_col", "halo_gcol", "recv_counts", "send_counts", "send_idx", "row_kind"):
                assert np.array_equal(np.asarray(D[k]), np.asarray(L[k])), (rank, k)
                assert np.array_equal(np.asarray(D[k]), np.asarray(O[k])), (rank, k, "oracle")
            assert (D["nloc"], D["nnz_loc"], D["nhalo"], D["nsend"]) == (L["nloc"], L["nnz_loc"], L["nhalo"], L["nsend"])
        else:
            nz, ny, nx, steps = 9, 6, 10, 3
            full = oracle.fill((nz + 2) * (ny + 2) * (nx + 2), 7, np.float32)
            want = oracle.jacobi7(full, nz, ny, nx, 1, 1, steps).reshape(nz + 2, ny + 2, nx + 2)
            b = hp.partition_slabs(nz, world)
            z0, z1 = int(b[rank]), int(b[rank + 1])
            cur = full.reshape(nz + 2, ny + 2, nx + 2)[z0:z1 + 2].copy()  # own planes + 1 ghost each side
            nxt = cur.copy()
            for _ in range(steps):
                nxt = oracle.jacobi7(cur.ravel(), z1 - z0, ny, nx, 1, 1).reshape(cur.shape)
                # exchange the new boundary planes (full padded planes, as the device path does)
                reqs = []
                lo, hi = rank - 1, rank + 1
                if lo >= 0:
                    reqs.append(dist.isend(torch.from_numpy(nxt[1].copy()), lo))
                    tl = torch.empty(ny + 2, nx + 2)
                    reqs.append(dist.irecv(tl, lo))
                if hi < world:
                    reqs.append(dist.isend(torch.from_numpy(nxt[-2].copy()), hi))
                    th = torch.empty(ny + 2, nx + 2)
                    reqs.append(dist.irecv(th, hi))
                for r in reqs:
                    r.wait()
                if lo >= 0:
                    nxt[0] = tl.numpy()
                if hi < world:
                    nxt[-1] = th.numpy()
                cur = nxt
            assert np.array_equal(cur[1:-1, 1:-1, 1:-1], want[z0 + 1:z1 + 1, 1:-1, 1:-1])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,world", [("spmv_op27", 2), ("spmv_rmat", 2), ("spmv_rmat", 3),
                                        ("jacobi", 2), ("jacobi", 3), ("spmv_op27", 4), ("spmv_rmat", 4)])
def test_partitioned_equals_global(case, world):
    mp.spawn(_worker, args=(world, _free_port(), case), nprocs=world, join=True)
