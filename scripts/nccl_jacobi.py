"""NCCL-transport Jacobi over z-slabs (512^3 per GPU): CUDA-graph replay vs eager launches, NCCL CTA cap.
Run under torchrun; env HPCX_NCCL_GRAPHS / HPCX_NCCL_MAX_CTAS select the variant."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp  # noqa: E402
from libhpc_b200 import dist as hd  # noqa: E402
from libhpc_b200 import synth  # noqa: E402

rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = hp.Context(local)
hd.init_comm(ctx, rank, world)
n, T = 512, 50
stream = torch.cuda.Stream()
for variant, gw, name in ((hp.VARIANT_TMA, 1, "one sweep"), (hp.VARIANT_TEMPORAL, 2, "two sweeps")):
    f = synth.field(7, (n + 2 * gw) ** 3, np.float32, first=rank)
    ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, gw), hp.DeviceGrid(ctx, (n, n, n), np.float32, gw)
    ga.upload(f)
    gb.upload(f)
    for transport in (0, 1):
        plan = hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, variant, transport=transport)
        with torch.cuda.stream(stream):
            plan.sweep(T, stream)
            plan.sweep(T, stream)
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(3):
                plan.sweep(T, stream)
            e1.record(stream)
            torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            us = float(t) * 1e3 / (3 * T)
            print(f"graphs={os.environ.get('HPCX_NCCL_GRAPHS', '1')} max_ctas={os.environ.get('HPCX_NCCL_MAX_CTAS', '4')} "
                  f"{name:10s} transport {transport}: {us:7.1f} us/sweep  {n ** 3 * world / us / 1e3:7.1f} Gcell/s", flush=True)
        plan.destroy()
    ga.destroy()
    gb.destroy()
ctx.close()
dist.destroy_process_group()
