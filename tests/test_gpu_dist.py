"""Distributed paths on the GPU box: world size 1 in-process, and world size 2..N under
torchrun when the box has more than one GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
SCRIPT = os.path.join(ROOT, "tests", "dist_gpu_check.py")


def test_dist_paths_world_1():
    env = dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    p = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, env=env, timeout=900)
    assert p.returncode == 0 and "DIST OK world=1" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dist_paths_multi_gpu(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port",
                        str(29500 + world), SCRIPT], capture_output=True, text=True, timeout=1200)
    assert p.returncode == 0 and f"DIST OK world={world}" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]


def test_peer_flag_timeout_traps_and_is_reported():
    """hpcx_comm_set_timeout: a wait that expires traps (no stale halo is read) and the caller sees
    HPCX_NCCL_ERROR; without a bound the wait lasts (dist_gpu_check.py delays a rank by 5 s)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = os.path.join(ROOT, "tests", "dist_timeout_check.py")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29511", script],
                       capture_output=True, text=True, timeout=600)
    assert "TIMEOUT OK" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]
