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
