"""N > 1 on real GPUs: torchrun (one rank per GPU, NCCL) over tests/multi_gpu_worker.py.

Skipped on a single-GPU box (run it with `gpurun --gpus 2 -- python -m pytest
tests/test_gpu_multi_rank.py -m gpu`).  The CPU/gloo twin of the host logic is
tests/test_sharding.py."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2])
def test_torchrun_nccl_sharding_and_dcp(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    tail = (res.stdout + res.stderr)[-4000:]
    assert res.returncode == 0, tail
    assert "MULTI_GPU_OK" in res.stdout, tail
