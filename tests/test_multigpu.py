"""Real multi-GPU run of the peer-memory path (CUDA IPC arenas, NVLink stores, flag protocol)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_native_path_across_gpus(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "mp_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    tail = (out.stdout + out.stderr)[-4000:]
    assert out.returncode == 0 and "MULTIGPU OK" in out.stdout, tail
