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
    log_dir = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(log_dir):
        with open(os.path.join(log_dir, "mp_worker_w%d.log" % world), "w") as f:
            f.write(out.stdout + "\n==== stderr ====\n" + out.stderr)
    fails = [ln for ln in out.stdout.splitlines() if "FAIL" in ln]
    tail = "\n".join(fails[:40]) + "\n" + (out.stdout + out.stderr)[-3000:]
    assert out.returncode == 0 and "MULTIGPU OK" in out.stdout, tail
