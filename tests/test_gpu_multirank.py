"""Multi-GPU parity of the CUDA path: the fleet sharded by node over 2 GPUs (one process per GPU, NCCL), bodies
concatenated in node order on rank 0, compared byte for byte with the CPU oracle's body of the whole fleet
(tests/multirank_worker.py).  Needs two visible GPUs: skipped on a one-GPU box; run with `gpurun --gpus 2`."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_node_sharded_fleet_on_two_gpus_equals_the_oracle(kgo):
    from kgx_b200 import build
    build.build()
    n = min(torch.cuda.device_count(), 8)
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(HERE, "multirank_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    sys.stdout.write(res.stdout[-3000:])
    assert res.returncode == 0, res.stderr[-3000:]
    assert "MULTIRANK PARITY OK" in res.stdout
