"""NCCL all-gather of the generated transitions on >= 2 GPUs: the gathered records must equal the
single-GPU rollout of all rows bit for bit (tools/nccl_gather_check.py, launched with torchrun).  Skipped on a
one-GPU box; the CPU twin of the host logic is tests/test_sharding_gloo.py."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_gpu_gather_equals_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "nccl_gather_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("equal_to_single_gpu=True") == 3, r.stdout
    assert r.stdout.count("dense_equal=True") == 3, r.stdout          # DensePoolGatherer, all three modes
