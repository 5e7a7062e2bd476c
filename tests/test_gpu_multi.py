"""Camera-batch DP on real GPUs (needs >= 2 visible devices; skipped on a 1-GPU box): tools/dp_check.py under
torchrun, once per exchange transport.  The 1-GPU emulation of the same protocol is tests/test_gpu_exchange.py, the
host logic over gloo is tests/test_dp_gloo.py."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _port():
  with socket.socket() as s:
    s.bind(("127.0.0.1", 0))
    return s.getsockname()[1]


@pytest.mark.parametrize("transport", ["auto", "peer-replicated", "nccl", "dense"])
def test_two_gpu_dp_equals_sequential(transport):
  if torch.cuda.device_count() < 2:
    pytest.skip("needs 2 GPUs")
  cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
         "--master-port", str(_port()), os.path.join(ROOT, "tools", "dp_check.py"), transport]
  r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
  assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
