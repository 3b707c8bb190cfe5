"""csm_stack2d_broadcast on two GPUs (tools/broadcast_check.py under torchrun); skipped on a
single-GPU box.  profiles/r02_broadcast_2gpu.log holds the run on a 2-GPU box."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_stack_broadcast_over_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29517",
                          os.path.join(ROOT, "tools", "broadcast_check.py")],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    doc = json.loads(line)
    assert doc["csm_stack2d_broadcast"] == "ok" and doc["ranks"] == 2
