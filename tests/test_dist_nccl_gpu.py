"""Row-sharded TSQR over NCCL on real GPUs (SURVEY.md section 8e / BASELINE config C4): launches
tests/dist_nccl_check.py with one rank per GPU. Skipped on boxes with a single GPU (the exchange
logic itself is covered on CPU by tests/test_dist_gloo.py, the local steps by
test_tsqr_gpu.py::test_row_sharded_virtual_ranks)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_sharded_nccl(cuda):
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs")
    P = 2 if ngpu < 4 else (4 if ngpu < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={P}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "dist_nccl_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert "DIST_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
