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


def test_library_sharded_driver_single_rank(cuda, oracle):
    """tsqr_b200_tsqr_sharded with a one-rank communicator (NCCL inside the library, no exchange partner): both forms
    must reproduce the single-GPU entry point -- bit for bit in the preconditioned-Cholesky form (same stages)."""
    import ctypes
    import numpy as np
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    buf = ctypes.create_string_buffer(128)
    assert lib.tsqr_b200_nccl_unique_id(buf) == 0, lib.tsqr_b200_last_error()
    h = lib.tsqr_b200_sharded_init(buf.raw, 1, 0)
    assert h, lib.tsqr_b200_last_error()
    try:
        for mode, m, n, exact in [("fp32_tc_cor", 1 << 17, 64, True), ("fp32_notc", 30000, 16, False), ("fp32_tc_cor", 5000, 48, False)]:
            a = tq.to_colmajor(torch.from_numpy(oracle.uniform_matrix(m, n, seed=n)).cuda())
            q = tq.colmajor_empty(m, n, torch.float32)
            r = torch.zeros((n, n), device="cuda").t()
            ws = torch.empty(lib.tsqr_b200_sharded_workspace_bytes(m, n, 1), dtype=torch.uint8, device="cuda")
            code = lib.tsqr_b200_tsqr_sharded(h, tq.MODES[mode], q.data_ptr(), m, r.data_ptr(), n, a.data_ptr(), m, m, n,
                                              ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream)
            assert code == 0, lib.tsqr_b200_last_error()
            q1, r1 = tq.tsqr(a, mode)
            torch.cuda.synchronize()
            if exact:
                assert torch.equal(q, q1) and torch.equal(r, r1)
            else:
                q64, r64 = q.cpu().numpy().astype(np.float64), r.cpu().numpy().astype(np.float64)
                assert oracle.orthogonality(q64) < 5e-6 and oracle.residual(q64, r64, a.cpu().numpy().astype(np.float64)) < 5e-6
    finally:
        lib.tsqr_b200_sharded_destroy(h)


def test_two_devices_in_one_process(cuda):
    """One process driving two GPUs one after the other: everything the library keeps per device (kernel attributes for
    dynamic shared memory, the constant-memory image slots and their events, the SM count, the static scratch address)
    must follow the current device. Same input, same bits on both devices."""
    import numpy as np
    import torch
    import tsqr_gpu_b200 as tq
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    outs = []
    try:
        for dev in (0, 1):
            torch.cuda.set_device(dev)
            for (m, n, mode, reorth) in [(1 << 16, 64, "fp32_tc_cor", False), (20000, 32, "fp32_notc", False),
                                         (1 << 15, 128, "tf32_tc_cor", True)]:
                g = torch.Generator(device=f"cuda:{dev}").manual_seed(5)
                a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device=f"cuda:{dev}").uniform_(-1, 1, generator=g).t())
                q, r = tq.qr(a, mode, reorth)
                torch.cuda.synchronize(dev)
                assert bool(torch.isfinite(q).all()) and float(tq.orthogonality(q, mode)) < 5e-6
                outs.append((q.cpu(), r.cpu()))
    finally:
        torch.cuda.set_device(0)
    k = len(outs) // 2
    for (q0, r0), (q1, r1) in zip(outs[:k], outs[k:]):
        assert torch.equal(q0, q1) and torch.equal(r0, r1)
