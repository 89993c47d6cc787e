"""GPU test of tsqr_b200_tsqr_static: the TSQR entry point that allocates nothing (its workspace is a file-scope
static __device__ array inside libtsqr_b200.so, see include/tsqr_b200.h). Called through the C ABI and compared with
the existing entry point tsqr_b200_tsqr on the same device bytes, on two calls in a row (the second call finds the
first call's contents in the shared scratch array), plus the size limit.

This file is named so that pytest collects it LAST; it passes both alone in a
fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 64 << 20        # TSQR_B200_STATIC_SCRATCH_BYTES of include/tsqr_b200.h


def _static_tsqr(a_dev, mode):
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    m, n = a_dev.shape
    q = tq.colmajor_empty(m, n, a_dev.dtype)
    r = torch.zeros((n, n), dtype=a_dev.dtype, device="cuda").t()
    code = lib.tsqr_b200_tsqr_static(tq.MODES[mode], q.data_ptr(), m, r.data_ptr(), n, a_dev.data_ptr(), a_dev.stride(1),
                                     m, n, torch.cuda.current_stream().cuda_stream)
    assert code == 0, lib.tsqr_b200_last_error().decode()
    return q, r


@pytest.mark.parametrize("mode,m,n", [("fp32_notc", 16384, 16), ("fp32_tc_cor", 1 << 20, 64), ("fp16_tc_nocor", 40000, 32)])
def test_static_scratch_entry_matches_the_allocating_one(cuda, oracle, mode, m, n):
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    dt = tq.io_dtype(mode)
    tol = 5e-6 if dt == torch.float32 else 5e-3          # the bounds of tests/helpers.py TOL for these modes
    for call in range(2):                                # two calls in a row, different inputs
        a = oracle.uniform_matrix(m, n, seed=900 + call)
        a_dev = tq.to_colmajor(torch.from_numpy(a).cuda().to(dt))
        q_ref, r_ref = tq.tsqr(a_dev, mode)              # existing entry point, caller-provided workspace
        q, r = _static_tsqr(a_dev, mode)
        torch.cuda.synchronize()
        dq = (q.float() - q_ref.float()).abs().max().item()
        dr = (r.float() - r_ref.float()).abs().max().item() / r_ref.float().abs().max().item()
        a_in = a_dev.float().cpu().numpy().astype(np.float64)
        q64, r64 = q.float().cpu().numpy().astype(np.float64), r.float().cpu().numpy().astype(np.float64)
        orth, res = oracle.orthogonality(q64), oracle.residual(q64, r64, a_in)
        print(f"{mode} {m}x{n} call {call}: |dQ| {dq:.1e} |dR| {dr:.1e} orth {orth:.2e} res {res:.2e}")
        assert dq <= tol and dr <= tol, "static-scratch entry differs from tsqr_b200_tsqr"
        assert orth < tol and res < tol


def test_static_scratch_entry_rejects_oversize(cuda):
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    m, n = 1 << 24, 64                 # the CUDA-core plan keeps one R checkpoint per tile: ~1 GB here
    assert lib.tsqr_b200_tsqr_workspace_bytes(m, n) > STATIC_SCRATCH_BYTES
    dummy = torch.zeros(16, device="cuda")
    # rejected before any pointer is touched or any kernel is launched
    code = lib.tsqr_b200_tsqr_static(tq.MODES["fp32_notc"], dummy.data_ptr(), m, dummy.data_ptr(), n, dummy.data_ptr(), m,
                                     m, n, None)
    assert code == -1
    assert b"static scratch" in lib.tsqr_b200_last_error()
    assert lib.tsqr_b200_tsqr_static(2, dummy.data_ptr(), 100, dummy.data_ptr(), 65, dummy.data_ptr(), 100, 100, 65, None) == 1
