"""GPU parity tests of the tensor-core (tcgen05/TMEM) TSQR passes through the C ABI, against host
LAPACK (R after sign normalisation, orthogonality, residual; definitions of reference
src/test.cu:146-165, src/validation.cu:43-80). Tolerances: fp32-class for the hi/lo corrected
modes, tf32/half operand rounding (2^-11) for the *_nocor modes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

COR = ["fp32_tc_cor", "tf32_tc_cor"]
NOCOR = ["fp32_tc_nocor", "tf32_tc_nocor", "fp16_tc_nocor"]


def _metrics(oracle, q, r, a_in):
    assert np.isfinite(q).all() and np.isfinite(r).all()
    assert np.abs(np.tril(r, -1)).max() == 0.0
    _, rl = oracle.lapack_qr(a_in, np.float64)
    return oracle.orthogonality(q), oracle.residual(q, r, a_in), oracle.r_rel_diff(r, rl)


@pytest.mark.parametrize("mode", COR)
@pytest.mark.parametrize("m,n", [(1024, 64), (4096, 64), (5000, 64), (16384, 16), (20000, 32), (3000, 48), (2500, 7),
                                 (70001, 64), (1 << 18, 32)])
def test_corrected_modes_match_lapack(cuda, oracle, mode, m, n):
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=200 + n)
    q, r, a_in = run_qr(a, mode, entry="tsqr")
    orth, res, rd = _metrics(oracle, q, r, a_in)
    print(f"{mode} {m}x{n}: orth {orth:.2e} res {res:.2e} R {rd:.2e}")
    assert orth < 5e-6 and res < 5e-6 and rd < 2e-5


@pytest.mark.parametrize("mode", NOCOR)
@pytest.mark.parametrize("m,n", [(4096, 64), (20000, 32), (16384, 16)])
def test_uncorrected_modes(cuda, oracle, mode, m, n):
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=300 + n)
    q, r, a_in = run_qr(a, mode, entry="tsqr")
    orth, res, rd = _metrics(oracle, q, r, a_in)
    print(f"{mode} {m}x{n}: orth {orth:.2e} res {res:.2e} R {rd:.2e}")
    assert orth < 5e-3 and res < 5e-3 and rd < 5e-3


@pytest.mark.parametrize("cond", [1e3, 1e6])
def test_conditioned_inputs_tc(cuda, oracle, cond):
    from helpers import run_qr
    a = oracle.latms_matrix(20000, 64, cond, seed=12)
    q, r, a_in = run_qr(a, "fp32_tc_cor", entry="tsqr")
    orth, res, _ = _metrics(oracle, q, r, a_in)
    print(f"cond {cond:g}: orth {orth:.2e} res {res:.2e}")
    assert orth < 5e-6 and res < 5e-6


def test_leading_dimensions_in_place_and_zero_column(cuda, oracle):
    import tsqr_gpu_b200 as tq
    m, n = 9211, 40
    a = oracle.uniform_matrix(m, n, seed=3)
    a[:, 5] = 0.0
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda(), ld=m + 37)
    a_copy = a_dev.clone()
    q = tq.colmajor_empty(m, n, torch.float32, ld=m + 5)
    qq, rr = tq.tsqr(a_dev, "fp32_tc_cor", q=q)
    torch.cuda.synchronize()
    assert torch.equal(a_dev, a_copy), "tsqr must not modify A"
    q64, r64 = qq.cpu().numpy().astype(np.float64), rr.cpu().numpy().astype(np.float64)
    assert np.isfinite(q64).all() and np.isfinite(r64).all()
    assert oracle.residual(q64, r64, a.astype(np.float64)) < 5e-6
    assert r64[5, 5] == 0.0
    # in place: Q over A
    q2, r2 = tq.tsqr(a_dev, "fp32_tc_cor", q=a_dev)
    torch.cuda.synchronize()
    assert torch.equal(q2[:, :n].cpu(), qq[:, :n].cpu()) and torch.equal(r2.cpu(), rr.cpu())
