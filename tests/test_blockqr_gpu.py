"""GPU parity tests of BlockQR (n > 64) through the C ABI against LAPACK and the C restatement
of the reference's block_qr_core / block_qr_reorthogonalization_core."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n", [(4096, 128), (5000, 200), (2048, 65), (1024, 512)])
@pytest.mark.parametrize("reorth", [False, True])
def test_blockqr_uniform(cuda, oracle, m, n, reorth):
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=n)
    q, r, a_in = run_qr(a, "fp32_notc", reorth)
    assert np.abs(np.tril(r, -1)).max() == 0.0
    orth, res = oracle.orthogonality(q), oracle.residual(q, r, a_in)
    st, qo, ro = oracle.ref_blockqr(a, reorth)
    assert st == 0
    # no worse than 2x the reference algorithm on the same input (small floor for fp32 noise)
    assert orth <= max(2 * oracle.orthogonality(qo), 2e-6), (orth, oracle.orthogonality(qo))
    assert res <= max(2 * oracle.residual(qo, ro, a), 2e-6), (res, oracle.residual(qo, ro, a))
    _, rl = oracle.lapack_qr(a_in, np.float64)
    assert oracle.r_rel_diff(r, rl) < 2e-5


def test_blockqr_conditioned_reorth(cuda, oracle):
    from helpers import run_qr
    a = oracle.latms_matrix(8192, 160, 1e4, seed=5)
    q0, r0, a_in = run_qr(a, "fp32_notc", False)
    q1, r1, _ = run_qr(a, "fp32_notc", True)
    _, qo0, ro0 = oracle.ref_blockqr(a, False)
    _, qo1, ro1 = oracle.ref_blockqr(a, True)
    assert oracle.orthogonality(q0) <= max(2 * oracle.orthogonality(qo0), 2e-6)
    assert oracle.orthogonality(q1) <= max(2 * oracle.orthogonality(qo1), 2e-6)
    assert oracle.orthogonality(q1) < 5e-6  # re-orthogonalisation restores O(eps)
    assert oracle.residual(q1, r1, a_in) < 5e-6 and oracle.residual(q0, r0, a_in) < 5e-6


def test_blockqr_half_io(cuda, oracle):
    from helpers import run_qr
    a = oracle.uniform_matrix(4096, 128, seed=8)
    q, r, a_in = run_qr(a, "fp16_tc_nocor", True)
    assert oracle.orthogonality(q) < 5e-3 and oracle.residual(q, r, a_in) < 5e-3


def test_blockqr_overwrites_a_only_when_wide(cuda, oracle):
    import tsqr_gpu_b200 as tq
    a = tq.to_colmajor(torch.from_numpy(oracle.uniform_matrix(2000, 64, seed=1)).cuda())
    a0 = a.clone()
    tq.qr(a, "fp32_notc")
    torch.cuda.synchronize()
    assert torch.equal(a, a0)


@pytest.mark.parametrize("mode", ["fp32_tc_cor", "tf32_tc_cor"])
@pytest.mark.parametrize("m,n", [(4096, 128), (5000, 200), (2048, 65), (6000, 320), (5001, 130)])
@pytest.mark.parametrize("reorth", [False, True])
def test_blockqr_tensor_core_modes(cuda, oracle, mode, m, n, reorth):
    """n > 64 in a *_tc_cor mode: TSQR panels and both projection GEMMs (tcgen05, hi/lo corrected)."""
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=n + 1)
    q, r, a_in = run_qr(a, mode, reorth)
    assert np.isfinite(q).all() and np.isfinite(r).all()
    orth, res = oracle.orthogonality(q), oracle.residual(q, r, a_in)
    st, qo, ro = oracle.ref_blockqr(a, reorth)
    print(f"{mode} {m}x{n} reorth={reorth}: orth {orth:.2e} (ref alg {oracle.orthogonality(qo):.2e}) res {res:.2e}")
    assert orth <= max(2 * oracle.orthogonality(qo), 3e-6), (orth, oracle.orthogonality(qo))
    assert res <= max(2 * oracle.residual(qo, ro, a), 3e-6), (res, oracle.residual(qo, ro, a))
    _, rl = oracle.lapack_qr(a_in, np.float64)
    assert oracle.r_rel_diff(r, rl) < 3e-5


def test_blockqr_tensor_core_nocor(cuda, oracle):
    from helpers import run_qr
    a = oracle.uniform_matrix(8192, 192, seed=4)
    for mode in ("fp32_tc_nocor", "tf32_tc_nocor", "fp16_tc_nocor"):
        q, r, a_in = run_qr(a, mode, True)
        assert oracle.orthogonality(q) < 5e-3 and oracle.residual(q, r, a_in) < 5e-3, mode


@pytest.mark.parametrize("reorth", [False, True])
def test_blockqr_tensor_core_wider_than_the_tmem_accumulators(cuda, oracle, reorth):
    """n = 704: the projection of the last panels runs over 576 and 640 columns of Q1, more than the 8 accumulator
    blocks (512 TMEM columns) one tc_tn_kernel launch holds; the driver must split Q1 into groups."""
    from helpers import run_qr
    a = oracle.uniform_matrix(4096, 704, seed=70)
    q, r, a_in = run_qr(a, "fp32_tc_cor", reorth)
    assert np.isfinite(q).all() and np.isfinite(r).all()
    orth, res = oracle.orthogonality(q), oracle.residual(q, r, a_in)
    print(f"4096x704 reorth={reorth}: orth {orth:.2e} res {res:.2e}")
    assert res < 5e-6
    assert orth < (5e-6 if reorth else 2e-4)
    _, rl = oracle.lapack_qr(a_in, np.float64)
    assert oracle.r_rel_diff(r, rl) < (3e-5 if reorth else 1e-3)


@pytest.mark.parametrize("mode,reorth", [("fp32_tc_cor", False), ("tf32_tc_cor", True), ("fp32_tc_nocor", True), ("fp16_tc_nocor", True)])
def test_blockqr_many_tiles_per_cta(cuda, mode, reorth):
    """2^17 x 256: 2048 row tiles of 64 over 148 CTAs, i.e. 14 tiles per CTA in the projection kernels (the shapes above give
    every CTA one tile). The first tensor-memory form of the coefficient product passed those and hung here: its two
    consumer groups shared one landing ring and a parity wait cannot tell phase k from phase k - 2. Metrics on the device."""
    import torch
    import tsqr_gpu_b200 as tq
    m, n = 1 << 17, 256
    g = torch.Generator(device="cuda").manual_seed(17)
    a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t().to(tq.io_dtype(mode)))
    a0 = a.clone()
    q, r = tq.qr(a, mode, reorth)
    torch.cuda.synchronize()
    orth, res = float(tq.orthogonality(q, mode)), float(tq.residual(a0, q, r, mode))
    print(f"{mode} {m}x{n} reorth={reorth}: orthogonality {orth:.2e} residual {res:.2e}")
    tol = 5e-6 if mode.endswith("_cor") else 5e-3
    assert orth < tol and res < tol
