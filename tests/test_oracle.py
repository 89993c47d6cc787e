"""CPU tests of the oracles themselves (no GPU): host LAPACK through the reference's metric
definitions, the C restatement of the reference algorithm (oracle/ref_tsqr.c) and the numpy
restatement of the new chunk algorithm (oracle/ts_flat.py). The reference holds no golden
vectors, so the oracles are pinned against each other and against the
defining properties of a QR factorisation."""
import numpy as np
import pytest


def test_lapack_oracle_is_a_qr(oracle):
    a = oracle.uniform_matrix(4096, 16, seed=1)
    q, r = oracle.lapack_qr(a)
    assert oracle.orthogonality(q) < 1e-6 and oracle.residual(q, r, a) < 1e-6
    q64, r64 = oracle.lapack_qr(a, np.float64)
    assert oracle.orthogonality(q64) < 1e-14
    assert oracle.r_rel_diff(r, r64) < 1e-5
    qn, rn = np.linalg.qr(a.astype(np.float64))
    assert oracle.r_rel_diff(r64, rn) < 1e-12


def test_metric_definitions(oracle):
    q = np.eye(8, 4)
    assert oracle.orthogonality(q) == 0.0
    q[0, 0] = 1.0 + 1e-3
    # ||Q^T Q - I||_F / sqrt(n), reference src/validation.cu:43-80
    assert abs(oracle.orthogonality(q) - (2e-3 + 1e-6) / 2.0) < 1e-9
    rr = np.triu(np.arange(16, dtype=np.float64).reshape(4, 4) + 1.0)
    assert oracle.residual(np.eye(8, 4), rr, np.vstack([rr, np.zeros((4, 4))])) == 0.0
    qs, rs = oracle.sign_normalize(-np.eye(4), -np.triu(np.ones((4, 4))))
    assert (np.diag(rs) > 0).all() and np.allclose(qs, np.eye(4))


@pytest.mark.parametrize("m,n", [(16384, 16), (1000, 16), (33, 7), (32, 16)])
def test_reference_restatement_matches_lapack(oracle, m, n):
    """oracle/ref_tsqr.c (reference algorithm, fp32_notc) against LAPACK: BASELINE config 1 shape"""
    a = oracle.uniform_matrix(m, n, seed=0x5EED0001)
    q, r = oracle.ref_tsqr16(a)
    _, rl = oracle.lapack_qr(a, np.float64)
    assert oracle.orthogonality(q) < 5e-6 and oracle.residual(q, r, a) < 5e-6
    assert oracle.r_rel_diff(r, rl) < 1e-5


def test_reference_blockqr_restatement(oracle):
    a = oracle.uniform_matrix(2048, 48, seed=3)
    for reorth in (False, True):
        st, q, r = oracle.ref_blockqr(a, reorth)
        assert st == 0
        assert oracle.orthogonality(q) < 1e-5 and oracle.residual(q, r, a) < 1e-5


@pytest.mark.parametrize("m,n,chunks", [(700, 16, 1), (1500, 64, 1), (3000, 32, 4), (1024, 7, 3), (40, 16, 1)])
def test_chunk_algorithm_restatement(oracle, m, n, chunks):
    """oracle/ts_flat.py (head QR + flat tree of [R; tile] steps, T^-1 = striu(V^T V) + diag) is a QR
    and agrees with LAPACK up to row signs -- in fp64 to 1e-12, i.e. the formulas are exact."""
    from oracle import ts_flat
    a = oracle.uniform_matrix(m, n, seed=5).astype(np.float64)
    q, r = ts_flat.tsqr(a, chunks)
    _, rl = oracle.lapack_qr(a, np.float64)
    assert oracle.orthogonality(q) < 1e-13 and oracle.residual(q, r, a) < 1e-13
    assert oracle.r_rel_diff(r, rl) < 1e-12
    assert np.abs(np.tril(r, -1)).max() == 0.0


def test_chunk_algorithm_conditioned_and_degenerate(oracle):
    from oracle import ts_flat
    a = oracle.latms_matrix(2000, 32, 1e6, seed=2).astype(np.float64)
    q, r = ts_flat.tsqr(a, 2)
    assert oracle.orthogonality(q) < 1e-13 and oracle.residual(q, r, a) < 1e-13
    a = oracle.uniform_matrix(600, 16, seed=9).astype(np.float64)
    a[:, 4] = 0.0  # zero column: v = 0, tau = 2 in both passes
    q, r = ts_flat.tsqr(a, 1)
    assert np.isfinite(q).all() and oracle.residual(q, r, a) < 1e-13 and r[4, 4] == 0.0
    # fp32 arithmetic: the property the GPU tests rely on (orthogonality independent of cond)
    a32 = oracle.latms_matrix(4000, 32, 1e6, seed=4)
    q, r = ts_flat.tsqr(a32, 2, dtype=np.float32)
    assert oracle.orthogonality(q.astype(np.float64)) < 5e-6


def test_latms_condition_number(oracle):
    a = oracle.latms_matrix(500, 16, 1e4, seed=1).astype(np.float64)
    s = np.linalg.svd(a, compute_uv=False)
    assert 0.5e4 < s[0] / s[-1] < 2e4


def test_pc_tsqr_restatement_matches_lapack_quality(oracle):
    """numpy restatement of the preconditioned-Cholesky path (oracle/pc_tsqr.py), fp32 arithmetic: LAPACK-class
    orthogonality and residual whatever cond(A) is, and the certificate separates the inputs the sample
    cannot represent."""
    from oracle import pc_tsqr as pc
    for a in (oracle.uniform_matrix(1 << 15, 64, seed=1), oracle.latms_matrix(1 << 15, 64, 1e6, seed=2),
              oracle.uniform_matrix(20000, 7, seed=3), oracle.latms_matrix(1 << 15, 32, 1e4, seed=5)):
        q, r, est, ok = pc.pc_tsqr(a)
        assert ok and est < 2.0
        q64, r64, a64 = q.astype(np.float64), r.astype(np.float64), a.astype(np.float64)
        assert oracle.orthogonality(q64) < 5e-7 and oracle.residual(q64, r64, a64) < 5e-7
        assert np.abs(np.tril(r64, -1)).max() == 0.0
        _, rl = oracle.lapack_qr(a64, np.float64)
        assert oracle.r_rel_diff(r64, rl) < 2e-5
    m, n = 1 << 15, 64
    bad = []
    a = oracle.uniform_matrix(m, n, seed=9); a *= np.logspace(0, -8, m).astype(np.float32)[:, None]; bad.append(a)
    a = oracle.uniform_matrix(m, n, seed=9); a[pc.sample_rows(m, n)] = 0.0; bad.append(a)
    a = oracle.uniform_matrix(m, n, seed=9); a[:, 11] = 0.0; bad.append(a)
    a = oracle.uniform_matrix(m, n, seed=9); a[:, 20] = a[:, 3]; bad.append(a)
    for a in bad:
        assert not pc.pc_tsqr(a)[3]
