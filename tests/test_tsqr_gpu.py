"""GPU parity tests of the TSQR path (n <= 64) through the C ABI, against the CPU oracles:
host LAPACK sgeqrf+sorgqr (R after sign normalisation, orthogonality, residual) and the C
restatement of the reference algorithm (metrics no worse than 2x the reference's value)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FP32_EXACT = ["fp32_notc", "fp32_tc_cor", "tf32_tc_cor", "mixed_tc_cor_emu", "tf32_tc_cor_emu"]
ALL_MODES = ["fp16_notc", "fp16_tc_nocor", "fp32_notc", "fp32_tc_cor", "fp32_tc_nocor", "mixed_tc_cor_emu",
             "tf32_tc_cor", "tf32_tc_cor_emu", "tf32_tc_nocor", "tf32_tc_nocor_emu"]


def _check(oracle, q, r, a_in, mode, r_tol=None):
    from helpers import TOL
    orth, res = oracle.orthogonality(q), oracle.residual(q, r, a_in)
    to, tr = TOL[mode]
    assert np.isfinite(q).all() and np.isfinite(r).all()
    assert orth < to, f"orthogonality {orth:.3e} (mode {mode})"
    assert res < tr, f"residual {res:.3e} (mode {mode})"
    assert np.abs(np.tril(r, -1)).max() == 0.0
    if r_tol is not None:
        _, rl = oracle.lapack_qr(a_in, np.float64)
        d = oracle.r_rel_diff(r, rl)
        assert d < r_tol, f"R differs from LAPACK by {d:.3e}"
    return orth, res


@pytest.mark.parametrize("m,n", [(16384, 16), (1000, 16), (9211, 16), (300, 7), (20, 16), (33, 1), (5000, 32),
                                 (4097, 48), (3000, 64), (70000, 64), (64, 64), (131072, 24)])
def test_fp32_notc_matches_lapack(cuda, oracle, m, n):
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=100 + n)
    q, r, a_in = run_qr(a, "fp32_notc", entry="tsqr")
    # BASELINE.json: "fp32_notc agrees to ~1e-5 relative" with LAPACK
    _check(oracle, q, r, a_in, "fp32_notc", r_tol=1e-5)


@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("m,n", [(16384, 16), (40000, 64)])
def test_all_modes(cuda, oracle, mode, m, n):
    from helpers import run_qr
    a = oracle.uniform_matrix(m, n, seed=7)
    q, r, a_in = run_qr(a, mode, entry="qr")
    _check(oracle, q, r, a_in, mode, r_tol=1e-5 if mode in FP32_EXACT else 5e-3)


def test_config1_vs_reference_restatement(cuda, oracle):
    """BASELINE config 1: M=16384 N=16 fp32_notc. Metrics no worse than 2x the reference
    algorithm's (C restatement) on the same input; R equal up to row signs."""
    from helpers import run_qr
    a = oracle.uniform_matrix(16384, 16, seed=0x5EED0001)
    q, r, a_in = run_qr(a, "fp32_notc")
    qo, ro = oracle.ref_tsqr16(a)
    assert oracle.orthogonality(q) <= 2 * oracle.orthogonality(qo)
    assert oracle.residual(q, r, a_in) <= 2 * oracle.residual(qo, ro, a)
    assert oracle.r_rel_diff(r, ro) < 1e-5
    qn, _ = oracle.sign_normalize(q, r)
    qon, _ = oracle.sign_normalize(qo, ro)
    assert np.abs(qn - qon).max() < 1e-5


@pytest.mark.parametrize("cond", [1e3, 1e6])
@pytest.mark.parametrize("n", [16, 64])
def test_conditioned_inputs(cuda, oracle, cond, n):
    """latms-style inputs: Householder TSQR keeps orthogonality independent of cond(A)."""
    from helpers import run_qr
    a = oracle.latms_matrix(20000, n, cond, seed=11)
    q, r, a_in = run_qr(a, "fp32_notc")
    orth, res = _check(oracle, q, r, a_in, "fp32_notc")
    ql, rl = oracle.lapack_qr(a)
    assert orth <= max(4 * oracle.orthogonality(ql), 2e-6)
    assert res <= max(4 * oracle.residual(ql, rl, a), 2e-6)


def test_leading_dimensions_and_no_clobber(cuda, oracle):
    import tsqr_gpu_b200 as tq
    m, n = 3001, 40
    a = oracle.uniform_matrix(m, n, seed=3)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda(), ld=m + 37)
    a_copy = a_dev.clone()
    q = tq.colmajor_empty(m, n, torch.float32, ld=m + 5)
    r = torch.full((n, n + 3), 7.0, device="cuda").t()[:n]  # shape (n, n), stride (1, n + 3)
    qq, rr = tq.tsqr(a_dev, "fp32_notc", q=q, r=r)
    torch.cuda.synchronize()
    assert torch.equal(a_dev, a_copy), "tsqr must not modify A"
    _check(oracle, qq.cpu().numpy().astype(np.float64), rr.cpu().numpy().astype(np.float64), a.astype(np.float64),
           "fp32_notc", r_tol=1e-5)


def test_zero_and_duplicate_columns(cuda, oracle):
    """the reference produces NaN (0/0) for a zero column (src/tcqr32x16.cu:125); here the
    reflector degenerates to the identity and the factorisation stays finite."""
    from helpers import run_qr
    a = oracle.uniform_matrix(5000, 16, seed=5)
    a[:, 3] = 0.0
    a[:, 9] = a[:, 2]
    q, r, a_in = run_qr(a, "fp32_notc")
    assert np.isfinite(q).all() and np.isfinite(r).all()
    assert oracle.residual(q, r, a_in) < 5e-6
    assert abs(r[3, 3]) == 0.0 and abs(r[9, 9]) < 1e-4


def test_invalid_sizes(cuda):
    import tsqr_gpu_b200 as tq
    a = tq.colmajor_empty(8, 16, torch.float32)
    with pytest.raises(tq.InvalidMatrixSize):
        tq.qr(a, "fp32_notc")
    b = tq.Buffer().allocate(100, 10)
    with pytest.raises(RuntimeError):
        b.allocate(100, 10)
    with pytest.raises(TypeError):
        tq.qr(tq.colmajor_empty(100, 8, torch.float16), "fp32_notc")


def test_bitwise_reproducible(cuda, oracle):
    from helpers import run_qr
    a = oracle.uniform_matrix(50000, 32, seed=9)
    q1, r1, _ = run_qr(a, "fp32_notc")
    q2, r2, _ = run_qr(a, "fp32_notc")
    assert np.array_equal(q1, q2) and np.array_equal(r1, r2)


@pytest.mark.parametrize("P", [2, 4, 8])
def test_row_sharded_virtual_ranks(cuda, oracle, P):
    """row-sharded TSQR (reduce -> gather R -> root QR -> build) emulated on one GPU with P
    slabs; must be a QR of the full matrix."""
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    m, n = 8192 * P, 64
    a = oracle.uniform_matrix(m, n, seed=21)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda())
    mp = m // P
    slabs = [tq.to_colmajor(a_dev[p * mp:(p + 1) * mp]) for p in range(P)]
    wss = [torch.empty(lib.tsqr_b200_tsqr_workspace_bytes(mp, n), dtype=torch.uint8, device="cuda") for _ in range(P)]
    rq = [tq.tsqr_reduce(slabs[p], "fp32_notc", wss[p]) for p in range(P)]
    rs = [x[0] for x in rq]
    stacked = tq.to_colmajor(torch.cat(rs, dim=0))
    qroot, r = tq.tsqr(stacked, "fp32_notc")
    qs = [tq.tsqr_build(slabs[p], "fp32_notc", wss[p], b=tq.to_colmajor(qroot[p * n:(p + 1) * n]), q=rq[p][1]) for p in range(P)]
    torch.cuda.synchronize()
    q = torch.cat(qs, dim=0).cpu().numpy().astype(np.float64)
    _check(oracle, q, r.cpu().numpy().astype(np.float64), a.astype(np.float64), "fp32_notc", r_tol=1e-5)


def test_small_qr_batched(cuda, oracle):
    import tsqr_gpu_b200 as tq
    m, n, batch = 32 * 9 - 5, 16, 9
    a = oracle.uniform_matrix(m, n, seed=31)
    starts = np.array([m * i // batch for i in range(batch + 1)], dtype=np.int32)
    q, r = tq.qr32x16_batched(tq.to_colmajor(torch.from_numpy(a).cuda()), torch.from_numpy(starts).cuda(), n)
    torch.cuda.synchronize()
    q, r = q.cpu().numpy().astype(np.float64), r.cpu().numpy().astype(np.float64)
    for i in range(batch):
        s, e = starts[i], starts[i + 1]
        qi, ri = q[s:e], r[n * i:n * i + n]
        assert oracle.orthogonality(qi) < 2e-6
        assert oracle.residual(qi, ri, a[s:e]) < 2e-6
        qo, ro = oracle.ref_qr32x16(a[s:e])
        assert oracle.r_rel_diff(ri, ro) < 1e-5
