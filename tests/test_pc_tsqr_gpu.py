"""GPU tests of the preconditioned-Cholesky TSQR passes (tsqr_gpu_b200/csrc/pc_tsqr.cu), the fast path of the
*_tc_* modes (tensor-core passes) and of fp32_notc / fp16_notc (the same passes on the CUDA cores) for m >= 16384:
  * probe: the two passes with caller-supplied R_pre / C against fp64 numpy -- isolates the TMA tile load,
    the tensor-memory A operand of the substitution / Q products and the 128-byte-swizzled Gram operand;
  * end to end through the C ABI against host LAPACK on uniform and kappa = 1e6 inputs (fast path taken);
  * inputs the row sample cannot represent (graded rows, leading zero rows, a zero column): the certificate
    must fail on the device and the Householder passes must deliver the same accuracy.
Metric definitions: reference src/test.cu:146-165 (residual), src/validation.cu:43-80 (orthogonality)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _np_of(n):
    return 16 if n <= 16 else (32 if n <= 32 else 64)


def _probe_inputs(m, n, seed):
    rng = np.random.default_rng(seed)
    NP = _np_of(n)
    a = rng.uniform(-1, 1, (m, n)).astype(np.float32)
    rpre = np.eye(NP, dtype=np.float32)
    r = np.triu(rng.uniform(-1, 1, (n, n))).astype(np.float32)
    r[np.arange(n), np.arange(n)] = np.sign(r[np.arange(n), np.arange(n)] + 0.5) * (2.0 + rng.uniform(0, 1, n))
    rpre[:n, :n] = r
    c = np.zeros((NP, NP), dtype=np.float32)
    c[:n, :n] = rng.uniform(-1, 1, (n, n))
    c[np.arange(n, NP), np.arange(n, NP)] = 1.0
    return a, rpre, c


@pytest.mark.parametrize("use_tma", [True, False])
@pytest.mark.parametrize("mode,m,n,lda_pad", [
    ("fp32_tc_cor", 16384, 64, 0), ("fp32_tc_cor", 20000, 64, 4), ("tf32_tc_cor", 16500, 32, 0),
    ("fp32_tc_cor", 16400, 16, 8), ("fp32_tc_cor", 17000, 40, 0), ("fp32_tc_nocor", 16384, 64, 0),
    ("fp16_tc_nocor", 16384, 64, 8), ("fp32_tc_cor", 1 << 18, 64, 0),
    ("fp32_notc", 20000, 64, 4), ("fp32_notc", 16500, 32, 0), ("fp32_notc", 17000, 10, 4), ("fp16_notc", 16384, 48, 8)])
def test_probe_passes_match_numpy(cuda, mode, m, n, lda_pad, use_tma):
    import tsqr_gpu_b200 as tq
    a, rpre, c = _probe_inputs(m, n, seed=m + n)
    dt = tq.io_dtype(mode)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda().to(dt), ld=m + lda_pad)
    a64 = a_dev.float().cpu().numpy().astype(np.float64)
    q, gsum = tq.pc_probe(a_dev, mode, torch.from_numpy(rpre).cuda(), torch.from_numpy(c).cuda(), use_tma=use_tma)
    torch.cuda.synchronize()
    NP = _np_of(n)
    z = np.linalg.solve(rpre[:n, :n].astype(np.float64).T, a64.T).T            # Z = A R_pre^-1
    qref = z @ c[:n, :n].astype(np.float64)
    cor = mode in ("fp32_tc_cor", "tf32_tc_cor")
    tol = 2e-6 if (cor or mode == "fp32_notc") else 3e-3
    qerr = np.abs(q.float().cpu().numpy()[:, :n] - qref).max() / np.abs(qref).max()
    g = gsum.cpu().numpy()
    hh, lh = g[:NP, :NP], g[NP:, :NP]
    gram = 0.5 * (hh + hh.T) + ((lh + lh.T) if cor else 0.0)
    gref = z.T @ z
    gerr = np.abs(gram[:n, :n] - gref).max() / np.abs(gref).max()
    print(f"{mode} {m}x{n} tma={use_tma}: Q err {qerr:.2e}, Gram err {gerr:.2e}")
    assert qerr < tol, "pass 2 (tile load, substitution products, Q product)"
    assert gerr < tol, "pass 1 (swizzled Gram operand, accumulation)"


def _run(oracle, a, mode):
    import tsqr_gpu_b200 as tq
    m, n = a.shape
    dt = tq.io_dtype(mode)
    a_dev = tq.to_colmajor(torch.from_numpy(np.ascontiguousarray(a)).cuda().to(dt))
    a_in = a_dev.float().cpu().numpy().astype(np.float64)
    buf = tq.Buffer().allocate(m, n, True)
    q, r = tq.tsqr(a_dev, mode, buffer=buf)
    torch.cuda.synchronize()
    st, est = tq.pc_status(buf, m, n, mode)
    q64, r64 = q.float().cpu().numpy().astype(np.float64), r.float().cpu().numpy().astype(np.float64)
    assert np.isfinite(q64).all() and np.isfinite(r64).all()
    assert np.abs(np.tril(r64, -1)).max() == 0.0
    _, rl = oracle.lapack_qr(a_in, np.float64)
    return st, est, oracle.orthogonality(q64), oracle.residual(q64, r64, a_in), oracle.r_rel_diff(r64, rl)


@pytest.mark.parametrize("mode", ["fp32_tc_cor", "tf32_tc_cor", "fp32_notc"])
@pytest.mark.parametrize("m,n", [(16384, 64), (70001, 64), (1 << 18, 32), (40000, 16), (1 << 17, 48), (30000, 7)])
def test_fast_path_matches_lapack(cuda, oracle, mode, m, n):
    a = oracle.uniform_matrix(m, n, seed=500 + n)
    st, est, orth, res, rd = _run(oracle, a, mode)
    print(f"{mode} {m}x{n}: status {st} est {est:.3f} orth {orth:.2e} res {res:.2e} R {rd:.2e}")
    assert st == 1, "homogeneous rows must pass the certificate"
    assert orth < 2e-6 and res < 2e-6 and rd < 2e-5


@pytest.mark.parametrize("mode", ["fp32_tc_nocor", "tf32_tc_nocor", "fp16_tc_nocor", "fp16_notc"])
def test_fast_path_uncorrected_modes(cuda, oracle, mode):
    a = oracle.uniform_matrix(1 << 17, 64, seed=77)
    st, est, orth, res, rd = _run(oracle, a, mode)
    print(f"{mode}: status {st} est {est:.3f} orth {orth:.2e} res {res:.2e} R {rd:.2e}")
    assert st == 1 and orth < 5e-3 and res < 5e-3 and rd < 5e-3


@pytest.mark.parametrize("mode", ["fp32_tc_nocor", "fp16_tc_nocor"])
def test_fast_path_uncorrected_modes_conditioned(cuda, oracle, mode):
    """kappa = 1e6 in a mode whose products carry 2^-11 errors: Q stays orthogonal to the mode's level only if both
    passes compute bit-identical Z = A R_pre^-1 (their difference would be amplified by cond(R_pre))."""
    a = oracle.latms_matrix(1 << 17, 64, 1e6, seed=33)
    st, est, orth, res, _ = _run(oracle, a, mode)
    print(f"{mode} cond 1e6: status {st} est {est:.3f} orth {orth:.2e} res {res:.2e}")
    assert orth < 5e-3 and res < 5e-3


@pytest.mark.parametrize("mode", ["fp32_tc_cor", "fp32_notc"])
@pytest.mark.parametrize("cond", [1e3, 1e6])
def test_fast_path_conditioned(cuda, oracle, cond, mode):
    a = oracle.latms_matrix(1 << 17, 64, cond, seed=31)
    st, est, orth, res, _ = _run(oracle, a, mode)
    print(f"{mode} cond {cond:g}: status {st} est {est:.3f} orth {orth:.2e} res {res:.2e}")
    assert st == 1 and orth < 2e-6 and res < 2e-6


@pytest.mark.parametrize("mode", ["fp32_tc_cor", "fp32_notc"])
@pytest.mark.parametrize("case", ["graded_rows", "zero_head", "zero_column", "rank_deficient"])
def test_certificate_rejects_and_householder_takes_over(cuda, oracle, case, mode):
    m, n = 1 << 16, 64
    a = oracle.uniform_matrix(m, n, seed=9)
    if case == "graded_rows":
        a *= np.logspace(0, -8, m).astype(np.float32)[:, None]
    elif case == "zero_head":        # every sampled row is zero
        from oracle import pc_tsqr as pc
        a[pc.sample_rows(m, n)] = 0.0
    elif case == "zero_column":
        a[:, 11] = 0.0
    else:
        a[:, 20] = a[:, 3]
    st, est, orth, res, _ = _run(oracle, a, mode)
    print(f"{mode} {case}: status {st} est {est} orth {orth:.2e} res {res:.2e}")
    if case in ("zero_head", "zero_column"):   # the other two may or may not pass the certificate (a duplicated column
        assert st == 0, "the certificate must reject this input"   # leaves a rounding-level pivot: a valid QR either way)
    assert res < 5e-6
    if case != "rank_deficient":     # Q of a rank-deficient matrix is still orthogonal with Householder reflectors
        assert orth < 5e-6


def test_in_place_and_reduce_build_split(cuda, oracle):
    import tsqr_gpu_b200 as tq
    m, n = 50000, 64
    a = oracle.uniform_matrix(m, n, seed=4)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda())
    q, r = tq.tsqr(a_dev, "fp32_tc_cor")
    from tsqr_gpu_b200._lib import lib
    ws = torch.empty(lib.tsqr_b200_tsqr_workspace_bytes(m, n), dtype=torch.uint8, device="cuda")
    # reduce / build with B = a rotation: Q_local * B
    r_loc, qbuf = tq.tsqr_reduce(a_dev, "fp32_tc_cor", ws)
    b = torch.linalg.qr(torch.randn(n, n, device="cuda", dtype=torch.float64))[0].float()
    q2 = tq.tsqr_build(a_dev, "fp32_tc_cor", ws, b=tq.to_colmajor(b), q=qbuf)
    torch.cuda.synchronize()
    assert torch.allclose(r_loc, r, atol=0, rtol=0)
    ref = q.double() @ b.double()
    assert (q2.double() - ref).abs().max().item() < 2e-6
    # in place: Q over A
    a2 = a_dev.clone()
    q3, r3 = tq.tsqr(a2, "fp32_tc_cor", q=a2)
    torch.cuda.synchronize()
    assert torch.equal(q3, q) and torch.equal(r3, r)
