"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by tsqr_gpu_b200/).

numpy restatement of the preconditioned-Cholesky TSQR that tsqr_gpu_b200/csrc/pc_tsqr.cu runs for the *_tc_*
modes (the executable specification of that path, like oracle/ts_flat.py is for the Householder passes):

    sample : h = 8 NP rows of A (h/16 clusters of 16 rows) -> R_pre (Householder QR of the sample)
    pass 1 : Z = A R_pre^-1 (substitution), G = Z^T Z      (per 128-row tile, fp32; partial sums in fp64)
    finish : R_z = chol(G); est = |G|_F |G^-1|_F / n; R = R_z R_pre; C = R_z^-1
    pass 2 : Q = Z C

It replaces the reference's leaf QR + R tree + backward products (src/tcqr32x16.cu:1373-1469,
src/tsqr.cu:1102-1259) for tall matrices whose rows the sample represents; `est` is the on-device certificate:
est = 1 for G ~ c I and n * est >= cond_2(G). PARITY: this file restates the NEW algorithm, not the reference;
it is pinned against host LAPACK through the reference's metric definitions (tests/test_oracle.py)."""
import numpy as np
import scipy.linalg as sl

MAX_EST = 6.0   # kPcMaxEst of csrc/pc_tsqr.h


def padded(n):
    return 16 if n <= 16 else (32 if n <= 32 else 64)


def sample_rows(m, n):
    """h = 8 NP rows: h / 16 clusters of 16 consecutive rows spread evenly over the matrix (pc_sample_kernel)"""
    h = 8 * padded(n)
    ncl = h // 16
    t = np.arange(h, dtype=np.int64)
    return ((t // 16) * (m - 16)) // max(ncl - 1, 1) + (t % 16)


def pc_tsqr(a, work=np.float32, tile=128):
    """returns q, r, est, ok. `work` is the arithmetic type of the two passes (float32 on the device)."""
    m, n = a.shape
    rp = np.linalg.qr(a[sample_rows(m, n)].astype(work), mode="r").astype(work)
    with np.errstate(all="ignore"):
        try:
            z = sl.solve_triangular(rp.T, a.astype(work).T, lower=True, check_finite=False).T.astype(work)
        except np.linalg.LinAlgError:          # exactly singular R_pre: the device gets inf/NaN and rejects
            return None, None, np.inf, False
        g = np.zeros((n, n), dtype=np.float64)
        for i in range(0, m, tile * 16):          # fp32 sums over 16 tiles, fp64 across them (the device's scheme)
            blk = z[i:i + tile * 16]
            g += (blk.T @ blk).astype(np.float64)
        ok = bool(np.isfinite(g).all())
        est = np.inf
        rz = np.eye(n)
        if ok:
            try:
                rz = np.linalg.cholesky(g).T
                gi = np.linalg.inv(g)
                est = np.linalg.norm(g) * np.linalg.norm(gi) / n
            except np.linalg.LinAlgError:
                ok = False
        ok = ok and est <= MAX_EST
        if not ok:
            return None, None, est, False
        r = (rz @ rp.astype(np.float64)).astype(work)
        c = np.linalg.inv(rz).astype(work)
        q = (z @ c).astype(work)
    return q, r, float(est), True
