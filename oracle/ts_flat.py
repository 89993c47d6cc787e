"""ORACLE -- TEST INFRASTRUCTURE ONLY (imported by tests/ only).

numpy restatement of the NEW library's chunk algorithm (tsqr_gpu_b200/csrc/ts_pass.h, TcParams),
the B200 replacement of the reference's leaf-QR + binary tree (reference src/tsqr.cu:1064-1279,
src/tcqr32x16.cu:1373-1469). It pins the mathematics the CUDA kernels implement:

  reduce: head = plain Householder QR of the first n rows (R0, V0); every following tile of
          `tile` rows is folded in as [R ; A_tile] with structured reflectors [e_j ; v_j]
          (v_j = x_j / (alpha - beta), tau_j = 1 + |alpha| / norm); the tails v_j are kept (V).
  build : tiles in reverse: S = V^T V, T^-1 = striu(S) + diag((1 + S_jj) / 2), T^-1 X = B,
          Q_tile = -V X, B <- B - X; finally Q_head = H_0 ... H_{n-1} B with
          tau_j = 2 / (1 + |v_j|^2).
PARITY PINNING: checked against host LAPACK in tests/test_oracle.py (orthogonality, residual and
R up to row signs); the reference has no golden vectors for this path.
"""
import numpy as np


def _hh(alpha, g):
    nx2 = alpha * alpha + g
    if nx2 <= 0:
        return 0.0, 2.0, 0.0
    norm = np.sqrt(nx2)
    beta = -np.copysign(norm, alpha)
    return beta, 1.0 + abs(alpha) / norm, 1.0 / (alpha - beta)


def reduce_chunk(a, tile=128, dtype=np.float64):
    """a: (rows, n), rows >= 1. Returns (R (n, n), V (rows, n)) in the layout the kernels write:
    rows [0, n) hold the strictly lower part of V0, the rest the reflector tails of each tile."""
    a = np.array(a, dtype=dtype)
    rows, n = a.shape
    head = np.zeros((n, n), dtype)
    head[:min(n, rows)] = a[:n]
    V = np.zeros_like(a)
    for j in range(n):
        beta, tau, rden = _hh(head[j, j], float(head[j + 1:, j] @ head[j + 1:, j]))
        v = head[j + 1:, j] * rden
        for c in range(j + 1, n):
            coef = tau * (head[j, c] + v @ head[j + 1:, c])
            head[j, c] -= coef
            head[j + 1:, c] -= coef * v
        head[j, j] = beta
        head[j + 1:, j] = v
    R = np.triu(head)
    V[:min(n, rows)] = np.tril(head, -1)[:min(n, rows)]
    for r0 in range(n, rows, tile):
        t = a[r0:r0 + tile].copy()
        for j in range(n):
            beta, tau, rden = _hh(R[j, j], float(t[:, j] @ t[:, j]))
            v = t[:, j] * rden
            for c in range(j + 1, n):
                coef = tau * (R[j, c] + v @ t[:, c])
                R[j, c] -= coef
                t[:, c] -= coef * v
            R[j, j] = beta
            t[:, j] = v
        V[r0:r0 + tile] = t
    return R, V


def build_chunk(V, B=None, tile=128, dtype=np.float64):
    """Q (rows, n) = Q_chunk [B ; 0] from the stored reflectors."""
    V = np.array(V, dtype=dtype)
    rows, n = V.shape
    B = np.eye(n, dtype=dtype) if B is None else np.array(B, dtype=dtype)
    Q = np.zeros_like(V)
    starts = list(range(n, rows, tile))
    for r0 in reversed(starts):
        v = V[r0:r0 + tile]
        S = v.T @ v
        Tinv = np.triu(S, 1) + np.diag((1.0 + np.diag(S)) / 2.0)
        X = np.linalg.solve(Tinv, B)
        Q[r0:r0 + tile] = -v @ X
        B = B - X
    h = min(n, rows)
    V0 = np.zeros((n, n), dtype)
    V0[:h] = V[:h]
    for j in range(n - 1, -1, -1):
        u = V0[:, j].copy()
        u[:j] = 0
        u[j] = 1.0
        tau = 2.0 / (u @ u)
        B = B - tau * np.outer(u, u @ B)
    Q[:h] = B[:h]
    return Q


def tsqr(a, nchunks=1, tile=128, dtype=np.float64):
    """two-level TSQR with the chunk algorithm: returns (Q, R)."""
    a = np.asarray(a)
    m, n = a.shape
    bounds = [m * i // nchunks for i in range(nchunks + 1)]
    parts = [reduce_chunk(a[bounds[i]:bounds[i + 1]], tile, dtype) for i in range(nchunks)]
    if nchunks == 1:
        return build_chunk(parts[0][1], None, tile, dtype), parts[0][0]
    stacked = np.vstack([p[0] for p in parts])
    Rtop, Vtop = reduce_chunk(stacked, tile, dtype)
    Qtop = build_chunk(Vtop, None, tile, dtype)
    Q = np.vstack([build_chunk(parts[i][1], Qtop[i * n:(i + 1) * n], tile, dtype) for i in range(nchunks)])
    return Q, Rtop
