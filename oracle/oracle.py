"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module, and only as the checker -- never the product path (tsqr_gpu_b200/).

Contents
  * lapack_qr            host LAPACK sgeqrf+sorgqr (dgeqrf+dorgqr for the fp64 truth) through
                         scipy's bundled OpenBLAS: the independent CPU oracle named by
                         BASELINE.json ("host LAPACK geqrf+orgqr on the same inputs").
  * ref_tsqr16 / ref_blockqr / ref_qr32x16
                         ctypes binding of oracle/ref_tsqr.c, the plain-C restatement of the
                         reference's fp32_notc algorithm (citations in that file).
  * orthogonality / residual / sign_normalize / r_rel_diff
                         the reference's metric definitions: residual ||QR-A||_F/||A||_F
                         (reference src/test.cu:146-165), orthogonality ||Q^T Q - I||_F/sqrt(n)
                         in fp64 (reference src/validation.cu:43-80), sign-insensitive Q/R
                         comparison (reference src/test_compare.hpp:236-255).
  * uniform_matrix / latms_matrix
                         seeded input generators: uniform(-1,1) (reference src/test.cu:97,
                         121-126) and LAPACK slatms-style A = U diag(s) V^T with prescribed
                         condition number (reference src/latms.cu:8-121, src/test_cond.cu:31-50).

PARITY PINNING: the reference holds no golden vectors and has no CPU path, so neither oracle
can be pinned against reference outputs in the CPU container. See oracle/ref_tsqr.c header.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def _load():
    if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "ref_tsqr.c")):
        build()
    lib = ctypes.CDLL(_LIB)
    fp = ctypes.POINTER(ctypes.c_float)
    sz = ctypes.c_size_t
    lib.ref_qr32x16.argtypes = [fp, sz, fp, sz, fp, sz, ctypes.c_uint, ctypes.c_uint]
    lib.ref_qr32x16.restype = None
    lib.ref_tsqr16.argtypes = [fp, sz, fp, sz, fp, sz, sz, sz]
    lib.ref_tsqr16.restype = None
    lib.ref_blockqr.argtypes = [fp, sz, fp, sz, fp, sz, sz, sz, ctypes.c_int]
    lib.ref_blockqr.restype = ctypes.c_int
    lib.ref_batch_size_log2.argtypes = [sz]
    lib.ref_batch_size_log2.restype = sz
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def _fptr(x):
    return x.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _colmajor(a):
    return np.asfortranarray(a, dtype=np.float32)


def ref_qr32x16(a):
    a = _colmajor(a)
    m, n = a.shape
    q = np.zeros((m, n), np.float32, order="F")
    r = np.zeros((n, n), np.float32, order="F")
    lib().ref_qr32x16(_fptr(q), m, _fptr(r), n, _fptr(a), m, m, n)
    return q, r


def ref_tsqr16(a):
    a = _colmajor(a)
    m, n = a.shape
    assert n <= 16
    q = np.zeros((m, n), np.float32, order="F")
    r = np.zeros((n, n), np.float32, order="F")
    lib().ref_tsqr16(_fptr(q), m, _fptr(r), n, _fptr(a), m, m, n)
    return q, r


def ref_blockqr(a, reorth=False):
    a = _colmajor(a).copy(order="F")
    m, n = a.shape
    q = np.zeros((m, n), np.float32, order="F")
    r = np.zeros((n, n), np.float32, order="F")
    st = lib().ref_blockqr(_fptr(q), m, _fptr(r), n, _fptr(a), m, m, n, 1 if reorth else 0)
    return st, q, r


def set_lapack_threads(n):
    try:
        from threadpoolctl import threadpool_limits
        return threadpool_limits(limits=n)
    except Exception:
        return None


def lapack_qr(a, dtype=np.float32):
    """thin QR through LAPACK geqrf + orgqr"""
    from scipy.linalg import lapack
    a = np.asfortranarray(a, dtype=dtype)
    m, n = a.shape
    geqrf, orgqr = (lapack.sgeqrf, lapack.sorgqr) if dtype == np.float32 else (lapack.dgeqrf, lapack.dorgqr)
    qr_, tau, _, info = geqrf(a, overwrite_a=False)
    assert info == 0
    r = np.triu(qr_[:n, :n])
    q, _, info = orgqr(qr_, tau, overwrite_a=True)
    assert info == 0
    return q[:, :n], r


# ---------------------------------------------------------------- metrics
def orthogonality(q):
    q = np.asarray(q, dtype=np.float64)
    n = q.shape[1]
    return float(np.linalg.norm(q.T @ q - np.eye(n)) / np.sqrt(n))


def residual(q, r, a):
    q = np.asarray(q, dtype=np.float64)
    r = np.triu(np.asarray(r, dtype=np.float64))
    a = np.asarray(a, dtype=np.float64)
    return float(np.linalg.norm(q @ r - a) / np.linalg.norm(a))


def sign_normalize(q, r):
    """flip signs so that diag(R) >= 0 (R rows, Q columns)"""
    r = np.triu(np.asarray(r, dtype=np.float64))
    s = np.where(np.diag(r) < 0, -1.0, 1.0)
    qn = None if q is None else np.asarray(q, dtype=np.float64) * s[None, :]
    return qn, r * s[:, None]


def r_rel_diff(r1, r2):
    _, a = sign_normalize(None, r1)
    _, b = sign_normalize(None, r2)
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


# ---------------------------------------------------------------- inputs
def uniform_matrix(m, n, seed, dtype=np.float32):
    rng = np.random.default_rng(seed)
    return np.asfortranarray(rng.uniform(-1.0, 1.0, size=(m, n)).astype(dtype))


def latms_matrix(m, n, cond, seed, dtype=np.float32):
    """A = U diag(s) V^T, U (m x n), V (n x n) Haar-like from QR of Gaussians, singular values
    geometric from 1 down to 1/cond (LAPACK slatms MODE=3)."""
    rng = np.random.default_rng(seed)
    u, _ = np.linalg.qr(rng.standard_normal((m, n)))
    v, _ = np.linalg.qr(rng.standard_normal((n, n)))
    s = cond ** (-np.arange(n) / max(n - 1, 1))
    return np.asfortranarray(((u * s[None, :]) @ v.T).astype(dtype))


def slatms_matrix(m, n, cond, iseed=(1, 2, 3, 5)):
    """REAL LAPACK slatms (scipy's bundled OpenBLAS, symbol scipy_slatms_): MODE = 3 (geometric singular values
    from 1 to 1/cond), DMAX = 1, full bandwidth, fixed ISEED -- the generator BASELINE.json's north_star names
    ("latms-generated inputs"; the reference's own stand-in is src/latms.cu:8-121 + src/test_cond.cu:31-50)."""
    import glob
    import scipy
    libs = glob.glob(os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs", "libscipy_openblas*"))
    if not libs:
        raise RuntimeError("scipy's bundled OpenBLAS not found")
    fn = ctypes.CDLL(libs[0]).scipy_slatms_
    fn.restype = None
    a = np.zeros((m, n), dtype=np.float32, order="F")
    d = np.zeros(min(m, n), dtype=np.float32)
    work = np.zeros(3 * max(m, n), dtype=np.float32)
    seed = (ctypes.c_int * 4)(*iseed)                      # ISEED(4) must be odd
    info = ctypes.c_int(0)
    ci = lambda v: ctypes.byref(ctypes.c_int(v))
    cf = lambda v: ctypes.byref(ctypes.c_float(v))
    vp = lambda x: x.ctypes.data_as(ctypes.c_void_p)
    one = ctypes.c_size_t(1)
    fn(ci(m), ci(n), ctypes.c_char_p(b"N"), seed, ctypes.c_char_p(b"N"), vp(d), ci(3), cf(float(cond)), cf(1.0),
       ci(m - 1), ci(n - 1), ctypes.c_char_p(b"N"), vp(a), ci(m), vp(work), ctypes.byref(info), one, one, one)
    if info.value != 0:
        raise RuntimeError(f"slatms failed: info = {info.value}")
    return a
