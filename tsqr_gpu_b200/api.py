"""Python mirror of mtk::qr::qr / mtk::tsqr::tsqr16 / mtk::qr::buffer (reference
src/blockqr.hpp:59-175, src/tsqr.hpp:49-140) on top of the C ABI.

Matrices are torch CUDA tensors in column-major layout: shape (m, n), stride (1, ld).
`to_colmajor` / `colmajor_empty` make such tensors. Everything runs on the current torch
CUDA stream and never synchronises the host.
"""
import ctypes

import torch

from ._lib import lib

# numeric values of mtk::qr::compute_mode (reference src/tsqr.hpp:9-20)
MODES = {
    "fp16_notc": 0, "fp16_tc_nocor": 1, "fp32_notc": 2, "fp32_tc_cor": 3, "fp32_tc_nocor": 4,
    "mixed_tc_cor_emu": 5, "tf32_tc_cor": 6, "tf32_tc_cor_emu": 7, "tf32_tc_nocor": 8, "tf32_tc_nocor_emu": 9,
}
MODE_NAMES = {v: k for k, v in MODES.items()}
MAX_TSQR_N = 64


class InvalidMatrixSize(ValueError):
    """mtk::qr::error_invalid_matrix_size (state 1, reference src/blockqr.hpp:27-29)."""


def _mode(mode):
    if isinstance(mode, str):
        return MODES[mode]
    mode = int(mode)
    if mode not in MODE_NAMES:
        raise ValueError(f"unknown compute mode {mode}")
    return mode


def io_dtype(mode):
    return torch.float16 if lib.tsqr_b200_io_bytes(_mode(mode)) == 2 else torch.float32


def colmajor_empty(m, n, dtype, device="cuda", ld=None):
    ld = m if ld is None else ld
    return torch.empty((n, ld), dtype=dtype, device=device).t()[:m]


def to_colmajor(x, ld=None):
    """copy of a 2-D tensor in column-major storage (stride (1, ld))"""
    m, n = x.shape
    out = colmajor_empty(m, n, x.dtype, x.device, ld)
    out.copy_(x)
    return out


def _check_colmajor(x, name):
    if x.dim() != 2 or not x.is_cuda:
        raise ValueError(f"{name}: expected a 2-D CUDA tensor")
    if x.shape[0] > 1 and x.stride(0) != 1:
        raise ValueError(f"{name}: expected column-major storage (stride(0) == 1), got strides {x.stride()}")
    ld = x.stride(1) if x.shape[1] > 1 else max(x.shape[0], 1)
    if ld < x.shape[0]:
        raise ValueError(f"{name}: leading dimension {ld} < rows {x.shape[0]}")
    return ld


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _raise(code):
    if code == 0:
        return
    if code == 1:
        raise InvalidMatrixSize("invalid matrix size (need 0 < n <= m)")
    raise RuntimeError(f"tsqr_b200 error {code}: {lib.tsqr_b200_last_error().decode()}")


class Buffer:
    """Work buffer, the analogue of mtk::qr::buffer<mode, Reorthogonalize>::allocate(m, n).
    One buffer per in-flight call. allocate twice -> RuntimeError like the reference."""

    def __init__(self):
        self.ws = None
        self.total_memory_size = 0

    def allocate(self, m, n, reorthogonalize=True, device="cuda"):
        if self.ws is not None:
            raise RuntimeError("The buffer has been already allocated")
        nbytes = lib.tsqr_b200_qr_workspace_bytes(m, n, 1 if reorthogonalize else 0)
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.total_memory_size = nbytes
        return self

    def destroy(self):
        self.ws = None
        self.total_memory_size = 0

    def get_device_memory_size(self):
        return self.total_memory_size


def _workspace(buffer, m, n, reorth, device):
    if buffer is None:
        buffer = Buffer().allocate(m, n, reorth, device)
    elif buffer.ws is None:
        buffer.allocate(m, n, reorth, device)
    return buffer


def qr(a, mode="fp32_notc", reorthogonalize=False, buffer=None, q=None, r=None):
    """mtk::qr::qr<mode, reorthogonalize>: returns (q, r). `a` is overwritten when n > 64."""
    md = _mode(mode)
    m, n = a.shape
    if n > m or m == 0 or n == 0:
        raise InvalidMatrixSize("invalid matrix size (need 0 < n <= m)")
    if a.dtype != io_dtype(md):
        raise TypeError(f"mode {MODE_NAMES[md]} takes {io_dtype(md)} matrices, got {a.dtype}")
    lda = _check_colmajor(a, "a")
    if q is None:
        q = colmajor_empty(m, n, a.dtype, a.device)
    if r is None:
        r = torch.zeros((n, n), dtype=a.dtype, device=a.device).t()
    ldq = _check_colmajor(q, "q")
    ldr = _check_colmajor(r, "r")
    buffer = _workspace(buffer, m, n, reorthogonalize, a.device)
    code = lib.tsqr_b200_qr(md, 1 if reorthogonalize else 0, q.data_ptr(), ldq, r.data_ptr(), ldr, a.data_ptr(), lda,
                            m, n, buffer.ws.data_ptr(), buffer.ws.numel(), _stream())
    _raise(code)
    return q, r


def tsqr(a, mode="fp32_notc", buffer=None, q=None, r=None):
    """mtk::tsqr::tsqr16<mode> widened to n <= 64: returns (q, r); `a` is not modified."""
    md = _mode(mode)
    m, n = a.shape
    if a.dtype != io_dtype(md):
        raise TypeError(f"mode {MODE_NAMES[md]} takes {io_dtype(md)} matrices, got {a.dtype}")
    lda = _check_colmajor(a, "a")
    if q is None:
        q = colmajor_empty(m, n, a.dtype, a.device)
    if r is None:
        r = torch.zeros((n, n), dtype=a.dtype, device=a.device).t()
    ldq = _check_colmajor(q, "q")
    ldr = _check_colmajor(r, "r")
    if buffer is None:
        nbytes = lib.tsqr_b200_tsqr_workspace_bytes(m, n)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=a.device)
    else:
        ws = _workspace(buffer, m, n, True, a.device).ws
    code = lib.tsqr_b200_tsqr(md, q.data_ptr(), ldq, r.data_ptr(), ldr, a.data_ptr(), lda, m, n, ws.data_ptr(),
                              ws.numel(), _stream())
    _raise(code)
    return q, r


def tsqr_reduce(a, mode, ws, q=None):
    """step 1 of the row-sharded TSQR: local slab -> (local R (n x n float32, column-major), q).
    The same q must be handed to tsqr_build: the tensor-core modes keep the reflectors in it."""
    md = _mode(mode)
    m, n = a.shape
    lda = _check_colmajor(a, "a")
    if q is None:
        q = colmajor_empty(m, n, a.dtype, a.device)
    ldq = _check_colmajor(q, "q")
    r_local = torch.empty((n, n), dtype=torch.float32, device=a.device).t()
    code = lib.tsqr_b200_tsqr_reduce(md, a.data_ptr(), lda, m, n, r_local.data_ptr(), q.data_ptr(), ldq,
                                     ws.data_ptr(), ws.numel(), _stream())
    _raise(code)
    return r_local, q


def tsqr_build(a, mode, ws, b=None, q=None):
    """step 3 of the row-sharded TSQR: Q_local * B (B n x n float32 column-major or None)"""
    md = _mode(mode)
    m, n = a.shape
    lda = _check_colmajor(a, "a")
    if q is None:
        q = colmajor_empty(m, n, a.dtype, a.device)
    ldq = _check_colmajor(q, "q")
    bptr, ldb = None, 0
    if b is not None:
        if b.dtype != torch.float32:
            raise TypeError("b must be float32")
        ldb = _check_colmajor(b, "b")
        bptr = b.data_ptr()
    code = lib.tsqr_b200_tsqr_build(md, q.data_ptr(), ldq, a.data_ptr(), lda, m, n, bptr, ldb, ws.data_ptr(),
                                    ws.numel(), _stream())
    _raise(code)
    return q


def qr32x16_batched(a, starts, n, mode="fp32_notc"):
    """mtk::tcqr::qr32x16_batched: `a` (m x n col-major), `starts` int32 tensor of batch+1 row offsets.
    Returns (q (m x n), r ((n*batch) x n))."""
    md = _mode(mode)
    m = a.shape[0]
    batch = starts.numel() - 1
    lda = _check_colmajor(a, "a")
    q = colmajor_empty(m, n, a.dtype, a.device)
    r = colmajor_empty(n * batch, n, a.dtype, a.device)
    code = lib.tsqr_b200_qr32x16_batched(md, q.data_ptr(), m, r.data_ptr(), n * batch, a.data_ptr(), lda, m, n, batch,
                                         starts.data_ptr(), _stream())
    _raise(code)
    return q, r


_cudart = None


def _memcpy2d_async(dst, src, stream):
    """Strided 2-D copy between a pinned host tensor and a device tensor (both (rows, cols) with unit
    column stride) as ONE cudaMemcpy2DAsync on `stream`: torch's copy_ would stage a non-contiguous
    host slice through pageable memory."""
    global _cudart
    if _cudart is None:
        _cudart = ctypes.CDLL("libcudart.so.12")
        _cudart.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                              ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
        _cudart.cudaMemcpy2DAsync.restype = ctypes.c_int
    if dst.shape != src.shape or dst.dtype != src.dtype or dst.stride(1) != 1 or src.stride(1) != 1:
        raise ValueError("memcpy2d: need equal shapes/dtypes and unit inner strides")
    esz = dst.element_size()
    rows, cols = dst.shape
    kind = 4  # cudaMemcpyDefault
    err = _cudart.cudaMemcpy2DAsync(dst.data_ptr(), dst.stride(0) * esz, src.data_ptr(), src.stride(0) * esz,
                                    cols * esz, rows, kind, stream.cuda_stream)
    if err != 0:
        raise RuntimeError(f"cudaMemcpy2DAsync failed with error {err}")


def slab_bounds(m, n, nslabs):
    """Row ranges [(lo, hi), ...] of the HostPipeline slabs: contiguous, covering [0, m), heights a multiple
    of the 128-row tile (except the last), at least 1024 rows and at least n rows each."""
    ms = -(-m // max(1, nslabs))
    ms = max(1024, n, -(-ms // 128) * 128)
    bounds = [(lo, min(m, lo + ms)) for lo in range(0, m, ms)]
    if len(bounds) > 1 and bounds[-1][1] - bounds[-1][0] < max(n, 1024):
        lo, _ = bounds[-2]                       # fold a short tail into the previous slab
        bounds[-2:] = [(lo, m)]
    return bounds


class HostPipeline:
    """TSQR (n <= 64) of a matrix that lives in pinned HOST memory, overlapped with its own transfers.

    The rows are cut into `nslabs` slabs. Slab s is copied to the device on a copy stream and reduced
    (tsqr_reduce) as soon as it has arrived; the slab R factors are stacked and factored once (the same
    root step as the row-sharded multi-GPU path, tsqr_gpu_b200/dist.py); then every slab is built
    (tsqr_build with its block of the root Q) and copied back while the next slab is being built.
    The result is the QR factorisation of the whole matrix (R n x n float32).
    Host matrices are column-major, given as (n, m) row-major pinned tensors."""

    def __init__(self, m, n, mode, nslabs=8, device="cuda"):
        self.m, self.n, self.mode = m, n, mode
        md = _mode(mode)
        self.dtype = io_dtype(md)
        if n > 64 or n > m:
            raise InvalidMatrixSize("HostPipeline takes 0 < n <= 64, n <= m")
        self.bounds = slab_bounds(m, n, nslabs)
        self.a = colmajor_empty(m, n, self.dtype, device)
        self.q = colmajor_empty(m, n, self.dtype, device)
        self.ws = [torch.empty(lib.tsqr_b200_tsqr_workspace_bytes(hi - lo, n), dtype=torch.uint8, device=device)
                   for lo, hi in self.bounds]
        self.h2d, self.d2h = torch.cuda.Stream(device), torch.cuda.Stream(device)
        S = len(self.bounds)
        self.ev_in = [torch.cuda.Event() for _ in range(S)]
        self.ev_out = [torch.cuda.Event() for _ in range(S)]
        self.ev_start, self.ev_done = torch.cuda.Event(), torch.cuda.Event()

    def run(self, a_host, q_host, r_host, group=None, world_size=1, rank=0):
        """a_host, q_host: (n, m) pinned; r_host: (n, n) float32 pinned (receives R column-major).
        Asynchronous: everything is ordered after the current stream's work and the current stream
        waits for the last copy; synchronise it before reading the host buffers.
        world_size > 1: this rank's matrix is one row shard of a matrix spread over `world_size` ranks (same m and
        slab count on every rank); the only exchange is one all-gather of the slab R factors (torch.distributed)."""
        from .dist import stack_r_factors
        n, cur = self.n, torch.cuda.current_stream()
        S = len(self.bounds)
        self.ev_start.record(cur)
        self.h2d.wait_event(self.ev_start)
        self.d2h.wait_event(self.ev_start)
        with torch.cuda.stream(self.h2d):
            for s, (lo, hi) in enumerate(self.bounds):
                _memcpy2d_async(self.a[lo:hi].t(), a_host[:, lo:hi], self.h2d)
                self.ev_in[s].record(self.h2d)
        rs = []
        for s, (lo, hi) in enumerate(self.bounds):
            cur.wait_event(self.ev_in[s])
            r_s, _ = tsqr_reduce(self.a[lo:hi], self.mode, self.ws[s], q=self.q[lo:hi])
            rs.append(r_s.t())                                   # R_s^T row-major = R_s column-major storage
        local = torch.stack(rs)                                  # (S, n, n)
        if world_size > 1:
            import torch.distributed as dist
            flat = torch.empty((world_size * S * n, n), dtype=torch.float32, device=local.device)
            dist.all_gather_into_tensor(flat, local.reshape(S * n, n).contiguous(), group=group)
            local = flat.view(world_size * S, n, n)
        stacked = stack_r_factors(local)
        q_root, r = tsqr(stacked, "fp32_notc")                   # redundant on every rank: identical bits
        for s, (lo, hi) in enumerate(self.bounds):
            k = rank * S + s
            tsqr_build(self.a[lo:hi], self.mode, self.ws[s], b=q_root[k * n:(k + 1) * n], q=self.q[lo:hi])
            self.ev_out[s].record(cur)
            self.d2h.wait_event(self.ev_out[s])
            _memcpy2d_async(q_host[:, lo:hi], self.q[lo:hi].t(), self.d2h)
        with torch.cuda.stream(self.d2h):
            r_host.copy_(r.t(), non_blocking=True)
            self.ev_done.record(self.d2h)
        cur.wait_event(self.ev_done)
        return r


class NativeHostPipeline:
    """The library's own host pipeline (tsqr_b200_host_pipeline_*, csrc/host_pipeline.cu): same slabbing and overlap as
    HostPipeline, but device buffers, copy streams, events and the whole call sequence live in C++ behind the C ABI.
    Host matrices are column-major, given as (n, m) row-major pinned tensors; r_host is (n, n) float32 and receives R
    column-major (so r_host.t() is R)."""

    def __init__(self, m, n, mode, nslabs=8):
        self.m, self.n, self.mode = m, n, mode
        self.dtype = io_dtype(mode)
        if n > 64 or n > m or n < 1:
            raise InvalidMatrixSize("NativeHostPipeline takes 0 < n <= 64, n <= m")
        self.handle = lib.tsqr_b200_host_pipeline_create(_mode(mode), m, n, nslabs)
        if not self.handle:
            raise RuntimeError("tsqr_b200_host_pipeline_create failed: " + lib.tsqr_b200_last_error().decode())
        self.nslabs = lib.tsqr_b200_host_pipeline_slabs(self.handle)

    def run(self, a_host, q_host, r_host):
        n, m = self.n, self.m
        for name, t in (("a_host", a_host), ("q_host", q_host)):
            if t.shape != (n, m) or t.dtype != self.dtype or t.stride(1) != 1 or t.is_cuda:
                raise ValueError(f"{name}: need a host ({n}, {m}) {self.dtype} tensor with unit inner stride")
        if r_host.shape != (n, n) or r_host.dtype != torch.float32 or r_host.stride(1) != 1 or r_host.is_cuda:
            raise ValueError("r_host: need a host (n, n) float32 tensor with unit inner stride")
        _raise(lib.tsqr_b200_host_pipeline_run(self.handle, a_host.data_ptr(), a_host.stride(0), q_host.data_ptr(), q_host.stride(0),
                                               r_host.data_ptr(), r_host.stride(0), _stream()))

    def close(self):
        if getattr(self, "handle", None):
            lib.tsqr_b200_host_pipeline_destroy(self.handle)
            self.handle = None

    def __del__(self):
        self.close()


def pc_status(ws, m, n, mode):
    """Which path produced the last TSQR result held in workspace `ws` (a uint8 tensor, or a Buffer):
    (1, est) preconditioned-Cholesky fast path, (0, est) Householder fallback chosen on the device,
    (-1, None) this (mode, m) never takes the fast path. Synchronises the current stream."""
    ws = ws.ws if isinstance(ws, Buffer) else ws
    est = ctypes.c_float(0.0)
    code = lib.tsqr_b200_pc_status(_mode(mode), m, n, ws.data_ptr(), _stream(), ctypes.byref(est))
    if code < -1:
        _raise(code)
    return code, (float(est.value) if code >= 0 else None)


def pc_probe(a, mode, rpre, c, use_tma=True):
    """tests: the two preconditioned-Cholesky passes with caller-supplied factors (see include/tsqr_b200.h).
    Returns (q = (a rpre^-1) c, gsum (2 NP x NP float64, row-major))."""
    md = _mode(mode)
    m, n = a.shape
    NP = 16 if n <= 16 else (32 if n <= 32 else 64)
    lda = _check_colmajor(a, "a")
    q = colmajor_empty(m, n, a.dtype, a.device)
    gsum = torch.zeros((2 * NP, NP), dtype=torch.float64, device=a.device)
    ws = torch.empty(lib.tsqr_b200_tsqr_workspace_bytes(m, n), dtype=torch.uint8, device=a.device)
    assert rpre.shape == (NP, NP) and c.shape == (NP, NP) and rpre.dtype == torch.float32 and c.dtype == torch.float32
    rp = to_colmajor(rpre)
    cc = to_colmajor(c)
    code = lib.tsqr_b200_pc_probe(md, a.data_ptr(), lda, m, n, rp.data_ptr(), cc.data_ptr(), q.data_ptr(), m,
                                  gsum.data_ptr(), 1 if use_tma else 0, ws.data_ptr(), ws.numel(), _stream())
    _raise(code)
    return q, gsum


def orthogonality(q, mode="fp32_notc"):
    """|Q^T Q - I|_F / sqrt(n) in fp64 on the device (reference src/validation.cu:43-80). Returns a 0-d float64 tensor."""
    m, n = q.shape
    ldq = _check_colmajor(q, "q")
    out = torch.zeros((), dtype=torch.float64, device=q.device)
    ws = torch.empty(lib.tsqr_b200_validation_workspace_bytes(n), dtype=torch.uint8, device=q.device)
    _raise(lib.tsqr_b200_orthogonality(_mode(mode), q.data_ptr(), ldq, m, n, out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return out


def residual(a, q, r, mode="fp32_notc"):
    """|A - Q R|_F / |A|_F in fp64 on the device (reference src/test.cu:146-165). Returns a 0-d float64 tensor."""
    m, n = a.shape
    lda, ldq, ldr = _check_colmajor(a, "a"), _check_colmajor(q, "q"), _check_colmajor(r, "r")
    out = torch.zeros((), dtype=torch.float64, device=a.device)
    ws = torch.empty(lib.tsqr_b200_validation_workspace_bytes(n), dtype=torch.uint8, device=a.device)
    _raise(lib.tsqr_b200_residual(_mode(mode), a.data_ptr(), lda, q.data_ptr(), ldq, r.data_ptr(), ldr, m, n, out.data_ptr(),
                                  ws.data_ptr(), ws.numel(), _stream()))
    return out


def latms(m, n, cond, seed=0, device="cuda"):
    """conditioned m x n float32 test matrix generated on the device (reference src/latms.cu:9-121), column-major"""
    a = colmajor_empty(m, n, torch.float32, device)
    ws = torch.empty(lib.tsqr_b200_latms_workspace_bytes(m, n), dtype=torch.uint8, device=device)
    _raise(lib.tsqr_b200_latms(a.data_ptr(), m, m, n, float(cond), int(seed), ws.data_ptr(), ws.numel(), _stream()))
    return a


def launch_count():
    return int(lib.tsqr_b200_launch_count())
