"""Row-sharded multi-GPU TSQR (one process per GPU, torch.distributed for the plumbing).

Rank p owns rows [p*M/P, (p+1)*M/P) of A and of Q. The only exchange step is one all-gather of
the P local N x N R factors (16 KiB per rank at N = 64); every rank then factors the stacked
(P*N) x N matrix redundantly -- same kernel, same reduction order, hence identical bits on all
ranks -- and applies its own N x N block of that root factor to its slab. No bulk data ever
crosses NVLink. The reference is single-GPU; this is the new path
BASELINE.json's north_star item (4) asks for.
"""
import torch
import torch.distributed as dist


def stack_r_factors(gathered):
    """gathered: (P, n, n) tensor holding R_p^T row-major (= R_p column-major). Returns the
    stacked (P*n) x n matrix as a column-major (P*n, n) tensor."""
    P, n, _ = gathered.shape
    # storage [k][p][i] = R_p[i, k]  ->  column k of the stacked matrix is contiguous
    return gathered.permute(1, 0, 2).contiguous().view(n, P * n).t()


def tsqr_sharded(a_local, mode, ws, group=None, q=None, world_size=None, rank=None, ops=None):
    """TSQR of the row-sharded matrix whose local slab is `a_local` (m_p x n, column-major).
    Returns (q_local, r) with r the full n x n factor (float32, identical on all ranks).
    `ops` (tests only) replaces the three local steps -- an object with tsqr_reduce / tsqr /
    tsqr_build of the api module's signatures -- so that the exchange logic can be exercised
    on CPU tensors over gloo; the default is the CUDA library (no CPU fallback)."""
    if ops is None:
        from . import api
    else:
        api = ops
    n = a_local.shape[1]
    P = dist.get_world_size(group) if world_size is None else world_size
    p = dist.get_rank(group) if rank is None else rank
    r_local, q = api.tsqr_reduce(a_local, mode, ws, q=q)             # (n, n) column-major float32
    if P == 1:
        gathered = r_local.t().contiguous().view(1, n, n)
    else:
        flat = torch.empty((P * n, n), dtype=torch.float32, device=a_local.device)   # 2-D: accepted by nccl and gloo
        dist.all_gather_into_tensor(flat, r_local.t().contiguous(), group=group)
        gathered = flat.view(P, n, n)
    stacked = stack_r_factors(gathered)
    q_root, r = api.tsqr(stacked, "fp32_notc")                       # redundant root QR
    b = q_root[p * n:(p + 1) * n]                                    # column-major view, ld = P*n
    q = api.tsqr_build(a_local, mode, ws, b=b, q=q)
    return q, r


class ShardedTSQR:
    """ctypes mirror of the library's own row-sharded driver (tsqr_b200_tsqr_sharded, csrc/sharded.cu): the exchange is
    done by NCCL inside the library on the current CUDA stream. torch.distributed is used once, to hand rank 0's
    ncclUniqueId to the other ranks."""

    def __init__(self, group=None):
        import ctypes
        from ._lib import lib
        self.lib = lib
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        buf = ctypes.create_string_buffer(128)
        if self.rank == 0 and lib.tsqr_b200_nccl_unique_id(buf) != 0:
            raise RuntimeError(lib.tsqr_b200_last_error().decode())
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        t = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev)
        dist.broadcast(t, src=0, group=group)
        self._id = bytes(t.cpu().tolist())
        self.handle = lib.tsqr_b200_sharded_init(self._id, self.world, self.rank)
        if not self.handle:
            raise RuntimeError("tsqr_b200_sharded_init failed: " + lib.tsqr_b200_last_error().decode())
        self._ws = {}

    def tsqr(self, a_local, mode, q=None, r=None):
        from . import api
        md = api._mode(mode)
        m, n = a_local.shape
        lda = api._check_colmajor(a_local, "a")
        if q is None:
            q = api.colmajor_empty(m, n, a_local.dtype, a_local.device)
        if r is None:
            r = torch.zeros((n, n), dtype=a_local.dtype, device=a_local.device).t()
        key = (m, n)
        if key not in self._ws:
            self._ws[key] = torch.empty(self.lib.tsqr_b200_sharded_workspace_bytes(m, n, self.world), dtype=torch.uint8,
                                        device=a_local.device)
        ws = self._ws[key]
        code = self.lib.tsqr_b200_tsqr_sharded(self.handle, md, q.data_ptr(), api._check_colmajor(q, "q"), r.data_ptr(),
                                               api._check_colmajor(r, "r"), a_local.data_ptr(), lda, m, n, ws.data_ptr(),
                                               ws.numel(), torch.cuda.current_stream().cuda_stream)
        api._raise(code)
        return q, r

    def close(self):
        if self.handle:
            self.lib.tsqr_b200_sharded_destroy(self.handle)
            self.handle = None
