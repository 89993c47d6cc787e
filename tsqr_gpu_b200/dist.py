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
