"""world_size-2 CPU test (gloo) of the row-sharded TSQR exchange step in tsqr_gpu_b200/dist.py:
local reduce -> all-gather of the n x n R factors -> redundant root QR -> local build. The three
local steps are played by the numpy oracle (tests may use it as a stand-in; the product path is
the CUDA library), so what is tested is the gather/stack/slice logic that runs on every rank."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


class _OracleOps:
    """api-shaped local steps on CPU tensors, backed by oracle/ts_flat.py"""

    def __init__(self):
        self.V = None

    @staticmethod
    def _cm(x):
        return torch.from_numpy(np.asfortranarray(x).T.copy()).t()  # column-major (m, n) tensor

    def tsqr_reduce(self, a, mode, ws, q=None):
        from oracle import ts_flat
        R, self.V = ts_flat.reduce_chunk(a.numpy().astype(np.float64))
        return self._cm(R.astype(np.float32)), q

    def tsqr(self, stacked, mode):
        from oracle import ts_flat
        Q, R = ts_flat.tsqr(stacked.numpy().astype(np.float64), 1)
        return self._cm(Q.astype(np.float32)), self._cm(R.astype(np.float32))

    def tsqr_build(self, a, mode, ws, b=None, q=None):
        from oracle import ts_flat
        return self._cm(ts_flat.build_chunk(self.V, None if b is None else b.numpy().astype(np.float64)).astype(np.float32))


def _worker(rank, world, port, m, n, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as o
    from tsqr_gpu_b200.dist import tsqr_sharded
    a = o.uniform_matrix(m, n, seed=77)
    mp_ = m // world
    slab = torch.from_numpy(np.asfortranarray(a[rank * mp_:(rank + 1) * mp_]).T.copy()).t()
    q_local, r = tsqr_sharded(slab, "fp32_tc_cor", None, ops=_OracleOps())
    gathered = [torch.empty_like(q_local.contiguous()) for _ in range(world)]
    dist.all_gather(gathered, q_local.contiguous())
    rs = [torch.empty_like(r.contiguous()) for _ in range(world)]
    dist.all_gather(rs, r.contiguous())
    if rank == 0:
        q = torch.cat(gathered, dim=0).numpy().astype(np.float64)
        r64 = r.numpy().astype(np.float64)
        out.put((o.orthogonality(q), o.residual(q, r64, a), all(torch.equal(x, rs[0]) for x in rs)))
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n", [16, 64])
def test_row_sharded_exchange_two_ranks(n):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    world, m = 2, 2048
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, m, n, out)) for r in range(world)]
    for p in procs:
        p.start()
    orth, res, same = out.get(timeout=90)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert orth < 5e-6 and res < 5e-6, (orth, res)
    assert same, "R must be bit-identical on every rank (redundant root factorisation)"


def test_stack_r_factors_layout():
    from tsqr_gpu_b200.dist import stack_r_factors
    P, n = 3, 4
    rs = [torch.arange(n * n, dtype=torch.float32).reshape(n, n) + 100 * p for p in range(P)]  # R_p (row-major view)
    gathered = torch.stack([r.t().contiguous() for r in rs])       # each rank sends R_p^T row-major = R_p column-major
    st = stack_r_factors(gathered)
    assert st.shape == (P * n, n) and st.stride(0) == 1
    assert torch.equal(st, torch.cat(rs, dim=0))
