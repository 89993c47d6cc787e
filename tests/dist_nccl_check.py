"""Multi-GPU parity check of the row-sharded TSQR (BASELINE config C4 shape, scaled down), run as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 --master-port N tests/dist_nccl_check.py
one rank per GPU over NCCL. Rank 0 gathers Q and checks orthogonality / residual / R-vs-LAPACK with
the CPU oracle; every rank checks that its R is bit-identical to rank 0's. Prints DIST_OK on success."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    from tsqr_gpu_b200.dist import tsqr_sharded, ShardedTSQR
    from oracle import oracle as o
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    sharded = ShardedTSQR()       # the library's own driver: NCCL inside libtsqr_b200.so (csrc/sharded.cu)
    cases = [("py", "fp32_tc_cor", 1 << 16, 64, 5e-6, None), ("py", "fp32_notc", 20000, 32, 5e-6, None),
             ("py", "fp32_tc_cor", 1 << 15, 48, 5e-6, None), ("py", "fp16_tc_nocor", 1 << 14, 64, 5e-3, None),
             # library driver: preconditioned-Cholesky form (broadcast + all-reduce) ...
             ("lib", "fp32_tc_cor", 1 << 16, 64, 5e-6, None), ("lib", "tf32_tc_cor", 20000, 32, 5e-6, None),
             ("lib", "fp16_tc_nocor", 1 << 15, 64, 5e-3, None),
             # ... its general form (all-gather + root factorisation): CUDA-core mode, short slabs, failed certificate
             ("lib", "fp32_notc", 20000, 32, 5e-6, None), ("lib", "fp32_tc_cor", 4096, 64, 5e-6, None),
             ("lib", "fp32_tc_cor", 1 << 15, 64, 5e-6, "zero_column")]
    for impl, mode, mp, n, tol, twist in cases:
        full = o.uniform_matrix(mp * world, n, seed=77)              # same on every rank
        if twist == "zero_column":
            full[:, 5] = 0.0
        dt = tq.io_dtype(mode)
        a_loc = tq.to_colmajor(torch.from_numpy(full[rank * mp:(rank + 1) * mp]).cuda().to(dt))
        if impl == "py":
            ws = torch.empty(lib.tsqr_b200_tsqr_workspace_bytes(mp, n), dtype=torch.uint8, device="cuda")
            q_loc, r = tsqr_sharded(a_loc, mode, ws)
        else:
            q_loc, r = sharded.tsqr(a_loc, mode)
        torch.cuda.synchronize()
        rc = r.float().contiguous()
        r0 = rc.clone()
        dist.broadcast(r0, src=0)
        same = bool(torch.equal(r0, rc))
        qs = [torch.empty((mp, n), dtype=torch.float32, device="cuda") for _ in range(world)]
        dist.all_gather(qs, q_loc.float().contiguous())
        if rank == 0:
            q = torch.cat(qs, 0).cpu().numpy().astype(np.float64)
            a_in = torch.from_numpy(full).to(dt).float().numpy().astype(np.float64)
            rr = r.float().cpu().numpy().astype(np.float64)
            orth, res = o.orthogonality(q), o.residual(q, rr, a_in)
            _, rl = o.lapack_qr(a_in, np.float64)
            rd = o.r_rel_diff(rr, rl)
            good = orth < tol and res < tol and (twist is not None or rd < (1e-5 if tol < 1e-4 else 5e-3))
            print(f"[dist P={world}] {impl} {mode} {mp}x{n}/rank {twist or ''}: orth {orth:.2e} res {res:.2e} R-vs-LAPACK {rd:.2e} {'ok' if good else 'FAIL'}", flush=True)
            ok = ok and good
        flag = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0 and int(flag.item()) != 1:
            print("R differs between ranks", flush=True)
            ok = False
    sharded.close()
    if rank == 0:
        print("DIST_OK" if ok else "DIST_FAIL", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
