#!/usr/bin/env python
"""Full-size parity against the reference (BASELINE configs C3 and, with --c5, C5). Default: conditioned inputs, M = 2^24, N = 64, kappa = 1e6 (geometric singular
values, A = U diag(s) V^T with U, V orthonormal from QR of seeded normals, built on the device in fp64 and
rounded to the mode's io type). For every C3 mode: this library and -- when oracle/_ref/libref_tcqr.so is
present -- the reference's own sm_100 build factor the SAME bytes; orthogonality ||Q^T Q - I||_F / sqrt(n)
(reference src/validation.cu:43-80) and residual ||A - QR||_F / ||A||_F (src/test.cu:146-165) are evaluated
in fp64 on the device, slab by slab. The north-star bound is: ours <= 2 x the reference's value per mode.
Prints one JSON line per mode."""
import ctypes, json, os, sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tsqr_gpu_b200 as tq  # noqa: E402


def conditioned(m, n, cond, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    u = torch.empty((m, n), dtype=torch.float64, device="cuda").normal_(generator=g)
    u, _ = torch.linalg.qr(u)
    v = torch.empty((n, n), dtype=torch.float64, device="cuda").normal_(generator=g)
    v, _ = torch.linalg.qr(v)
    s = torch.logspace(0, -torch.log10(torch.tensor(float(cond))).item(), n, dtype=torch.float64, device="cuda")
    return (u * s) @ v.t()


def metrics(q, r, a):
    """q, a: (m, n) column-major io tensors; r: (n, n). fp64 accumulation over row slabs."""
    m, n = a.shape
    r64 = torch.triu(r.double())
    gram = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    num = torch.zeros((), dtype=torch.float64, device="cuda"); den = torch.zeros((), dtype=torch.float64, device="cuda")
    step = 1 << 20
    for i in range(0, m, step):
        qs = q[i:i + step].double(); as_ = a[i:i + step].double()
        gram += qs.t() @ qs
        num += ((qs @ r64 - as_) ** 2).sum(); den += (as_ ** 2).sum()
    orth = float((gram - torch.eye(n, dtype=torch.float64, device="cuda")).norm() / n ** 0.5)
    return orth, float((num / den).sqrt())


def main():
    pos = [x for x in sys.argv[1:] if not x.startswith('--')]
    m = 1 << int(pos[0]) if pos else 1 << 24
    n, cond = 64, 1e6
    c5 = "--c5" in sys.argv                # BASELINE config C5: BlockQR 2^22 x 512 tf32_tc_cor with reorthogonalisation
    if c5:
        m, n = (1 << int(pos[0]) if pos else 1 << 22), 512
    ref = None
    so = os.path.join(ROOT, "oracle", "_ref", "libref_tcqr.so")
    if os.path.exists(so):
        ref = ctypes.CDLL(so)
        vp, sz = ctypes.c_void_p, ctypes.c_size_t
        ref.ref_gpu_qr.argtypes = [ctypes.c_int, ctypes.c_int, vp, sz, vp, sz, vp, sz, sz, sz, vp]
        ref.ref_gpu_qr.restype = ctypes.c_int
    uniform = "--uniform" in sys.argv or c5      # uniform(-1, 1) inputs (reference src/test.cu:97), all ten compute modes
    if uniform:
        g = torch.Generator(device="cuda"); g.manual_seed(20261105)
        a64 = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t()
        cond = None
    else:
        a64 = conditioned(m, n, cond, seed=20261105)
    for mode in (["tf32_tc_cor"] if c5 else list(tq.MODES) if uniform else ["fp32_tc_cor", "tf32_tc_cor", "fp16_tc_nocor"]):
        dt = tq.io_dtype(mode)
        a0 = tq.to_colmajor(a64.to(dt))
        q = tq.colmajor_empty(m, n, dt); r = torch.zeros((n, n), dtype=dt, device="cuda").t()
        tq.qr(a0.clone() if c5 else a0, mode, c5, None, q, r)     # n > 64 overwrites its input
        torch.cuda.synchronize()
        orth, res = metrics(q, r, a0)
        out = {"config": "C5" if c5 else "C3-uniform" if uniform else "C3-cond", "reorthogonalize": c5, "m": m, "n": n, "cond": cond, "mode": mode, "orthogonality": orth, "residual": res}
        if ref is not None:
            import math
            for reorth, tag in (((1, "reference_reorth"),) if c5 else ((0, "reference"), (1, "reference_reorth"))):   # mtk::qr::qr<mode, false / true>
                a = a0.clone(); r.zero_()
                ref.ref_gpu_qr(tq.MODES[mode], reorth, q.data_ptr(), m, r.data_ptr(), n, a.data_ptr(), m, m, n,
                               torch.cuda.current_stream().cuda_stream)
                torch.cuda.synchronize()
                ro, rr = metrics(q, r, a0)
                ok = None if (math.isnan(ro) or math.isnan(rr)) else bool(orth <= 2 * ro and res <= 2 * rr)
                out.update({tag + "_orthogonality": ro, tag + "_residual": rr, "within_2x_of_" + tag: ok})
                del a
        print(json.dumps(out), flush=True)
        del a0, q
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
