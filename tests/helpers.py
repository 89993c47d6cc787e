"""shared helpers for the GPU parity tests"""
import numpy as np
import torch

import tsqr_gpu_b200 as tq

# tolerance classes per compute mode: (orthogonality, residual) absolute bounds used when the
# reference's own number is not available; the fp32-exact classes are ~20x fp32 eps * sqrt-ish
# growth, the fp16 io classes are bounded by half rounding of Q and R (2^-11).
TOL = {
    "fp32_notc": (5e-6, 5e-6), "fp32_tc_cor": (5e-6, 5e-6), "tf32_tc_cor": (5e-6, 5e-6),
    "mixed_tc_cor_emu": (5e-6, 5e-6), "tf32_tc_cor_emu": (5e-6, 5e-6),
    "fp32_tc_nocor": (5e-3, 5e-3), "tf32_tc_nocor": (5e-3, 5e-3), "tf32_tc_nocor_emu": (5e-3, 5e-3),
    "fp16_notc": (5e-3, 5e-3), "fp16_tc_nocor": (5e-3, 5e-3),
}


def run_qr(a_np, mode, reorth=False, lda=None, entry="qr"):
    """a_np: (m, n) float32 numpy. Returns q, r as float64 numpy (converted from io dtype) and
    the io-rounded input that was actually factored."""
    dt = tq.io_dtype(mode)
    a_t = torch.from_numpy(np.ascontiguousarray(a_np)).to("cuda").to(dt)
    a_dev = tq.to_colmajor(a_t, ld=lda)
    a_in = a_dev.clone().float().cpu().numpy().astype(np.float64)
    if entry == "qr":
        q, r = tq.qr(a_dev, mode, reorth)
    else:
        q, r = tq.tsqr(a_dev, mode)
    torch.cuda.synchronize()
    return q.float().cpu().numpy().astype(np.float64), r.float().cpu().numpy().astype(np.float64), a_in


def device_metrics(q, r, a):
    """orthogonality ||Q^T Q - I||_F / sqrt(n) (reference src/validation.cu:43-80) and residual
    ||A - Q R||_F / ||A||_F (reference src/test.cu:146-165) of device matrices of any height, accumulated in
    fp64 on the device slab by slab. q, a: (m, n) column-major tensors of the io type; r: (n, n)."""
    m, n = a.shape
    r64 = torch.triu(r.double())
    gram = torch.zeros((n, n), dtype=torch.float64, device=a.device)
    num = torch.zeros((), dtype=torch.float64, device=a.device)
    den = torch.zeros((), dtype=torch.float64, device=a.device)
    step = 1 << 20 if n <= 64 else 1 << 18
    for i in range(0, m, step):
        qs, as_ = q[i:i + step].double(), a[i:i + step].double()
        gram += qs.t() @ qs
        num += ((qs @ r64 - as_) ** 2).sum()
        den += (as_ ** 2).sum()
    orth = float((gram - torch.eye(n, dtype=torch.float64, device=a.device)).norm() / n ** 0.5)
    return orth, float((num / den).sqrt())


def conditioned_device(m, n, cond, seed):
    """A = U diag(s) V^T on the device in fp64 (geometric singular values 1 .. 1/cond, slatms MODE = 3; the
    construction of reference src/latms.cu:8-121 with the singular values of src/test_cond.cu:31-50)."""
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    u = torch.empty((m, n), dtype=torch.float64, device="cuda").normal_(generator=g)
    u, _ = torch.linalg.qr(u)
    v = torch.empty((n, n), dtype=torch.float64, device="cuda").normal_(generator=g)
    v, _ = torch.linalg.qr(v)
    s = torch.logspace(0, -float(np.log10(cond)), n, dtype=torch.float64, device="cuda")
    return (u * s) @ v.t()


def uniform_device(m, n, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t()
