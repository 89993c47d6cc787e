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
