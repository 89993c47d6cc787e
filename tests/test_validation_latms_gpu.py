"""GPU tests of the SURVEY section 8(f) rows built in round 2, through the C ABI:
  f2  tsqr_b200_orthogonality / tsqr_b200_residual  (fp64 on the device; reference src/validation.cu:43-80, src/test.cu:146-165)
      against the oracle's numpy definitions of the same metrics;
  f1  tsqr_b200_latms (reference src/latms.cu:9-121, singular values of src/test_cond.cu:31-50): prescribed spectrum,
      determinism in the seed;
and of the small-QR entries that had no test (rows a5/a6): half io and the single-tile form of qr32x16."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,m,n", [("fp32_notc", 5000, 16), ("fp32_tc_cor", 70001, 64), ("fp32_notc", 3000, 200),
                                      ("fp16_tc_nocor", 20000, 64)])
def test_device_metrics_match_the_oracle_definitions(cuda, oracle, mode, m, n):
    import tsqr_gpu_b200 as tq
    dt = tq.io_dtype(mode)
    a = oracle.uniform_matrix(m, n, seed=n)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda().to(dt), ld=m + 3)
    q, r = tq.qr(a_dev.clone() if n > 64 else a_dev, mode, n > 64)
    # perturb Q a little so that the metric is not at the rounding floor of the factorisation
    q[7, 3] += 1e-3
    orth = float(tq.orthogonality(q, mode))
    res = float(tq.residual(a_dev, q, r, mode))
    q64 = q.float().cpu().numpy().astype(np.float64)
    r64 = r.float().cpu().numpy().astype(np.float64)
    a64 = a_dev.float().cpu().numpy().astype(np.float64)
    o_ref, r_ref = oracle.orthogonality(q64), oracle.residual(q64, r64, a64)
    print(f"{mode} {m}x{n}: orthogonality {orth:.6e} (oracle {o_ref:.6e}) residual {res:.6e} (oracle {r_ref:.6e})")
    assert abs(orth - o_ref) <= 1e-9 * max(1.0, o_ref) + 1e-12 * n
    assert abs(res - r_ref) <= 1e-9 * max(1.0, r_ref) + 1e-12


def test_device_metrics_at_full_height(cuda):
    import tsqr_gpu_b200 as tq
    from helpers import device_metrics, uniform_device
    a = tq.to_colmajor(uniform_device(1 << 22, 64, 99))
    q, r = tq.qr(a, "fp32_tc_cor")
    orth, res = float(tq.orthogonality(q, "fp32_tc_cor")), float(tq.residual(a, q, r, "fp32_tc_cor"))
    o2, r2 = device_metrics(q, r, a)
    print(f"2^22 x 64: {orth:.4e} / {res:.4e} (torch fp64: {o2:.4e} / {r2:.4e})")
    assert abs(orth - o2) < 1e-10 and abs(res - r2) < 1e-10


@pytest.mark.parametrize("m,n,cond", [(1 << 17, 64, 1e6), (20000, 32, 1e3), (4096, 7, 10.0), (1 << 15, 128, 1e4), (6000, 200, 1e2)])
def test_latms_spectrum_and_determinism(cuda, m, n, cond):
    import tsqr_gpu_b200 as tq
    a = tq.latms(m, n, cond, seed=11)
    b = tq.latms(m, n, cond, seed=11)
    c = tq.latms(m, n, cond, seed=12)
    torch.cuda.synchronize()
    assert torch.equal(a, b), "same seed, same matrix"
    assert not torch.equal(a, c)
    s = torch.linalg.svdvals(a.double()).cpu().numpy()
    want = cond ** (-np.arange(n) / max(n - 1, 1))
    rel = np.abs(s - want) / want
    print(f"latms {m}x{n} cond {cond:g}: s[0] {s[0]:.6f} s[-1] {s[-1]:.3e} max rel dev {rel.max():.2e}")
    assert rel.max() < (2e-2 if cond >= 1e6 else 1e-3)      # the smallest of 64 values at cond 1e6 sits at fp32 resolution
    assert abs(s[0] / s[-1] / cond - 1.0) < 2e-2


def test_latms_input_through_the_fast_path(cuda):
    import tsqr_gpu_b200 as tq
    m, n = 1 << 18, 64
    a = tq.latms(m, n, 1e6, seed=3)
    buf = tq.Buffer().allocate(m, n, True)
    q, r = tq.tsqr(a, "fp32_tc_cor", buffer=buf)
    st, est = tq.pc_status(buf, m, n, "fp32_tc_cor")
    orth, res = float(tq.orthogonality(q, "fp32_tc_cor")), float(tq.residual(a, q, r, "fp32_tc_cor"))
    print(f"latms kappa=1e6 -> status {st} est {est:.3f} orthogonality {orth:.2e} residual {res:.2e}")
    assert st == 1 and orth < 2e-6 and res < 2e-6


@pytest.mark.parametrize("mode", ["fp16_notc", "fp16_tc_nocor"])
def test_qr32x16_batched_half_io(cuda, oracle, mode):
    import tsqr_gpu_b200 as tq
    m, n, batch = 1000, 16, 40
    a = oracle.uniform_matrix(m, n, seed=41)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda().half())
    a_in = a_dev.float().cpu().numpy().astype(np.float64)
    starts = np.array([m * i // batch for i in range(batch + 1)], dtype=np.int32)
    q, r = tq.qr32x16_batched(a_dev, torch.from_numpy(starts).cuda(), n, mode)
    torch.cuda.synchronize()
    q, r = q.float().cpu().numpy().astype(np.float64), r.float().cpu().numpy().astype(np.float64)
    for i in range(batch):
        s, e = starts[i], starts[i + 1]
        assert oracle.orthogonality(q[s:e]) < 5e-3 and oracle.residual(q[s:e], r[n * i:n * i + n], a_in[s:e]) < 5e-3


@pytest.mark.parametrize("rows,n", [(32, 16), (20, 9), (17, 16)])
def test_qr32x16_single_tile(cuda, oracle, rows, n):
    """the single-tile form (mtk::tcqr::qr32x16, reference src/tcqr32x16.cu:1534-1581): batch 1, no start positions"""
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    a = oracle.uniform_matrix(rows, n, seed=rows)
    a_dev = tq.to_colmajor(torch.from_numpy(a).cuda())
    q = tq.colmajor_empty(rows, n, torch.float32)
    r = tq.colmajor_empty(n, n, torch.float32)
    code = lib.tsqr_b200_qr32x16_batched(tq.MODES["fp32_notc"], q.data_ptr(), rows, r.data_ptr(), n, a_dev.data_ptr(), rows,
                                         rows, n, 1, None, torch.cuda.current_stream().cuda_stream)
    assert code == 0
    torch.cuda.synchronize()
    q64, r64 = q.cpu().numpy().astype(np.float64), r.cpu().numpy().astype(np.float64)
    assert oracle.orthogonality(q64) < 2e-6 and oracle.residual(q64, r64, a) < 2e-6
    qo, ro = oracle.ref_qr32x16(a)                       # the C restatement of the reference's leaf QR
    assert oracle.r_rel_diff(r64, ro) < 1e-5
