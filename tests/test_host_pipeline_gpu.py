"""GPU test of the host-buffer pipeline (tsqr_gpu_b200.HostPipeline): row slabs copied in, reduced,
one root factorisation, built and copied out on overlapping streams must give the same factorisation
quality as the one-call path, against host LAPACK (metric definitions of reference src/test.cu:146-165,
src/validation.cu:43-80)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,m,n,nslabs", [("fp32_tc_cor", 1 << 16, 64, 8), ("fp32_tc_cor", 70001, 32, 5),
                                             ("tf32_tc_cor", 40000, 16, 3), ("fp32_notc", 20000, 16, 4),
                                             ("fp16_tc_nocor", 1 << 15, 64, 4)])
@pytest.mark.parametrize("impl", ["python", "native"])
def test_host_pipeline_matches_lapack(cuda, oracle, mode, m, n, nslabs, impl):
    import tsqr_gpu_b200 as tq
    a = oracle.uniform_matrix(m, n, seed=900 + n)
    dt = tq.io_dtype(tq.MODES[mode])
    a_host = torch.from_numpy(np.ascontiguousarray(a.T)).to(dt).pin_memory()      # (n, m) = column-major m x n
    q_host = torch.empty((n, m), dtype=dt).pin_memory()
    r_host = torch.empty((n, n), dtype=torch.float32).pin_memory()
    if impl == "native":    # tsqr_b200_host_pipeline_* (csrc/host_pipeline.cu): the same pipeline behind the C ABI
        pipe = tq.NativeHostPipeline(m, n, mode, nslabs=nslabs)
        pipe.bounds = tq.api.slab_bounds(m, n, nslabs)
        assert pipe.nslabs == len(pipe.bounds)
    else:
        pipe = tq.HostPipeline(m, n, mode, nslabs=nslabs)
    assert pipe.bounds[0][0] == 0 and pipe.bounds[-1][1] == m
    for _ in range(2):                      # second run reuses the buffers and events
        q_host.zero_()
        pipe.run(a_host, q_host, r_host)
        torch.cuda.synchronize()
    a_in = a_host.float().numpy().T.astype(np.float64)
    q = q_host.float().numpy().T.astype(np.float64)
    r = r_host.numpy().T.astype(np.float64)
    assert np.isfinite(q).all() and np.isfinite(r).all()
    _, rl = oracle.lapack_qr(a_in, np.float64)
    orth, res, rd = oracle.orthogonality(q), oracle.residual(q, r, a_in), oracle.r_rel_diff(r, rl)
    print(f"pipeline {mode} {m}x{n}/{len(pipe.bounds)} slabs: orth {orth:.2e} res {res:.2e} R {rd:.2e}")
    if mode == "fp16_tc_nocor":
        assert orth < 5e-3 and res < 5e-3 and rd < 5e-3
    elif mode == "fp32_notc":
        assert orth < 5e-6 and res < 5e-6 and rd < 1e-5
    else:
        assert orth < 5e-6 and res < 5e-6 and rd < 2e-5


def test_native_pipeline_equals_python_pipeline(cuda, oracle):
    """same slabs, same kernels, same order: the two drivers must agree to the bit; strided host buffers (ld > m)"""
    import tsqr_gpu_b200 as tq
    m, n, ld = 50000, 64, 50048
    a = oracle.uniform_matrix(m, n, seed=77)
    a_full = torch.zeros((n, ld), dtype=torch.float32).pin_memory()
    a_full[:, :m] = torch.from_numpy(np.ascontiguousarray(a.T))
    outs = []
    for cls in (tq.HostPipeline, tq.NativeHostPipeline):
        q_full = torch.full((n, ld), 7.0, dtype=torch.float32).pin_memory()
        r_host = torch.empty((n, n), dtype=torch.float32).pin_memory()
        pipe = cls(m, n, "fp32_tc_cor", nslabs=3)
        pipe.run(a_full[:, :m], q_full[:, :m], r_host)
        torch.cuda.synchronize()
        assert bool((q_full[:, m:] == 7.0).all()), "the padding of the host buffer must stay untouched"
        outs.append((q_full[:, :m].clone(), r_host.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_native_pipeline_argument_errors(cuda):
    import tsqr_gpu_b200 as tq
    with pytest.raises(tq.InvalidMatrixSize):
        tq.NativeHostPipeline(1000, 65, "fp32_notc")
    pipe = tq.NativeHostPipeline(4096, 16, "fp32_notc", nslabs=2)
    bad = torch.empty((16, 4095), dtype=torch.float32)
    with pytest.raises(ValueError):
        pipe.run(bad, bad, torch.empty((16, 16), dtype=torch.float32))
    pipe.close()
