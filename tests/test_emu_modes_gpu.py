"""The three _emu compute modes (reference src/matmul.hpp:26-209): the tensor-core products emulated
on the CUDA cores -- operands cut to tf32 (half for mixed_tc_cor_emu), one or three fp32 sums per product.
  * tf32_tc_nocor_emu must show tf32-level errors: NOT the bits of fp32_notc, in the error class of the real tf32_tc_nocor (and within 2x of the
    reference's own tf32_tc_nocor_emu: tests/test_reference_gpu.py);
  * tf32_tc_cor_emu and mixed_tc_cor_emu must be back at fp32 level (the correction terms restore the precision).
The comparison with the reference's own build of the same modes is tests/test_reference_gpu.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(oracle, a, mode, reorth=False):
    from helpers import run_qr
    q, r, a_in = run_qr(a, mode, reorth)
    assert np.isfinite(q).all() and np.isfinite(r).all()
    return q, r, oracle.orthogonality(q), oracle.residual(q, r, a_in)


@pytest.mark.parametrize("m,n", [(8192, 64), (20000, 32), (3000, 16)])
def test_emu_modes_have_their_own_numerics(cuda, oracle, m, n):
    a = oracle.uniform_matrix(m, n, seed=n + 3)
    q0, r0, o0, e0 = _run(oracle, a, "fp32_notc")
    q1, r1, o1, e1 = _run(oracle, a, "tf32_tc_nocor_emu")
    q2, r2, o2, e2 = _run(oracle, a, "tf32_tc_cor_emu")
    q3, r3, o3, e3 = _run(oracle, a, "mixed_tc_cor_emu")
    _, _, ot, et = _run(oracle, a, "tf32_tc_nocor")
    print(f"{m}x{n}: fp32_notc {o0:.2e}/{e0:.2e} | nocor_emu {o1:.2e}/{e1:.2e} (real tf32_tc_nocor {ot:.2e}/{et:.2e}) | "
          f"cor_emu {o2:.2e}/{e2:.2e} | mixed_cor_emu {o3:.2e}/{e3:.2e}")
    assert not np.array_equal(q1, q0) and not np.array_equal(r1, r0), "tf32_tc_nocor_emu must not be an alias of fp32_notc"
    assert not np.array_equal(q2, q0) and not np.array_equal(q3, q0)
    assert 1e-5 < max(o1, e1) < 5e-3, "tf32-level error expected"
    # every product of the sweep is cut here, the tcgen05 path keeps its panel sweep in fp32: same class, not same value
    assert o1 <= 10 * max(ot, 2e-4) and e1 <= 10 * max(et, 2e-4)
    assert o2 < 5e-6 and e2 < 5e-6 and o3 < 5e-6 and e3 < 5e-6, "the corrected emulations are fp32-class"


def test_emu_modes_blockqr(cuda, oracle):
    a = oracle.uniform_matrix(4096, 160, seed=8)
    for mode, tol in (("tf32_tc_nocor_emu", 5e-3), ("tf32_tc_cor_emu", 5e-6), ("mixed_tc_cor_emu", 5e-6)):
        _, _, o, e = _run(oracle, a, mode, True)
        print(f"blockqr {mode}: {o:.2e} / {e:.2e}")
        assert o < tol and e < tol
