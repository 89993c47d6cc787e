"""Every thread shape of the CUDA-core reduce pass (csrc/cs_reduce.cu: 4x16, 8x16, 4x32 row segments x
column lanes) must give the same factorisation quality. The shape is read once per process from
TSQR_B200_CS_SHAPE, so each one runs in its own interpreter. Checked against host LAPACK with the metric
definitions of reference src/test.cu:146-165 and src/validation.cu:43-80."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from oracle import oracle as o
from helpers import run_qr
worst = 0.0
for (m, n, mode) in [(40000, 64, "fp32_tc_cor"), (9000, 48, "tf32_tc_cor"), (30000, 32, "fp32_tc_cor"),
                     (20000, 16, "fp32_tc_cor"), (1 << 15, 64, "fp16_tc_nocor")]:
    a = o.uniform_matrix(m, n, seed=77 + n)
    q, r, a_in = run_qr(a, mode, entry="tsqr")
    _, rl = o.lapack_qr(a_in, np.float64)
    orth, res, rd = o.orthogonality(q), o.residual(q, r, a_in), o.r_rel_diff(r, rl)
    tol = 5e-3 if mode == "fp16_tc_nocor" else 5e-6
    print(mode, m, n, orth, res, rd)
    assert np.isfinite(q).all() and orth < tol and res < tol and rd < 4 * tol, (mode, m, n, orth, res, rd)
print("SHAPE_OK")
"""


@pytest.mark.parametrize("shape", ["4x16", "8x16", "4x32"])
def test_reduce_shape(cuda, shape):
    env = dict(os.environ, TSQR_B200_CS_SHAPE=shape)
    out = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT)], capture_output=True, text=True,
                         timeout=600, cwd=ROOT, env=env)
    assert "SHAPE_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-3000:]
