"""Known-answer test of the tcgen05/TMEM operand layouts (tsqr_gpu_b200/csrc/tc05.cuh): the
GPU products are compared with float64 numpy on the same inputs. 1-term products see tf32
operand rounding (2^-11 relative per operand), 3-term (hi/lo corrected) products must be
fp32-accurate."""
import numpy as np
import pytest
import torch

from tsqr_gpu_b200._lib import lib

pytestmark = pytest.mark.gpu


def _cm(x):  # column-major device copy, returned flat
    return torch.from_numpy(np.asfortranarray(x).ravel(order="F").copy()).cuda()


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("corrected", [0, 1])
def test_mma_selftest(cuda, corrected, variant):
    rng = np.random.default_rng(7)
    A = rng.uniform(-1, 1, (128, 64)).astype(np.float32)
    B = rng.uniform(-1, 1, (128, 64)).astype(np.float32)
    X = rng.uniform(-1, 1, (64, 64)).astype(np.float32)
    dA, dB, dX = _cm(A), _cm(B), _cm(X)
    outs = [torch.zeros(128 * 64, dtype=torch.float32, device="cuda") for _ in range(4)]
    rc = lib.tsqr_b200_selftest_mma(corrected, variant, dA.data_ptr(), dB.data_ptr(), dX.data_ptr(),
                                    *[o.data_ptr() for o in outs], None)
    assert rc == 0
    torch.cuda.synchronize()
    D1, D2, D3, D4 = [o.cpu().numpy().reshape(64, 128).T.astype(np.float64) for o in outs]
    A64, B64, X64 = A.astype(np.float64), B.astype(np.float64), X.astype(np.float64)
    E1 = A64.T @ B64
    E2 = -A64 @ X64
    E3 = A64[:, 16:64].T @ B64[:, 48:64]
    E4 = B64[:, 0:48] + A64[:, 16:32] @ X64[16:32, 16:64]
    tol = 2e-5 if corrected else 2e-2
    errs = {
        "gram": np.abs(D1[:64, :] - E1).max(),
        "update_neg": np.abs(D2 - E2).max(),
        "gram_offset": np.abs(D3[:48, :16] - E3).max(),
        "update_acc": np.abs(D4[:, :48] - E4).max(),
    }
    keep = ("gram", "gram_offset") if variant == 0 else ("update_neg", "update_acc")
    errs = {k: errs[k] for k in keep}
    print(variant, corrected, errs)
    for k, v in errs.items():
        assert v < tol, (k, v, errs)
