"""CSV bench driver: the headers are the reference's byte for byte (src/test.cu:51-63,
src/test_cond.cu:109-112), the rows parse with the same column count, the numbers are sane."""
import io
import math

import pytest

REF_ACCURACY_HEAD = "m,n,rand_range,type,compute_mode,reorthogonalization,residual,residual_variance,orthogonality,orthogonality_variance"
REF_SPEED_HEAD = "m,n,rand_range,type,compute_mode,reorthogonalization,elapsed_time,tflops,working_memory_size"
REF_COND_HEAD = "m,n,cond,type,compute_mode,reorthogonalization,residual,residual_deviation,orthogonality,orthogonality_deviation"


def test_headers_and_flop_model():
    from tsqr_gpu_b200 import csv_driver as d
    assert d.ACCURACY_HEAD == REF_ACCURACY_HEAD and d.SPEED_HEAD == REF_SPEED_HEAD and d.COND_HEAD == REF_COND_HEAD
    # reference src/test.cu:310-326 by hand for m = 1024, n = 16: batch 32, 32 leaf QRs of 32 x 16, 31 merges
    qr_c = lambda m, n: 2 * n * (m * m * n + m ** 3)
    want = 32 * qr_c(32, 16) + 31 * qr_c(32, 16) + 31 * 4 * 16 ** 3 + 4 * 16 * 16 * 1024
    assert d.reference_complexity(1024, 16) == want
    assert d.reference_complexity(1024, 32) == 2 * want + 2 * 2 * 16 * 16 * 1024
    assert d.type_name("fp16_tc_nocor") == "half" and d.type_name("tf32_tc_cor_emu") == "float"


def _rows(text, head):
    lines = text.strip().splitlines()
    assert lines[0].startswith(head)
    ncol = len(lines[0].split(","))
    rows = [ln.split(",") for ln in lines[1:]]
    assert rows and all(len(r) == ncol for r in rows)
    return rows


@pytest.mark.gpu
def test_rows_on_device(cuda):
    from tsqr_gpu_b200 import csv_driver as d
    out = io.StringIO()
    d.accuracy("fp32_tc_cor", False, [(32768, 64, 1.0), (4096, 128, 4.0)], C=3, out=out)
    for r in _rows(out.getvalue(), REF_ACCURACY_HEAD):
        assert r[3] == "float" and r[4] == "fp32_tc_cor" and r[5] == "0"
        assert 0 < float(r[6]) < 5e-6 and 0 < float(r[8]) < 5e-6 and float(r[7]) >= 0 and float(r[9]) >= 0
    out = io.StringIO()
    d.accuracy_cond("fp16_tc_nocor", True, [(16384, 64, 1024.0), (8192, 128, 16.0)], C=2, out=out)
    for r in _rows(out.getvalue(), REF_COND_HEAD):
        assert r[3] == "half" and r[5] == "1" and 0 < float(r[6]) < 5e-3
    out = io.StringIO()
    d.speed("fp32_tc_cor", False, [(1 << 20, 64, 1.0), (1 << 15, 128, 1.0)], C=3, out=out)
    rows = _rows(out.getvalue(), REF_SPEED_HEAD)
    for r in rows:
        t = float(r[6])
        assert 1e-6 < t < 1.0 and math.isclose(float(r[7]), d.reference_complexity(int(r[0]), int(r[1])) / t / 1024.0 ** 4, rel_tol=1e-4)
        assert int(r[8]) > 0 and float(r[9]) > 0 and 0 < float(r[11]) < 1.0
