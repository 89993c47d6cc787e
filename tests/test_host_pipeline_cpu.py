"""CPU checks of the host-pipeline slab partition (tsqr_gpu_b200.api.slab_bounds): the slabs tile the row
range exactly, every slab is tall enough to be factored on its own (>= n rows, >= 1024 rows when m allows),
and slab heights are multiples of the 128-row tile except for the last one."""
import pytest

from tsqr_gpu_b200.api import slab_bounds


@pytest.mark.parametrize("m,n,nslabs", [(1 << 24, 64, 8), (70001, 32, 5), (40000, 16, 3), (1024, 64, 8), (1500, 64, 8),
                                        (2047, 16, 2), (100, 64, 4), (1 << 20, 64, 1), (129 * 1024 + 5, 64, 16)])
def test_slab_bounds_partition(m, n, nslabs):
    b = slab_bounds(m, n, nslabs)
    assert b[0][0] == 0 and b[-1][1] == m
    assert all(b[i][1] == b[i + 1][0] for i in range(len(b) - 1))
    assert len(b) <= max(1, nslabs)
    for i, (lo, hi) in enumerate(b):
        assert hi > lo
        if len(b) > 1:
            assert hi - lo >= max(n, 1024)
        if i + 1 < len(b):
            assert (hi - lo) % 128 == 0
