"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/tsqr_b200.h declares, and the C++ template instantiations carry the same mangled
names as the reference's static library. No compute calls."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tsqr_gpu_b200", "libtsqr_b200.so")


def _exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", LIB], text=True)
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def test_library_loads_and_header_symbols_exported():
    from tsqr_gpu_b200 import _lib
    header = open(os.path.join(ROOT, "include", "tsqr_b200.h")).read()
    declared = set(re.findall(r"\b(tsqr_b200_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    exported = _exported()
    assert declared <= exported, f"missing: {declared - exported}"
    assert declared == set(_lib.SIGNATURES), "ctypes table and header disagree"
    assert _lib.lib.tsqr_b200_version() >= 100


def test_io_bytes_and_invalid_modes():
    from tsqr_gpu_b200 import _lib
    assert [_lib.lib.tsqr_b200_io_bytes(m) for m in range(10)] == [2, 2, 4, 4, 4, 4, 4, 4, 4, 4]
    assert _lib.lib.tsqr_b200_io_bytes(10) == 0 and _lib.lib.tsqr_b200_io_bytes(-1) == 0
    # argument validation happens before any CUDA work
    assert _lib.lib.tsqr_b200_qr(2, 0, None, 0, None, 0, None, 0, 4, 8, None, 0, None) == 1   # n > m
    assert _lib.lib.tsqr_b200_qr(2, 0, None, 0, None, 0, None, 0, 0, 0, None, 0, None) == 1
    assert _lib.lib.tsqr_b200_qr(42, 0, None, 0, None, 0, None, 0, 8, 4, None, 0, None) == -1
    assert _lib.lib.tsqr_b200_tsqr(2, None, 0, None, 0, None, 0, 100, 65, None, 0, None) == 1


def test_mangled_cxx_symbols_match_reference_abi():
    """tests/golden/reference_symbols.txt = `nm` of the reference's own objects (tcqr32x16.o,
    tsqr.o, blockqr.o) compiled from /root/reference by oracle/build_ref.sh: 20 qr + 10 tsqr16
    + 22 tcqr instantiations + 7 size functions. The new library must export exactly that set."""
    golden = set(open(os.path.join(ROOT, "tests", "golden", "reference_symbols.txt")).read().split())
    assert len(golden) == 59
    mine = {s for s in _exported() if re.match(r"_ZN3mtk(2qr|4tsqr|4tcqr)", s)}
    assert golden <= mine, f"missing: {sorted(golden - mine)}"
    assert mine == golden, f"extra mtk symbols: {sorted(mine - golden)}"


def test_workspace_sizes_are_host_logic():
    from tsqr_gpu_b200 import _lib
    l = _lib.lib
    small = l.tsqr_b200_tsqr_workspace_bytes(16384, 16)
    big = l.tsqr_b200_tsqr_workspace_bytes(1 << 24, 64)
    assert 0 < small < big
    # checkpoints are the dominant term: N(N+1)/2 floats per 128-row tile at N = 64
    assert big < (1 << 24) * 64 * 4 * 0.30
    assert l.tsqr_b200_qr_workspace_bytes(1 << 22, 512, 1) > l.tsqr_b200_qr_workspace_bytes(1 << 22, 64, 1)
    assert l.tsqr_b200_qr_workspace_bytes(10, 20, 0) == 256  # invalid size -> nothing to allocate
