"""The drop-in boundary seen from C++: tests/cxx/consumer.cu is compiled with nvcc against include/tsqr/*.hpp and
linked against libtsqr_b200.so through the mtk:: symbols only (reference README.md:60-103 sample, buffer& overloads,
allocate_host, double-allocate throw, host-blocking calls). The CPU test compiles and links it (no device needed);
the GPU test runs it and the program checks Q and R itself."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cxx", "consumer.cu")
OUT_DIR = os.path.join(ROOT, "tests", "cxx", "build")
EXE = os.path.join(OUT_DIR, "consumer")
LIB_DIR = os.path.join(ROOT, "tsqr_gpu_b200")


def _build():
    os.makedirs(OUT_DIR, exist_ok=True)
    lib = os.path.join(LIB_DIR, "libtsqr_b200.so")
    assert os.path.exists(lib), "build libtsqr_b200.so first (python -c 'import __graft_entry__ as g; g.build()')"
    if os.path.exists(EXE) and os.path.getmtime(EXE) > max(os.path.getmtime(SRC), os.path.getmtime(lib)):
        return
    cmd = ["/usr/local/cuda/bin/nvcc", "-std=c++14", "-O2", "-gencode", "arch=compute_100a,code=sm_100a",
           "-I", os.path.join(ROOT, "include", "tsqr"), SRC, "-o", EXE, "-L", LIB_DIR, "-ltsqr_b200", "-lcublas",
           "-Xlinker", "-rpath", "-Xlinker", LIB_DIR]
    subprocess.check_call(cmd)


def test_consumer_compiles_and_links():
    _build()
    nm = subprocess.run(["nm", "-u", "-C", EXE], capture_output=True, text=True, check=True).stdout
    assert "mtk::qr::qr<" in nm and "mtk::tsqr::tsqr16<" in nm     # bound through the reference's own symbols
    assert "tsqr_b200_" not in nm                                   # and not through the C ABI


@pytest.mark.gpu
def test_consumer_runs(cuda):
    _build()
    p = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    print(p.stdout[-3000:], p.stderr[-2000:])
    assert p.returncode == 0 and "consumer OK" in p.stdout
