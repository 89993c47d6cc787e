"""tsqr_gpu_b200 -- B200-native tall-skinny QR behind the mtk::qr / mtk::tsqr API of
enp1s0/tsqr-gpu. The product is libtsqr_b200.so (C++/CUDA, C ABI in include/tsqr_b200.h);
this package is the Python mirror of that interface used by the tests and bench.py."""
from .api import (MODES, MODE_NAMES, Buffer, InvalidMatrixSize, colmajor_empty, io_dtype, launch_count, qr, tsqr,
                  tsqr_build, tsqr_reduce, qr32x16_batched, to_colmajor, HostPipeline, NativeHostPipeline, pc_status, pc_probe, orthogonality, residual, latms)

__all__ = ["MODES", "MODE_NAMES", "Buffer", "InvalidMatrixSize", "colmajor_empty", "io_dtype", "launch_count", "qr",
           "tsqr", "tsqr_build", "tsqr_reduce", "qr32x16_batched", "to_colmajor", "HostPipeline", "NativeHostPipeline", "pc_status", "pc_probe", "orthogonality", "residual", "latms"]
