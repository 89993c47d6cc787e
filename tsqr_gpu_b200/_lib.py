"""ctypes binding of the C ABI in include/tsqr_b200.h (libtsqr_b200.so, built in-tree).

There is no fallback: if the CUDA library is missing the import fails loudly.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TSQR_B200_LIB: load another build of the same library (same-box A/B measurements of kernel variants)
LIB_PATH = os.environ.get("TSQR_B200_LIB") or os.path.join(_HERE, "libtsqr_b200.so")

_c = ctypes
_sz = ctypes.c_size_t
_vp = ctypes.c_void_p

# name -> (restype, argtypes); must list every symbol include/tsqr_b200.h declares
SIGNATURES = {
    "tsqr_b200_version": (_c.c_int, []),
    "tsqr_b200_last_error": (_c.c_char_p, []),
    "tsqr_b200_io_bytes": (_sz, [_c.c_int]),
    "tsqr_b200_qr_workspace_bytes": (_sz, [_sz, _sz, _c.c_int]),
    "tsqr_b200_tsqr_workspace_bytes": (_sz, [_sz, _sz]),
    "tsqr_b200_qr": (_c.c_int, [_c.c_int, _c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _sz, _vp]),
    "tsqr_b200_tsqr": (_c.c_int, [_c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _sz, _vp]),
    "tsqr_b200_tsqr_reduce": (_c.c_int, [_c.c_int, _vp, _sz, _sz, _sz, _vp, _vp, _sz, _vp, _sz, _vp]),
    "tsqr_b200_tsqr_build": (_c.c_int, [_c.c_int, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _sz, _vp, _sz, _vp]),
    "tsqr_b200_qr32x16_batched": (_c.c_int, [_c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _c.c_uint, _c.c_uint, _sz, _vp, _vp]),
    "tsqr_b200_launch_count": (_c.c_ulonglong, []),
    "tsqr_b200_profile_enable": (None, [_c.c_int]),
    "tsqr_b200_profile_read": (_c.c_int, [_vp, _vp, _c.c_int]),
    "tsqr_b200_mode_uses_tensor_cores": (_c.c_int, [_c.c_int, _sz]),
    "tsqr_b200_pc_status": (_c.c_int, [_c.c_int, _sz, _sz, _vp, _vp, _vp]),
    "tsqr_b200_pc_probe": (_c.c_int, [_c.c_int, _vp, _sz, _sz, _sz, _vp, _vp, _vp, _sz, _vp, _c.c_int, _vp, _sz, _vp]),
    "tsqr_b200_tsqr_static": (_c.c_int, [_c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _sz, _sz, _vp]),
    "tsqr_b200_validation_workspace_bytes": (_sz, [_sz]),
    "tsqr_b200_orthogonality": (_c.c_int, [_c.c_int, _vp, _sz, _sz, _sz, _vp, _vp, _sz, _vp]),
    "tsqr_b200_residual": (_c.c_int, [_c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _vp, _sz, _vp]),
    "tsqr_b200_latms_workspace_bytes": (_sz, [_sz, _sz]),
    "tsqr_b200_latms": (_c.c_int, [_vp, _sz, _sz, _sz, _c.c_float, _c.c_ulonglong, _vp, _sz, _vp]),
    "tsqr_b200_nccl_unique_id": (_c.c_int, [_vp]),
    "tsqr_b200_host_pipeline_create": (_vp, [_c.c_int, _sz, _sz, _c.c_int]),
    "tsqr_b200_host_pipeline_slabs": (_c.c_int, [_vp]),
    "tsqr_b200_host_pipeline_run": (_c.c_int, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp]),
    "tsqr_b200_host_pipeline_destroy": (None, [_vp]),
    "tsqr_b200_sharded_init": (_vp, [_vp, _c.c_int, _c.c_int]),
    "tsqr_b200_sharded_attach": (_vp, [_vp, _c.c_int, _c.c_int]),
    "tsqr_b200_sharded_destroy": (None, [_vp]),
    "tsqr_b200_sharded_workspace_bytes": (_sz, [_sz, _sz, _c.c_int]),
    "tsqr_b200_tsqr_sharded": (_c.c_int, [_vp, _c.c_int, _vp, _sz, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _sz, _vp]),
    "tsqr_b200_selftest_mma": (_c.c_int, [_c.c_int, _c.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C tsqr_gpu_b200/csrc` (there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()
