#!/usr/bin/env python
"""Times the BASELINE.json configurations (C1..C5) on one GPU: this library
and, when oracle/_ref/libref_tcqr.so is present, the reference's own sm_100 build, same device
buffers, CUDA events, median of `reps` calls. Prints one JSON line per configuration."""
import ctypes, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tsqr_gpu_b200 as tq  # noqa: E402


def flops(m, n):
    return 4.0 * m * n * n - 4.0 / 3.0 * n ** 3


def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    ref = None
    so = os.path.join(ROOT, "oracle", "_ref", "libref_tcqr.so")
    if os.path.exists(so) and "--no-ref" not in sys.argv:
        ref = ctypes.CDLL(so)
        vp, sz = ctypes.c_void_p, ctypes.c_size_t
        ref.ref_gpu_qr.argtypes = [ctypes.c_int, ctypes.c_int, vp, sz, vp, sz, vp, sz, sz, sz, vp]
        ref.ref_gpu_qr.restype = ctypes.c_int
    cfgs = [("C1", 16384, 16, "fp32_notc", False), ("C2", 1 << 20, 32, "fp32_tc_cor", False),
            ("C3a", 1 << 24, 64, "fp16_tc_nocor", False), ("C3b", 1 << 24, 64, "fp32_tc_cor", False),
            ("C3c", 1 << 24, 64, "tf32_tc_cor", False), ("C3d", 1 << 24, 64, "fp32_notc", False),
            ("C5", 1 << 22, 512, "tf32_tc_cor", True)]
    only = [a for a in sys.argv[1:] if not a.startswith("--")]
    for name, m, n, mode, reorth in cfgs:
        if only and name not in only:
            continue
        dt = tq.io_dtype(mode)
        g = torch.Generator(device="cuda"); g.manual_seed(1234)
        a0 = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).to(dt).t()
        a = tq.colmajor_empty(m, n, dt); q = tq.colmajor_empty(m, n, dt)
        r = torch.zeros((n, n), dtype=dt, device="cuda").t()
        buf = tq.Buffer().allocate(m, n, reorth)
        def mine():
            if n > 64: a.copy_(a0)
            tq.qr(a if n > 64 else a0, mode, reorth, buf, q, r)
        reps = 5 if m * n >= (1 << 28) else 20
        ms = timed(mine, reps)
        if n > 64:
            ms -= timed(lambda: a.copy_(a0), reps)
        esz = 2 if dt == torch.float16 else 4
        out = {"config": name, "m": m, "n": n, "mode": mode, "reorth": reorth, "ms": ms,
               "gflops": flops(m, n) / ms / 1e6, "hbm_gbs_alg": (3.0 * m * n + n * n) * esz / ms / 1e6}
        if ref is not None:
            md = tq.MODES[mode]
            stream = torch.cuda.current_stream().cuda_stream
            def theirs():
                a.copy_(a0); r.zero_()
                ref.ref_gpu_qr(md, int(reorth), q.data_ptr(), m, r.data_ptr(), n, a.data_ptr(), m, m, n, stream)
            try:
                rms = timed(theirs, max(2, reps // 4)) - timed(lambda: (a.copy_(a0), r.zero_()), reps)
                out["reference_ms"] = rms
                out["speedup"] = rms / ms
            except Exception as ex:  # noqa: BLE001
                out["reference_ms"] = f"failed: {ex}"
        print(json.dumps(out), flush=True)
        del a, q, a0, buf
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
