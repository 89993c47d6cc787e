"""Benchmark / accuracy driver printing the reference's CSV rows, so that the reference's
plotting scripts (scripts/*/mk_*graph.py) read results of this library unchanged.

  accuracy       header + rows of reference src/test.cu:51-57, 213-227 (mean and variance over C runs)
  speed          header + rows of reference src/test.cu:60-63, 328-336; `tflops` uses the reference's own operation count
                 (src/test.cu:310-326, with 1024^4 as "tera") so the column compares with numbers it printed; the
                 standard count 2mn^2 - 2n^3/3, the algorithmic HBM GB/s and the roofline fraction are ADDED columns
  accuracy_cond  header + rows of reference src/test_cond.cu:109-112, 129-248 (mean and standard deviation)

Everything is measured on the device: matrices are made there (uniform by torch's generator, conditioned by
tsqr_b200_latms), the metrics are the fp64 device kernels of csrc/validate.cu, times are CUDA events.

  python -m tsqr_gpu_b200.csv_driver speed --mode fp32_tc_cor --m-log2 20 24 --n 64 -C 8
"""
import argparse
import json
import math
import os
import sys

import torch

from . import api

ACCURACY_HEAD = "m,n,rand_range,type,compute_mode,reorthogonalization,residual,residual_variance,orthogonality,orthogonality_variance"
SPEED_HEAD = "m,n,rand_range,type,compute_mode,reorthogonalization,elapsed_time,tflops,working_memory_size"
SPEED_EXTRA = ",std_gflops,hbm_gbs,hbm_roofline_frac,n_gpus"
COND_HEAD = "m,n,cond,type,compute_mode,reorthogonalization,residual,residual_deviation,orthogonality,orthogonality_deviation"


def type_name(mode):
    """reference src/test.cu:20-23"""
    return "half" if api.io_dtype(mode) == torch.float16 else "float"


def reference_complexity(m, n):
    """the operation count the reference divides by (src/test.cu:310-326; batch size of src/tsqr.cu:39-44)"""
    batch = 1 << (max(5, math.ceil(math.log2(m))) - 5)
    qr_c = lambda mm, nn: 2 * nn * (mm * mm * nn + mm * mm * mm)
    tsqr_c = lambda mm, nn: batch * qr_c(mm // batch, nn) + (batch - 1) * qr_c(2 * nn, nn) + (batch - 1) * 4 * nn ** 3 + 4 * nn * nn * mm
    total = 0
    for i in range((n + 15) // 16):
        total += tsqr_c(m, min(16, n - 16 * i)) + 2 * 2 * 16 * 16 * i * m
    return total


def standard_flops(m, n):
    return 2.0 * m * n * n - 2.0 * n ** 3 / 3.0


def hbm_peak_gbs():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        with open(os.path.join(root, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"])
    except (OSError, ValueError, KeyError):
        return 6650.0   # the fallback /opt/skills/guides/B200_PROFILING.md states (same rule as bench.py)


def _uniform(m, n, rand_range, mode, gen):
    a = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-rand_range, rand_range, generator=gen)
    return a.to(api.io_dtype(mode)).t()


def _conditioned(m, n, cond, mode, seed):
    return api.to_colmajor(api.latms(m, n, cond, seed).to(api.io_dtype(mode)))


def _mean_var(xs):
    mean = sum(xs) / len(xs)
    return mean, sum((x - mean) ** 2 for x in xs) / len(xs)


def _metrics(make, mode, reorth, C):
    res, orth = [], []
    for c in range(C):
        a = make(c)
        keep = a.clone() if a.shape[1] > 64 else a   # BlockQR overwrites its input
        q, r = api.qr(a, mode, reorth)
        res.append(float(api.residual(keep, q, r, mode)))
        orth.append(float(api.orthogonality(q, mode)))
    return _mean_var(res), _mean_var(orth)


def accuracy(mode, reorth, configs, C=16, out=sys.stdout, seed=0):
    """configs: (m, n, rand_range) tuples"""
    print(ACCURACY_HEAD, file=out, flush=True)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    for m, n, rr in configs:
        (res, res_v), (orth, orth_v) = _metrics(lambda c: api.to_colmajor(_uniform(m, n, rr, mode, gen)), mode, reorth, C)
        print(f"{m},{n},{rr:g},{type_name(mode)},{mode},{int(reorth)},{res:.6e},{res_v:.6e},{orth:.6e},{orth_v:.6e}", file=out, flush=True)


def accuracy_cond(mode, reorth, configs, C=16, out=sys.stdout, seed=0):
    """configs: (m, n, cond) tuples"""
    print(COND_HEAD, file=out, flush=True)
    for m, n, cond in configs:
        (res, res_v), (orth, orth_v) = _metrics(lambda c: _conditioned(m, n, cond, mode, seed + c), mode, reorth, C)
        print(f"{m},{n},{cond:g},{type_name(mode)},{mode},{int(reorth)},{res:.6e},{math.sqrt(res_v):.6e},{orth:.6e},{math.sqrt(orth_v):.6e}",
              file=out, flush=True)


def speed(mode, reorth, configs, C=16, out=sys.stdout, seed=0, extra=True):
    """configs: (m, n, rand_range) tuples; elapsed_time in seconds per call like the reference"""
    print(SPEED_HEAD + (SPEED_EXTRA if extra else ""), file=out, flush=True)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    peak = hbm_peak_gbs()
    for m, n, rr in configs:
        a = api.to_colmajor(_uniform(m, n, rr, mode, gen))
        src = a.clone() if n > 64 else None
        buf = api.Buffer().allocate(m, n, reorth)
        q = api.colmajor_empty(m, n, a.dtype)
        r = torch.zeros((n, n), dtype=a.dtype, device="cuda").t()
        api.qr(a, mode, reorth, buf, q, r)   # "for cache" (src/test.cu:292-300)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        total_ms = 0.0
        for _ in range(C):
            if src is not None:
                a.copy_(src)   # BlockQR overwrote it; the restore is outside the timed region
            e0.record()
            api.qr(a, mode, reorth, buf, q, r)
            e1.record()
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
        t = total_ms * 1e-3 / C
        row = f"{m},{n},{rr:g},{type_name(mode)},{mode},{int(reorth)},{t:.6e},{reference_complexity(m, n) / t / 1024.0 ** 4:.6e},{buf.get_device_memory_size()}"
        if extra:
            esz = 2 if type_name(mode) == "half" else 4
            gbs = (3.0 * m * n + n * n) * esz / t / 1e9
            row += f",{standard_flops(m, n) / t / 1e9:.6e},{gbs:.6e},{gbs / peak:.4f},1"
        print(row, file=out, flush=True)
        buf.destroy()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("test", choices=["accuracy", "speed", "accuracy_cond"])
    ap.add_argument("--mode", default="fp32_tc_cor", choices=sorted(api.MODES))
    ap.add_argument("--reorth", action="store_true")
    ap.add_argument("--m-log2", type=int, nargs="+", default=[15])
    ap.add_argument("--n", type=int, nargs="+", default=[64])
    ap.add_argument("--param", type=float, nargs="+", default=None, help="rand_range (accuracy, speed) or cond (accuracy_cond) values")
    ap.add_argument("-C", type=int, default=16)
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args(argv)
    params = args.param or ([1.0] if args.test != "accuracy_cond" else [float(1 << c) for c in range(2, 16)])
    configs = [(1 << ml, n, p) for ml in args.m_log2 for n in args.n for p in params if n <= (1 << ml)]
    {"accuracy": accuracy, "speed": speed, "accuracy_cond": accuracy_cond}[args.test](args.mode, args.reorth, configs, args.C, seed=args.seed)


if __name__ == "__main__":
    main()
