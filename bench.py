#!/usr/bin/env python
"""Headline benchmark: TSQR GFLOP/s (+ HBM GB/s vs roofline) at M = 2^24 rows per GPU, N = 64,
fp32_tc_cor -- BASELINE.json's metric -- on 1..8 B200s of one node.

  python bench.py --gpus N --steps K --warmup W            # this library
  python bench.py --impl reference --gpus N ...            # the reference's own GPU build
  (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

One "step" = one QR factorisation (Q and R) of the synthetic matrix through the public API
(mtk::qr::qr<mode,false> semantics via the C ABI). Flops are the standard geqrf+orgqr count
4 M N^2 - 4/3 N^3 for every implementation.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="new", choices=["new", "reference"])
    ap.add_argument("--m-log2", type=int, default=24, help="rows per GPU = 2^this")
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--mode", default="fp32_tc_cor")
    ap.add_argument("--reorth", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-sequential", action="store_true", help="e2e as H2D -> qr -> D2H on one stream (no HostPipeline)")
    return ap.parse_args()


def config_of(args, P):
    """the same keys and wording for both arms (the driver compares them); the reference is single-GPU, so at
    P > 1 its arm still factors one GPU's share"""
    m, n = 1 << args.m_log2, args.n
    esz = 2 if args.mode in ("fp16_notc", "fp16_tc_nocor") else 4
    return {"workload": f"TSQR M=2^{args.m_log2} per GPU x N={n} {args.mode} "
                        f"(mtk::qr::qr<{args.mode},{str(args.reorth).lower()}>), column-major, uniform(-1,1) generated on device",
            "m_per_gpu": m, "m_total": m * P, "n": n, "mode": args.mode,
            "parallelism": (f"row-sharded x{P}: tsqr_b200_tsqr_sharded, NCCL inside the library (broadcast of the {n}x{n} preconditioner + "
                            f"all-reduce of the fp64 Gram sums)") if P > 1 else "single GPU",
            "l2": f"inputs ({m * n * esz / 2**30:.1f} GiB per GPU) are larger than the 126 MB L2; no flush needed"}


def flops(m, n):
    return 4.0 * m * n * n - 4.0 / 3.0 * n ** 3


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_input(m, n, dtype, seed):
    """uniform(-1,1), generated on the device, column-major (m x n view of an (n, m) buffer)"""
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    buf = torch.empty((n, m), dtype=torch.float32, device="cuda")
    buf.uniform_(-1.0, 1.0, generator=g)
    return buf.to(dtype).t()


def cpu_lapack_baseline(n, budget_s=12.0):
    """multithreaded host LAPACK sgeqrf+sorgqr on a bounded sample of the same workload"""
    from oracle import oracle as o
    cores = len(os.sched_getaffinity(0))
    m_s = 1 << 22                     # full size or a 2^22-row slab, labelled
    a = o.uniform_matrix(m_s, n, seed=1)
    o.lapack_qr(a[: 1 << 16])  # warm the library
    t0 = time.perf_counter(); o.lapack_qr(a); dt = time.perf_counter() - t0
    reps, total = 1, dt
    while total < budget_s and reps < 3:
        t0 = time.perf_counter(); o.lapack_qr(a); d = time.perf_counter() - t0
        total += d; reps += 1; dt = min(dt, d)
    return {"value": flops(m_s, n) / dt / 1e9, "unit": "GFLOP/s", "cores": cores, "kind": "port",
            "sample": f"host LAPACK sgeqrf+sorgqr (scipy OpenBLAS) on an M=2^22 x N={n} fp32 slab of the workload, "
                      f"best of {reps}, {dt * 1e3:.0f} ms"}


def run_reference(args, rank, world):
    """reference arm: the UNMODIFIED reference (mtk::qr::qr of enp1s0/tsqr-gpu built for sm_100 in
    oracle/_ref) on the same metric/config. The reference is single-GPU and has no CPU path, so
    rank 0 runs it on its GPU; if the build is missing, the host-LAPACK port is timed instead."""
    if rank != 0:
        return
    # nothing of this repo's product is imported or mapped in this arm: the helpers are local
    modes = {"fp16_notc": 0, "fp16_tc_nocor": 1, "fp32_notc": 2, "fp32_tc_cor": 3, "fp32_tc_nocor": 4,
             "mixed_tc_cor_emu": 5, "tf32_tc_cor": 6, "tf32_tc_cor_emu": 7, "tf32_tc_nocor": 8, "tf32_tc_nocor_emu": 9}
    m, n = 1 << args.m_log2, args.n
    md = modes[args.mode]
    so = os.path.join(ROOT, "oracle", "_ref", "libref_tcqr.so")
    base = {"impl": "reference", "metric": "tsqr_gflops", "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic"}
    cfg = config_of(args, 1)
    if not os.path.exists(so) or not torch.cuda.is_available():
        cb = cpu_lapack_baseline(n)
        cb["sample"] = "reference GPU build unavailable; " + cb["sample"]
        line = dict(base, value=cb["value"], ms_per_step=None, config=cfg, cpu_baseline=cb,
                    e2e={"value": cb["value"], "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        print(json.dumps(line)); return
    lib = ctypes.CDLL(so)
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    lib.ref_gpu_qr.argtypes = [ctypes.c_int, ctypes.c_int, vp, sz, vp, sz, vp, sz, sz, sz, vp]
    lib.ref_gpu_qr.restype = ctypes.c_int
    dt = torch.float16 if md in (0, 1) else torch.float32
    a0 = make_input(m, n, dt, 0x5EED0003)
    a = torch.empty((n, m), dtype=dt, device="cuda").t()
    q = torch.empty((n, m), dtype=dt, device="cuda").t()
    r = torch.zeros((n, n), dtype=dt, device="cuda").t()
    stream = torch.cuda.current_stream().cuda_stream
    times = []
    for it in range(args.warmup + args.steps):
        a.copy_(a0)  # the reference overwrites A for n > 16 (src/blockqr.cu:107-116): restore, untimed
        r.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st = lib.ref_gpu_qr(md, int(args.reorth), q.data_ptr(), m, r.data_ptr(), n, a.data_ptr(), m, m, n, stream)
        e1.record(); torch.cuda.synchronize()
        if st != 0:
            print(json.dumps({"impl": "reference", "unavailable": f"reference qr returned {st}"})); return
        if it >= args.warmup:
            times.append(e0.elapsed_time(e1))
    ms = float(np.mean(times))
    val = flops(m, n) / (ms * 1e-3) / 1e9
    balg = (3.0 * m * n + n * n) * (2 if dt == torch.float16 else 4)
    line = dict(base, value=val, ms_per_step=ms, config=cfg,
                hbm_gbs=balg / (ms * 1e-3) / 1e9,
                cpu_baseline={"value": val, "unit": "GFLOP/s", "cores": 0, "kind": "reference",
                              "sample": "reference's own GPU build (sm_100, oracle/_ref) on rank 0's B200, full workload; "
                                        "the reference has no CPU path"},
                e2e={"value": val, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(line))


def main():
    args = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank)
        run_reference(args, rank, world)
        return
    import torch.distributed as dist
    import tsqr_gpu_b200 as tq
    from tsqr_gpu_b200._lib import lib
    from tsqr_gpu_b200.dist import ShardedTSQR

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    P = world
    m, n = 1 << args.m_log2, args.n      # rows PER GPU (weak scaling: total rows = P * m)
    dt = tq.io_dtype(args.mode)
    esz = 2 if dt == torch.float16 else 4
    a = make_input(m, n, dt, 0x5EED0003 + rank)
    a_pristine = a.clone() if n > 64 else None
    q = tq.colmajor_empty(m, n, dt)
    r = torch.zeros((n, n), dtype=dt, device="cuda").t()
    buf = tq.Buffer().allocate(m, n, True)

    sharded = ShardedTSQR() if P > 1 else None     # the library's own row-sharded driver (NCCL inside libtsqr_b200.so)

    def step():
        if P > 1:
            return sharded.tsqr(a, args.mode, q=q, r=r)
        if a_pristine is not None:
            a.copy_(a_pristine)
        return tq.qr(a, args.mode, args.reorth, buf, q, r)

    def barrier():
        if P > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    lib.tsqr_b200_profile_enable(1)
    launches0 = tq.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = tq.launch_count() - launches0
    lib.tsqr_b200_profile_enable(0)
    clocks = sampler.stop() if rank == 0 else None
    ms_total = e0.elapsed_time(e1)
    if P > 1:
        tt = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms = ms_total / args.steps
    # per-launch durations of the level passes (CUDA events on the launching stream)
    cap = 64 * args.steps + 64
    labels = (ctypes.c_int * cap)(); dur = (ctypes.c_float * cap)()
    cnt = lib.tsqr_b200_profile_read(labels, dur, cap)
    per = {}
    for i in range(cnt):
        per.setdefault(labels[i], []).append(dur[i])
    # (row-sharded runs also log the small root factorisation under the same labels: the slab's own
    # launches are the `steps` longest ones of a label)
    top = lambda v: sorted(v)[-args.steps:] if v else [0.0]
    t_lab = {lb: float(np.mean(top(v))) for lb, v in per.items()}
    t_all = sum(float(np.sum(v)) for v in per.values()) / max(args.steps, 1)

    # sanity of the result that was just timed: ||Q^T Q - I||_F / sqrt(n) (fp32 on device)
    orth_probe = None
    if P == 1:
        qt = q.t().float() if q.t().is_contiguous() else q.t().contiguous().float()
        g = torch.zeros((n, n), device="cuda", dtype=torch.float64)
        step_rows = 1 << 20
        for i in range(0, m, step_rows):
            blk = qt[:, i:i + step_rows].double()
            g += blk @ blk.t()
        orth_probe = float((g - torch.eye(n, device="cuda", dtype=torch.float64)).norm() / n ** 0.5)
        del g, qt

    # ---- BASELINE config C4 (strong scaling): M = 2^26 rows in total, row-sharded over the P GPUs of this run
    c4 = None
    if P > 1 and n <= 64 and not args.reorth:
        m4 = (1 << 26) // P
        del q
        a4 = make_input(m4, n, dt, 0x5EED0004 + rank)
        q4 = tq.colmajor_empty(m4, n, dt)
        r4 = torch.zeros((n, n), dtype=dt, device="cuda").t()
        for _ in range(3):
            sharded.tsqr(a4, args.mode, q=q4, r=r4)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k4 = max(3, min(args.steps, 10))
        g0.record()
        for _ in range(k4):
            sharded.tsqr(a4, args.mode, q=q4, r=r4)
        g1.record()
        barrier()
        tt = torch.tensor([g0.elapsed_time(g1) / k4], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms4 = float(tt.item())
        balg4 = (3.0 * m4 * n + n * n) * esz
        c4 = {"workload": f"C4: TSQR M=2^26 total x N={n} {args.mode}, row-sharded x{P} (strong scaling)", "m_total": 1 << 26,
              "m_per_gpu": m4, "ms_per_step": ms4, "steps": k4, "gflops": flops(1 << 26, n) / (ms4 * 1e-3) / 1e9,
              "hbm_gbs_per_gpu": balg4 / (ms4 * 1e-3) / 1e9}
        del a4, q4
        sharded._ws.pop((m4, n), None)
        torch.cuda.empty_cache()
        q = tq.colmajor_empty(m, n, dt)

    # ---- end-to-end through the public API with HOST buffers (H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        a_host = torch.empty((n, m), dtype=dt, pin_memory=True)
        q_host = torch.empty((n, m), dtype=dt, pin_memory=True)
        r_host = torch.empty((n, n), dtype=torch.float32 if P > 1 else dt, pin_memory=True)
        a_host.copy_(a.t() if a_pristine is None else a_pristine.t())
        torch.cuda.synchronize()

        # single GPU, n <= 64: the public host-buffer call overlaps the slab transfers with the two passes
        # n <= 64: the public host-buffer call overlaps the slab transfers with the two passes (row-sharded at P > 1:
        # every rank pipelines its own shard, one all-gather of the slab R factors)
        # (one GPU: the library's own C entry tsqr_b200_host_pipeline_run; row-sharded: the Python twin with its all-gather)
        pipe = None
        if n <= 64 and not args.reorth and not args.e2e_sequential:
            pipe = tq.NativeHostPipeline(m, n, args.mode, nslabs=8) if P == 1 else tq.HostPipeline(m, n, args.mode, nslabs=8)
        if pipe is not None:
            r_host = torch.empty((n, n), dtype=torch.float32, pin_memory=True)

        def e2e_step():
            if pipe is not None and P == 1:
                pipe.run(a_host, q_host, r_host)
                return
            if pipe is not None:
                pipe.run(a_host, q_host, r_host, world_size=P, rank=rank)
                return
            a.t().copy_(a_host, non_blocking=True)
            if P > 1:
                _, rr = sharded.tsqr(a, args.mode, q=q, r=r)
            else:
                _, rr = tq.qr(a, args.mode, args.reorth, buf, q, r)
            q_host.copy_(q.t(), non_blocking=True)
            r_host.copy_(rr.t(), non_blocking=True)

        e2e_steps = max(1, min(args.steps, 3))
        e2e_step(); barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(e2e_steps):
            e2e_step()
        f1.record()
        barrier()
        ems = f0.elapsed_time(f1)
        # the floor of this number: the same bytes moved with no factorisation in between (every rank at once; all of A must
        # be on the device before any of Q can leave, so the two directions cannot overlap within one factorisation)
        barrier()
        f0.record()
        a.t().copy_(a_host, non_blocking=True)
        q_host.copy_(q.t(), non_blocking=True)
        f1.record()
        barrier()
        copy_ms = f0.elapsed_time(f1)
        if P > 1:
            tt = torch.tensor([copy_ms], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            copy_ms = float(tt.item())
        if P > 1:
            tt = torch.tensor([ems], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ems = float(tt.item())
        ems /= e2e_steps
        e2e = {"value": flops(m * P, n) / (ems * 1e-3) / 1e9, "unit": "GFLOP/s", "ms_per_step": ems,
               "steps": e2e_steps, "h2d_bytes_per_step": m * n * esz * P,
               "d2h_bytes_per_step": (m * n * esz + n * n * 4) * P, "copies_only_ms": copy_ms,
               "api": (("tsqr_b200_host_pipeline_run (C ABI, csrc/host_pipeline.cu)" if P == 1 else "tsqr_gpu_b200.HostPipeline.run")
                       + ": 8 row slabs per GPU, H2D | reduce and build | D2H overlapped on copy streams"
                       if pipe is not None else "H2D, tsqr_gpu_b200.qr / dist.ShardedTSQR.tsqr, D2H on one stream")}

    if rank == 0:
        peak, peak_src = peaks()
        balg_step = (3.0 * m * n + n * n) * esz            # per GPU
        # per-kernel algorithmic bytes: the level-0 reduce pass reads A once
        # (M*N*s), the level-0 build pass reads the slab again and writes Q (2*M*N*s); the reflector
        # store/load of the tensor-core path is implementation traffic, not algorithmic bytes.
        balg_reduce, balg_build = 1.0 * m * n * esz, 2.0 * m * n * esz
        tc = lib.tsqr_b200_mode_uses_tensor_cores(tq.MODES[args.mode], m) != 0 and n <= 64
        notc = args.mode.endswith("_notc")   # the fast path with CUDA-core passes
        # label -> (kernel, algorithmic bytes per launch): first pass reads A once, second pass reads it again
        # and writes Q; anything else a kernel moves is implementation traffic
        kernels = {
            101: (("pc_gram_notc_kernel" if notc else "pc_gram_kernel") + " (pass 1: Z = A R_pre^-1, Gram accumulation: read A once)", balg_reduce, "pc1"),
            103: (("pc_apply_notc_kernel" if notc else "pc_apply_kernel") + " (pass 2: Z again, Q = Z C: read A, write Q)", balg_build, "pc2"),
            0: (("cs_reduce_kernel" if tc else "ts_pass_kernel<REDUCE>") + " level 0 (R-reduce pass: read A once)", balg_reduce, "reduce"),
            1: (("tc_build_kernel" if tc else "ts_pass_kernel<BUILD>") + " level 0 (Q-build pass: read the slab, write Q)", balg_build, "build"),
        }
        dom_label = max(kernels, key=lambda lb: t_lab.get(lb, 0.0))
        dom = {"kernel": kernels[dom_label][0], "bytes": kernels[dom_label][1], "ms": t_lab.get(dom_label, float("nan")),
               "key": kernels[dom_label][2]}
        names = {100: "pc_sample", 101: "pc_pass1", 102: "pc_finish", 103: "pc_pass2", 0: "householder_reduce_level0",
                 1: "householder_build_level0"}
        others = {names.get(lb, f"level{lb // 2}_{'build' if lb % 2 else 'reduce'}"): t for lb, t in sorted(t_lab.items())
                  if lb != dom_label}
        others["all_launches_total"] = t_all
        if not dom["ms"] > 0.0:   # no labelled launch seen: say so instead of printing NaN
            raise RuntimeError("bench.py: the timed region recorded no labelled pass launches (profile scopes missing on this path)")
        achieved = dom["bytes"] / (dom["ms"] * 1e-3) / 1e9
        traffic = None
        try:  # dram__bytes_read.sum + dram__bytes_write.sum of that kernel from the committed ncu capture
            with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
                tr = json.load(f)
            traffic = tr.get(f"{dom['key']}:{args.mode}:m{args.m_log2}:n{n}")
        except Exception:
            traffic = None
        line = {
            "metric": "tsqr_gflops", "value": flops(m * P, n) / (ms * 1e-3) / 1e9, "unit": "GFLOP/s",
            "n_gpus": P, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(args, P),
            "hbm_gbs_per_gpu": balg_step / (ms * 1e-3) / 1e9,
            "hbm_frac_step": balg_step / (ms * 1e-3) / 1e9 / peak,
            "roofline": {"bound": "hbm", "kernel": dom["kernel"],
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": dom["bytes"],
                         "ms_per_launch": dom["ms"], "other_kernels_ms": others},
            "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e,
            "orthogonality_probe": orth_probe,
        }
        if c4 is not None:
            c4["hbm_frac_per_gpu"] = c4["hbm_gbs_per_gpu"] / peak
            line["c4"] = c4
        if P == 1 and not args.no_cpu:
            try:
                line["cpu_baseline"] = cpu_lapack_baseline(n)
            except Exception as ex:  # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": "GFLOP/s", "cores": 0, "kind": "port", "sample": f"failed: {ex}"}
        print(json.dumps(line))
    if P > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
