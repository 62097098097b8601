#!/usr/bin/env python
"""bench.py -- headline benchmark of the RandBLAS sketching hot path on B200.

Workload (BASELINE.json configs[1], "C2"): SASO SparseSkOp<double>, vec_nnz = 8, d = 2048, left-sketching a dense
A of 2^20 x 2048 doubles (lskges): B = S A. One "step" = one sketch_general call over one resident A.
  value : GFLOP/s = 2 * vec_nnz * m * n / t with A and B resident in HBM (CUDA events, max over ranks)
  e2e   : same metric through the host-pointer C-ABI call (rbx_host_sketch_sparseop_f64): pinned host A -> HBM,
          sketch, B -> host, all inside the timed region
  roofline : algorithmic bytes (A read once + B written) / kernel time vs the measured HBM copy bandwidth
  cpu_baseline : the reference's own lskges (oracle/_ref: unmodified headers + shims) on this box's host cores, on a
          bounded row-sample of the same workload
N > 1: column-sharded weak scaling -- every rank regenerates the same S from the shared RNGState and sketches its own
2^20 x 2048 slab of a 2^20 x (2048 N) matrix; no collective on the data path.
`--impl reference` times the reference's CPU path (rank 0 only).
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

D, VEC_NNZ, KEY = 2048, 8, 1997
M_FULL, N_COLS = 1 << 20, 2048
REF_ROWS = 1 << 17          # rows of A per step of the `--impl reference` arm (1/8 of the workload, ~0.5 s of CPU work)


def flops(m, n):
    return 2.0 * VEC_NNZ * m * n


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc = index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            t = [x.strip() for x in line.split(",")]
            if len(t) < 7:
                continue
            try:
                sm.append(float(t[0]))
                mx.append(float(t[1]))
            except ValueError:
                continue
            for nm, v in zip(names, t[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_leg(threads_note=True, budget_rows=None):
    """the reference's lskges (fill_sparse + COO sort + per-column CSC matvec) on host cores, bounded sample"""
    import numpy as np
    import oracle_lib as ol
    if ol.REF is not None:
        kind, lib = "reference", ol.REF
    else:
        kind, lib = "port", None
    m = budget_rows or (1 << 18)
    n = N_COLS
    A = _cpu_sample(np, m, n)
    B = np.zeros(D * n)
    t0 = time.perf_counter()
    rc = ol.sketch_sparseop("f64", "L", "C", "N", "N", D, n, m, 1.0, D, m, VEC_NNZ, "S", [0, 0, 0, 0], [KEY, 0], 0, 0, A, m, 0.0, B, D,
                            which="ref" if kind == "reference" else "orc")
    dt = time.perf_counter() - t0
    assert rc == 0
    cores = int(lib.ref_omp_threads()) if lib is not None else (os.cpu_count() or 1)
    return {"value": flops(m, n) / dt / 1e9, "unit": "GFLOP/s", "cores": cores, "kind": kind, "seconds": dt,
            "sample": f"same operator family and n={n}, d={D}, vec_nnz={VEC_NNZ}; m={m} rows of A ({m / M_FULL:.4f} of the workload), "
                      f"fill_sparse + lskges timed together, OpenMP threads={cores}"}


_SAMPLE = {}


def _cpu_sample(np, m, n):
    """synthetic Gaussian rows of A for the CPU legs (generated once per process, not timed)"""
    if (m, n) not in _SAMPLE:
        _SAMPLE.clear()
        _SAMPLE[(m, n)] = np.random.default_rng(99).standard_normal(m * n)
    return _SAMPLE[(m, n)]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    leg = None
    for i in range(args.warmup + args.steps):
        leg = cpu_reference_leg(budget_rows=REF_ROWS)
        if i >= args.warmup:
            vals.append(leg)
    v = statistics.mean(x["value"] for x in vals)
    ms = statistics.mean(x["seconds"] for x in vals) * 1e3
    leg["value"] = v
    line = {"impl": "reference", "metric": "sketch_gflops", "value": v, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": "C2: SASO SparseSkOp<double> vec_nnz=8 d=2048, left sketch of dense A (lskges); CPU sample m=%d rows x 2048 cols per step" % REF_ROWS,
                                             "d": D, "n": N_COLS, "m_sample": REF_ROWS, "vec_nnz": VEC_NNZ},
            "cpu_baseline": {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": v, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--m", type=int, default=M_FULL, help="rows of A (default: the full C2 size)")
    ap.add_argument("--layout", default="C", choices=["C", "R"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-warmup", type=int, default=3, help="untimed end-to-end calls (the first calls grow the staging pools)")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist
    import randblas_b200 as rb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    m, n, lay = args.m, N_COLS, args.layout
    lda = m if lay == "C" else n
    ldb = D if lay == "C" else n
    # synthetic A: this rank's column slab of a Gaussian m x (n * world) matrix, generated on the device (key 99)
    A = torch.empty(m * n, dtype=torch.float64, device=dev)
    Dist = rb.DenseDist(m, n * world, rb.DenseDistName.Gaussian, rb.MajorAxis.Long)
    rb.fill_dense(rb.Layout(lay), Dist, m, n, 0, rank * n, A, rb.RNGState(99))
    B = torch.zeros(D * n, dtype=torch.float64, device=dev)
    S = rb.SparseSkOp(rb.SparseDist(D, m, VEC_NNZ), KEY)

    def step():
        rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, D, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = rb.lib.rbx_launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for e0, e1 in evs:
        e0.record()
        step()
        e1.record()
    t_all1.record()
    barrier()
    launches = rb.lib.rbx_launch_count() - l0
    total_ms = t_all0.elapsed_time(t_all1)
    per_step = [e0.elapsed_time(e1) for e0, e1 in evs]
    clocks = sampler.stop()
    tm = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    total_ms = float(tm.item())
    ms_per_step = total_ms / args.steps
    value = flops(m, n) * world / (ms_per_step * 1e-3) / 1e9

    # roofline of the dominant kernel (the SASO sketch): per-launch algorithmic bytes / mean per-step device time
    hbm, peak_src = peaks()
    alg_bytes = 8.0 * m * n + 8.0 * D * n
    k_ms = statistics.mean(per_step)
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "saso_traffic.json")
    if os.path.exists(tp) and m == M_FULL:
        tj = json.load(open(tp)).get("ColMajor" if lay == "C" else "RowMajor")
        if tj:
            traffic, traffic_src = tj["dram_bytes"], tj["source"]
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": traffic,
                "traffic_source": traffic_src, "peak_source": peak_src, "kernel_ms": k_ms, "algorithmic_bytes": alg_bytes,
                "kernel": "saso_tile_kernel<double> (+ saso_prep_kernel and saso_reduce_kernel, < 2% of the step, inside the same events)"}

    # end to end through the host-pointer C ABI (rank-local; N>1: every rank stages its own slab concurrently)
    e2e = None
    if args.e2e_steps > 0:
        Ah = torch.empty(m * n, dtype=torch.float64, pin_memory=True)
        Ah.copy_(A)
        Bh = torch.zeros(D * n, dtype=torch.float64, pin_memory=True)
        barrier()
        ts = []
        for i in range(args.e2e_warmup + args.e2e_steps):
            barrier()
            t0 = time.perf_counter()
            rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, D, n, m, 1.0, S, 0, 0, Ah, lda, 0.0, Bh, ldb)
            chk = float(Bh[0])          # the device->host read of the result
            barrier()
            if i >= args.e2e_warmup:
                ts.append(time.perf_counter() - t0)
        te = torch.tensor([max(ts) if False else statistics.mean(ts)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": flops(m, n) * world / float(te.item()) / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": 8 * m * n * world,
               "d2h_bytes_per_step": 8 * D * n * world, "ms_per_step": float(te.item()) * 1e3, "steps": args.e2e_steps, "ms_steps": [round(t * 1e3, 1) for t in ts], "checksum0": chk}
        # device path and host path agree
        assert torch.allclose(Bh, B.cpu(), rtol=1e-12, atol=1e-9)
        del Ah

    extras = None
    cpu = None
    if rank == 0:
        if world == 1:
            cpu = cpu_reference_leg()
        if not args.no_extras and world == 1:
            extras = run_extras(rb, torch, dev, hbm)
    if rank == 0:
        line = {"metric": "sketch_gflops", "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "C2: SASO SparseSkOp<double> vec_nnz=8, d=2048, left sketch (lskges) of dense A 2^20 x 2048 per GPU",
                           "d": D, "n_per_gpu": n, "m": m, "vec_nnz": VEC_NNZ, "layout": "ColMajor" if lay == "C" else "RowMajor",
                           "sharding": "column-sharded, S regenerated per rank, no collective", "l2": "inputs (17.2 GB) exceed L2; no flush needed"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline}
        if cpu is not None:
            line["cpu_baseline"] = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if extras is not None:
            line["extras"] = extras
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_extras(rb, torch, dev, hbm):
    """secondary lines of the same metric family: fill_dense Gsamples/s (C3 shape class) and the dense Gaussian sketch (C1)"""
    out = {}

    def timeit(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    n = 32768                      # 2^30 floats (4.3 GB), the C3 generator at 1/4 size per launch
    buf = torch.empty(n * n, dtype=torch.float32, device=dev)
    for fam in ("G", "U"):
        Dd = rb.DenseDist(n, n, fam, "L")
        ms = timeit(lambda: rb.fill_dense(Dd, buf, rb.RNGState(0)))
        out[f"fill_dense_{fam}_f32"] = {"gsamples_per_s": n * n / ms / 1e6, "ms": ms, "hbm_frac": 4.0 * n * n / (ms * 1e-3) / 1e9 / hbm, "entries": n * n}
    del buf
    d, nn, m = 1024, 1024, 65536   # C1
    A = torch.empty(m * nn, dtype=torch.float32, device=dev)
    rb.fill_dense(rb.DenseDist(m, nn), A, rb.RNGState(99))
    B = torch.zeros(d * nn, dtype=torch.float32, device=dev)
    S = rb.DenseSkOp(rb.DenseDist(d, m), 0)
    ms = timeit(lambda: rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, nn, m, 1.0, S, 0, 0, A, m, 0.0, B, d), reps=3)
    out["sketch_dense_C1_f32"] = {"gflops": 2.0 * d * nn * m / ms / 1e6, "ms": ms}
    return out


if __name__ == "__main__":
    main()
