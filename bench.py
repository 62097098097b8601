#!/usr/bin/env python
"""bench.py -- headline benchmark of the RandBLAS sketching hot path on B200.

Headline workload (BASELINE.json configs[1], "C2"): SASO SparseSkOp<double>, vec_nnz = 8, d = 2048, left-sketching a dense
A of 2^20 x 2048 doubles (lskges): B = S A. One "step" = one sketch_general call over one resident A.
  value : GFLOP/s = 2 * vec_nnz * m * n / t with A and B resident in HBM (CUDA events, max over ranks)
  e2e   : same metric through the host-pointer C-ABI call (rbx_host_sketch_sparseop_f64): pinned host A -> HBM,
          sketch, B -> host, all inside the timed region
  roofline : algorithmic bytes (A read once + B written) / kernel time vs the measured HBM copy bandwidth
  cpu_baseline : the reference's own fill_sparse + lskges (oracle/_ref: unmodified headers + shims) on this box's host
          cores, on the full workload
  extras : the other four BASELINE configs (C1, C3, C4, C5-slab), each with value / roofline / cpu_baseline / e2e, and the
          tensor / FP64 denominators measured in this run (cuBLAS TF32 and FP64 GEMM 8192^3)
N > 1: column-sharded weak scaling -- every rank regenerates the same S from the shared RNGState and sketches its own
2^20 x 2048 slab of a 2^20 x (2048 N) matrix; no collective on the data path. The same line carries
`extras.C5_row_sharded`: the row-sharded dense FP64 sketch with its NCCL reduce-scatter inside the timed region (strong
scaling over a fixed slab of C5), the one communicating path the north star names.
`--impl reference` times the reference's CPU path on the same config (rank 0 only; all host cores whatever
OMP_NUM_THREADS the launcher exported).
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


def flops(m, n):
    return 2.0 * VEC_NNZ * m * n


def c2_config(m, lay, world):
    """the `config` object of the JSON line: identical in the GPU arm and in the reference arm"""
    return {"workload": "C2: SASO SparseSkOp<double> vec_nnz=8, d=2048, left sketch (lskges) of dense A 2^20 x 2048 per GPU",
            "d": D, "n_per_gpu": N_COLS, "m": m, "vec_nnz": VEC_NNZ, "layout": "ColMajor" if lay == "C" else "RowMajor",
            "n_gpus": world, "sharding": "column-sharded, S regenerated per rank, no collective",
            "l2": "inputs (17.2 GB) exceed L2; no flush needed"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def use_all_host_cores():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU legs want every core of the box. Must run
    before the oracle libraries (libgomp) are loaded; ref_set_threads() covers a runtime that is already initialised."""
    n = host_cores()
    os.environ["OMP_NUM_THREADS"] = str(n)
    os.environ["OPENBLAS_NUM_THREADS"] = str(n)
    os.environ.pop("OMP_THREAD_LIMIT", None)
    return n


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


# ---- CPU legs: the reference's own implementation (oracle/_ref) on the box's host cores ------------------------------------------
# The only place outside tests/ and smoke() that loads anything under oracle/; never inside a CUDA-timed region.
_ORACLE = {}


def oracle():
    if not _ORACLE:
        cores = use_all_host_cores()
        import numpy as np
        import oracle_lib as ol
        kind = "reference" if ol.REF is not None else "port"
        if ol.REF is not None and hasattr(ol.REF, "ref_set_threads"):
            ol.REF.ref_set_threads(ctypes.c_int(cores))
        threads = int(ol.REF.ref_omp_threads()) if ol.REF is not None else cores
        _ORACLE.update(np=np, ol=ol, kind=kind, which="ref" if kind == "reference" else "orc", cores=threads)
    return _ORACLE


def host_gaussian(np, count, dtype, seed=99):
    """synthetic Gaussian host data, generated once per call and not timed: a 2^22-sample block repeated (the CPU paths'
    time does not depend on the values)"""
    block = np.random.default_rng(seed).standard_normal(min(count, 1 << 22)).astype(dtype)
    out = np.empty(count, dtype)
    for o in range(0, count, len(block)):
        k = min(len(block), count - o)
        out[o:o + k] = block[:k]
    return out


def _leg(o, value, unit, sample, seconds):
    return {"value": value, "unit": unit, "cores": o["cores"], "kind": o["kind"], "sample": sample, "seconds": seconds}


_C2_HOST = {}


def cpu_c2(m=M_FULL, n=N_COLS):
    """fill_sparse + lskges (sparse_skops.hh:389-413, skge.hh:485-510) timed together, as one sketch_general call runs them"""
    o = oracle(); np, ol = o["np"], o["ol"]
    if _C2_HOST.get("shape") != (m, n):                       # kept across the steps of the reference arm (17.2 GB at full size)
        _C2_HOST.clear()
        _C2_HOST.update(shape=(m, n), A=host_gaussian(np, m * n, np.float64))
    A = _C2_HOST["A"]
    B = np.zeros(D * n)
    t0 = time.perf_counter()
    rc = ol.sketch_sparseop("f64", "L", "C", "N", "N", D, n, m, 1.0, D, m, VEC_NNZ, "S", [0, 0, 0, 0], [KEY, 0], 0, 0, A, m, 0.0, B, D, which=o["which"])
    dt = time.perf_counter() - t0
    assert rc == 0
    frac = "the full workload" if m == M_FULL else f"{m / M_FULL:.4f} of the workload's rows"
    return _leg(o, flops(m, n) / dt / 1e9, "GFLOP/s", f"one rank's slab, m={m} x n={n}, d={D}, vec_nnz={VEC_NNZ} ({frac}); fill_sparse + lskges timed "
                f"together, OpenMP threads={o['cores']}", dt)


def cpu_c1():
    """sketch_general = fill_dense + blas::gemm (skge.hh:173-215; sgemm = scipy-OpenBLAS through oracle/shims/blas.hh)"""
    o = oracle(); np, ol = o["np"], o["ol"]
    d, n, m = 1024, 1024, 65536
    A = host_gaussian(np, m * n, np.float32)
    B = np.zeros(d * n, np.float32)
    t0 = time.perf_counter()
    rc = ol.sketch_dense("f32", "L", "C", "N", "N", d, n, m, 1.0, d, m, "G", "L", [0, 0, 0, 0], [0, 0], 0, 0, A, m, 0.0, B, d, which=o["which"])
    dt = time.perf_counter() - t0
    assert rc == 0
    return _leg(o, 2.0 * d * n * m / dt / 1e9, "GFLOP/s", f"the full C1 config (fill_dense of S 1024 x 65536 + sgemm), threads={o['cores']}", dt)


def cpu_c3(fam, cols=16384):
    """fill_dense (dense_skops.hh:486-532) of a 65536 x `cols` column window of the 65536 x 65536 operator (natural ColMajor:
    the window is contiguous), 2^30 of the 2^32 entries by default"""
    o = oracle(); np, ol = o["np"], o["ol"]
    lib, pre = ol._lib(o["which"])
    n = 65536
    out = np.empty(n * cols, np.float32)
    nxt = ol.u32([0] * 4)
    t0 = time.perf_counter()
    rc = getattr(lib, pre + "fill_dense_f32")(ol.ch("C"), ol.c_i64(n), ol.c_i64(n), ol.ch(fam), ol.ch("L"), ol.c_i64(n), ol.c_i64(cols), ol.c_i64(0),
                                              ol.c_i64(0), ol.ptr(out), ol.u32([0] * 4), ol.u32([0, 0]), nxt)
    dt = time.perf_counter() - t0
    assert rc == 0
    return _leg(o, n * cols / dt / 1e9, "Gsamples/s", f"{n * cols} of the 2^32 entries (the first {cols} columns of the same operator), threads={o['cores']}", dt)


def cpu_c4(rows=100000):
    """sketch_sparse = fill_dense of the needed window of S + lsksp3 (sksp.hh:147-192) on the first `rows` rows of A
    (contraction axis sampled: the fill and the SpMM both shrink by the same factor)"""
    o = oracle(); np, ol = o["np"], o["ol"]
    d, n, per_row = 512, 1000000, 100
    rng = np.random.default_rng(7)
    cols = np.sort(rng.integers(0, n, size=(rows, per_row), dtype=np.int64), axis=1).reshape(-1)
    vals = host_gaussian(np, rows * per_row, np.float64, seed=8)
    rowptr = np.arange(0, rows * per_row + 1, per_row, dtype=np.int64)
    B = np.zeros(d * n)
    t0 = time.perf_counter()
    rc = ol.sketch_sparsedata("f64", "L", "C", "N", "N", d, n, rows, 1.0, d, n, "G", "L", [0, 0, 0, 0], [0, 0], 0, 0, "R", rows, n, vals, rowptr, cols,
                              0, 0, 0.0, B, d, which=o["which"])
    dt = time.perf_counter() - t0
    assert rc == 0
    return _leg(o, 2.0 * d * rows * per_row / dt / 1e9, "GFLOP/s", f"the first {rows} of the 10^6 rows of A (nnz = {rows * per_row}), S window 512 x {rows}, "
                f"B 512 x 10^6; threads={o['cores']}", dt)


def cpu_c5(slab=32768):
    """dense FP64 left sketch of a row slab of C5 (fill_dense of S' 8192 x slab + dgemm); the work is exactly linear in m"""
    o = oracle(); np, ol = o["np"], o["ol"]
    d, n = 8192, 4096
    A = host_gaussian(np, slab * n, np.float64)
    B = np.zeros(d * n)
    t0 = time.perf_counter()
    rc = ol.sketch_dense("f64", "L", "C", "N", "N", d, n, slab, 1.0, d, 1 << 23, "G", "L", [0, 0, 0, 0], [0, 0], 0, 0, A, slab, 0.0, B, d, which=o["which"])
    dt = time.perf_counter() - t0
    assert rc == 0
    return _leg(o, 2.0 * d * n * slab / dt / 1e9, "GFLOP/s", f"a slab of {slab} of the 2^23 rows of A (1/{(1 << 23) // slab} of the flops; the full "
                f"config extrapolates linearly to {dt * ((1 << 23) / slab):.0f} s), threads={o['cores']}", dt)


def pick(leg):
    return {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the headline path, same config, all host cores"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    legs = []
    for i in range(args.warmup + args.steps):
        leg = cpu_c2(args.m, N_COLS)
        if i >= args.warmup:
            legs.append(leg)
    v = statistics.mean(x["value"] for x in legs)
    ms = statistics.mean(x["seconds"] for x in legs) * 1e3
    base = pick(legs[-1])
    base["value"] = v
    line = {"impl": "reference", "metric": "sketch_gflops", "value": v, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": c2_config(args.m, args.layout, world), "cpu_baseline": base,
            "e2e": {"value": v, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---- GPU arm -----------------------------------------------------------------------------------------------------------------------
def cuda_time(torch, fn, reps, warm=1):
    """median device time of fn() in ms (CUDA events on the current stream, synchronised on both sides)"""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def wall_time(torch, fn, reps, warm=1):
    """mean wall-clock seconds of fn() (host-pointer calls: the copies are part of the call), device idle on both sides"""
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return statistics.mean(ts)


def measure_denominators(torch, dev):
    """cuBLAS TF32 and FP64 GEMM 8192^3 on this GPU, in this run (MEASURED_PEAKS.json holds neither)"""
    n = 8192
    out = {}
    for name, dt, tf32 in (("tf32_tflops", torch.float32, True), ("fp64_tflops", torch.float64, False)):
        a = torch.randn(n, n, dtype=dt, device=dev)
        b = torch.randn(n, n, dtype=dt, device=dev)
        c = torch.empty(n, n, dtype=dt, device=dev)
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        best = min(cuda_time(torch, lambda: torch.matmul(a, b, out=c), 1, warm=2) for _ in range(6))
        torch.backends.cuda.matmul.allow_tf32 = old
        out[name] = 2.0 * n ** 3 / best / 1e9
        del a, b, c
    out["how"] = "torch.matmul 8192^3 (cuBLAS), best of 6, CUDA events; tf32 with allow_tf32, fp64 = DGEMM"
    return out


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
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs (kernel experiments)")
    ap.add_argument("--c4-n", type=int, default=1000000)
    ap.add_argument("--c5-slab", type=int, default=65536)
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
    e2e, Ah = None, None
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
        te = torch.tensor([statistics.mean(ts)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": flops(m, n) * world / float(te.item()) / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": 8 * m * n * world,
               "d2h_bytes_per_step": 8 * D * n * world, "ms_per_step": float(te.item()) * 1e3, "steps": args.e2e_steps,
               "ms_steps": [round(t * 1e3, 1) for t in ts], "checksum0": chk, "host_memory": "pinned (torch pin_memory)"}
        # device path and host path agree
        assert torch.allclose(Bh, B.cpu(), rtol=1e-12, atol=1e-9)
        if world == 1 and not args.no_extras:
            # what a caller of the reference has: pageable memory (new T[]). The driver stages pageable copies through its own
            # pinned bounce buffers, so the same call runs at a fraction of the link rate; reported beside the pinned figure.
            Ap = torch.empty(m * n, dtype=torch.float64)
            Ap.copy_(Ah)
            Bp = torch.zeros(D * n, dtype=torch.float64)
            tp = wall_time(torch, lambda: rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, D, n, m, 1.0, S, 0, 0, Ap, lda, 0.0, Bp, ldb), 1, warm=1)
            e2e["pageable"] = {"value": flops(m, n) / tp / 1e9, "unit": "GFLOP/s", "ms_per_step": tp * 1e3,
                               "note": "same call with pageable (unpinned) host buffers, as a caller's new T[] would be"}
            assert torch.equal(Bp, Bh)
            del Ap, Bp

    extras = {}
    if world == 1 and not args.no_extras:
        # RowMajor data and the float variant of the same workload (same kernel family)
        extras.update(c2_variants(rb, torch, dev, m, hbm))
    del A, B
    torch.cuda.empty_cache()
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        # the reference on the host cores, full workload; its A is the pinned copy of the device's A when there is one
        if Ah is not None:
            _C2_HOST.update(shape=(m, n), A=Ah.numpy())
        cpu = cpu_c2(m, n)
        _C2_HOST.clear()
    if world == 1 and not args.no_extras:
        extras.update(run_extras(rb, torch, dev, hbm, Ah, args))
    if world > 1 and not args.no_extras:
        extras.update(run_sharded_c5(rb, torch, dist, dev, world, rank, args))
    if rank == 0:
        line = {"metric": "sketch_gflops", "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": c2_config(m, lay, world), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline}
        if cpu is not None:
            line["cpu_baseline"] = pick(cpu)
        if extras:
            line["extras"] = extras
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def c2_variants(rb, torch, dev, m, hbm):
    """the same SASO workload with row-major data, and in single precision (half the bytes per gather), device-resident"""
    out = {}
    n = N_COLS
    for name, dt, lay in (("C2_rowmajor_f64", torch.float64, "R"), ("C2_colmajor_f32", torch.float32, "C"), ("C2_rowmajor_f32", torch.float32, "R")):
        A = torch.empty(m * n, dtype=dt, device=dev)
        rb.fill_dense(rb.Layout(lay), rb.DenseDist(m, n), m, n, 0, 0, A, rb.RNGState(99))
        B = torch.zeros(D * n, dtype=dt, device=dev)
        S = rb.SparseSkOp(rb.SparseDist(D, m, VEC_NNZ), KEY, dtype=dt)
        lda, ldb = (m, D) if lay == "C" else (n, n)
        ms = cuda_time(torch, lambda: rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, D, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb), 5, warm=2)
        byts = A.element_size() * (m * n + D * n)
        out[name] = {"value": flops(m, n) / ms / 1e6, "unit": "GFLOP/s", "ms": ms,
                     "roofline": {"bound": "hbm", "achieved": byts / ms / 1e6, "peak": hbm, "unit": "GB/s", "frac": byts / ms / 1e6 / hbm, "traffic": None,
                                  "algorithmic_bytes": byts}}
        del A, B
    return out


def run_extras(rb, torch, dev, hbm, pinned, args):
    """the other BASELINE configs on one GPU: value (device-resident), roofline, cpu_baseline (the reference on the host
    cores, bounded sample stated), e2e (host-pointer call, copies inside the timed region)"""
    out = {}
    den = measure_denominators(torch, dev)
    out["denominators"] = den
    cpu_on = not args.no_cpu
    L, O = rb.Layout, rb.Op

    def host_buf(count, dtype):
        """pinned host memory for the e2e legs: a view of the headline's pinned buffer when it is large enough"""
        nbytes = count * torch.empty(0, dtype=dtype).element_size()
        if pinned is not None and pinned.numel() * 8 >= nbytes:
            return pinned.view(torch.uint8)[:nbytes].view(dtype)
        return torch.empty(count, dtype=dtype, pin_memory=True)

    # ---- C1: Gaussian DenseSkOp<float> d = 1024, A 65536 x 1024 (tcgen05 3xTF32) ----
    d, n, m = 1024, 1024, 65536
    A = torch.empty(m * n, dtype=torch.float32, device=dev)
    rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    B = torch.zeros(d * n, dtype=torch.float32, device=dev)
    S = rb.DenseSkOp(rb.DenseDist(d, m), 0)
    fl = 2.0 * d * n * m
    ms = cuda_time(torch, lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d), 10, warm=2)
    # the same data read as a RowMajor 65536 x 1024 matrix (unit stride along n): k-major copy route of the tcgen05 path
    ms_r = cuda_time(torch, lambda: rb.sketch_general(L.RowMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, n, 0.0, B, n), 10, warm=2)
    Su = rb.DenseSkOp(rb.DenseDist(d, m, "U"), 0)
    ms_u = cuda_time(torch, lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, Su, 0, 0, A, m, 0.0, B, d), 10, warm=2)
    peak = den["tf32_tflops"] / 3
    Ah, Bh = host_buf(m * n, torch.float32), torch.zeros(d * n, dtype=torch.float32, pin_memory=True)
    Ah.copy_(A)
    te = wall_time(torch, lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, Ah, m, 0.0, Bh, d), 3, warm=2)
    out["C1"] = {"workload": "Gaussian DenseSkOp<float> d=1024, left sketch of 65536 x 1024 A (ColMajor), operator never stored",
                 "value": fl / ms / 1e6, "unit": "GFLOP/s", "ms": ms, "dtype": "f32 (3xTF32 split, FP32 accumulate)",
                 "roofline": {"bound": "tensor", "achieved": fl / ms / 1e9, "peak": peak, "unit": "TFLOP/s", "frac": fl / ms / 1e9 / peak, "traffic": None,
                              "peak_source": "cuBLAS TF32 GEMM 8192^3 measured in this run / 3 (three MMAs per useful product)"},
                 "uniform_family": {"value": fl / ms_u / 1e6, "ms": ms_u, "frac": fl / ms_u / 1e9 / peak},
                 "rowmajor_data": {"value": fl / ms_r / 1e6, "ms": ms_r, "frac": fl / ms_r / 1e9 / peak},
                 "e2e": {"value": fl / te / 1e9, "unit": "GFLOP/s", "ms": te * 1e3, "h2d_bytes_per_step": 4 * m * n, "d2h_bytes_per_step": 4 * d * n}}
    if cpu_on:
        out["C1"]["cpu_baseline"] = pick(cpu_c1())
    del A, B, S, Su

    # ---- C3: fill_dense of 2^32 floats (65536 x 65536), Gaussian and Uniform ----
    n = 65536
    buf = torch.empty(n * n, dtype=torch.float32, device=dev)
    hb = host_buf(n * n, torch.float32)
    for fam in ("G", "U"):
        Dd = rb.DenseDist(n, n, fam, "L")
        ms = cuda_time(torch, lambda: rb.fill_dense(Dd, buf, rb.RNGState(0)), 5, warm=1)
        te = wall_time(torch, lambda: rb.fill_dense(Dd, hb, rb.RNGState(0)), 2, warm=1)
        key = "C3_gaussian" if fam == "G" else "C3_uniform"
        out[key] = {"workload": f"fill_dense of 2^32 float entries, DenseDist(65536, 65536, {'Gaussian' if fam == 'G' else 'Uniform'})",
                    "value": n * n / ms / 1e6, "unit": "Gsamples/s", "ms": ms, "dtype": "f32",
                    "roofline": {"bound": "hbm", "achieved": 4.0 * n * n / ms / 1e6, "peak": hbm, "unit": "GB/s", "frac": 4.0 * n * n / ms / 1e6 / hbm,
                                 "traffic": None, "algorithmic_bytes": 4.0 * n * n},
                    "e2e": {"value": n * n / te / 1e9, "unit": "Gsamples/s", "ms": te * 1e3, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4 * n * n}}
        if cpu_on:
            out[key]["cpu_baseline"] = pick(cpu_c3(fam))
    del buf

    # ---- C4: sketch_sparse, Gaussian d = 512, CSR n x n with 100 nnz per row ----
    n, d, per_row = args.c4_n, 512, 100
    nnz = n * per_row
    cols = torch.empty(nnz, dtype=torch.int64, device=dev)
    rb.repeated_fisher_yates(rb.RNGState(7), per_row, n, n, cols)        # the reference's own without-replacement sampler
    cols = cols.view(n, per_row).sort(dim=1).values.reshape(-1).contiguous()
    vals = torch.empty(nnz, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(nnz, 1), vals, rb.RNGState(8))
    rowptr = torch.arange(0, nnz + 1, per_row, dtype=torch.int64, device=dev)
    Asp = rb.CSRMatrix(n, n, vals, rowptr, cols)
    S = rb.DenseSkOp(rb.DenseDist(d, n), 0, dtype=torch.float64)
    B = torch.zeros(d * n, dtype=torch.float64, device=dev)
    ms = cuda_time(torch, lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, Asp, 0, 0, 0.0, B, d), 3, warm=1)
    byts = nnz * 16.0 + (n + 1) * 8.0 + d * n * 8.0
    fl = 2.0 * d * nnz
    Ahs = rb.CSRMatrix(n, n, vals.cpu().pin_memory(), rowptr.cpu().pin_memory(), cols.cpu().pin_memory())
    Bh = host_buf(d * n, torch.float64)
    te = wall_time(torch, lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, Ahs, 0, 0, 0.0, Bh, d), 2, warm=1)
    out["C4"] = {"workload": f"sketch_sparse: Gaussian DenseSkOp<double> d=512 x CSR {n} x {n}, 100 nnz/row (nnz = {nnz})" + ("" if n == 1000000 else " -- REDUCED n"),
                 "value": fl / ms / 1e6, "unit": "GFLOP/s", "ms": ms, "dtype": "f64",
                 "roofline": {"bound": "hbm", "achieved": byts / ms / 1e6, "peak": hbm, "unit": "GB/s", "frac": byts / ms / 1e6 / hbm, "traffic": None,
                              "algorithmic_bytes": byts, "fp64_frac": fl / ms / 1e9 / den["fp64_tflops"],
                              "note": "compulsory bytes = nnz*16 + (rows+1)*8 + d*n*8 (SURVEY 8d); fp64_frac = flops against the measured DGEMM rate"},
                 "e2e": {"value": fl / te / 1e9, "unit": "GFLOP/s", "ms": te * 1e3, "h2d_bytes_per_step": nnz * 16 + (n + 1) * 8, "d2h_bytes_per_step": 8 * d * n}}
    if cpu_on:
        out["C4"]["cpu_baseline"] = pick(cpu_c4(50000))
    del Asp, Ahs, B, vals, cols, rowptr

    # ---- C5: double, Gaussian d = 8192, n = 4096, one GPU, a slab of the 2^23 rows (work exactly linear in m) ----
    d, n, m = 8192, 4096, args.c5_slab
    A = torch.empty(m * n, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    B = torch.zeros(d * n, dtype=torch.float64, device=dev)
    S = rb.DenseSkOp(rb.DenseDist(d, 1 << 23), 0, dtype=torch.float64)
    fl = 2.0 * d * n * m
    ms = cuda_time(torch, lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 1.0, B, d), 3, warm=1)
    Ah, Bh = host_buf(m * n, torch.float64), torch.zeros(d * n, dtype=torch.float64, pin_memory=True)
    Ah.copy_(A)
    te = wall_time(torch, lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, Ah, m, 0.0, Bh, d), 2, warm=1)
    out["C5_slab"] = {"workload": f"Gaussian DenseSkOp<double> d=8192, left sketch of a {m}-row slab of the 2^23 x 4096 A (1/{(1 << 23) // m} of the "
                                  f"config; extrapolated full problem {ms * ((1 << 23) / m) / 1e3:.1f} s on one GPU)",
                      "value": fl / ms / 1e6, "unit": "GFLOP/s", "ms": ms, "dtype": "f64 (DMMA)",
                      "roofline": {"bound": "tensor", "achieved": fl / ms / 1e9, "peak": den["fp64_tflops"], "unit": "TFLOP/s",
                                   "frac": fl / ms / 1e9 / den["fp64_tflops"], "traffic": None, "peak_source": "cuBLAS DGEMM 8192^3 measured in this run"},
                      "e2e": {"value": fl / te / 1e9, "unit": "GFLOP/s", "ms": te * 1e3, "h2d_bytes_per_step": 8 * m * n, "d2h_bytes_per_step": 8 * d * n}}
    if cpu_on:
        out["C5_slab"]["cpu_baseline"] = pick(cpu_c5(32768))
    return out


def run_sharded_c5(rb, torch, dist, dev, world, rank, args):
    """C5 on N GPUs, strong scaling over a fixed slab: row-sharded (each rank sketches its rows with co_s = its start, then
    ONE NCCL reduce-scatter of the 8192 x 4096 partials, inside the timed region) and column-sharded (no collective)."""
    from randblas_b200 import sharded
    d, n, m, m_full = 8192, 4096, args.c5_slab, 1 << 23
    S = rb.DenseSkOp(rb.DenseDist(d, m_full), 0, dtype=torch.float64)
    fl = 2.0 * d * n * m
    L, O = rb.Layout, rb.Op

    def timed(fn, reps=3):
        fn()
        best = []
        for _ in range(reps):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best.append(float(t.item()))
        return min(best)

    out = {}
    r0, ml = sharded.shard_range(m, world, rank, 16)
    A = torch.empty(ml * n, dtype=torch.float64, device=dev)
    rb.fill_dense(L.ColMajor, rb.DenseDist(m, n), ml, n, r0, 0, A, rb.RNGState(99))
    B = torch.zeros(d * n, dtype=torch.float64, device=dev)
    comm = sharded.Comm(dist, dev)                           # rbx_comm_init over an id broadcast through torch.distributed
    piece = torch.empty(d * n // world, dtype=torch.float64, device=dev)
    ms = timed(lambda: comm.sketch_dense_row_sharded(L.ColMajor, O.NoTrans, d, n, m, 1.0, S, A, ml, 0.0, B, d, piece, align=16))
    ms_rs = timed(lambda: comm.reduce_scatter(B, piece))
    out["C5_row_sharded"] = {"workload": f"C5 slab of {m} rows split over {world} GPUs (strong scaling), partial sketch per rank + ncclReduceScatter "
                                         f"of the {d} x {n} doubles through the C ABI (rbx_shard_sketch_dense_f64)",
                             "value": fl / ms / 1e6, "unit": "GFLOP/s", "ms": ms, "n_gpus": world, "scaling": "strong",
                             "reduce_scatter_ms": ms_rs, "reduce_scatter_bytes": 8 * d * n,
                             "reduce_scatter_busbw_gbs": 8.0 * d * n * (world - 1) / world / ms_rs / 1e6}
    comm.destroy()
    del A, B, piece
    c0, nl = sharded.shard_range(n, world, rank, 128)
    A = torch.empty(m * nl, dtype=torch.float64, device=dev)
    rb.fill_dense(L.ColMajor, rb.DenseDist(m, n), m, nl, 0, c0, A, rb.RNGState(99))
    B = torch.zeros(d * nl, dtype=torch.float64, device=dev)
    ms = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, nl, m, 1.0, S, 0, 0, A, m, 1.0, B, d))
    out["C5_column_sharded"] = {"workload": f"C5 slab of {m} rows, the {n} columns split over {world} GPUs (strong scaling), no collective",
                                "value": fl / ms / 1e6, "unit": "GFLOP/s", "ms": ms, "n_gpus": world, "scaling": "strong"}
    return out


if __name__ == "__main__":
    main()
