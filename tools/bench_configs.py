#!/usr/bin/env python
"""Secondary measurements: every BASELINE.json config on one B200 (C2 is bench.py's headline and is only summarised here).

One JSON object per config on stdout, device-resident operands, CUDA events, median of `--reps` after a warm-up.
Algorithmic work and rooflines follow SURVEY.md 8(d); HBM and BF16 peaks come from MEASURED_PEAKS.json, the FP64 peak
(not in that file) is the nominal 40 TFLOP/s and is labelled as such.
  C1  Gaussian DenseSkOp<float> d=1024, A 65536 x 1024 ColMajor (tcgen05 3xTF32 path)
  C3  fill_dense 2^32 floats, Gaussian and Uniform
  C4  sketch_sparse, Gaussian d=512, CSR n x n with 100 nnz/row (n = --c4-n, default 10^5; 10^6 is the full config)
  C5  double, d=8192, n=4096, a slab of m' rows of the 2^23 (work is linear in m; the slab is stated in the output)
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                      # noqa: E402
import randblas_b200 as rb        # noqa: E402


def timeit(fn, reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--c4-n", type=int, default=100000)
    ap.add_argument("--c5-slab", type=int, default=1 << 15)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    only = set(args.only.split(",")) if args.only else None
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    hbm = float(peaks["hbm_gbs"])
    dev = torch.device("cuda", 0)
    out = []

    def emit(d):
        print(json.dumps(d), flush=True)
        out.append(d)

    if not only or "C1" in only:
        d, n, m = 1024, 1024, 65536
        A = torch.empty(m * n, dtype=torch.float32, device=dev)
        rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
        B = torch.zeros(d * n, dtype=torch.float32, device=dev)
        for fam in ("G", "U"):
            S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 0)
            ms = timeit(lambda: rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d), args.reps)
            tf = 2.0 * d * n * m / ms / 1e9
            emit({"config": "C1", "family": fam, "ms": ms, "tflops": tf, "flops": 2.0 * d * n * m,
                  "roofline": {"bound": "tensor", "peak_note": "3xTF32: useful peak = TF32 dense / 3; TF32 dense = cuBLAS TF32 GEMM 8192^3 measured on this pool (746.4 TFLOP/s, tools/micro/cublas_peaks.py)",
                               "peak_tflops": 746.4 / 3, "frac": tf / (746.4 / 3)},
                  "kernel": "sketch_dense_tc_kernel (tcgen05) + tc_reduce_kernel"})
        del A, B

    if not only or "C3" in only:
        n = 65536
        buf = torch.empty(n * n, dtype=torch.float32, device=dev)
        for fam in ("G", "U"):
            D = rb.DenseDist(n, n, fam, "L")
            ms = timeit(lambda: rb.fill_dense(D, buf, rb.RNGState(0)), args.reps)
            gs = n * n / ms / 1e6
            emit({"config": "C3", "family": fam, "entries": n * n, "ms": ms, "gsamples_per_s": gs,
                  "roofline": {"bound": "hbm", "achieved_gbs": 4.0 * n * n / ms / 1e6, "peak_gbs": hbm, "frac": 4.0 * n * n / ms / 1e6 / hbm},
                  "kernel": "fill_dense_kernel"})
        del buf

    if not only or "C4" in only:
        n, d, per_row = args.c4_n, 512, 100
        nnz = n * per_row
        # 100 nonzeros per row, columns from the reference's own without-replacement sampler, values Gaussian
        cols = torch.empty(nnz, dtype=torch.int64, device=dev)
        rb.repeated_fisher_yates(rb.RNGState(7), per_row, n, n, cols)
        cols = cols.view(n, per_row).sort(dim=1).values.reshape(-1).contiguous()
        vals = torch.empty(nnz, dtype=torch.float64, device=dev)
        rb.fill_dense(rb.DenseDist(nnz, 1), vals, rb.RNGState(8))
        rowptr = torch.arange(0, nnz + 1, per_row, dtype=torch.int64, device=dev)
        A = rb.CSRMatrix(n, n, vals, rowptr, cols)
        S = rb.DenseSkOp(rb.DenseDist(d, n), 0, dtype=torch.float64)
        B = torch.zeros(d * n, dtype=torch.float64, device=dev)
        ms = timeit(lambda: rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, n, 1.0, S, 0, 0, A, 0, 0, 0.0, B, d), max(2, args.reps // 2))
        byts = nnz * 16.0 + (n + 1) * 8.0 + d * n * 8.0
        emit({"config": "C4", "n": n, "full_config_n": 1000000, "nnz": nnz, "d": d, "ms": ms, "gflops": 2.0 * d * nnz / ms / 1e6,
              "roofline": {"bound": "hbm", "algorithmic_bytes": byts, "achieved_gbs": byts / ms / 1e6, "peak_gbs": hbm, "frac": byts / ms / 1e6 / hbm},
              "kernel": "spd_group_by_column (CSR -> column segments) + fill_dense_kernel (float operator slab) + spdata_gather_kernel"})
        del A, B, vals, cols

    if not only or "C5" in only:
        d, n, m = 8192, 4096, args.c5_slab
        A = torch.empty(m * n, dtype=torch.float64, device=dev)
        rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
        B = torch.zeros(d * n, dtype=torch.float64, device=dev)
        S = rb.DenseSkOp(rb.DenseDist(d, 1 << 23), 0, dtype=torch.float64)
        ms = timeit(lambda: rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 1.0, B, d), max(2, args.reps // 2))
        tf = 2.0 * d * n * m / ms / 1e9
        emit({"config": "C5", "slab_rows": m, "full_rows": 1 << 23, "ms_slab": ms, "tflops": tf, "extrapolated_full_s": ms * ((1 << 23) / m) / 1e3,
              "roofline": {"bound": "fp64", "peak_tflops": 37.1, "peak_note": "DMMA.8x8x4 register-resident issue rate measured on this pool's B200 (tools/micro/fp64_peak.cu; cuBLAS DGEMM 8192^3 = 35.5)", "frac": tf / 37.1},
              "kernel": "fill_dense_kernel (operator k-panels) + dmma_gemm_kernel"})
    return out


if __name__ == "__main__":
    main()
