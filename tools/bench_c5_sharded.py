#!/usr/bin/env python
"""BASELINE config 5 on N GPUs of one box: double A (a slab of m' rows of the 2^23 x 4096 matrix), Gaussian d = 8192,
column-sharded (no communication) and row-sharded (partial sketch per GPU + one NCCL reduce_scatter of the 8192 x 4096
partials over NVLink). One JSON line per mode on rank 0; times are device times (CUDA events), max over ranks.

  python tools/bench_c5_sharded.py                          # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
      tools/bench_c5_sharded.py --slab 32768

The work is exactly linear in m (the operator's counters advance with co_s), so the slab result extrapolates to the full
2^23 rows; the slab size is stated in the output. "strong" scaling: the slab is fixed and split across the ranks.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
import torch.distributed as dist                # noqa: E402
import randblas_b200 as rb                      # noqa: E402
from randblas_b200 import sharded               # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slab", type=int, default=32768)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--pretend-world", type=int, default=0, help="one process times rank 0's column shard of a G-GPU run")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    d, n, m, m_full = 8192, 4096, args.slab, 1 << 23
    S = rb.DenseSkOp(rb.DenseDist(d, m_full), 0, dtype=torch.float64)
    flops = 2.0 * d * n * m

    def timed(fn):
        fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ts = []
        for _ in range(args.reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        return min(ts)

    # column-sharded: this rank's n / world columns of the slab (column-major, lda = m), generated on the device
    pw = args.pretend_world or world
    c0, nl = sharded.shard_range(n, pw, rank, 128)
    A = torch.empty(m * nl, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(m, n), m, nl, 0, c0, A, rb.RNGState(99))
    B = torch.zeros(d * nl, dtype=torch.float64, device=dev)
    ms = timed(lambda: sharded.sketch_general_column_sharded(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, nl, m, 1.0, S, A, m, 1.0, B, d))
    if rank == 0:
        print(json.dumps({"config": "C5", "mode": "column-sharded", "n_gpus": pw, "slab_rows": m, "ms_slab": ms, "tflops_aggregate": flops / ms / 1e9,
                          "tflops_per_gpu": flops / ms / 1e9 / pw, "pretend": bool(args.pretend_world), "extrapolated_full_s": ms * (m_full / m) / 1e3, "collective": None}), flush=True)
    del A, B
    if args.pretend_world:
        return

    # row-sharded: this rank's m / world rows of the slab (all n columns), then one reduce_scatter of the d x n partials
    r0, ml = sharded.shard_range(m, world, rank, 16)
    A = torch.empty(ml * n, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(m, n), ml, n, r0, 0, A, rb.RNGState(99))
    B = torch.zeros(d * n, dtype=torch.float64, device=dev)
    ms = timed(lambda: sharded.sketch_general_row_sharded(rb.Layout.ColMajor, rb.Op.NoTrans, d, n, m, 1.0, S, A, ml, 0.0, B, d, align=16,
                                                          scatter=world > 1))
    if rank == 0:
        print(json.dumps({"config": "C5", "mode": "row-sharded", "n_gpus": world, "slab_rows": m, "ms_slab": ms, "tflops_aggregate": flops / ms / 1e9,
                          "tflops_per_gpu": flops / ms / 1e9 / world, "collective": "reduce_scatter of 8192 x 4096 doubles (268 MB) per slab"
                          if world > 1 else None,
                          "note": "the full problem runs ONE collective after all of a rank's row slabs; per-slab timing overstates its share"}),
              flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
