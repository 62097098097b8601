"""Launched with torchrun on N >= 2 GPUs (not collected by pytest):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py
Row-sharded SASO and dense sketches over NCCL (all_reduce and reduce_scatter) against the one-GPU result, and the
column-sharded form against slices of it. Prints one JSON line on rank 0.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import randblas_b200 as rb                      # noqa: E402
from randblas_b200 import sharded                # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    comm = sharded.Comm(dist, dev)                # the library's own communicator (rbx_comm_init), id handed over by torch.distributed
    out["nccl_version"] = comm.nccl_version()
    for kind in ("sparse", "dense"):
        d, n, m = (2048, 512, 65536 + 448) if kind == "sparse" else (256, 128, 8192 + 12)
        A = torch.empty(m * n, dtype=torch.float64, device=dev)
        rb.fill_dense(rb.Layout.RowMajor, rb.DenseDist(m, n), m, n, 0, 0, A, rb.RNGState(99))
        S = rb.SparseSkOp(rb.SparseDist(d, m, 8), 1997) if kind == "sparse" else rb.DenseSkOp(rb.DenseDist(d, m), 0, dtype=torch.float64)
        B0 = torch.ones(d * n, dtype=torch.float64, device=dev)
        want = B0.clone()
        rb.sketch_general(rb.Layout.RowMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, S, 0, 0, A, n, -0.5, want, n)
        scale = float(want.abs().max())
        start, ml = sharded.shard_range(m, world, rank, 64)
        A_loc = A[start * n:(start + ml) * n]
        B = B0.clone()
        sharded.sketch_general_row_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B, n, align=64)
        e1 = float((B - want).abs().max()) / scale
        B2 = B0.clone()
        part = sharded.sketch_general_row_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B2, n, align=64, scatter=True)
        per = want.numel() // world
        e2 = float((part - want[rank * per:(rank + 1) * per]).abs().max()) / scale
        c0, nl = sharded.shard_range(n, world, rank, 16)
        A_col = A.view(m, n)[:, c0:c0 + nl].contiguous().view(-1)
        B_col = torch.ones(d * nl, dtype=torch.float64, device=dev)
        sharded.sketch_general_column_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, nl, m, 1.5, S, A_col, nl, -0.5, B_col, nl)
        e3 = float((B_col.view(d, nl) - want.view(d, n)[:, c0:c0 + nl]).abs().max()) / scale
        # the same two row-sharded forms through the C ABI: rbx_shard_sketch_* issue the NCCL collective themselves
        fn = comm.sketch_sparse_row_sharded if kind == "sparse" else comm.sketch_dense_row_sharded
        B3 = B0.clone()
        fn(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B3, n, align=64)
        e4 = float((B3 - want).abs().max()) / scale
        B4, piece = B0.clone(), torch.empty(per, dtype=torch.float64, device=dev)
        fn(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B4, n, piece=piece, align=64)
        e5 = float((piece - want[rank * per:(rank + 1) * per]).abs().max()) / scale
        bitwise = bool(torch.equal(B3, B)) and bool(torch.equal(piece, part))       # same partials, same NCCL reduction
        errs = torch.tensor([e1, e2, e3, e4, e5, 0.0 if bitwise else 1.0], dtype=torch.float64, device=dev)
        dist.all_reduce(errs, op=dist.ReduceOp.MAX)
        out[kind] = {"row_allreduce_relerr": errs[0].item(), "row_reducescatter_relerr": errs[1].item(), "column_relerr": errs[2].item(),
                     "cabi_allreduce_relerr": errs[3].item(), "cabi_reducescatter_relerr": errs[4].item(),
                     "cabi_equals_torch_distributed_bitwise": errs[5].item() == 0.0}
        assert errs[:5].max().item() < 1e-11, out
    comm.destroy()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
