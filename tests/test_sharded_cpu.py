"""Multi-rank host logic on CPU: world_size-2 `gloo` process groups.

The local compute of each rank is the ORACLE here (there is no GPU in this test); what is under test is the sharding
arithmetic of randblas_b200.sharded: shard ranges, the co_s offset per rank, beta applied once, the collective.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_local(kind):
    import oracle_lib as ol
    import randblas_b200 as rb

    def run(layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, lda, beta, B, ldb):
        lay = "C" if layout == rb.Layout.ColMajor else "R"
        oS = "N" if opS == rb.Op.NoTrans else "T"
        oA = "N" if opA == rb.Op.NoTrans else "T"
        An, Bn = A.numpy(), B.numpy()
        ctr, key = list(S.seed_state.counter), list(S.seed_state.key)
        if kind == "sparse":
            rc = ol.sketch_sparseop("f64", "L", lay, oS, oA, d, n, m, alpha, S.dist.n_rows, S.dist.n_cols, S.dist.vec_nnz, "S", ctr, key,
                                    ro_s, co_s, An, lda, beta, Bn, ldb)
        else:
            rc = ol.sketch_dense("f64", "L", lay, oS, oA, d, n, m, alpha, S.dist.n_rows, S.dist.n_cols, "G", "L", ctr, key,
                                 ro_s, co_s, An, lda, beta, Bn, ldb)
        assert rc == 0
    return run


def _worker(rank, world, port, kind, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import randblas_b200 as rb
        from randblas_b200 import sharded
        d, n, m = 16, 10, 203                      # m does not divide evenly: ragged shards
        rng = np.random.default_rng(5)
        A_full = rng.standard_normal((m, n))       # row-major m x n
        B0 = rng.standard_normal((d, n))
        if kind == "sparse":
            S = rb.SparseSkOp(rb.SparseDist(d, m, 4), 1997)
        else:
            S = rb.DenseSkOp(rb.DenseDist(d, m), 7, dtype=torch.float64)
        local = _oracle_local(kind)
        # reference result on one rank: B = 1.5 * S * A - 0.5 * B0
        want = torch.from_numpy(B0.copy()).reshape(-1)
        local(rb.Layout.RowMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, S, 0, 0, torch.from_numpy(A_full.copy()).reshape(-1), n, -0.5, want, n)

        # --- shard ranges cover [0, m) exactly once, aligned
        ranges = [sharded.shard_range(m, world, r, 4) for r in range(world)]
        assert ranges[0][0] == 0 and sum(x[1] for x in ranges) == m
        assert all(ranges[i][0] + ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
        assert all(x[0] % 4 == 0 for x in ranges)

        # --- row-sharded: contraction axis split, one all_reduce
        start, ml = sharded.shard_range(m, world, rank, 4)
        A_loc = torch.from_numpy(A_full[start:start + ml].copy()).reshape(-1)
        B = torch.from_numpy(B0.copy()).reshape(-1)
        sharded.sketch_general_row_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B, n, align=4, local_sketch=local)
        err_row = float((B - want).abs().max())

        # --- row-sharded with reduce_scatter: each rank ends with its slice of the flattened result
        B2 = torch.from_numpy(B0.copy()).reshape(-1)
        out = sharded.sketch_general_row_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, d, n, m, 1.5, S, A_loc, n, -0.5, B2, n, align=4,
                                                 local_sketch=local, scatter=True)
        per = want.numel() // world
        err_rs = float((out - want[rank * per:(rank + 1) * per]).abs().max())

        # --- column-sharded: disjoint column slabs, no communication; gather only to check
        c0, nl = sharded.shard_range(n, world, rank, 1)
        A_col = torch.from_numpy(np.ascontiguousarray(A_full[:, c0:c0 + nl])).reshape(-1)
        B_col = torch.from_numpy(np.ascontiguousarray(B0[:, c0:c0 + nl])).reshape(-1)
        sharded.sketch_general_column_sharded(rb.Layout.RowMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, nl, m, 1.5, S, A_col, nl, -0.5, B_col, nl,
                                              local_sketch=local)
        err_col = float((B_col.reshape(d, nl) - want.reshape(d, n)[:, c0:c0 + nl]).abs().max())
        errs = torch.tensor([err_row, err_rs, err_col], dtype=torch.float64)
        dist.all_reduce(errs, op=dist.ReduceOp.MAX)
        if rank == 0:
            q.put(errs.tolist())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["sparse", "dense"])
def test_world2_gloo_sharded_sketch(kind):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    err_row, err_rs, err_col = q.get()
    assert err_row < 1e-11 and err_rs < 1e-11      # different summation order than the one-rank result
    assert err_col < 1e-13
