"""Multi-GPU forms of the sketching path: one process per GPU, torch.distributed for the plumbing.

Both forms rest on the fact that a RandBLAS operator is a pure function of (distribution, RNGState): every rank can
regenerate exactly the slice of S it needs through the (ro_s, co_s) sub-matrix arguments the reference already has.

* column-sharded  B[:, cols_r] = S * A[:, cols_r]         -- no communication at all; each rank calls
                                                              `sketch_general` on its own column slab.
* row-sharded     B = sum_r S[:, rows_r] * A[rows_r, :]   -- each rank sketches its row slab with co_s = start_r
                                                              (left sketch, opS = NoTrans), then ONE collective sums the
                                                              d x n partials (NCCL all_reduce / reduce_scatter over NVLink).
"""
import ctypes

from . import api
from ._capi import lib


def shard_range(total, world, rank, align=1):
    """[start, start + length) of `total` items owned by `rank` (contiguous, sizes differ by at most `align`)."""
    s, n = ctypes.c_int64(), ctypes.c_int64()
    lib.rbx_shard_range(int(total), int(world), int(rank), int(align), ctypes.byref(s), ctypes.byref(n))
    return s.value, n.value


def _dist():
    import torch.distributed as dist
    return dist


def sketch_general_column_sharded(layout, opS, opA, d, n_local, m, alpha, S, A_local, lda, beta, B_local, ldb, ro_s=0, co_s=0,
                                  local_sketch=None):
    """Left sketch of this rank's column slab (n_local columns of op(A)). Nothing is exchanged: S is regenerated per rank."""
    fn = local_sketch or api.sketch_general
    return fn(layout, opS, opA, d, n_local, m, alpha, S, ro_s, co_s, A_local, lda, beta, B_local, ldb)


def sketch_general_row_sharded(layout, opA, d, n, m_total, alpha, S, A_local, lda, beta, B, ldb, group=None, ro_s=0, co_s=0,
                               align=1, local_sketch=None, scatter=False):
    """B = alpha * S[ro_s:ro_s+d, co_s:co_s+m_total] * op(A) + beta * B with the contraction axis split across ranks.

    A_local holds this rank's `m_local` rows of op(A) (shard_range(m_total, world, rank, align)). B is a torch tensor
    holding the d x n result (ld `ldb`); on return every rank holds the full sum (all_reduce), or -- with scatter=True and
    a dense B (ldb equal to the leading extent) -- its 1/world slice of the flattened result via reduce_scatter_tensor.
    beta is applied on rank 0 only so that beta * B enters the sum once.
    """
    dist = _dist()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    start, m_local = shard_range(m_total, world, rank, align)
    fn = local_sketch or api.sketch_general
    my_beta = beta if rank == 0 else 0.0
    if m_local > 0:
        fn(layout, api.Op.NoTrans, opA, d, n, m_local, alpha, S, ro_s, co_s + start, A_local, lda, my_beta, B, ldb)
    else:
        B.mul_(my_beta)
    if world == 1:
        return B
    if scatter:
        import torch
        assert B.numel() % world == 0, "reduce_scatter needs the flattened result to divide evenly"
        out = torch.empty(B.numel() // world, dtype=B.dtype, device=B.device)
        dist.reduce_scatter_tensor(out, B.view(-1), op=dist.ReduceOp.SUM, group=group)
        return out
    dist.all_reduce(B, op=dist.ReduceOp.SUM, group=group)
    return B
