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


class Comm:
    """The library's own NCCL communicator (rbx_comm_*, include/rbx.h) and the row-sharded sketches that run through it
    (rbx_shard_sketch_*): partial sketch per rank + ONE ncclReduceScatter / ncclAllReduce issued by the C ABI on the sketch's
    stream. torch.distributed is used once, to hand rank 0's 128-byte NCCL id to the other ranks; any other channel works
    (pass `unique_id` / `world` / `rank` explicitly)."""

    def __init__(self, dist=None, device=None, group=None, unique_id=None, world=None, rank=None):
        import torch
        if device is not None:
            torch.cuda.set_device(device)
        if unique_id is None:
            dist = dist or _dist()
            world, rank = dist.get_world_size(group), dist.get_rank(group)
            box = [None]
            if rank == 0:
                buf = (ctypes.c_uint8 * 128)()
                api.check(lib.rbx_comm_unique_id(buf), "rbx_comm_unique_id")
                box[0] = bytes(buf)
            dist.broadcast_object_list(box, src=0, group=group)
            unique_id = box[0]
        self.world, self.rank = int(world), int(rank)
        self._h = ctypes.c_void_p()
        ida = (ctypes.c_uint8 * 128).from_buffer_copy(unique_id)
        api.check(lib.rbx_comm_init(ctypes.byref(self._h), self.world, self.rank, ida), "rbx_comm_init")

    def nccl_version(self):
        v = ctypes.c_int()
        lib.rbx_comm_info(self._h, None, None, ctypes.byref(v))
        return v.value

    def destroy(self):
        if self._h:
            api.check(lib.rbx_comm_destroy(self._h), "rbx_comm_destroy")
            self._h = ctypes.c_void_p()

    def reduce_scatter(self, B, piece):
        api._call(f"rbx_comm_reduce_scatter_{api._suf(B.dtype)}", self._h, api._ptr(B), B.numel(), api._ptr(piece), api._stream(B))
        return piece

    def all_reduce(self, B):
        api._call(f"rbx_comm_all_reduce_{api._suf(B.dtype)}", self._h, api._ptr(B), B.numel(), api._stream(B))
        return B

    def sketch_dense_row_sharded(self, layout, opA, d, n, m_total, alpha, S, A_local, lda, beta, B, ldb, piece=None, align=1, ro_s=0, co_s=0):
        """B (d x n) = alpha * S[ro_s:, co_s:co_s+m_total] * op(A) + beta * B, op(A)'s rows split over the ranks (DenseSkOp S)"""
        D, dt = S.dist, B.dtype
        api._call(f"rbx_shard_sketch_dense_{api._suf(dt)}", self._h, api._c(layout), api._c(opA), d, n, m_total, api._scalar(dt, alpha),
                  D.n_rows, D.n_cols, api._c(D.family), api._c(D.major_axis), api._u32(S.seed_state.counter), api._u32(S.seed_state.key),
                  ro_s, co_s, api._ptr(A_local), lda, api._scalar(dt, beta), api._ptr(B), ldb, api._ptr(piece), align, api._stream(B))
        return piece if piece is not None else B

    def sketch_sparse_row_sharded(self, layout, opA, d, n, m_total, alpha, S, A_local, lda, beta, B, ldb, piece=None, align=1, ro_s=0, co_s=0):
        """the same with a SparseSkOp (SASO / LASO) regenerated per rank from its seed state"""
        D, dt = S.dist, B.dtype
        api._call(f"rbx_shard_sketch_sparseop_{api._suf(dt)}", self._h, api._c(layout), api._c(opA), d, n, m_total, api._scalar(dt, alpha),
                  D.n_rows, D.n_cols, D.vec_nnz, api._c(D.major_axis), api._u32(S.seed_state.counter), api._u32(S.seed_state.key),
                  ro_s, co_s, api._ptr(A_local), lda, api._scalar(dt, beta), api._ptr(B), ldb, api._ptr(piece), align, api._stream(B))
        return piece if piece is not None else B
