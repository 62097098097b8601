"""CPU tests of the drop-in boundary: librbx.so loads, exports every symbol include/rbx.h declares, the host-side
state algebra is exact, argument checks mirror the reference, and compute entry points refuse to run without a GPU
(no CPU fallback)."""
import ctypes

import numpy as np
import pytest
import torch

import oracle_lib as ol


def test_library_exports_every_declared_symbol(rb):
    assert len(rb.PROTOS) >= 40
    for name in rb.PROTOS:
        assert hasattr(rb.lib, name), name
    assert rb.lib.rbx_version().startswith(b"rbx")


def test_counter_algebra_matches_oracle(rb):
    rng = np.random.default_rng(0)
    for _ in range(200):
        ctr = [int(x) for x in rng.integers(0, 2**32, 4)]
        if rng.random() < 0.3:
            ctr[0] = ctr[1] = 0xFFFFFFFF
        n = int(rng.integers(0, 2**63)) * 2 + int(rng.integers(0, 2))
        assert rb.RNGState(1, ctr).incr(n).counter == ol.ctr_add(ctr, n)


def test_next_state_and_layout(rb):
    for (r, c) in [(7, 20), (20, 7), (8, 8), (1024, 65536), (8192, 1 << 23), (65536, 65536)]:
        for ma in (rb.MajorAxis.Long, rb.MajorAxis.Short):
            D = rb.DenseDist(r, c, rb.DenseDistName.Gaussian, ma)
            S = rb.DenseSkOp(D, rb.RNGState(5, [0xFFFFFFF0, 0xFFFFFFFF, 0, 0]))
            assert S.next_state.counter == ol.dense_next_state(r, c, ma.value, S.seed_state.counter)
            assert rb.dist_to_layout(D).value == ol.ORC.orc_dist_to_layout(ctypes.c_int64(r), ctypes.c_int64(c), ol.ch(ma.value)).to_bytes(1, "little").decode()
    # sparse next_state reproduces sparse_skops.hh:115-126 as written (min dim for Short, max for Long)
    S = rb.SparseSkOp(rb.SparseDist(7, 20, 3, rb.MajorAxis.Short), 0)
    assert S.next_state.counter == [21, 0, 0, 0]
    S = rb.SparseSkOp(rb.SparseDist(7, 20, 3, rb.MajorAxis.Long), 0)
    assert S.next_state.counter == [60, 0, 0, 0]


def test_shard_range_partitions(rb):
    for total in (0, 1, 7, 4096, 1 << 23, 1000003):
        for world in (1, 2, 3, 8):
            for align in (1, 4, 128):
                got = []
                for r in range(world):
                    s, n = ctypes.c_int64(), ctypes.c_int64()
                    rb.lib.rbx_shard_range(total, world, r, align, ctypes.byref(s), ctypes.byref(n))
                    got.append((s.value, n.value))
                    assert s.value % align == 0 or s.value == total
                assert got[0][0] == 0 and sum(n for _, n in got) == total
                for (s0, n0), (s1, _) in zip(got, got[1:]):
                    assert s0 + n0 == s1


def test_argument_checks_raise_like_the_reference(rb):
    buf = torch.empty(16, dtype=torch.float32)
    D = rb.DenseDist(4, 4)
    with pytest.raises(rb.Error):            # dense_skops.hh:488-489
        rb.fill_dense(rb.Layout.ColMajor, D, 4, 4, 1, 0, buf, rb.RNGState(0))
    with pytest.raises(ValueError):          # dense_skops.hh:515-517
        rb.fill_dense(rb.DenseDist(4, 4, rb.DenseDistName.BlackBox), buf, rb.RNGState(0))
    S = rb.DenseSkOp(rb.DenseDist(3, 10), 0)
    A = torch.zeros(10 * 2)
    B = torch.zeros(3 * 2)
    with pytest.raises(rb.Error):            # skge.hh:200-201: lda >= rows_A
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 3, 2, 10, 1.0, S, 0, 0, A, 9, 0.0, B, 3)
    with pytest.raises(rb.Error):            # skge.hh:1104-1111: full-operator overload checks dims
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 3, 2, 9, 1.0, S, A, 10, 0.0, B, 3)
    Ss = rb.SparseSkOp(rb.SparseDist(3, 10, 4), 0, dtype=torch.float32)
    with pytest.raises(rb.Error):            # sparse_skops.hh:63: vec_nnz > dim_major
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 3, 2, 10, 1.0, Ss, 0, 0, A, 10, 0.0, B, 3)


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback(rb):
    buf = torch.empty(16, dtype=torch.float32)
    with pytest.raises(rb.RbxError) as e:
        rb.fill_dense(rb.DenseDist(4, 4), buf, rb.RNGState(0))
    assert e.value.code == -3
    S = rb.DenseSkOp(rb.DenseDist(3, 10), 0)
    with pytest.raises(rb.RbxError):
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 3, 2, 10, 1.0, S, 0, 0, torch.zeros(20), 10, 0.0, torch.zeros(6), 3)
