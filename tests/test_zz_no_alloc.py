"""rbx_sketch_vector_noalloc (include/rbx.h): the allocation-free sketch_vector, called through the C ABI and compared
with the existing entry point (rbx_sketch_dense_* with n = 1) on two calls in a row. Both run the same GEMV kernels in
the same chunking and summation order, so the outputs must be bit-identical (tolerance 0, tighter than the reference's
componentwise bound the other sketch tests use). Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [  # dtype, opS, d, m, offsets, family, major axis, incx, incy
    ("f32", "N", 300, 5000, (0, 0), "G", "L", 1, 1),
    ("f64", "N", 300, 5000, (3, 5), "U", "L", 2, 3),
    ("f32", "T", 290, 3000, (2, 7), "G", "L", 1, 2),
    ("f64", "N", 300, 2500, (1, 2), "G", "S", 1, 1),
    ("f32", "T", 130, 4099, (0, 1), "U", "S", 3, 1),
]


def _u32(v):
    return (ctypes.c_uint32 * len(v))(*v)


def _existing(rb, dt, opS, d, m, alpha, Dr, Dc, fam, ma, ctr, key, off, x, incx, beta, y, incy, stream):
    """sketch_vector as the package issues it today: the (RowMajor, n = 1) left sketch, skve.hh:152-176"""
    p0, p2 = (m, d) if opS == "T" else (d, m)
    sc = ctypes.c_float if dt == "f32" else ctypes.c_double
    return getattr(rb.lib, f"rbx_sketch_dense_{dt}")(b"L", b"R", opS.encode(), b"N", p0, 1, p2, sc(alpha), Dr, Dc, fam.encode(), ma.encode(),
                                                     ctr, key, ctypes.c_void_p(0), b"C", off[0], off[1], ctypes.c_void_p(x.data_ptr()), incx,
                                                     sc(beta), ctypes.c_void_p(y.data_ptr()), incy, stream)


def _noalloc(rb, dt, opS, d, m, alpha, Dr, Dc, fam, ma, ctr, key, off, x, incx, beta, y, incy, stream):
    return rb.lib.rbx_sketch_vector_noalloc(b"f" if dt == "f32" else b"d", opS.encode(), d, m, ctypes.c_double(alpha), Dr, Dc, fam.encode(),
                                            ma.encode(), ctr, key, off[0], off[1], ctypes.c_void_p(x.data_ptr()), incx,
                                            ctypes.c_double(beta), ctypes.c_void_p(y.data_ptr()), incy, stream)


@pytest.mark.parametrize("case", CASES, ids=[f"{c[0]}-{c[1]}-{c[5]}{c[6]}" for c in CASES])
def test_noalloc_matches_existing_entry_point_twice(rb, cuda, case):
    dt, opS, d, m, off, fam, ma, incx, incy = case
    tdt = torch.float32 if dt == "f32" else torch.float64
    Dr, Dc = d + off[0] + 1, m + off[1] + 2
    lx, ly = (m, d) if opS == "N" else (d, m)
    g = torch.Generator(device="cpu").manual_seed(d * 7 + m)
    ctr, key = _u32([0, 0, 0, 0]), _u32([5, 0])
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    # 0.5 and -0.25 are exact in float, so the float entry point sees the same alpha / beta through its double arguments
    for call, (alpha, beta) in enumerate([(0.5, -0.25), (-0.25, 0.0)]):
        x = torch.randn(lx * incx, generator=g, dtype=tdt).to(cuda)
        y0 = torch.randn(ly * incy, generator=g, dtype=tdt).to(cuda)
        y_old, y_new = y0.clone(), y0.clone()
        assert _existing(rb, dt, opS, d, m, alpha, Dr, Dc, fam, ma, ctr, key, off, x, incx, beta, y_old, incy, stream) == 0
        l0 = rb.lib.rbx_launch_count()
        assert _noalloc(rb, dt, opS, d, m, alpha, Dr, Dc, fam, ma, ctr, key, off, x, incx, beta, y_new, incy, stream) == 0
        assert rb.lib.rbx_launch_count() - l0 == 2          # partial sums + ordered reduce, no helper kernels
        torch.cuda.synchronize()
        a, b = y_old.cpu().numpy(), y_new.cpu().numpy()
        assert np.isfinite(a).all() and not np.array_equal(a, y0.cpu().numpy())
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"call {call}: outputs differ"


def test_noalloc_rejects_sizes_beyond_the_static_scratch(rb, cuda):
    """4 MiB of partial sums: ceil(K / 1024) * M * sizeof(T) with generator rows along the output (wide operator, Long axis)"""
    d, m = 8192, 1 << 20                    # 1024 chunks x 8192 rows x 4 bytes = 32 MiB
    x = torch.zeros(m, dtype=torch.float32, device=cuda)
    y = torch.full((d,), 7.0, dtype=torch.float32, device=cuda)
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    rc = _noalloc(rb, "f32", "N", d, m, 1.0, d, m, "U", "L", _u32([0, 0, 0, 0]), _u32([1, 0]), (0, 0), x, 1, 0.0, y, 1, stream)
    assert rc == -6 and b"capacity" in rb.lib.rbx_strerror(rc)
    torch.cuda.synchronize()
    assert bool((y == 7.0).all())           # rejected before anything was launched
    # argument errors are still the reference's: negative stride, operator window outside the distribution
    assert _noalloc(rb, "f32", "N", 8, 8, 1.0, 8, 8, "G", "L", _u32([0, 0, 0, 0]), _u32([1, 0]), (1, 0), x, 1, 0.0, y, 1, stream) == -1
    assert _noalloc(rb, "f32", "N", 8, 8, 1.0, 8, 8, "G", "L", _u32([0, 0, 0, 0]), _u32([1, 0]), (0, 0), x, 0, 0.0, y, 1, stream) == -1
