"""randblas_b200: the RandBLAS sketching hot path on B200 (sm_100a) behind RandBLAS's own interface.

Importing this package loads randblas_b200/librbx.so (hand-written CUDA kernels + the C ABI of
include/rbx.h). There is no CPU implementation and no fallback: a missing library raises ImportError.
"""
from ._capi import lib, PROTOS, RbxError          # noqa: F401
from .api import *                                 # noqa: F401,F403
from .api import Error                             # noqa: F401
