"""conjugate_gradient_b200 -- Python host-side mirror of the C ABI in include/cgb200.h.

The product is libcgb200.so (hand-written sm_100a CUDA kernels + a C host
solver) and the cg_b200 command line; this module only binds that library with
ctypes for the tests and bench.py.  There is no Python or CPU implementation
behind it: importing works without a GPU (so the symbol table can be checked),
but every computing call needs a B200 and raises CgbError otherwise, and a
missing library raises ImportError instead of falling back to anything.
"""
from ._binding import (  # noqa: F401
    CG_CLASSIC, CG_FUSED_REDUCTIONS, CGB_EPSILON, COMM_AUTO, COMM_NCCL, COMM_PEER, GEMV_AUTO, GEMV_BULK, GEMV_LDG, GEN_GRAM, GEN_KAPPA_LOG2, GEN_ROUND4, GEN_SPECTRUM, HEADER_SYMBOLS, LIB_PATH,
    CgbError, Context, Options, Stats, build, conjugrad_files, conjugrad_host, device_count, lib,
    read_dimensions, read_values, solve_generated, unique_id, write_values,
)
