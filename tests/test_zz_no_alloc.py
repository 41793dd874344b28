"""cgb_matvec_noalloc: the matVec with the fused dot for callers that may not allocate device memory
(include/cgb200.h; no counterpart in the reference).  Its scratch is a
static device array shared by all calls, so: its output must equal the existing entry point's (cgb_op_matvec, which
cudaMallocs its scratch per call) on two calls IN A ROW with different shapes and data -- the second call finds the
first call's partials, ticket and flags in the array -- and sizes beyond the stated maximum must be refused.

This file is collected last on purpose (a failure here hides nothing after it), and
passes both alone in a fresh process and after the rest of the suite.  The test's own buffers are torch CUDA tensors;
the entry point itself allocates nothing."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cg():
    import conjugate_gradient_b200 as m
    assert m.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return m


def noalloc(cg, torch, A, v, w):
    """One call through the C ABI on device tensors; returns (y, dot) on the host."""
    from conjugate_gradient_b200 import _binding as B
    m, n = A.shape
    dA = torch.from_numpy(A).cuda()
    dv = torch.from_numpy(v).cuda()
    dy = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
    dw = torch.from_numpy(w).cuda() if w is not None else None
    ddot = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda") if w is not None else None
    torch.cuda.synchronize()
    rc = B.lib().cgb_matvec_noalloc(dA.data_ptr(), dv.data_ptr(), dw.data_ptr() if w is not None else None, dy.data_ptr(),
                                    ddot.data_ptr() if w is not None else None, m, n, None)       # legacy default stream
    assert rc == 0, B.lib().cgb_last_error().decode()
    torch.cuda.synchronize()
    return dy.cpu().numpy(), (float(ddot.cpu()[0]) if w is not None else None)


def test_two_calls_in_a_row_match_the_allocating_entry_point(cg):
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(20260)
    with cg.Context(0) as ctx:
        # two shapes with different grids and vector widths; the second is smaller, so stale partials of the first
        # lie beyond its own in the shared scratch array
        for (m, n) in ((3000, 4096), (517, 1026)):
            A = rng.standard_normal((m, n)); v = rng.standard_normal(n); w = rng.standard_normal(m)
            y_ref, dot_ref = ctx.op_matvec(A, v, w)                 # existing entry point: same kernel, cudaMalloc'ed scratch
            y, dot = noalloc(cg, torch, A, v, w)
            scale = np.abs(A) @ np.abs(v)
            assert np.all(np.abs(y - y_ref) <= 1e-15 * np.sqrt(n) * scale + 1e-300)      # the tolerance of test_matvec_and_fused_dot
            assert abs(dot - dot_ref) <= 1e-13 * (np.abs(w) @ np.abs(y_ref)) + 1e-300
            assert y.tobytes() == y_ref.tobytes() and dot == dot_ref                      # in fact the same bits: same plan, same order
            # and once more on the same data right away: the array keeps the last call's contents, the result must not change
            y2, dot2 = noalloc(cg, torch, A, v, w)
            assert y2.tobytes() == y.tobytes() and dot2 == dot
            y3, none = noalloc(cg, torch, A, v, None)                                    # without the fused dot (A * x0 at init)
            assert none is None and y3.tobytes() == y.tobytes()


def test_sizes_beyond_the_static_scratch_and_host_pointers_are_refused(cg):
    import torch
    from conjugate_gradient_b200 import _binding as B
    torch.cuda.set_device(0)
    L = B.lib()
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = d.data_ptr()
    too_many = 262144 + 1                                            # CGB_NOALLOC_MAX_ROWS + 1; refused before anything is read
    assert L.cgb_matvec_noalloc(p, p, None, p, None, too_many, 4, None) == -1
    assert "CGB_NOALLOC_MAX_ROWS" in L.cgb_last_error().decode()
    h = np.zeros(8)
    assert L.cgb_matvec_noalloc(h.ctypes.data, p, None, p, None, 2, 4, None) == -1       # a host pointer: it copies nothing
    assert "device memory" in L.cgb_last_error().decode()
    assert L.cgb_matvec_noalloc(p, p, p, p, None, 2, 4, None) == -1                       # w without a place for the dot
    # and a good call still works afterwards
    A = np.arange(8.0).reshape(2, 4); v = np.ones(4)
    y, _ = noalloc(cg, torch, A, v, None)
    assert y.tolist() == [6.0, 22.0]
