"""Out-of-bounds-write check of every kernel (compute-sanitizer is not available on the build pool):
with CGB_DEBUG_GUARDS=1 each device array of a context sits between two 256-byte guard bands that
cgb_solve() and the cgb_op_* calls read back before returning.  Shapes are the awkward ones: odd n,
ragged last CTA / warp / vector, n below one tile, every GEMV variant, both recurrences, the
symmetric product, text ingest, and the row-sharded forms of all of them."""
import os

import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cg():
    import conjugate_gradient_b200 as m
    assert m.device_count() >= 1
    os.environ["CGB_DEBUG_GUARDS"] = "1"
    m.lib().cgb_shutdown()                      # team contexts are re-created with the bands on
    yield m
    m.lib().cgb_shutdown()
    del os.environ["CGB_DEBUG_GUARDS"]


@pytest.fixture()
def ctx(cg):
    with cg.Context(0) as c:
        yield c


def residual(x, A, b):
    return np.linalg.norm(b - A @ x) / np.linalg.norm(b)


SHAPES = [1, 2, 3, 5, 31, 33, 127, 129, 255, 257, 300, 511, 513, 1000, 1023, 1025, 2047, 2049, 3000]


@pytest.mark.parametrize("n", SHAPES)
def test_solve_leaves_guard_bands_untouched(cg, ctx, n):
    for opts in ({}, dict(use_graph=0), dict(cg_variant=cg.CG_FUSED_REDUCTIONS), dict(gemv_variant=cg.GEMV_BULK),
                 dict(gemv_rows=8, gemv_unroll=2, gemv_threads=512), dict(gemv_vec=1), dict(gemv_vec=2), dict(guard=1)):
        ctx.generate(n, 3 + n)
        ctx.solve(**opts)                        # raises CgbError(-6) if a band was overwritten
        ctx.check_guards()
    A, b = ctx.op_generate(n, 3 + n)
    assert residual(ctx.read_x(), A, b) < 1e-6


@pytest.mark.parametrize("what,offset", [(3, 64), (3, -1), (5, 64 + 31), (2, 64), (1, -32)])
def test_guard_check_reports_an_overwrite(cg, what, offset):
    # the check itself: one element stored just outside an array is caught and named
    n = 64
    with cg.Context(0) as c:
        c.generate(n, 1)
        c.solve()
        c.check_guards()
        c.debug_poke(what, offset, 1.0)
        with pytest.raises(cg.CgbError) as e:
            c.check_guards()
        assert e.value.code == -6 and "was overwritten" in str(e.value)
        with pytest.raises(cg.CgbError):
            c.solve()                           # and every later solve reports it too


def test_guard_bands_are_off_by_default(cg):
    n = 64
    os.environ.pop("CGB_DEBUG_GUARDS")
    try:
        with cg.Context(0) as off:
            off.generate(n, 1)
            off.solve()
            with pytest.raises(cg.CgbError) as e:
                off.check_guards()
            assert "guard bands are off" in str(e.value)
    finally:
        os.environ["CGB_DEBUG_GUARDS"] = "1"


@pytest.mark.parametrize("n", [512, 1024, 1536, 2560])
def test_symmetric_product_and_uploads_leave_guard_bands_untouched(cg, ctx, n):
    A, b = O.spd_with_condition(n, 10.0, seed=n)
    for opts in (dict(symmetric=2), dict(symmetric=2, cg_variant=cg.CG_FUSED_REDUCTIONS), dict(symmetric=2, gemv_unroll=1), {}):
        x = np.zeros(n)
        ctx.conjugrad_rank(A, b, x, **opts)
        ctx.check_guards()
        assert residual(x, A, b) < 1e-6
    ctx.load_host(A, b, np.zeros(n))
    ctx.solve(symmetric=1)
    ctx.check_guards()


@pytest.mark.parametrize("m,n", [(1, 1), (3, 7), (5, 1000), (17, 513), (100, 2049), (33, 4096), (257, 255)])
def test_operators_leave_guard_bands_untouched(cg, ctx, m, n):
    rng = np.random.default_rng(m * n)
    A, v, w = rng.random((m, n)), rng.random(n), rng.random(m)
    for opts in ({}, dict(gemv_variant=cg.GEMV_BULK), dict(gemv_vec=1), dict(gemv_rows=8, gemv_unroll=2, gemv_threads=512)):
        y, d = ctx.op_matvec(A, v, w=w, **opts)          # the operator checks its own scratch arrays
        assert np.allclose(y, A @ v, rtol=1e-12)
    ctx.op_init_residual(w, y)
    ctx.op_xr_update(1.0, 2.0, w.copy(), w.copy(), w.copy(), y)
    ctx.op_p_update(1.0, 0.5, 1e-6, w.copy(), w.copy())
    ctx.check_guards()


def test_text_ingest_leaves_guard_bands_untouched(cg, ctx, tmp_path):
    for n in (1, 7, 129, 300):
        A, b = O.spd_with_condition(n, 5.0, seed=n) if n > 1 else (np.array([[2.0]]), np.array([1.0]))
        A = np.round(A, 4); A = 0.5 * (A + A.T); b = np.round(b, 4)
        fa, fb, fx = (str(tmp_path / f"{k}{n}.txt") for k in "Abx")
        np.savetxt(fa, A.ravel(), fmt="%.17g"); np.savetxt(fb, b, fmt="%.17g"); np.savetxt(fx, np.zeros(n), fmt="%.1f")
        ctx.load_text(n, fa, fb, fx)
        ctx.check_guards()
        ctx.solve()
        assert residual(ctx.read_x(), A, b) < 1e-6


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_sharded_solves_leave_guard_bands_untouched(cg, gpus):
    if cg.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    for per_rank in (1, 3, 33, 127, 250, 512):
        n = per_rank * gpus
        for opts in (dict(comm=cg.COMM_PEER), dict(comm=cg.COMM_NCCL), dict(comm=cg.COMM_PEER, cg_variant=cg.CG_FUSED_REDUCTIONS),
                     dict(comm=cg.COMM_PEER, use_graph=0)):
            x, st = cg.solve_generated(n, 9, gpus=gpus, **opts)      # every rank's solve checks its own bands
            assert st["final_residual"] < 1e-6
    for n in (512 * gpus, 1024 * gpus):
        for opts in (dict(symmetric=1), dict(symmetric=1, cg_variant=cg.CG_FUSED_REDUCTIONS)):
            x, st = cg.solve_generated(n, 9, gpus=gpus, **opts)
            assert st["gemv_variant"] == 3 and st["final_residual"] < 1e-6
        A, b = O.spd_with_condition(n, 10.0, seed=gpus)
        x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus, symmetric=2)
        assert residual(x, A, b) < 1e-6
