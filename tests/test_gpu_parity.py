"""GPU parity tests: the CUDA path, called through the C ABI (libcgb200.so),
against the CPU oracle and the golden vectors of the real reference.

Tolerances (north star): relative solution error <= 1e-10 against the FP64
reference's own final iterate, iteration count within +-1, final residual below
the reference's tolerance (1e-6).  Integer / bit-level items (generator,
element-wise updates with the reference's rounding points) are bit-exact.
"""
import numpy as np
import pytest

import _oracle as O
from conftest import unhex

pytestmark = pytest.mark.gpu
EPS = 1.0e-6
RTOL_X = 1.0e-10


@pytest.fixture(scope="module")
def cg():
    import conjugate_gradient_b200 as m
    assert m.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return m


@pytest.fixture(scope="module")
def ctx(cg):
    with cg.Context(0) as c:
        yield c


def relerr(x, ref):
    return np.linalg.norm(x - ref) / max(np.linalg.norm(ref), 1e-300)


def case_inputs(c):
    n = c["n"]
    if "A" in c:
        return unhex(c["A"]).reshape(n, n), unhex(c["b"]), unhex(c["x0"])
    x0 = O.gen_b(n, c["x0_seed"]) if "x0_seed" in c else np.zeros(n)
    return O.gen_A(n, c["seed"], c["flags"]), O.gen_b(n, c["seed"], c["flags"]), x0


# ----------------------------------------------------------------- generator
@pytest.mark.parametrize("n,row0,m,flags", [(8, 0, 8, 0), (300, 0, 300, 0), (1024, 512, 256, 0),
                                            (513, 7, 100, 0), (256, 0, 256, 1),
                                            (300, 0, 300, 7 << 8), (1024, 1000, 24, (10 << 8) | 1), (1, 0, 1, 5 << 8)])
def test_generator_bit_identical_to_cpu_twin(ctx, golden, n, row0, m, flags):
    seed = golden["seed"]
    A, b = ctx.op_generate(n, seed, flags, row0, m)
    assert A.tobytes() == O.gen_A(n, seed, flags, row0, m).tobytes()
    assert b.tobytes() == O.gen_b(n, seed, flags, row0, m).tobytes()
    if n == 8 and flags == 0:
        assert A.ravel().tobytes() == unhex(golden["generator"]["A8"]).tobytes()


def test_generator_large_indices(ctx):
    # 64-bit index arithmetic: element indices beyond 2^32 (n = 70000 -> n*n = 4.9e9)
    n, seed = 70000, 11
    A, b = ctx.op_generate(n, seed, 0, row0=69990, m=4)
    assert A.tobytes() == O.gen_A(n, seed, 0, 69990, 4).tobytes()
    assert b.tobytes() == O.gen_b(n, seed, 0, 69990, 4).tobytes()


# ----------------------------------------------------------------- single kernels
GEMV_CFGS = [dict(), dict(gemv_variant=1, gemv_vec=2), dict(gemv_variant=1, gemv_rows=8, gemv_unroll=1, gemv_threads=128),
             dict(gemv_variant=1, gemv_rows=2, gemv_unroll=4, gemv_threads=256),
             dict(gemv_variant=2), dict(gemv_variant=2, gemv_rows=4, gemv_stages=4, gemv_tile_cols=512),
             dict(gemv_variant=2, gemv_rows=8, gemv_stages=3, gemv_tile_cols=1024, gemv_threads=256)]


@pytest.mark.parametrize("m,n", [(1, 1), (2, 2), (4, 4), (3, 5), (7, 33), (64, 1000), (130, 1026),
                                 (257, 2048), (512, 4096), (100, 8192), (2048, 2048)])
@pytest.mark.parametrize("cfg", range(len(GEMV_CFGS)))
def test_matvec_and_fused_dot(ctx, m, n, cfg):
    rng = np.random.default_rng(m * 7919 + n)
    A = rng.standard_normal((m, n)); v = rng.standard_normal(n); w = rng.standard_normal(m)
    y, dot = ctx.op_matvec(A, v, w, **GEMV_CFGS[cfg])
    y_ref = O.matvec(A, v)                                   # parallel_cg.c:172-184 order
    scale = np.abs(A) @ np.abs(v)                            # rounding-error scale per row
    assert np.all(np.abs(y - y_ref) <= 1e-15 * np.sqrt(n) * scale + 1e-300)
    dot_ref = O.vecvec(w, y_ref)
    assert abs(dot - dot_ref) <= 1e-13 * (np.abs(w) @ np.abs(y_ref)) + 1e-300
    y2 = ctx.op_matvec(A, v, None, **GEMV_CFGS[cfg])         # w == NULL path (A*x0 at init)
    assert y2.tobytes() == y.tobytes()                       # deterministic, same order


def test_matvec_is_run_to_run_deterministic(ctx):
    rng = np.random.default_rng(5)
    A = rng.standard_normal((1000, 4096)); v = rng.standard_normal(4096); w = rng.standard_normal(1000)
    for cfg in (dict(), dict(gemv_variant=2)):
        outs = [ctx.op_matvec(A, v, w, **cfg) for _ in range(3)]
        assert all(o[0].tobytes() == outs[0][0].tobytes() and o[1] == outs[0][1] for o in outs)


@pytest.mark.parametrize("m", [1, 2, 5, 255, 256, 257, 4096, 100003])
def test_vector_kernels_keep_reference_rounding_points(ctx, m):
    rng = np.random.default_rng(m)
    b = rng.standard_normal(m); Ax = rng.standard_normal(m)
    r, p, rr = ctx.op_init_residual(b, Ax)                   # parallel_cg.c:284-286
    assert r.tobytes() == (b - Ax).tobytes() and p.tobytes() == r.tobytes()
    assert abs(rr - O.vecvec(r, r)) <= 1e-13 * rr
    x = rng.standard_normal(m); Ap = rng.standard_normal(m)
    rho, pAp = 3.25, 7.5
    alpha = rho / pAp
    x2, r2, rr2 = ctx.op_xr_update(rho, pAp, x, r, p, Ap)    # parallel_cg.c:298-309
    assert x2.tobytes() == (x + p * alpha).tobytes()         # product rounded, then add: no FMA
    assert r2.tobytes() == (r - Ap * alpha).tobytes()
    assert abs(rr2 - O.vecvec(r2, r2)) <= 1e-13 * rr2
    p2, stopped = ctx.op_p_update(rho, 1.5, EPS, p, r2)      # parallel_cg.c:314-322
    assert not stopped and p2.tobytes() == (r2 + p * (1.5 / rho)).tobytes()
    p3, stopped = ctx.op_p_update(rho, 0.9e-12, EPS, p, r2)  # sqrt(beta) < eps: break, p untouched
    assert stopped and p3.tobytes() == p.tobytes()
    p4, stopped = ctx.op_p_update(rho, 1.0e-12, EPS, p, r2)  # sqrt(beta) == eps is NOT < eps
    assert not stopped


# ----------------------------------------------------------------- whole solves
@pytest.mark.parametrize("tag", ["g1", "g2", "g3", "g4"])
@pytest.mark.parametrize("variant", [1, 2])
def test_solve_matches_reference_golden(cg, golden, tag, variant):
    c = golden["cases"][tag]
    A, b, x0 = case_inputs(c)
    x, st = cg.conjugrad_host(A, b, x0, gpus=1, gemv_variant=variant)
    want = unhex(c["x"]["serial_f64"])                       # the REAL reference, float->double build
    xo, so = O.conjugrad(A, b, x0)
    assert relerr(x, want) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1
    assert st["final_residual"] < EPS
    # FP32 reference as shipped: loose cross-check
    assert relerr(x, unhex(c["x"]["serial_f32"])) <= 2e-5


def test_known_answers_of_shipped_fixtures(cg, golden):
    A, b, x0 = case_inputs(golden["cases"]["g1"])
    x, st = cg.conjugrad_host(A, b, x0)
    np.testing.assert_allclose(x, [2 / 3, 1 / 3], rtol=1e-14)
    assert st["break_k"] == 1 and st["iterations"] == 2
    A, b, x0 = case_inputs(golden["cases"]["g2"])
    x, st = cg.conjugrad_host(A, b, x0)
    np.testing.assert_allclose(x, [-1, 1, -1, 1], rtol=1e-12)
    assert st["break_k"] == 3


@pytest.mark.parametrize("n,graph,chunk", [(96, 1, 0), (1000, 1, 2), (1001, 1, 8), (1536, 0, 0), (4096, 1, 0)])
def test_generated_solve_vs_oracle(ctx, n, graph, chunk):
    seed = 1234 + n
    ctx.generate(n, seed)
    st = ctx.solve(use_graph=graph, check_every=chunk)
    x = ctx.read_x()
    xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n))
    assert relerr(x, xo) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["break_k"] == st["iterations"] - 1
    assert st["final_residual"] < EPS
    assert abs(st["rsold0"] - so["rsold0"]) <= 1e-12 * so["rsold0"]


@pytest.mark.parametrize("n,kappa,variant", [(512, 30.0, 1), (512, 30.0, 2), (1000, 30.0, 1)])
def test_long_running_solve_on_ill_conditioned_matrix(ctx, n, kappa, variant):
    # ~47 iterations (generateSPDmatrix-style systems stop after ~5): the device-side stop
    # flag, the rr[] ping-pong and the graph chunking over many replays.  kappa = 30 because
    # the 1e-10 bar is only meaningful where the reference agrees with itself: at kappa = 30
    # its serial and SPMD summation orders differ by ~1e-15 in x, at kappa = 100 (86
    # iterations) already by 1e-10..2e-10 (measured with the oracle, both orders).
    A, b = O.spd_with_condition(n, kappa, seed=n)
    ctx.load_host(A, b, np.zeros(n))
    st = ctx.solve(gemv_variant=variant)
    xo, so = O.conjugrad(A, b, np.zeros(n))
    assert so["iterations"] > 40
    assert relerr(ctx.read_x(), xo) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS


@pytest.mark.parametrize("n,k", [(1024, 4), (2048, 7)])
def test_generated_spread_family_solve(cg, ctx, n, k):
    # the on-device generator's harder family (cond ~ 2^k): tens of iterations, no host matrix.
    # (diagonally dominant: the reference's serial and SPMD orders agree to ~4e-16 even at k = 7)
    seed, flags = 17, cg.GEN_KAPPA_LOG2(k)
    ctx.generate(n, seed, flags)
    st = ctx.solve()
    xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n))
    assert so["iterations"] > 25
    assert relerr(ctx.read_x(), xo) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS


# ----------------------------------------------------------------- fused-reductions CG
@pytest.mark.parametrize("n,k,graph,sym", [(512, 0, 1, 0), (1000, 0, 0, 0), (2048, 7, 1, 0), (1024, 4, 1, 1), (4096, 0, 1, 0)])
def test_fused_reductions_variant_matches_reference_and_its_own_oracle(cg, ctx, n, k, graph, sym):
    seed, flags = 70 + n, cg.GEN_KAPPA_LOG2(k)
    ctx.generate(n, seed, flags)
    st = ctx.solve(cg_variant=cg.CG_FUSED_REDUCTIONS, use_graph=graph, symmetric=sym)
    x = ctx.read_x()
    A, b = O.gen_A(n, seed, flags), O.gen_b(n, seed, flags)
    xc, sc = O.conjugrad(A, b, np.zeros(n))                  # the reference's recurrence
    xf, sf = O.cg_fused(A, b, np.zeros(n))                   # the variant's CPU statement
    assert st["cg_variant"] == 1 and st["exit_reason"] == 1
    assert relerr(x, xc) <= RTOL_X and relerr(x, xf) <= RTOL_X
    assert abs(st["iterations"] - sc["iterations"]) <= 1 and abs(st["iterations"] - sf["iterations"]) <= 1
    assert st["break_k"] == st["iterations"] - 1 and st["final_residual"] < EPS
    # fixed-iteration and max_iter modes count like the classic path
    ctx.set_x0(np.zeros(n))
    st = ctx.solve(cg_variant=1, ignore_stop=1, max_iter=3, use_graph=graph)
    x3, _ = O.cg_fused(A, b, np.zeros(n), max_iter=3, ignore_stop=True)
    assert st["iterations"] == 3 and st["break_k"] == -1 and relerr(ctx.read_x(), x3) <= RTOL_X
    ctx.set_x0(np.zeros(n))
    st = ctx.solve(cg_variant=1, max_iter=2, eps=1e-300, use_graph=graph)
    assert st["iterations"] == 2 and st["exit_reason"] == 2


def test_fused_reductions_variant_guards_and_shipped_fixture(cg, ctx, golden):
    c = golden["cases"]["g2"]
    A, b, x0 = unhex(c["A"]).reshape(4, 4), unhex(c["b"]), unhex(c["x0"])
    ctx.load_host(A, b, x0)
    st = ctx.solve(cg_variant=1)
    np.testing.assert_allclose(ctx.read_x(), [-1, 1, -1, 1], rtol=1e-12)
    ctx.load_host(np.zeros((4, 4)), np.ones(4), np.zeros(4))
    st = ctx.solve(cg_variant=1, guard=1)
    assert st["iterations"] == 0 and st["exit_reason"] == 3
    ctx.load_host(np.eye(2) * 2, np.array([2.0, 4.0]), np.array([1.0, 2.0]))
    st = ctx.solve(cg_variant=1, guard=1)
    assert st["iterations"] == 0 and st["exit_reason"] == 4


# ----------------------------------------------------------------- symmetric product
@pytest.mark.parametrize("n,k,ru", [(512, 0, 0), (1024, 0, 1), (2048, 7, 2), (1536, 4, 4)])
def test_symmetric_upper_tile_product_solves_like_the_reference(cg, ctx, n, k, ru):
    # matVec from the upper 512 x 512 tiles only (half the bytes): same solution, same stop
    seed, flags = 41 + n, cg.GEN_KAPPA_LOG2(k)
    ctx.generate(n, seed, flags)
    st = ctx.solve(symmetric=1, gemv_unroll=ru)
    x = ctx.read_x()
    xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n))
    assert st["gemv_variant"] == 3 and relerr(x, xo) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS
    assert st["gemv_bytes"] == 8.0 * (0.5 * n * (n + 1) + 2 * n)
    # bitwise reproducible (fixed-order partial sums, no atomics)
    ctx.set_x0(np.zeros(n))
    ctx.solve(symmetric=1, gemv_unroll=ru)
    np.testing.assert_array_equal(ctx.read_x(), x)


def test_symmetric_product_matches_the_full_product_elementwise(ctx):
    n, seed = 2048, 3
    ctx.generate(n, seed)
    ctx.solve(ignore_stop=1, max_iter=1)
    ap_full = ctx.debug_read(5, 0, n)
    ctx.set_x0(np.zeros(n))
    ctx.solve(ignore_stop=1, max_iter=1, symmetric=1)
    ap_sym = ctx.debug_read(5, 0, n)
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    want = A @ b                                             # first direction p0 = b - A 0 = b
    assert np.abs(ap_sym - want).max() <= 1e-13 * np.abs(want).max()
    assert np.abs(ap_sym - ap_full).max() <= 1e-13 * np.abs(want).max()


def test_symmetric_option_checks_its_preconditions(cg, ctx):
    n = 512
    A, b = O.spd_with_condition(n, 30.0, seed=2)
    ctx.load_host(A, b, np.zeros(n))                         # host matrix: verified on the device first
    st = ctx.solve(symmetric=1)
    xo, so = O.conjugrad(A, b, np.zeros(n))
    assert st["gemv_variant"] == 3 and relerr(ctx.read_x(), xo) <= RTOL_X
    A2 = A.copy(); A2[3, 400] = np.nextafter(A2[3, 400], 1e9)       # one ulp off in one element
    ctx.load_host(A2, b, np.zeros(n))
    with pytest.raises(cg.CgbError) as e:
        ctx.solve(symmetric=1)
    assert e.value.code == -1 and "not exactly symmetric" in str(e.value)
    assert ctx.solve()["gemv_variant"] != 3                  # the full product still serves it
    ctx.generate(1000, 1)
    with pytest.raises(cg.CgbError) as e:
        ctx.solve(symmetric=1)
    assert e.value.code == -1 and "multiple of 512" in str(e.value)


def test_asserted_symmetric_upload_sends_only_the_tiles_the_product_reads(cg, ctx):
    # cgb_options.symmetric = 2 through the reference-facing call: tiles below the tile diagonal
    # never leave the host (poisoned here with NaN to prove it), h2d bytes = n (n + 512) / 2 + 2n
    n, T = 2048, 512
    A, b = O.spd_with_condition(n, 30.0, seed=5)
    xo, so = O.conjugrad(A, b, np.zeros(n))
    P = A.copy()
    for I in range(n // T):
        P[I * T:(I + 1) * T, :I * T] = np.nan
    x = np.zeros(n)
    st = ctx.conjugrad_rank(P, b, x, symmetric=2)
    assert st["gemv_variant"] == 3 and relerr(x, xo) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS
    assert st["h2d_bytes"] == 8 * (n * (n + T) // 2 + 2 * n)
    with pytest.raises(cg.CgbError) as e:                    # the full product has nothing to read
        ctx.solve()
    assert e.value.code == -6 and "cgb_load_host_symmetric" in str(e.value)
    ctx.set_x0(np.zeros(n))
    assert ctx.solve(symmetric=1, cg_variant=cg.CG_FUSED_REDUCTIONS)["gemv_variant"] == 3
    assert relerr(ctx.read_x(), xo) <= RTOL_X
    # shapes the symmetric product does not take fall back to the full upload and the full product
    n2 = 700
    A2, b2 = O.spd_with_condition(n2, 10.0, seed=6)
    x2 = np.zeros(n2)
    with pytest.raises(cg.CgbError):
        ctx.conjugrad_rank(A2, b2, x2, symmetric=2)
    with pytest.raises(cg.CgbError) as e:
        ctx.load_host_symmetric(A2, b2, x2)
    assert e.value.code == -1 and "multiple of 512" in str(e.value)
    ctx.load_host(A, b, np.zeros(n))                         # a full load clears the restriction
    assert st["h2d_bytes"] < ctx.solve()["h2d_bytes"] == 8 * (n * n + 2 * n)


def test_nonzero_initial_guess_and_restart(ctx):
    n, seed = 768, 9
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    x0 = np.random.default_rng(1).standard_normal(n)
    ctx.load_host(A, b, x0)
    st = ctx.solve()
    x = ctx.read_x()
    xo, so = O.conjugrad(A, b, x0)
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    # the solution fed back as initial guess (the reference's restart input): already converged
    ctx.set_x0(x)
    st2 = ctx.solve()
    xo2, so2 = O.conjugrad(A, b, x)
    assert st2["iterations"] <= 2 and abs(st2["iterations"] - so2["iterations"]) <= 1
    assert relerr(ctx.read_x(), xo2) <= RTOL_X


def test_fixed_iteration_mode_matches_oracle_iterates(ctx):
    # bench mode: stop rule evaluated, not acted on; compare at equal iteration counts
    n, seed, K = 512, 77, 3
    ctx.generate(n, seed)
    st = ctx.solve(ignore_stop=1, max_iter=K)
    xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n), max_iter=K, ignore_stop=True)
    assert st["iterations"] == K == so["iterations"] and st["break_k"] == -1
    assert relerr(ctx.read_x(), xo) <= RTOL_X
    assert abs(st["final_beta"] - so["final_beta"]) <= 1e-6 * so["final_beta"] + 1e-30


def test_max_iter_bound_and_odd_counts(ctx):
    n, seed = 400, 3
    ctx.generate(n, seed)
    for K in (1, 2, 3):
        st = ctx.solve(max_iter=K, eps=1e-300)
        xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n), eps=1e-300, max_iter=K)
        assert st["iterations"] == K and st["break_k"] == -1
        assert relerr(ctx.read_x(), xo) <= RTOL_X
        ctx.set_x0(np.zeros(n))


def test_exact_initial_guess_mirrors_reference_nan_behaviour(ctx):
    # x0 already exact: rho = 0 -> alpha = 0/0; the reference has no guard and
    # iterates NaNs to the loop bound; mirrored, bounded here.
    A = np.array([[2.0, 0.0], [0.0, 2.0]]); b = np.array([2.0, 4.0]); x0 = np.array([1.0, 2.0])
    ctx.load_host(A, b, x0)
    st = ctx.solve()
    xo, so = O.conjugrad(A, b, x0)
    assert st["iterations"] == so["iterations"] == 2 and st["break_k"] == -1
    assert np.isnan(ctx.read_x()).all() and np.isnan(xo).all()


def test_guard_option_ends_the_solve_where_the_reference_iterates_nans(cg, ctx):
    # not in the reference: opt-in guards, default stays the mirror above
    A = np.array([[2.0, 0.0], [0.0, 2.0]]); b = np.array([2.0, 4.0]); x0 = np.array([1.0, 2.0])
    ctx.load_host(A, b, x0)
    st = ctx.solve(guard=1)
    assert st["iterations"] == 0 and st["exit_reason"] == 4          # CGB_EXIT_EXACT_X0
    np.testing.assert_array_equal(ctx.read_x(), x0)
    # p.Ap = 0 (A = 0 is not SPD; it is the cheapest way to hit the division): alpha = inf
    ctx.load_host(np.zeros((4, 4)), np.ones(4), np.zeros(4))
    st = ctx.solve(guard=1)
    assert st["iterations"] == 0 and st["exit_reason"] == 3          # CGB_EXIT_BREAKDOWN
    np.testing.assert_array_equal(ctx.read_x(), np.zeros(4))
    # ordinary exits are labelled too, and the guard changes nothing on a healthy system
    n, seed = 512, 5
    ctx.generate(n, seed)
    st_plain = ctx.solve()
    x_plain = ctx.read_x()
    ctx.set_x0(np.zeros(n))
    st_guard = ctx.solve(guard=1)
    assert st_plain["exit_reason"] == st_guard["exit_reason"] == 1   # CGB_EXIT_STOP_RULE
    np.testing.assert_array_equal(ctx.read_x(), x_plain)
    ctx.set_x0(np.zeros(n))
    assert ctx.solve(max_iter=2, eps=1e-300)["exit_reason"] == 2     # CGB_EXIT_MAX_ITER


def test_errors(cg, ctx):
    with pytest.raises(cg.CgbError) as e:
        cg.conjugrad_host(np.eye(6), np.ones(6), np.zeros(6), gpus=4)
    assert e.value.code in (-4, -2)          # "6 is not divisible by 4" (or < 4 GPUs on the box)
    with cg.Context(0) as c2:
        with pytest.raises(cg.CgbError) as e:
            c2.solve()
        assert e.value.code == -6


# ----------------------------------------------------------------- full-size properties
def test_full_size_gemv_properties(ctx):
    """n = 16384 (config 3): linearity and symmetry checks that need no oracle pass,
    plus an oracle check on a sample of rows."""
    n, seed = 16384, 20170209
    ctx.generate(n, seed)
    st = ctx.solve()
    x = ctx.read_x()
    assert st["final_residual"] < EPS and 2 <= st["iterations"] <= 8
    # residual recomputed on the CPU for sampled rows: b_i - A_i x small
    rows = [0, 1, 777, 8191, 16383]
    for i in rows:
        Ai = O.gen_A(n, seed, 0, i, 1)[0]
        assert abs(O.gen_b(n, seed, 0, i, 1)[0] - Ai @ x) < 1e-6
    # oracle iterate at the same iteration count on this size takes ~1 s per matVec
    xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n))
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1


def test_baseline_size_vs_oracle_65536(ctx):
    """BASELINE.json configs[3] (n = 65536, 34 GB), the size the metric is quoted on, against the CPU oracle itself:
    the threaded row-block form of the port (parallel_cg.c:283-323, bit-identical to the real reference wherever the
    real reference can run) regenerates the same matrix from the seed and streams it ~0.3 s per iteration on 16
    cores.  Converged solve: relative error of x <= 1e-10, iteration count within +-1, residual below eps; and x
    after 10 fixed iterations (the bench's mode), where the GEMV's summation order over 65536 columns has had ten
    products to show.  The symmetric product and bitwise reproducibility are checked on the same resident matrix."""
    n, seed = 65536, 20170209
    info = ctx.device_info()
    if info["total_bytes"] < 8 * n * n + (4 << 30):
        pytest.skip("GPU memory too small for this size")
    import os
    threads = 1
    while threads * 2 <= (os.cpu_count() or 1) and threads < 64:
        threads *= 2
    ctx.generate(n, seed)
    st = ctx.solve()
    x = ctx.read_x()
    xo, so = O.team_solve_generated(n, seed, threads)
    assert st["final_residual"] < EPS and st["exit_reason"] == 1
    assert abs(st["iterations"] - so["iterations"]) <= 1
    if st["iterations"] == so["iterations"]:
        assert relerr(x, xo) <= RTOL_X, relerr(x, xo)
        assert np.sqrt(so["final_beta"]) < EPS
    ctx.set_x0(np.zeros(n))
    st10 = ctx.solve(ignore_stop=1, max_iter=10)
    x10 = ctx.read_x()
    xo10, so10 = O.team_solve_generated(n, seed, threads, ignore_stop=True, max_iter=10)
    assert relerr(x10, xo10) <= RTOL_X, relerr(x10, xo10)
    assert st10["first_hit_k"] == so10["first_hit_k"]
    # the symmetric product on the same matrix: same solution
    ctx.set_x0(np.zeros(n))
    sts = ctx.solve(symmetric=1)
    assert sts["gemv_variant"] == 3 and abs(sts["iterations"] - so["iterations"]) <= 1
    assert relerr(ctx.read_x(), xo) <= RTOL_X
    # run-to-run bitwise reproducibility of the fixed-iteration mode
    ctx.set_x0(np.zeros(n))
    ctx.solve(ignore_stop=1, max_iter=10)
    np.testing.assert_array_equal(ctx.read_x(), x10)
    ctx.generate(512, 1)                      # release the big block for the tests that follow


def test_baseline_size_131072_through_size_independent_properties(ctx):
    """BASELINE.json configs[4] (137 GB): the oracle would need 137 GB of host memory and ~6 s per product on the
    bench box's 16 cores (196 GB of RAM there, but not on every box), so this size -- and only this one -- is held
    to size-independent properties: the solve converges in the iteration count the family always needs; b - A x is
    small on sampled rows recomputed on the CPU from the generator twin; the symmetric product gives the same
    solution; a fixed-iteration run reproduces itself bit for bit.  n = 65536 has the oracle pass above."""
    n = 131072
    info = ctx.device_info()
    if info["total_bytes"] < 8 * n * n + (4 << 30):
        pytest.skip("GPU memory too small for this size")
    seed = 20170209
    ctx.generate(n, seed)
    st = ctx.solve()
    x = ctx.read_x()
    assert st["final_residual"] < EPS and st["exit_reason"] == 1 and 2 <= st["iterations"] <= 8
    for i in (0, 1, 4097, n // 2 + 3, n - 1):
        Ai = O.gen_A(n, seed, 0, i, 1)[0]
        assert abs(O.gen_b(n, seed, 0, i, 1)[0] - Ai @ x) < 1e-6
    ctx.set_x0(np.zeros(n))
    sts = ctx.solve(symmetric=1)
    assert sts["gemv_variant"] == 3 and abs(sts["iterations"] - st["iterations"]) <= 1
    assert relerr(ctx.read_x(), x) <= RTOL_X
    ctx.set_x0(np.zeros(n))
    ctx.solve(ignore_stop=1, max_iter=3)
    x3 = ctx.read_x()
    ctx.set_x0(np.zeros(n))
    ctx.solve(ignore_stop=1, max_iter=3)
    np.testing.assert_array_equal(ctx.read_x(), x3)
    ctx.generate(512, 1)                      # release the big block for the tests that follow


def test_cli_on_shipped_fixture(cg, golden_dir):
    """The reference CLI surface: prog A.txt b.txt x0.txt + dimensions.txt; stdout lines of
    serialConjugate.c:58,249-251 and printer()'s format behind --print-x."""
    import subprocess
    from conjugate_gradient_b200 import _binding as B
    g = golden_dir
    r = subprocess.run([B.CLI_PATH, f"{g}/g1_matrixA.txt", f"{g}/g1_vectorb.txt", f"{g}/g1_x0.txt",
                        "--dims", f"{g}/g1_dimensions.txt", "--print-x", "--stats"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "Computing cg of matrix size : 4"
    assert lines[1] == "0.666667  " and lines[2] == "0.333333  "
    assert lines[3].startswith("average clock execution time in seconds: ")
    assert "iterations: 2" in lines[4]


def test_cli_generated_and_restart_file(cg, tmp_path):
    import subprocess
    from conjugate_gradient_b200 import _binding as B
    xf = tmp_path / "x.txt"
    r = subprocess.run([B.CLI_PATH, "--gen", "1024", "--seed", "5", "--write-x", str(xf), "--stats"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    x = cg.read_values(str(xf), 1024)
    xo, so = O.conjugrad(O.gen_A(1024, 5), O.gen_b(1024, 5), np.zeros(1024))
    assert relerr(x, xo) <= RTOL_X
    assert f"iterations: {so['iterations']}" in r.stdout


def test_cli_emitted_files_run_record_and_restart(cg, tmp_path):
    # the generator emits generateSPDmatrix.m-format files; the three-file
    # CLI solves them (parallel text ingest), writes a JSON run record and a restart vector
    import json
    import subprocess
    from conjugate_gradient_b200 import _binding as B
    n, seed, pre = 600, 8, str(tmp_path / "sys")
    r = subprocess.run([B.CLI_PATH, "--gen", str(n), "--seed", str(seed), "--emit-files", pre],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    A = O.gen_A(n, seed, O.GEN_ROUND4); b = O.gen_b(n, seed, O.GEN_ROUND4)
    assert open(pre + "_matrix.txt").readline() == "%.4f\n" % A[0, 0]
    np.testing.assert_array_equal(cg.read_values(pre + "_matrix.txt", n * n), A.ravel())
    np.testing.assert_array_equal(cg.read_values(pre + "_vector.txt", n), b)
    assert open(pre + "_x0.txt").readline() == "0.0\n" and cg.read_dimensions(pre + "_dimensions.txt") == (n, n, n, 1)
    xf, rec = str(tmp_path / "x.txt"), str(tmp_path / "run.json")
    files = [pre + "_matrix.txt", pre + "_vector.txt", pre + "_x0.txt", "--dims", pre + "_dimensions.txt"]
    r = subprocess.run([B.CLI_PATH] + files + ["--write-x", xf, "--run-record", rec],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    xo, so = O.conjugrad(A, b, np.zeros(n))
    assert relerr(cg.read_values(xf, n), xo) <= RTOL_X
    d = json.load(open(rec))
    assert d["n"] == n and d["exit"] == "stop_rule" and abs(d["iterations"] - so["iterations"]) <= 1
    assert d["final_residual"] < EPS and d["source"] == "files" and d["iterations_per_second"] > 0
    # restart: the written solution as initialguess.txt
    r = subprocess.run([B.CLI_PATH, files[0], files[1], xf, "--dims", files[4], "--run-record", rec],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and json.load(open(rec))["iterations"] <= 2


# ----------------------------------------------------------------- further generator families
@pytest.mark.parametrize("n,row0,m,k,r4", [(64, 0, 64, 4, 0), (300, 0, 300, 10, 0), (1024, 512, 256, 7, 0), (513, 7, 100, 3, 1), (2, 0, 2, 1, 0)])
def test_prescribed_spectrum_generator_bit_identical_to_cpu_twin(cg, ctx, n, row0, m, k, r4):
    flags = cg.GEN_SPECTRUM | cg.GEN_KAPPA_LOG2(k) | r4
    A, b = ctx.op_generate(n, 77, flags, row0, m)
    assert A.tobytes() == O.gen_A(n, 77, flags, row0, m).tobytes()
    assert b.tobytes() == O.gen_b(n, 77, flags, row0, m).tobytes()


@pytest.mark.parametrize("n,row0,m,k", [(64, 0, 64, 0), (200, 0, 200, 5), (1000, 0, 1000, 0), (1024, 256, 300, 3), (77, 10, 50, 0)])
def test_gram_generator_matches_cpu_twin_and_is_exactly_symmetric(cg, ctx, n, row0, m, k):
    """A = R R' + n diag(d) from the FP64 tensor pipe (cg_gram.cu) against the CPU twin's plain loops: <= 1e-13
    relative per entry (the instruction's internal summation order is the hardware's); bitwise symmetric."""
    flags = cg.GEN_GRAM | cg.GEN_KAPPA_LOG2(k)
    A, b = ctx.op_generate(n, 5, flags, row0, m)
    want = O.gen_A(n, 5, flags, row0, m)
    assert np.max(np.abs(A - want) / np.abs(want)) <= 1e-13
    assert b.tobytes() == O.gen_b(n, 5, flags, row0, m).tobytes()
    if m == n:
        assert np.array_equal(A, A.T)
    else:                                          # a row block: its square part against the transposed block of the other rows
        full, _ = ctx.op_generate(n, 5, flags, 0, n)
        assert np.array_equal(full, full.T) and full[row0:row0 + m].tobytes() == A.tobytes()


@pytest.mark.parametrize("family,k,lo", [("spectrum", 7, 60), ("spectrum", 10, 150), ("gram", 0, 4), ("gram", 3, 10)])
def test_solves_on_the_further_families_match_the_oracle(cg, ctx, family, k, lo):
    """Iteration-bound parity runs: the matrix the GPU generated (read back, so both sides see the same bits even for
    the gram family) through the oracle's serial CG against the resident solve; full and symmetric products."""
    n, seed = 1024, 9
    flags = (cg.GEN_SPECTRUM if family == "spectrum" else cg.GEN_GRAM) | cg.GEN_KAPPA_LOG2(k)
    ctx.generate(n, seed, flags)
    A = ctx.debug_read(0, 0, n * n).reshape(n, n)
    b = ctx.debug_read(1, 0, n)
    xo, so = O.conjugrad(A, b, np.zeros(n))
    assert so["iterations"] >= lo, so["iterations"]
    for sym in (0, 1):
        ctx.set_x0(np.zeros(n))
        st = ctx.solve(symmetric=sym)              # symmetric=1 verifies a_ij == a_ji on the device first
        assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS
        if st["iterations"] == so["iterations"]:
            assert relerr(ctx.read_x(), xo) <= RTOL_X
