"""CPU-only: pins the oracle (oracle/cg_oracle.c) against the reference.

1. against golden vectors produced by the REAL reference (tests/golden/golden.json,
   made by tests/golden/make_golden.py from oracle/_ref) -- runs everywhere;
2. bit for bit against the live oracle/_ref build when it is present (build
   container; on the GPU box the prebuilt files travel with the snapshot).
"""
import numpy as np
import pytest

import _oracle as O
from conftest import unhex

EPS = 1.0e-6  # serialConjugate.c:28


def case_inputs(c):
    n = c["n"]
    if "A" in c:
        return unhex(c["A"]).reshape(n, n), unhex(c["b"]), unhex(c["x0"])
    x0 = O.gen_b(n, c["x0_seed"]) if "x0_seed" in c else np.zeros(n)
    return O.gen_A(n, c["seed"], c["flags"]), O.gen_b(n, c["seed"], c["flags"]), x0


def same_bits(x, want):
    """bit-for-bit, except that a NaN matches any NaN (float.hex() keeps neither sign nor payload)"""
    nx, nw = np.isnan(x), np.isnan(want)
    return (nx == nw).all() and x[~nx].tobytes() == want[~nw].tobytes()


# g1-g2: the shipped fixtures; g3-g4: generateSPDmatrix-style (3-5 iterations); g5-g6: cond ~ 2^k,
# hundreds of iterations; g7: rho = 0 breakdown (tests/golden/make_golden.py)
@pytest.mark.parametrize("tag", ["g1", "g2", "g3", "g4", "g5", "g6", "g7"])
@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_oracle_matches_reference_golden_bit_for_bit(golden, tag, prec):
    c = golden["cases"][tag]
    dt = np.float64 if prec == "f64" else np.float32
    A, b, x0 = (a.astype(dt) for a in case_inputs(c))
    x, st = O.conjugrad(A, b, x0, eps=EPS)
    want = unhex(c["x"][f"serial_{prec}"], dt)
    assert same_bits(x, want)
    for P in c["ranks"]:
        xs, _ = O.conjugrad(A, b, x0, eps=EPS, ranks=P)
        want = unhex(c["x"][f"spmd{P}_{prec}"], dt)
        assert same_bits(xs, want), (tag, prec, P)


def test_long_runs_and_breakdown_behave_like_the_reference(golden):
    # g5: the FP64 reference converges after ~500 iterations, the FP32 one never meets sqrt(beta) < 1e-6 and
    # leaves through the loop bound k < ROWS (serialConjugate.c:211) -- the golden x above is that last iterate
    A, b, x0 = case_inputs(golden["cases"]["g5"])
    _, s64 = O.conjugrad(A, b, x0)
    _, s32 = O.conjugrad(A.astype(np.float32), b.astype(np.float32), x0.astype(np.float32))
    assert 400 < s64["iterations"] < 512 and s64["break_k"] == s64["iterations"] - 1
    assert s32["iterations"] == 512 and s32["break_k"] == -1
    # g6: ~150 iterations from a non-zero guess; the answer is the true solution to the stop tolerance
    A, b, x0 = case_inputs(golden["cases"]["g6"])
    assert np.abs(x0).max() > 0.5
    x, st = O.conjugrad(A, b, x0)
    assert 100 < st["iterations"] < 200 and np.linalg.norm(A @ x - b) < 1e-5
    # g7: x0 exact -> rho = 0 -> alpha = 0/0: NaNs to the loop bound, on every rank count (parallel_cg.c:298,318)
    c = golden["cases"]["g7"]
    A, b, x0 = case_inputs(c)
    assert (A @ x0 == b).all()
    x, st = O.conjugrad(A, b, x0)
    assert np.isnan(x).all() and st["iterations"] == 4 and st["break_k"] == -1 and st["rsold0"] == 0.0
    assert all(v == ["nan"] * 4 for v in c["x"].values())


def test_shipped_fixtures_known_answers(golden):
    # G1 -> (2/3, 1/3) at k=1 with beta = 0; G2 -> (-1,1,-1,1) at k=3
    A, b, x0 = case_inputs(golden["cases"]["g1"])
    x, st = O.conjugrad(A, b, x0)
    assert st["break_k"] == 1 and st["final_beta"] == 0.0
    np.testing.assert_allclose(x, [2 / 3, 1 / 3], rtol=1e-15)
    A, b, x0 = case_inputs(golden["cases"]["g2"])
    x, st = O.conjugrad(A, b, x0)
    assert st["break_k"] == 3
    np.testing.assert_allclose(x, [-1, 1, -1, 1], rtol=1e-13)
    np.testing.assert_allclose(x, np.linalg.solve(A, b), rtol=1e-13)


def test_generated_iteration_counts(golden):
    # FP64 breaks at k=4 (n=512) and k=3 (n=2048)
    for tag, k64 in (("g3", 4), ("g4", 3)):
        A, b, x0 = case_inputs(golden["cases"][tag])
        x, st = O.conjugrad(A, b, x0)
        assert st["break_k"] == k64
        xs = np.linalg.solve(A, b)
        assert np.linalg.norm(x - xs) / np.linalg.norm(xs) < 1e-7


# 16384 is the largest size SURVEY.md 8(c) names for this check (2.1 GB of FP64 matrix, a few seconds): the
# 64-bit-index port and the real int-indexed reference still address the same elements there
@pytest.mark.parametrize("n,ranks", [(2, 1), (2, 2), (4, 4), (512, 8), (2048, 4), (16384, 8)])
@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_oracle_vs_live_reference(n, ranks, prec):
    if not (O.ref_available("serial", prec, n) and O.ref_available("parallel", prec, n)):
        pytest.skip("oracle/_ref not built (needs /root/reference; run make -C oracle)")
    dt = np.float64 if prec == "f64" else np.float32
    rng = np.random.default_rng(n * 31 + ranks)
    if n <= 4:
        R = rng.random((n, n)); A = (0.5 * (R + R.T) + n * np.eye(n)).astype(dt)
        b = rng.random(n).astype(dt); x0 = rng.random(n).astype(dt)
    else:
        A = O.gen_A(n, 777 + n).astype(dt); b = O.gen_b(n, 777 + n).astype(dt)
        x0 = (0.01 * rng.random(n)).astype(dt)      # non-zero guess: exercises A*x0
    x, _ = O.conjugrad(A, b, x0)
    assert x.tobytes() == O.ref_conjugrad(A, b, x0).tobytes()
    xs, _ = O.conjugrad(A, b, x0, ranks=ranks)
    assert xs.tobytes() == O.ref_conjugrad(A, b, x0, ranks=ranks).tobytes()


@pytest.mark.parametrize("n,ranks,k", [(4, 2, 0), (4, 4, 0), (512, 8, 0), (512, 4, 12), (2048, 8, 7)])
@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_point_to_point_program_is_the_same_arithmetic(n, ranks, k, prec):
    # point-to-point_cg.c:397-504 (SURVEY.md 8a-1's "p2p twin"): the real program over the shim's MPI_Send /
    # MPI_Recv gives the bits of the collective program and of the oracle -- its emulated reduce adds the
    # partials in rank order at rank 0 (:339-359) -- up to runs of 512 iterations.  So the SPMD golden vectors
    # and `conjugrad_mpi` stand for it as well.
    if not (O.ref_available("p2p", prec, n) and O.ref_available("parallel", prec, n)):
        pytest.skip("oracle/_ref not built (needs /root/reference; run make -C oracle)")
    dt = np.float64 if prec == "f64" else np.float32
    if n == 4:
        rng = np.random.default_rng(3)
        R = rng.random((n, n)); A = (0.5 * (R + R.T) + n * np.eye(n)).astype(dt)
        b = rng.random(n).astype(dt); x0 = rng.random(n).astype(dt)
    else:
        f = O.GEN_KAPPA_LOG2(k)
        A = O.gen_A(n, 5, f).astype(dt); b = O.gen_b(n, 5, f).astype(dt); x0 = np.zeros(n, dt)
    xp = O.ref_conjugrad(A, b, x0, ranks=ranks, program="p2p")
    assert xp.tobytes() == O.ref_conjugrad(A, b, x0, ranks=ranks).tobytes()
    xo, so = O.conjugrad(A, b, x0, ranks=ranks)
    assert xp.tobytes() == xo.tobytes()
    if k == 12:
        assert so["iterations"] > 400


def test_threaded_team_equals_sequential_spmd():
    n, seed = 512, 99
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    for P in (1, 2, 8):
        xt, st = O.team_solve_generated(n, seed, P)
        xs, ss = O.conjugrad(A, b, np.zeros(n), ranks=P)
        assert xt.tobytes() == xs.tobytes()
        assert st["iterations"] == ss["iterations"] and st["final_beta"] == ss["final_beta"]
    with pytest.raises(ValueError):
        O.team_solve_generated(510, seed, 4)       # parallel_cg.c:86-90


def test_fixed_iteration_mode_keeps_iterating():
    n, seed = 256, 5
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    x, st = O.conjugrad(A, b, np.zeros(n), max_iter=12, ignore_stop=True)
    assert st["iterations"] == 12 and st["break_k"] == -1 and 0 <= st["first_hit_k"] < 12


def test_generator_known_answers(golden):
    g = golden["generator"]
    seed = golden["seed"]
    for k, hx in g["u"].items():
        u = O.lib().orc_gen_sample_u(seed, int(k))
        assert u.hex() == hx and 0.0 <= u < 1.0
    assert O.gen_A(8, seed).ravel().tobytes() == unhex(g["A8"]).tobytes()
    assert O.gen_b(8, seed).tobytes() == unhex(g["b8"]).tobytes()
    assert O.gen_A(8, seed, O.GEN_ROUND4).ravel().tobytes() == unhex(g["A8_round4"]).tobytes()


def test_generator_construction_matches_generateSPDmatrix():
    # generateSPDmatrix.m:4-17: A = 0.5(R+R') + n I with R ~ U(0,1); b ~ U(0,1)
    n, seed = 300, 4242
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    assert (A == A.T).all()
    off = A[~np.eye(n, dtype=bool)]
    assert off.min() > 0.0 and off.max() < 1.0 and abs(off.mean() - 0.5) < 0.01
    d = np.diag(A)
    assert d.min() >= n and d.max() < n + 1
    assert b.min() >= 0.0 and b.max() < 1.0
    w = np.linalg.eigvalsh(A)
    assert w[0] > 0 and w[-1] / w[0] < 2.0          # kappa ~ 1.5
    # row blocks generated independently agree with the whole matrix
    assert (O.gen_A(n, seed, row0=100, m=50) == A[100:150]).all()
    # the %.4f round trip of generateSPDmatrix.m:29
    Ar = O.gen_A(n, seed, O.GEN_ROUND4)
    assert np.abs(Ar - A).max() <= 0.5e-4 + 1e-12 and (Ar == Ar.T).all()
    assert np.allclose(Ar * 1e4, np.rint(Ar * 1e4), atol=1e-6)


def test_generator_spread_family_is_spd_with_prescribed_condition():
    # n*I -> n*diag(1 + (2^k - 1) t^2); k = 0 is the reference's matrix
    n, seed = 256, 99
    assert O.gen_A(n, seed, O.GEN_KAPPA_LOG2(0)).tobytes() == O.gen_A(n, seed).tobytes()
    iters = []
    for k in (4, 7):
        A = O.gen_A(n, seed, O.GEN_KAPPA_LOG2(k))
        assert (A == A.T).all()
        d = np.diag(A)
        assert abs(d[0] - n) < 1.0 and abs(d[-1] - n * 2.0 ** k) < 1.0 and (np.diff(d) > -1.0).all()
        w = np.linalg.eigvalsh(A)
        assert w[0] > 0 and 0.5 * 2 ** k < w[-1] / w[0] < 2.0 * 2 ** k
        assert (O.gen_A(n, seed, O.GEN_KAPPA_LOG2(k), row0=64, m=32) == A[64:96]).all()
        x, st = O.conjugrad(A, O.gen_b(n, seed), np.zeros(n))
        assert st["break_k"] >= 0 and np.linalg.norm(A @ x - O.gen_b(n, seed)) < 1e-5
        iters.append(st["iterations"])
    assert iters[0] > 20 and iters[1] > 2 * iters[0]      # ~sqrt(kappa) growth: iteration-bound workloads


def test_fused_reductions_oracle_agrees_with_the_pinned_classic_oracle(golden):
    # Chronopoulos-Gear CG has no counterpart in the reference: its CPU statement is validated here
    # against the classic oracle (which IS pinned): same iterates in exact arithmetic
    for n, seed, flags in ((512, 3, 0), (1024, 9, O.GEN_KAPPA_LOG2(4)), (768, 5, O.GEN_KAPPA_LOG2(7))):
        A, b = O.gen_A(n, seed, flags), O.gen_b(n, seed, flags)
        xc, sc = O.conjugrad(A, b, np.zeros(n))
        xf, sf = O.cg_fused(A, b, np.zeros(n))
        assert abs(sf["iterations"] - sc["iterations"]) <= 1 and sf["break_k"] >= 0
        assert np.linalg.norm(xf - xc) / np.linalg.norm(xc) <= 1e-10
        assert sf["rsold0"] == sc["rsold0"]
    c = golden["cases"]["g2"]                                # the shipped 4x4 fixture
    A, b, x0 = unhex(c["A"]).reshape(4, 4), unhex(c["b"]), unhex(c["x0"])
    xf, sf = O.cg_fused(A, b, x0)
    np.testing.assert_allclose(xf, [-1, 1, -1, 1], rtol=1e-12)
    xk, sk = O.cg_fused(A, b, x0, max_iter=2, eps=1e-300)
    xc, _ = O.conjugrad(A, b, x0, max_iter=2, eps=1e-300)
    np.testing.assert_allclose(xk, xc, rtol=1e-12)


@pytest.mark.parametrize("n", [4, 512, 2048])
@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_building_blocks_against_the_real_functions(n, prec):
    """matVec / vecVec of the REAL build (serialConjugate.c:109-120,145-155, exported by oracle/_ref's library form)
    against the oracle's restatements, bit for bit: these two carry all of the path's rounding (strict
    left-to-right accumulation, the product rounded before each add)."""
    import ctypes as C
    import os
    if not O.ref_available("serial", prec, n):
        pytest.skip("oracle/_ref not built")
    dt, ct = (np.float64, C.c_double) if prec == "f64" else (np.float32, C.c_float)
    L = C.CDLL(os.path.join(O.REF_DIR, f"libserial_{prec}_n{n}.so"))
    vp = C.c_void_p
    L.matVec.argtypes = [vp, vp, vp]; L.matVec.restype = None
    L.vecVec.argtypes = [vp, vp]; L.vecVec.restype = ct
    rng = np.random.default_rng(n)
    A = (rng.standard_normal((n, n)) * 10.0 ** rng.integers(-2, 3, (n, n))).astype(dt)
    v = rng.standard_normal(n).astype(dt); w = rng.standard_normal(n).astype(dt)
    y = np.full(n, np.nan, dtype=dt)
    L.matVec(y.ctypes.data, A.ctypes.data, v.ctypes.data)
    assert y.tobytes() == O.matvec(A, v).tobytes()
    assert dt(L.vecVec(v.ctypes.data, w.ctypes.data)).tobytes() == dt(O.vecvec(v, w)).tobytes()
    # ... and these bits are NOT what a fused or pairwise evaluation gives (the test has teeth)
    if n >= 512:
        assert y.tobytes() != (A @ v).astype(dt).tobytes()


# ----------------------------------------------------------------- the two further generator families
@pytest.mark.parametrize("n,k", [(64, 4), (256, 7), (301, 10)])
def test_prescribed_spectrum_family_has_exactly_the_stated_eigenvalues(n, k):
    """ORC_GEN_SPECTRUM: A = H diag(lambda) H with lambda_i = n (1 + (2^k - 1) t_i^2): symmetric bit for bit, and its
    eigenvalues (LAPACK) are the stated ones."""
    A = O.gen_A(n, 11, O.GEN_SPECTRUM | O.GEN_KAPPA_LOG2(k))
    assert np.array_equal(A, A.T)
    t = np.arange(n) / (n - 1)
    want = np.sort(n * (1.0 + (2.0 ** k - 1.0) * t * t))
    got = np.sort(np.linalg.eigvalsh(A))
    assert np.max(np.abs(got - want) / want) < 1e-11
    # rows generated on their own equal the rows of the whole matrix (what a rank of a row-sharded run generates)
    assert O.gen_A(n, 11, O.GEN_SPECTRUM | O.GEN_KAPPA_LOG2(k), 5, 7).tobytes() == A[5:12].tobytes()


def test_prescribed_spectrum_family_is_iteration_bound():
    n = 512
    counts = []
    for k in (4, 7, 10):
        A = O.gen_A(n, 3, O.GEN_SPECTRUM | O.GEN_KAPPA_LOG2(k))
        x, st = O.conjugrad(A, O.gen_b(n, 3), np.zeros(n))
        assert np.linalg.norm(A @ x - O.gen_b(n, 3)) < 1e-5
        counts.append(st["iterations"])
    assert counts[0] > 20 and counts[1] > 2 * counts[0] and counts[2] > 2 * counts[1], counts


@pytest.mark.parametrize("n,k", [(96, 0), (200, 5)])
def test_gram_family_is_R_Rt_plus_shift(n, k):
    """ORC_GEN_GRAM: generateSPDmatrix.m:12's `A = A*A'` followed by :17's `+ n*eye(n)` (with the spread diagonal)."""
    flags = O.GEN_GRAM | O.GEN_KAPPA_LOG2(k)
    A = O.gen_A(n, 21, flags)
    R = np.array([[O.lib().orc_gen_sample_u(21, i * n + j) for j in range(n)] for i in range(n)])
    t = np.arange(n) / (n - 1)
    d = 1.0 + (2.0 ** k - 1.0) * t * t if k else np.ones(n)
    want = R @ R.T + np.diag(n * d)
    assert np.array_equal(A, A.T)
    assert np.max(np.abs(A - want) / np.abs(want)) < 1e-13
    assert np.all(np.linalg.eigvalsh(A) > 0)
    assert O.gen_A(n, 21, flags, 10, 3).tobytes() == A[10:13].tobytes()
