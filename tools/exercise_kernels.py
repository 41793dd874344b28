"""Every kernel of libcgb200.so once or twice at small sizes, through the C ABI, no torch; results
are checked against numpy (|b - Ax| / |b|).  A quick whole-library exercise for a new box:

    python tools/exercise_kernels.py [ranks]

ranks > 1 runs the row-sharded paths (threads of this process, one GPU each, peer exchange).
(compute-sanitizer is closed on the build pool, so memory-safety evidence is this plus the parity
tests' bounds: odd shapes, ragged tails, NaN-poisoned regions that must never be read.)"""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import conjugate_gradient_b200 as cg  # noqa: E402


def spd(n, seed):
    rng = np.random.default_rng(seed)
    R = rng.random((n, n))
    return 0.5 * (R + R.T) + n * np.eye(n), rng.random(n)


def check(x, A, b, what):
    res = np.linalg.norm(b - A @ x) / np.linalg.norm(b)
    assert res < 1e-6, (what, res)
    print(f"  ok {what}: |b - Ax|/|b| = {res:.2e}", flush=True)


def one_rank():
    with cg.Context(0) as ctx:
        # generator + streaming-load GEMV (several shapes), bulk ring, L2-keep, stream and graph modes
        for n, opts in [(300, {}), (1024, dict(gemv_variant=cg.GEMV_BULK)), (1000, dict(use_graph=0)),
                        (2048, dict(gemv_rows=8, gemv_unroll=2, gemv_threads=512)),
                        (4096, dict(gemv_l2_keep_pct=60)), (513, dict(guard=1)),
                        (1024, dict(cg_variant=cg.CG_FUSED_REDUCTIONS)),
                        (1024, dict(symmetric=1)), (1536, dict(symmetric=1, cg_variant=cg.CG_FUSED_REDUCTIONS)),
                        (2048, dict(symmetric=1, gemv_unroll=2)), (768, dict(ignore_stop=1, max_iter=40))]:
            ctx.generate(n, 7 + n, cg.GEN_KAPPA_LOG2(3) if n % 2 == 0 else 0)
            st = ctx.solve(**opts)
            A, b = ctx.op_generate(n, 7 + n, cg.GEN_KAPPA_LOG2(3) if n % 2 == 0 else 0)
            check(ctx.read_x(), A, b, f"n={n} {opts} it={st['iterations']}")
        # host matrices: full upload, asserted-symmetric upload, device symmetry check
        A, b = spd(1024, 1)
        x = np.zeros(1024)
        ctx.conjugrad_rank(A, b, x)
        check(x, A, b, "host upload")
        x[:] = 0
        ctx.conjugrad_rank(A, b, x, symmetric=2)
        check(x, A, b, "asserted-symmetric upload")
        ctx.load_host(A, b, np.zeros(1024))
        ctx.solve(symmetric=1)
        check(ctx.read_x(), A, b, "verified-symmetric")
        # single-kernel operators
        v = np.random.default_rng(2).random(1024)
        y, d = ctx.op_matvec(A, v, w=v)
        assert np.allclose(y, A @ v, rtol=1e-13) and np.isclose(d, v @ (A @ v), rtol=1e-13)
        ctx.op_init_residual(b, y)
        ctx.op_xr_update(1.0, 2.0, v.copy(), v.copy(), v.copy(), y)
        ctx.op_p_update(1.0, 0.5, 1e-6, v.copy(), v.copy())
        ctx.time_gemv(warmup=1, reps=2)
        print("  ok operators", flush=True)
        # text ingest on the GPU
        with tempfile.TemporaryDirectory() as d_:
            n = 256
            A, b = spd(n, 3)
            A = np.round(A, 4); b = np.round(b, 4)
            fa, fb, fx = (os.path.join(d_, f) for f in ("A.txt", "b.txt", "x.txt"))
            np.savetxt(fa, A.ravel(), fmt="%.4f"); np.savetxt(fb, b, fmt="%.4f"); np.savetxt(fx, np.zeros(n), fmt="%.1f")
            ctx.load_text(n, fa, fb, fx)
            ctx.solve()
            check(ctx.read_x(), A, b, "text ingest")


def ranks(g):
    for n, opts in [(1024, {}), (2048, dict(cg_variant=cg.CG_FUSED_REDUCTIONS)), (2048, dict(symmetric=1)),
                    (1024 * g, dict(symmetric=1, cg_variant=cg.CG_FUSED_REDUCTIONS)), (1000 * g, dict(comm=cg.COMM_NCCL))]:
        x, st = cg.solve_generated(n, 5, gpus=g, **opts)
        with cg.Context(0) as ctx:
            A, b = ctx.op_generate(n, 5)
        check(x, A, b, f"ranks={g} n={n} {opts} comm={st['comm']} it={st['iterations']}")
    A, b = spd(512 * g, 4)
    x, st = cg.conjugrad_host(A, b, np.zeros(512 * g), gpus=g, symmetric=2)
    check(x, A, b, f"ranks={g} asserted-symmetric upload")


if __name__ == "__main__":
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    one_rank() if g == 1 else ranks(g)
    print("exercise_kernels: all checks passed", flush=True)
