#!/usr/bin/env python
"""Regenerates tests/golden/golden.json and the g1_*/g2_* text fixtures.

Run in the build container, where /root/reference exists and
oracle/build_ref.sh has compiled the REAL reference into oracle/_ref:

    make -C oracle && python tests/golden/make_golden.py

Every expected vector below is the output of the reference's own conjugrad()
(serialConjugate.c:180-259 / parallel_cg.c:248-345 over oracle/mpi_shim), never
of the oracle restatement.  The GPU box has no /root/reference: tests there
read only the committed JSON.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as O  # noqa: E402

REF = os.environ.get("CG_REFERENCE_DIR", "/root/reference")
SEED = 20170209  # the reference's date stamp (serialConjugate.c:17), used as default seed


def read_vals(name):
    with open(os.path.join(REF, name), "rb") as f:
        txt = f.read().decode("utf-8", "ignore").replace("﻿", " ")
    return [float(t) for t in txt.split()]


def hexes(a):
    return [float(v).hex() for v in np.asarray(a, dtype=np.float64)]


def ref_all(A, b, x0, ranks_list):
    out = {}
    for prec, dt in (("f64", np.float64), ("f32", np.float32)):
        Ad, bd, xd = A.astype(dt), b.astype(dt), x0.astype(dt)
        out[f"serial_{prec}"] = hexes(O.ref_conjugrad(Ad, bd, xd))
        for P in ranks_list:
            xs = O.ref_conjugrad(Ad, bd, xd, ranks=P)
            out[f"spmd{P}_{prec}"] = hexes(xs)
            if O.ref_available("p2p", prec, A.shape[0]):      # point-to-point_cg.c: the same bits, not stored twice
                xp = O.ref_conjugrad(Ad, bd, xd, ranks=P, program="p2p")
                assert np.array_equal(xs, xp, equal_nan=True), (prec, P)
    return out


def main():
    gold = {"seed": SEED, "cases": {}, "generator": {}}
    # --- shipped fixtures (G1, G2)
    for tag, fa, fb, fx, n, ranks in (
            ("g1", "matrixA.txt", "vectorb.txt", "initialguess.txt", 2, [1, 2]),
            ("g2", "matrixA1.txt", "vectorb1.txt", "X0.txt", 4, [1, 2, 4])):
        A = np.array(read_vals(fa)).reshape(n, n)
        b = np.array(read_vals(fb)); x0 = np.array(read_vals(fx))[:n]
        for nm, arr in (("matrixA", A.ravel()), ("vectorb", b), ("x0", x0)):
            with open(os.path.join(HERE, f"{tag}_{nm}.txt"), "w") as f:
                f.writelines(f"{v:.1f}\n" for v in arr)
        with open(os.path.join(HERE, f"{tag}_dimensions.txt"), "w") as f:
            f.write(f"{n}\n{n}\n{n}\n1\n")
        gold["cases"][tag] = {"n": n, "A": hexes(A.ravel()), "b": hexes(b), "x0": hexes(x0),
                              "ranks": ranks, "x": ref_all(A, b, x0, ranks)}
    # --- generator-style systems (G3, G4): inputs are regenerated from (seed, n)
    for tag, n in (("g3", 512), ("g4", 2048)):
        A = O.gen_A(n, SEED); b = O.gen_b(n, SEED); x0 = np.zeros(n)
        ranks = [1, 2, 4, 8]
        gold["cases"][tag] = {"n": n, "seed": SEED, "flags": 0, "ranks": ranks,
                              "x": ref_all(A, b, x0, ranks)}
    # --- long runs (SURVEY.md 8f-2 family, cond ~ 2^k): hundreds of iterations of the REAL reference pin the
    #     recurrence itself (alpha, beta/rho, the stale-p exit), not just its first steps.
    #     g5: FP64 stops after ~500 iterations, FP32 never meets the test and runs into the loop bound k < ROWS
    #     (serialConjugate.c:211); g6: non-zero initial guess, so A*x0 matters (serialConjugate.c:209-210)
    for tag, n, k, ranks, x0_seed in (("g5", 512, 12, [1, 2, 4, 8], None), ("g6", 2048, 7, [1, 8], SEED + 1)):
        flags = O.GEN_KAPPA_LOG2(k)
        A = O.gen_A(n, SEED, flags); b = O.gen_b(n, SEED, flags)
        x0 = np.zeros(n) if x0_seed is None else O.gen_b(n, x0_seed)
        c = {"n": n, "seed": SEED, "flags": flags, "ranks": ranks, "x": ref_all(A, b, x0, ranks)}
        if x0_seed is not None:
            c["x0_seed"] = x0_seed
        gold["cases"][tag] = c
    # --- no breakdown guard (parallel_cg.c:298,318): x0 already exact -> rho = 0 -> alpha = 0/0; the reference
    #     carries NaNs to the loop bound.  NaN payload/sign is not portable through float.hex(): "nan" marks them.
    A = np.array([[4.0, 1, 0, 0], [1, 4, 1, 0], [0, 1, 4, 1], [0, 0, 1, 4]]); x0 = np.array([1.0, 2, 3, 4]); b = A @ x0
    gold["cases"]["g7"] = {"n": 4, "A": hexes(A.ravel()), "b": hexes(b), "x0": hexes(x0), "ranks": [1, 2, 4],
                           "x": ref_all(A, b, x0, [1, 2, 4])}
    # --- generator known answers (pins the CPU twin; the GPU generator is
    #     compared with the twin bit for bit in the gpu tests)
    ks = [0, 1, 2, 3, 1000, 2**31, 2**40 + 12345]
    gold["generator"]["u"] = {str(k): float(O.lib().orc_gen_sample_u(SEED, k)).hex() for k in ks}
    A8 = O.gen_A(8, SEED); b8 = O.gen_b(8, SEED)
    gold["generator"]["A8"] = hexes(A8.ravel()); gold["generator"]["b8"] = hexes(b8)
    A8r = O.gen_A(8, SEED, O.GEN_ROUND4)
    gold["generator"]["A8_round4"] = hexes(A8r.ravel())
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(gold, f, indent=0)
    print("wrote golden.json:", {k: v["n"] for k, v in gold["cases"].items()})


if __name__ == "__main__":
    main()
