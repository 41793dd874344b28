"""Sweep of the L2-residency share of the small-n GEMV (cgb_options.gemv_l2_keep_pct) on one B200."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import conjugate_gradient_b200 as cg

def main():
    sizes = [int(s) for s in (sys.argv[1:] or ["4096", "5120", "6144"])]
    with cg.Context(0) as ctx:
        for n in sizes:
            ctx.generate(n, 7)
            iters = 3000 if n <= 4096 else 1500
            for pct in (-1, 20, 35, 50, 60, 70, 80, 90, 100):
                best = 0.0
                for rep in range(2):
                    st = ctx.solve(ignore_stop=1, max_iter=iters, gemv_l2_keep_pct=pct)
                    best = max(best, st["iterations"] / (st["loop_ms"] * 1e-3))
                us = 1e6 / best
                print(f"n={n} keep_pct={pct:4d} keep_rows={st['gemv_keep_rows']:6d} {best:9.1f} it/s {us:7.2f} us/iter "
                      f"eff {8.0*n*n/us/1e6:6.2f} TB/s cfg r{st['gemv_rows']} u{st['gemv_unroll']} t{st['gemv_threads']}", flush=True)

main()
