#!/usr/bin/env python
"""Sweeps the GEMV launch configurations on a B200 (run under gpurun) and prints a table.
large n: cgb_time_gemv (CUDA events around single launches, A >> L2);
small n: iterations/s of the graph-replayed fixed-iteration solve (what the user sees)."""
import itertools
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import conjugate_gradient_b200 as cg

sizes = [int(a) for a in sys.argv[1:]] or [4096, 16384, 65536]
ldg = [dict(gemv_variant=1, gemv_rows=r, gemv_unroll=u, gemv_threads=t, gemv_vec=v)
       for r, u, t, v in itertools.product((1, 2, 4, 8), (1, 2, 4), (128, 256, 512), (2, 4)) if r * u * v <= 64]
bulk = [dict(gemv_variant=2, gemv_rows=r, gemv_tile_cols=c, gemv_stages=s, gemv_threads=t)
        for r, c, s, t in ((4, 512, 4, 128), (4, 512, 8, 128), (8, 512, 4, 128), (8, 512, 5, 128), (8, 256, 8, 128),
                           (4, 1024, 4, 256), (8, 1024, 2, 256), (8, 1024, 3, 256), (16, 256, 5, 128), (16, 512, 3, 128))]
out = []
with cg.Context(0) as ctx:
    for n in sizes:
        ctx.generate(n, 20170209)
        gb = 8.0 * (n * n + 2 * n) / 1e9
        res = []
        for cfg in ldg + bulk:
            try:
                if n >= 16384:
                    avg, mn = ctx.time_gemv(warmup=2, reps=6, **cfg)
                    res.append((mn, avg, cfg))
                else:
                    st = ctx.solve(ignore_stop=1, max_iter=400, **cfg)
                    per = st["loop_ms"] / 400
                    res.append((per, per, cfg))
            except cg.CgbError as e:
                print("skip", cfg, e)
        res.sort(key=lambda t: t[0])
        print(f"== n={n}  ({gb:.3f} GB per GEMV)")
        for mn, avg, cfg in res[:12] + res[-3:]:
            print(f"  {mn*1e3:9.1f} us  avg {avg*1e3:9.1f} us  {gb/mn*1e3/1e3:7.2f} TB/s-equiv  {cfg}")
        out.append({"n": n, "results": [{"ms_min": a, "ms_avg": b, **c} for a, b, c in res]})
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gpurun_out", "tune_gemv.json"), "w"))
