#!/usr/bin/env python
"""Small systems on one GPU: rows per CTA of the L2-keeping product against wave quantisation.
The three shapes hold the same 16 doubles in flight per thread (R x U = 4x1, 2x2, 1x4; 128 threads, 7 CTAs per SM =
1036 slots); a finer R means more, shorter CTAs.  Graph replay, 2000 fixed iterations, it/s."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import conjugate_gradient_b200 as cg

sizes = [int(a) for a in sys.argv[1:]] or [1024, 2048, 3072, 4096, 4608, 5120, 5632, 6144, 6656, 7168, 7680]
with cg.Context(0) as ctx:
    print("n      default        (4,1,128)      (2,2,128)      (1,4,128)     grid4 grid2 grid1")
    for n in sizes:
        ctx.generate(n, 20170209)
        row = []
        for shape in (None, (4, 1), (2, 2), (1, 4)):
            o = {} if shape is None else dict(gemv_rows=shape[0], gemv_unroll=shape[1], gemv_threads=128)
            best = 0.0
            for _ in range(3):
                ctx.set_x0(np.zeros(n))
                st = ctx.solve(ignore_stop=1, max_iter=2000, **o)
                best = max(best, 2000 / (st["loop_ms"] * 1e-3))
            row.append((best, st["gemv_rows"], st["gemv_unroll"], st["gemv_threads"], st["gemv_keep_rows"]))
        print(f"{n:5d} " + " ".join(f"{v:9.0f}({r},{u},{t})" for v, r, u, t, k in row) + f"   {-(-n // 4)} {-(-n // 2)} {n}  keep_rows {row[0][4]}", flush=True)
