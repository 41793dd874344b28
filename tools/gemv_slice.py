#!/usr/bin/env python
"""One rank's row block as a stand-alone product: y = A_blk v with the fused dot, A_blk of shape m x n
(default 8192 x 65536 = the per-GPU slice of the headline system on 8 GPUs, SURVEY.md 8(d) "ncu evidence").
Runs the plain streaming kernel through cgb_op_matvec `reps` times; meant to be wrapped in
  ncu --set full --clock-control none --import-source on -k regex:rowblock_matvec -c 2 -o gpurun_out/slice python tools/gemv_slice.py
so that dram__bytes_read.sum can be set against the slice's algorithmic bytes 8 (m n + n + m).
usage: gemv_slice.py [m n reps rows unroll threads]   (rows unroll threads = 4 2 256: the shape of the exchange-fused kernel)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import conjugate_gradient_b200 as cg

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
shape = dict(gemv_rows=int(sys.argv[4]), gemv_unroll=int(sys.argv[5]), gemv_threads=int(sys.argv[6])) if len(sys.argv) > 6 else {}
rng = np.random.default_rng(1)
A = rng.random((m, n)); v = rng.random(n); w = rng.random(m)
with cg.Context(0) as ctx:
    for _ in range(reps):
        y, dot = ctx.op_matvec(A, v, w, **shape)
ref = A @ v
print(f"m={m} n={n} algorithmic bytes {8.0 * (m * n + n + m):.0f}  max rel err vs numpy {np.abs(y - ref).max() / np.abs(ref).max():.2e}  "
      f"dot rel err {abs(dot - ref @ w) / abs(ref @ w):.2e}")
