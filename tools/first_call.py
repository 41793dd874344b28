#!/usr/bin/env python
"""First-call cost of the staged upload: a caller shaped like the reference's main() calls conjugrad() ONCE, so the
pinned ring's allocation is part of its bill.  For each (n, threads, chunk): a fresh context, the first cgb_load_host
from malloc'ed memory timed, then a second one (ring cached).  Also: what cudaHostAlloc of the ring costs by size."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import conjugate_gradient_b200 as cg
from conjugate_gradient_b200 import _binding as B

torch.cuda.init(); torch.zeros(1, device="cuda")
for mb in (4, 16, 32, 64, 128):
    t0 = time.perf_counter(); t = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True); dt = time.perf_counter() - t0
    print(f"pin {mb} MiB: {dt*1e3:.2f} ms", flush=True); del t
L = B.lib()
for n in ([int(a) for a in sys.argv[1:]] or [2048, 4096, 8192, 16384]):
    A = np.random.default_rng(1).random((n, n)); z = np.zeros(n)
    for env in ({}, {"CGB_STAGE": "0"}, {"CGB_STAGE_CHUNK_KB": "1024"}, {"CGB_STAGE_CHUNK_KB": "2048"}, {"CGB_STAGE_CHUNK_KB": "4096"},
                {"CGB_STAGE_CHUNK_KB": "1024", "CGB_STAGE_THREADS": "4"}, {"CGB_STAGE_CHUNK_KB": "512", "CGB_STAGE_THREADS": "8"},
                {"CGB_STAGE_CHUNK_KB": "2048", "CGB_STAGE_THREADS": "4"}, {"CGB_STAGE_CHUNK_KB": "1664", "CGB_STAGE_THREADS": "4"},
                {"CGB_STAGE_CHUNK_KB": "2048", "CGB_STAGE_THREADS": "3"}):
        for k in ("CGB_STAGE", "CGB_STAGE_CHUNK_KB", "CGB_STAGE_THREADS"):
            os.environ.pop(k, None)
        os.environ.update(env)
        with cg.Context(0) as c:
            c.generate(n, 1)                                   # device arrays exist: only the upload is timed
            ts = []
            for _ in range(5):
                t0 = time.perf_counter(); c.load_host_ptr(n, A.ctypes.data, z.ctypes.data, z.ctypes.data); ts.append(time.perf_counter() - t0)
        print(f"n={n} {env or 'default'}: first {ts[0]*1e3:.1f} ms ({8*n*n/ts[0]/1e9:.1f} GB/s)  later {min(ts[1:])*1e3:.1f} ms ({8*n*n/min(ts[1:])/1e9:.1f} GB/s)  all {[round(t*1e3,1) for t in ts]}", flush=True)
