set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "symmetric" 2>&1 | tail -12
python - <<'PY'
import sys
sys.path.insert(0, "tests")
import numpy as np
import conjugate_gradient_b200 as cg
with cg.Context(0) as ctx:
    for n in (8192, 16384, 65536):
        ctx.generate(n, 7)
        gb_full = 8.0 * (n * n + 2 * n)
        avg, mn = ctx.time_gemv(warmup=2, reps=6)
        print(f"n={n} full  {mn:.4f} ms  {gb_full/mn/1e6:.0f} GB/s")
        for ru in (1, 2, 4):
            avg, mn = ctx.time_gemv(warmup=2, reps=6, symmetric=1, gemv_unroll=ru)
            gb = 8.0 * (0.5 * n * (n + 1) + 2 * n)
            print(f"n={n} symv ru={ru} {mn:.4f} ms  {gb/mn/1e6:.0f} GB/s on declared bytes  speedup {0:.2f}", flush=True)
PY
python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(l['value'], l['graph_replay'], l.get('symmetric_variant'))"
