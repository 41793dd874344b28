set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "symmetric" 2>&1 | tail -6
for sp in 1 2 4; do
CGB_SYMV_SPLIT=$sp python - <<'PY'
import os, sys
sys.path.insert(0, "tests")
import conjugate_gradient_b200 as cg
with cg.Context(0) as ctx:
    for n in (8192, 16384, 32768, 65536):
        ctx.generate(n, 7)
        for ru in (2, 4):
            avg, mn = ctx.time_gemv(warmup=2, reps=6, symmetric=1, gemv_unroll=ru)
            gb = 8.0 * (0.5 * n * (n + 1) + 2 * n)
            print(f"split={os.environ['CGB_SYMV_SPLIT']} n={n} ru={ru} {mn:.4f} ms {gb/mn/1e6:.0f} GB/s declared", flush=True)
PY
done
