set -x
timeout 900 python -m pytest tests/test_gpu_ingest.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -12
# ingest throughput: n = 8192 in generateSPDmatrix.m's format (the reference's largest published size)
python - <<'PY'
import time, os, sys, numpy as np
sys.path.insert(0, "tests")
import conjugate_gradient_b200 as cg
n = 8192
pre = "/tmp/sys8192"
t = time.time()
os.system(f"conjugate_gradient_b200/lib/cg_b200 --gen {n} --seed 1 --emit-files {pre} > /tmp/emit.log 2>&1")
print("emit+solve seconds", round(time.time() - t, 2), "file MB", os.path.getsize(pre + "_matrix.txt") / 1e6)
with cg.Context(0) as ctx:
    for rep in range(3):
        t = time.time(); ctx.load_text(n, pre + "_matrix.txt", pre + "_vector.txt", pre + "_x0.txt"); dt = time.time() - t
        print(f"GPU ingest n={n}: {dt:.3f} s  {n*n/dt/1e6:.1f} Mvalues/s  {os.path.getsize(pre + '_matrix.txt')/dt/1e9:.2f} GB/s of text")
    t = time.time(); v = cg.read_values(pre + "_matrix.txt", n * n); dt = time.time() - t
    print(f"host parallel parser: {dt:.3f} s  {n*n/dt/1e6:.1f} Mvalues/s ({os.cpu_count()} cores)")
    assert ctx.debug_read(0, 0, n * n).tobytes() == v.tobytes()
t = time.time()
os.system(f"conjugate_gradient_b200/lib/cg_b200 {pre}_matrix.txt {pre}_vector.txt {pre}_x0.txt --dims {pre}_dimensions.txt --stats")
print("CLI wall seconds (files, 1 GPU)", round(time.time() - t, 2))
PY
