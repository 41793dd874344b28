"""Diagnosis: GPU text ingest against the host parser and the generator's CPU twin, per staging-chunk size."""
import os, subprocess, sys, tempfile
import numpy as np
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import conjugate_gradient_b200 as cg
import _oracle as O
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = tempfile.mkdtemp(); pre = os.path.join(d, "sys")
cli = os.path.join(os.path.dirname(cg.__file__), "lib", "cg_b200")
subprocess.run([cli, "--gen", str(n), "--seed", "1", "--emit-files", pre], check=True, stdout=subprocess.DEVNULL)
fa, fb, fx = (pre + s for s in ("_matrix.txt", "_vector.txt", "_x0.txt"))
raw = open(fa, "rb").read()
print("text bytes", len(raw), "head", raw[:40])
want = O.gen_A(n, 1, O.GEN_ROUND4).ravel()
v = cg.read_values(fa, n * n)
print("host parser vs generator twin: mismatches", int((v != want).sum()))
a = np.frombuffer(raw, np.uint8)
starts = np.flatnonzero((a != 10) & np.concatenate(([True], a[:-1] == 10)))
with cg.Context(0) as ctx:
    for mib in (8, 8, 8, 32, 32, 2, 16):
        os.environ["CGB_INGEST_CHUNK_MIB"] = str(mib)
        ctx.load_text(n, fa, fb, fx)
        got = ctx.debug_read(0, 0, n * n)
        bad = np.flatnonzero(got != v)
        print("chunk", mib, "MiB: mismatches", len(bad))
        for i in bad[:6]:
            print("   idx", int(i), "row", int(i) // n, "col", int(i) % n, "gpu", repr(float(got[i])), "host", repr(float(v[i])),
                  "byte", int(starts[i]), "byte mod chunk", int(starts[i]) % (mib << 20), "text", raw[starts[i] - 8:starts[i] + 12])
        if len(bad):
            print("   last bad idx", int(bad[-1]), "distinct rows", len(set((bad // n).tolist())))
