"""Whole programs on the same files, as a user of the reference runs them: `prog A.txt b.txt x0.txt`.

The REAL reference programs (oracle/_ref: serialConjugate.c as shipped in FP32 and as float->double, parallel_cg.c
over the MPI shim) against the drop-in CLI on one GPU, wall clock from exec to exit, on the n x n system the CLI's
generator emits in generateSPDmatrix.m's "%.4f" format.  The reference's own report (results.xlsx) has 15.6 s of
its 16.0 s at n = 8192 in reading the text -- this is the run that shows what became of it.
usage: python tools/whole_program_table.py [n]  ->  a markdown table"""
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import conjugate_gradient_b200 as cg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
ref_dir = os.path.join(ROOT, "oracle", "_ref")
cli = os.path.join(os.path.dirname(cg.__file__), "lib", "cg_b200")
d = tempfile.mkdtemp()
pre = os.path.join(d, "sys")
subprocess.run([cli, "--gen", str(n), "--seed", "20170209", "--emit-files", pre], check=True, stdout=subprocess.DEVNULL)
files = [pre + s for s in ("_matrix.txt", "_vector.txt", "_x0.txt")]
text_mb = os.path.getsize(files[0]) / 1e6


def own_timers(out):
    """the seconds the program prints about itself (serialConjugate.c:249-251, parallel_cg.c:334-335)"""
    return "; ".join(f"{m.group(1).strip()} {float(m.group(2)):.4f}" for m in re.finditer(r"([A-Za-z ]+time[A-Za-z ]*):\s*([0-9.eE+-]+)", out))


rows = []


def run(label, argv, env=None):
    t = time.perf_counter()
    r = subprocess.run(argv, capture_output=True, text=True, timeout=600, env=dict(os.environ, **(env or {})))
    el = time.perf_counter() - t
    rows.append((label, el, r.returncode, own_timers(r.stdout)))


cores = os.cpu_count()
for prec in ("f32", "f64"):
    exe = os.path.join(ref_dir, f"serial_{prec}_n{n}")
    if os.path.exists(exe):
        run(f"REAL serialConjugate.c ({'as shipped, float' if prec == 'f32' else 'float->double'}), 1 core", [exe] + files)
for np_ in (2, 8):
    exe = os.path.join(ref_dir, f"parallel_f64_n{n}")
    if os.path.exists(exe) and np_ <= cores:
        run(f"REAL parallel_cg.c (float->double) over the MPI shim, {np_} ranks", [exe] + files, {"CGSHIM_NP": str(np_)})
for rep in range(2):
    run("**cg_b200 (libcgb200), 1 B200**" + (" (second run)" if rep else ""), [cli] + files + ["--dims", pre + "_dimensions.txt"])

print(f"n = {n}: matrix file {text_mb:.0f} MB of `%.4f` text; host cores {cores}; wall clock from exec to exit")
print()
print("| program | wall seconds | exit | the program's own timers |")
print("|---|---|---|---|")
for label, el, rc, own in rows:
    print(f"| {label} | {el:.3f} | {rc} | {own} |")
