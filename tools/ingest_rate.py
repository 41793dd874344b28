"""Text-ingest rate on one GPU: an n x n matrix in generateSPDmatrix.m's "%.4f\\n" format
through cgb_load_text, with the library's own phase stamps (CGB_INGEST_TIMING) and a sweep of the staging-chunk
size (CGB_INGEST_CHUNK_MIB).  usage: python tools/ingest_rate.py [n] [chunk MiB ...]  ->  one JSON line"""
import json
import os
import subprocess
import sys
import tempfile
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import conjugate_gradient_b200 as cg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
chunks = [int(a) for a in sys.argv[2:]] or [8, 32, 4, 16, 8]
d = tempfile.mkdtemp()
pre = os.path.join(d, "sys")
cli = os.path.join(os.path.dirname(cg.__file__), "lib", "cg_b200")
t = time.time()
subprocess.run([cli, "--gen", str(n), "--seed", "1", "--emit-files", pre], check=True, stdout=subprocess.DEVNULL)
out = {"n": n, "cores": os.cpu_count(), "emit_and_solve_seconds": round(time.time() - t, 3),
       "text_bytes": os.path.getsize(pre + "_matrix.txt"), "runs": []}
os.environ["CGB_INGEST_TIMING"] = "1"
os.environ["CGB_LOCATE_NO_CACHE"] = "1"      # every load counts the file's tokens itself, as a fresh process would
fa, fb, fx = (pre + s for s in ("_matrix.txt", "_vector.txt", "_x0.txt"))
with cg.Context(0) as ctx:
    for mib in chunks:
        os.environ["CGB_INGEST_CHUNK_MIB"] = str(mib)
        best = None
        for rep in range(3):
            t = time.time(); ctx.load_text(n, fa, fb, fx); dt = time.time() - t
            best = dt if best is None else min(best, dt)
        out["runs"].append({"chunk_mib": mib, "seconds_best_of_3": round(best, 4), "mvalues_per_s": round(n * n / best / 1e6, 1),
                            "text_gbs": round(out["text_bytes"] / best / 1e9, 2)})
    del os.environ["CGB_INGEST_CHUNK_MIB"]
    t = time.time(); v = cg.read_values(fa, n * n); dt = time.time() - t
    out["host_parser"] = {"seconds": round(dt, 4), "mvalues_per_s": round(n * n / dt / 1e6, 1)}
    ctx.load_text(n, fa, fb, fx)
    assert ctx.debug_read(0, 0, n * n).tobytes() == v.tobytes(), "GPU ingest and the host parser disagree"
    out["bit_identical_to_host_parser"] = True
t = time.time()
subprocess.run([cli, fa, fb, fx, "--dims", pre + "_dimensions.txt", "--stats"], check=True, stdout=subprocess.DEVNULL)
out["cli_wall_seconds_files_1gpu"] = round(time.time() - t, 3)
print(json.dumps(out))
