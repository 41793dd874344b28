"""The reference's CPU programs timed beside the GPU path on the same box and the same seeded systems
(SURVEY.md 8d "CPU baseline timed beside it").  n = 2048: the REAL reference binaries (oracle/_ref, built
by oracle/build_ref.sh from /root/reference) where present, else the bit-identical port; n = 8192 (the
reference's largest published size, results.xlsx: serial CG 0.562 s, 8 ranks 0.234 s): the port.
Writes a markdown table to stdout."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import _oracle as O
import conjugate_gradient_b200 as cg

SEED = 20170209
cores = os.cpu_count()
print(f"host cores: {cores}; CPU model: " + next((l.split(':')[1].strip() for l in open('/proc/cpuinfo') if l.startswith('model name')), '?'))
print()
print("| n | program | precision | ranks/threads | CG seconds (converging solve) | iterations | it/s |")
print("|---|---|---|---|---|---|---|")
for n in (2048, 8192):
    A, b = O.gen_A(n, SEED), O.gen_b(n, SEED)
    for prec, dt in (("f32", np.float32), ("f64", np.float64)):
        Ap, bp = A.astype(dt), b.astype(dt)
        t = time.perf_counter(); x, st = O.conjugrad(Ap, bp, np.zeros(n, dt)); el = time.perf_counter() - t
        print(f"| {n} | serialConjugate.c conjugrad (port, 64-bit indices) | {prec} | 1 | {st['loop_seconds']:.4f} | {st['iterations']} | {st['iterations']/st['loop_seconds']:.1f} |")
        if O.ref_available("serial", prec, n):
            t = time.perf_counter(); O.ref_conjugrad(Ap, bp, np.zeros(n, dt)); el = time.perf_counter() - t
            print(f"| {n} | serialConjugate.c conjugrad (REAL reference, oracle/_ref; wall incl. its 6 mallocs) | {prec} | 1 | {el:.4f} | {st['iterations']} | {st['iterations']/el:.1f} |")
    for ranks in (2, 4, 8, 16):
        if ranks > cores or n % ranks:
            continue
        _, st = O.team_solve_generated(n, SEED, ranks)
        print(f"| {n} | parallel_cg.c conjugrad, per-rank loops on threads (port) | f64 | {ranks} | {st['loop_seconds']:.4f} | {st['iterations']} | {st['iterations']/st['loop_seconds']:.1f} |")
    with cg.Context(0) as ctx:
        ctx.generate(n, SEED)
        ctx.solve()
        ctx.set_x0(np.zeros(n))
        st = ctx.solve()
        print(f"| {n} | **libcgb200, 1 B200, device-resident** | f64 | 1 GPU | {st['solve_seconds']:.6f} | {st['iterations']} | {st['iterations']/(st['loop_ms']*1e-3):.0f} |")
        hA = np.ascontiguousarray(A); x = np.zeros(n)
        t = time.perf_counter(); st = ctx.conjugrad_rank(hA, b, x); el = time.perf_counter() - t
        print(f"| {n} | **libcgb200, 1 B200, host buffers in/out (malloc'ed), FIRST call of the context (pins its staging ring)** | f64 | 1 GPU | {el:.4f} | {st['iterations']} | {st['iterations']/el:.1f} |")
        x[:] = 0
        t = time.perf_counter(); st = ctx.conjugrad_rank(hA, b, x); el = time.perf_counter() - t
        print(f"| {n} | **libcgb200, 1 B200, host buffers in/out (malloc'ed), second call** | f64 | 1 GPU | {el:.4f} | {st['iterations']} | {st['iterations']/el:.1f} |")
