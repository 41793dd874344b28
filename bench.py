#!/usr/bin/env python
"""bench.py -- CG iterations/sec and GEMV GB/s, FP64 dense SPD, on N B200s.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` (N > 1 under
torchrun, one rank per GPU) prints ONE JSON line on rank 0 and exits non-zero if
the parity block of that line failed.

A *step* is one conjugate-gradient iteration (matVec + fused vector phase +, for
N > 1, the p all-gather and the two scalar all-reduces) of the seeded synthetic
SPD system of order n (default n = 65536, BASELINE.json configs[3], the size the
metric is quoted on; A is row-sharded over the N GPUs, so scaling is "strong").
Iterations are run in the solver's fixed-iteration mode (the stop rule is
evaluated and recorded but not acted on) because generateSPDmatrix-style
matrices converge in ~5 iterations.

  value    K / (CUDA-event time of the K-iteration loop), A resident in HBM,
           max over ranks; the loop is replayed from the captured CUDA graph, as
           cgb_solve() runs it by default.  `roofline` comes from an instrumented
           pass of the same K iterations right after it (one launch per kernel, an
           event pair around every GEMV): `instrumented_pass` is that pass's it/s.
  e2e      the same K iterations through the reference-facing C-ABI call
           cgb_conjugrad_rank() on MALLOC'ED host buffers -- what a caller shaped
           like the reference's main() passes (memAllocate(), serialConjugate.c:
           62-64): upload of A, b, x0 through the library's pinned staging ring,
           init, K iterations, download of x -- wall clock, max over ranks.
           `e2e_pinned` is the same call on cudaHostAlloc'ed buffers (the DMA
           ceiling of this box) and `e2e.h2d_frac_of_pinned` relates the two.
  parity   rank 0 reads x back after K' = min(K, 10) fixed iterations and after a
           converging solve and compares both with the CPU oracle's row-block SPMD
           form on the same generated system (parallel_cg.c:283-323): relative
           error of x <= 1e-10, iteration count within +-1, residual below 1e-6.
           Runs at every N, so a SCALE run proves the row-sharded exchange right.
  cpu_baseline / --impl reference
           the reference's row-block MPI algorithm (parallel_cg.c:248-345) on the
           host cores, as the 64-bit-index port in oracle/ (kind "port"): the real
           source's `int` indexing cannot address n > 46340.  `cpu_baseline.rows`
           (N = 1) adds the REAL reference binaries of oracle/_ref at n = 8192, the
           largest size the reference published: serialConjugate.c in FP32 as shipped
           and in FP64 on one core, parallel_cg.c on 2/4/8 ranks over the MPI shim,
           each timed by the program's own timer (serialConjugate.c:208,249;
           parallel_cg.c:278-279,328-329).
  configs1_n4096, configs2_n16384, configs4_n131072
           the other BASELINE.json sizes that fit, device-resident, same K-iteration
           fixed run (n = 4096: 2000 iterations from the captured CUDA graph).
  symmetric_variant, fused_reductions_variant
           the two opt-in variants of SURVEY.md 8f-4 beside the headline -- the symmetric product declares its own
           algorithmic bytes and carries its own `e2e` (asserted-symmetric upload).
  config   names the workload only and is identical in both arms; what each arm
           ran it with is under `implementation`.
"""
import argparse
import json
import os
import shutil
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SEED = 20170209
METRIC = "cg_iterations_per_sec_fp64_dense_spd"
UNIT = "iterations/s"
PARITY_TOL = 1.0e-10
EPS = 1.0e-6


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--matrix-n", dest="n", type=int, default=int(os.environ.get("CGB_BENCH_N", 65536)))
    ap.add_argument("--gemv", type=int, default=int(os.environ.get("CGB_BENCH_GEMV", 0)))
    ap.add_argument("--comm", default=os.environ.get("CGB_BENCH_COMM", "auto"), choices=["auto", "nccl", "peer"],
                    help="N > 1: per-iteration exchange (auto = fused peer-memory pushes when the GPUs map each other)")
    ap.add_argument("--rows", type=int, default=0, help="GEMV rows per CTA (0 = library default)")
    ap.add_argument("--unroll", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--l2-keep", type=int, default=0, help="small n: %% of L2 holding rows of A across iterations (0 default, -1 off)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip cpu_baseline (the parity block still runs the oracle)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs1/2/4 blocks")
    ap.add_argument("--no-variants", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0)
    ap.add_argument("--cpu-rows", action="store_true", help=argparse.SUPPRESS)     # child process of the N = 1 line
    return ap.parse_args()


def peak_hbm_gbs():
    """Roofline denominator: the driver-measured copy bandwidth, else the recipe's fallback."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(n, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per GEMV launch from the committed `ncu --set full`
    capture of this command (profiles/ncu_traffic.json, written from the .ncu-rep), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
        e = t.get(f"n{n}_g{world}")
        return (float(e["bytes"]), e["source"]) if e else (None, None)
    except Exception:
        return None, None


def workload_name(n):
    """BASELINE.json names four synthetic sizes; anything else is an off-list size asked for with --matrix-n."""
    idx = {4096: 1, 16384: 2, 65536: 3, 131072: 4}.get(n)
    return f"seeded synthetic SPD n={n} FP64, fixed-iteration CG ({'configs[%d]' % idx if idx else 'off-list size'})"


def workload_config(n, world):
    """`config` of the JSON line: the workload only, identical in both arms (--impl b200 / reference)."""
    m = n // max(world, 1)
    return {"workload": workload_name(n), "n": n, "seed": SEED, "generator": "A = 0.5 (R + R') + n I, b = rand, x0 = 0 (generateSPDmatrix.m:4-17), splitmix64 counter stream",
            "mode": "fixed iteration count (stop rule evaluated, not acted on)",
            "l2": (f"A is {8.0 * m * n / 1e9:.2f} GB per GPU at --gpus {world}, far larger than the 126 MB L2: no flush needed"
                   if 8.0 * m * n > 3 * 126e6 else
                   f"A is {8.0 * m * n / 1e6:.0f} MB per GPU at --gpus {world}, within a few L2 sizes: every iteration re-reads the same "
                   "matrix, so L2 residency is part of the workload (DESIGN.md 4.2) and is NOT flushed")}


def gemv_bytes(m, n):
    """Algorithmic bytes of one row-block GEMV launch: A once, p once, Ap once."""
    return 8.0 * (m * n + n + m)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc = None
        self.path = f"/tmp/cgb_clocks_{os.getpid()}.csv"
        if shutil.which("nvidia-smi"):
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=self.f, stderr=subprocess.DEVNULL)
            t_end = time.time() + 10.0                  # nvidia-smi takes ~1 s to deliver its first sample
            while time.time() < t_end and os.path.getsize(self.path) == 0:
                time.sleep(0.05)
                self.f.flush()

    @staticmethod
    def _ts(s):
        import datetime
        return datetime.datetime.strptime(s.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()

    def mark(self):
        """Wall-clock mark; samples between two marks are the ones taken during the timed region."""
        return time.time()

    def stop(self, t_lo=None, t_hi=None):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [t.strip() for t in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                ts = self._ts(parts[0])
                if t_lo is not None and not (t_lo - 0.03 <= ts <= t_hi + 0.03):
                    continue
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        try:
            os.remove(self.path)
        except OSError:
            pass
        return out


# --------------------------------------------------------------------------- CPU arm
def cpu_threads_for(n, want=0):
    cores = os.cpu_count() or 1
    p = 1
    while p * 2 <= (want or cores) and n % (p * 2) == 0:
        p *= 2
    return p, cores


def run_cpu_port(n, iters, threads, warmup=0, want_x=False, ignore_stop=True):
    """The reference's row-block SPMD algorithm (parallel_cg.c:283-323) on host threads:
    oracle/cg_oracle.c orc_team_solve_generated -- one thread per rank, the same
    per-rank loops, rank-ordered all-reduce.  Returns (iterations/s, stats, x or None)."""
    import _oracle as O
    if warmup > 0:
        O.team_solve_generated(n, SEED, threads, ignore_stop=True, max_iter=warmup, want_x=False)
    x, st = O.team_solve_generated(n, SEED, threads, ignore_stop=ignore_stop, max_iter=iters, want_x=want_x)
    return st["iterations"] / st["loop_seconds"], st, x


REFERENCE_BUDGET_S = 150.0


def reference_iterations(n, threads, steps, gbs_per_thread=5.0):
    """Iterations the CPU arm runs for --steps K: K itself unless streaming 8 n^2 bytes K times
    through `threads` cores (a conservative 5 GB/s each; measured 6.9-7.3 on the bench box)
    would exceed REFERENCE_BUDGET_S -- then the largest count that fits (at least 3)."""
    per_iter = 8.0 * n * n / (threads * gbs_per_thread * 1e9)
    cap = max(3, int(REFERENCE_BUDGET_S / max(per_iter, 1e-9)))
    return min(steps, cap)


def reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host
    cores, same metric/config.  Rank 0 only; other ranks exit 0 without work."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    n = args.n
    threads, cores = cpu_threads_for(n, args.cpu_threads)
    # bounded: every step streams 8*n*n bytes through the host cores (~0.3 s at n = 65536 on 16
    # threads), so a very large --steps is cut to what fits REFERENCE_BUDGET_S and reported as such
    t0 = time.time()
    asked, W = max(1, args.steps), max(0, args.warmup)
    iters = reference_iterations(n, threads, asked)
    w_run = min(W, reference_iterations(n, threads, max(W, 1))) if W > 0 else 0
    val, st, _ = run_cpu_port(n, iters, threads, warmup=w_run)
    m = n // threads
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": asked, "warmup": W, "ms_per_step": 1e3 / val, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(n, args.gpus),
        "implementation": {"parallelism": f"{threads} host threads, row-block SPMD (parallel_cg.c)",
                           "steps_run": iters, "warmup_run": w_run},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"full n={n} matrix ({8.0 * n * n / 1e9:.1f} GB), {iters} iterations"
                                   + (f" (of the {asked} asked for: bounded to ~{REFERENCE_BUDGET_S:.0f} s of CPU work)" if iters < asked else "") + ", "
                                   f"{threads} of {cores} host cores; port because the reference's int "
                                   "indexing overflows above n=46340 (serialConjugate.c:78,117); the port is "
                                   "bit-identical to oracle/_ref at n<=16384 (tests/test_oracle.py)",
                         "gemv_gbs_per_rank": gemv_bytes(m, n) * val / 1e9,
                         "gen_seconds": st["gen_seconds"]},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_seconds": time.time() - t0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---- the REAL reference binaries (oracle/_ref), timed by their own timers: child process of the N = 1 line
class CaptureStdoutFd:
    """File-descriptor-level capture of stdout: the reference's conjugrad() printf()s its own timing
    (serialConjugate.c:249-251, parallel_cg.c:334-335), from this process or from a forked rank."""

    def __enter__(self):
        import ctypes
        import tempfile
        sys.stdout.flush()
        self._libc = ctypes.CDLL(None)
        self._saved = os.dup(1)
        self._tmp = tempfile.TemporaryFile()
        os.dup2(self._tmp.fileno(), 1)
        self.text = ""
        return self

    def __exit__(self, *exc):
        self._libc.fflush(None)
        os.dup2(self._saved, 1)
        os.close(self._saved)
        self._tmp.seek(0)
        self.text = self._tmp.read().decode("utf-8", "replace")
        self._tmp.close()
        return False


def reference_timer_seconds(text):
    """The seconds the reference printed: '... execution time in seconds: %f' (first such line)."""
    import re
    m = re.search(r"(?:average clock|cg method) execution time in seconds:\s*([0-9.eE+-]+)", text)
    return float(m.group(1)) if m else None


def cpu_rows_child(n=8192):
    """Runs in a fresh interpreter (no CUDA context to fork): the real serialConjugate.c conjugrad()
    in FP32 as shipped and in FP64 on one core, and the real parallel_cg.c conjugrad() on 2/4/8 ranks
    over the MPI shim, on the generated system of order n.  Prints one JSON list."""
    import numpy as np
    import _oracle as O
    rows = []
    cores = os.cpu_count() or 1
    A, b = O.gen_A(n, SEED), O.gen_b(n, SEED)
    for prec, dt in (("f32", np.float32), ("f64", np.float64)):
        if not O.ref_available("serial", prec, n):
            continue
        Ap, bp = A.astype(dt), b.astype(dt)
        _, so = O.conjugrad(Ap, bp, np.zeros(n, dt))                 # bit-identical port: the iteration count
        t0 = time.perf_counter()
        with CaptureStdoutFd() as cap:
            O.ref_conjugrad(Ap, bp, np.zeros(n, dt))
        wall = time.perf_counter() - t0
        sec = reference_timer_seconds(cap.text)
        rows.append({"name": f"serial_{prec}_1core", "program": "serialConjugate.c conjugrad()" + (" as shipped (float)" if prec == "f32" else " float->double"),
                     "kind": "reference", "n": n, "cores": 1, "precision": prec, "iterations": so["iterations"],
                     "seconds": sec if sec else wall, "timer": "the program's own clock() (serialConjugate.c:208,249)" if sec else "wall clock around the call",
                     "value": so["iterations"] / (sec if sec else wall), "unit": UNIT})
        del Ap, bp
    if O.ref_available("parallel", "f64", n):
        for ranks in (2, 4, 8):
            if ranks > cores or n % ranks:
                continue
            _, so = O.conjugrad(A, b, np.zeros(n), ranks=ranks)
            t0 = time.perf_counter()
            with CaptureStdoutFd() as cap:
                O.ref_conjugrad(A, b, np.zeros(n), ranks=ranks)
            wall = time.perf_counter() - t0
            sec = reference_timer_seconds(cap.text)
            rows.append({"name": f"spmd_np{ranks}", "program": "parallel_cg.c conjugrad() float->double, ranks = processes over oracle/mpi_shim",
                         "kind": "reference", "n": n, "cores": ranks, "precision": "f64", "iterations": so["iterations"],
                         "seconds": sec if sec else wall, "timer": "the program's own MPI_Wtime between barriers (parallel_cg.c:278-279,328-329)" if sec else "wall clock around the call",
                         "value": so["iterations"] / (sec if sec else wall), "unit": UNIT})
    print(json.dumps(rows), flush=True)
    return 0


def cpu_reference_rows():
    """The child above, bounded; [] when oracle/_ref has no n = 8192 build."""
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-rows"], capture_output=True, text=True, timeout=240)
        lines = [l for l in r.stdout.splitlines() if l.startswith("[")]
        return json.loads(lines[-1]) if r.returncode == 0 and lines else []
    except Exception:
        return []


# --------------------------------------------------------------------------- B200 arm
def b200_arm(args):
    import numpy as np
    import torch
    import conjugate_gradient_b200 as cg

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
        args.gpus = world
    n, K, W = args.n, max(1, args.steps), max(0, args.warmup)
    if n % world != 0:
        raise SystemExit(f"{n} is not divisible by {world}")
    m = n // world
    if cg.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; libcgb200 has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    uid = None
    if world > 1:
        import datetime
        import torch.distributed as dist
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world, timeout=datetime.timedelta(minutes=30))
        box = [cg.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return float(v)
        t = torch.tensor([float(v)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ctx = cg.Context(device=local, rank=rank, ranks=world, uid=uid)
    info = ctx.device_info()
    ctx.generate(n, SEED)                                   # each GPU writes its own row block in HBM
    gopts = dict(gemv_variant=args.gemv, comm={"auto": 0, "nccl": 1, "peer": 2}[args.comm],
                 gemv_rows=args.rows, gemv_unroll=args.unroll, gemv_threads=args.threads,
                 gemv_l2_keep_pct=args.l2_keep)

    # ---- warm-up: W untimed iterations (also builds plans, first-touches everything), then one untimed
    # pass of the K-iteration graph run so that capture / instantiation / first upload of the graph are out
    if W > 0:
        ctx.solve(ignore_stop=1, max_iter=W, time_gemv=1, **gopts)
        ctx.set_x0(np.zeros(n))
        ctx.solve(ignore_stop=1, max_iter=K, **gopts)
        ctx.set_x0(np.zeros(n))
    # ---- timed region: exactly K iterations, device-resident, as cgb_solve() runs them by default -- replayed
    # from the captured CUDA graph; CUDA events on the solver stream around the loop, max over ranks
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    t_lo = time.time()
    stg = ctx.solve(ignore_stop=1, max_iter=K, **gopts)
    barrier()
    t_hi = time.time()
    clocks = sampler.stop(t_lo, t_hi) if sampler else None
    loop_ms = max_over_ranks(stg["loop_ms"])
    value = K / (loop_ms * 1e-3)
    # ---- the same K iterations once more, instrumented: launched one by one with an event pair around every
    # GEMV (event pairs cannot be read out of a replayed graph) -> the kernel's average duration for `roofline`
    ctx.set_x0(np.zeros(n))
    barrier()
    st = ctx.solve(ignore_stop=1, max_iter=K, time_gemv=1, **gopts)
    barrier()
    inst_loop_ms = max_over_ranks(st["loop_ms"])
    gemv_ms = max_over_ranks(st["gemv_ms_avg"])
    # ---- time to solution, parity mode (stop rule active), device-resident
    ctx.set_x0(np.zeros(n))
    barrier()
    stc = ctx.solve(**gopts)
    conv_seconds = max_over_ranks(stc["solve_seconds"])
    x_conv = ctx.read_x() if rank == 0 else None

    symmetric = fused = None
    if not getattr(args, "no_variants", False):
        # ---- opt-in variant: matVec from the upper tiles of the (exactly symmetric) matrix only.
        # Reported beside the headline, never as the headline: its algorithmic bytes are HALF the reference's.
        if n % (512 * world) == 0 and st["comm"] in (0, 2):
            ctx.set_x0(np.zeros(n))
            barrier()
            sts = ctx.solve(ignore_stop=1, max_iter=K, time_gemv=1, symmetric=1, **gopts)
            ctx.set_x0(np.zeros(n))
            barrier()
            stsg = ctx.solve(ignore_stop=1, max_iter=K, symmetric=1, **gopts)
            barrier()
            symmetric = {"value": K / (max_over_ranks(sts["loop_ms"]) * 1e-3), "unit": UNIT,
                         "graph_replay": K / (max_over_ranks(stsg["loop_ms"]) * 1e-3), "gemv_ms_avg": max_over_ranks(sts["gemv_ms_avg"]),
                         "declared_bytes_per_launch": sts["gemv_bytes"],
                         "achieved_gbs_on_declared_bytes": sts["gemv_bytes"] / (sts["gemv_ms_avg"] * 1e-3) / 1e9,
                         "note": "cgb_options.symmetric=1: 8(n(n+1)/2 + 2n) bytes per product instead of 8(n*n + 2n); "
                                 "exact symmetry is verified on the device before use"}
        # ---- opt-in recurrence: Chronopoulos-Gear CG, two kernels and one exposed reduction per iteration
        if st["comm"] in (0, 2):
            ctx.set_x0(np.zeros(n))
            barrier()
            stf = ctx.solve(ignore_stop=1, max_iter=K, cg_variant=1, **gopts)
            barrier()
            fused = {"value": K / (max_over_ranks(stf["loop_ms"]) * 1e-3), "unit": UNIT,
                     "note": "cgb_options.cg_variant=1, full product, graph replay; with symmetric=1 as well: see fused_symmetric"}
            if symmetric:
                ctx.set_x0(np.zeros(n))
                barrier()
                stfs = ctx.solve(ignore_stop=1, max_iter=K, cg_variant=1, symmetric=1, **gopts)
                barrier()
                fused["fused_symmetric"] = K / (max_over_ranks(stfs["loop_ms"]) * 1e-3)

    peak, peak_src = peak_hbm_gbs()
    gbytes = gemv_bytes(m, n)
    traffic, traffic_src = ncu_traffic(n, world)
    achieved = gbytes / (gemv_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": loop_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(n, world),
        "implementation": {
            "rows_per_gpu": m, "parallelism": f"row-block x{world}" if world > 1 else "single GPU",
            "exchange": {0: "none (one rank)", 1: "nccl all-gather + 2 all-reduce per iteration",
                         2: "fused peer-memory pushes over NVLink inside the solver kernels"}[st["comm"]],
            "launch": {"programmatic_dependent_launch": int(stg.get("pdl", 0)), "graph_iterations": stg["graph_chunk"]},
            "gemv_kernel": {k: st[k] for k in ("gemv_variant", "gemv_rows", "gemv_unroll", "gemv_threads",
                                               "gemv_vec", "gemv_stages", "gemv_tile_cols", "gemv_grid")}},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / 8000.0,
                     "kernel": "cgb_rowblock_matvec (A*p with fused p.Ap)", "bytes_per_launch": gbytes,
                     "gemv_ms_avg": gemv_ms, "gemv_ms_min": st["gemv_ms_min"], "launches_timed": st["gemv_timed"],
                     "gemv_share_of_step": gemv_ms * K / inst_loop_ms,
                     "measured_in": "the instrumented pass of the same K iterations (one launch per kernel, event pair around "
                                    "every GEMV; with the peer exchange the GEMV's time includes its wait for the peers' slices)"},
        "clocks": clocks,
        "gpu_launches": int(stg["kernel_launches"]) - 4,
        "timed_region": {"mode": "CUDA-graph replay, the default of cgb_solve()", "iterations_per_graph": stg["graph_chunk"],
                         "kernels_per_iteration": (int(stg["kernel_launches"]) - 4) / K},
        "instrumented_pass": {"value": K / (inst_loop_ms * 1e-3), "unit": UNIT, "ms_per_step": inst_loop_ms / K,
                              "kernels_per_iteration": (int(st["kernel_launches"]) - 4) / K},
        "time_to_solution": {"seconds": conv_seconds, "iterations": stc["iterations"],
                             "final_residual": stc["final_residual"], "mode": "device-resident, stop rule active"},
        "device": info["name"],
    }
    if symmetric:
        line["symmetric_variant"] = symmetric
    if fused:
        line["fused_reductions_variant"] = fused
    env = (rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist)
    return ctx, line, env, (stc, x_conv)


def relerr(x, ref):
    import numpy as np
    return float(np.linalg.norm(x - ref) / max(np.linalg.norm(ref), 1e-300))


def parity_leg(ctx, line, env, conv, args):
    """Every rank solves (the solve is collective); rank 0 compares with the CPU oracle's row-block SPMD
    form on the same generated system and also takes `cpu_baseline` from the fixed-iteration oracle run."""
    import numpy as np
    rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist = env
    Kp = min(K, 10)
    ctx.set_x0(np.zeros(n))
    barrier()
    stp = ctx.solve(ignore_stop=1, max_iter=Kp, **gopts)
    x_fix = ctx.read_x() if rank == 0 else None
    ok = True
    if rank == 0:
        stc, x_conv = conv
        threads, cores = cpu_threads_for(n, args.cpu_threads)
        iters = min(Kp, reference_iterations(n, threads, Kp))
        t0 = time.time()
        val, so_fix, xo_fix = run_cpu_port(n, iters, threads, want_x=True)
        _, so_conv, xo_conv = run_cpu_port(n, n, threads, want_x=True, ignore_stop=False)
        e_fix = relerr(x_fix, xo_fix) if iters == Kp else None
        d_it = int(stc["iterations"]) - int(so_conv["iterations"])
        e_conv = relerr(x_conv, xo_conv)
        res_gpu, res_cpu = float(stc["final_residual"]), float(np.sqrt(so_conv["final_beta"]))
        ok = (e_fix is None or e_fix <= PARITY_TOL) and abs(d_it) <= 1 and res_gpu < EPS and (d_it != 0 or e_conv <= PARITY_TOL)
        line["parity"] = {
            "ok": bool(ok), "tol": PARITY_TOL, "oracle": f"oracle/cg_oracle.c orc_team_solve_generated on {threads} host threads "
                                                        "(parallel_cg.c:283-323 loop for loop, 64-bit indices, rank-ordered all-reduce)",
            "fixed_run": {"iterations": iters, "rel_err_x": e_fix,
                          "first_hit_k": {"gpu": int(stp["first_hit_k"]), "oracle": int(so_fix["first_hit_k"])}},
            "converging_solve": {"rel_err_x": e_conv, "iterations": {"gpu": int(stc["iterations"]), "oracle": int(so_conv["iterations"])},
                                 "final_residual": {"gpu": res_gpu, "oracle": res_cpu, "eps": EPS}},
            "n": n, "ranks": world, "seconds": time.time() - t0}
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = {
                "value": val, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"full n={n} matrix, {iters} iterations after A*x0, {threads} of {cores} host cores "
                          f"(oracle/cg_oracle.c orc_team_solve_generated: parallel_cg.c's per-rank loops on threads)",
                "gemv_gbs_per_rank": gemv_bytes(n // threads, n) * val / 1e9,
                "gen_seconds": so_fix["gen_seconds"]}
    barrier()
    return ok


def e2e_leg(ctx, line, env):
    """K iterations through the reference-facing C-ABI entry point on HOST buffers
    (cgb_conjugrad_rank = parallel_cg.c's conjugrad(local_A, local_b, local_x, m, rank, np)):
    H2D of the rank's A block, b block and x0, init, K iterations, D2H of x.  Headline `e2e`: buffers from
    malloc, as the reference's main() allocates them; `e2e_pinned`: the same call on cudaHostAlloc'ed memory."""
    import ctypes as C
    import numpy as np
    import torch
    from conjugate_gradient_b200 import _binding as B
    rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist = env
    L = B.lib()
    pA = torch.empty(m * n, dtype=torch.float64, pin_memory=True)
    B._check(L.cgb_debug_read(ctx._h, 0, 0, m * n, pA.data_ptr()))      # the generated block, to the host
    hA = np.empty(m * n, dtype=np.float64)                               # malloc / mmap: pageable, like memAllocate()
    hA[:] = pA.numpy()
    hb = np.empty(m, dtype=np.float64)
    B._check(L.cgb_debug_read(ctx._h, 1, 0, m, hb.ctypes.data))
    hx = np.zeros(n, dtype=np.float64)
    st = B.Stats()
    h2d = 8.0 * (m * n + m + n)

    def call(a_ptr, opts):
        hx[:] = 0.0
        barrier()
        t0 = time.perf_counter()
        B._check(L.cgb_conjugrad_rank(ctx._h, a_ptr, hb.ctypes.data, hx.ctypes.data, m, C.byref(opts), C.byref(st)))
        torch.cuda.synchronize()
        return max_over_ranks(time.perf_counter() - t0)

    def measure(a_ptr, opts, what):
        call(a_ptr, opts)                                          # warm-up call (staging ring, graphs)
        dt = call(a_ptr, opts)
        h2d_s = max_over_ranks(st.h2d_seconds)
        return {"value": K / dt, "unit": UNIT, "h2d_bytes_per_step": float(st.h2d_bytes) / K, "d2h_bytes_per_step": 8.0 * n / K,
                "seconds_per_call": dt, "h2d_seconds": h2d_s, "h2d_gbs_per_gpu": float(st.h2d_bytes) / max(h2d_s, 1e-12) / 1e9,
                "solve_seconds_this_rank": float(st.solve_seconds), "iterations": int(st.iterations),
                "h2d_path": {0: "none", 1: "direct DMA from pinned memory", 2: "staged: host threads pack a ring of pinned slots (cg_stage.c)",
                             3: "driver's pageable path", 4: "registered in place"}.get(int(st.h2d_path), "?"),
                "api": what}

    o = B.make_options(ignore_stop=1, max_iter=K, **gopts)
    e2e = measure(hA.ctypes.data, o, "cgb_conjugrad_rank (C ABI) on malloc'ed host buffers, one upload of A per call")
    pin = measure(pA.data_ptr(), o, "cgb_conjugrad_rank (C ABI) on cudaHostAlloc'ed host buffers")
    e2e["h2d_frac_of_pinned"] = e2e["h2d_gbs_per_gpu"] / max(pin["h2d_gbs_per_gpu"], 1e-12)
    e2e["note"] = ("upload-bound: seconds_per_call ~ h2d_seconds + the device-resident solve; e2e_pinned is this box's DMA ceiling for "
                   f"{world} concurrent upload(s), h2d_frac_of_pinned relates the malloc'ed path to it")
    assert abs(h2d - float(st.h2d_bytes)) < 1.0
    line["e2e"], line["e2e_pinned"] = e2e, pin
    if "symmetric_variant" in line:
        # the same call with the matrix asserted symmetric (symmetric = 2): the tiles the product
        # never reads stay on the host, so about half the bytes cross PCIe
        o3 = B.make_options(ignore_stop=1, max_iter=K, symmetric=2, **gopts)
        line["symmetric_variant"]["e2e"] = measure(hA.ctypes.data, o3, "cgb_conjugrad_rank with cgb_options.symmetric = 2, malloc'ed host buffers")
    # the converging solve a user of the reference CLI runs: upload + ~5 iterations + download
    dt2 = call(hA.ctypes.data, B.make_options(**gopts))
    line["time_to_solution"]["e2e_seconds_host_buffers"] = dt2
    line["time_to_solution"]["e2e_iterations"] = int(st.iterations)
    del hA, pA, hb, hx


def fixed_run_block(ctx, env, n2, K2, graph_only=False, extra=None):
    """Device-resident K2-iteration run of another BASELINE.json size on the same context."""
    import numpy as np
    rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist = env
    m2 = n2 // world
    ctx.generate(n2, SEED)
    ctx.solve(ignore_stop=1, max_iter=3, time_gemv=1, **gopts)
    ctx.set_x0(np.zeros(n2))
    barrier()
    stt = ctx.solve(ignore_stop=1, max_iter=min(K2, 1000), time_gemv=1, **gopts)
    ctx.set_x0(np.zeros(n2))
    barrier()
    stg = ctx.solve(ignore_stop=1, max_iter=K2, **gopts)
    barrier()
    peak, _ = peak_hbm_gbs()
    gms = max_over_ranks(stt["gemv_ms_avg"])
    loop = max_over_ranks(stg["loop_ms"])
    out = {"workload": workload_name(n2), "n": n2, "rows_per_gpu": m2, "steps": K2,
           "value": K2 / (loop * 1e-3), "unit": UNIT, "ms_per_step": loop / K2, "timing": "CUDA-graph replay, CUDA events, max over ranks",
           "stream_mode_value": min(K2, 1000) / (max_over_ranks(stt["loop_ms"]) * 1e-3),
           "gemv_ms_avg": gms, "gemv_gbs": gemv_bytes(m2, n2) / (gms * 1e-3) / 1e9,
           "roofline_frac": gemv_bytes(m2, n2) / (gms * 1e-3) / 1e9 / peak,
           "gemv_keep_rows": int(stg.get("gemv_keep_rows", 0)), "pdl": int(stg.get("pdl", 0)), "graph_iterations": stg["graph_chunk"]}
    if extra:
        out.update(extra(ctx, stg, loop))
    return out


def configs_leg(ctx, line, env, args):
    """The other BASELINE.json sizes on the same GPUs: n = 131072 (the north star's scaling
    target is quoted on it), n = 16384 (configs[2]) and, on one GPU, n = 4096 (latency-bound, graph path)."""
    import numpy as np
    rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist = env
    info = ctx.device_info()
    if n != 131072 and 131072 % world == 0 and 8.0 * 131072 * 131072 / world + (6 << 30) <= info["total_bytes"]:
        line["configs4_n131072"] = fixed_run_block(ctx, env, 131072, min(K, 20))
    if n != 16384 and 16384 % world == 0:
        line["configs2_n16384"] = fixed_run_block(ctx, env, 16384, 200)
    if n != 4096 and world == 1:
        def extra(c, stg, loop):
            # every iteration streams 8 (n^2 + 2n) algorithmic bytes; what exceeds the HBM peak came out of L2
            peak, _ = peak_hbm_gbs()
            eff = gemv_bytes(4096, 4096) / (loop / 2000 * 1e-3) / 1e9
            c.set_x0(np.zeros(4096))
            stf = c.solve(ignore_stop=1, max_iter=2000, cg_variant=1, **gopts)
            return {"us_per_iteration": loop / 2000 * 1e3, "effective_gbs_whole_iteration": eff,
                    "kept_in_l2_fraction_planned": int(stg.get("gemv_keep_rows", 0)) / 4096.0,
                    "min_fraction_of_bytes_from_l2": max(0.0, 1.0 - peak / eff) if eff > 0 else None,
                    "fused_reductions_value": 2000 / (stf["loop_ms"] * 1e-3)}
        line["configs1_n4096"] = fixed_run_block(ctx, env, 4096, 2000, extra=extra)


def main():
    args = parse_args()
    if args.cpu_rows:
        return cpu_rows_child()
    if args.impl == "reference":
        return reference_arm(args)
    ctx, line, env, conv = b200_arm(args)
    rank, world = env[0], env[1]
    ok = True
    if not args.no_parity:
        ok = parity_leg(ctx, line, env, conv, args)
    if not args.no_e2e:
        e2e_leg(ctx, line, env)
    if not args.no_configs:
        configs_leg(ctx, line, env, args)
    ctx.close()
    if rank == 0 and world == 1 and not args.no_cpu and "cpu_baseline" in line:
        rows = cpu_reference_rows()
        if rows:
            line["cpu_baseline"]["rows"] = rows
            line["cpu_baseline"]["rows_note"] = ("the REAL reference binaries (oracle/_ref: the reference's own sources through sed for ROWS/COLS and "
                                                 "float->double, gcc -O2) on the generated system of order 8192, the largest size the reference "
                                                 "published (results.xlsx: serial 0.562 s, 8 ranks 0.234 s); converging solves, it/s = iterations / "
                                                 "the program's own timer")
    if env[9] is not None:
        env[9].barrier()
        env[9].destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
        if not ok:
            print("bench.py: PARITY FAILED against the CPU oracle -- see the line's `parity` block", file=sys.stderr)
            return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
