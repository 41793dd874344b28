#!/usr/bin/env python
"""bench.py -- CG iterations/sec and GEMV GB/s, FP64 dense SPD, on N B200s.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` (N > 1 under
torchrun, one rank per GPU) prints ONE JSON line on rank 0.

A *step* is one conjugate-gradient iteration (matVec + fused vector phase +, for
N > 1, the p all-gather and the two scalar all-reduces) of the seeded synthetic
SPD system of order n (default n = 65536, BASELINE.json configs[3], the size the
metric is quoted on; A is row-sharded over the N GPUs, so scaling is "strong").
Iterations are run in the solver's fixed-iteration mode (the stop rule is
evaluated and recorded but not acted on) because generateSPDmatrix-style
matrices converge in ~5 iterations.

  value    K / (CUDA-event time of the K-iteration loop), A resident in HBM,
           max over ranks.  The timed run records an event pair around every
           GEMV launch, which gives `roofline` live from the same region.
  e2e      the same K iterations through the reference-facing C-ABI call
           cgb_conjugrad_rank() on pinned HOST buffers: upload of A, b, x0,
           init, K iterations, download of x -- wall clock, max over ranks.
  cpu_baseline / --impl reference
           the reference's row-block MPI algorithm (parallel_cg.c:248-345) on the
           host cores, as the 64-bit-index port in oracle/ (kind "port"): the real
           source's `int` indexing cannot address n > 46340, and it can only run
           to convergence (~4 iterations, A*x0 and its mallocs inside its timer),
           not K fixed iterations.  The port is bit-identical to the real build
           up to n = 16384 (tests/test_oracle.py); the real binaries of oracle/_ref
           are timed beside it in tools/cpu_vs_gpu_table.py (profiles/r01_cpu_vs_gpu.md).
  graph_replay, symmetric_variant, fused_reductions_variant
           the same K iterations replayed from the captured CUDA graph (what a
           converging solve runs), and the two opt-in variants of SURVEY.md 8f-4
           beside the headline -- the symmetric product declares its own
           algorithmic bytes and carries its own `e2e` (asserted-symmetric upload).
  config.launch
           whether the iteration kernels used programmatic dependent launch, and
           the iterations per graph.
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
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0)
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


def run_cpu_port(n, iters, threads, warmup=0):
    """The reference's row-block SPMD algorithm (parallel_cg.c:283-323) on host threads:
    oracle/cg_oracle.c orc_team_solve_generated -- one thread per rank, the same
    per-rank loops, rank-ordered all-reduce.  Returns (iterations/s, stats)."""
    import _oracle as O
    if warmup > 0:
        O.team_solve_generated(n, SEED, threads, ignore_stop=True, max_iter=warmup, want_x=False)
    _, st = O.team_solve_generated(n, SEED, threads, ignore_stop=True, max_iter=iters, want_x=False)
    return st["iterations"] / st["loop_seconds"], st


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
    iters, asked = reference_iterations(n, threads, max(1, args.steps)), max(1, args.steps)
    val, st = run_cpu_port(n, iters, threads, warmup=min(args.warmup, 1))
    m = n // threads
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": iters, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 / val, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(n), "n": n,
                   "seed": SEED, "parallelism": f"{threads} host threads, row-block SPMD (parallel_cg.c)"},
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


# --------------------------------------------------------------------------- B200 arm
def b200_arm(args):
    import numpy as np
    import torch
    import conjugate_gradient_b200 as cg
    from conjugate_gradient_b200 import _binding as B

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
        import torch.distributed as dist
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world)
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

    # ---- warm-up: W untimed iterations (also builds plans, first-touches everything)
    if W > 0:
        ctx.solve(ignore_stop=1, max_iter=W, time_gemv=1, **gopts)
        ctx.set_x0(np.zeros(n))
    # ---- timed region: exactly K iterations, device-resident, event pair around every GEMV
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    t_lo = time.time()
    st = ctx.solve(ignore_stop=1, max_iter=K, time_gemv=1, **gopts)
    barrier()
    t_hi = time.time()
    clocks = sampler.stop(t_lo, t_hi) if sampler else None
    loop_ms = max_over_ranks(st["loop_ms"])
    gemv_ms = max_over_ranks(st["gemv_ms_avg"])
    value = K / (loop_ms * 1e-3)
    # ---- the same K iterations replayed from the captured CUDA graph (what a solve uses)
    ctx.set_x0(np.zeros(n))
    barrier()
    stg = ctx.solve(ignore_stop=1, max_iter=K, **gopts)
    barrier()
    graph_value = K / (max_over_ranks(stg["loop_ms"]) * 1e-3)
    # ---- time to solution, parity mode (stop rule active), device-resident
    ctx.set_x0(np.zeros(n))
    barrier()
    stc = ctx.solve(**gopts)
    conv_seconds = max_over_ranks(stc["solve_seconds"])

    # ---- opt-in variant: matVec from the upper tiles of the (exactly symmetric) matrix only.
    # Reported beside the headline, never as the headline: its algorithmic bytes are HALF the reference's.
    symmetric = None
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
    fused = None
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
        "config": {"workload": workload_name(n), "n": n,
                   "seed": SEED, "rows_per_gpu": m, "parallelism": f"row-block x{world}" if world > 1 else "single GPU",
                   "exchange": {0: "none (one rank)", 1: "nccl all-gather + 2 all-reduce per iteration",
                                2: "fused peer-memory pushes over NVLink inside the solver kernels"}[st["comm"]],
                   "l2": (f"A block is {8.0 * m * n / 1e9:.2f} GB per GPU, far larger than the 126 MB L2: no flush needed"
                          if 8.0 * m * n > 3 * 126e6 else
                          f"A block is {8.0 * m * n / 1e6:.0f} MB per GPU, within a few L2 sizes: every iteration re-reads the same "
                          "matrix, so L2 residency is part of the workload (DESIGN.md 4.2) and is NOT flushed"),
                   "launch": {"programmatic_dependent_launch": int(stg.get("pdl", 0)), "graph_iterations": stg["graph_chunk"]},
                   "gemv_kernel": {k: st[k] for k in ("gemv_variant", "gemv_rows", "gemv_unroll", "gemv_threads",
                                                      "gemv_vec", "gemv_stages", "gemv_tile_cols", "gemv_grid")}},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / 8000.0,
                     "kernel": "gemv (A*p with fused p.Ap)", "bytes_per_launch": gbytes,
                     "gemv_ms_avg": gemv_ms, "gemv_ms_min": st["gemv_ms_min"], "launches_timed": st["gemv_timed"],
                     "gemv_share_of_step": gemv_ms * K / loop_ms},
        "clocks": clocks,
        "gpu_launches": int(st["kernel_launches"]) - 4,
        "graph_replay": {"value": graph_value, "unit": UNIT, "iterations_per_graph": stg["graph_chunk"]},
        "time_to_solution": {"seconds": conv_seconds, "iterations": stc["iterations"],
                             "final_residual": stc["final_residual"], "mode": "device-resident, stop rule active"},
        "device": info["name"],
    }
    if symmetric:
        line["symmetric_variant"] = symmetric
    if fused:
        line["fused_reductions_variant"] = fused
    return ctx, line, (rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist)


def parse_cpulist(txt):
    """sysfs cpulist ("0-15,32-47") -> set of CPU numbers; empty set for anything malformed."""
    out = set()
    try:
        for part in txt.strip().split(","):
            if not part:
                continue
            lo, _, hi = part.partition("-")
            out.update(range(int(lo), int(hi or lo) + 1))
    except ValueError:
        return set()
    return out


class NumaLocal:
    """User-side placement of the caller's host buffers: while they are allocated and pinned (first touch
    decides the NUMA node of the pages) the process runs on the CPUs next to its GPU -- sysfs `local_cpulist`
    of the GPU's PCI function -- so that every rank's upload crosses its own PCIe root and not the socket
    interconnect (8 ranks uploading at once measured 23 GB/s each against 55 GB/s alone).  The previous
    affinity is restored on exit; any missing piece (no sysfs entry, a VM without topology, a container
    cpuset that excludes those CPUs) leaves the process as it was."""

    def __init__(self, pci_domain, pci_bus, pci_device, sysfs="/sys/bus/pci/devices"):
        self.path = os.path.join(sysfs, f"{pci_domain:04x}:{pci_bus:02x}:{pci_device:02x}.0", "local_cpulist")
        self.old = None
        self.bound = 0

    def __enter__(self):
        if os.environ.get("CGB_BENCH_NUMA", "1") == "0":          # A/B switch: default placement
            return self
        try:
            with open(self.path) as f:
                local = parse_cpulist(f.read())
            cur = os.sched_getaffinity(0)
            want = local & cur
            if want and want != cur:
                os.sched_setaffinity(0, want)
                self.old, self.bound = cur, len(want)
        except (OSError, AttributeError, ValueError):
            self.old = None
        return self

    def __exit__(self, *exc):
        if self.old is not None:
            try:
                os.sched_setaffinity(0, self.old)
            except OSError:
                pass
        return False


def e2e_leg(ctx, line, env):
    """K iterations through the reference-facing C-ABI entry point on HOST buffers
    (cgb_conjugrad_rank = parallel_cg.c's conjugrad(local_A, local_b, local_x, m, rank, np)):
    H2D of the rank's A block, b block and x0 from pinned memory, init, K iterations, D2H of x."""
    import ctypes as C
    import numpy as np
    import torch
    from conjugate_gradient_b200 import _binding as B
    rank, world, local, n, m, K, barrier, max_over_ranks, gopts, dist = env
    L = B.lib()
    props = torch.cuda.get_device_properties(local)
    with NumaLocal(getattr(props, "pci_domain_id", 0), getattr(props, "pci_bus_id", 0), getattr(props, "pci_device_id", 0)) as numa:
        hA = torch.empty(m * n, dtype=torch.float64, pin_memory=True)
        hb = torch.empty(m, dtype=torch.float64, pin_memory=True)
        hx = torch.zeros(n, dtype=torch.float64, pin_memory=True)
        B._check(L.cgb_debug_read(ctx._h, 0, 0, m * n, hA.data_ptr()))  # the generated block, to the host (first touch)
        B._check(L.cgb_debug_read(ctx._h, 1, 0, m, hb.data_ptr()))
        cpus_bound = numa.bound
    o = B.make_options(ignore_stop=1, max_iter=K, **gopts)
    st = B.Stats()

    def one_call():
        hx.zero_()
        barrier()
        t0 = time.perf_counter()
        B._check(L.cgb_conjugrad_rank(ctx._h, hA.data_ptr(), hb.data_ptr(), hx.data_ptr(), m, C.byref(o), C.byref(st)))
        torch.cuda.synchronize()
        return max_over_ranks(time.perf_counter() - t0)

    one_call()                                             # warm-up call (pins, graphs)
    dt = one_call()
    h2d = 8.0 * (m * n + m + n)
    line["e2e"] = {"value": K / dt, "unit": UNIT, "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": 8.0 * n / K,
                   "seconds_per_call": dt, "h2d_seconds": max_over_ranks(st.h2d_seconds),
                   "h2d_gbs_per_gpu": h2d / max(st.h2d_seconds, 1e-12) / 1e9,
                   "solve_seconds": max_over_ranks(st.solve_seconds), "iterations": int(st.iterations),
                   "api": "cgb_conjugrad_rank (C ABI, pinned host buffers, one upload of A per call)",
                   "host_buffers": (f"pinned, allocated on the {cpus_bound} CPUs local to the GPU (sysfs local_cpulist)"
                                    if cpus_bound else "pinned, default placement (no narrower GPU-local CPU set on this box)")}
    if "symmetric_variant" in line:
        # the same call with the matrix asserted symmetric (symmetric = 2): the tiles the product
        # never reads stay on the host, so about half the bytes cross PCIe
        o3 = B.make_options(ignore_stop=1, max_iter=K, symmetric=2, **gopts)

        def sym_call():
            hx.zero_()
            barrier()
            t0 = time.perf_counter()
            B._check(L.cgb_conjugrad_rank(ctx._h, hA.data_ptr(), hb.data_ptr(), hx.data_ptr(), m, C.byref(o3), C.byref(st)))
            torch.cuda.synchronize()
            return max_over_ranks(time.perf_counter() - t0)

        sym_call()
        dt3 = sym_call()
        line["symmetric_variant"]["e2e"] = {"value": K / dt3, "unit": UNIT, "h2d_bytes_per_step": float(st.h2d_bytes) / K,
                                            "d2h_bytes_per_step": 8.0 * n / K, "seconds_per_call": dt3,
                                            "h2d_seconds": max_over_ranks(st.h2d_seconds),
                                            "api": "cgb_conjugrad_rank with cgb_options.symmetric = 2"}
    # the converging solve a user of the reference CLI runs: upload + ~5 iterations + download
    o2 = B.make_options(**gopts)
    hx.zero_()
    barrier()
    t0 = time.perf_counter()
    B._check(L.cgb_conjugrad_rank(ctx._h, hA.data_ptr(), hb.data_ptr(), hx.data_ptr(), m, C.byref(o2), C.byref(st)))
    dt2 = max_over_ranks(time.perf_counter() - t0)
    line["time_to_solution"]["e2e_seconds_host_buffers"] = dt2
    line["time_to_solution"]["e2e_iterations"] = int(st.iterations)
    del hA, hb, hx


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    ctx, line, env = b200_arm(args)
    rank, world = env[0], env[1]
    if not args.no_e2e:
        e2e_leg(ctx, line, env)
    ctx.close()
    if rank == 0 and world == 1 and not args.no_cpu:
        # reported baseline, not the target: the reference's SPMD algorithm on the host cores
        threads, cores = cpu_threads_for(args.n, args.cpu_threads)
        iters = min(10, reference_iterations(args.n, threads, 10))      # ~3 s of wall at n = 65536 on 16 threads
        val, st = run_cpu_port(args.n, iters, threads)
        line["cpu_baseline"] = {
            "value": val, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"full n={args.n} matrix, {iters} iterations after A*x0, {threads} of {cores} host cores "
                      f"(oracle/cg_oracle.c orc_team_solve_generated: parallel_cg.c's per-rank loops on threads)",
            "gemv_gbs_per_rank": gemv_bytes(args.n // threads, args.n) * val / 1e9,
            "gen_seconds": st["gen_seconds"]}
    if env[9] is not None:
        env[9].barrier()
        env[9].destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
