"""CPU tests of bench.py's host-side pieces: the JSON line of the reference arm carries every key
the driver's contract names, the bounded-sample rule of the CPU arm, the algorithmic-bytes
figure of SURVEY.md 8(d), the workload names, and the nvidia-smi clock-sample parser."""
import json
import os
import subprocess
import sys

import pytest

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_algorithmic_bytes_match_the_survey_table():
    # GEMV = 8 (m n + n + m); G = 1 values 134.28 MB / 2.1477 GB / 34.361 GB / 137.441 GB
    assert bench.gemv_bytes(4096, 4096) == pytest.approx(134.28e6, rel=1e-4)
    assert bench.gemv_bytes(16384, 16384) == pytest.approx(2.1477e9, rel=1e-4)
    assert bench.gemv_bytes(65536, 65536) == pytest.approx(34.361e9, rel=1e-4)
    assert bench.gemv_bytes(131072, 131072) == pytest.approx(137.441e9, rel=1e-4)
    assert bench.gemv_bytes(8192, 65536) == 8.0 * (8192 * 65536 + 65536 + 8192)       # one of 8 row blocks


def test_workload_names_follow_baseline_configs():
    with open(os.path.join(ROOT, "BASELINE.json")) as f:
        configs = json.load(f)["configs"]
    for n, idx in ((4096, 1), (16384, 2), (65536, 3), (131072, 4)):
        assert f"n={n} " in configs[idx]
        assert f"configs[{idx}]" in bench.workload_name(n) and f"n={n} " in bench.workload_name(n)
    assert "off-list" in bench.workload_name(1000)


def test_cpu_arm_is_a_bounded_sample():
    assert bench.reference_iterations(65536, 16, 20) == 20            # the default run is not cut
    big = bench.reference_iterations(65536, 16, 100000)
    assert 3 <= big < 100000
    assert 8.0 * 65536 ** 2 * big / (16 * 5.0e9) <= bench.REFERENCE_BUDGET_S
    assert bench.reference_iterations(131072, 1, 1000) == 5          # 27.5 s an iteration on one core
    assert bench.reference_iterations(1 << 20, 1, 1000) == 3         # never fewer than 3 iterations
    assert bench.reference_iterations(2048, 8, 500) == 500


def test_thread_count_is_a_power_of_two_dividing_n():
    cores = os.cpu_count() or 1
    for n in (65536, 4096, 6, 7, 24):
        t, c = bench.cpu_threads_for(n)
        assert c == cores and 1 <= t <= cores and n % t == 0 and t & (t - 1) == 0
    assert bench.cpu_threads_for(65536, want=4)[0] == 4
    assert bench.cpu_threads_for(7, want=8)[0] == 1


def test_reference_arm_line_has_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--matrix-n", "1024",
                        "--steps", "4", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                            # ONE JSON line
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
              "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == bench.UNIT
    assert d["steps"] == 4 and d["dtype"] == "f64" and d["vs_baseline"] is None and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] == pytest.approx(1e3 / d["value"])
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] and cb["cores"] >= 1 and "1024" in cb["sample"]
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["config"] == bench.workload_config(1024, 1)              # the same dict the B200 arm emits
    assert d["warmup"] == 1 and d["implementation"]["steps_run"] == 4


def test_reference_timer_line_is_parsed():
    assert bench.reference_timer_seconds("Computing\naverage clock execution time in seconds: 0.562000\n") == pytest.approx(0.562)
    assert bench.reference_timer_seconds("cg method execution time in seconds: 0.234\nclock execution time in seconds: 16.0\n") == pytest.approx(0.234)
    assert bench.reference_timer_seconds("nothing") is None


def test_cpu_rows_child_uses_the_real_reference_timers():
    import _oracle as O
    if not O.ref_available("serial", "f64", 8192):
        pytest.skip("oracle/_ref has no n = 8192 build here")
    rows = bench.cpu_reference_rows()
    names = [r["name"] for r in rows]
    assert names[:2] == ["serial_f32_1core", "serial_f64_1core"] and "spmd_np2" in names
    for r in rows:
        assert r["kind"] == "reference" and r["n"] == 8192 and r["value"] > 0 and "own" in r["timer"]


def test_clock_sample_parser(tmp_path):
    s = bench.ClockSampler.__new__(bench.ClockSampler)

    class Done:
        def terminate(self): pass
        def wait(self, timeout=None): return 0
        def kill(self): pass
    s.proc = Done()
    s.path = str(tmp_path / "clocks.csv")
    with open(s.path, "w") as f:
        f.write("2026/11/04 10:00:00.000, 1200, 1965, 300.0, Not Active, Not Active, Not Active, Not Active\n"   # before
                "2026/11/04 10:00:01.000, 1965, 1965, 900.0, Not Active, Not Active, Not Active, Active\n"
                "2026/11/04 10:00:01.020, 1950, 1965, 910.0, Not Active, Not Active, Not Active, Not Active\n"
                "2026/11/04 10:00:01.040, 1965, 1965, 905.0, Not Active, Not Active, Not Active, Not Active\n"
                "garbage line\n"
                "2026/11/04 10:00:05.000, 800, 1965, 200.0, Active, Not Active, Not Active, Not Active\n")       # after
    s.f = open(s.path)
    lo = bench.ClockSampler._ts("2026/11/04 10:00:01.000")
    out = s.stop(lo, lo + 0.05)
    assert out == {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "reasons": ["sw_power_cap"], "samples": 3}


# The two stub tests below patch torch and the binding; they run in a fresh interpreter so that nothing they
# replace can leak into other test modules (and so that they do not depend on which NCCL copy earlier modules
# of a full run have mapped -- see _binding._share_nccl_with_torch and tests/test_abi.py).
IN_CHILD = os.environ.get("CGB_STUB_CHILD") == "1"


def test_fresh_interpreter_runs_both_legs_on_stubs():
    if IN_CHILD:
        pytest.skip("already the child")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-p", "no:cacheprovider", os.path.abspath(__file__), "-k", "stub_"],
                       capture_output=True, text=True, timeout=600, env=dict(os.environ, CGB_STUB_CHILD="1"), cwd=ROOT)
    assert r.returncode == 0 and "2 passed" in r.stdout, r.stdout[-3000:] + r.stderr[-1000:]


@pytest.mark.skipif(not IN_CHILD, reason="runs inside test_fresh_interpreter_runs_both_legs_on_stubs")
def test_stub_e2e_leg_runs_end_to_end_against_a_stub_library(monkeypatch):
    """Executes every line of bench.e2e_leg on this GPU-less machine: the library is replaced by a stub
    that records the calls (no arithmetic happens, nothing here is a parity claim) -- a typo in the leg
    that produces the headline `e2e` object would otherwise only show on the GPU box."""
    import ctypes as C
    import types
    import torch
    from conjugate_gradient_b200 import _binding as B

    calls = []

    class StubLib:
        def cgb_debug_read(self, h, what, off, count, dst):
            calls.append(("read", what, count)); return 0

        def cgb_conjugrad_rank(self, h, A, b, x, m, o, st):
            opt = C.cast(o, C.POINTER(B.Options)).contents
            s = C.cast(st, C.POINTER(B.Stats)).contents
            s.iterations = opt.max_iter if opt.ignore_stop else 4
            s.h2d_seconds, s.solve_seconds = 0.5, 0.1
            s.h2d_bytes = 8 * (m * 64 + m + 64) // (2 if opt.symmetric == 2 else 1)
            s.h2d_path = 1 if A in pinned_ptrs else 2
            calls.append(("solve", int(opt.symmetric), int(opt.ignore_stop), int(s.h2d_path)))
            return 0

        def cgb_default_options(self, o):
            return None

    real_empty = torch.empty
    pinned_ptrs = set()

    def fake_empty(*a, pin_memory=False, **k):
        t = real_empty(*a, **k)
        if pin_memory:
            pinned_ptrs.add(t.data_ptr())
            keep.append(t)
        return t
    keep = []
    monkeypatch.setattr(torch, "empty", fake_empty)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.setattr(B, "lib", lambda: StubLib())
    n = m = 64
    K = 20
    line = {"symmetric_variant": {}, "time_to_solution": {}}
    env = (0, 1, 0, n, m, K, lambda: None, float, {}, None)
    bench.e2e_leg(types.SimpleNamespace(_h=None), line, env)
    e = line["e2e"]
    assert e["unit"] == bench.UNIT and e["value"] > 0 and e["iterations"] == K
    assert e["h2d_bytes_per_step"] == 8.0 * (m * n + m + n) / K and e["d2h_bytes_per_step"] == 8.0 * n / K
    assert "malloc" in e["api"] and e["h2d_path"].startswith("staged") and e["h2d_frac_of_pinned"] == pytest.approx(1.0)
    assert line["e2e_pinned"]["h2d_path"].startswith("direct") and line["e2e_pinned"]["value"] > 0
    assert line["symmetric_variant"]["e2e"]["value"] > 0
    assert line["time_to_solution"]["e2e_iterations"] == 4
    # A and b read back once; two calls on malloc'ed buffers (the headline), two on pinned ones, two
    # asserted-symmetric calls and one converging solve on malloc'ed buffers again
    assert [c for c in calls if c[0] == "read"] == [("read", 0, m * n), ("read", 1, m)]
    assert [c[1:] for c in calls if c[0] == "solve"] == [(0, 1, 2), (0, 1, 2), (0, 1, 1), (0, 1, 1), (2, 1, 2), (2, 1, 2), (0, 0, 2)]


@pytest.mark.skipif(not IN_CHILD, reason="runs inside test_fresh_interpreter_runs_both_legs_on_stubs")
def test_stub_b200_arm_line_against_a_stub_context(monkeypatch):
    """The same for the device-resident leg: bench.b200_arm assembled on a stub Context (fixed timings, no
    arithmetic) must produce every key of the driver's contract with the stated meaning."""
    import types
    import torch
    import conjugate_gradient_b200 as cg

    class StubCtx:
        def __init__(self, **kw):
            self.solves = []

        def device_info(self):
            return {"name": "stub"}

        def generate(self, n, seed):
            self.n = n

        def set_x0(self, x):
            pass

        def read_x(self):
            import numpy as np
            return np.zeros(self.n)

        def close(self):
            pass

        def solve(self, **kw):
            self.solves.append(kw)
            K = kw.get("max_iter", 4)
            return {"loop_ms": 5.0 * K, "gemv_ms_avg": 4.5, "gemv_ms_min": 4.4, "gemv_timed": K, "kernel_launches": 3 * K + 4,
                    "comm": 0, "pdl": 1, "graph_chunk": 2, "solve_seconds": 0.02, "iterations": K, "final_residual": 1e-8,
                    "gemv_bytes": 1.0e9, "gemv_variant": 1, "gemv_rows": 8, "gemv_unroll": 2, "gemv_threads": 512, "gemv_vec": 4,
                    "gemv_stages": 0, "gemv_tile_cols": 0, "gemv_grid": 8192}

    monkeypatch.setattr(cg, "Context", StubCtx)
    monkeypatch.setattr(cg, "device_count", lambda: 1)
    monkeypatch.setattr(torch.cuda, "set_device", lambda i: None)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.setattr(bench, "ClockSampler", lambda i: types.SimpleNamespace(
        stop=lambda lo, hi: {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "reasons": [], "samples": 5}))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    args = types.SimpleNamespace(gpus=1, steps=20, warmup=3, n=65536, gemv=0, comm="auto", rows=0, unroll=0, threads=0, l2_keep=0)
    ctx, line, env, conv = bench.b200_arm(args)
    assert conv[0]["iterations"] == 4 and conv[1].shape == (65536,)
    assert line["config"] == bench.workload_config(65536, 1) and "gemv_kernel" in line["implementation"]
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "roofline", "clocks", "gpu_launches"):
        assert k in line, k
    assert line["value"] == pytest.approx(20 / 0.1) and line["ms_per_step"] == pytest.approx(5.0)
    assert line["steps"] == 20 and line["warmup"] == 3 and line["n_gpus"] == 1 and line["dtype"] == "f64"
    assert "configs[3]" in line["config"]["workload"] and "model" not in line["config"]
    r = line["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["bytes_per_launch"] == bench.gemv_bytes(65536, 65536)
    assert r["achieved"] == pytest.approx(r["bytes_per_launch"] / 4.5e-3 / 1e9) and r["frac"] == pytest.approx(r["achieved"] / r["peak"])
    assert line["gpu_launches"] == 3 * 20                     # the 4 set-up launches are outside the K iterations
    # warm-up (W stream-mode iterations, one pass of the graph run), the timed graph run, the instrumented pass
    assert ctx.solves[0]["max_iter"] == 3 and ctx.solves[0]["time_gemv"] == 1
    assert ctx.solves[1]["max_iter"] == 20 and "time_gemv" not in ctx.solves[1]
    assert ctx.solves[2]["max_iter"] == 20 and "time_gemv" not in ctx.solves[2]
    assert ctx.solves[3]["max_iter"] == 20 and ctx.solves[3]["time_gemv"] == 1
    assert line["instrumented_pass"]["value"] == pytest.approx(line["value"]) and "graph" in line["timed_region"]["mode"]
    assert "symmetric_variant" in line and "fused_reductions_variant" in line
    json.dumps(line)                                         # serialisable as ONE line


def test_parity_leg_accepts_the_oracle_and_rejects_a_wrong_answer():
    """bench.parity_leg on a stand-in context that answers with the oracle's own x (serial order: differs from the
    threaded SPMD order in the last bits only) must report ok; the same x with one entry off by 1e-6 must not."""
    import types
    import numpy as np
    import _oracle as O
    n, K = 512, 20
    A, b = O.gen_A(n, bench.SEED), O.gen_b(n, bench.SEED)
    x_fix, s_fix = O.conjugrad(A, b, np.zeros(n), ignore_stop=True, max_iter=10)
    x_conv, s_conv = O.conjugrad(A, b, np.zeros(n))

    class Ctx:
        def __init__(self, bump):
            self.bump = bump

        def set_x0(self, x):
            pass

        def solve(self, **kw):
            return {"first_hit_k": s_fix["first_hit_k"], "iterations": 10}

        def read_x(self):
            x = x_fix.copy()
            x[7] += self.bump
            return x

    args = types.SimpleNamespace(cpu_threads=2, no_cpu=False)
    env = (0, 1, 0, n, n, K, lambda: None, float, {}, None)
    conv = ({"iterations": s_conv["iterations"], "final_residual": float(np.sqrt(s_conv["final_beta"]))}, x_conv)
    line = {}
    assert bench.parity_leg(Ctx(0.0), line, env, conv, args) is True
    p = line["parity"]
    assert p["ok"] and p["fixed_run"]["iterations"] == 10 and p["fixed_run"]["rel_err_x"] <= 1e-12
    assert p["converging_solve"]["iterations"]["gpu"] == p["converging_solve"]["iterations"]["oracle"]
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] == 2
    line = {}
    assert bench.parity_leg(Ctx(1e-6), line, env, conv, args) is False and line["parity"]["ok"] is False
