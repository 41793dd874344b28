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


def test_numa_local_binds_and_restores(tmp_path):
    assert bench.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert bench.parse_cpulist("") == set() and bench.parse_cpulist("a-b") == set()
    before = os.sched_getaffinity(0)
    dev = tmp_path / "0000:1b:00.0"
    dev.mkdir()
    first = min(before)
    (dev / "local_cpulist").write_text(f"{first}\n")
    with bench.NumaLocal(0, 0x1b, 0, sysfs=str(tmp_path)) as nl:
        if len(before) > 1:
            assert os.sched_getaffinity(0) == {first} and nl.bound == 1
        else:
            assert nl.bound == 0                                   # nothing narrower to bind to
    assert os.sched_getaffinity(0) == before
    # no sysfs entry, or a CPU list outside the container's cpuset: the process is left alone
    with bench.NumaLocal(0, 0x99, 0, sysfs=str(tmp_path)) as nl:
        assert nl.bound == 0 and os.sched_getaffinity(0) == before
    (dev / "local_cpulist").write_text("100000-100003\n")
    with bench.NumaLocal(0, 0x1b, 0, sysfs=str(tmp_path)) as nl:
        assert nl.bound == 0 and os.sched_getaffinity(0) == before
    assert os.sched_getaffinity(0) == before


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
            s.h2d_seconds, s.solve_seconds, s.h2d_bytes = 0.5, 0.1, 8 * (m * 64) // (2 if opt.symmetric == 2 else 1)
            calls.append(("solve", int(opt.symmetric), int(opt.ignore_stop)))
            return 0

        def cgb_default_options(self, o):
            return None

    real_empty, real_zeros = torch.empty, torch.zeros
    monkeypatch.setattr(torch, "empty", lambda *a, pin_memory=False, **k: real_empty(*a, **k))
    monkeypatch.setattr(torch, "zeros", lambda *a, pin_memory=False, **k: real_zeros(*a, **k))
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.setattr(torch.cuda, "get_device_properties",
                        lambda i: types.SimpleNamespace(pci_domain_id=0, pci_bus_id=0x1b, pci_device_id=0))
    monkeypatch.setattr(B, "lib", lambda: StubLib())
    before = os.sched_getaffinity(0)
    n = m = 64
    K = 20
    line = {"symmetric_variant": {}, "time_to_solution": {}}
    env = (0, 1, 0, n, m, K, lambda: None, float, {}, None)
    bench.e2e_leg(types.SimpleNamespace(_h=None), line, env)
    e = line["e2e"]
    assert e["unit"] == bench.UNIT and e["value"] > 0 and e["iterations"] == K
    assert e["h2d_bytes_per_step"] == 8.0 * (m * n + m + n) / K and e["d2h_bytes_per_step"] == 8.0 * n / K
    assert "host_buffers" in e and "api" in e
    assert line["symmetric_variant"]["e2e"]["value"] > 0
    assert line["time_to_solution"]["e2e_iterations"] == 4
    # A and b read back once; two plain calls, two asserted-symmetric calls, one converging solve
    assert [c for c in calls if c[0] == "read"] == [("read", 0, m * n), ("read", 1, m)]
    assert [c[1:] for c in calls if c[0] == "solve"] == [(0, 1), (0, 1), (2, 1), (2, 1), (0, 0)]
    assert os.sched_getaffinity(0) == before


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
    ctx, line, env = bench.b200_arm(args)
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
    assert ctx.solves[0]["max_iter"] == 3 and ctx.solves[1]["max_iter"] == 20 and ctx.solves[1]["time_gemv"] == 1
    assert "symmetric_variant" in line and "fused_reductions_variant" in line
    json.dumps(line)                                         # serialisable as ONE line
