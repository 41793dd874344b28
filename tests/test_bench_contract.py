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
