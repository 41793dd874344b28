"""CPU-only: the plain-C part of the product path (cg_io.c: multi-threaded exact text parser, writer, the
token locator shared with the GPU ingest) rebuilt from source under AddressSanitizer + UBSan and under
ThreadSanitizer and driven by tests/csrc/host_io_sanitize.c (every value checked against strtod, ranges,
short / empty / malformed files).  The sanitizers' own verdict is part of the exit status."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = [os.path.join(ROOT, "tests", "csrc", "host_io_sanitize.c"),
       os.path.join(ROOT, "conjugate_gradient_b200", "csrc", "cg_io.c")]
CC = shutil.which("gcc") or "gcc"


def build(tmp_path, name, flags):
    exe = str(tmp_path / name)
    r = subprocess.run([CC, "-O1", "-g", "-std=gnu11", "-fno-omit-frame-pointer", *flags, *SRC, "-o", exe, "-lpthread", "-lm"],
                       capture_output=True, text=True)
    if r.returncode != 0 and ("cannot find" in r.stderr or "unrecognized" in r.stderr):
        pytest.skip(f"this gcc has no {flags} runtime: {r.stderr.strip().splitlines()[-1]}")
    assert r.returncode == 0, r.stderr
    return exe


def run(exe, tmp_path, env_extra):
    env = dict(os.environ, **env_extra)
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=600, env=env)
    if r.returncode != 0 and "unexpected memory mapping" in r.stderr:
        pytest.skip("ThreadSanitizer cannot map its shadow memory under this kernel's ASLR settings")
    if r.returncode != 0 and "LeakSanitizer has encountered a fatal error" in r.stderr:
        # leak checking needs ptrace, which some sandboxes refuse: keep the address / UB checks
        env["ASAN_OPTIONS"] = "detect_leaks=0"
        r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=600, env=env)
    return r


def test_host_text_path_under_asan_ubsan(tmp_path):
    exe = build(tmp_path, "io_asan", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"])
    r = run(exe, tmp_path, {"ASAN_OPTIONS": "detect_leaks=1:abort_on_error=0", "UBSAN_OPTIONS": "print_stacktrace=1"})
    assert r.returncode == 0 and "host_io_sanitize: ok" in r.stdout, r.stderr[-3000:]
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr and "LeakSanitizer" not in r.stderr


def test_host_text_path_when_no_thread_can_be_created(tmp_path):
    # pthread_create renamed to a stub that always answers EAGAIN: every span / formatter job runs inline
    exe = build(tmp_path, "io_nothreads", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
                                           "-Dpthread_create=cgb_test_failing_create", "-DCGB_TEST_THREADS_FAIL"])
    r = run(exe, tmp_path, {"ASAN_OPTIONS": "detect_leaks=1"})
    assert r.returncode == 0 and "host_io_sanitize: ok" in r.stdout, r.stderr[-3000:]
    assert "AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr


def test_host_text_path_under_tsan(tmp_path):
    exe = build(tmp_path, "io_tsan", ["-fsanitize=thread"])
    r = run(exe, tmp_path, {"TSAN_OPTIONS": "halt_on_error=0:exitcode=66"})
    assert r.returncode == 0 and "host_io_sanitize: ok" in r.stdout, r.stderr[-3000:]
    assert "WARNING: ThreadSanitizer" not in r.stderr


def test_cli_argument_handling_under_asan_ubsan(tmp_path, golden_dir):
    """cg_main.c (the reference's `prog A.txt b.txt x0.txt` surface) rebuilt with the sanitizers and linked to
    the normal libcgb200.so: every path that ends before a GPU is needed -- usage errors, the dimensions.txt
    lookup, non-square / mismatching sizes, missing files, the divisibility rule."""
    import conjugate_gradient_b200 as cg
    libdir = os.path.dirname(cg.LIB_PATH)
    exe = str(tmp_path / "cg_b200_asan")
    r = subprocess.run([CC, "-O1", "-g", "-std=gnu11", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
                        os.path.join(ROOT, "conjugate_gradient_b200", "csrc", "cg_main.c"), "-o", exe,
                        f"-L{libdir}", "-lcgb200", f"-Wl,-rpath,{libdir}", "-lpthread", "-lm"], capture_output=True, text=True)
    if r.returncode != 0 and "cannot find -lasan" in r.stderr:
        pytest.skip("no libasan")
    assert r.returncode == 0, r.stderr
    g = lambda f: os.path.join(golden_dir, f)
    files = [g("g2_matrixA.txt"), g("g2_vectorb.txt"), g("g2_x0.txt")]
    bad_dims = tmp_path / "dims_bad.txt"; bad_dims.write_text("4\n3\n4\n1\n")
    b_dims = tmp_path / "dims_b.txt"; b_dims.write_text("4\n4\n5\n1\n")
    cases = [
        ([], 2, "requires four (4) files"),
        (files[:2], 2, "requires four (4) files"),
        (files + ["--gpus", "0"], 2, ">= 1"),
        (files + ["--eps", "0"], 2, "positive number"),
        (files + ["--seed", "1e3"], 2, "unsigned integer"),
        (files + ["--dims"], 2, "needs a value"),
        (["--gen", "64"] + files, 2, "takes no input files"),
        (files + ["--dims", str(bad_dims)], 1, "4 and 3 must be same size"),                 # serialConjugate.c:53-57
        (files + ["--dims", str(b_dims)], 1, "do not match A"),
        (files + ["--dims", g("g2_dimensions.txt"), "--gpus", "3"], 1, "4 is not divisible by 3"),   # parallel_cg.c:86-90
        (["/nonexistent/A.txt", "/nonexistent/b.txt", "/nonexistent/x.txt"], 1, "Could not open file"),
        (["--gen", "10", "--gpus", "4"], 1, "10 is not divisible by 4"),
        (["x" * 5000, "y", "z"], 1, "Could not open file"),                                   # a path longer than the buffers
    ]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0")
    for args, code, msg in cases:
        r = subprocess.run([exe] + args, capture_output=True, text=True, env=env, timeout=120)
        out = r.stdout + r.stderr
        assert "AddressSanitizer" not in out and "runtime error" not in out, (args[:2], out[-2000:])
        assert r.returncode == code and msg in out, (args, r.returncode, out[-500:])


def test_cli_random_argument_vectors_under_asan(tmp_path, golden_dir):
    """300 seeded random argument vectors from the CLI's own vocabulary (plus junk) against the sanitizer build:
    whatever the combination, the process ends with status 0, 1 or 2 and no sanitizer report."""
    import random
    import conjugate_gradient_b200 as cg
    libdir = os.path.dirname(cg.LIB_PATH)
    exe = str(tmp_path / "cg_b200_asan")
    r = subprocess.run([CC, "-O1", "-g", "-std=gnu11", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
                        os.path.join(ROOT, "conjugate_gradient_b200", "csrc", "cg_main.c"), "-o", exe,
                        f"-L{libdir}", "-lcgb200", f"-Wl,-rpath,{libdir}", "-lpthread", "-lm"], capture_output=True, text=True)
    if r.returncode != 0 and "cannot find -lasan" in r.stderr:
        pytest.skip("no libasan")
    assert r.returncode == 0, r.stderr
    g = lambda f: os.path.join(golden_dir, f)
    flags = ["--print-x", "--stats", "--round4", "--guard", "--symmetric", "--dims", "--write-x", "--run-record", "--emit-files",
             "--gpus", "--eps", "--max-iter", "--gen", "--seed", "--bench-iters", "--gemv", "--cg", "--kappa-log2", "--comm", "--bogus"]
    values = ["0", "1", "2", "3", "16", "-1", "1e-6", "x", "", "classic", "fused", "peer", "nccl", "auto", "99999999999999999999",
              g("g2_dimensions.txt"), g("g2_matrixA.txt"), g("g2_vectorb.txt"), g("g2_x0.txt"), str(tmp_path / "out.txt"), "/nonexistent/f"]
    rnd = random.Random(20170209)
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0")
    for _ in range(300):
        argv = [rnd.choice(flags + values) for _ in range(rnd.randint(0, 9))]
        r = subprocess.run([exe] + argv, capture_output=True, text=True, env=env, timeout=120, cwd=str(tmp_path))
        out = r.stdout + r.stderr
        assert "AddressSanitizer" not in out and "runtime error" not in out, (argv, out[-1500:])
        assert r.returncode in (0, 1, 2), (argv, r.returncode, out[-500:])
