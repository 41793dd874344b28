"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports
every symbol include/cgb200.h declares, the host-side text I/O behaves like the
reference's initialize()/printer(), and every computing entry point fails
loudly (no fallback) when there is no GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import conjugate_gradient_b200 as cg
from conjugate_gradient_b200 import _binding as B


def test_library_exports_every_declared_symbol():
    L = cg.lib()
    assert len(cg.HEADER_SYMBOLS) >= 30
    missing = [s for s in cg.HEADER_SYMBOLS if not hasattr(L, s)]
    assert not missing, missing
    assert L.cgb_version() == 100


def test_struct_layouts_match_header():
    # sizeof as the C compiler sees them (a tiny probe compiled against the header)
    src = '#include <stdio.h>\n#include "cgb200.h"\nint main(){printf("%zu %zu\\n", sizeof(cgb_options), sizeof(cgb_stats));return 0;}\n'
    exe = "/tmp/cgb_sizeof_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(B.ROOT, "include"), "-o", exe],
                   input=src.encode(), check=True)
    so, ss = map(int, subprocess.run([exe], capture_output=True, check=True).stdout.split())
    assert so == C.sizeof(cg.Options) and ss == C.sizeof(cg.Stats)


def test_no_oracle_in_product():
    # the product must not link, import or execute anything under oracle/
    out = subprocess.run(["ldd", cg.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out
    for root, _, files in os.walk(os.path.join(B.ROOT, "conjugate_gradient_b200")):
        for f in files:
            if f.endswith((".py", ".c", ".cu", ".h", ".cuh", "Makefile")):
                txt = open(os.path.join(root, f), errors="ignore").read()
                assert "liboracle" not in txt and "_oracle" not in txt and "oracle/" not in txt, f


def test_text_inputs(tmp_path, golden_dir):
    assert cg.read_dimensions(os.path.join(golden_dir, "g1_dimensions.txt")) == (2, 2, 2, 1)
    v = cg.read_values(os.path.join(golden_dir, "g2_matrixA.txt"), 16)
    assert v[0] == 7.0 and v[7] == -4.0 and v[15] == 6.0
    # separators "%f%*c" tolerates, CRLF, a BOM (the reference's initialguess1.txt has one)
    p = tmp_path / "mixed.txt"
    p.write_bytes(b"1.5,2.5;3e0\r\n-4.25 \t5\n\xef\xbb\xbf6.0\n")
    np.testing.assert_array_equal(cg.read_values(str(p), 6), [1.5, 2.5, 3.0, -4.25, 5.0, 6.0])
    with pytest.raises(cg.CgbError):           # short file: reported, not silently garbage
        cg.read_values(str(p), 7)
    with pytest.raises(cg.CgbError):           # missing file ("Could not open file")
        cg.read_values(str(tmp_path / "nope.txt"), 1)
    q = tmp_path / "bad.txt"
    q.write_text("1.0\nabc\n")
    with pytest.raises(cg.CgbError):
        cg.read_values(str(q), 2)
    # generateSPDmatrix.m:28-43 format round trip
    w = tmp_path / "w.txt"
    cg.write_values(str(w), [0.12344, 2048.5], 4)
    assert w.read_text() == "0.1234\n2048.5000\n"
    cg.write_values(str(w), [1 / 3], -1)
    assert cg.read_values(str(w), 1)[0] == 1 / 3


def test_cli_messages_without_gpu(golden_dir):
    exe = B.CLI_PATH
    r = subprocess.run([exe, "a", "b"], capture_output=True, text=True)
    assert r.returncode != 0 and "requires four (4) files" in r.stdout      # serialConjugate.c:50
    r = subprocess.run([exe, os.path.join(golden_dir, "g2_matrixA.txt"), os.path.join(golden_dir, "g2_vectorb.txt"),
                        os.path.join(golden_dir, "g2_x0.txt"), "--dims", os.path.join(golden_dir, "g2_dimensions.txt"),
                        "--gpus", "3"], capture_output=True, text=True)
    assert r.returncode != 0 and "4 is not divisible by 3" in r.stdout       # parallel_cg.c:88
    r = subprocess.run([exe, "/nonexistent/A.txt", "/nonexistent/b.txt", "/nonexistent/x.txt"], capture_output=True, text=True)
    assert r.returncode != 0 and "Could not open file" in r.stdout           # serialConjugate.c:103


@pytest.mark.parametrize("args,msg", [
    (["--gpus", "0"], "--gpus takes an integer >= 1"),          # was a division by zero in the n % gpus test
    (["--gpus", "two"], "--gpus takes an integer >= 1"),
    (["--eps", "-1"], "--eps takes a positive number"),
    (["--max-iter", "x"], "--max-iter takes an integer >= 1"),
    (["--gen", "0"], "--gen takes an integer >= 1"),
    (["--kappa-log2", "99"], "at most 63"),
    (["--cg", "pipelined"], "--cg takes classic or fused"),
    (["--comm", "mpi"], "--comm takes nccl, peer or auto"),
    (["--no-such-flag", "1"], "unknown option"),
    (["--dims"], "needs a value"),
    (["--seed", "12x"], "--seed takes an unsigned integer"),
    (["--seed", "-5"], "--seed takes an unsigned integer"),
    (["--gemv", "7"], "--gemv takes 0"),
    (["--gemv", "bulk"], "--gemv takes an integer"),
    (["--gen", "64"], "takes no input files"),                  # was a silent exit 2
    (["--max-iter", "99999999999999999999"], "--max-iter takes an integer"),   # strtoll saturates: not a number we were given
])
def test_cli_usage_errors(golden_dir, args, msg):
    files = [os.path.join(golden_dir, f) for f in ("g2_matrixA.txt", "g2_vectorb.txt", "g2_x0.txt")]
    r = subprocess.run([B.CLI_PATH] + files + args, capture_output=True, text=True)
    assert r.returncode == 2 and msg in r.stderr, (r.returncode, r.stderr)


def test_cli_refuses_orders_whose_square_overflows():
    r = subprocess.run([B.CLI_PATH, "--gen", "4000000000"], capture_output=True, text=True)
    assert r.returncode == 1 and "out of range" in r.stderr and "Computing cg" not in r.stdout   # printed n*n = -2446744073709551616
    r = subprocess.run([B.CLI_PATH, "--gen", "99999999999999999999"], capture_output=True, text=True)
    assert r.returncode == 2 and "--gen takes an integer" in r.stderr


def test_compute_fails_loudly_without_gpu():
    if cg.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(cg.CgbError):
        cg.Context(0)
    with pytest.raises(cg.CgbError):
        cg.conjugrad_host(np.eye(2), np.ones(2), np.zeros(2))


def test_invalid_arguments_are_status_codes_not_crashes(golden_dir):
    """Every entry point answers a bad argument with a CGB_ERR_* code and a message (the reference
    would exit(0) or crash: serialConjugate.c:48-57,78-82); nothing here needs a GPU."""
    import ctypes as C
    L = cg.lib()
    A, b, x = np.eye(2), np.ones(2), np.zeros(2)
    o, st = B.make_options(), B.Stats()
    ARG, NODEV, NOTDIV, STATE, IO = -1, -2, -4, -6, -7
    host = L.cgb_conjugrad_host_opt
    for n in (0, -1):
        for gpus in (-1, 0, 1, 3, 100000):
            assert host(A.ctypes.data, b.ctypes.data, x.ctypes.data, n, gpus, C.byref(o), C.byref(st)) == ARG
    assert host(None, b.ctypes.data, x.ctypes.data, 2, 1, C.byref(o), C.byref(st)) == ARG
    # orders whose n * n * 8 bytes would wrap size_t are refused before anything is sized from them
    for n in (2 ** 24 + 1, 2 ** 31, 2 ** 40, 2 ** 62):
        assert host(A.ctypes.data, b.ctypes.data, x.ctypes.data, n, 1, C.byref(o), C.byref(st)) == ARG
        assert L.cgb_solve_generated(n, 1, 0, 1, C.byref(o), None, C.byref(st)) == ARG
        assert L.cgb_conjugrad_files(b"a", b"b", b"c", n, 1, C.byref(o), None, C.byref(st)) == ARG
    assert host(A.ctypes.data, b.ctypes.data, x.ctypes.data, 2, 3, C.byref(o), C.byref(st)) == NOTDIV   # parallel_cg.c:86-90
    assert b"2 is not divisible by 3" in L.cgb_last_error()
    assert host(A.ctypes.data, b.ctypes.data, x.ctypes.data, 2, 17, C.byref(o), C.byref(st)) == ARG
    assert L.cgb_solve_generated(0, 1, 0, 1, C.byref(o), None, C.byref(st)) == ARG
    assert L.cgb_solve_generated(4, 1, 0, 3, C.byref(o), None, C.byref(st)) == NOTDIV
    files = [os.fsencode(os.path.join(golden_dir, f)) for f in ("g1_matrixA.txt", "g1_vectorb.txt", "g1_x0.txt")]
    assert L.cgb_conjugrad_files(None, None, None, 2, 1, C.byref(o), x.ctypes.data, C.byref(st)) == ARG
    assert L.cgb_conjugrad_files(*files, 0, 1, C.byref(o), x.ctypes.data, C.byref(st)) == ARG
    assert L.cgb_conjugrad_files(*files, 2, 3, C.byref(o), x.ctypes.data, C.byref(st)) == NOTDIV
    h = C.c_void_p()
    assert L.cgb_create(None, 0) == ARG
    assert L.cgb_create_rank(C.byref(h), 0, 3, 2, None) == ARG
    assert L.cgb_create_rank(C.byref(h), 0, 0, 2, None) == ARG
    assert L.cgb_destroy(None) == 0
    for call in (lambda: L.cgb_solve(None, None, None),
                 lambda: L.cgb_load_host(None, 2, A.ctypes.data, b.ctypes.data, x.ctypes.data),
                 lambda: L.cgb_load_host_symmetric(None, 2, A.ctypes.data, b.ctypes.data, x.ctypes.data),
                 lambda: L.cgb_load_text(None, 2, b"a", b"b", b"c"),
                 lambda: L.cgb_generate(None, 2, 1, 0),
                 lambda: L.cgb_read_x(None, x.ctypes.data),
                 lambda: L.cgb_set_x0(None, x.ctypes.data),
                 lambda: L.cgb_debug_check_guards(None),
                 lambda: L.cgb_debug_poke(None, 1, 0, 1.0),
                 lambda: L.cgb_debug_read(None, 0, 0, 1, x.ctypes.data),
                 lambda: L.cgb_unique_id(None)):
        assert call() == ARG
        assert len(L.cgb_last_error()) > 0
    assert L.cgb_last_stats(None) == STATE
    d = (C.c_int64 * 4)()
    assert L.cgb_read_dimensions(None, d) == IO
    vb = files[1]
    assert L.cgb_read_values(None, x.ctypes.data, 2, None) == ARG
    assert L.cgb_read_values(vb, None, 2, None) == ARG
    assert L.cgb_read_values(vb, x.ctypes.data, -3, None) == ARG
    assert L.cgb_read_values(vb, x.ctypes.data, 0, None) == 0
    assert L.cgb_write_values(None, x.ctypes.data, 2, 4) == IO
    assert L.cgb_write_values(b"/nonexistent_dir/x.txt", x.ctypes.data, 2, 4) == IO
    assert L.cgb_count_values(None) == -1
    assert L.cgb_shutdown() == 0 and L.cgb_release_rank() == 0


def test_torch_imports_after_the_library_is_loaded():
    # libcgb200.so and torch both want a `libnccl.so.2`; a process holds one.  Loading the library first must
    # not break a later `import torch` (the binding maps torch's bundled copy when there is one), and the
    # reverse order -- bench.py's -- must keep working; CGB_NCCL_PRELOAD=0 opts out.
    import sys
    code = ("import conjugate_gradient_b200 as cg; L = cg.lib(); import torch; "
            "print('ok', L.cgb_version(), sum('libnccl' in l for l in open('/proc/self/maps')) > 0)")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=B.ROOT, timeout=600)
    assert r.returncode == 0 and "ok 100 True" in r.stdout, r.stderr[-2000:]
    code = ("import torch; import conjugate_gradient_b200 as cg; L = cg.lib(); "
            "print('ok', L.cgb_version(), len({l.split()[-1] for l in open('/proc/self/maps') if 'libnccl' in l}))")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=B.ROOT, timeout=600)
    assert r.returncode == 0 and "ok 100 1" in r.stdout, r.stderr[-2000:]          # ONE copy of NCCL mapped
    code = "import conjugate_gradient_b200 as cg; cg.lib(); print([l.split()[-1] for l in open('/proc/self/maps') if 'libnccl' in l][0])"
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=B.ROOT, timeout=600,
                       env=dict(os.environ, CGB_NCCL_PRELOAD="0"))
    assert r.returncode == 0 and "site-packages" not in r.stdout, r.stdout + r.stderr[-1000:]
