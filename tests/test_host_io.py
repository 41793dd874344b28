"""CPU tests of the host text path (cg_io.c): the parser that replaces initialize()'s
fscanf("%f%*c") loop (serialConjugate.c:85-105, parallel_cg.c:147-168) for b, x0 and irregular
files, and the writer of generateSPDmatrix.m's file format (:28-43).  The parser is
multi-threaded (the file is cut at separators and the pieces are converted in parallel), so the
properties are checked on files long enough to be split as well as on short ones.  Bit-exactness
bar: every value equals Python's float() of the same token (correctly rounded, as strtod)."""
import os
import time

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import conjugate_gradient_b200 as cg


def parse(tmp_path, text, count, name="f.txt"):
    p = tmp_path / name
    p.write_bytes(text if isinstance(text, bytes) else text.encode())
    return cg.read_values(str(p), count)


@pytest.mark.parametrize("fmt", ["%.4f", "%.1f", "%.17g", "%.6e", "%+.10E", "%g"])
def test_parser_matches_float_bit_for_bit(tmp_path, fmt):
    rng = np.random.default_rng(7)
    v = np.concatenate([rng.random(2000), rng.standard_normal(2000) * 1e3, rng.standard_normal(500) * 1e-300,
                        rng.standard_normal(500) * 1e300, [0.0, -0.0, 1.0, 4096.5, 5e-324, 1.7976931348623157e308]])
    toks = [fmt % x for x in v]
    got = parse(tmp_path, "\n".join(toks) + "\n", len(toks))
    want = np.array([float(t) for t in toks])
    assert got.tobytes() == want.tobytes()


def test_generator_file_format_round_trip(tmp_path):
    # generateSPDmatrix.m:29,35,41: "%.4f\n" for A and b, "%.1f\n" for x
    rng = np.random.default_rng(3)
    n = 300
    A = rng.random((n, n)); A = 0.5 * (A + A.T) + n * np.eye(n)
    fa, fx = str(tmp_path / "A.txt"), str(tmp_path / "x.txt")
    cg.write_values(fa, A, 4)
    lines = open(fa).read().split("\n")
    assert lines[-1] == "" and len(lines) == n * n + 1
    assert lines[0] == "%.4f" % A[0, 0] and lines[n * n - 1] == "%.4f" % A[-1, -1]
    back = cg.read_values(fa, n * n).reshape(n, n)
    assert np.abs(back - A).max() <= 0.5e-4 + 1e-12
    assert back.tobytes() == np.array([float("%.4f" % a) for a in A.ravel()]).reshape(n, n).tobytes()
    cg.write_values(fx, np.zeros(n), 1)
    assert open(fx).read() == "0.0\n" * n
    assert cg.lib().cgb_count_values(os.fsencode(fa)) == n * n


def test_large_file_is_split_between_threads_consistently(tmp_path):
    # 2 M tokens of uneven length: piece boundaries fall inside runs of separators and next to signs
    rng = np.random.default_rng(11)
    n = 2_000_000
    v = rng.standard_normal(n) * 10.0 ** rng.integers(-8, 8, n)
    seps = np.array(["\n", " ", ",", "\r\n", "\t", ";", " \n "])[rng.integers(0, 7, n)]
    toks = ["%.12g" % x for x in v]
    text = "".join(t + s for t, s in zip(toks, seps))
    got = parse(tmp_path, text, n)
    assert got.tobytes() == np.array([float(t) for t in toks]).tobytes()
    assert cg.lib().cgb_count_values(os.fsencode(str(tmp_path / "f.txt"))) == n
    with pytest.raises(cg.CgbError) as e:
        parse(tmp_path, text, n + 1)
    assert f"found {n}" in str(e.value)


@pytest.mark.parametrize("text,count,want", [
    (b"1 2 3", 3, [1, 2, 3]),                                   # no trailing newline
    (b"\n\n  1.5\n\n\n2.5   \n", 2, [1.5, 2.5]),                 # blank lines, trailing blanks
    (b"\xef\xbb\xbf0.5\n0.25\n", 2, [0.5, 0.25]),               # BOM first (the reference's initialguess1.txt)
    (b"1e3,1E-3;+7,-.5,5.", 5, [1000.0, 0.001, 7.0, -0.5, 5.0]),  # exponent forms, bare dot
    (b"inf -inf\n", 2, [np.inf, -np.inf]),
    (b"0x10\n", 1, [16.0]),                                      # strtod grammar: what "%lf" accepts
    (b"3 4 5 6", 2, [3, 4]),                                     # more tokens than asked for: the first `count`
])
def test_token_rules(tmp_path, text, count, want):
    np.testing.assert_array_equal(parse(tmp_path, text, count), np.array(want, dtype=float))


@pytest.mark.parametrize("text,count", [(b"", 1), (b"\n \n", 1), (b"1.0 2.0", 3), (b"1.0 abc 2.0", 3), (b"1.0 2.0x 3.0", 3)])
def test_short_or_malformed_files_are_errors_not_garbage(tmp_path, text, count):
    # serialConjugate.c:92-98 ignores fscanf's return value and leaves the tail uninitialised
    with pytest.raises(cg.CgbError) as e:
        parse(tmp_path, text, count)
    assert e.value.code == -3 or "expected" in str(e.value)


def test_dimensions_file(tmp_path):
    p = tmp_path / "dimensions.txt"
    p.write_text("4096\n4096\n4096\n1\n")
    assert cg.read_dimensions(str(p)) == (4096, 4096, 4096, 1)
    p.write_text("2 2\r\n2 1")
    assert cg.read_dimensions(str(p)) == (2, 2, 2, 1)
    p.write_text("2\n2\n2\n")
    with pytest.raises(cg.CgbError):
        cg.read_dimensions(str(p))
    with pytest.raises(cg.CgbError):
        cg.read_dimensions(str(tmp_path / "missing.txt"))


finite = st.floats(allow_nan=False, allow_infinity=False, width=64)


@settings(max_examples=200, deadline=None)
@given(st.lists(finite, min_size=1, max_size=60),
       st.lists(st.sampled_from(["\n", " ", ",", "\r\n", "\t", ";", "\n\n", " , "[1:2]]), min_size=60, max_size=60),
       st.sampled_from(["%r", "%.17g", "%.4f", "%.3e", "%+.8g"]))
def test_parser_property_any_doubles_any_separators(tmp_path_factory, values, seps, fmt):
    toks = [(repr(x) if fmt == "%r" else fmt % x) for x in values]
    text = "".join(t + s for t, s in zip(toks, seps))
    d = tmp_path_factory.mktemp("prop")
    got = parse(d, text, len(toks))
    assert got.tobytes() == np.array([float(t) for t in toks]).tobytes()


def test_writer_round_trips_every_double(tmp_path):
    rng = np.random.default_rng(5)
    bits = rng.integers(0, 2 ** 63, 20000, dtype=np.int64)
    v = bits.view(np.float64)
    v = v[np.isfinite(v)]
    f = str(tmp_path / "rt.txt")
    cg.write_values(f, v, -1)                                    # "%.17g"
    assert cg.read_values(f, v.shape[0]).tobytes() == v.tobytes()


# ---- the pieces the GPU ingest path shares with the host parser (cg_io.h) ----------------------
import ctypes as C


class Text(C.Structure):
    _fields_ = [("p", C.c_void_p), ("len", C.c_size_t), ("mapped", C.c_int)]


def _io():
    L = cg.lib()
    L.cgb_text_open.argtypes = [C.c_char_p, C.POINTER(Text)]; L.cgb_text_open.restype = C.c_int
    L.cgb_text_close.argtypes = [C.POINTER(Text)]; L.cgb_text_close.restype = None
    L.cgb_text_locate.argtypes = [C.POINTER(Text), C.c_int64, C.c_int64, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    L.cgb_text_locate.restype = C.c_int64
    L.cgb_token_to_double.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_double)]; L.cgb_token_to_double.restype = C.c_int
    return L


@pytest.mark.parametrize("ranks", [1, 2, 3, 8])
def test_locating_a_ranks_rows_in_the_text(tmp_path, ranks):
    """cgb_load_text: every rank finds the BYTES of its own m x n values without parsing the rest
    (what MPI_Scatter of the parsed matrix does in parallel_cg.c:112-115).  The ranges must start
    at the first byte of the rank's first value, contain exactly its values, and tile the file."""
    L = _io()
    rng = np.random.default_rng(ranks)
    n = 24 * ranks
    vals = rng.standard_normal(n * n) * 10.0 ** rng.integers(-3, 4, n * n)
    seps = np.array(["\n", " ", ",", "\r\n", "\t", "\n\n"])[rng.integers(0, 6, n * n)]
    toks = ["%.6g" % x for x in vals]
    raw = ("  \n" + "".join(t + s for t, s in zip(toks, seps))).encode()        # leading blanks too
    f = tmp_path / "A.txt"
    f.write_bytes(raw)
    tx = Text()
    assert L.cgb_text_open(os.fsencode(str(f)), C.byref(tx)) == 0
    try:
        m = n // ranks
        prev_end = None
        for r in range(ranks):
            beg, end = C.c_size_t(), C.c_size_t()
            total = L.cgb_text_locate(C.byref(tx), r * m * n, m * n, C.byref(beg), C.byref(end))
            assert total == n * n
            piece = raw[beg.value:end.value].decode()
            got = piece.replace(",", " ").split()
            assert got == toks[r * m * n:(r + 1) * m * n]
            assert not chr(raw[beg.value]).isspace() and raw[beg.value:beg.value + len(toks[r * m * n])].decode() == toks[r * m * n]
            if prev_end is not None:
                assert prev_end <= beg.value and raw[prev_end:beg.value].decode().replace(",", " ").strip() == ""
            prev_end = end.value
        # asking for more values than the file holds: the count says so
        beg, end = C.c_size_t(), C.c_size_t()
        assert L.cgb_text_locate(C.byref(tx), 0, n * n + 5, C.byref(beg), C.byref(end)) == n * n
    finally:
        L.cgb_text_close(C.byref(tx))


@pytest.mark.parametrize("threads", [1, 3, 8])
def test_locate_on_several_spans_against_a_byte_level_reference(tmp_path, threads, monkeypatch):
    """cgb_text_locate counts token starts 16 bytes at a time (SSE2) per 1 MiB block, the blocks shared out between
    the host threads, and seeks the rank's first / one-past-last token inside one block by skipping whole groups: tokens of 1..40 bytes, runs of every separator
    the reader tolerates (initialize()'s "%f%*c", serialConjugate.c:92-98), tokens that straddle group and span
    borders, a file of a few MiB so that there ARE several spans; checked against numpy on the raw bytes."""
    L = _io()
    monkeypatch.setenv("CGB_IO_THREADS", str(threads))
    rng = np.random.default_rng(100 + threads)
    ntok = 260_000
    lens = rng.choice([1, 2, 3, 6, 7, 15, 16, 17, 31, 40], ntok)
    digits = np.frombuffer(b"0123456789", np.uint8)
    seps = np.frombuffer(b" \t\n\r,;\f\v", np.uint8)
    pieces = [b"\n \t"]                                                       # leading separators
    seplens = rng.choice([1, 1, 1, 2, 3, 17, 33], ntok)
    body = digits[rng.integers(0, 10, int(lens.sum()))]
    sepb = seps[rng.integers(0, 8, int(seplens.sum()))]
    bo = np.concatenate(([0], np.cumsum(lens))); so = np.concatenate(([0], np.cumsum(seplens)))
    for k in range(ntok):
        pieces.append(body[bo[k]:bo[k + 1]].tobytes()); pieces.append(sepb[so[k]:so[k + 1]].tobytes())
    raw = b"".join(pieces)
    if threads == 3: raw = raw.rstrip(b" \t\n\r,;\f\v")                        # last token ends the file
    assert len(raw) > 3 << 20
    a = np.frombuffer(raw, np.uint8)
    is_sep = np.isin(a, seps)
    starts = np.flatnonzero(~is_sep & np.concatenate(([True], is_sep[:-1])))
    assert len(starts) == ntok
    f = tmp_path / "tokens.txt"
    f.write_bytes(raw)
    tx = Text()
    assert L.cgb_text_open(os.fsencode(str(f)), C.byref(tx)) == 0
    try:
        beg, end = C.c_size_t(), C.c_size_t()
        cases = [(0, ntok), (0, 1), (ntok - 1, 1), (ntok // 2, ntok // 2), (ntok // 3, ntok // 3), (5, 0)]
        cases += [(int(x), int(c)) for x, c in zip(rng.integers(0, ntok - 1000, 40), rng.integers(1, 1000, 40))]
        for first, count in cases:
            assert L.cgb_text_locate(C.byref(tx), first, count, C.byref(beg), C.byref(end)) == ntok
            assert beg.value == starts[first], (first, count)
            want_end = starts[first] if count == 0 else (starts[first + count] if first + count < ntok else len(raw))
            assert end.value == want_end, (first, count)
        assert L.cgb_text_locate(C.byref(tx), 10, ntok, C.byref(beg), C.byref(end)) == ntok     # past the end: the count says so
        assert beg.value == end.value == len(raw)
    finally:
        L.cgb_text_close(C.byref(tx))


def test_locate_defers_files_with_a_bom_to_the_sequential_parser(tmp_path):
    L = _io()
    f = tmp_path / "bom.txt"
    f.write_bytes(b"\xef\xbb\xbf1.0\n2.0\n")
    tx = Text()
    assert L.cgb_text_open(os.fsencode(str(f)), C.byref(tx)) == 0
    beg, end = C.c_size_t(), C.c_size_t()
    assert L.cgb_text_locate(C.byref(tx), 0, 2, C.byref(beg), C.byref(end)) == -1
    L.cgb_text_close(C.byref(tx))
    np.testing.assert_array_equal(cg.read_values(str(f), 2), [1.0, 2.0])       # which handles it


@pytest.mark.parametrize("tok,ok", [("1.5", True), ("-0.0001", True), ("+3", True), ("1e-7", True), ("6.02E+23", True),
                                    (".5", True), ("5.", True), ("123456789012345678901234567890", True),
                                    ("1-2", False), ("abc", False), ("1.2.3", False), ("", False), ("--1", False),
                                    ("1e", False), ("e5", False)])
def test_single_token_conversion(tok, ok):
    L = _io()
    b = tok.encode()
    buf = C.create_string_buffer(b, len(b) + 1)
    base = C.addressof(buf)
    out = C.c_double(-777.0)
    r = L.cgb_token_to_double(C.cast(base, C.c_char_p), C.cast(base + len(b), C.c_char_p), C.byref(out))
    assert bool(r) == ok
    if ok:
        assert out.value == float(tok)


def test_appending_writer_builds_a_matrix_file_block_by_block(tmp_path):
    # --emit-files writes the row blocks of a generated matrix in turn (cgb_write_values_mode, append)
    L = cg.lib()
    L.cgb_write_values_mode.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_int, C.c_int]
    L.cgb_write_values_mode.restype = C.c_int
    rng = np.random.default_rng(2)
    A = rng.random((6, 5))
    f = str(tmp_path / "blocks.txt")
    for k, blk in enumerate(np.array_split(A, 3)):
        blk = np.ascontiguousarray(blk)
        assert L.cgb_write_values_mode(os.fsencode(f), blk.ctypes.data, blk.size, 4, 1 if k else 0) == 0
    assert open(f).read() == "".join("%.4f\n" % a for a in A.ravel())


@pytest.mark.parametrize("sep", ["\n", "\r\n", " ", ",", "\t", " \n", ";"])
@pytest.mark.parametrize("fmt", ["%.4f", "%.1f", "%.17g", "%e"])
def test_parser_agrees_with_the_real_initialize(tmp_path, sep, fmt):
    """The REAL reader -- initialize() of serialConjugate.c:85-105 in the float->double build of oracle/_ref
    (fscanf "%lf%*c", ROWS = 4) -- and cgb_read_values give the same doubles on every separator the reference's
    "one arbitrary character after each number" rule tolerates."""
    import ctypes as C
    import _oracle as O
    if not O.ref_available("serial", "f64", 4):
        pytest.skip("oracle/_ref not built")
    L = C.CDLL(os.path.join(O.REF_DIR, "libserial_f64_n4.so"))
    L.initialize.argtypes = [C.c_void_p, C.c_char_p, C.c_int]; L.initialize.restype = None
    rng = np.random.default_rng(len(sep) * 31 + len(fmt))
    for cols in (1, 4):                                        # a vector and the 4 x 4 matrix
        vals = (rng.random(4 * cols) - 0.3) * 10.0 ** rng.integers(-3, 6, 4 * cols)
        path = tmp_path / f"v{cols}.txt"
        path.write_text(sep.join(fmt % v for v in vals) + sep)
        ref = np.full(4 * cols, np.nan)
        L.initialize(ref.ctypes.data, os.fsencode(str(path)), cols)
        got = cg.read_values(str(path), 4 * cols)
        assert got.tobytes() == ref.tobytes()
        assert got.tobytes() == np.array([float(fmt % v) for v in vals]).tobytes()


def test_print_format_is_the_real_printers(tmp_path):
    """printer() of serialConjugate.c:260-276 (disabled at its call site, :252) run from the real build: the CLI's
    --print-x lines (cgb_print_vector, asserted in tests/test_gpu_parity.py::test_cli_on_shipped_fixture) are its bytes."""
    import subprocess
    import sys
    import _oracle as O
    if not O.ref_available("serial", "f64", 4):
        pytest.skip("oracle/_ref not built")
    code = ("import ctypes as C, numpy as np\n"
            f"L = C.CDLL({os.path.join(O.REF_DIR, 'libserial_f64_n4.so')!r})\n"
            "L.printer.argtypes = [C.c_void_p, C.c_int, C.c_int]; L.printer.restype = None\n"
            "v = np.array([2 / 3, 1 / 3, -1.0, 12345.678901]); L.printer(v.ctypes.data, 4, 1); L.printer(v.ctypes.data, 2, 2)\n"
            "C.CDLL(None).fflush(None)\n")
    real = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, check=True).stdout
    assert real.splitlines()[:2] == ["0.666667  ", "0.333333  "]
    # the product's formatter on the same data, through a tiny C probe (cg_io.c is plain C)
    src = tmp_path / "probe.c"
    src.write_text('#include "cg_io.h"\nint main(void){double v[4]={2.0/3,1.0/3,-1.0,12345.678901};'
                   'cgb_print_vector(stdout,v,4,1);cgb_print_vector(stdout,v,2,2);return 0;}\n')
    exe = tmp_path / "probe"
    csrc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "conjugate_gradient_b200", "csrc")
    subprocess.run(["gcc", str(src), os.path.join(csrc, "cg_io.c"), "-I", csrc, "-o", str(exe), "-lpthread", "-lm"], check=True)
    ours = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    assert ours == real


def test_host_thread_counts_follow_the_cpu_affinity_mask():
    """io_threads(): a process pinned to two CPUs (taskset, a container's cpuset) gets a copy team of two -- its
    workers spin between chunks and must not outnumber the CPUs they may run on (8 spinning on 2 CPUs measured
    4x slower than 2)."""
    import subprocess
    import sys
    if len(os.sched_getaffinity(0)) < 2:
        pytest.skip("needs two CPUs")
    two = sorted(os.sched_getaffinity(0))[:2]
    code = ("import os, ctypes as C\n"
            f"os.sched_setaffinity(0, {{{two[0]}, {two[1]}}})\n"
            "import conjugate_gradient_b200 as cg\n"
            "L = cg.lib(); L.cgb_copy_pool_threads.argtypes = [C.c_size_t]; L.cgb_copy_pool_threads.restype = C.c_int\n"
            "print(L.cgb_copy_pool_threads(8 << 20), L.cgb_copy_pool_threads(1 << 20))\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                       env={k: v for k, v in os.environ.items() if k != "CGB_IO_THREADS"})
    assert r.returncode == 0, r.stderr[-1500:]
    assert r.stdout.split() == ["2", "2"]


def test_ranks_of_one_process_share_the_counting_pass(tmp_path, monkeypatch):
    """cgb_text_locate: `cg_b200 --gpus G` runs its ranks as threads and each needs the token counts of the whole
    file; the first one counts, the others wait and reuse its spans (one pass instead of G).  The shared entry is
    for the same UNCHANGED file only: a rewrite (same path, same size) is counted again, and
    CGB_LOCATE_NO_CACHE=1 turns the sharing off."""
    import threading
    L = _io()
    L.cgb_text_locate_passes.argtypes = []; L.cgb_text_locate_passes.restype = C.c_ulong
    monkeypatch.delenv("CGB_LOCATE_NO_CACHE", raising=False)
    ranks, n = 8, 512
    rng = np.random.default_rng(5)

    def write(path):
        toks = ["%.4f" % v for v in rng.random(n * n)]
        raw = ("\n".join(toks) + "\n").encode()
        path.write_bytes(raw)
        a = np.frombuffer(raw, np.uint8)
        return raw, np.flatnonzero((a != 10) & np.concatenate(([True], a[:-1] == 10)))

    f = tmp_path / "A.txt"
    raw, starts = write(f)
    assert len(raw) > (1 << 20)                                                  # several spans
    m = n // ranks

    def rank_main(r, out):
        tx = Text()
        assert L.cgb_text_open(os.fsencode(str(f)), C.byref(tx)) == 0            # every rank maps the file itself
        beg, end = C.c_size_t(), C.c_size_t()
        total = L.cgb_text_locate(C.byref(tx), r * m * n, m * n, C.byref(beg), C.byref(end))
        out[r] = (total, beg.value, end.value)
        L.cgb_text_close(C.byref(tx))

    def all_ranks():
        out = [None] * ranks
        th = [threading.Thread(target=rank_main, args=(r, out)) for r in range(ranks)]
        before = L.cgb_text_locate_passes()
        for t in th: t.start()
        for t in th: t.join()
        return out, L.cgb_text_locate_passes() - before

    def check(out, starts, raw):
        for r, (total, beg, end) in enumerate(out):
            assert total == n * n and beg == starts[r * m * n]
            assert end == (starts[(r + 1) * m * n] if r + 1 < ranks else len(raw))

    out, passes = all_ranks()
    check(out, starts, raw)
    assert passes == 1                                                           # eight ranks, one pass
    out, passes = all_ranks()
    check(out, starts, raw)
    assert passes == 0                                                           # the same unchanged file again
    time.sleep(0.01)
    raw2, starts2 = write(f)                                                     # same path, same size, other digits
    assert len(raw2) == len(raw)
    out, passes = all_ranks()
    check(out, starts2, raw2)
    assert passes == 1
    monkeypatch.setenv("CGB_LOCATE_NO_CACHE", "1")
    out, passes = all_ranks()
    check(out, starts2, raw2)
    assert passes == ranks
