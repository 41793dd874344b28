"""GPU text ingest (cg_ingest.cu / cg_ingest.c) against the host parser: the reference's
initialize() (serialConjugate.c:85-105, "%f%*c") replaced by tokenising and decimal->binary
conversion on the device.  Bar: bit-identical doubles to strtod (= Python float()), the same
accepted separators, and loud errors for short or missing files."""
import os

import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cg():
    import conjugate_gradient_b200 as m
    assert m.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return m


@pytest.fixture(scope="module")
def ctx(cg):
    with cg.Context(0) as c:
        yield c


def write_tokens(path, tokens, sep="\n"):
    with open(path, "w", newline="") as f:
        f.write(sep.join(tokens) + sep)


def check_load(cg, ctx, tmp_path, n, a_tokens, sep="\n", raw=None):
    fa, fb, fx = (str(tmp_path / s) for s in ("A.txt", "b.txt", "x.txt"))
    if raw is None:
        write_tokens(fa, a_tokens, sep)
    else:
        open(fa, "wb").write(raw)
    b = np.arange(n) * 0.25 - 1.0
    x0 = np.linspace(-1, 1, n)
    cg.write_values(fb, b, -1)
    cg.write_values(fx, x0, -1)
    ctx.load_text(n, fa, fb, fx)
    want = np.array([float(t) for t in a_tokens], dtype=np.float64)
    got = ctx.debug_read(0, 0, n * n)
    assert got.tobytes() == want[: n * n].tobytes()
    assert ctx.debug_read(1, 0, n).tobytes() == b.tobytes()
    assert ctx.debug_read(2, 0, n).tobytes() == x0.tobytes()


def test_generator_format_file_is_parsed_bit_exactly(cg, ctx, tmp_path):
    n = 300                                                   # generateSPDmatrix.m:28-30 "%.4f\n"
    A = O.gen_A(n, 3, O.GEN_ROUND4)
    check_load(cg, ctx, tmp_path, n, ["%.4f" % v for v in A.ravel()])


def test_number_forms_and_separators(cg, ctx, tmp_path):
    forms = ["1.5", "-2.25", "+3", "0", "-0.0", ".5", "5.", "000.250", "1e3", "1.5E-3", "2e+2", "-7.125e0",
             "123456789012345", "0.000000000000001", "9.87654321e10", "4096.0001", "1e22", "1e-22", "12345.6789e-5",
             # outside the exact fast path: resolved by strtod on the host, still bit-identical
             "0.1234567890123456789", "1e23", "1e-300", "1.7976931348623157e308", "4.9e-324", "123456789012345678",
             "inf", "-inf", "1e400"]
    n = 8
    toks = [forms[i % len(forms)] for i in range(n * n)]
    check_load(cg, ctx, tmp_path, n, toks)
    check_load(cg, ctx, tmp_path, n, toks, sep=" ")
    check_load(cg, ctx, tmp_path, n, toks, sep="\r\n")
    check_load(cg, ctx, tmp_path, n, toks, sep=",\t")
    check_load(cg, ctx, tmp_path, n, toks, sep=" ;  \n\n")


def test_seventeen_digit_files_use_the_exception_path_and_the_overflow_fallback(cg, ctx, tmp_path):
    rng = np.random.default_rng(5)
    for n in (100, 300):           # 10 000 exceptions (list); 90 000 > list capacity (host re-parse)
        v = rng.standard_normal(n * n)
        check_load(cg, ctx, tmp_path, n, [repr(float(t)) for t in v])


def test_bom_and_irregular_tokens_follow_the_sequential_parser(cg, ctx, tmp_path):
    n = 3
    toks = ["%d.5" % i for i in range(9)]
    raw = ("\n".join(toks[:4]) + "\n").encode() + b"\xef\xbb\xbf" + ("\n".join(toks[4:]) + "\n").encode()
    check_load(cg, ctx, tmp_path, n, toks, raw=raw)           # the reference ships a file with a BOM
    # "1-2" is two values for strtod-style parsing ("1", "-2"), one token for a tokeniser
    raw = b"1-2 3 4 5 6 7 8 9\n"
    check_load(cg, ctx, tmp_path, n, ["1", "-2", "3", "4", "5", "6", "7", "8", "9"], raw=raw)


def test_multi_chunk_file(cg, ctx, tmp_path):
    n = 2200                                                  # ~43 MB of text: six 8 MiB chunks (it was two of 32 MiB)
    fa, fb, fx = (str(tmp_path / s) for s in ("A.txt", "b.txt", "x.txt"))
    A = (np.random.default_rng(1).random(n * n) * 2048.0).round(4)
    cg.write_values(fa, A, 4)
    assert os.path.getsize(fa) > (32 << 20)
    cg.write_values(fb, np.ones(n), 4)
    cg.write_values(fx, np.zeros(n), 1)
    ctx.load_text(n, fa, fb, fx)
    want = cg.read_values(fa, n * n)                          # the host parser (itself checked against float())
    assert ctx.debug_read(0, 0, n * n).tobytes() == want.tobytes()
    # extra values after the n*n-th are ignored, as fscanf's fixed count does (serialConjugate.c:92-98)
    ctx.load_text(n - 1, fa, fb, fx)
    assert ctx.debug_read(0, 0, (n - 1) * (n - 1)).tobytes() == want[: (n - 1) * (n - 1)].tobytes()


def test_every_staging_chunk_size_gives_the_host_parsers_doubles(cg, ctx, tmp_path, monkeypatch):
    """The bytes of a chunk reach the GPU through a pinned buffer filled by several host threads.  Round 2 found
    that copy dropping up to 7 tail bytes of a chunk whose length / threads was a whole number of pages (8 MiB + 3):
    the GPU then parsed whatever the recycled pinned buffer held there ("0.5618" read as "0.5").  Fresh buffers are
    zero-filled and zeros go to the exception list, so a single load hid it; loads in a row with different chunk
    sizes (CGB_INGEST_CHUNK_MIB) recycle pinned memory that holds text and do not."""
    n = 2200
    fa, fb, fx = (str(tmp_path / s) for s in ("A.txt", "b.txt", "x.txt"))
    A = (np.random.default_rng(7).random(n * n) * 64.0).round(4)
    cg.write_values(fa, A, 4)
    cg.write_values(fb, np.ones(n), 4)
    cg.write_values(fx, np.zeros(n), 1)
    want = cg.read_values(fa, n * n).tobytes()
    for mib in (32, 8, 1, 2, 4, 8, 1):
        monkeypatch.setenv("CGB_INGEST_CHUNK_MIB", str(mib))
        ctx.load_text(n, fa, fb, fx)
        assert ctx.debug_read(0, 0, n * n).tobytes() == want, f"chunks of {mib} MiB"


def test_short_and_missing_files_are_errors(cg, ctx, tmp_path):
    fa, fb, fx = (str(tmp_path / s) for s in ("A.txt", "b.txt", "x.txt"))
    cg.write_values(fa, np.ones(15), 4)
    cg.write_values(fb, np.ones(4), 4)
    cg.write_values(fx, np.zeros(4), 1)
    with pytest.raises(cg.CgbError) as e:
        ctx.load_text(4, fa, fb, fx)
    assert e.value.code == -7 and "expected 16 values, found 15" in str(e.value)
    with pytest.raises(cg.CgbError) as e:
        ctx.load_text(4, str(tmp_path / "nope.txt"), fb, fx)
    assert e.value.code == -7 and "Could not open file" in str(e.value)
    cg.write_values(fb, np.ones(3), 4)
    cg.write_values(fa, np.ones(16), 4)
    with pytest.raises(cg.CgbError):
        ctx.load_text(4, fa, fb, fx)


@pytest.mark.parametrize("gpus", [1, 2, 4])
def test_files_entry_point_matches_the_oracle(cg, golden_dir, tmp_path, gpus):
    if cg.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    g = golden_dir
    x, st = cg.conjugrad_files(f"{g}/g2_matrixA.txt", f"{g}/g2_vectorb.txt", f"{g}/g2_x0.txt", 4, gpus=gpus)
    np.testing.assert_allclose(x, [-1, 1, -1, 1], rtol=1e-12)          # the shipped 4x4 fixture
    n, seed = 512, 12
    A, b = O.gen_A(n, seed, O.GEN_ROUND4), O.gen_b(n, seed, O.GEN_ROUND4)
    fa, fb, fx = (str(tmp_path / s) for s in ("A.txt", "b.txt", "x.txt"))
    cg.write_values(fa, A, 4); cg.write_values(fb, b, 4); cg.write_values(fx, np.zeros(n), 1)
    x, st = cg.conjugrad_files(fa, fb, fx, n, gpus=gpus)
    xo, so = O.conjugrad(A, b, np.zeros(n), ranks=None if gpus == 1 else gpus)
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= 1e-10
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["h2d_seconds"] > 0
