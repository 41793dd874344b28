"""ctypes binding of libcgb200.so (include/cgb200.h).  See the package docstring."""
import ctypes as C
import os
import re
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "lib", "libcgb200.so")
CLI_PATH = os.path.join(PKG_DIR, "lib", "cg_b200")
HEADER = os.path.join(ROOT, "include", "cgb200.h")

CGB_EPSILON = 1.0e-6          # serialConjugate.c:28


GEN_ROUND4, GEN_SPECTRUM, GEN_GRAM = 1, 2, 4


def GEN_KAPPA_LOG2(k):
    """cgb200.h CGB_GEN_KAPPA_LOG2: generator family with condition number ~2^k."""
    return (int(k) & 0x3F) << 8


GEMV_AUTO, GEMV_LDG, GEMV_BULK = 0, 1, 2
COMM_AUTO, COMM_NCCL, COMM_PEER = 0, 1, 2


class CgbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libcgb200 error {code}: {msg}")
        self.code = code


class Options(C.Structure):
    _fields_ = [("eps", C.c_double), ("max_iter", C.c_int64), ("ignore_stop", C.c_int),
                ("use_graph", C.c_int), ("check_every", C.c_int), ("time_gemv", C.c_int),
                ("gemv_variant", C.c_int), ("gemv_rows", C.c_int), ("gemv_unroll", C.c_int),
                ("gemv_threads", C.c_int), ("gemv_vec", C.c_int), ("gemv_stages", C.c_int),
                ("gemv_tile_cols", C.c_int), ("comm", C.c_int), ("guard", C.c_int), ("gemv_l2_keep_pct", C.c_int), ("symmetric", C.c_int), ("cg_variant", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("n", C.c_int64), ("m", C.c_int64), ("rank", C.c_int), ("ranks", C.c_int),
                ("iterations", C.c_int64), ("break_k", C.c_int64), ("first_hit_k", C.c_int64),
                ("final_beta", C.c_double), ("final_residual", C.c_double), ("rsold0", C.c_double),
                ("solve_seconds", C.c_double), ("loop_ms", C.c_double), ("init_ms", C.c_double),
                ("h2d_seconds", C.c_double), ("d2h_seconds", C.c_double),
                ("gemv_ms_avg", C.c_double), ("gemv_ms_min", C.c_double), ("gemv_timed", C.c_int64),
                ("gemv_bytes", C.c_double), ("kernel_launches", C.c_int64),
                ("gemv_variant", C.c_int), ("gemv_rows", C.c_int), ("gemv_unroll", C.c_int),
                ("gemv_threads", C.c_int), ("gemv_vec", C.c_int), ("gemv_stages", C.c_int),
                ("gemv_tile_cols", C.c_int), ("gemv_grid", C.c_int), ("graph_chunk", C.c_int),
                ("comm", C.c_int), ("exit_reason", C.c_int), ("cg_variant", C.c_int),
                ("gemv_keep_rows", C.c_int64), ("h2d_bytes", C.c_int64), ("pdl", C.c_int), ("h2d_path", C.c_int)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


CG_CLASSIC, CG_FUSED_REDUCTIONS = 0, 1


def header_symbols():
    """Every function name declared in include/cgb200.h."""
    with open(HEADER) as f:
        txt = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    names = re.findall(r"\b(cgb_[a-z0-9_]+|conjugrad(?:_mpi)?)\s*\(", txt)
    return sorted(set(names))


HEADER_SYMBOLS = header_symbols()


def build(force=False):
    """make -C csrc: nvcc (sm_100a) for the kernels, gcc for the C host code."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.run(["make", "-s", "-C", os.path.join(PKG_DIR, "csrc")], check=True)
    return LIB_PATH


_lib = None


def _share_nccl_with_torch():
    """libcgb200.so links `libnccl.so.2`; PyTorch wheels bundle their own, usually newer, library under
    the same soname (site-packages/nvidia/nccl/lib) and need its symbols.  A process can hold only one
    `libnccl.so.2`, and the first one loaded serves everybody: with the system copy loaded first, a later
    `import torch` fails on an undefined NCCL symbol.  So when no NCCL is mapped yet and the bundled copy
    exists, load that one first -- it is the copy `bench.py` (which imports torch first) has always run on.
    CGB_NCCL_PRELOAD=0 keeps the system library.  Returns the path loaded, or None."""
    if os.environ.get("CGB_NCCL_PRELOAD", "1") == "0":
        return None
    try:
        with open("/proc/self/maps") as f:
            if "libnccl.so" in f.read():
                return None
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                return cand
    except (OSError, ImportError, ValueError, AttributeError):
        pass
    return None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `make -C conjugate_gradient_b200/csrc` "
                          "(or __graft_entry__.build()); there is no fallback implementation")
    _share_nccl_with_torch()
    L = C.CDLL(LIB_PATH)
    dp, i64, u64, vp = C.POINTER(C.c_double), C.c_int64, C.c_uint64, C.c_void_p
    op, sp = C.POINTER(Options), C.POINTER(Stats)
    sig = {
        "cgb_version": ([], C.c_int), "cgb_device_count": ([], C.c_int),
        "cgb_last_error": ([], C.c_char_p), "cgb_default_options": ([op], None),
        "cgb_create": ([C.POINTER(vp), C.c_int], C.c_int), "cgb_unique_id": ([vp], C.c_int),
        "cgb_create_rank": ([C.POINTER(vp), C.c_int, C.c_int, C.c_int, vp], C.c_int),
        "cgb_destroy": ([vp], C.c_int),
        "cgb_load_host": ([vp, i64, vp, vp, vp], C.c_int),
        "cgb_load_host_symmetric": ([vp, i64, vp, vp, vp], C.c_int),
        "cgb_debug_check_guards": ([vp], C.c_int),
        "cgb_debug_poke": ([vp, C.c_int, i64, C.c_double], C.c_int),
        "cgb_debug_probe_read": ([vp, vp, i64], i64),
        "cgb_matvec_noalloc": ([vp, vp, vp, vp, vp, i64, i64, vp], C.c_int),
        "cgb_debug_rendezvous_id": ([C.c_int, C.c_int, vp], C.c_int),
        "cgb_load_text": ([vp, i64, C.c_char_p, C.c_char_p, C.c_char_p], C.c_int),
        "cgb_conjugrad_files": ([C.c_char_p, C.c_char_p, C.c_char_p, i64, C.c_int, op, vp, sp], C.c_int),
        "cgb_generate": ([vp, i64, u64, C.c_uint], C.c_int), "cgb_set_x0": ([vp, vp], C.c_int),
        "cgb_solve": ([vp, op, sp], C.c_int), "cgb_read_x": ([vp, vp], C.c_int),
        "cgb_conjugrad_host": ([vp, vp, vp, i64, C.c_double, i64, C.c_int, sp], C.c_int),
        "cgb_conjugrad_host_opt": ([vp, vp, vp, i64, C.c_int, op, sp], C.c_int),
        "cgb_solve_generated": ([i64, u64, C.c_uint, C.c_int, op, vp, sp], C.c_int),
        "cgb_shutdown": ([], C.c_int),
        "cgb_conjugrad_rank": ([vp, vp, vp, vp, i64, op, sp], C.c_int),
        "cgb_set_dimension": ([i64], None), "conjugrad": ([vp, vp, vp], None),
        "conjugrad_mpi": ([vp, vp, vp, C.c_int, C.c_int, C.c_int], None),
        "cgb_last_stats": ([sp], C.c_int), "cgb_release_rank": ([], C.c_int),
        "cgb_read_dimensions": ([C.c_char_p, C.POINTER(i64)], C.c_int),
        "cgb_read_values": ([C.c_char_p, vp, i64, C.POINTER(i64)], C.c_int),
        "cgb_count_values": ([C.c_char_p], i64),
        "cgb_write_values": ([C.c_char_p, vp, i64, C.c_int], C.c_int),
        "cgb_op_matvec": ([vp, vp, vp, vp, vp, dp, i64, i64, op], C.c_int),
        "cgb_op_init_residual": ([vp, vp, vp, vp, vp, dp, i64], C.c_int),
        "cgb_op_xr_update": ([vp, C.c_double, C.c_double, vp, vp, vp, vp, dp, i64], C.c_int),
        "cgb_op_p_update": ([vp, C.c_double, C.c_double, C.c_double, vp, vp, C.POINTER(C.c_int), i64], C.c_int),
        "cgb_op_generate": ([vp, i64, i64, i64, u64, C.c_uint, vp, vp], C.c_int),
        "cgb_time_gemv": ([vp, op, C.c_int, C.c_int, C.c_int, dp, dp], C.c_int),
        "cgb_debug_read": ([vp, C.c_int, i64, i64, vp], C.c_int),
        "cgb_device_info": ([vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(i64), C.POINTER(i64), C.c_char_p], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(L, name)          # AttributeError = the library lacks a declared symbol
        fn.argtypes, fn.restype = args, res
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise CgbError(rc, lib().cgb_last_error().decode("utf-8", "replace"))


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        assert a.shape == shape, (a.shape, shape)
    return a


def device_count():
    return lib().cgb_device_count()


def unique_id():
    buf = C.create_string_buffer(128)
    _check(lib().cgb_unique_id(buf))
    return buf.raw


def make_options(**kw):
    o = Options()
    lib().cgb_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(f"unknown option {k}")
        setattr(o, k, v)
    return o


class Context:
    """One rank on one GPU (cgb_ctx).  rank/ranks/uid mirror parallel_cg.c's
    MPI_Comm_rank / MPI_Comm_size; uid is the NCCL id shared by the launcher."""

    def __init__(self, device=0, rank=0, ranks=1, uid=None):
        self._h = C.c_void_p()
        if ranks == 1:
            _check(lib().cgb_create(C.byref(self._h), device))
        else:
            idbuf = C.create_string_buffer(uid, 128)
            _check(lib().cgb_create_rank(C.byref(self._h), device, rank, ranks, idbuf))
        self.rank, self.ranks, self.device = rank, ranks, device
        self.n = 0

    def close(self):
        if self._h:
            lib().cgb_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- problem set-up
    def load_host(self, A_rows, b_rows, x0_full):
        x0 = _f64(x0_full)
        n = x0.shape[0]
        m = n // self.ranks
        A = A_rows if (isinstance(A_rows, np.ndarray) and A_rows.dtype == np.float64 and A_rows.flags.c_contiguous) else _f64(A_rows)
        assert A.size == m * n, "A_rows must be the rank's (n/ranks) x n block"
        b = _f64(b_rows, (m,))
        _check(lib().cgb_load_host(self._h, n, A.ctypes.data, b.ctypes.data, x0.ctypes.data))
        self.n = n

    def load_host_symmetric(self, A_rows, b_rows, x0_full):
        """load_host for a matrix asserted symmetric: only the tiles the symmetric product reads travel."""
        x0 = _f64(x0_full)
        n = x0.shape[0]
        m = n // self.ranks
        A = A_rows if (isinstance(A_rows, np.ndarray) and A_rows.dtype == np.float64 and A_rows.flags.c_contiguous) else _f64(A_rows)
        assert A.size == m * n, "A_rows must be the rank's (n/ranks) x n block"
        b = _f64(b_rows, (m,))
        _check(lib().cgb_load_host_symmetric(self._h, n, A.ctypes.data, b.ctypes.data, x0.ctypes.data))
        self.n = n

    def load_host_ptr(self, n, A_ptr, b_ptr, x0_ptr):
        """Raw host pointers (e.g. pinned torch tensors' data_ptr())."""
        _check(lib().cgb_load_host(self._h, n, A_ptr, b_ptr, x0_ptr))
        self.n = n

    def load_text(self, n, file_a, file_b, file_x):
        """The reference's three text files; this rank's rows of A are parsed on the GPU."""
        _check(lib().cgb_load_text(self._h, n, os.fsencode(file_a), os.fsencode(file_b), os.fsencode(file_x)))
        self.n = n

    def generate(self, n, seed, flags=0):
        _check(lib().cgb_generate(self._h, n, seed, flags))
        self.n = n

    def set_x0(self, x0_full):
        x0 = _f64(x0_full, (self.n,))
        _check(lib().cgb_set_x0(self._h, x0.ctypes.data))

    # ---- solver
    def solve(self, **opts):
        o = make_options(**opts)
        st = Stats()
        _check(lib().cgb_solve(self._h, C.byref(o), C.byref(st)))
        return st.as_dict()

    def read_x(self, out=None):
        x = np.empty(self.n, dtype=np.float64) if out is None else out
        _check(lib().cgb_read_x(self._h, x.ctypes.data))
        return x

    def conjugrad_rank(self, local_A, local_b, local_x, **opts):
        """parallel_cg.c:248 conjugrad(local_A, local_b, local_x, m, rank, np): local_x in/out."""
        A = _f64(local_A); b = _f64(local_b)
        assert isinstance(local_x, np.ndarray) and local_x.dtype == np.float64
        o = make_options(**opts)
        st = Stats()
        _check(lib().cgb_conjugrad_rank(self._h, A.ctypes.data, b.ctypes.data, local_x.ctypes.data,
                                        b.shape[0], C.byref(o), C.byref(st)))
        self.n = local_x.shape[0]
        return st.as_dict()

    # ---- single-kernel operators
    def op_matvec(self, A, v, w=None, **opts):
        A = _f64(A); v = _f64(v)
        m, n = A.shape
        y = np.empty(m, dtype=np.float64)
        dot = C.c_double(0.0)
        wd = None if w is None else _f64(w, (m,))
        o = make_options(**opts)
        _check(lib().cgb_op_matvec(self._h, A.ctypes.data, v.ctypes.data,
                                   None if wd is None else wd.ctypes.data, y.ctypes.data,
                                   C.byref(dot), m, n, C.byref(o)))
        return (y, dot.value) if w is not None else y

    def op_init_residual(self, b, Ax):
        b = _f64(b); Ax = _f64(Ax)
        m = b.shape[0]
        r = np.empty(m); p = np.empty(m); rr = C.c_double()
        _check(lib().cgb_op_init_residual(self._h, b.ctypes.data, Ax.ctypes.data, r.ctypes.data,
                                          p.ctypes.data, C.byref(rr), m))
        return r, p, rr.value

    def op_xr_update(self, rho, pAp, x, r, p, Ap):
        x = np.array(x, dtype=np.float64); r = np.array(r, dtype=np.float64)
        p = _f64(p); Ap = _f64(Ap); rr = C.c_double()
        _check(lib().cgb_op_xr_update(self._h, rho, pAp, x.ctypes.data, r.ctypes.data, p.ctypes.data,
                                      Ap.ctypes.data, C.byref(rr), x.shape[0]))
        return x, r, rr.value

    def op_p_update(self, rho, beta, eps, p, r):
        p = np.array(p, dtype=np.float64); r = _f64(r); stopped = C.c_int()
        _check(lib().cgb_op_p_update(self._h, rho, beta, eps, p.ctypes.data, r.ctypes.data,
                                     C.byref(stopped), p.shape[0]))
        return p, bool(stopped.value)

    def op_generate(self, n, seed, flags=0, row0=0, m=None, want_A=True, want_b=True):
        m = n if m is None else m
        A = np.empty((m, n), dtype=np.float64) if want_A else None
        b = np.empty(m, dtype=np.float64) if want_b else None
        _check(lib().cgb_op_generate(self._h, n, row0, m, seed, flags,
                                     A.ctypes.data if want_A else None, b.ctypes.data if want_b else None))
        return A, b

    # ---- measurement
    def time_gemv(self, warmup=3, reps=10, flush_l2=False, **opts):
        o = make_options(**opts)
        avg, mn = C.c_double(), C.c_double()
        _check(lib().cgb_time_gemv(self._h, C.byref(o), warmup, reps, int(flush_l2), C.byref(avg), C.byref(mn)))
        return avg.value, mn.value

    def check_guards(self):
        """CGB_DEBUG_GUARDS=1: raise CgbError if a kernel wrote outside one of the context's device arrays."""
        _check(lib().cgb_debug_check_guards(self._h))

    def debug_poke(self, what, offset, value):
        _check(lib().cgb_debug_poke(self._h, what, offset, value))

    def probe_read(self):
        """CGB_PROBE=1: the exchange kernels' per-CTA time stamps of their last launch, as an (N, 8) uint64 array."""
        words = 8 * (16384 + 2 * 592)
        out = np.zeros(words, dtype=np.uint64)
        got = lib().cgb_debug_probe_read(self._h, out.ctypes.data, words)
        if got < 0:
            _check(int(got))
        return out[:got].reshape(-1, 8)

    def debug_read(self, what, offset, count):
        out = np.empty(count, dtype=np.float64)
        _check(lib().cgb_debug_read(self._h, what, offset, count, out.ctypes.data))
        return out

    def device_info(self):
        dev, sms, fr, tot = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
        name = C.create_string_buffer(64)
        _check(lib().cgb_device_info(self._h, C.byref(dev), C.byref(sms), C.byref(fr), C.byref(tot), name))
        return {"device": dev.value, "sm_count": sms.value, "free_bytes": fr.value,
                "total_bytes": tot.value, "name": name.value.decode()}


def conjugrad_host(A, b, x0, gpus=1, **opts):
    """serialConjugate.c:180 conjugrad(A, b, x) with run-time n; returns (x, stats)."""
    A = _f64(A); b = _f64(b)
    x = np.array(x0, dtype=np.float64, copy=True)
    o = make_options(**opts)
    st = Stats()
    _check(lib().cgb_conjugrad_host_opt(A.ctypes.data, b.ctypes.data, x.ctypes.data, x.shape[0], gpus,
                                        C.byref(o), C.byref(st)))
    return x, st.as_dict()


def conjugrad_files(file_a, file_b, file_x, n, gpus=1, **opts):
    """`prog matrixA.txt vectorb.txt initialguess.txt` (serialConjugate.c:62-68) on `gpus` GPUs."""
    x = np.empty(n, dtype=np.float64)
    o = make_options(**opts)
    st = Stats()
    _check(lib().cgb_conjugrad_files(os.fsencode(file_a), os.fsencode(file_b), os.fsencode(file_x), n, gpus,
                                     C.byref(o), x.ctypes.data, C.byref(st)))
    return x, st.as_dict()


def solve_generated(n, seed, gpus=1, flags=0, want_x=True, **opts):
    x = np.empty(n, dtype=np.float64) if want_x else None
    o = make_options(**opts)
    st = Stats()
    _check(lib().cgb_solve_generated(n, seed, flags, gpus, C.byref(o), x.ctypes.data if want_x else None, C.byref(st)))
    return x, st.as_dict()


def read_dimensions(path):
    d = (C.c_int64 * 4)()
    _check_io(lib().cgb_read_dimensions(os.fsencode(path), d), path)
    return tuple(d)


def read_values(path, count):
    out = np.empty(count, dtype=np.float64)
    got = C.c_int64()
    rc = lib().cgb_read_values(os.fsencode(path), out.ctypes.data, count, C.byref(got))
    if rc != 0:
        raise CgbError(rc, f"{path}: expected {count} values, found {got.value}")
    return out


def write_values(path, v, decimals=-1):
    v = _f64(v).ravel()
    _check_io(lib().cgb_write_values(os.fsencode(path), v.ctypes.data, v.shape[0], decimals), path)


def _check_io(rc, path):
    if rc != 0:
        raise CgbError(rc, f"cannot read/write {path}")
