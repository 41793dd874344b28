"""ctypes bindings for the CPU oracle (oracle/liboracle.so) and, when present,
the real reference compiled into oracle/_ref.  TEST INFRASTRUCTURE ONLY: imported
by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by
the product package."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
GEN_ROUND4, GEN_SPECTRUM, GEN_GRAM = 1, 2, 4


class OrcStats(C.Structure):
    _fields_ = [("iterations", C.c_int64), ("break_k", C.c_int64),
                ("first_hit_k", C.c_int64), ("final_beta", C.c_double),
                ("rsold0", C.c_double), ("seconds", C.c_double),
                ("loop_seconds", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"], check=True)


def lib():
    """liboracle.so, built on demand (gcc only; no GPU needed)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(path):
        build_oracle()
    L = C.CDLL(path)
    dp, fp, i64, u64 = C.POINTER(C.c_double), C.POINTER(C.c_float), C.c_int64, C.c_uint64
    for sfx, ptr, real in (("f64", dp, C.c_double), ("f32", fp, C.c_float)):
        getattr(L, f"orc_matvec_{sfx}").argtypes = [ptr, ptr, ptr, i64, i64]
        getattr(L, f"orc_matvec_{sfx}").restype = None
        getattr(L, f"orc_vecvec_{sfx}").argtypes = [ptr, ptr, i64]
        getattr(L, f"orc_vecvec_{sfx}").restype = real
        getattr(L, f"orc_conjugrad_{sfx}").argtypes = [ptr, ptr, ptr, i64, C.c_double, i64, C.c_int, C.POINTER(OrcStats)]
        getattr(L, f"orc_conjugrad_{sfx}").restype = None
        getattr(L, f"orc_conjugrad_spmd_{sfx}").argtypes = [ptr, ptr, ptr, i64, i64, C.c_double, i64, C.c_int, C.POINTER(OrcStats)]
        getattr(L, f"orc_conjugrad_spmd_{sfx}").restype = None
    L.orc_gen_fill_A.argtypes = [dp, i64, i64, i64, u64, C.c_uint]
    L.orc_gen_fill_A.restype = None
    L.orc_gen_fill_b.argtypes = [dp, i64, i64, i64, u64, C.c_uint]
    L.orc_gen_fill_b.restype = None
    L.orc_gen_sample_u.argtypes = [u64, u64]
    L.orc_gen_sample_u.restype = C.c_double
    L.orc_team_solve_generated.argtypes = [i64, u64, C.c_uint, i64, C.c_double, i64, C.c_int, dp, C.POINTER(OrcStats), dp]
    L.orc_team_solve_generated.restype = C.c_int
    L.orc_time_matvec_sample.argtypes = [i64, i64, u64, C.c_int, dp]
    L.orc_time_matvec_sample.restype = C.c_double
    _lib = L
    return L


def _ptr(a):
    ct = C.c_double if a.dtype == np.float64 else C.c_float
    return a.ctypes.data_as(C.POINTER(ct))


def GEN_KAPPA_LOG2(k):
    """oracle/spd_gen_ref.h ORC_GEN_KAPPA_LOG2: the spread-diagonal family, cond(A) ~ 2^k."""
    return (int(k) & 0x3F) << 8


def gen_A(n, seed, flags=0, row0=0, m=None):
    m = n if m is None else m
    A = np.empty((m, n), dtype=np.float64)
    lib().orc_gen_fill_A(_ptr(A), n, row0, m, seed, flags)
    return A


def gen_b(n, seed, flags=0, row0=0, m=None):
    m = n if m is None else m
    b = np.empty(m, dtype=np.float64)
    lib().orc_gen_fill_b(_ptr(b), n, row0, m, seed, flags)
    return b


def matvec(A, v):
    A = np.ascontiguousarray(A); v = np.ascontiguousarray(v)
    sfx = "f64" if A.dtype == np.float64 else "f32"
    y = np.empty(A.shape[0], dtype=A.dtype)
    getattr(lib(), f"orc_matvec_{sfx}")(_ptr(y), _ptr(A), _ptr(v), A.shape[0], A.shape[1])
    return y


def vecvec(a, b):
    sfx = "f64" if a.dtype == np.float64 else "f32"
    return getattr(lib(), f"orc_vecvec_{sfx}")(_ptr(a), _ptr(b), a.shape[0])


def conjugrad(A, b, x0, eps=1.0e-6, max_iter=None, ignore_stop=False, ranks=None):
    """Oracle solve.  ranks=None -> serial form (serialConjugate.c:180-259);
    ranks=P -> row-block SPMD form (parallel_cg.c:248-345).  Returns (x, stats)."""
    A = np.ascontiguousarray(A); b = np.ascontiguousarray(b)
    x = np.array(x0, dtype=A.dtype, copy=True)
    n = A.shape[0]
    sfx = "f64" if A.dtype == np.float64 else "f32"
    st = OrcStats()
    mi = n if max_iter is None else max_iter
    if ranks is None:
        getattr(lib(), f"orc_conjugrad_{sfx}")(_ptr(A), _ptr(b), _ptr(x), n, eps, mi, int(ignore_stop), C.byref(st))
    else:
        assert n % ranks == 0
        getattr(lib(), f"orc_conjugrad_spmd_{sfx}")(_ptr(A), _ptr(b), _ptr(x), n, ranks, eps, mi, int(ignore_stop), C.byref(st))
    return x, st.as_dict()


def cg_fused(A, b, x0, eps=1.0e-6, max_iter=None, ignore_stop=False):
    """Chronopoulos-Gear CG (oracle/cg_oracle.c orc_cg_fused_f64): the CPU statement of the product's
    cg_variant = CGB_CG_FUSED_REDUCTIONS.  Returns (x, stats)."""
    A = np.ascontiguousarray(A, dtype=np.float64); b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.array(x0, dtype=np.float64, copy=True)
    n = A.shape[0]
    st = OrcStats()
    f = lib().orc_cg_fused_f64
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int64, C.c_int, C.POINTER(OrcStats)]
    f.restype = None
    f(A.ctypes.data, b.ctypes.data, x.ctypes.data, n, eps, n if max_iter is None else max_iter, int(ignore_stop), C.byref(st))
    return x, st.as_dict()


def team_solve_generated(n, seed, ranks, flags=0, eps=1.0e-6, max_iter=None, ignore_stop=False, want_x=True):
    x = np.empty(n, dtype=np.float64) if want_x else None
    st = OrcStats(); gs = C.c_double(0.0)
    rc = lib().orc_team_solve_generated(n, seed, flags, ranks, eps, n if max_iter is None else max_iter,
                                        int(ignore_stop), _ptr(x) if want_x else None, C.byref(st), C.byref(gs))
    if rc != 0:
        raise ValueError(f"{n} is not divisible by {ranks}")
    d = st.as_dict(); d["gen_seconds"] = gs.value
    return x, d


def spd_with_condition(n, kappa, seed):
    """Dense SPD test matrix with a log-spaced spectrum in [1, kappa] (Q diag(w) Q') and a
    random right-hand side: CG needs ~sqrt(kappa) * ln(1/eps) iterations on it, unlike the
    kappa ~ 1.5 systems of generateSPDmatrix.m that converge in ~5."""
    rng = np.random.default_rng(seed)
    Q, _ = np.linalg.qr(rng.standard_normal((n, n)))
    A = (Q * np.logspace(0.0, np.log10(kappa), n)) @ Q.T
    return np.ascontiguousarray(0.5 * (A + A.T)), rng.standard_normal(n)


# ---------------------------------------------------------------- real reference
def ref_available(kind, prec, n):
    return os.path.exists(os.path.join(REF_DIR, f"lib{kind}_{prec}_n{n}.so"))


def ref_conjugrad(A, b, x0, ranks=None, program="parallel"):
    """Run the REAL reference's conjugrad() (compiled by oracle/build_ref.sh) on
    in-memory arrays.  Precision follows A.dtype; n must have been built.
    ranks=None: serialConjugate.c; ranks=P: parallel_cg.c, or with program="p2p"
    point-to-point_cg.c (same signature, collectives emulated with sends and receives)."""
    A = np.ascontiguousarray(A); b = np.ascontiguousarray(b)
    n = A.shape[0]
    prec = "f64" if A.dtype == np.float64 else "f32"
    x = np.array(x0, dtype=A.dtype, copy=True)
    vp = C.c_void_p
    if ranks is None:
        L = C.CDLL(os.path.join(REF_DIR, f"libserial_{prec}_n{n}.so"))
        L.conjugrad.argtypes = [vp, vp, vp]; L.conjugrad.restype = None
        L.conjugrad(A.ctypes.data, b.ctypes.data, x.ctypes.data)
    else:
        L = C.CDLL(os.path.join(REF_DIR, f"lib{program}_{prec}_n{n}.so"))
        L.cgshim_run_conjugrad.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int]
        L.cgshim_run_conjugrad.restype = C.c_int
        fn = C.cast(L.conjugrad, vp)
        rc = L.cgshim_run_conjugrad(fn, A.ctypes.data, b.ctypes.data, x.ctypes.data, n, ranks, A.dtype.itemsize)
        if rc != 0:
            raise ValueError(f"{n} is not divisible by {ranks}")
    return x
