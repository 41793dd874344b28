"""CPU tests of the launch-plan logic (cg_kernels.cu: cgk_gemv_make_plan, cgk_gemv_make_plan_comm,
cgk_symv_pick_split): pure host code that decides which hand-written kernel shape a problem gets.
The shapes themselves were picked from B200 sweeps (profiles/r01_tune_gemv.md, r01_l2_keep_sweep.txt,
r01_deferred_reductions.txt); these tests pin the decisions, the vector-width fallbacks for odd /
unaligned shapes and the L2-keep window, so that a later edit cannot silently change what runs."""
import ctypes as C

import pytest

import conjugate_gradient_b200 as cg

LDG, BULK = 1, 2
L2 = 126 * 1024 * 1024 + 0          # B200: cudaDeviceProp.l2CacheSize is about 126 MiB


class Plan(C.Structure):
    _fields_ = [("variant", C.c_int), ("rows_per_cta", C.c_int), ("unroll", C.c_int), ("threads", C.c_int),
                ("vec", C.c_int), ("stages", C.c_int), ("tile_cols", C.c_int), ("grid", C.c_int),
                ("smem_bytes", C.c_int), ("keep_pct", C.c_int), ("l2_bytes", C.c_longlong), ("keep_rows", C.c_longlong)]


def make(m, n, a=0x7f0000000000, v=0x7f1000000000, comm=False, **req):
    L = cg.lib()
    f = L.cgk_gemv_make_plan_comm if comm else L.cgk_gemv_make_plan
    f.argtypes = [C.POINTER(Plan), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int]
    f.restype = C.c_int
    p = Plan(**req)
    assert f(C.byref(p), m, n, a, v, 148) == 0
    return p


def shape(p):
    return (p.rows_per_cta, p.unroll, p.threads, p.vec)


@pytest.mark.parametrize("n,want", [(65536, (8, 2, 512, 4)), (131072, (8, 2, 512, 4)), (16384, (2, 1, 256, 4)),
                                    (8192, (2, 1, 256, 4)), (4096, (4, 1, 128, 4)), (1024, (4, 1, 128, 4))])
def test_default_shapes_per_size_class(n, want):
    p = make(n, n, l2_bytes=L2)
    assert p.variant == LDG and shape(p) == want
    assert p.grid == -(-n // p.rows_per_cta)                    # one CTA per R rows: ceil(m / R)


def test_vector_width_follows_shape_and_alignment():
    assert make(1000, 1000).vec == 4                            # n % 4 == 0, 32-byte aligned bases
    assert shape(make(1002, 1002)) == (4, 1, 128, 2)            # n % 4 != 0: 128-bit loads
    assert shape(make(1001, 1001)) == (4, 1, 128, 1)            # odd n: scalar loads, still a CUDA kernel
    assert make(1000, 1000, a=0x7f0000000010).vec == 2          # A only 16-byte aligned
    assert make(1000, 1000, v=0x7f1000000008).vec == 1          # v only 8-byte aligned
    assert make(1000, 1000, vec=2).vec == 2 and make(1000, 1000, vec=1).vec == 1   # explicit narrower request
    assert make(1001, 1001, vec=4).vec == 1                     # a wider request than the shape allows is reduced
    p = make(3, 300)                                            # fewer rows than R
    assert p.rows_per_cta <= 3 and p.grid >= 1


def test_l2_keep_window():
    # on when A is between 0.8 and 3 L2 sizes; keep_rows = 60 % of L2 / (8 n), a multiple of R
    p = make(4096, 4096, l2_bytes=L2)                           # 134 MB
    assert p.keep_rows == (int(L2 * 0.6 / (8 * 4096)) // p.rows_per_cta) * p.rows_per_cta and 2300 < p.keep_rows < 2500
    assert make(2048, 2048, l2_bytes=L2).keep_rows == 0         # 33 MB: resident by itself
    assert make(8192, 8192, l2_bytes=L2).keep_rows == 0         # 537 MB > 3 L2: measured 1 % slower with keeping
    assert make(6144, 6144, l2_bytes=L2).keep_rows > 0          # 302 MB
    assert make(4096, 4096, l2_bytes=L2, keep_pct=-1).keep_rows == 0            # opt out
    assert make(4096, 4096, l2_bytes=0).keep_rows == 0                          # no L2 information: plain streaming
    q = make(4096, 4096, l2_bytes=L2, keep_pct=30)
    assert 0 < q.keep_rows < p.keep_rows
    assert make(2048, 4096, l2_bytes=L2).keep_rows == 0         # a 2-rank row block of the same matrix: 67 MB


def test_explicit_requests_and_fallbacks():
    p = make(8192, 8192, rows_per_cta=4, unroll=2, threads=256)
    assert shape(p) == (4, 2, 256, 4)
    L = cg.lib()
    bad = Plan(rows_per_cta=3, unroll=1, threads=128)           # no such instantiation
    assert L.cgk_gemv_make_plan(C.byref(bad), 512, 512, 0x7f0000000000, 0x7f1000000000, 148) != 0
    # the bulk-copy ring cannot tile n % tile_cols != 0: the streaming-load kernel takes over
    p = make(1000, 1000, variant=BULK)
    assert p.variant == LDG


@pytest.mark.parametrize("m,n,want", [(32768, 65536, (4, 2, 256, 4)), (8192, 65536, (4, 2, 256, 4)),
                                      (8192, 16384, (4, 2, 256, 4)), (2048, 4096, (4, 1, 128, 4))])
def test_comm_kernel_shapes(m, n, want):
    p = make(m, n, comm=True)
    assert shape(p) == want and p.grid == -(-m // p.rows_per_cta)


def test_comm_kernel_vector_width_needs_block_boundaries_aligned():
    assert make(1002, 2004, comm=True).vec == 2                 # m % 4 != 0: block boundaries not 32-byte aligned
    assert make(1001, 2002, comm=True).vec == 1
    assert make(1000, 2000, comm=True).vec == 4


def test_symmetric_product_split():
    L = cg.lib()
    L.cgk_symv_pick_split.argtypes = [C.c_int64, C.c_int]; L.cgk_symv_pick_split.restype = C.c_int
    assert L.cgk_symv_pick_split(8256, 148) == 1                # n = 65536 on one GPU: 8256 tiles
    assert L.cgk_symv_pick_split(136, 148) == 1                 # >= 0.9 x SMs
    assert L.cgk_symv_pick_split(100, 148) == 2                 # >= SMs / 2
    assert L.cgk_symv_pick_split(36, 148) == 4                  # n = 4096: few tiles, 4 CTAs each


# ----------------------------------------------------------------- staging ring of malloc'ed host buffers (cg_stage.c)
MiB, GiB = 1 << 20, 1 << 30
STAGED, DRIVER = 2, 3


def ring_plan(total, ranks, team, have=0):
    L = cg.lib()
    L.cgb_stage_plan.argtypes = [C.c_size_t, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                 C.POINTER(C.c_size_t)]
    L.cgb_stage_plan.restype = None
    mode, T, chunk = C.c_int(), C.c_int(), C.c_size_t()
    L.cgb_stage_plan(total, ranks, team, (1 if ranks > 1 else 8) * MiB, have, C.byref(mode), C.byref(T), C.byref(chunk))
    return mode.value, T.value, chunk.value


def test_staging_ring_shapes_are_the_measured_ones():
    """cgb_stage_plan: what the ring of pinned slots looks like per upload size and rank count.  Every line is a
    configuration measured on B200 boxes (profiles/r02_h2d_*.json, r02_first_call_slots.txt); the N > 1 lines in
    particular cannot be re-measured cheaply, so an edit that changes them must fail here."""
    # one rank, the bench size (34 GB): the full ring, 8 threads x 2 slots x 8 MiB
    assert ring_plan(8 * 65536 * 65536, 1, 8) == (STAGED, 8, 8 * MiB)
    # one rank, asserted-symmetric upload of the same system (17 GB): still the full ring
    assert ring_plan(8 * 65536 * 65537 // 2, 1, 8) == (STAGED, 8, 8 * MiB)
    # one rank, one-shot sizes: total / 40 capped at 16 MiB, a thread per 4 MiB of ring
    m, T, ch = ring_plan(8 * 16384 * 16384, 1, 8)
    assert (m, T, ch) == (STAGED, 4, 2 * MiB)
    m, T, ch = ring_plan(8 * 8192 * 8192, 1, 8)                # 537 MB -> 13.4 MB of ring
    assert m == STAGED and T == 4 and 1.5 * MiB <= ch <= 2 * MiB and ch % 65536 == 0
    m, T, ch = ring_plan(8 * 4096 * 4096, 1, 8)                # 134 MB -> 3.4 MB of ring: two threads
    assert m == STAGED and T == 2 and 768 * 1024 <= ch <= MiB
    # below 64 MiB: the driver's own path
    assert ring_plan(8 * 2048 * 2048, 1, 8)[0] == DRIVER
    # a context that already holds a larger ring uses it
    assert ring_plan(8 * 8192 * 8192, 1, 8, have=128 * MiB) == (STAGED, 8, 8 * MiB)
    # several ranks uploading at once: 1 MiB slots (8 ranks on 32 CPUs: 4 threads each; 2 ranks: 8 threads each)
    assert ring_plan(8 * 8192 * 65536, 8, 4) == (STAGED, 4, 1 * MiB)
    assert ring_plan(8 * 16384 * 65536, 4, 8) == (STAGED, 8, 1 * MiB)
    assert ring_plan(8 * 32768 * 65536, 2, 8) == (STAGED, 8, 1 * MiB)
    assert ring_plan(8 * 16384 * 131072, 8, 4) == (STAGED, 4, 1 * MiB)
    # several ranks, a small system: the ring shrinks with the bytes, the team does not fall below 256 KiB slots
    m, T, ch = ring_plan(8 * 1024 * 8192, 8, 4)                # 67 MB per rank -> 2 MiB of ring
    assert m == STAGED and T * 2 * ch <= 2 * MiB and ch >= 256 * 1024
