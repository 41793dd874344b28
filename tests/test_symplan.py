"""CPU check of the row-sharded symmetric product's work split (cg_symv.cu, cgk_symplan): for every
rank count and block size the tiles the ranks read, together with their mirror images, cover the
matrix exactly once; every rank reads (about) half of its block; the sender / receiver tables agree.
A numpy emulation of the three kernels (tile, push, finish) on those tables reproduces A @ p."""
import ctypes as C

import numpy as np
import pytest

import conjugate_gradient_b200 as cg

T = 512


class SymPlan(C.Structure):
    _fields_ = [("nblk", C.c_int), ("blk_rank", C.c_int * 8), ("row_end", C.c_int * 8), ("col_begin", C.c_int * 8),
                ("nsrc", C.c_int), ("src_rank", C.c_int * 8), ("src_col_begin", C.c_int * 8)]


def plan_and_tiles(rank, ranks, mt):
    L = cg.lib()
    L.cgk_symv_make_plan.argtypes = [C.POINTER(SymPlan), C.c_int, C.c_int, C.c_int64]
    L.cgk_symv_make_plan.restype = None
    L.cgk_symv_tile_list_rank.argtypes = [C.POINTER(SymPlan), C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_int)]
    L.cgk_symv_tile_list_rank.restype = C.c_int64
    sp = SymPlan()
    m = mt * T
    L.cgk_symv_make_plan(C.byref(sp), rank, ranks, m)
    cnt = L.cgk_symv_tile_list_rank(C.byref(sp), rank, m, 1, None)
    buf = (C.c_int * (2 * cnt))()
    assert L.cgk_symv_tile_list_rank(C.byref(sp), rank, m, 1, buf) == cnt
    assert L.cgk_symv_tile_list_rank(C.byref(sp), rank, m, 4, None) == 4 * cnt
    return sp, [(buf[2 * k], buf[2 * k + 1]) for k in range(cnt)]


@pytest.mark.parametrize("ranks", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("mt", [1, 2, 3, 4])
def test_tiles_and_mirrors_cover_the_matrix_exactly_once(ranks, mt):
    assert cg.lib().cgk_symv_tile() == T
    ntl = ranks * mt
    cover = np.zeros((ntl, ntl), dtype=int)
    per_rank = []
    plans = []
    for r in range(ranks):
        sp, tiles = plan_and_tiles(r, ranks, mt)
        plans.append(sp)
        per_rank.append(len(tiles))
        for il, j in tiles:
            ig = r * mt + il
            assert 0 <= il < mt and 0 <= j < ntl
            cover[ig, j] += 1
            if ig != j:
                cover[j, ig] += 1
        # own-block tiles come first (their p needs no exchange)
        own = [k for k, (il, j) in enumerate(tiles) if j // mt == r]
        assert own == list(range(len(own)))
    assert (cover == 1).all()
    # balance: half of the block, up to the diagonal and the odd split
    ideal = ranks * mt * mt / 2.0
    assert max(per_rank) <= ideal + mt and min(per_rank) >= ideal - mt * mt / 2.0 - mt
    # whoever reads a block of my columns is in my source table with the same column range
    for r in range(ranks):
        senders = {}
        for g in range(ranks):
            for b in range(plans[g].nblk):
                if plans[g].blk_rank[b] == r:
                    senders[g] = plans[g].col_begin[b]
        mine = {plans[r].src_rank[q]: plans[r].src_col_begin[q] for q in range(plans[r].nsrc)}
        assert mine == senders


@pytest.mark.parametrize("ranks,mt", [(1, 2), (2, 1), (2, 2), (3, 1), (4, 2), (8, 1)])
def test_numpy_emulation_of_the_three_kernels_gives_A_times_p(ranks, mt):
    tt = 4                                  # a small tile edge: the split only depends on tile indices
    m, n = mt * tt, ranks * mt * tt
    rng = np.random.default_rng(ranks * 10 + mt)
    A = rng.standard_normal((n, n)); A = 0.5 * (A + A.T)
    p = rng.standard_normal(n)
    y = np.zeros(n)
    recv = np.zeros((ranks, ranks, m))      # recv[owner][source][local column]
    plans, rowpart, colpart = [], [], []
    for r in range(ranks):
        sp, tiles = plan_and_tiles(r, ranks, mt)
        plans.append(sp)
        rp = np.zeros((ranks * mt, m)); cp = np.zeros((mt, n))
        blk = A[r * m:(r + 1) * m]
        for il, j in tiles:                 # symv_tile_comm
            tile = blk[il * tt:(il + 1) * tt, j * tt:(j + 1) * tt]
            rp[j, il * tt:(il + 1) * tt] = tile @ p[j * tt:(j + 1) * tt]
            if r * mt + il != j:
                cp[il, j * tt:(j + 1) * tt] = tile.T @ p[r * m + il * tt: r * m + (il + 1) * tt]
        rowpart.append(rp); colpart.append(cp)
        for b in range(sp.nblk):            # symv_push
            h = sp.blk_rank[b]
            lo = sp.col_begin[b] * tt
            recv[h, r, lo:] = colpart[r][:sp.row_end[b], h * m + lo:(h + 1) * m].sum(axis=0)
    for r in range(ranks):                  # symv_finish_comm
        sp = plans[r]
        for i in range(m):
            il = i // tt
            s = rowpart[r][r * mt + il:(r + 1) * mt, i].sum()
            for b in range(sp.nblk):
                if il < sp.row_end[b]:
                    h = sp.blk_rank[b]
                    s += rowpart[r][h * mt + sp.col_begin[b]:(h + 1) * mt, i].sum()
            s += colpart[r][:il, r * m + i].sum()
            for q in range(sp.nsrc):
                if il >= sp.src_col_begin[q]:
                    s += recv[r, sp.src_rank[q], i]
            y[r * m + i] = s
    np.testing.assert_allclose(y, A @ p, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("ranks", [2, 3, 4, 8])
@pytest.mark.parametrize("mt", [1, 2, 4, 16])
@pytest.mark.parametrize("split", [1, 2, 4])
def test_finish_kernels_split_the_partial_list_exactly_once(ranks, mt, split):
    """cgb_halfmat_finish_comm_kernel (cg_symv.cu) walks the flat list [own block's row partials | each foreign
    block's | own column partials] with 8 warps at stride 8, carrying the offset from one segment into
    the next (k -= cnt).  Restated here: every term is visited by exactly one warp, whatever the
    segment lengths."""
    NW = 8
    for r in range(ranks):
        sp, _ = plan_and_tiles(r, ranks, mt)
        for il in range(mt):
            want = [("row", r * mt + il + k) for k in range(mt - il)]
            for b in range(sp.nblk):
                if il < sp.row_end[b]:
                    want += [("row", sp.blk_rank[b] * mt + sp.col_begin[b] + k) for k in range(mt - sp.col_begin[b])]
            want += [("col", k) for k in range(il * split)]
            got = []
            for wid in range(NW):
                k = wid
                cnt, j0 = mt - il, r * mt + il
                while k < cnt:
                    got.append(("row", j0 + k)); k += NW
                k -= cnt
                for b in range(sp.nblk):
                    if il < sp.row_end[b]:
                        cnt, j0 = mt - sp.col_begin[b], sp.blk_rank[b] * mt + sp.col_begin[b]
                        while k < cnt:
                            got.append(("row", j0 + k)); k += NW
                        k -= cnt
                cnt = il * split
                while k < cnt:
                    got.append(("col", k)); k += NW
            assert sorted(got) == sorted(want) and len(set(got)) == len(got)


def test_grid_helpers_of_the_deferred_reductions():
    L = cg.lib()
    L.cgk_symv_finish_grid.argtypes = [C.c_int64]; L.cgk_symv_finish_grid.restype = C.c_int
    L.cgk_vec_grid.argtypes = [C.c_int64]; L.cgk_vec_grid.restype = C.c_int
    assert L.cgk_symv_finish_grid(4096) == 4096 // 32       # 32 rows per CTA
    assert L.cgk_vec_grid(1) == 1 and L.cgk_vec_grid(256) == 1 and L.cgk_vec_grid(257) == 2
    assert L.cgk_vec_grid(10 ** 9) == 592                    # capped: the consumer adds at most 592 partials
