"""Multi-GPU parity (row-sharded A; the all-gather of p and the all-reduce of the dot scalars of
parallel_cg.c:287-313 either as NCCL calls or fused into the kernels as peer-memory pushes).  Skipped on boxes with fewer GPUs than the case needs."""
import os
import subprocess
import sys

import numpy as np
import pytest

import _oracle as O
from conftest import unhex

pytestmark = pytest.mark.gpu
EPS, RTOL_X = 1.0e-6, 1.0e-10


@pytest.fixture(scope="module")
def cg():
    import conjugate_gradient_b200 as m
    return m


def need(cg, g):
    if cg.device_count() < g:
        pytest.skip(f"needs {g} GPUs")


def relerr(x, ref):
    return np.linalg.norm(x - ref) / np.linalg.norm(ref)


COMMS = [1, 2]          # CGB_COMM_NCCL, CGB_COMM_PEER


@pytest.mark.parametrize("comm", COMMS)
@pytest.mark.parametrize("gpus", [2, 4, 8])
@pytest.mark.parametrize("tag", ["g2", "g3", "g4"])
def test_sharded_solve_matches_reference_mpi_golden(cg, golden, gpus, tag, comm):
    need(cg, gpus)
    c = golden["cases"][tag]
    n = c["n"]
    if n % gpus or gpus not in c["ranks"]:
        pytest.skip("rank count not in the golden set")
    if "A" in c:
        A, b, x0 = unhex(c["A"]).reshape(n, n), unhex(c["b"]), unhex(c["x0"])
    else:
        A, b, x0 = O.gen_A(n, c["seed"]), O.gen_b(n, c["seed"]), np.zeros(n)
    x, st = cg.conjugrad_host(A, b, x0, gpus=gpus, comm=comm)
    want = unhex(c["x"][f"spmd{gpus}_f64"])       # the REAL parallel_cg.c (double build) over the MPI shim
    xo, so = O.conjugrad(A, b, x0, ranks=gpus)
    assert relerr(x, want) <= RTOL_X
    assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS
    assert st["ranks"] == gpus and st["m"] == n // gpus and st["comm"] == comm


@pytest.mark.parametrize("comm", COMMS)
@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_long_running_sharded_solve(cg, gpus, comm):
    # ~47 iterations: epochs, slot parity and the replayed graph of the fused exchange.
    # kappa = 30: see test_gpu_parity.py (at kappa = 100 the reference's own serial and SPMD
    # orders already differ by more than the 1e-10 bar)
    need(cg, gpus)
    n = 512
    A, b = O.spd_with_condition(n, 30.0, seed=5)
    for use_graph in (1, 0):
        x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus, comm=comm, use_graph=use_graph)
        xo, so = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
        assert so["iterations"] > 40 and st["comm"] == comm
        assert relerr(x, xo) <= RTOL_X
        assert abs(st["iterations"] - so["iterations"]) <= 1 and st["final_residual"] < EPS


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_merged_vector_kernel_gives_the_two_kernel_bits(cg, gpus, monkeypatch):
    """CGB_XRP=1: xr_update, the r.r all-reduce and p_update as one launch (cgb_xrp_update_kernel).  Same
    arithmetic in the same order as the two-kernel form, so the iterates agree bit for bit -- over a long
    run (epochs, both slot parities, the break in the middle of a captured graph) and in fixed-iteration mode."""
    need(cg, gpus)
    n, seed, flags = 2048, 23, cg.GEN_KAPPA_LOG2(7)
    xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n), ranks=gpus)
    monkeypatch.setenv("CGB_XRP", "0")
    x2, st2 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, comm=2)
    f2, sf2 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, comm=2, ignore_stop=1, max_iter=11)
    monkeypatch.setenv("CGB_XRP", "1")
    for use_graph in (1, 0):
        x1, st1 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, comm=2, use_graph=use_graph)
        assert st1["iterations"] == st2["iterations"] and st1["exit_reason"] == 1 and so["iterations"] > 80
        np.testing.assert_array_equal(x1, x2)
        assert relerr(x1, xo) <= RTOL_X and abs(st1["iterations"] - so["iterations"]) <= 1
        # one launch less per iteration
        assert st1["kernel_launches"] < st2["kernel_launches"]
    f1, sf1 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, comm=2, ignore_stop=1, max_iter=11)
    assert sf1["iterations"] == 11
    np.testing.assert_array_equal(f1, f2)


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_pageable_rank_blocks_go_through_the_staging_ring(cg, gpus):
    """cgb_conjugrad_host on malloc'ed arrays, several ranks uploading at once: every rank's block goes through its
    own staging ring (cg_stage.c) and the solve sees exactly the caller's matrix."""
    need(cg, gpus)
    n, seed = 4096, 19                       # 16 MiB per rank at 8 ranks: above the small-copy cut-off
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus)
    xo, so = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    assert st["h2d_path"] in (2, 3), st["h2d_path"]


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_peer_exchange_is_bitwise_reproducible_and_rank_ordered(cg, gpus):
    # the mailbox sums add the partials in rank order on every rank: two runs agree bit for bit,
    # and so do the iteration counts of all ranks (else the solve would dead-lock or diverge)
    need(cg, gpus)
    n, seed = 2048, 77
    x1, st1 = cg.solve_generated(n, seed, gpus=gpus, comm=2, ignore_stop=1, max_iter=9)
    x2, st2 = cg.solve_generated(n, seed, gpus=gpus, comm=2, ignore_stop=1, max_iter=9)
    assert st1["iterations"] == st2["iterations"] == 9
    np.testing.assert_array_equal(x1, x2)
    assert st1["final_beta"] == st2["final_beta"]


@pytest.mark.parametrize("comm", COMMS)
@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_generated_sharded_solve(cg, gpus, comm):
    need(cg, gpus)
    n, seed = 4096, 31
    x, st = cg.solve_generated(n, seed, gpus=gpus, comm=comm)
    xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n), ranks=gpus)
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    x1, st1 = cg.solve_generated(n, seed, gpus=1)
    assert relerr(x, x1) <= RTOL_X
    # fixed-iteration mode across ranks: every rank stops at the same count
    x, st = cg.solve_generated(n, seed, gpus=gpus, ignore_stop=1, max_iter=7, comm=comm)
    assert st["iterations"] == 7 and st["comm"] == comm


@pytest.mark.parametrize("comm", COMMS)
@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_generated_spread_family_sharded(cg, gpus, comm):
    # ~100 iterations on the on-device generator's cond ~ 2^7 family, every GPU its own rows
    need(cg, gpus)
    n, seed, flags = 2048, 23, cg.GEN_KAPPA_LOG2(7)
    x, st = cg.solve_generated(n, seed, gpus=gpus, flags=flags, comm=comm)
    xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n), ranks=gpus)
    assert so["iterations"] > 80
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    assert st["final_residual"] < EPS and st["comm"] == comm


@pytest.mark.parametrize("gpus,n,k", [(2, 2048, 7), (2, 1024, 0), (4, 4096, 4), (4, 2048, 7), (8, 4096, 4), (8, 8192, 0)])
def test_symmetric_product_row_sharded(cg, gpus, n, k):
    # every rank reads half of its row block; column sums travel to their owners over NVLink
    need(cg, gpus)
    seed, flags = 50 + gpus, cg.GEN_KAPPA_LOG2(k)
    x, st = cg.solve_generated(n, seed, gpus=gpus, flags=flags, symmetric=1)
    xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n), ranks=gpus)
    assert st["gemv_variant"] == 3 and st["comm"] == 2
    assert relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    assert st["final_residual"] < EPS
    x2, st2 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, symmetric=1)
    np.testing.assert_array_equal(x, x2)                       # fixed-order sums on every rank


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_symmetric_option_on_ranks_verifies_loaded_matrices_across_the_ranks(cg, gpus):
    """symmetric = 1 on a matrix LOADED on several ranks: no rank holds both a_ij and a_ji of an off-diagonal block,
    so the ranks compare 64-bit fingerprints of their blocks (cg_symv.cu: cgk_symmetry_hash) and their diagonal blocks
    exactly.  A symmetric matrix is accepted and solved with the half-bytes product; one flipped bit anywhere -- inside
    a diagonal block, or in a block only another rank can mirror -- is refused on every rank, as is a NaN."""
    need(cg, gpus)
    n = 512 * gpus
    A, b = O.spd_with_condition(n, 30.0, seed=8)
    x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus, symmetric=1)
    xo, so = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
    assert st["gemv_variant"] == 3 and relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    m = n // gpus
    for (i, j) in ((3, 40), (5, m + 7), (n - 1, 0)):            # own diagonal block; a block pair of two ranks; the far corner
        B = A.copy()
        B[i, j] = np.nextafter(B[i, j], np.inf)
        with pytest.raises(cg.CgbError) as e:
            cg.conjugrad_host(B, b, np.zeros(n), gpus=gpus, symmetric=1)
        assert "not exactly symmetric" in str(e.value), str(e.value)
    B = A.copy()
    B[7, m + 1] = B[m + 1, 7] = np.nan
    with pytest.raises(cg.CgbError):
        cg.conjugrad_host(B, b, np.zeros(n), gpus=gpus, symmetric=1)
    x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus, symmetric=2)     # the caller's assertion still skips the check
    assert st["gemv_variant"] == 3 and relerr(x, xo) <= RTOL_X
    with pytest.raises(cg.CgbError):                            # NCCL transport: not available for this product
        cg.solve_generated(n, 1, gpus=gpus, symmetric=1, comm=1)


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_asserted_symmetric_upload_on_ranks_reads_only_each_ranks_tiles(cg, gpus):
    # symmetric = 2 through the reference-facing call: a rank uploads the tiles of its tile list only;
    # everything no rank reads is poisoned with NaN in the host matrix and must not matter
    need(cg, gpus)
    from test_symplan import plan_and_tiles, T
    mt = 2
    n = gpus * mt * T
    A, b = O.spd_with_condition(n, 30.0, seed=20 + gpus)
    xo, so = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
    read = np.zeros((n // T, n // T), dtype=bool)
    per_rank = 0
    for r in range(gpus):
        tiles = plan_and_tiles(r, gpus, mt)[1]
        per_rank = max(per_rank, len(tiles))
        for il, j in tiles:
            read[r * mt + il, j] = True
    P = A.copy()
    for I in range(n // T):
        for J in range(n // T):
            if not read[I, J]:
                P[I * T:(I + 1) * T, J * T:(J + 1) * T] = np.nan
    assert read.sum() == (n // T) * (n // T + 1) // 2
    x, st = cg.conjugrad_host(P, b, np.zeros(n), gpus=gpus, symmetric=2)
    assert st["gemv_variant"] == 3 and relerr(x, xo) <= RTOL_X and abs(st["iterations"] - so["iterations"]) <= 1
    assert st["h2d_bytes"] <= 8 * (per_rank * T * T + n + n // gpus)


@pytest.mark.parametrize("gpus,n,k,sym", [(2, 2048, 7, 0), (2, 2048, 4, 1), (4, 2048, 7, 0), (4, 4096, 0, 1), (8, 4096, 4, 0), (8, 4096, 4, 1)])
def test_fused_reductions_variant_sharded(cg, gpus, n, k, sym):
    # Chronopoulos-Gear recurrence: r is the exchanged vector, one exposed reduction per iteration
    need(cg, gpus)
    seed, flags = 90 + gpus, cg.GEN_KAPPA_LOG2(k)
    x, st = cg.solve_generated(n, seed, gpus=gpus, flags=flags, cg_variant=1, symmetric=sym)
    A, b = O.gen_A(n, seed, flags), O.gen_b(n, seed, flags)
    xc, sc = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
    assert st["cg_variant"] == 1 and st["comm"] == 2
    assert relerr(x, xc) <= RTOL_X and abs(st["iterations"] - sc["iterations"]) <= 1 and st["final_residual"] < EPS
    x2, st2 = cg.solve_generated(n, seed, gpus=gpus, flags=flags, cg_variant=1, symmetric=sym)
    np.testing.assert_array_equal(x, x2)
    with pytest.raises(cg.CgbError):
        cg.solve_generated(n, seed, gpus=gpus, cg_variant=1, comm=1)


def test_not_divisible(cg):
    need(cg, 2)
    with pytest.raises(cg.CgbError) as e:
        cg.conjugrad_host(np.eye(5), np.ones(5), np.zeros(5), gpus=2)
    assert e.value.code == -4 and "5 is not divisible by 2" in str(e.value)    # parallel_cg.c:88


def test_cli_two_gpus(cg, golden_dir):
    need(cg, 2)
    from conjugate_gradient_b200 import _binding as B
    g = golden_dir
    r = subprocess.run([B.CLI_PATH, f"{g}/g2_matrixA.txt", f"{g}/g2_vectorb.txt", f"{g}/g2_x0.txt",
                        "--dims", f"{g}/g2_dimensions.txt", "--gpus", "2", "--print-x"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "Computing cg of matrix size : 16"
    xs = [float(l) for l in lines[1:5]]
    np.testing.assert_allclose(xs, [-1, 1, -1, 1], atol=1e-6)
    assert lines[5].startswith("cg method execution time in seconds: ")
    assert lines[6].startswith("collective data distribution time in seconds: ")
    assert lines[7].startswith("clock execution time in seconds: ")


def test_torchrun_bench_two_ranks(cg):
    need(cg, 2)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(root, "bench.py"),
                        "--gpus", "2", "--steps", "4", "--warmup", "3", "--matrix-n", "8192"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["value"] > 0 and line["e2e"]["value"] > 0
    # one process per GPU: the windows are mapped through CUDA IPC
    assert "peer" in line["implementation"]["exchange"]
    assert line["parity"]["ok"] and line["parity"]["ranks"] == 2 and line["parity"]["fixed_run"]["rel_err_x"] <= 1e-10


@pytest.mark.parametrize("step", ["ALLOC", "LOAD", "PREPARE"])
@pytest.mark.parametrize("gpus", [2, 4])
def test_a_rank_failing_alone_takes_the_team_with_it_instead_of_hanging(cg, gpus, step, monkeypatch):
    """a rank that fails a local step (no device memory for its block, a failed copy, no workspace)
    used to leave its peers inside the next NCCL call for ever.  The ranks now agree on a status before every
    collective step: with the failure injected on the last rank every rank returns the error, promptly, and the
    next call -- the team is rebuilt -- solves normally."""
    need(cg, gpus)
    import time
    n, seed = 2048, 3
    A, b = O.gen_A(n, seed), O.gen_b(n, seed)
    monkeypatch.setenv(f"CGB_DEBUG_FAIL_{step}", str(gpus - 1))
    t0 = time.time()
    with pytest.raises(cg.CgbError) as e:
        cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus)
    assert time.time() - t0 < 60.0
    assert e.value.code == -5 and "injected failure" in str(e.value), str(e.value)
    monkeypatch.delenv(f"CGB_DEBUG_FAIL_{step}")
    x, st = cg.conjugrad_host(A, b, np.zeros(n), gpus=gpus)
    xo, so = O.conjugrad(A, b, np.zeros(n), ranks=gpus)
    assert relerr(x, xo) <= RTOL_X and st["ranks"] == gpus


@pytest.mark.parametrize("gpus", [2, 4, 8])
@pytest.mark.parametrize("family", ["spectrum", "gram"])
def test_further_generator_families_row_sharded(cg, gpus, family):
    """Every GPU generates its own rows of the prescribed-spectrum / R R' matrix; the sharded solve (full and
    symmetric product: the blocks of different GPUs must agree bit for bit across the diagonal) against the oracle."""
    need(cg, gpus)
    n, seed = 4096, 13
    flags = (cg.GEN_SPECTRUM | cg.GEN_KAPPA_LOG2(8)) if family == "spectrum" else (cg.GEN_GRAM | cg.GEN_KAPPA_LOG2(5))
    x1, st1 = cg.solve_generated(n, seed, gpus=1, flags=flags)
    x, st = cg.solve_generated(n, seed, gpus=gpus, flags=flags)
    assert abs(st["iterations"] - st1["iterations"]) <= 1 and st["final_residual"] < EPS and st["iterations"] > 15
    if st["iterations"] == st1["iterations"]:
        assert relerr(x, x1) <= RTOL_X
    xs, sts = cg.solve_generated(n, seed, gpus=gpus, flags=flags, symmetric=1)
    assert sts["gemv_variant"] == 3 and abs(sts["iterations"] - st1["iterations"]) <= 1
    if sts["iterations"] == st1["iterations"]:
        assert relerr(xs, x1) <= RTOL_X
    if family == "spectrum":
        xo, so = O.conjugrad(O.gen_A(n, seed, flags), O.gen_b(n, seed, flags), np.zeros(n), ranks=gpus)
        assert abs(st["iterations"] - so["iterations"]) <= 1
        if st["iterations"] == so["iterations"]:
            assert relerr(x, xo) <= RTOL_X
