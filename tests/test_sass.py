"""CPU-only, static evidence about the machine code in libcgb200.so (cuobjdump on the built
library; nothing is executed): the library carries sm_100a code only, the headline GEMV streams
A with the 256-bit evict-first loads DESIGN.md §4.1 names, the bulk variant really uses the TMA
engine (UBLKCP + mbarrier SYNCS), the small-n kernel keeps rows with evict-last, the iteration's
kernels carry the programmatic-dependent-launch instructions, the dot products use no FP64
atomics (bitwise reproducible sums), nothing touches a tensor core (a single-vector GEMV is not a
contraction) and the hot kernels do not spill.  Mnemonics as listed in B200_PROFILING.md."""
import re
import shutil
import subprocess

import pytest

import conjugate_gradient_b200 as cg

CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
pytestmark = pytest.mark.skipif(shutil.which(CUOBJDUMP) is None, reason="cuobjdump not installed")


@pytest.fixture(scope="module")
def sass():
    """{demangled kernel name without the argument list: SASS text of the function}"""
    txt = subprocess.run([CUOBJDUMP, "-sass", cg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"^\s*Function : (\S+)\s*$", txt, flags=re.M)
    mangled = parts[1::2]
    names = subprocess.run(["c++filt"], input="\n".join(mangled), capture_output=True, text=True, check=True).stdout.split("\n")
    out = {}
    for nm, body in zip(names, parts[2::2]):
        nm = re.sub(r"^void ", "", nm)
        nm = nm[:nm.index("(")] if "(" in nm else nm
        out[nm.replace(" ", "")] = body
    return out


def ops(body):
    """opcode (with modifiers) of every instruction line"""
    return re.findall(r"^\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body, flags=re.M)


def test_only_sm_100a_code():
    out = subprocess.run([CUOBJDUMP, "-lelf", cg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    elfs = re.findall(r"ELF file\s+\d+: (\S+)", out)
    assert elfs and all(e.endswith(".sm_100a.cubin") for e in elfs), elfs
    ptx = subprocess.run([CUOBJDUMP, "-lptx", cg.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_9" not in ptx and "sm_8" not in ptx


def test_every_kernel_family_is_present(sass):
    fam = {re.sub(r"<.*", "", k) for k in sass}
    want = {"cgb_rowblock_matvec_ldg", "cgb_rowblock_matvec_keep", "cgb_rowblock_matvec_comm", "cgb_rowblock_matvec_bulk",
            "cgb_xr_update_kernel", "cgb_p_update_kernel", "cgb_xrp_update_kernel", "cgb_init_residual_kernel", "cgb_after_init_kernel", "cgb_reset_kernel",
            "cgb_cg2_step_kernel", "cgb_cg2_step_one_kernel", "cgb_generate_A_kernel", "cgb_generate_b_kernel",
            "cgb_halfmat_tile_kernel", "cgb_halfmat_tile_comm_kernel", "cgb_halfmat_finish_kernel", "cgb_halfmat_finish_comm_kernel",
            "cgb_halfmat_push_kernel", "cgb_symmetry_check_kernel", "cgb_ingest_scan_kernel", "cgb_ingest_count_kernel",
            "cgb_ingest_parse_kernel", "cgb_place_values_kernel", "cgb_l2_flush_kernel"}
    assert want <= fam, want - fam


def test_headline_gemv_streams_with_256_bit_evict_first_loads(sass):
    body = sass["cgb_rowblock_matvec_ldg<8,2,512,4>"]          # the n = 65536 default (profiles/r01_summary.md §2)
    o = ops(body)
    wide = [x for x in o if x.startswith("LDG.") and ".256" in x]
    # R * U = 16 independent 32-byte loads of A in flight per thread, all no-allocate + evict-first
    assert sum(x == "LDG.E.NA.EFL2.256.CONSTANT" for x in wide) >= 16, wide
    assert not any(x.startswith(("LDL", "STL")) for x in o), "the headline GEMV spills"
    assert sum(x.startswith("DFMA") for x in o) >= 64
    assert any(x.startswith("SHFL.BFLY") for x in o)     # warp-shuffle row reduction


def test_default_gemv_shapes_do_not_spill(sass):
    # the shapes cgk_gemv_make_plan picks by itself (tests/test_plans.py); the register-capped keep
    # kernel and a few selectable 512-thread shapes carry 8-24 bytes of stack (profiles/r01_kernel_resources.txt)
    for name in ("cgb_rowblock_matvec_ldg<8,2,512,4>", "cgb_rowblock_matvec_ldg<2,1,256,4>", "cgb_rowblock_matvec_ldg<4,1,128,4>",
                 "cgb_rowblock_matvec_ldg<4,1,128,2>", "cgb_rowblock_matvec_ldg<4,1,128,1>", "cgb_rowblock_matvec_keep<2,1,256,4>"):
        assert not any(x.startswith(("LDL", "STL")) for x in ops(sass[name])), name


def test_every_vw4_gemv_uses_256_bit_loads(sass):
    seen = 0
    for name, body in sass.items():
        m = re.match(r"cgb_rowblock_matvec(_ldg|_keep|_comm)<(\d+),(\d+),(\d+),(\d+)>", name)
        if not m:
            continue
        seen += 1
        o = ops(body)
        if m.group(5) == "4":
            assert any(x.startswith("LDG.") and ".256" in x for x in o), name
        if m.group(1) == "_keep":       # rows meant to survive in L2: evict-last beside evict-first
            assert any(".ELL2." in x for x in o) and any(".EFL2." in x for x in o), name
    assert seen >= 20, seen


def test_bulk_variant_uses_the_tma_engine(sass):
    bulk = {k: v for k, v in sass.items() if k.startswith("cgb_rowblock_matvec_bulk<")}
    assert len(bulk) >= 8
    for name, body in bulk.items():
        o = ops(body)
        assert any(x.startswith("UBLKCP") for x in o), name                      # cp.async.bulk
        assert any(x.startswith("SYNCS.ARRIVE.TRANS64") for x in o), name        # mbarrier expect_tx
        assert any(x.startswith("SYNCS.PHASECHK.TRANS64.TRYWAIT") for x in o), name
        assert any(x.startswith("LDS.128") or x.startswith("LDS.U.128") or x == "LDS.128" for x in o), name


def test_iteration_kernels_carry_programmatic_dependent_launch(sass):
    # everything cg_kernels.cu / cg_symv.cu launch through cgk_launch (the attribute is only legal on
    # kernels that order themselves with griddepcontrol.wait)
    for name in ("cgb_xr_update_kernel<false>", "cgb_p_update_kernel<false>", "cgb_cg2_step_one_kernel", "cgb_cg2_step_kernel<true>",
                 "cgb_rowblock_matvec_keep<4,1,128,4>", "cgb_rowblock_matvec_comm<4,2,256,4>", "cgb_halfmat_tile_kernel<4>",
                 "cgb_halfmat_finish_kernel", "cgb_halfmat_push_kernel"):
        o = ops(sass[name])
        assert "ACQBULK" in o, name        # griddepcontrol.wait
        assert "PREEXIT" in o, name        # griddepcontrol.launch_dependents
        assert o.index("ACQBULK") < o.index("PREEXIT"), name     # wait first: completion stays transitive (cg_device.cuh)
    # the bandwidth-bound GEMV of large systems is launched the ordinary way and carries neither
    o = ops(sass["cgb_rowblock_matvec_ldg<8,2,512,4>"])
    assert "ACQBULK" not in o and "PREEXIT" not in o


def test_no_fp64_atomics_and_tensor_cores_only_where_there_is_a_contraction(sass):
    """The CG iteration is HBM-bound vector work: no tensor instruction anywhere on it.  The one dense contraction of
    the package -- the R R' generator family (generateSPDmatrix.m:12), an FP64 SYRK -- runs on the FP64 tensor pipe."""
    for name, body in sass.items():
        for x in ops(body):
            assert not re.match(r"(ATOM|ATOMG|ATOMS|RED)\..*F64", x), (name, x)
            if not name.startswith("cgb_gram_rows_kernel"):
                assert not re.match(r"(HMMA|DMMA|IMMA|QMMA|UTCHMMA|UTCQMMA|UTCIMMA|HGMMA)", x), (name, x)
    gram = ops(sass["cgb_gram_rows_kernel"])
    assert sum(x.startswith("DMMA") for x in gram) >= 32 and not any(x.startswith(("LDL", "STL")) for x in gram)


def test_peer_exchange_uses_system_scope_accesses(sass):
    # the kernels that push into peer windows fence at system scope, the consumers poll with sys loads
    assert any(x == "MEMBAR.SC.SYS" for x in ops(sass["cgb_p_update_kernel<true>"]))
    assert any(x == "MEMBAR.SC.SYS" for x in ops(sass["cgb_xrp_update_kernel"]))
    for name in ("cgb_xr_update_kernel<true>", "cgb_p_update_kernel<true>", "cgb_xrp_update_kernel", "cgb_rowblock_matvec_comm<4,2,256,4>"):
        assert any(x.endswith("STRONG.SYS") and x.startswith("LDG") for x in ops(sass[name])), name
    # the exchange-fused product keeps all R * U = 8 wide loads of A of a tile in flight
    o = ops(sass["cgb_rowblock_matvec_comm<4,2,256,4>"])
    assert sum(x == "LDG.E.NA.EFL2.256.CONSTANT" for x in o) >= 8
    # ... and the one-rank forms pay no system-scope fence
    for name in ("cgb_xr_update_kernel<false>", "cgb_p_update_kernel<false>", "cgb_rowblock_matvec_ldg<8,2,512,4>", "cgb_cg2_step_one_kernel"):
        assert "MEMBAR.SC.SYS" not in ops(sass[name]), name
