#!/usr/bin/env python
"""Static evidence from the built library (cuobjdump -sass, no GPU needed): per kernel, the memory / sync
instructions that DESIGN.md names.  Writes a table to stdout (committed as profiles/r01_sass_evidence.txt);
tests/test_sass.py asserts the same facts."""
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import conjugate_gradient_b200 as cg

txt = subprocess.run(["cuobjdump", "-sass", cg.LIB_PATH], capture_output=True, text=True, check=True).stdout
parts = re.split(r"^\s*Function : (\S+)\s*$", txt, flags=re.M)
names = subprocess.run(["c++filt"], input="\n".join(parts[1::2]), capture_output=True, text=True, check=True).stdout.split("\n")
COLS = [("LDG.256 evict-first", r"LDG\.E(\.NA)?\.EFL2\.256"), ("LDG.256 evict-last", r"LDG\.E(\.NA)?\.ELL2\.256"),
        ("LDG.256 other", r"LDG\.E(\.NA)?(\.ENL2)?\.256"), ("LDG.128", r"LDG\.E(\.NA)?\.128"), ("UBLKCP (TMA)", r"UBLKCP"),
        ("SYNCS (mbarrier)", r"SYNCS\."), ("DFMA", r"DFMA"), ("SHFL", r"SHFL\."), ("ACQBULK/PREEXIT (PDL)", r"ACQBULK|PREEXIT"),
        ("MEMBAR.SYS", r"MEMBAR\.SC\.SYS"), ("LD/ST .SYS", r"(LDG|STG)\.E(\.64)?\.STRONG\.SYS"), ("ATOMG int", r"ATOMG\.E\.(ADD|EXCH)\.STRONG"),
        ("FP64 atomics", r"(ATOM|ATOMG|RED)\.[A-Z.]*F64"), ("tensor-core ops", r"\b(HMMA|DMMA|IMMA|QMMA|UTC[A-Z]*MMA)"),
        ("LDL/STL (spill)", r"\b(LDL|STL)")]
print("# cuobjdump -sass of conjugate_gradient_b200/lib/libcgb200.so (sm_100a only: " +
      ", ".join(re.findall(r"ELF file\s+\d+: (\S+)", subprocess.run(["cuobjdump", "-lelf", cg.LIB_PATH], capture_output=True, text=True).stdout)) + ")")
print("# instruction counts per kernel; made by tools/sass_evidence.py, asserted by tests/test_sass.py")
print("kernel | " + " | ".join(c for c, _ in COLS))
rows = []
for nm, body in zip(names, parts[2::2]):
    nm = re.sub(r"^void ", "", nm)
    nm = nm[:nm.index("(")] if "(" in nm else nm
    ops = re.findall(r"^\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body, flags=re.M)
    rows.append((nm, [sum(1 for o in ops if re.match(rx, o)) for _, rx in COLS]))
keep = lambda nm: not nm.startswith("cgb_rowblock_matvec_ldg<") or nm.replace(" ", "") in (
    "cgb_rowblock_matvec_ldg<8,2,512,4>", "cgb_rowblock_matvec_ldg<2,1,256,4>", "cgb_rowblock_matvec_ldg<4,1,128,4>", "cgb_rowblock_matvec_ldg<4,1,128,2>", "cgb_rowblock_matvec_ldg<4,1,128,1>")
for nm, counts in sorted(rows):
    if keep(nm):
        print(nm + " | " + " | ".join(str(c) for c in counts))
print(f"# ({sum(1 for nm, _ in rows if not keep(nm))} further selectable cgb_rowblock_matvec_ldg<R,U,NT,VW> shapes omitted; the five listed are the ones the plan picks by itself)")
