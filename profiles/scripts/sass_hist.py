"""Opcode histogram of the decode-iteration function of libfec_b200.so (cuobjdump -sass), per pipe.
Usage: python profiles/scripts/sass_hist.py [path/to/libfec_b200.so] > profiles/r02_sass_row_hist.txt
The function holds one unrolled copy of the row code per (degree, state location); the totals divided by the edges those copies
cover (sum of the instantiated degrees) give instructions per edge and thread (two checks)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "srsran-5g_b200/libfec_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
FMA = {"HFMA2", "HADD2", "HMUL2", "IMAD", "FFMA", "FADD", "FMUL"}
ALU = {"LOP3", "SHF", "PRMT", "HMNMX2", "HSET2", "HSETP2", "ISETP", "SEL", "IADD3", "IADD", "LEA", "VIADD", "VIMNMX", "VIADDMNMX",
       "MOV", "IABS", "POPC", "FLO", "BREV", "SGXT", "BMSK", "IMNMX", "FMNMX", "FSETP", "FSEL", "LOP", "SHL", "SHR", "PLOP3",
       "P2R", "R2P", "VHMNMX", "VIMNMX3", "CS2R", "HMNMX2.XORSIGN"}
LSU = {"LDS", "STS", "LDG", "STG", "LD", "ST", "LDSM", "ATOMS", "ATOMG", "RED", "LDL", "STL", "LDC"}
TM = {"LDTM", "STTM"}
kern = [f for f in re.split(r"\n\s*Function : ", sass)[1:] if f.startswith("_Z20fec_cb_packed_kernel")][0]
lines = []
for line in kern.split("\n"):
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)(.*?);", line)
    if m:
        lines.append((int(m.group(1), 16), m.group(3), m.group(4)))
# The iteration loop is inlined into the kernel: its code is the span of the row loop -- from the target of the backward branch
# that closes the loop behind the last HSET2 (only the row bodies use that opcode: "edge e was the min1 edge") to that branch.
idx = [i for i, (_, op, _) in enumerate(lines) if op.startswith("HSET2")]
first, last = idx[0], idx[-1]


def bra_target(i):
    t = re.findall(r"0x([0-9a-f]+)", lines[i][2])
    return int(t[-1], 16) if t else None


hi_i = next(i for i in range(last, len(lines)) if lines[i][1].startswith("BRA") and bra_target(i) is not None and
            bra_target(i) <= lines[first][0])
lo_i = next(i for i in range(len(lines)) if lines[i][0] >= bra_target(hi_i))
lo, hi = lines[lo_i][0], lines[hi_i][0] + 16
body = [(a_, op) for a_, op, _ in lines[lo_i:hi_i + 1]]
ops = collections.Counter(op for _, op in body)
by = collections.Counter()
for op, n in ops.items():
    base = op.split(".")[0]
    pipe = "fma" if base in FMA else "alu" if base in ALU else "lsu" if base in LSU else "tmem" if base in TM else \
        "uniform" if base.startswith("U") or base in ("R2UR", "S2UR") else "control" if base in (
            "BAR", "BRA", "BRX", "BSSY", "BSYNC", "EXIT", "RET", "CALL", "NOP", "WARPSYNC", "DEPBAR", "BPT", "JMP", "ERRBAR") else "other"
    by[pipe] += n
total = sum(ops.values())
edges = int(sys.argv[2]) if len(sys.argv) > 2 else 0
print("fec_packed_iteration: SASS range 0x%x-0x%x of fec_cb_packed_kernel, %d instructions" % (lo, hi, total))
if edges:
    print("one unrolled copy of the row per instantiation; the instantiated degrees sum to %d edges -> %.1f instructions per edge "
          "and thread (two checks)" % (edges, total / edges))
for pipe, n in by.most_common():
    print("  pipe %-8s %6d  %5.1f %%%s" % (pipe, n, 100.0 * n / total, ("  %.1f per edge" % (n / edges)) if edges else ""))
print("opcodes:")
for op, n in ops.most_common():
    print("  %-28s %6d%s" % (op, n, ("  %.2f" % (n / edges)) if edges else ""))
