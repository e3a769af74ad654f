#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of libibp_b200.so (cuobjdump -sass): the instructions that show which hardware paths a
kernel uses — UBLKCP (1-D bulk copy, TMA engine), UTMALDG (tensor-map TMA), SYNCS (mbarrier), FFMA2/FMUL2 (packed fp32),
HFMA2/F2FP (half), BAR, LDS/STS, LDG, ATOM/RED.  `python tools/sass_counts.py > profiles/r2_sass_counts.txt`."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "improved-body-parts_b200", "libibp_b200.so")
WATCH = ["UBLKCP", "UTMALDG", "SYNCS", "FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL", "FADD", "DADD", "DMUL", "DFMA", "HADD2", "F2FP", "BAR",
         "LDS", "STS", "LDG", "STG", "LDGSTS", "ATOMG", "ATOMS", "RED", "REDUX", "SHFL", "VOTE", "MUFU", "ACQBULK", "NANOSLEEP"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kern, counts, total, arch = None, collections.OrderedDict(), {}, set()
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            kern = re.sub(r"\(.*", "", kern).replace("ibp::", "")
            counts[kern] = collections.Counter()
            total[kern] = 0
            continue
        m = re.search(r"arch = (sm_\w+)", line)
        if m:
            arch.add(m.group(1))
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and kern:
            counts[kern][m.group(1)] += 1
            total[kern] += 1
    print("libibp_b200.so — cuobjdump -sass, arch %s; static instruction counts per kernel" % ", ".join(sorted(arch)))
    used = [w for w in WATCH if any(c[w] for c in counts.values())]
    print("%-52s %7s " % ("kernel", "total") + " ".join("%7s" % w for w in used))
    for k, c in counts.items():
        print("%-52s %7d " % (k[:52], total[k]) + " ".join("%7d" % c[w] for w in used))
    print("\nNo FFMA contraction is allowed on the scored expressions (-fmad=false + explicit _rn intrinsics): FFMA that appear are the\n"
          "compiler's own division / reciprocal / sqrt sequences and the packed add written as fma(x, 1, y) (k_multiscale.cuh: add2).")


if __name__ == "__main__":
    sys.exit(main())
