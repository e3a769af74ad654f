#!/usr/bin/env python
"""Dynamic SASS opcode histogram and the heaviest instructions of one kernel launch from an .ncu-rep (--import-source on).

    python tools/ncu_sass.py rep.ncu-rep kernel_regex [launch_index] [units]      units = divide counts by this (e.g. CTAs)
"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    units = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv", "--kernel-name", "regex:" + kern],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, launch = None, -1
    per, wf, samples = collections.Counter(), collections.Counter(), collections.Counter()
    tot = tot_wf = tot_s = 0
    heavy = []
    for r in rows:
        if r and r[0] == "Kernel Name":
            launch += 1
            continue
        if r and r[0] == "Address":
            hdr = r
            continue
        if hdr is None or launch != which or len(r) != len(hdr):
            continue
        parts = r[1].split()
        op = (parts[1] if parts[0].startswith("@") else parts[0]).split(".")[0]
        n = int(r[hdr.index("Instructions Executed")])
        w = int(r[hdr.index("L1 Wavefronts Shared")] or 0)
        sm = int(r[hdr.index("# Samples")] or 0)
        per[op] += n
        wf[op] += w
        samples[op] += sm
        tot += n
        tot_wf += w
        tot_s += sm
        heavy.append((sm, n, w, r[1].strip()))
    print("launch %d: %.0f warp-instructions, %.0f shared wavefronts, %d samples (all / %g)" % (which, tot / units, tot_wf / units, tot_s, units))
    for op, n in per.most_common(32):
        print("%-10s %6.2f%%i %10.1f   wavefronts %9.1f   samples %5.1f%%" % (op, 100.0 * n / tot, n / units, wf[op] / units, 100.0 * samples[op] / max(tot_s, 1)))
    print("heaviest by samples:")
    for sm, n, w, src in sorted(heavy, reverse=True)[:40]:
        print("%5.1f%%s %10.1f inst %9.1f wf  %s" % (100.0 * sm / max(tot_s, 1), n / units, w / units, src[:90]))


if __name__ == "__main__":
    main()
