#!/usr/bin/env python
"""Per-source-line warp-stall samples of one kernel from an .ncu-rep (needs -lineinfo and --import-source on).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep limbs_kernel [top_n] [launch_index]
"""
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                          "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    cur_file, hdr, launch = None, None, -1
    agg = {}
    stall_cols = None
    seen_files_in_launch = set()
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            if cur_file in seen_files_in_launch or launch < 0:
                launch += 1
                seen_files_in_launch = set()
            seen_files_in_launch.add(cur_file)
            continue
        if len(r) > 6 and r[0] == "Line No":
            hdr = r
            stall_cols = [(i, c) for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
            continue
        if hdr is None or len(r) != len(hdr) or launch != which:
            continue
        if r[0] == "" or not r[0].isdigit():
            continue
        try:
            s = int(r[hdr.index("# Samples")])
            inst = int(r[hdr.index("Instructions Executed")])
        except ValueError:
            continue
        key = (cur_file, int(r[0]))
        e = agg.setdefault(key, [0, 0, r[1].strip()[:100], {}])
        e[0] += s
        e[1] += inst
        for i, c in stall_cols:
            try:
                v = int(r[i])
            except ValueError:
                v = 0
            if v:
                e[3][c] = e[3].get(c, 0) + v
    tot = sum(e[0] for e in agg.values()) or 1
    toti = sum(e[1] for e in agg.values()) or 1
    print("kernel %s launch %d: %d samples, %d warp-instructions" % (kern, which, tot, toti))
    for (f, ln), e in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        st = sorted(e[3].items(), key=lambda kv: -kv[1])[:3]
        print("%5.1f%% %5.1f%%i %-16s:%-4d %-90s %s" % (100.0 * e[0] / tot, 100.0 * e[1] / toti, f, ln, e[2],
                                                      " ".join("%s=%d" % (k[6:], v) for k, v in st)))


if __name__ == "__main__":
    main()
