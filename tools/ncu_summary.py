#!/usr/bin/env python
"""Transposed metric table (metric rows x kernel columns) of the first launch of each named kernel in one or more
.ncu-rep files, and the dram traffic entries for profiles/r1_traffic.json.

    python tools/ncu_summary.py out.csv traffic.json rep1.ncu-rep[:regex[,regex...]] [rep2.ncu-rep:regex ...]
"""
import csv
import io
import json
import re
import subprocess
import sys

SHORT = {"peaks_stream_kernel": "peaks", "index_bands_kernel": "index", "limbs_stream_kernel": "limbs", "fuse_multiscale_kernel": "fuse",
         "py_connections_kernel": "py_connections", "py_humans_kernel": "py_humans", "limbs_kernel": "limbs_oneshot",
         "peaks_kernel": "peaks_oneshot"}


def raw(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    return rows[0], rows[1], rows[2:]


def main():
    out_csv, out_json = sys.argv[1], sys.argv[2]
    cols = []  # (kernel name, {metric: (unit, value)})
    for spec in sys.argv[3:]:
        rep, _, pats = spec.partition(":")
        pats = [re.compile(p) for p in pats.split(",") if p] or [re.compile(".")]
        names, units, rows = raw(rep)
        ki = names.index("Kernel Name")
        seen = set()
        for r in rows:
            k = r[ki]
            if k in seen or not any(p.search(k) for p in pats):
                continue
            seen.add(k)
            cols.append((k, {n: (u, v) for n, u, v in zip(names, units, r)}))
    metrics = []
    for _, m in cols:
        for n in m:
            if n not in metrics and (n in ("Kernel Name", "Block Size", "Grid Size") or "__" in n):
                metrics.append(n)
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [k for k, _ in cols])
        for n in metrics:
            unit = next((m[n][0] for _, m in cols if n in m), "")
            w.writerow([n, unit] + [m.get(n, ("", ""))[1] for _, m in cols])
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tscale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    kernels = {}
    for k, m in cols:
        short = next((v for key, v in SHORT.items() if key in k), k)
        def val(n, table):
            u, v = m[n]
            return float(v.replace(",", "")) * table[u]
        kernels[short] = {"kernel": k, "dram_bytes_read": val("dram__bytes_read.sum", scale), "dram_bytes_write": val("dram__bytes_write.sum", scale),
                          "duration_us_under_ncu": val("gpu__time_duration.sum", tscale)}
    try:
        cur = json.load(open(out_json))
    except Exception:
        cur = {"kernels": {}}
    cur["source"] = "ncu --set full --clock-control none, bench.py --steps 1-5 --warmup 1-3, one launch per kernel; see r1_ncu_full_summary.csv"
    cur.pop("workload", None)
    cur.setdefault("kernels", {}).update(kernels)
    json.dump(cur, open(out_json, "w"), indent=1)
    print(json.dumps(kernels, indent=1))


if __name__ == "__main__":
    main()
