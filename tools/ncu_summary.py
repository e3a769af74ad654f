#!/usr/bin/env python
"""Transposed metric table (metric rows x kernel columns) of the first launch of each named kernel in one or more
.ncu-rep files, and the dram traffic entries bench.py reads (profiles/rN_traffic.json: {"<workload>@<batch>": {stage: ...}}).

    python tools/ncu_summary.py out.csv traffic.json rep1.ncu-rep[:workload@batch[:regex,regex...]] [rep2.ncu-rep:...]
"""
import csv
import io
import json
import re
import subprocess
import sys

# bench.py's stage names; checked in this order (first match wins)
SHORT = [("peaks_stream_kernel", "peaks"), ("index_bands_kernel", "index"), ("limbs_stream3_kernel", "limbs"), ("limbs_stream_kernel", "limbs"), ("fuse_multiscale_kernel", "fuse"),
         ("fuse_strip_kernel", "fuse_strip"), ("fuse_linear_kernel", "fuse_linear"), ("gauss_smooth_kernel", "smooth"),
         ("py_connections_kernel", "py_connections"), ("py_humans_kernel", "py_humans"), ("limbs_kernel", "limbs"), ("peaks_kernel", "peaks")]


def raw(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    return rows[0], rows[1], rows[2:]


def main():
    out_csv, out_json = sys.argv[1], sys.argv[2]
    cols = []  # (workload key, kernel name, {metric: (unit, value)})
    for spec in sys.argv[3:]:
        parts = spec.split(":")
        rep, wkey = parts[0], (parts[1] if len(parts) > 1 else "")
        pats = [re.compile(p) for p in (parts[2].split(",") if len(parts) > 2 else []) if p] or [re.compile(r"ibp|peaks|limbs|fuse|gauss|py_|index_bands")]
        names, units, rows = raw(rep)
        ki = names.index("Kernel Name")
        seen = set()
        for r in rows:
            k = r[ki]
            if k in seen or not any(p.search(k) for p in pats):
                continue
            seen.add(k)
            cols.append((wkey, k, {n: (u, v) for n, u, v in zip(names, units, r)}))
    metrics = []
    for _, _, m in cols:
        for n in m:
            if n not in metrics and (n in ("Kernel Name", "Block Size", "Grid Size") or "__" in n):
                metrics.append(n)
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + ["%s %s" % (wk, k) for wk, k, _ in cols])
        for n in metrics:
            unit = next((m[n][0] for _, _, m in cols if n in m), "")
            w.writerow([n, unit] + [m.get(n, ("", ""))[1] for _, _, m in cols])
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tscale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    try:
        cur = json.load(open(out_json))
    except Exception:
        cur = {}
    for wk, k, m in cols:
        short = next((v for key, v in SHORT if key in k), k)
        if "__half" in k:
            short += "_f16"

        def val(n, table):
            u, v = m[n]
            return float(v.replace(",", "")) * table[u]
        e = {"kernel": k, "dram_bytes_read": val("dram__bytes_read.sum", scale), "dram_bytes_write": val("dram__bytes_write.sum", scale),
             "duration_us_under_ncu": val("gpu__time_duration.sum", tscale)}
        cur.setdefault(wk or "unkeyed", {}).setdefault(short, e)  # the first launch seen of a stage stays
    cur["source"] = "ncu --set full --clock-control none over tools/profile_driver.py (tools/gpu_profile_pass.sh), first launch of each kernel per workload; metric table beside this file"
    json.dump(cur, open(out_json, "w"), indent=1)
    print(json.dumps({k: list(v) for k, v in cur.items() if isinstance(v, dict)}, indent=1))


if __name__ == "__main__":
    main()
