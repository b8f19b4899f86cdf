"""Summarise the ncu launch list of scripts/bench_cv1.py (gpu__time_duration.sum + dram__bytes_{read,write}.sum per launch)
into a per-kernel table: launches, mean duration, mean DRAM bytes, achieved DRAM GB/s against the measured HBM peak.

    python scripts/cv1_launch_summary.py gpurun_out/<tag>_launches.csv profiles/<name>.txt
"""
import collections
import csv
import json
import os
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0,
        "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = [r for r in csv.reader(open(src)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    per = collections.defaultdict(dict)
    names = {}
    for r in rows[1:]:
        if len(r) < len(hdr):
            continue
        lid = r[ix["ID"]]
        names[lid] = r[ix["Kernel Name"]].split("(")[0]
        grid = r[ix["Grid Size"]] if "Grid Size" in ix else ""
        val = float(r[ix["Metric Value"]].replace(",", "")) * UNIT.get(r[ix["Metric Unit"]], 1.0)
        per[lid][r[ix["Metric Name"]]] = val
        per[lid]["grid"] = grid
    agg = collections.OrderedDict()
    for lid, m in per.items():
        key = (names[lid], m.get("grid", ""))
        a = agg.setdefault(key, [0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += m.get("gpu__time_duration.sum", 0.0)
        a[2] += m.get("dram__bytes_read.sum", 0.0)
        a[3] += m.get("dram__bytes_write.sum", 0.0)
    peak = None
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        try:
            j = json.load(open(pk))
            peak = j.get("hbm_gbs") or j.get("hbm_gbps")
        except Exception:
            peak = None
    lines = [f"ncu launch list summary of {src} (cold-cache, serialised launches; --clock-control none)",
             f"HBM peak used for the fraction: {peak} GB/s (MEASURED_PEAKS.json)" if peak else "HBM peak: MEASURED_PEAKS.json not readable",
             "", f"{'kernel':44s} {'grid':>18s} {'n':>4s} {'us/launch':>10s} {'MB read':>9s} {'MB written':>10s} {'DRAM GB/s':>10s} {'frac':>6s}"]
    for (name, grid), (n, t, rd, wr) in agg.items():
        us = t / n * 1e6
        gbps = (rd + wr) / t / 1e9 if t > 0 else 0.0
        frac = f"{gbps / peak:.2f}" if peak else "-"
        lines.append(f"{name[-44:]:44s} {grid:>18s} {n:4d} {us:10.1f} {rd / n / 1e6:9.2f} {wr / n / 1e6:10.2f} {gbps:10.0f} {frac:>6s}")
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
