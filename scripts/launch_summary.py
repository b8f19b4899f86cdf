"""Summarise an `ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv` launch list
(read here, without a GPU) into the per-kernel table kept under profiles/.

    python scripts/launch_summary.py gpurun_out/<tag>_launches.csv profiles/<name>.txt ["header line"]
"""
import collections
import csv
import re
import sys

UNIT = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(path, out, header=""):
    rows = list(csv.reader(open(path, errors="replace")))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[start]
    ki, mi, ui, vi, idi = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Unit", "Metric Value", "ID"))
    per = collections.defaultdict(lambda: [0, 0.0, 0.0])      # name -> [launches, ms, dram bytes]
    seen = set()
    for r in rows[start + 1:]:
        if len(r) <= vi:
            continue
        name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("mossdec::<unnamed>::", "").replace("mossdec::(anonymous namespace)::", "")
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        if r[mi] == "gpu__time_duration.sum":
            per[name][1] += v * UNIT.get(r[ui], 1e-6)
            if r[idi] not in seen:
                seen.add(r[idi])
                per[name][0] += 1
        elif r[mi].startswith("dram__bytes"):
            per[name][2] += v * BYTES.get(r[ui], 1.0)
    total = sum(v[1] for v in per.values())
    lines = [header] if header else []
    lines.append(f"total kernel time {total:.1f} ms over {sum(v[0] for v in per.values())} launches")
    lines.append("")
    lines.append(f"{'kernel':<58}{'launches':>9}{'total ms':>11}{'share':>8}{'mean us':>10}{'DRAM MB/launch':>16}{'DRAM GB/s':>11}")
    for name, (n, ms, by) in sorted(per.items(), key=lambda kv: -kv[1][1]):
        gbs = by / (ms / 1e3) / 1e9 if ms > 0 and by > 0 else 0.0
        lines.append(f"{name[-58:]:<58}{n:>9}{ms:>11.2f}{100 * ms / total:>7.1f}%{1e3 * ms / max(n, 1):>10.1f}"
                     f"{(by / max(n, 1) / 1e6 if by else 0):>16.1f}{gbs:>11.0f}")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:45]))


if __name__ == "__main__":
    main(*sys.argv[1:4])
