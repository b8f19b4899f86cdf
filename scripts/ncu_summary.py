"""Summarise an `ncu --set full` report (read here, without a GPU) into the text files kept under profiles/.

    python scripts/ncu_summary.py gpurun_out/<tag>_full.ncu-rep profiles/<name>.txt [--traffic profiles/ncu_dram_traffic.json]

Per captured launch: duration, tensor-pipe utilisation, issue-slot utilisation, DRAM bytes read / written, L2 and L1
throughput, registers, and the top warp-stall reasons of the hottest source lines.  With --traffic the mean
`dram__bytes_read.sum + dram__bytes_write.sum` per launch of every kernel is written as JSON; bench.py reports it as
`roofline.traffic`.
"""
import collections
import csv
import io
import json
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots active %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of max"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/TEX throughput %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("smsp__inst_executed.sum", "warp instructions"),
]
UNIT_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def short(name):
    for k in ("attention_ws_kernel", "attention_tc_kernel", "ffn_tc_kernel", "gemm_tc_kernel"):
        if k in name:
            return k + (name[name.index(k) + len(k):name.index("(")] if "<" in name[name.index(k):name.index("(")] else "")
    return name.split("(")[0][-48:]


def main():
    rep, out_path = sys.argv[1], sys.argv[2]
    traffic_path = sys.argv[sys.argv.index("--traffic") + 1] if "--traffic" in sys.argv else None
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = [f"ncu --set full --clock-control none summary of {rep}", ""]
    traffic = collections.defaultdict(list)
    for r in rows[2:]:
        name = short(r[idx["Kernel Name"]])
        lines.append(f"[launch {r[idx['ID']]}] {name}")
        rd = wr = 0.0
        for key, label in METRICS:
            if key not in idx:
                continue
            v, u = r[idx[key]], units[idx[key]]
            lines.append(f"    {label:28s} {v} {u}")
            if key == "dram__bytes_read.sum":
                rd = float(v.replace(",", "")) * UNIT_BYTES.get(u, 1)
            if key == "dram__bytes_write.sum":
                wr = float(v.replace(",", "")) * UNIT_BYTES.get(u, 1)
        traffic[name.split("<")[0]].append(rd + wr)
        lines.append("")
    # stall reasons per kernel (source page, all launches of the report in order)
    src = ncu_csv(rep, "source")
    blocks, cur = [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = {"name": short(r[1]), "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    seen = set()
    for b in blocks:
        if not b["rows"] or b["name"] in seen:
            continue
        seen.add(b["name"])
        h = b["rows"][0]
        ii = {n: i for i, n in enumerate(h)}
        stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
        tot = collections.Counter()
        for r in b["rows"][1:]:
            for s in stalls:
                try:
                    tot[s] += int(r[ii[s]])
                except (ValueError, IndexError):
                    pass
        T = sum(tot.values()) or 1
        lines.append(f"[warp stall sampling] {b['name']}: " + ", ".join(f"{s[6:]} {v / T * 100:.1f}%" for s, v in tot.most_common(7)))
        top = sorted(b["rows"][1:], key=lambda r: -int(r[ii["# Samples"]] or 0))[:8]
        for r in top:
            lines.append(f"    {int(r[ii['# Samples']] or 0):6d} samples  {r[ii['Source']].strip()[:90]}")
        lines.append("")
    open(out_path, "w").write("\n".join(lines) + "\n")
    if traffic_path:
        json.dump({k: sum(v) / len(v) for k, v in traffic.items()}, open(traffic_path, "w"), indent=1)
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()
