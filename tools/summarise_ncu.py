"""Summarise `ncu -i X.ncu-rep --page raw --csv` into a markdown table and a JSON of per-launch DRAM traffic.

usage: ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv
       python tools/summarise_ncu.py raw.csv profiles/r1_ncu_full.md profiles/r1_traffic.json
"""
import csv
import json
import re
import sys

METRICS = [
    ("gpu__time_duration.sum", "ms", 1e-6),
    ("dram__bytes_read.sum", "DRAM rd MB", None),
    ("dram__bytes_write.sum", "DRAM wr MB", None),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %", 1),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads / inst", 1),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe %", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("smsp__sass_inst_executed_op_local_ld.sum", "local ld Minst", 1e-6),
    ("smsp__sass_inst_executed_op_local_st.sum", "local st Minst", 1e-6),
]


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)


def main():
    raw, md, js = sys.argv[1:4]
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    names = [m for m, _, _ in METRICS if m in col]
    out = []
    traffic = {}
    for r in data:
        k = re.sub(r"\((SceneView|tc::SceneView).*$", "", r[col["Kernel Name"]])
        k = re.sub(r"^.*::(?=k_)", "", k).replace("void ", "")
        rec = {"kernel": k, "grid": r[col["Grid Size"]], "block": r[col["Block Size"]]}
        for m, label, scale in METRICS:
            if m not in col:
                rec[label] = None
                continue
            v, u = r[col[m]], units[col[m]]
            if scale is None:
                rec[label] = to_bytes(v, u) / 1e6
            else:
                x = float(v.replace(",", ""))
                if m == "gpu__time_duration.sum":
                    x *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
                elif scale != 1:
                    x *= scale
                rec[label] = x
        out.append(rec)
        t = traffic.setdefault(k, {"launches": 0, "dram_bytes": 0.0, "ms": 0.0})
        t["launches"] += 1
        t["dram_bytes"] += (rec["DRAM rd MB"] + rec["DRAM wr MB"]) * 1e6
        t["ms"] += rec["ms"]
    labels = [lab for _, lab, _ in METRICS]
    with open(md, "w") as f:
        f.write(f"source: `{raw}` (ncu --set full --clock-control none; replayed, cold caches: use for shares and "
                "counters, not absolute times)\n\n")
        f.write("| kernel | grid | " + " | ".join(labels) + " |\n|---|---|" + "---:|" * len(labels) + "\n")
        for rec in out:
            f.write(f"| `{rec['kernel']}` | {rec['grid']} | " + " | ".join(
                "–" if rec[lab] is None else f"{rec[lab]:.2f}" for lab in labels) + " |\n")
    for t in traffic.values():
        t["dram_bytes_per_launch"] = t["dram_bytes"] / t["launches"]
        t["ms_per_launch"] = t["ms"] / t["launches"]
    json.dump(traffic, open(js, "w"), indent=1)
    print(open(md).read())


if __name__ == "__main__":
    main()
