"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals and shares.

usage: python tools/summarise_launches.py gpurun_out/launches_r1.csv [passes] > profiles/r1_launches.md
`passes` = number of hot-path passes (tc_eval + rows) the profiled command ran, to print per-pass times.
"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    passes = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"]
        name = re.sub(r"\(.*$", "", name)
        name = name.replace("tc::<unnamed>::", "").replace("void ", "")
        name = re.sub(r"cub::CUB_\w+::", "cub::", name)
        if len(name) > 70:
            name = name[:67] + "..."
        rows.append((name, float(r["Metric Value"]) / 1e6, r["Grid Size"], r["Block Size"]))
    agg = OrderedDict()
    for name, ms, grid, block in rows:
        a = agg.setdefault(name, [0, 0.0, 0.0, grid, block])
        a[0] += 1
        a[1] += ms
        a[2] = max(a[2], ms)
        if ms == a[2]:
            a[3], a[4] = grid, block
    total = sum(a[1] for a in agg.values())
    print(f"source: `{path}` — {len(rows)} launches, {total:.2f} ms of kernel time under ncu "
          f"(serialised, cold caches; shares matter, not absolutes), {passes} hot-path pass(es)\n")
    print("| kernel | launches | total ms | ms / pass | share | largest launch ms | grid × block of largest |")
    print("|---|---:|---:|---:|---:|---:|---|")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {a[0]} | {a[1]:.3f} | {a[1] / passes:.3f} | {100 * a[1] / total:.1f}% | {a[2]:.3f} | {a[3]} × {a[4]} |")


if __name__ == "__main__":
    main()
