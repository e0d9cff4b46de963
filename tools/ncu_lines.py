#!/usr/bin/env python3
"""Top source lines / stall reasons of the kernels in an ncu report's source page.
usage: ncu_lines.py <source.csv> [n_lines]   (csv from: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass)"""
import csv
import sys


def num(x):
    try:
        return float(x)
    except Exception:
        return 0.0


rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
fp = fn = hdr = None
data = []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fp = r[1]
        continue
    if r[0] == "Function Name":
        fn = r[1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr and r[0].isdigit():
        data.append((fp, fn, dict(zip(hdr, r))))
for kern in sorted({d[1] for d in data}):
    sel = [d for d in data if d[1] == kern and d[0]]
    tot = sum(num(d[2]['# Samples']) for d in sel) or 1.0
    stall = {}
    for d in sel:
        for k, v in d[2].items():
            if k.startswith('stall_') and 'Not Issued' not in k:
                stall[k] = stall.get(k, 0) + num(v)
    print("=====", kern[:70], "samples", tot)
    print("  stalls:", ", ".join("%s %.1f%%" % (k[6:], 100 * v / tot) for k, v in sorted(stall.items(), key=lambda kv: -kv[1])[:9]))
    inst = sum(num(d[2]['Instructions Executed']) for d in sel)
    thr = sum(num(d[2]['Thread Instructions Executed']) for d in sel)
    print("  warp instructions %.3g, threads/instruction %.1f" % (inst, thr / max(inst, 1)))
    sel.sort(key=lambda d: -num(d[2]['# Samples']))
    for d in sel[:top]:
        m = d[2]
        st = {k: num(v) for k, v in m.items() if k.startswith('stall_') and 'Not Issued' not in k}
        t3 = sorted(st.items(), key=lambda kv: -kv[1])[:2]
        ie = num(m['Instructions Executed'])
        print("  %5.1f%% %s:%s inst %.3g thr %.1f | %s" % (100 * num(m['# Samples']) / tot, d[0].split('/')[-1], m['Line No'], ie,
              num(m['Thread Instructions Executed']) / max(ie, 1), ", ".join("%s %d" % (k[6:], v) for k, v in t3)))
