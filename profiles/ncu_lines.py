#!/usr/bin/env python
"""Aggregate `ncu --page source --csv --print-source cuda,sass` output per CUDA source line:
warp instructions executed, average active threads, stall samples. usage: ncu_lines.py file.csv [topN]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# the file is a sequence of per-file tables: "File Path",path / "Function Name",... / header / rows
agg = collections.OrderedDict()
cur_file, hdr = None, None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    d = dict(zip(hdr, r))
    try:
        ie = float(d["Instructions Executed"] or 0); te = float(d["Thread Instructions Executed"] or 0); ss = float(d["# Samples"] or 0)
    except (ValueError, KeyError):
        continue
    if not d["Line No"].strip(): continue  # SASS rows repeat their source line's counts
    key = (cur_file, d["Line No"])
    a = agg.setdefault(key, [0.0, 0.0, 0.0, d.get("Source", "")[:110]])
    a[0] += ie; a[1] += te; a[2] += ss
tot_i = sum(a[0] for a in agg.values()); tot_t = sum(a[1] for a in agg.values()); tot_s = sum(a[2] for a in agg.values())
print(f"total warp-inst {tot_i:.3e} thread-inst {tot_t:.3e} avg threads {tot_t/max(tot_i,1):.2f} samples {tot_s:.0f}")
print(f"{'file:line':24s} {'inst%':>6s} {'thr':>5s} {'smp%':>6s}  source")
for k, a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{k[0]+':'+k[1]:24s} {100*a[0]/tot_i:6.2f} {a[1]/max(a[0],1):5.1f} {100*a[2]/max(tot_s,1):6.2f}  {a[3].strip()}")
