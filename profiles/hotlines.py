#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` per CUDA source line.
usage: hotlines.py file.csv [top]   (first kernel result in the file only)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname, hdr, agg, seen_kernels = None, None, {}, 0
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name":
        seen_kernels += 1
        if seen_kernels > 1: break
        continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr) or r[0] == "": continue
    ie, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    key = (fname or "?", int(r[0]), r[1].strip()[:100])
    a = agg.setdefault(key, [0, 0, 0])
    a[0] += float(r[ie] or 0); a[1] += float(r[it] or 0); a[2] += float(r[isamp] or 0)
ti = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print(f"total warp-inst {ti:.4g}  samples {ts:.0f}  lines {len(agg)}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    thr = a[1] / a[0] if a[0] else 0
    print(f"{k[0][:14]:14s}:{k[1]:4d} inst {100*a[0]/ti:5.1f}%  samp {100*a[2]/ts:5.1f}%  thr {thr:4.1f}  {k[2]}")
