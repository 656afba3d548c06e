#!/usr/bin/env python
"""Bucket the per-line instruction counts of hotlines.py's input by enclosing function.
usage: hotfuncs.py src_cs.csv  (run from the repo root)"""
import csv, re, sys
files = {f: open(f"deep-rl-battlespace_b200/csrc/{f}").read().split("\n") for f in ("tick.cuh", "battlespace.cu")}
def func_of(fname, line):
    src = files[fname]
    for k in range(line - 1, -1, -1):
        m = re.match(r"^(?:__device__|__global__|template|static|inline).*?(\w+)\s*\(", src[k])
        if m and not src[k].startswith("template <"): return m.group(1)
        m = re.match(r"^__global__ void.*?(\w+)\(", src[k])
        if m: return m.group(1)
    return "?"
rows = list(csv.reader(open(sys.argv[1])))
hdr, agg = None, {}
for r in rows:
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr) or r[0] == "": continue
    ln, text = int(r[0]), r[1].strip()
    fname = None
    for f, src in files.items():
        if ln <= len(src) and src[ln - 1].strip() == text: fname = f
    if fname is None: key = "(libdevice/other)"
    else: key = f"{fname}:{func_of(fname, ln)}"
    a = agg.setdefault(key, [0, 0, 0])
    a[0] += float(r[hdr.index("Instructions Executed")] or 0); a[1] += float(r[hdr.index("Thread Instructions Executed")] or 0)
    a[2] += float(r[hdr.index("# Samples")] or 0)
ti = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:45s} inst {100*a[0]/ti:5.1f}%  samp {100*a[2]/ts:5.1f}%  thr {a[1]/max(a[0],1):4.1f}")
