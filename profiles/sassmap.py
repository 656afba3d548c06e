#!/usr/bin/env python
"""Per-SASS-instruction attribution of an `ncu --page source --csv --print-source cuda,sass` dump.
With inlining ncu lists one SASS instruction under every source line of its inline stack; this script
keeps each address once, under the most specific line (the one that owns the fewest instructions), and
prints instructions per warp-tick by line and by function.
usage: sassmap.py file.csv warp_ticks [--lines] [--sass LO HI]"""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1])))
W = float(sys.argv[2])
fname, hdr, cur, seen, first_path = None, None, None, 0, None
line_addrs = collections.defaultdict(list)   # (file, line, text) -> [addr]
addr_info = {}                               # addr -> (sass, inst, thr_inst, samples)
order = []
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name" or r[0] == "Function Name": continue
    if r[0] == "File Path":
        if first_path is None: first_path = r[1]
        elif r[1] == first_path: break          # second launch of the same kernel
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; ie, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples"); continue
    if hdr is None or len(r) != len(hdr): continue
    if r[0] != "":
        cur = (fname or "?", int(r[0]), r[1].strip()[:120]); continue
    addr = r[2]
    if addr in ("", "-"): continue
    line_addrs[cur].append(addr)
    if addr not in addr_info:
        order.append(addr)
    fl = lambda x: float(x) if x not in ("", "-") else 0.0
    addr_info[addr] = (r[3].strip(), fl(r[ie]), fl(r[it]), fl(r[isamp]))
owner = {}
for key, addrs in line_addrs.items():
    for a in addrs:
        if a not in owner or len(line_addrs[owner[a]]) > len(addrs):
            owner[a] = key
tot = sum(v[1] for v in addr_info.values()); tots = sum(v[3] for v in addr_info.values())
print(f"unique SASS {len(addr_info)}  warp-inst {tot:.4g}  per warp-tick {tot/W:.1f}  samples {tots:.0f}")
bl = collections.defaultdict(lambda: [0, 0, 0, 0])
for a, key in owner.items():
    s, i, t, sm = addr_info[a]
    b = bl[key]; b[0] += i; b[1] += t; b[2] += sm; b[3] += 1
if "--lines" in sys.argv:
    for key, b in sorted(bl.items(), key=lambda kv: (kv[0][0], kv[0][1])):
        if b[0] == 0: continue
        print(f"{key[0][:14]:14s}:{key[1]:4d} {b[0]/W:6.1f}/wt {100*b[0]/tot:4.1f}% samp {100*b[2]/tots:4.1f}% thr {b[1]/b[0]:4.1f} n{b[3]:3d}  {key[2]}")
if "--sass" in sys.argv:
    k = sys.argv.index("--sass"); lo, hi = int(sys.argv[k + 1]), int(sys.argv[k + 2])
    f = sys.argv[k + 3] if len(sys.argv) > k + 3 else None
    for a in order:
        key = owner[a]
        if lo <= key[1] <= hi and (f is None or key[0].startswith(f)):
            s, i, t, sm = addr_info[a]
            print(f"{key[0][:10]}:{key[1]:4d} {i/W:5.2f} thr {t/i if i else 0:4.1f} smp {sm:5.0f}  {s}")
if "--funcs" in sys.argv:
    # enclosing function of a source line: the nearest non-indented line above it that opens a body
    import os
    srcdir = sys.argv[sys.argv.index("--funcs") + 1]
    fmap = {}
    def enclosing(fn, ln):
        if fn not in fmap:
            path = os.path.join(srcdir, fn)
            heads = []
            if os.path.exists(path):
                for i, l in enumerate(open(path), 1):
                    if l[:1] not in (" ", "\t", "#", "/", "}", "\n") and "(" in l:
                        m = re.search(r"([A-Za-z_][A-Za-z0-9_]*)\s*\(", l.split("(")[0] + "(")
                        heads.append((i, m.group(1) if m else l.strip()[:30]))
            fmap[fn] = heads
        name = fn
        for i, h in fmap[fn]:
            if i <= ln: name = h
            else: break
        return name
    bf = collections.defaultdict(lambda: [0, 0, 0])
    for key, b in bl.items():
        f = enclosing(key[0], key[1]) if key else "?"
        bf[f][0] += b[0]; bf[f][1] += b[1]; bf[f][2] += b[2]
    for f, b in sorted(bf.items(), key=lambda kv: -kv[1][0]):
        if b[0] == 0: continue
        print(f"{f[:28]:28s} {b[0]/W:7.1f}/wt {100*b[0]/tot:5.1f}%  samp {100*b[2]/tots:5.1f}%  thr {b[1]/b[0]:4.1f}")
