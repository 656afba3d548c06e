#!/usr/bin/env python
"""Static SASS per source line for one kernel of `nvdisasm -g file.cubin` output.
usage: sass_by_line.py all_g.sass kernel_substring [opcode_regex]"""
import re, sys, collections
txt = open(sys.argv[1]).read().split("\n")
kern, pat = sys.argv[2], (re.compile(sys.argv[3]) if len(sys.argv) > 3 else None)
inside, cur = False, None
cnt, hit = collections.Counter(), collections.Counter()
for l in txt:
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m: inside = kern in m.group(1); continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
        cnt[cur] += 1
        if pat and pat.search(l): hit[cur] += 1
print("instructions", sum(cnt.values()))
src = hit if pat else cnt
for k, v in sorted(src.items(), key=lambda kv: (str(kv[0][0]), kv[0][1])): print(k, v)
