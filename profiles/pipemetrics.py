#!/usr/bin/env python
"""Pipe utilisation, shared-memory bank conflicts and branch / shared-atomic counts of the first kernel result in an
.ncu-rep (`ncu --set full`); appended to the *_raw.txt summaries.  usage: pipemetrics.py X.ncu-rep"""
import csv, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
pat = re.compile(r"^sm__(inst_executed_pipe_(fp64|lsu|alu|fma|xu|cbu|adu|uniform)|pipe_(fp64|fma|fmaheavy)_cycles_active)\.avg\.pct_of_peak_sustained_active$"
                 r"|^l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_(ld|st|atom))?\.sum$"
                 r"|^l1tex__data_pipe_lsu_wavefronts\.avg\.pct_of_peak_sustained_elapsed$"
                 r"|^smsp__inst_executed_op_(branch|shared_atom)\.sum$")
for i, k in enumerate(hdr):
    if pat.search(k): print(f"  {k:90s} {r[i]:>16s} {units[i]}")
