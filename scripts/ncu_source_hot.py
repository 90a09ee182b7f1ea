#!/usr/bin/env python
"""Hot source lines of a kernel from an ncu report captured with --import-source on (built with -lineinfo):
   python scripts/ncu_source_hot.py report.ncu-rep [top]   -> file:line, warp instructions, avg active lanes, stall samples"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; hdr = None; agg = {}
for row in csv.reader(out.splitlines()):
    if not row: continue
    if row[0] == "File Path": cur = row[1].split("/")[-1]; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or len(row) < len(hdr) or row[2] != "-": continue   # keep the per-source-line summary rows
    d = dict(zip(hdr, row))
    try:
        ie = int(d["Instructions Executed"]); te = int(d["Thread Instructions Executed"]); ss = int(d["# Samples"])
    except ValueError:
        continue
    agg[(cur, int(row[0]))] = (ie, te, ss, row[1].strip()[:110])
tot_i = sum(v[0] for v in agg.values()); tot_s = sum(v[2] for v in agg.values())
print(f"total warp instructions {tot_i:,}  stall samples {tot_s:,}")
byfile = collections.Counter()
for (f, l), v in agg.items(): byfile[f] += v[0]
print("by file:", ", ".join(f"{f} {100*c/tot_i:.1f}%" for f, c in byfile.most_common()))
for (f, l), v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f}:{l:<4d} inst {100*v[0]/tot_i:5.1f}%  lanes {v[1]/max(v[0],1):5.1f}  samples {100*v[2]/max(tot_s,1):5.1f}%  | {v[3]}")
