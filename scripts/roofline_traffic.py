#!/usr/bin/env python
"""profiles/roofline_traffic.json from an `ncu --set full` report of k_seed2 captured on bench.py's own workload:
   python scripts/roofline_traffic.py gpurun_out/<tag>_k_seed2.ncu-rep <pairs_per_step> <ref_mb>"""
import csv, json, subprocess, sys
rep, pairs, ref_mb = sys.argv[1], int(sys.argv[2]), float(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines())); h, u, row = r[0], r[1], r[2]
def val(name):
    i = h.index(name); v = float(row[i].replace(",", "")); unit = u[i].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[unit]
rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
j = {"pairs_per_step": pairs, "ref_mb": ref_mb, "dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr,
     "note": f"dram__bytes_read.sum + dram__bytes_write.sum of one k_seed2 launch, ncu --set full on bench.py --pairs {pairs} ({rep.split('/')[-1]})"}
json.dump(j, open("profiles/roofline_traffic.json", "w"), indent=1); print(j)
