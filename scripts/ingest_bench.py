#!/usr/bin/env python
"""Throughput of the native FASTQ ingest (csrc/ingest.cu) on this host: pairs/s for plain and gzip-compressed input, next to the
GPU's pairs/s so the end-to-end bottleneck is visible.   python scripts/ingest_bench.py [pairs=1000000] [threads=nproc]"""
import gzip, json, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import Ingest

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
thr = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 4)
wd = "/tmp/arachne_ingest_bench"; os.makedirs(wd, exist_ok=True)
g = np.random.default_rng(1)
L = 150
bases = np.frombuffer(b"ACGT", np.uint8)
per_bc = 30
def write(path, mate):
    with open(path, "wb") as f:
        for c0 in range(0, n, 100000):
            m = min(100000, n - c0)
            seq = bases[g.integers(0, 4, (m, L))]
            qual = np.full((m, L), ord("I"), np.uint8)
            out = []
            for i in range(m):
                k = c0 + i
                out.append(b"@sim%d/%d\tBX:Z:A%02dC%02dB%02dD%02d-1\tVX:i:1\n" % (k, mate, (k // per_bc) % 96, (k // per_bc // 96) % 96, (k // per_bc // 9216) % 96, 1))
                out.append(seq[i].tobytes()); out.append(b"\n+\n"); out.append(qual[i].tobytes()); out.append(b"\n")
            f.write(b"".join(out))
p1, p2 = os.path.join(wd, "r1.fq"), os.path.join(wd, "r2.fq")
if not os.path.exists(p2) or os.path.getsize(p2) < n * 300:
    write(p1, 1); write(p2, 2)
res = {"pairs": n, "parser_threads": thr, "host_cores": os.cpu_count(), "plain_bytes": os.path.getsize(p1) + os.path.getsize(p2)}
for kind in ("plain", "gz"):
    a, b = (p1, p2) if kind == "plain" else (p1 + ".gz", p2 + ".gz")
    if kind == "gz" and not os.path.exists(b):
        for p in (p1, p2):
            subprocess.run("gzip -1 -k -f " + p, shell=True, check=True)
    t = time.perf_counter(); I = Ingest(a, b, n_threads=thr); got = units = 0
    while True:
        x = I.next(250000, copy=False)
        if x is None:
            break
        got += x["n_pairs"]; units += x["n_units"]
    dt = time.perf_counter() - t; I.close()
    assert got == n
    res[kind] = {"seconds": dt, "pairs_per_s": n / dt, "work_units": units}
print(json.dumps(res))
