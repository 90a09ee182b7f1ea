#!/usr/bin/env python
"""Throughput of the native FASTQ ingest (csrc/ingest.cu) on this host: pairs/s for plain and gzip-compressed input, next to the
GPU's pairs/s so the end-to-end bottleneck is visible.   python scripts/ingest_bench.py [pairs=1000000] [threads=nproc]"""
import gzip, json, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

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
def bgzf_write(src, dst, level=1, block=0xff00):
    """bgzip-style output (SAM spec 4.1): gzip members of <= 64 KB with the BC extra field, + the EOF marker."""
    import struct, zlib
    with open(src, "rb") as f, open(dst, "wb") as o:
        while True:
            d = f.read(block)
            if not d:
                break
            c = zlib.compressobj(level, zlib.DEFLATED, -15); comp = c.compress(d) + c.flush()
            o.write(struct.pack("<BBBBIBBHBBHH", 0x1f, 0x8b, 8, 4, 0, 0, 0xff, 6, 66, 67, 2, len(comp) + 25) + comp + struct.pack("<II", zlib.crc32(d), len(d)))
        o.write(bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000"))

import ctypes as C
from arachne_b200.lib import IngestBatch, lib_path
LIB = C.CDLL(lib_path())
LIB.arachne_ingest_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
LIB.arachne_ingest_next.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.POINTER(IngestBatch))]
LIB.arachne_ingest_batch_free.argtypes = [C.POINTER(IngestBatch)]
LIB.arachne_ingest_close.argtypes = [C.c_void_p]

def run(a, b):
    """The C entry points alone: batches are taken and released, nothing is copied into Python objects."""
    t0 = os.times(); t = time.perf_counter(); h = C.c_void_p(); got = units = 0
    assert LIB.arachne_ingest_open(a.encode(), b.encode(), thr, C.byref(h)) == 0
    while True:
        pb = C.POINTER(IngestBatch)()
        rc = LIB.arachne_ingest_next(h, 250000, C.byref(pb))
        if rc == 1:
            break
        assert rc == 0, rc
        got += pb.contents.n_pairs; units += pb.contents.n_units
        LIB.arachne_ingest_batch_free(pb)
    dt = time.perf_counter() - t; LIB.arachne_ingest_close(h); t1 = os.times()
    return got, units, dt, (t1.user - t0.user) + (t1.system - t0.system)

res = {"pairs": n, "parser_threads": thr, "host_cores": os.cpu_count(), "plain_bytes": os.path.getsize(p1) + os.path.getsize(p2),
       "note": "C entry points only (arachne_ingest_open / _next / _batch_free); best of 3; cpu_us_per_pair = process CPU time / pairs, all threads"}
for kind in ("plain", "gz", "bgzf"):
    ext = {"plain": "", "gz": ".gz", "bgzf": ".bgz"}[kind]
    a, b = p1 + ext, p2 + ext
    if kind == "gz" and not os.path.exists(b):
        for p in (p1, p2):
            subprocess.run("gzip -1 -k -f " + p, shell=True, check=True)
    if kind == "bgzf" and not os.path.exists(b):
        for p in (p1, p2):
            bgzf_write(p, p + ext)
    best = None
    for _ in range(3):
        got, units, dt, cpu = run(a, b)
        assert got == n
        if best is None or dt < best[0]:
            best = (dt, cpu, units)
    res[kind] = {"seconds": best[0], "pairs_per_s": n / best[0], "cpu_us_per_pair": 1e6 * best[1] / n, "work_units": best[2]}
print(json.dumps(res))
