#!/usr/bin/env python
"""Dev check on a GPU box: build the index of a large synthetic reference on the device, load it, and compare a sample
of read pairs against the reference's own code (oracle/_ref) loading the same files.  python scripts/dev_bigindex.py [Mb] [pairs]"""
import os, sys, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import ArachneCuda, simulate as S

mb = float(sys.argv[1]) if len(sys.argv) > 1 else 3100
npairs = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
wd = os.environ.get("ARACHNE_BENCH_DIR", "/tmp/arachne_bench"); os.makedirs(wd, exist_ok=True)
L = int(mb * 1e6)
t = time.time(); ctg = S.make_reference(L, 24 if L >= 24_000_000 else 1, seed=2); print("reference %.1fs" % (time.time() - t), flush=True)
p = os.path.join(wd, f"ref{L}")
t = time.time(); S.write_bns(p, ctg); print("write_bns %.1fs" % (time.time() - t), flush=True)
A = ArachneCuda()
t = time.time(); info = A.build_index(p, 0, 0); print("build_index %.1fs" % (time.time() - t), json.dumps(info), flush=True)
open(p + ".done", "w").write("ok")
t = time.time(); A.load_index(p); print("load_index %.1fs, device bytes %d" % (time.time() - t, A.lib.arachne_index_device_bytes(A.idx)), flush=True)
rs = S.simulate_pairs_fast(ctg, npairs, seed=12)
t = time.time(); res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=7); print("gpu batch %.2fs" % (time.time() - t), flush=True)
from oracle.refbwa import RefBwa, available
if available():
    t = time.time(); R = RefBwa().load(p); print("ref load %.1fs" % (time.time() - t), flush=True)
    t = time.time(); ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=os.cpu_count()); dt = time.time() - t
    print("ref batch %.2fs (%.0f pairs/s on %d threads)" % (dt, npairs / dt, os.cpu_count()), flush=True)
    ok = np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes()
    print("regions identical:", ok, "n_regs", len(ref["reg_off"]) and int(ref["reg_off"][-1]))
    a, e = res["alns"], ref["aln"]  # e: pos rid is_rev NM n_cigar cigar_off score sub
    same = all(np.array_equal(a[f], e[:, j]) for f, j in (("pos", 0), ("rid", 1), ("is_rev", 2), ("nm", 3), ("n_cigar", 4), ("score", 6), ("sub", 7)))
    cig = same and all(np.array_equal(res["cigar"][a["cigar_off"][i]:a["cigar_off"][i] + a["n_cigar"][i]], ref["cigar"][e[i, 5]:e[i, 5] + e[i, 4]]) for i in range(len(a)))
    print("alignments identical:", same, "cigars identical:", cig, "max pos", int(a["pos"].max()), "max rb", int(np.frombuffer(ref["regs"].tobytes(), dtype=np.int64)[::11].max()))
    g = np.random.default_rng(1); k = g.integers(0, 2 * A.l_pac, 5000).astype(np.uint64)
    print("sa identical:", np.array_equal(A.sa(k), np.array([R.sa(int(x)) for x in k], dtype=np.uint64)))
