#!/usr/bin/env python
"""Direct identity check of arachne_index_build against the reference's own `bwa index` (oracle/_ref) at a size where the reference
takes its rope path (bwtindex.c:96-121: above 50 Mb) and still finishes in minutes.  Writes md5 of .bwt / .sa from both builders.
   python scripts/index_identity.py [Mb=200] [out.json]"""
import hashlib, json, os, shutil, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import ArachneCuda, simulate as S
from oracle.refbwa import RefBwa

mb = float(sys.argv[1]) if len(sys.argv) > 1 else 200
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "index_identity.json")
wd = "/tmp/arachne_identity"; shutil.rmtree(wd, ignore_errors=True); os.makedirs(wd)
L = int(mb * 1e6)
ctg = S.make_reference(L, 24 if L >= 24_000_000 else 1, seed=2)
md5 = lambda p: hashlib.md5(open(p, "rb").read()).hexdigest()
res = {"ref_mb": mb, "contigs": len(ctg)}
fa = os.path.join(wd, "ref.fa"); S.write_fasta(fa, ctg)
t = time.time(); RefBwa().build_index(fa); res["bwa_index_seconds"] = time.time() - t
dev = os.path.join(wd, "dev"); S.write_bns(dev, ctg)
res["pac_identical"] = md5(fa + ".pac") == md5(dev + ".pac"); res["ann_identical"] = open(fa + ".ann").read().split("\n")[1:] == open(dev + ".ann").read().split("\n")[1:]
t = time.time(); res["device_build"] = ArachneCuda().build_index(dev, 0, 0); res["device_seconds_total"] = time.time() - t
for ext in ("bwt", "sa"):
    res[ext] = {"reference_md5": md5(fa + "." + ext), "device_md5": md5(dev + "." + ext), "bytes": os.path.getsize(fa + "." + ext)}
    res[ext]["identical"] = res[ext]["reference_md5"] == res[ext]["device_md5"]
res["ok"] = bool(res["pac_identical"] and res["bwt"]["identical"] and res["sa"]["identical"])
json.dump(res, open(out, "w"), indent=1); print(json.dumps(res))
sys.exit(0 if res["ok"] else 1)
