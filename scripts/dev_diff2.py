import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import util
from arachne_b200 import ArachneCuda
from oracle.refbwa import RefBwa
name = sys.argv[1]; os.makedirs("/tmp/w", exist_ok=True)
fa, ctg, rs = util.make_dataset(name, "/tmp/w", damaged=True)
R = RefBwa().load(fa); A = ArachneCuda().load_index(fa)
for p in (13, 17, 47):
    seq = rs.seq[rs.off[2*p]:rs.off[2*p+2]]; off = rs.off[2*p:2*p+3] - rs.off[2*p]
    res = A.align_batch(seq, off, None, stages=3)
    e1, e2 = R.pair(rs.read(2*p), rs.read(2*p+1))
    print("PAIR", p, "single-pair batch: gpu n =", np.diff(res["reg_off"]), "ref n =", len(e1), len(e2))
    print(res["regs"])
# growing prefixes: where does it start to differ?
for n in (14, 20, 48, 100, 400):
    seq = rs.seq[:rs.off[2*n]]; off = rs.off[:2*n+1]
    res = A.align_batch(seq, off, None, stages=3)
    ref = R.batch(seq, off, rescue=True, cigar=False, threads=8)
    print("prefix", n, "equal", np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes())
