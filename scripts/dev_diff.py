import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import util
from arachne_b200 import ArachneCuda
from oracle.refbwa import RefBwa
name = sys.argv[1]; os.makedirs("/tmp/w", exist_ok=True)
fa, ctg, rs = util.make_dataset(name, "/tmp/w", damaged=True)
R = RefBwa().load(fa); A = ArachneCuda().load_index(fa)
res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=3)
ref = R.batch(rs.seq, rs.off, rescue=True, cigar=False, threads=8)
nr = R.batch(rs.seq, rs.off, rescue=False, cigar=False, threads=8)
shown = 0
for r in range(len(ref["reg_off"]) - 1):
    a = res["regs"][res["reg_off"][r]:res["reg_off"][r + 1]]; e = ref["regs"][ref["reg_off"][r]:ref["reg_off"][r + 1]]
    if a.tobytes() != e.tobytes():
        n0 = nr["regs"][nr["reg_off"][r]:nr["reg_off"][r + 1]]
        print("READ", r, "pair", r // 2, "len", rs.off[r + 1] - rs.off[r]); print(" gpu", a); print(" ref", e); print(" pre", n0)
        m = r ^ 1
        print(" mate pre", nr["regs"][nr["reg_off"][m]:nr["reg_off"][m + 1]])
        shown += 1
        if shown >= 3: break
res1 = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=1)
print("stage1 equal:", np.array_equal(res1["reg_off"], nr["reg_off"]) and res1["regs"].tobytes() == nr["regs"].tobytes())
for r in range(len(nr["reg_off"]) - 1):
    a = res1["regs"][res1["reg_off"][r]:res1["reg_off"][r + 1]]; e = nr["regs"][nr["reg_off"][r]:nr["reg_off"][r + 1]]
    if a.tobytes() != e.tobytes():
        print("S1 READ", r); print(" gpu", a); print(" ref", e); print("".join("ACGTN"[c] for c in rs.read(r))); break
