import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import util
from arachne_b200 import ArachneCuda
from oracle.refbwa import RefBwa
name = sys.argv[1]; os.makedirs("/tmp/w", exist_ok=True)
fa, ctg, rs = util.make_dataset(name, "/tmp/w", damaged=True)
R = RefBwa().load(fa); A = ArachneCuda().load_index(fa)
ref = R.batch(rs.seq, rs.off, rescue=True, cigar=False, threads=8)
for it in range(3):
    res = A.align_batch(rs.seq, rs.off, None, stages=3)
    print("full equal", np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes(), len(res["regs"]), len(ref["regs"]))
