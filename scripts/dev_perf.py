"""Developer probe (not the bench): stage times of align_batch on a synthetic reference."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from arachne_b200 import simulate as S, ArachneCuda
from oracle.refbwa import RefBwa

L = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
NP = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
stages = int(sys.argv[3]) if len(sys.argv) > 3 else 1
os.makedirs("/tmp/w", exist_ok=True)
t = time.time()
ctg = S.make_reference(L, 1, seed=1)
fa = f"/tmp/w/ref{L}.fa"
if not os.path.exists(fa + ".bwt"):
    S.write_fasta(fa, ctg); RefBwa().build_index(fa)
print("index %.1fs" % (time.time() - t)); t = time.time()
rs = S.simulate_pairs_fast(ctg, NP, seed=5)
print("sim %.1fs" % (time.time() - t))
A = ArachneCuda().load_index(fa)
for it in range(3):
    t = time.time()
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=stages)
    dt = time.time() - t
    print("iter", it, "wall %.3fs" % dt, "pairs/s %.0f" % (NP / dt), {k: round(v, 2) for k, v in res["ms"].items()}, res["counters"], "regs", len(res["regs"]))
if len(sys.argv) > 4:
    R = RefBwa().load(fa); n = int(sys.argv[4])
    sub = rs.subset(0, n // 30)
    t = time.time(); ref = R.batch(sub.seq, sub.off, rescue=False, cigar=False, threads=os.cpu_count()); dt = time.time() - t
    print("cpu ref: %d pairs %.2fs -> %.0f pairs/s on %d threads" % (sub.n_pairs, dt, sub.n_pairs / dt, os.cpu_count()))
    r2 = A.align_batch(sub.seq, sub.off, stages=stages)
    print("parity:", np.array_equal(r2["reg_off"], ref["reg_off"]) and r2["regs"].tobytes() == ref["regs"].tobytes())
