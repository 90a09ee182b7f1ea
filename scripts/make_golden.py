#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference's own code (oracle/_ref = the vendored BWA C sources under
/root/reference compiled by oracle/Makefile, driven as src/gobwa/gobwa.go drives them).

The reference ships no golden vectors of its own, so these are "outputs of the reference itself run
here".  Run in the build container (where /root/reference exists):  python scripts/make_golden.py
The fixtures carry their own tiny reference + BWA index files, so every tier can replay them where /root/reference and
oracle/_ref are absent."""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import simulate as S
from oracle.refbwa import RefBwa
from tests import util

OUT = os.path.join(ROOT, "tests", "golden")


def ksw_kats(R):
    qs, ts, h0, w = util.extension_problems(2024, 1500)
    q, qo = util.pack(qs); t, to = util.pack(ts)
    ext = np.array([[*R.ksw_extend2(qs[i], ts[i], int(w[i]), eb, zd, int(h0[i]))] for (zd, eb) in ((100, 5), (0, 0), (20, 5)) for i in range(len(qs))], dtype=np.int32)
    g = np.random.default_rng(77)
    gq, gt, gw, gs, gc = [], [], [], [], []
    for _ in range(600):
        ql = int(g.integers(1, 200)); a = g.integers(0, 4, ql).astype(np.uint8)
        b = util.mutate(g, a, g.choice([0.0, 0.02, 0.1]), g.choice([0.0, 0.01, 0.05, 0.1]))
        if len(b) == 0:
            continue
        ww = int(max(abs(len(b) - ql) + 3, g.choice([3, 5, 10, 50, 100])))
        sc, cig = R.ksw_global2(a, b, ww)
        gq.append(a); gt.append(b); gw.append(ww); gs.append(sc); gc.append(np.array(cig, dtype=np.uint32))
    # local SW (mate-rescue kernel): u8 and i16 paths, with the flags mem_matesw passes (bwamem_pair.c:150-151)
    lq, lt, lx, lo = [], [], [], []
    for _ in range(600):
        ql = int(g.integers(30, 256)); a = g.integers(0, 4, ql).astype(np.uint8)
        core = util.mutate(g, a, g.choice([0.0, 0.03, 0.1]), g.choice([0.0, 0.02, 0.08]))
        b = np.concatenate([g.integers(0, 4, int(g.integers(0, 300))).astype(np.uint8), core, g.integers(0, 4, int(g.integers(0, 300))).astype(np.uint8)])
        xtra = 0x40000 | 0x80000 | (0x10000 if ql < 250 else 0) | 19  # KSW_XSUBO | KSW_XSTART | KSW_XBYTE | min_seed_len
        lq.append(a); lt.append(b); lx.append(xtra); lo.append(R.ksw_align2(a, b, xtra))
    d = {}
    d["ext_q"], d["ext_qo"] = q, qo; d["ext_t"], d["ext_to"] = t, to; d["ext_h0"], d["ext_w"] = h0, w; d["ext_out"] = ext
    d["glb_q"], d["glb_qo"] = util.pack(gq); d["glb_t"], d["glb_to"] = util.pack(gt); d["glb_w"] = np.array(gw, np.int32); d["glb_score"] = np.array(gs, np.int32)
    d["glb_cig"], d["glb_co"] = util.pack(gc) if gc else (np.zeros(0), np.zeros(1))
    d["glb_cig"] = np.concatenate(gc).astype(np.uint32)
    d["loc_q"], d["loc_qo"] = util.pack(lq); d["loc_t"], d["loc_to"] = util.pack(lt); d["loc_x"] = np.array(lx, np.int32); d["loc_out"] = np.array(lo, np.int32)
    np.savez_compressed(os.path.join(OUT, "ksw_kat.npz"), **d)
    print("ksw_kat:", len(qs) * 3, "extend,", len(gq), "global,", len(lq), "local")


def pipeline_golden(R, name, ref_kw, read_kw, damaged):
    with tempfile.TemporaryDirectory() as wd:
        ctg = S.make_reference(**ref_kw)
        fa = os.path.join(wd, name + ".fa")
        S.write_fasta(fa, ctg); R.build_index(fa); R.load(fa)
        rs = S.simulate_linked_reads(ctg, **read_kw)
        if damaged:
            rs = util.damage(rs, 0.3, 0.11, 5)
        d = {"seq": rs.seq, "off": rs.off, "bc_off": rs.bc_off, "l_pac": np.int64(R.l_pac)}
        for ext in ("bwt", "sa", "pac", "ann", "amb"):
            d["idx_" + ext] = np.frombuffer(open(fa + "." + ext, "rb").read(), dtype=np.uint8)
        prim, *L2, sl = R.bwt_meta()
        g = np.random.default_rng(3)
        ks = np.array([0, 1, prim - 1, prim, prim + 1, sl - 1, sl, 127, 128, 129, 255, 256] + [int(x) for x in g.integers(0, sl + 1, 500)], dtype=np.uint64)
        ls = np.minimum(ks + g.integers(0, 300, len(ks)).astype(np.uint64), np.uint64(sl))
        d["occ_k"], d["occ_l"] = ks, ls
        d["occ_out"] = np.array([sum(R.occ4(int(k), int(l)), []) for k, l in zip(ks, ls)], dtype=np.uint64)
        km1 = ks - np.uint64(1)
        d["occ_out_km1"] = np.array([sum(R.occ4(int(k), int(l)), []) for k, l in zip(km1, ls)], dtype=np.uint64)
        d["sa_out"] = np.array([R.sa(int(k)) for k in ks], dtype=np.uint64)
        # SMEM lists (bwt_smem1) from every 7th position of the first 40 reads
        sm_read, sm_x, sm_ret, sm_off, sm_iv = [], [], [], [0], []
        for r in range(min(40, 2 * rs.n_pairs)):
            q = rs.read(r)
            for x in range(0, len(q), 7):
                if q[x] > 3:
                    continue
                ret, iv = R.smem1(q, x)
                sm_read.append(r); sm_x.append(x); sm_ret.append(ret); sm_iv.append(iv); sm_off.append(sm_off[-1] + len(iv))
        d["smem_read"], d["smem_x"], d["smem_ret"], d["smem_off"] = map(np.array, (sm_read, sm_x, sm_ret, sm_off))
        d["smem_iv"] = np.concatenate(sm_iv) if sm_iv else np.zeros((0, 4), np.uint64)
        for tag, resc, cig in (("s1", False, False), ("s7", True, True)):
            ref = R.batch(rs.seq, rs.off, rescue=resc, cigar=cig, threads=4)
            d[tag + "_reg_off"] = ref["reg_off"]; d[tag + "_regs"] = ref["regs"].view(np.uint8)
            if cig:
                d[tag + "_aln"] = ref["aln"]; d[tag + "_cigar"] = ref["cigar"]
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, ":", rs.n_pairs, "pairs,", len(d["s7_regs"]) // 88, "regions,", os.path.getsize(os.path.join(OUT, name + ".npz")) >> 10, "KiB")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    R = RefBwa()
    ksw_kats(R)
    pipeline_golden(R, "tiny_rand", dict(total_len=60_000, n_contigs=2, seed=21), dict(n_barcodes=10, seed=31), damaged=True)
    pipeline_golden(R, "tiny_rep", dict(total_len=50_000, n_contigs=3, seed=22, repeat_frac=0.35, n_frac=0.003),
                    dict(n_barcodes=10, seed=32, sub=0.02, ins=0.002, dele=0.002), damaged=True)
