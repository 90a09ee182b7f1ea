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


def edge_problems():
    """Targeted known-answer cases: what a random generator reaches rarely or never.  Deterministic.
    -> (extend: list of (q, t, h0, w, zdrop, end_bonus)), (local: list of (q, t, xtra)), (global: list of (q, t, w))."""
    g = np.random.default_rng(4242)
    A = lambda n, c=0: np.full(n, c, np.uint8)
    rnd = lambda n: g.integers(0, 4, n).astype(np.uint8)
    cat = lambda *xs: np.concatenate([np.asarray(x, np.uint8) for x in xs])
    bands = [1, 3, 8, 16, 32, 64, 100, 128, 200, 256]
    ext = []
    def add(q, t, h0s=(1, 19, 60), ws=bands, modes=((100, 5), (0, 0), (20, 5), (10, 0))):
        for h0 in h0s:
            for w in ws:
                for zd, eb in modes:
                    ext.append((q, t, h0, w, zd, eb))
    add(A(50, 0), A(80, 1), h0s=(1, 3, 5, 30))                               # rows die at once / decay from h0 only
    good = rnd(60)
    add(cat(good, rnd(100)), cat(good, rnd(140)), ws=(16, 100, 256))         # good prefix then noise: z-drop and the score past the peak
    core = rnd(180)
    for gap in (7, 15, 40):                                                  # one long insertion / deletion: band growth, narrow bands cut it
        add(cat(core[:90], rnd(gap), core[90:]), core, h0s=(19,))
        add(core, cat(core[:90], rnd(gap), core[90:]), h0s=(19,))
        add(cat(core[:60], rnd(gap), core[60:120], core[120 + gap:]), core, h0s=(30,), ws=(8, 16, 64, 100))   # I block then D block
    qn = core.copy(); qn[::9] = 4; tn = core.copy(); tn[4::11] = 4
    add(qn, tn, h0s=(19,), ws=(16, 100)); add(A(40, 4), rnd(60), h0s=(5, 50), ws=(16, 100))                    # N in query / target / all-N query
    add(rnd(1), rnd(1), ws=(1, 100)); add(rnd(1), rnd(300), ws=(1, 100, 256)); add(rnd(250), rnd(1), ws=(1, 100))
    add(rnd(30), np.zeros(0, np.uint8), h0s=(19,), ws=(100,))                # empty target
    add(A(100, 0), cat(A(30, 0), A(50, 1), A(100, 0)), ws=(16, 64, 100, 256))      # columns die out, come back: stale eh[] entries
    add(np.tile([0, 1], 50), np.tile([0, 1], 70), ws=(16, 100, 256))         # tandem repeat: ties everywhere
    add(np.tile([0, 1, 2], 40), cat(np.tile([0, 1, 2], 20), [0, 1], np.tile([0, 1, 2], 30)), h0s=(19,), ws=(3, 16, 100))
    perfect = rnd(120)
    add(perfect, cat(perfect, rnd(40)), h0s=(1, 19, 130)); add(perfect, perfect[:100], h0s=(19,))   # to-end score vs local max, target ends early
    drift = cat(*[cat(core[20 * i:20 * i + 20], rnd(2)) for i in range(8)])
    add(drift, core, h0s=(19,), ws=(3, 8, 16, 32, 100))                      # steady drift off the diagonal: max_off
    loc = []
    for L in (248, 249, 250, 251):                                           # the u8 / i16 switch of mem_matesw (l_ms * a < 250)
        xtra = 0x40000 | 0x80000 | (0x10000 if L < 250 else 0) | 19
        q = rnd(L)
        for kind in range(10):
            if kind == 0: core_t = q.copy()
            elif kind == 1: core_t = cat(q[:100], rnd(7), q[100:])                       # >= 7-base insertion in the target
            elif kind == 2: core_t = cat(q[:100], q[109:])                               # 9-base deletion
            elif kind == 3: core_t = cat(q[:80], rnd(8), q[80:160], q[170:])             # I block and D block
            elif kind == 4: core_t = util.mutate(g, q, 0.1, 0.05)
            elif kind == 5: core_t = cat(q, rnd(50), q[:200])                            # a second, shorter copy: score2 / te2
            elif kind == 6: core_t = cat(q[:150], rnd(30), util.mutate(g, q, 0.03, 0.0)) # a weaker copy first, the better one later
            elif kind == 7: core_t = rnd(300)                                            # nothing to find: below the XSUBO threshold
            elif kind == 8: core_t = cat(A(300, 0)); q = A(L, 0)                         # homopolymer: every end position ties
            else: core_t = q[:40]                                                        # target shorter than the mate
            loc.append((q, cat(rnd(int(g.integers(0, 200))), core_t, rnd(int(g.integers(0, 200)))), xtra))
    glb = []
    for w in (3, 10, 50, 100):
        glb.append((cat(A(20, 0), A(20, 1)), cat(A(25, 0), A(20, 1)), w + 5))              # which A's leave: traceback tie preference
        glb.append((cat(A(25, 0), A(20, 1)), cat(A(20, 0), A(20, 1)), w + 5))
        glb.append((np.tile([0, 1], 30), np.tile([0, 1], 34), w + 8))                     # tandem repeat, deletion anywhere
        glb.append((np.tile([0, 1], 34), np.tile([0, 1], 30), w + 8))
        glb.append((cat(core[:50], A(6, 2), core[50:100]), cat(core[:50], core[50:100], A(6, 2)), w + 3))  # I and D of equal length
        glb.append((cat(core[:40], core[47:120]), cat(core[:80], core[87:120]), w + 3))
        glb.append((rnd(1), rnd(1), w)); glb.append((rnd(1), rnd(4), w + 3)); glb.append((rnd(5), rnd(1), w + 4))
        x = rnd(150); y = x.copy(); y[::5] = (y[::5] + 1) & 3
        glb.append((x, y, w)); glb.append((qn[:100], tn[:100], w))
    return ext, loc, glb


def ksw_edge_kats(R):
    ext, loc, glb = edge_problems()
    d = {}
    d["ext_q"], d["ext_qo"] = util.pack([e[0] for e in ext]); d["ext_t"], d["ext_to"] = util.pack([e[1] for e in ext])
    d["ext_par"] = np.array([e[2:] for e in ext], np.int32)  # h0, w, zdrop, end_bonus
    d["ext_out"] = np.array([[*R.ksw_extend2(q, t, w, eb, zd, h0)] for q, t, h0, w, zd, eb in ext], np.int32)
    d["loc_q"], d["loc_qo"] = util.pack([e[0] for e in loc]); d["loc_t"], d["loc_to"] = util.pack([e[1] for e in loc])
    d["loc_x"] = np.array([e[2] for e in loc], np.int32); d["loc_out"] = np.array([R.ksw_align2(*e) for e in loc], np.int32)
    d["glb_q"], d["glb_qo"] = util.pack([e[0] for e in glb]); d["glb_t"], d["glb_to"] = util.pack([e[1] for e in glb])
    d["glb_w"] = np.array([e[2] for e in glb], np.int32)
    res = [R.ksw_global2(*e) for e in glb]
    d["glb_score"] = np.array([r[0] for r in res], np.int32)
    d["glb_co"] = np.concatenate([[0], np.cumsum([len(r[1]) for r in res])]).astype(np.int64)
    d["glb_cig"] = np.concatenate([np.array(r[1], np.uint32) for r in res])
    np.savez_compressed(os.path.join(OUT, "ksw_edge.npz"), **d)
    print("ksw_edge:", len(ext), "extend,", len(loc), "local,", len(glb), "global,", os.path.getsize(os.path.join(OUT, "ksw_edge.npz")) >> 10, "KiB")


def introsort_cases():
    """Region arrays full of equal keys in the shapes that steer klib's introsort down its different paths (median of three, the
    16-element insertion-sort cut-off, the depth limit -> combsort).  `hash` carries a unique tag so the permutation can be read back."""
    from arachne_b200.lib import ALNREG
    g = np.random.default_rng(808)
    cases = []
    for n in [0, 1, 2, 3, 5, 16, 17, 18, 33, 64, 100, 257, 1000, 5000]:
        for shape in ("few", "equal", "sorted", "reversed", "organ", "saw", "random"):
            a = np.zeros(n, dtype=ALNREG)
            k = {"few": g.integers(0, 4, n), "equal": np.zeros(n, np.int64), "sorted": np.arange(n) // 3, "reversed": (n - np.arange(n)) // 3,
                 "organ": np.minimum(np.arange(n), n - 1 - np.arange(n)) // 2, "saw": np.arange(n) % 7, "random": g.integers(0, max(n, 1), n)}[shape]
            a["re"] = 1000 + k; a["score"] = 50 + (k % 11); a["rb"] = 900 + (k % 5); a["qb"] = k % 3
            a["hash"] = np.arange(n)
            cases.append(a)
    return cases


def introsort_golden(R):
    import ctypes as C
    cases = introsort_cases()
    L = R.lib
    out = {}
    tags = []
    for mode, fn in ((0, L.ks_introsort_mem_ars2), (1, L.ks_introsort_mem_ars)):
        fn.argtypes = [C.c_size_t, C.c_void_p]; fn.restype = None
        for a in cases:
            b = a.copy()
            fn(len(b), b.ctypes.data_as(C.c_void_p))
            tags.append(b["hash"].astype(np.int32))
    out["perm"], out["perm_off"] = np.concatenate(tags), np.concatenate([[0], np.cumsum([len(t) for t in tags])]).astype(np.int64)
    np.savez_compressed(os.path.join(OUT, "introsort.npz"), **out)
    print("introsort:", len(tags), "arrays,", os.path.getsize(os.path.join(OUT, "introsort.npz")) >> 10, "KiB")


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
    if len(sys.argv) > 1 and sys.argv[1] == "edge":   # only the targeted cases (the other fixtures stay as committed)
        ksw_edge_kats(R); introsort_golden(R); sys.exit(0)
    ksw_kats(R)
    ksw_edge_kats(R)
    introsort_golden(R)
    pipeline_golden(R, "tiny_rand", dict(total_len=60_000, n_contigs=2, seed=21), dict(n_barcodes=10, seed=31), damaged=True)
    pipeline_golden(R, "tiny_rep", dict(total_len=50_000, n_contigs=3, seed=22, repeat_frac=0.35, n_frac=0.003),
                    dict(n_barcodes=10, seed=32, sub=0.02, ins=0.002, dele=0.002), damaged=True)
