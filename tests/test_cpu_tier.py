"""CPU tier (pytest -m "not gpu"): the oracle against the golden vectors (outputs of the reference's own code,
scripts/make_golden.py) and, where it is built, against oracle/_ref directly; the host-side logic (host simulation of
the thread-sequential device code, barcode sharding under gloo with world_size 2); and the C ABI surface."""
import ctypes as C
import os
import re
import shutil
import subprocess
import sys

import numpy as np
import pytest

from tests import util

ROOT = util.ROOT


@pytest.fixture(scope="module")
def O():
    return util.oracle_lib()


@pytest.fixture(scope="module")
def HS():
    return util.HostSim()


def _problems(d, p):
    qo, to = d[p + "_qo"], d[p + "_to"]
    return [(d[p + "_q"][qo[i]:qo[i + 1]], d[p + "_t"][to[i]:to[i + 1]]) for i in range(len(qo) - 1)]


# ---- oracle restatement (oracle/restate_ksw.c) vs golden known answers from the reference ----
def test_oracle_ksw_extend_golden(O):
    d = util.load_golden("ksw_kat")
    pr = _problems(d, "ext"); n = len(pr)
    for s, (zd, eb) in enumerate(((100, 5), (0, 0), (20, 5))):
        for i, (q, t) in enumerate(pr):
            got, _ = util.orc_extend(O, q, t, int(d["ext_w"][i]), eb, zd, int(d["ext_h0"][i]))
            assert got == tuple(d["ext_out"][s * n + i]), (s, i)


def test_oracle_ksw_global_golden(O):
    d = util.load_golden("ksw_kat")
    co = np.zeros(len(d["glb_w"]) + 1, dtype=np.int64)
    pr = _problems(d, "glb")
    pos = 0
    for i, (q, t) in enumerate(pr):
        sc, cig = util.orc_global(O, q, t, int(d["glb_w"][i]))
        assert sc == d["glb_score"][i], i
        assert cig == list(d["glb_cig"][pos:pos + len(cig)]), i
        pos += len(cig)
    assert pos == len(d["glb_cig"])


def test_oracle_ksw_local_golden(O):
    d = util.load_golden("ksw_kat")
    for i, (q, t) in enumerate(_problems(d, "loc")):
        assert util.orc_local(O, q, t, int(d["loc_x"][i])) == tuple(d["loc_out"][i]), i


def test_oracle_edge_kats_golden(O):
    """Targeted known answers of the reference (tests/golden/ksw_edge.npz, scripts/make_golden.py::edge_problems): rows that die
    at once, z-drop after a good prefix, 7..40-base insertions / deletions under bands 1..256, N bases, 1-base and empty operands, stale
    eh[] columns, tandem repeats; the u8 / i16 switch of the mate-rescue kernel at 249 / 250 bases with >= 7-base gap blocks, second
    copies and homopolymers; global-alignment traceback ties."""
    d = util.load_golden("ksw_edge")
    n_zdrop = 0
    for i, (q, t) in enumerate(_problems(d, "ext")):
        h0, w, zd, eb = (int(x) for x in d["ext_par"][i])
        got, _ = util.orc_extend(O, q, t, w, eb, zd, h0)
        assert got == tuple(d["ext_out"][i]), (i, got, tuple(d["ext_out"][i]))
        if zd == 10:
            n_zdrop += got != util.orc_extend(O, q, t, w, eb, 0, h0)[0]
    assert n_zdrop >= 10  # the z-drop cases really take the z-drop exit
    for i, (q, t) in enumerate(_problems(d, "loc")):
        assert util.orc_local(O, q, t, int(d["loc_x"][i])) == tuple(d["loc_out"][i]), i
    assert {len(q) for q, _ in _problems(d, "loc")} == {248, 249, 250, 251}
    co = d["glb_co"]
    for i, (q, t) in enumerate(_problems(d, "glb")):
        sc, cig = util.orc_global(O, q, t, int(d["glb_w"][i]))
        assert sc == d["glb_score"][i] and cig == list(d["glb_cig"][co[i]:co[i + 1]]), i


def test_oracle_vs_reference_live(O):
    """Fresh random problems against the reference's compiled sources (skipped where oracle/_ref is absent)."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    R = RefBwa()
    qs, ts, h0, w = util.extension_problems(int.from_bytes(os.urandom(2), "little"), 800)
    for i in range(len(qs)):
        for zd, eb in ((100, 5), (0, 0)):
            got, _ = util.orc_extend(O, qs[i], ts[i], int(w[i]), eb, zd, int(h0[i]))
            assert got == R.ksw_extend2(qs[i], ts[i], int(w[i]), eb, zd, int(h0[i])), i
        if len(ts[i]) and len(qs[i]):
            ww = max(abs(len(ts[i]) - len(qs[i])) + 3, int(w[i]))
            assert util.orc_global(O, qs[i], ts[i], ww) == R.ksw_global2(qs[i], ts[i], ww), i


# ---- host simulation of the device logic vs golden pipeline outputs ----
@pytest.mark.parametrize("name", ["tiny_rand", "tiny_rep"])
def test_hostsim_pipeline_golden(HS, workdir, name):
    from arachne_b200.lib import ALNREG
    d, prefix = util.load_golden(name, workdir)
    HS.load(prefix)
    for i in range(len(d["occ_k"])):
        assert HS.occ(d["occ_k"][i], d["occ_l"][i]) == list(d["occ_out"][i])
        assert HS.occ(int(d["occ_k"][i]) - 1, d["occ_l"][i]) == list(d["occ_out_km1"][i])
        assert HS.sa(d["occ_k"][i]) == d["sa_out"][i]
    seq, off = d["seq"], d["off"]
    n_pairs = (len(off) - 1) // 2
    s1 = d["s1_regs"].view(ALNREG); s7 = d["s7_regs"].view(ALNREG)
    assert d["s1_regs"].tobytes() != d["s7_regs"].tobytes(), "fixture does not exercise mate rescue"
    n_cig = 0
    for p in range(n_pairs):
        r1, r2 = seq[off[2 * p]:off[2 * p + 1]], seq[off[2 * p + 1]:off[2 * p + 2]]
        for k, q in ((2 * p, r1), (2 * p + 1, r2)):
            exp = s1[d["s1_reg_off"][k]:d["s1_reg_off"][k + 1]]
            assert HS.align1(q).tobytes() == exp.tobytes(), ("align1", k)
        o1, o2 = HS.pair(r1, r2)
        for k, q, got in ((2 * p, r1, o1), (2 * p + 1, r2, o2)):
            lo, hi = d["s7_reg_off"][k], d["s7_reg_off"][k + 1]
            assert got.tobytes() == s7[lo:hi].tobytes(), ("pair", k)
            if p % 3 == 0:  # CIGAR / NM / pos for every third pair keeps the tier fast
                for j in range(lo, hi):
                    a, cig, _mm = HS.reg2aln(q, s7[j:j + 1])
                    e = d["s7_aln"][j]
                    assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (e[0], e[1], e[2], e[3], e[4], e[6], e[7]), j
                    assert np.array_equal(cig, d["s7_cigar"][e[5]:e[5] + e[4]]), j
                    n_cig += 1
    assert n_cig > 50


def test_go_math_rand_pinned(HS):
    """Row a17's noise source: Go's math/rand (ALFG 607 / 273 seeded through the Lehmer LCG) as the DEVICE computes it (csrc/gorand.cuh in the
    host build: closed-form stateless outputs, the materialised state, the switch between them at any point) and as the oracle computes it,
    against tests/golden/gorand_float64.json -- 700 Float64 outputs for 11 seeds (0, negative, beyond 2^31, the int64 extremes) from
    scripts/gen_go_rng_table.py, whose generator is itself checked against Go's published outputs for NewSource(1)."""
    import json
    rows = json.load(open(os.path.join(util.GOLDEN, "gorand_float64.json")))
    O = util.oracle_lib()
    O.orc_gorand_float64.argtypes = [C.c_int64, C.c_int, C.POINTER(C.c_double)]
    assert len(rows) == 11
    for row in rows:
        seed, exp = int(row[0]), [float.fromhex(x) for x in row[1:]]
        out = (C.c_double * len(exp))(); O.orc_gorand_float64(seed, len(exp), out)
        assert list(out) == exp, ("oracle", seed)
        for switch_at in (0, 1, 100, 272, 273, 10**9):
            assert HS.gorand_float64(seed, len(exp), switch_at) == exp, ("device code", seed, switch_at)
    one = [float.fromhex(x) for x in rows[1][1:3]]
    assert rows[1][0] == 1 and one == [0.6046602879796196, 0.9405090880450124]  # the values every Go programmer has seen


def test_hostsim_introsort_tie_permutations(HS):
    """The order klib's introsort leaves EQUAL keys in is observable (which of two equal-score regions survives de-duplication, which chain
    is extended first).  The product's ks_introsort (csrc/common.cuh) against the permutations the reference's own ks_introsort_mem_ars2 /
    _mem_ars produced on tie-heavy arrays of 0..5000 regions (tests/golden/introsort.npz, scripts/make_golden.py::introsort_cases)."""
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    from make_golden import introsort_cases
    d = util.load_golden("introsort")
    cases = introsort_cases(); k = 0
    for mode in (0, 1):
        for a in cases:
            got = HS.introsort_regs(mode, a)["hash"]
            assert np.array_equal(got, d["perm"][d["perm_off"][k]:d["perm_off"][k + 1]]), (mode, len(a))
            k += 1
    assert k == len(d["perm_off"]) - 1


def test_hostsim_chain_level_live(HS, workdir):
    """Rows a7 / a8 at their own level: the chains csrc/chain.cuh forms and keeps -- every chain's pos, contig, seed count, weight and
    `kept` class, every seed, frac_rep, in the reference's order -- against the reference's own mem_chain / mem_chain_flt (needs
    oracle/_ref).  An exact-repeat reference gives reads dozens of chains (B-tree splits, equal keys, equal weights: the tie
    permutations of the weight sort), the diverged-repeat one the shadowing / drop rules."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import simulate as S
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    sets = [(fa, [rs.seq[rs.off[i]:rs.off[i + 1]] for i in range(0, 1200)])]
    xfa = os.path.join(workdir, "xrep.fa")
    xctg = util.exact_repeat_reference()
    if not os.path.exists(xfa + ".bwt"):
        S.write_fasta(xfa, xctg); RefBwa().build_index(xfa)
    xrs = S.simulate_linked_reads(xctg, n_barcodes=30, seed=77, sub=0.01, ins=0.001, dele=0.001)
    sets.append((xfa, [xrs.seq[xrs.off[i]:xrs.off[i + 1]] for i in range(min(1200, 2 * xrs.n_pairs))]))
    most, n_multi, n_dropped = 0, 0, 0
    for prefix, reads in sets:
        R = RefBwa().load(prefix); HS.load(prefix)
        for k, q in enumerate(reads):
            raw = None
            for flt in (False, True):
                ech, esd, efr = R.mem_chain(q, do_flt=flt)
                gch, gsd, gfr = HS.chain(q, do_flt=flt)
                assert np.array_equal(gch, ech) and np.array_equal(gsd, esd), (prefix, k, flt, gch[:4], ech[:4])
                if len(ech):
                    assert gfr == efr, (prefix, k, gfr, efr)
                if not flt:
                    raw = len(ech)
                else:
                    n_dropped += raw > len(ech)
            most = max(most, raw); n_multi += raw > 9
    assert most >= 20 and n_multi >= 20 and n_dropped >= 20, (most, n_multi, n_dropped)  # B-tree deeper than its root; the filter really drops


def test_hostsim_pipeline_live(HS, workdir):
    """The host build of the chaining / extension-driver / de-duplication / mate-rescue code against the reference's own C code on
    freshly simulated pairs (needs oracle/_ref): wrecked mates that only rescue can place, and pairs re-arranged so that their hits
    fall into each of mem_infer_dir's four orientation classes (same strand, swapped order, both)."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    R = RefBwa().load(fa); HS.load(fa)
    n_pairs = min(800, rs.n_pairs)
    reads = [rs.seq[rs.off[i]:rs.off[i + 1]].copy() for i in range(2 * n_pairs)]
    rc = lambda s: np.where(s[::-1] < 4, 3 - s[::-1], 4).astype(np.uint8)
    for p in range(n_pairs):
        k = p % 8
        if k == 5: reads[2 * p + 1] = rc(reads[2 * p + 1])                       # both mates on one strand
        elif k == 6: reads[2 * p], reads[2 * p + 1] = reads[2 * p + 1], reads[2 * p]  # order swapped
        elif k == 7: reads[2 * p] = rc(reads[2 * p])                             # the other same-strand arrangement
    seq, off = util.pack(reads)
    ref = R.batch(seq, off, rescue=True, cigar=True, threads=4)
    plain = R.batch(seq, off, rescue=False, cigar=False, threads=4)
    assert ref["regs"].tobytes() != plain["regs"].tobytes(), "dataset does not exercise rescue"
    from arachne_b200.lib import ALNREG
    exp = ref["regs"].view(ALNREG); ro = ref["reg_off"]
    exp1 = plain["regs"].view(ALNREG); ro1 = plain["reg_off"]
    n_rescued = n_cig = n_gapped = n_warp_rescued = n_warp_cig = 0
    from arachne_b200 import ArachneCuda
    opt = ArachneCuda().opt
    for p in range(n_pairs):
        for k in (2 * p, 2 * p + 1):
            assert HS.align1(reads[k]).tobytes() == exp1[ro1[k]:ro1[k + 1]].tobytes(), ("align1", k)
        o1, o2 = HS.pair(reads[2 * p], reads[2 * p + 1])
        if p % 2 == 0:  # the same rescue on the device's own warp primitives (csrc/warp_ctx.cuh as a warp of 32 fibers: striped-semantics local SW, leader writes)
            w1, w2 = HS.pair_warp(opt, reads[2 * p], reads[2 * p + 1], exp1[ro1[2 * p]:ro1[2 * p + 1]], exp1[ro1[2 * p + 1]:ro1[2 * p + 2]])
            assert w1.tobytes() == exp[ro[2 * p]:ro[2 * p + 1]].tobytes() and w2.tobytes() == exp[ro[2 * p + 1]:ro[2 * p + 2]].tobytes(), ("warp rescue", p)
            n_warp_rescued += int(len(w1) != ro1[2 * p + 1] - ro1[2 * p]) + int(len(w2) != ro1[2 * p + 2] - ro1[2 * p + 1])
        for k, got in ((2 * p, o1), (2 * p + 1, o2)):
            assert got.tobytes() == exp[ro[k]:ro[k + 1]].tobytes(), ("pair", k)
            n_rescued += int(ro[k + 1] - ro[k] != ro1[k + 1] - ro1[k])
            if p % 2 == 0:  # mem_reg2aln + the GetAlignments walk of every region of every second pair (csrc/cigar.cuh, the warp tier's logic)
                for j in range(ro[k], ro[k + 1]):
                    a, cig, _mm = HS.reg2aln(reads[k], exp[j:j + 1])
                    e = ref["aln"][j]
                    assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (e[0], e[1], e[2], e[3], e[4], e[6], e[7]), j
                    assert np.array_equal(cig, ref["cigar"][e[5]:e[5] + e[4]]), j
                    if p % 8 == 0:  # the same region on the device's warp tier (WarpCtx: ksw_global2 striped over 32 lanes, leader backtrack)
                        wa, wcig, wmm = HS.reg2aln_warp(opt, reads[k], exp[j:j + 1])
                        assert all(wa[f] == a[f] for f in a.dtype.names if f not in ("cigar_off", "mm_off")) and np.array_equal(wcig, cig) and np.array_equal(wmm, _mm), ("warp tier", j)
                        n_warp_cig += 1
                    n_cig += 1; n_gapped += len(cig) > 1 and bool(np.any((cig & 0xf) != 0) and np.any((cig & 0xf) == 1) | np.any((cig & 0xf) == 2))
    assert n_rescued >= 10 and n_cig >= 500 and n_gapped >= 20 and n_warp_rescued >= 5 and n_warp_cig >= 100, (n_rescued, n_cig, n_gapped, n_warp_rescued, n_warp_cig)


def _seed_test_reads(ctg, rs, n_sim=160):
    """Simulated reads + exact substrings of both strands (direct comparison running into the read's end), an N inside an exact match,
    all-N, shorter than a seed, tandem repeat, homopolymer: what tests/test_gpu_parity.py::test_seed_intervals_every_shortcut uses."""
    g = np.random.default_rng(11)
    ref = np.concatenate([np.asarray(c[1], dtype=np.uint8) for c in ctg])
    reads = [rs.seq[rs.off[r]:rs.off[r + 1]] for r in range(0, min(len(rs.off) - 1, n_sim))]
    for k in range(30):
        p = int(g.integers(0, len(ref) - 260)); L = int(g.choice([30, 100, 150, 250]))
        s = ref[p:p + L].copy()
        if s.max() > 3:
            continue
        if k % 3 == 1:
            s = (3 - s)[::-1].copy()
        if k % 3 == 2:
            s[int(g.integers(20, L - 5))] = 4
        reads.append(s)
    reads += [np.full(40, 4, dtype=np.uint8), np.zeros(12, dtype=np.uint8), np.array([0, 1, 2, 3] * 30, dtype=np.uint8), np.zeros(150, dtype=np.uint8),
              np.zeros(0, dtype=np.uint8), np.array([0, 1] * 100, dtype=np.uint8)]
    return reads


@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
def test_hostsim_shipped_seed_kernel(HS, workdir, name):
    """Rows a3-a5 on the code the GPU runs: csrc/seed_kernel.cuh (k_pack_reads, k_seed2, k_sort_intv) and the index builders of
    csrc/index_kernels.cuh, compiled for the host through tests/hostsim/cuda_lane_shim.h and run one lane at a time (k_seed2's lanes are
    independent automata), against the reference's own bwt_smem1 / bwt_seed_strategy1 composed into mem_collect_intv -- with the k-mer
    tables / forward windows / direct text extension on and off, different numbers of shared list slots and different lanes of the block
    (the lane-strided layouts).  The bwt_extend counter (bench.py takes the roofline's algorithmic bytes from a pass with the shortcuts off)
    stays within 1 % when a shortcut replaces the calls it counts."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import ArachneCuda
    fa, ctg, rs = util.make_dataset(name, workdir)
    R = RefBwa().load(fa); HS.load(fa)
    reads = _seed_test_reads(ctg, rs)
    seq, off = util.pack(reads)
    opt = ArachneCuda().opt
    si, K = HS.seed_index(fa)
    assert K >= 8
    try:
        base = None
        for use_kmer, direct, slist, lane in [(0, 0, -1, 0), (1, 0, -1, 37), (0, 1, -1, 255), (1, 1, -1, 0), (5, 1, -1, 1), (1 | (3 << 8), 1, 4, 100), (1 | (14 << 8), 1, 16, 31), (1, 1, 0, 200)]:
            io, iv, st = HS.seed2(si, opt, seq, off, use_kmer, direct, slist, lane)
            if base is None:
                base = (io, iv, st)
                for r, q in enumerate(reads):
                    got = [tuple(int(v) for v in a) for a in iv[io[r]:io[r + 1]]]
                    assert got == util.ref_collect_intv(R, q), (name, r)
                assert st[0] > 100 * len(reads) and st[0] <= st[1] <= 2 * st[0]
            else:
                assert np.array_equal(io, base[0]) and np.array_equal(iv, base[1]), (name, use_kmer, direct, slist, lane)
            assert abs(st[0] - base[2][0]) <= 0.01 * base[2][0], (st, base[2])
    finally:
        HS.seed_index_free(si)


def _check_cigar_tiers(HS, R, opt, reads, what):
    """Every region of `reads` through the shipped CIGAR tiers on the host: pos / contig / strand / NM / score / sub / CIGAR words against the
    reference's mem_reg2aln, and the GetAlignments walk's fields (matches, mismatches, indels, clips, mismatch loci on both axes) against
    the warp tier's implementation of the same walk (csrc/cigar.cuh in the host build).  -> counts."""
    from arachne_b200.lib import ALNREG
    seq, off = util.pack(reads)
    ref = R.batch(seq, off, rescue=True, cigar=True, threads=4)
    regs = ref["regs"].view(ALNREG); ro = ref["reg_off"]
    alns, cig, mm, list3, cnt = HS.cigar_tiers(opt, seq, off, ro, regs)
    left = set(int(g) for g in list3)
    reg_read = np.repeat(np.arange(len(reads)), np.diff(ro))
    n_gapped = 0
    for g in range(len(regs)):
        if g in left:
            continue
        a, e = alns[g], ref["aln"][g]
        assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (e[0], e[1], e[2], e[3], e[4], e[6], e[7]), (what, g)
        c = cig[a["cigar_off"]:a["cigar_off"] + a["n_cigar"]]
        assert np.array_equal(c, ref["cigar"][e[5]:e[5] + e[4]]), (what, g)
        n_gapped += bool(np.any(((c & 0xf) == 1) | ((c & 0xf) == 2)))
        w, wc, wm = HS.reg2aln(reads[reg_read[g]], regs[g:g + 1])
        for f in ("matches", "mismatches", "indels", "soft_clipped", "soft_clipped_length", "n_mm"):
            assert a[f] == w[f], (what, g, f)
        assert np.array_equal(mm[a["mm_off"]:a["mm_off"] + 2 * a["n_mm"]], wm), (what, g)
    # what the lane tiers hand over, through the shipped wide-band kernel itself (k_cigar as a warp of fibers: queue, pools, leader stores)
    lst = list3[:200]
    order = {"rep600k": 1, "len100": 2}.get(what, 0)  # the lanes' scheduling order differs between the read sets (0 is the one that exposed reg2aln's missing barrier, on len250)
    HS.lib.hs_warp_order(order)
    try:
        kal, kcig, kmm = HS.k_cigar(opt, seq, off, ro, regs, lst, warp=len(reads) % 4, block=what != "len250")
    finally:
        HS.lib.hs_warp_order(0)
    for g in (int(x) for x in lst):
        a, e = kal[g], ref["aln"][g]
        assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (e[0], e[1], e[2], e[3], e[4], e[6], e[7]), (what, "k_cigar", g)
        assert np.array_equal(kcig[a["cigar_off"]:a["cigar_off"] + a["n_cigar"]], ref["cigar"][e[5]:e[5] + e[4]]), (what, "k_cigar", g)
        w, wc, wm = HS.reg2aln(reads[reg_read[g]], regs[g:g + 1])
        for f in ("matches", "mismatches", "indels", "soft_clipped", "soft_clipped_length", "n_mm"):
            assert a[f] == w[f], (what, "k_cigar", g, f)
        assert np.array_equal(kmm[a["mm_off"]:a["mm_off"] + 2 * a["n_mm"]], wm), (what, "k_cigar", g)
    cnt["regions"] = len(regs); cnt["gapped"] = n_gapped; cnt["k_cigar"] = len(lst)
    return cnt


def test_hostsim_shipped_cigar_tiers(HS, workdir):
    """Row a13 (+ the walk of a15) on the code the GPU runs for nearly every region: csrc/cigar_kernel.cuh's k_cigar_fast and k_cigar_band
    compiled for the host through tests/hostsim/cuda_lane_shim.h and run one lane at a time (a lane owns its regions, ring column and
    traceback slice), staged as csrc/batch.inl stages them.  Reads of 100 / 150 / 250 bases with substitutions and indels, wrecked mates."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import ArachneCuda, simulate as S
    opt = ArachneCuda().opt
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    R = RefBwa().load(fa); HS.load(fa)
    tot = dict(regions=0, gapped=0, tier2=0, tier3=0, band_cells=0, k_cigar=0)
    sets = [("rep600k", [rs.seq[rs.off[i]:rs.off[i + 1]] for i in range(600)])]
    for L, seed in ((100, 5), (250, 6)):
        r2 = S.simulate_linked_reads(ctg, n_barcodes=6, seed=seed, read_len=L, sub=0.03, ins=0.01, dele=0.01)
        sets.append(("len%d" % L, [r2.seq[r2.off[i]:r2.off[i + 1]] for i in range(min(300, 2 * r2.n_pairs))]))
    for what, reads in sets:
        c = _check_cigar_tiers(HS, R, opt, reads, what)
        for k in tot:
            tot[k] += c[k]
    assert tot["tier3"] < tot["regions"] // 2 and tot["regions"] > 1500 and tot["gapped"] > 300 and tot["tier2"] > 300 and tot["band_cells"] > 100000 and tot["k_cigar"] > 200, tot


def _intervals_of(R, reads):
    """mem_collect_intv of every read from the reference's own code, flattened the way the stage kernels take them."""
    iv = [util.ref_collect_intv(R, q) for q in reads]
    io = np.zeros(len(reads) + 1, np.int64); io[1:] = np.cumsum([len(x) for x in iv])
    flat = np.array([t for x in iv for t in x], dtype=np.uint64).reshape(-1, 4)
    return io, flat


def test_hostsim_shipped_stage_kernels_in_launch_order(HS, workdir):
    """Rows a6-a12 on the code the GPU runs, kernel by kernel in csrc/batch.inl's launch order: k_expand, k_sa, k_chain (csrc/stage_kernels.cuh),
    the extension rounds k_ext_drive / k_ext_binbase / k_ext_scatter / k_ext_dp on the PACKED reference / k_ext_roll and k_ext_finish
    (csrc/ext_stage.cuh), k_init_loc, k_rescue and k_gather_regs2 (csrc/pair_kernels.cuh), compiled for the host and run as blocks of fibers
    (tests/hostsim/hostsim_stages.cpp).  Region lists byte-identical to the reference's mem_align1_core (no rescue) and GoBwaMemMateSW (with),
    from every warp slot of the block and with either column codec of the DP kernel; pairs re-arranged into all four orientation classes."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import ArachneCuda
    from arachne_b200.lib import ALNREG
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    R = RefBwa().load(fa); HS.load(fa)
    n_pairs = min(240, rs.n_pairs)
    reads = [rs.seq[rs.off[i]:rs.off[i + 1]].copy() for i in range(2 * n_pairs)]
    rc = lambda s: np.where(s[::-1] < 4, 3 - s[::-1], 4).astype(np.uint8)
    for p in range(n_pairs):
        k = p % 8
        if k == 5: reads[2 * p + 1] = rc(reads[2 * p + 1])
        elif k == 6: reads[2 * p], reads[2 * p + 1] = reads[2 * p + 1], reads[2 * p]
        elif k == 7: reads[2 * p] = rc(reads[2 * p])
    reads[6] = np.full(40, 4, np.uint8)        # all N: no interval, no chain, no region
    reads[9] = reads[9][:12].copy()            # shorter than a seed
    seq, off = util.pack(reads)
    opt = ArachneCuda().opt
    io, flat = _intervals_of(R, reads)
    plain = R.batch(seq, off, rescue=False, cigar=False, threads=4)
    full = R.batch(seq, off, rescue=True, cigar=False, threads=4)
    assert full["regs"].tobytes() != plain["regs"].tobytes(), "dataset does not exercise rescue"
    # each configuration under another scheduling order of the emulated lanes (ascending, descending, rotating, random: cuda_warp_shim.h)
    # ... and the warp-cooperative kernels either as one of their block's warps or as the whole block, its warps sharing the queue counters
    for order, (rescue, warp, wide, block, exp) in zip((0, 1, 2, 7), ((False, 0, False, False, plain), (False, 2, True, True, plain), (True, 1, False, False, full),
                                                                     (True, 3, True, True, full))):
        HS.lib.hs_warp_order(order)
        try:
            ro, regs, info = HS.stages(opt, seq, off, io, flat, rescue=rescue, warp=warp, wide=wide, block=block)
        finally:
            HS.lib.hs_warp_order(0)
        assert np.array_equal(ro, exp["reg_off"]), (rescue, warp, wide)
        assert regs.tobytes() == exp["regs"].view(ALNREG).tobytes(), (rescue, warp, wide)
        assert info["rounds"] >= 3 and info["problems"] >= 2 * n_pairs and info["total_sa"] >= 2 * n_pairs, info
        assert (info["rescued"] >= 10) == rescue, info


def test_hostsim_shipped_extension_dp_kernel(HS, O):
    """Row a10 on the code the GPU runs: k_ext_dp (csrc/ext_stage.cuh) -- one problem per lane, row-synchronous, descriptor pool dealt out by
    shuffles, first rows filled by the whole warp, DPX cells on packed column words, both column codecs -- compiled for the host and run as
    one warp of 32 fibers (tests/hostsim/cuda_warp_shim.h: every collective is a rendezvous of the 32 lanes), against the oracle's
    ksw_extend2 on the random + adversarial problems of the GPU test, and against the REFERENCE's answers on the targeted edge cases of
    tests/golden/ksw_edge.npz (rows dying at once, z-drop, long gaps under bands 1..256, N bases, 1-base / empty operands, stale columns)."""
    from arachne_b200 import ArachneCuda
    opt = ArachneCuda().opt
    qs, ts, h0, w = util.extension_problems(101, 1500)
    for codec in ("byte", "wide"):
        if codec == "byte":
            hh = np.minimum(h0, 255 - np.array([len(x) for x in qs])).astype(np.int32)
            keep = [i for i in range(len(qs)) if hh[i] > 0]
        else:
            hh = h0; keep = list(range(len(qs)))
        q2 = [qs[i] for i in keep]; t2 = [ts[i] for i in keep]; h2 = hh[keep]; w2 = w[keep]
        for k, (zdrop, eb) in enumerate(((100, 5), (0, 0), (20, 5))):
            out, cells = HS.ext_dp(opt, q2, t2, h2, w2, end_bonus=eb, zdrop=zdrop, codec=1 if codec == "wide" else 0, warp=k + (8 if k == 2 else 0))  # k == 2: all four warps of the block
            tot = 0
            for i in range(len(q2)):
                exp, c = util.orc_extend(O, q2[i], t2[i], int(w2[i]), eb, zdrop, int(h2[i]))
                assert tuple(out[i]) == exp, (codec, zdrop, i, tuple(out[i]), exp)
                tot += c
            assert cells == tot   # the GCUPS numerator: the reference's own in-band count
    d = util.load_golden("ksw_edge")
    pr = _problems(d, "ext"); par = d["ext_par"]
    for zd, eb in sorted(set((int(p[2]), int(p[3])) for p in par)):
        sel = [i for i in range(len(pr)) if (int(par[i][2]), int(par[i][3])) == (zd, eb)]
        for codec in (0, 1):
            out, _ = HS.ext_dp(opt, [pr[i][0] for i in sel], [pr[i][1] for i in sel], par[sel, 0], par[sel, 1], end_bonus=eb, zdrop=zd, codec=codec, warp=3)
            for k, i in enumerate(sel):
                assert tuple(out[k]) == tuple(d["ext_out"][i]), ("edge", i, zd, eb, codec, tuple(out[k]), tuple(d["ext_out"][i]))


def test_hostsim_stage_entry_kernels(HS, O, workdir):
    """The kernels behind the stage entry points arachne_ksw_global_batch and arachne_occ_batch (k_ksw_global: ksw_global2 with CIGAR on the
    warp context; k_occ: bwt_2occ4) on the host: the REFERENCE's scores and CIGARs on the global-alignment edge cases of
    tests/golden/ksw_edge.npz, the oracle's on random problems (one warp and the whole block of four), occ counts of a golden index."""
    from arachne_b200 import ArachneCuda
    opt = ArachneCuda().opt
    d = util.load_golden("ksw_edge")
    pr = _problems(d, "glb"); co = d["glb_co"]
    keep = [i for i in range(len(pr)) if 1 <= len(pr[i][0]) <= 255]
    assert len(keep) >= 30
    for block in (False, True):
        sc, cigs = HS.k_ksw_global(opt, [pr[i][0] for i in keep], [pr[i][1] for i in keep], d["glb_w"][keep], warp=1, block=block)
        for k, i in enumerate(keep):
            assert sc[k] == d["glb_score"][i] and cigs[k] == list(d["glb_cig"][co[i]:co[i + 1]]), ("edge", i, block)
    qs, ts, _h0, w = util.extension_problems(77, 240)
    sel = [i for i in range(len(qs)) if len(ts[i]) > 0]
    ww = np.array([max(abs(len(ts[i]) - len(qs[i])) + 3, int(w[i])) for i in sel], np.int32)
    sc, cigs = HS.k_ksw_global(opt, [qs[i] for i in sel], [ts[i] for i in sel], ww, warp=2, block=True)
    for k, i in enumerate(sel):
        esc, ecig = util.orc_global(O, qs[i], ts[i], int(ww[k]))
        assert sc[k] == esc and cigs[k] == ecig, ("random", i)
    dg, prefix = util.load_golden("tiny_rep", workdir)
    HS.load(prefix)
    g = np.random.default_rng(4)
    sl = 8 * (len(dg["idx_pac"]) - 1) - 8   # a little under 2 * l_pac, the BWT's length
    ks = g.integers(0, sl, 300).astype(np.uint64); ls = np.minimum(ks + g.integers(0, 500, 300).astype(np.uint64), sl).astype(np.uint64)
    got = HS.k_occ(ks, ls)
    for i in range(len(ks)):
        assert [int(x) for x in got[i]] == HS.occ(int(ks[i]), int(ls[i])), i


def test_hostsim_allocation_free_entry_kernels(HS, O):
    """arachne_ksw_extend_static (include/arachne_cuda.h) without a GPU: its launch sequence -- control block cleared, k_xs_count, k_xs_base
    (block of 256), k_xs_queue, k_ext_dp<false, true>, k_xs_finish (csrc/xs_kernels.cuh, csrc/ext_stage.cuh) -- compiled for the host and run
    as blocks of fibers over a scratch that persists between calls like the static device array.  The CPU twin of tests/test_zz_no_alloc.py,
    against the oracle's ksw_extend2: two calls in a row (the second smaller, other parameters), refusals reported through the status word
    with the lowest refused index, the other problems still answered."""
    from arachne_b200 import ArachneCuda
    opt = ArachneCuda().opt
    qs, ts, h0, w = util.extension_problems(211, 1200)
    assert any(int(h0[i]) + len(qs[i]) > 255 for i in range(len(qs)))

    def expect(qs, ts, h0, w, eb, zd):
        rows, tot = [], 0
        for i in range(len(qs)):
            e, c = util.orc_extend(O, qs[i], ts[i], int(w[i]), eb, zd, int(h0[i]))
            rows.append(e); tot += c
        return np.array(rows, np.int32).reshape(-1, 6), tot

    exp, exp_cells = expect(qs, ts, h0, w, 5, 100)
    out, status, cells = HS.xs_static(opt, qs, ts, h0, w, 5, 100, warp=1)
    assert status == 0 and np.array_equal(out, exp) and cells == exp_cells
    # straight after: fewer and other problems, other parameters; nothing of call 1 may survive in the scratch
    sel = list(range(len(qs) - 1, 0, -3))[:300]
    qs2 = [qs[i] for i in sel]; ts2 = [ts[i][: max(0, len(ts[i]) - 7)] for i in sel]; h2 = h0[sel]; w2 = np.maximum(1, w[sel] // 2).astype(np.int32)
    exp2, exp_cells2 = expect(qs2, ts2, h2, w2, 0, 20)
    out2, status2, cells2 = HS.xs_static(opt, qs2, ts2, h2, w2, 0, 20, max_qlen=255, block=True)  # the whole block of four warps on one queue
    assert status2 == 0 and np.array_equal(out2, exp2) and cells2 == exp_cells2
    # refusals: h0 = 0 (ksw.c:385 asserts h0 > 0), a query longer than the declared bound, a score beyond the 13-bit codec; the lowest index is reported
    h_bad = h0.copy(); h_bad[17] = 0; h_bad[400] = 8191
    lens = np.array([len(x) for x in qs]); mq = int(np.sort(lens)[len(lens) // 2])
    too_long = [i for i in range(len(qs)) if lens[i] > mq]
    assert too_long
    out3, status3, _ = HS.xs_static(opt, qs, ts, h_bad, w, 5, 100, max_qlen=mq, warp=3)
    bad = sorted(set([17, 400] + too_long))
    assert status3 == bad[0] + 1
    for i in bad:
        assert tuple(out3[i]) == (-1,) * 6, i
    keep = [i for i in range(len(qs)) if i not in bad]
    assert np.array_equal(out3[keep], exp[keep])
    # an empty call, and empty targets (ksw.c:409 never enters its loop: the score stays h0)
    out4, status4, cells4 = HS.xs_static(opt, [], [], np.zeros(0, np.int32), np.zeros(0, np.int32))
    assert status4 == 0 and len(out4) == 0 and cells4 == 0
    out5, status5, _ = HS.xs_static(opt, qs[:40], [np.zeros(0, np.uint8)] * 40, h0[:40], w[:40])
    exp5, _ = expect(qs[:40], [np.zeros(0, np.uint8)] * 40, h0[:40], w[:40], 5, 100)
    assert status5 == 0 and np.array_equal(out5, exp5)


@pytest.mark.parametrize("name,rfa", [("rand1m", 1), ("rep600k", 1), ("rep600k", 0), ("stlfr", 1)])
def test_hostsim_shipped_placement_kernels(HS, workdir, name, rfa):
    """Rows a15-a20 on the code the GPU runs: csrc/place.cuh (candidates, tagBestAlignments with the device's Go PRNG, molecules, the
    optimizer sweep and MAPQ) and markDuplicates of csrc/records.cuh, compiled for the host and staged as csrc/place_host.inl stages them;
    k_place_rfa -- one warp per barcode, 23 block barriers, warp shuffles -- runs as 32 fibers whose barriers are rendezvous
    (tests/hostsim/hostsim_place.cpp).  Against the CPU restatement of the Go code with the tolerances of the GPU test (the Go half stays
    unpinned: both sides are restatements).  "stlfr": hundreds of pairs and dozens of molecules per barcode."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import ArachneCuda, simulate as S
    from arachne_b200.lib import ALNREG
    A = ArachneCuda(); opt = A.opt
    if name == "stlfr":
        fa, ctg, _ = util.make_dataset("rep600k", workdir, damaged=True)
        rs = S.simulate_linked_reads(ctg, n_barcodes=3, seed=41, style="stlfr", dash_suffix=True, mol_per_bc=14.0, pairs_per_mol=9.0, sub=0.01)
    else:
        fa, ctg, rs = util.make_dataset(name, workdir, damaged=True)
    R = RefBwa().load(fa); HS.load(fa)
    reads = [rs.seq[rs.off[i]:rs.off[i + 1]] for i in range(2 * rs.n_pairs)]
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    regs = ref["regs"].view(ALNREG)
    alns = HS.all_alns(opt, reads, rs.seq, rs.off, ref["reg_off"], regs)
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    run_rfa = np.full(len(rs.bc_off) - 1, rfa, np.uint8)
    res = HS.place(rs.off, ref["reg_off"], regs, alns, rs.bc_off, run_rfa, seeds)
    oc, ocands, olmp = util.PlaceOracle(fa).place(rs.seq, rs.off, rs.bc_off, run_rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    bad = util.place_mismatches(res, oc, ocands, olmp)
    assert not bad, (name, rfa, bad)
    act = ocands[ocands["active"] == 1]
    assert len(act) == 2 * rs.n_pairs
    if rfa:
        assert ocands["molecule_id"].max() >= 1 and act["active_molecule"].sum() > 0
    if name == "stlfr":
        assert np.diff(rs.bc_off).max() > 100


@pytest.mark.parametrize("rfa", [1, 0])
def test_hostsim_shipped_record_kernels(HS, workdir, rfa):
    """Row f1 on the code the GPU runs: markDuplicates, CheckSplitReads and the fields of every SAM record (csrc/records.cuh: k_rec_active,
    k_rec_dup_keys / _mark, k_rec_split, k_rec_emit) after the shipped placement kernels, all on the host (tests/hostsim/hostsim_place.cpp),
    against the oracle's restatement field by field -- on the GPU test's dataset (exact duplicate pairs, chimeric reads, noise reads, distant
    mates: duplicates, split alignments, unmapped / demoted / improper records)."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200 import ArachneCuda
    from arachne_b200.lib import ALNREG
    from tests.test_gpu_parity import _records_dataset
    A = ArachneCuda(); opt = A.opt
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    R = RefBwa().load(fa); HS.load(fa)
    rs = _records_dataset(ctg, 77)
    nb = len(rs.bc_off) - 1
    reads = [rs.seq[rs.off[i]:rs.off[i + 1]] for i in range(2 * rs.n_pairs)]
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    run_rfa = np.full(nb, rfa, np.uint8)
    attach = (np.arange(nb) % 5 != 0).astype(np.uint8)
    if rfa:
        attach[:] = 1
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    regs = ref["regs"].view(ALNREG)
    alns = HS.all_alns(opt, reads, rs.seq, rs.off, ref["reg_off"], regs)
    e = ref["aln"]
    assert np.array_equal(alns["n_cigar"], e[:, 4])
    alns["cigar_off"] = e[:, 5]                     # one pool for every region: the reference's (its words equal ours, checked elsewhere)
    res = HS.place(rs.off, ref["reg_off"], regs, alns, rs.bc_off, run_rfa, seeds, records=(attach, ref["cigar"]))
    oc, ocands, olmp, orecs = util.PlaceOracle(fa).place_records(rs.seq, rs.off, rs.bc_off, run_rfa, attach, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    bad = util.place_mismatches(res, oc, ocands, olmp)
    assert not bad, bad
    g = res["records"]
    assert len(g) == len(orecs) and np.array_equal(res["rec_off"][1:] - res["rec_off"][:-1], np.bincount(orecs["read"], minlength=2 * rs.n_pairs))
    for f in g.dtype.names:
        if f in ("_pad", "dm"):
            continue
        d = np.nonzero(g[f] != orecs[f])[0]
        assert len(d) == 0, (f, len(d), int(d[0]), g[d[0]], orecs[d[0]])
    assert np.allclose(g["dm"], orecs["dm"], rtol=1e-9, atol=1e-12)
    fl = g["flag"]
    assert (g["primary"] == 0).sum() > 20 and (fl & 0x400).sum() > 10 and (fl & 0x4).sum() > 20 and (fl & 0x8).sum() > 20 and ((fl & 0x2) == 0).sum() > 50
    assert (g["sa_reg"] >= 0).sum() > 40 and (g["has_xc"] == 1).sum() > 50 and (g["tlen"] != 0).sum() > 500


def test_hostsim_seed_intervals_live(HS, workdir):
    """mem_collect_intv restated from bwt_smem1 / bwt_seed_strategy1 calls on the reference's code (needs oracle/_ref)."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    d, prefix = util.load_golden("tiny_rep", workdir)
    R = RefBwa().load(prefix); HS.load(prefix)
    seq, off = d["seq"], d["off"]
    for r in range(0, 120):
        q = seq[off[r]:off[r + 1]]
        got = HS.seed(q)
        # mem_collect_intv (bwamem.c:114-162) composed from the reference's own primitives
        L = len(q); exp = []
        x = 0
        while x < L:
            if q[x] < 4:
                x, iv = R.smem1(q, x)
                exp += [tuple(int(v) for v in a) for a in iv if (int(a[3]) & 0xffffffff) - (int(a[3]) >> 32) >= 19]
            else:
                x += 1
        first = list(exp)
        for a in first:
            s, e = a[3] >> 32, a[3] & 0xffffffff
            if e - s < 28 or a[2] > 10:
                continue
            _, iv = R.smem1(q, (s + e) >> 1, a[2] + 1)
            exp += [tuple(int(v) for v in b) for b in iv if (int(b[3]) & 0xffffffff) - (int(b[3]) >> 32) >= 19]
        x = 0
        while x < L:
            if q[x] < 4:
                x, m = R.seed_strategy1(q, x, 19, 20)
                if m[2] > 0:
                    exp.append(tuple(m))
            else:
                x += 1
        exp.sort(key=lambda a: a[3])
        assert sorted(map(tuple, got.tolist()), key=lambda a: a[3]) == exp or [tuple(a) for a in got.tolist()] == exp, r
        assert [a[3] for a in got.tolist()] == [a[3] for a in exp], r


# ---- the gobwa-shaped layer's own logic (csrc/compat.cu is host code) over a mock batch backend served by the reference's C code ----
def test_compat_layer_batching_logic_on_mock_backend(workdir):
    """64 threads make per-pair GoBwaMemMateSW + per-candidate GoBwaSmithWaterman calls (the reference's worker goroutines): every region array,
    mem_aln_t field and CIGAR word equals the reference's own per-pair results; calls were batched and batches overlapped (two leaders at
    once); a region stays addressable until its block is released and not after; a failing batch gives every caller of that batch the error
    and leaves no block behind.  The same layer runs over the GPU in tests/test_gpu_parity.py; scripts/host_sanitize.sh runs THIS under TSan."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    from arachne_b200.lib import Compat, ArachneError
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    R = RefBwa().load(fa)
    n = min(rs.n_pairs, 1200)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    B = util.CompatMockBackend(fa)
    K = Compat(B, max_batch_pairs=48, max_wait_us=1500)
    try:
        def one(p):
            r1 = np.ascontiguousarray(rs.seq[rs.off[2 * p]:rs.off[2 * p + 1]]); r2 = np.ascontiguousarray(rs.seq[rs.off[2 * p + 1]:rs.off[2 * p + 2]])
            g1, g2, a1, a2, owner = K.mem_mate_sw(r1, r2, 25)
            alns = [K.reg2aln(base + 88 * k) for base, g in ((a1, g1), (a2, g2)) for k in range(len(g))]
            return p, g1, g2, alns, owner, (a1 or a2)
        with ThreadPoolExecutor(64) as ex:
            out = list(ex.map(one, range(n)))
        st = K.stats()
        assert st["n_calls"] == n and st["n_batches"] < n // 4 and 8 < st["max_batch"] <= 48 and st["n_live_blocks"] == n, st
        assert B.lib.mock_max_batches_in_flight() >= 2  # while one leader's batch runs, the next leader is already collecting / running
        e = ref["aln"]
        for p, g1, g2, alns, owner, _ in out:
            lo, mid, hi = (int(ref["reg_off"][2 * p + k]) for k in range(3))
            assert g1.tobytes() == ref["regs"][lo:mid].tobytes() and g2.tobytes() == ref["regs"][mid:hi].tobytes(), p
            assert len(alns) == hi - lo
            for k, a in enumerate(alns):
                x = e[lo + k]
                assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (x[0], x[1], x[2], x[3], x[4], x[6], x[7]), (p, k)
                assert np.array_equal(a["cigar"], ref["cigar"][x[5]:x[5] + x[4]]), (p, k)
        # lifetime: still addressable after everything else ran, gone once released; an address inside a block but off a record boundary is refused
        p, _, _, alns, owner, addr = next(o for o in out if o[5])
        again = K.reg2aln(addr)
        assert again["pos"] == alns[0]["pos"] and np.array_equal(again["cigar"], alns[0]["cigar"])
        with pytest.raises(ArachneError):
            K.reg2aln(addr + 8)
        K.release(owner)
        with pytest.raises(ArachneError):
            K.reg2aln(addr)
        for o in out:
            if o[4] is not owner:
                K.release(o[4])
        assert K.stats()["n_live_blocks"] == 0
        # a failing batch: every caller of it gets the error, nobody a block
        B.lib.mock_fail_next_batches(1)
        def attempt(p):
            try:
                one(p); return True
            except ArachneError:
                return False
        with ThreadPoolExecutor(16) as ex:
            ok = list(ex.map(attempt, range(16)))
        st2 = K.stats()
        assert not all(ok) and st2["n_live_blocks"] == sum(ok), (ok, st2)
        # GetSeq = bns_fetch_seq with its clamp
        got, b, en, rid = K.fetch_seq(100, 250, 400)
        assert (b, en, rid) == (100, 400, 0) and len(got) == 300 and got.max() <= 3
    finally:
        K.close(); B.close()


# ---- C ABI surface ----
# ---- index construction (csrc/index_build.cuh on the host backend) vs the reference's own `bwa index` output ----
@pytest.mark.parametrize("name", ["tiny_rand", "tiny_rep"])
def test_hostsim_index_build_golden(name):
    """The fixture carries the .pac/.bwt/.sa files the reference's bwa_idx_build wrote: ours must be byte-identical."""
    d = util.load_golden(name)
    for cap in (0, 4000):
        bwt, sa, st = util.hostsim_index_build(d["idx_pac"].tobytes(), int(d["l_pac"]), cap)
        assert bwt == d["idx_bwt"].tobytes(), (name, cap, "bwt")
        assert sa == d["idx_sa"].tobytes(), (name, cap, "sa")
        assert st["n_batches"] >= (1 if cap == 0 else 2)


def test_hostsim_index_build_live(tmp_path):
    """Stress references against oracle/_ref's bwa_idx_build run here; also pins simulate.write_bns to the reference's pack step."""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    deep = 0
    for name, ctg, cap in util.index_build_cases():
        fa = str(tmp_path / (name + ".fa"))
        util.S.write_fasta(fa, ctg)
        RefBwa().build_index(fa)
        l_pac = util.S.write_bns(str(tmp_path / (name + ".mine")), ctg)
        for ext in (".pac", ".ann", ".amb"):
            assert open(fa + ext, "rb").read() == open(str(tmp_path / (name + ".mine")) + ext, "rb").read(), (name, ext)
        bwt, sa, st = util.hostsim_index_build(open(fa + ".pac", "rb").read(), l_pac, cap)
        assert bwt == open(fa + ".bwt", "rb").read(), (name, "bwt")
        assert sa == open(fa + ".sa", "rb").read(), (name, "sa")
        deep = max(deep, st["rounds"])
    assert deep >= 10  # the homopolymer / tandem cases really went through many refinement rounds


def test_abi_exports_every_declared_symbol():
    from arachne_b200.lib import lib_path
    assert os.path.exists(lib_path()), "libarachne_cuda.so missing: run __graft_entry__.build()"
    lib = C.CDLL(lib_path())
    declared = []
    for h in sorted(os.listdir(os.path.join(ROOT, "include"))):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        declared += re.findall(r"\b((?:arachne|mem|bwa|bns|bwt)_[a-z0-9_]+)\s*\(", src)
    declared = sorted(set(declared))
    assert len(declared) >= 15
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing


def test_abi_defaults_and_errors_without_device():
    """Host-only entry points work without a GPU; compute entry points fail loudly (no CPU fallback)."""
    from arachne_b200 import ArachneCuda, ArachneError
    A = ArachneCuda()
    o = A.opt
    assert (o.a, o.b, o.o_del, o.e_del, o.o_ins, o.e_ins, o.w, o.zdrop, o.T, o.min_seed_len, o.max_occ, o.max_chain_gap) == (1, 4, 6, 1, 6, 1, 100, 100, 30, 19, 500, 10000)
    assert (o.pen_clip5, o.pen_clip3, o.max_mem_intv, o.split_width, o.max_matesw) == (5, 5, 20, 10, 50)
    assert abs(o.split_factor - 1.5) < 1e-6 and abs(o.mask_level_redun - 0.95) < 1e-6 and abs(o.drop_ratio - 0.5) < 1e-6
    assert list(o.mat) == [1, -4, -4, -4, -1, -4, 1, -4, -4, -1, -4, -4, 1, -4, -1, -4, -4, -4, 1, -1, -1, -1, -1, -1, -1]
    from oracle.refbwa import RefBwa, available
    if available():
        assert bytes(o)[:C.sizeof(o)] == RefBwa().opt_default()[:C.sizeof(o)], "arachne_mem_opt_t differs from mem_opt_init()"
    if A.device_count() == 0:
        q = np.zeros(10, np.uint8); qo = np.array([0, 10], np.int64)
        with pytest.raises(ArachneError):
            A.ksw_extend(q, qo, q, qo, np.array([10], np.int32), np.array([10], np.int32))
        with pytest.raises(ArachneError):
            A.load_index("/nonexistent/prefix")


# ---- barcode sharding: host logic, world_size 2 over gloo ----
def test_shard_barcodes_properties():
    from arachne_b200.shard import shard_barcodes
    g = np.random.default_rng(0)
    for _ in range(50):
        n = int(g.integers(0, 40)); sizes = g.integers(1, 400, n)
        bc = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        for w in (1, 2, 3, 4, 8):
            sh = shard_barcodes(bc, w)
            assert len(sh) == w and sh[0][0] == 0 and sh[-1][1] == n
            assert all(sh[i][1] == sh[i + 1][0] for i in range(w - 1)) and all(a <= b for a, b in sh)
            if n >= 4 * w:
                loads = [int(bc[b] - bc[a]) for a, b in sh]
                assert max(loads) - min(loads) <= 2 * int(sizes.max())


_GLOO_WORKER = r"""
import os, sys, pickle
sys.path.insert(0, {root!r})
import numpy as np, torch.distributed as dist
from tests import util
from arachne_b200.shard import shard_barcodes, shard_view, merge_results
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
d, prefix = util.load_golden("tiny_rand", {wd!r} + "/r%d" % rank) if os.makedirs({wd!r} + "/r%d" % rank, exist_ok=True) is None else None
HS = util.HostSim().load(prefix)
lo, hi = shard_barcodes(d["bc_off"], world)[rank]
seq, off, bc = shard_view(d["seq"], d["off"], d["bc_off"], lo, hi)
regs, reg_off = [], [0]
for p in range((len(off) - 1) // 2):
    o1, o2 = HS.pair(seq[off[2*p]:off[2*p+1]], seq[off[2*p+1]:off[2*p+2]])
    for o in (o1, o2):
        regs.append(o); reg_off.append(reg_off[-1] + len(o))
part = dict(reg_off=np.array(reg_off, dtype=np.int64), regs=np.concatenate(regs) if regs else np.zeros(0, dtype=util.ALNREG_DT()))
out = [None] * world
dist.gather_object(part, out if rank == 0 else None, dst=0)
if rank == 0:
    m = merge_results(out)
    from arachne_b200.lib import ALNREG
    ok = np.array_equal(m["reg_off"], d["s7_reg_off"]) and m["regs"].tobytes() == d["s7_regs"].tobytes()
    print("SHARD_MERGE_OK" if ok else "SHARD_MERGE_MISMATCH", flush=True)
dist.destroy_process_group()
"""


def test_sharded_two_ranks_gloo_matches_unsharded(workdir):
    util.HostSim()  # build once before the ranks race on make
    script = os.path.join(workdir, "gloo_worker.py")
    open(script, "w").write(_GLOO_WORKER.format(root=ROOT, wd=workdir))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(29500 + os.getpid() % 2000), script], capture_output=True, text=True, timeout=600)
    assert "SHARD_MERGE_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])


# ---- row f2: barcode-grouped FASTQ ingest (csrc/ingest.cu) against a line-by-line Python restatement of src/fastqreader/reader.go ----
def _reader_go(r1_lines, r2_lines):
    """ReadOneLine + ParseHeader + ReadBarcodeSet (reader.go:95-291) and worthRunningRFA (aligner.go:1018-1030) on in-memory lines."""
    import re
    recs = []
    for i in range(0, min(len(r1_lines), len(r2_lines)) - 3, 4):
        h = r1_lines[i][1:]                           # R1_line[1:], newline included
        # Go's regexp (RE2) \s is [\t\n\f\r ]; strings.Fields splits at unicode.IsSpace, which adds \v in the ASCII range
        bx = re.search(r"BX:Z:([^\t\n\f\r ]+)[\t\n\f\r ]", h)
        fields = lambda t: [x for x in re.split(r"[\t\n\v\f\r ]+", t) if x]
        if bx:
            _id = fields(h)[0]; name = _id[:len(_id) - 2] if len(_id) >= 2 else ""; bc = bx.group(1)
            vx = re.search(r"VX:i:([01])[\t\n\f\r ]", h); valid = bool(vx and vx.group(1) == "1")
        else:
            name, bc, valid = "", "", False
        f = fields(h[:-1])
        rg = f[-1] if len(f) >= 2 else ""
        recs.append((name, bc, valid, rg, r1_lines[i + 1][:-1], r1_lines[i + 3][:-1], r2_lines[i + 1][:-1], r2_lines[i + 3][:-1]))
    units, pos, last = [], 0, None
    while pos < len(recs):
        first = recs[pos][1]; idx = 0; full = None
        while True:
            if idx == 30000:
                full = False; break
            if pos + idx >= len(recs):
                full = True; break                    # EOF: deferred error, the unit is returned as a full barcode
            if recs[pos + idx][1] != first:
                full = True; break                    # the record is left pending
            if last is not None and first == last and idx >= 200:
                idx += 1; full = False; break         # "abnormal break": the record stays
            idx += 1
        units.append((pos, pos + idx, full)); pos += idx; last = first
    return recs, units


def _bgzf_bytes(data: bytes, block: int, level: int = 1) -> bytes:
    """`data` as a BGZF file (SAM spec 4.1): gzip members of at most `block` input bytes with the BC extra field + the EOF marker."""
    import struct, zlib
    out = []
    for i in range(0, len(data), block):
        d = data[i:i + block]
        c = zlib.compressobj(level, zlib.DEFLATED, -15); comp = c.compress(d) + c.flush()
        out.append(struct.pack("<BBBBIBBHBBHH", 0x1f, 0x8b, 8, 4, 0, 0, 0xff, 6, 66, 67, 2, len(comp) + 25) + comp + struct.pack("<II", zlib.crc32(d), len(d)))
    out.append(bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000"))
    return b"".join(out)


@pytest.mark.parametrize("gz", [False, True, "bgzf", "bgzf_small", "members"])
def test_ingest_matches_reader_go(tmp_path, gz):
    import gzip, hashlib
    from arachne_b200 import Ingest
    g = np.random.default_rng(9)
    # barcodes: ordinary ones, one above the 30000-record cap (-> 30000 + 201-record continuation units), a plain (no '-') one,
    # a header without BX (empty name and barcode), one with VX:i:0, lower-case / N / '-' bases, a last field that is a read group
    plan = [("A01C02B03D04-1", 7), ("A01C02B03D05-1", 3), ("NOBX", 2), ("A09C02B03D05", 6), ("BIG-1", 30000 + 201 + 201 + 17), ("A11C12B13D14-1", 5), ("ZZZ-1", 1)]
    r1, r2, k = [], [], 0
    for bc, n in plan:
        for _ in range(n):
            L1 = int(g.integers(30, 152)); L2 = int(g.integers(30, 152))
            s1 = "".join("ACGTNacgt-"[int(x)] for x in g.integers(0, 10 if k % 50 == 0 else 4, L1)); s2 = "".join("ACGT"[int(x)] for x in g.integers(0, 4, L2))
            if k % 97 == 5:                    # every byte value a sequence line can hold, in a shuffled order (the 16-at-a-time base conversion)
                s1 = "".join(chr(int(x)) for x in g.permutation([c for c in range(1, 256) if c != 10])[:L1])
            q1 = "".join(chr(33 + int(x)) for x in g.integers(0, 41, L1)); q2 = "".join(chr(33 + int(x)) for x in g.integers(0, 41, L2))
            tag = "" if bc == "NOBX" else "\tBX:Z:%s\tVX:i:%d" % (bc, 0 if k % 7 == 3 else 1)
            tail = "\tRG:sample:lib:%d" % (k % 3) if k % 2 else ""
            r1 += ["@read%d/1%s%s\n" % (k, tag, tail), s1 + "\n", "+\n", q1 + "\n"]; r2 += ["@read%d/2%s%s\n" % (k, tag, tail), s2 + "\n", "+\n", q2 + "\n"]
            k += 1
    p1, p2 = str(tmp_path / ("r1.fq" + (".gz" if gz else ""))), str(tmp_path / ("r2.fq" + (".gz" if gz else "")))
    for p, lines in ((p1, r1), (p2, r2)):
        text = "".join(lines).encode("latin-1")
        if gz in ("bgzf", "bgzf_small"):       # bgzip-style blocks (inflated by a team of threads); 0xff00 is bgzip's block size
            blob = _bgzf_bytes(text, 0xff00 if gz == "bgzf" else 777)
        elif gz == "members":                  # a gzip file of several members without the BC field: the zlib stream path
            blob = b"".join(gzip.compress(text[i:i + 200_000]) for i in range(0, len(text), 200_000))
        else:
            blob = gzip.compress(text, 1) if gz else text
        with open(p, "wb") as f:
            f.write(blob)
    recs, units = _reader_go(r1, r2)
    nt4 = np.full(256, 4, np.uint8)
    for i, c in enumerate("ACGT"):
        nt4[ord(c)] = i; nt4[ord(c) + 32] = i
    nt4[ord("-")] = 5
    I = Ingest(p1, p2, n_threads=3)
    try:
        got_units, pair = [], 0
        for target in (10, 1, 100000, 50, 10**9):     # different batch sizes: units never straddle batches
            while True:
                b = I.next(target)
                if b is None:
                    break
                assert b["unit_off"][0] == 0 and b["unit_off"][-1] == b["n_pairs"] and b["n_units"] >= 1
                for u in range(b["n_units"]):
                    lo, hi = int(b["unit_off"][u]), int(b["unit_off"][u + 1])
                    got_units.append((pair + lo, pair + hi, bool(b["unique"][u])))
                    first = recs[pair + lo]
                    assert b["barcodes"][u] == first[1] and bool(b["valid"][u]) == first[2]
                    assert bool(b["run_rfa"][u]) == (bool(b["unique"][u]) and "-" in first[1] and hi - lo >= 5)
                for i in range(b["n_pairs"]):
                    name, bc, valid, rg, s1, q1, s2, q2 = recs[pair + i]
                    assert b["names"][i] == name and b["read_groups"][i] == rg
                    assert b["rand_seed"][i] == int.from_bytes(hashlib.md5(name.encode()).digest()[:8], "little", signed=True)
                    o = b["off"][2 * i:2 * i + 3]
                    assert np.array_equal(b["seq"][o[0]:o[1]], nt4[np.frombuffer(s1.encode("latin-1"), np.uint8)]) and np.array_equal(b["seq"][o[1]:o[2]], nt4[np.frombuffer(s2.encode(), np.uint8)])
                    assert b["quals"][2 * i] == q1 and b["quals"][2 * i + 1] == q2
                pair += b["n_pairs"]
                if target != 10**9 and pair > 200:
                    break
        assert pair == len(recs)
        assert got_units == units, (got_units[:12], units[:12])
        sizes = [hi - lo for lo, hi, _ in units]
        assert 30000 in sizes and sizes.count(201) == 2 and any(not f for _, _, f in units)
    finally:
        I.close()


def test_ingest_bad_inputs_fail_loudly(tmp_path):
    """A truncated or corrupted BGZF file, a record that does not start with '@' and a missing file are errors, never silent short reads."""
    from arachne_b200 import Ingest
    from arachne_b200.lib import ArachneError
    rec = lambda k, m: "@r%d/%d\tBX:Z:A01C02B03D04-1\tVX:i:1\nACGTACGTAC\n+\nIIIIIIIIII\n" % (k, m)
    good1 = "".join(rec(k, 1) for k in range(3000)).encode(); good2 = "".join(rec(k, 2) for k in range(3000)).encode()
    blob = _bgzf_bytes(good1, 4096)
    cases = {"truncated": blob[:len(blob) // 2 + 11], "crc": blob[:200] + bytes([blob[200] ^ 0x55]) + blob[201:], "garbage_after": blob[:4000] + b"not a block at all, just text" * 3}
    p2 = str(tmp_path / "r2.fq"); open(p2, "wb").write(good2)
    for name, data in cases.items():
        p1 = str(tmp_path / (name + ".fq.gz")); open(p1, "wb").write(data)
        I = Ingest(p1, p2, n_threads=2)
        try:
            with pytest.raises(ArachneError):
                while I.next(1000) is not None:
                    pass
        finally:
            I.close()
    bad = str(tmp_path / "bad.fq"); open(bad, "wb").write(good1[:500] + b"oops\n" + good1[500:])
    I = Ingest(bad, p2, n_threads=2)
    try:
        with pytest.raises(ArachneError, match="malformed FASTQ"):
            while I.next(1000) is not None:
                pass
    finally:
        I.close()
    with pytest.raises(ArachneError):
        Ingest(str(tmp_path / "missing.fq"), p2)
    # the good pair, for contrast, reads to the end: 3000 pairs in one unit of one barcode
    ok = str(tmp_path / "ok.fq.gz"); open(ok, "wb").write(blob)
    I = Ingest(ok, p2, n_threads=2); n = 0
    try:
        while True:
            b = I.next(1000)
            if b is None:
                break
            n += b["n_pairs"]
    finally:
        I.close()
    assert n == 3000


# ---- rows a15-a20: does the order of equal positions in sort.Sort(ByPosition) (Go's unstable pdqsort, aligner.go:1477-1479) reach any output? ----
def test_place_tie_order_is_not_observable(workdir):
    """The device sorts the per-contig position lists stably; Go's pdqsort may order EQUAL positions differently once a list is longer than
    12.  Run the placement oracle three ways -- ties in insertion order (= the device), ties reversed, and a restatement of Go's pdqsort -- on
    barcodes whose lists are hundreds long with many equal positions (stLFR-style units on a repetitive reference, duplicated pairs), and
    require identical outputs.  (Exact Go execution stays unpinned: no toolchain; this pins that the unpinned freedom changes nothing here.)"""
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    R = RefBwa(); R.load(fa)
    from arachne_b200 import simulate as S
    rs = S.simulate_linked_reads(ctg, n_barcodes=3, seed=41, style="stlfr", dash_suffix=True, mol_per_bc=10.0, pairs_per_mol=20.0, sub=0.01)
    # equal positions on purpose: copy every fifth pair onto its successor
    seq = rs.seq.copy()
    for p in range(0, rs.n_pairs - 1, 5):
        a, b = rs.off[2 * p], rs.off[2 * p + 2]
        if rs.off[2 * p + 4] - b == b - a:
            seq[b:rs.off[2 * p + 4]] = seq[a:b]
    ref = R.batch(seq, rs.off, rescue=True, cigar=True, threads=4)
    P = util.PlaceOracle(fa)
    P.O.orc_set_tie_mode.argtypes = [C.c_int]
    seeds = np.array([P.name_seed(n) for n in rs.names], dtype=np.int64)
    rfa = np.ones(len(rs.bc_off) - 1, np.uint8)
    outs = []
    try:
        for mode in (0, 1, 2):
            P.O.orc_set_tie_mode(mode)
            oc, cands, lmp, recs = P.place_records(seq, rs.off, rs.bc_off, rfa, rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
            outs.append((oc.copy(), cands.copy(), lmp.copy(), recs.copy()))
    finally:
        P.O.orc_set_tie_mode(0)
    pos = outs[0][1]["pos"][outs[0][1]["pos"] >= 0]
    assert len(pos) - len(np.unique(pos)) > 50 and np.diff(rs.bc_off).max() > 100, "dataset has no long lists with equal positions"
    for m in (1, 2):
        assert np.array_equal(outs[0][0], outs[m][0])
        assert outs[0][1].tobytes() == outs[m][1].tobytes(), "tie mode %d changes a candidate field" % m
        assert np.array_equal(outs[0][2], outs[m][2]) and outs[0][3].tobytes() == outs[m][3].tobytes(), "tie mode %d changes a record" % m


def test_pack_step_matches_bwa_index(tmp_path):
    """arachne_index_pack against the reference's own `bwa index` pack step (bns_fasta2bntseq + bns_dump, compiled into oracle/_ref): .pac,
    .ann and .amb byte for byte on a FASTA with N runs (the lrand48 replacement), IUPAC codes, lower case, comments, short and empty lines,
    CRLF line ends, several contigs and lengths around the 4-base packing boundary."""
    import gzip
    from arachne_b200 import ArachneCuda
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    g = np.random.default_rng(21)
    for case in range(6):
        recs = []
        for c in range(int(g.integers(1, 6))):
            L = int(g.integers(60, 4000)) + case
            s = np.array(list("ACGT"))[g.integers(0, 4, L)]
            for _ in range(int(g.integers(0, 5))):
                a = int(g.integers(0, L - 5)); n = int(g.integers(1, 40))
                s[a:a + n] = g.choice(list("NnRYM"))
            low = g.random(L) < 0.2
            s[low] = np.char.lower(s[low])
            recs.append(("ctg%d_%d" % (case, c), ("a comment %d" % c) if c % 2 else "", "".join(s)))
        eol = "\r\n" if case == 3 else "\n"
        txt = ""
        for name, com, s in recs:
            txt += ">" + name + ((" " + com) if com else "") + eol
            w = int(g.integers(20, 90))
            for i in range(0, len(s), w):
                txt += s[i:i + w] + eol
                if i == 0 and case in (2, 3):
                    txt += eol  # an empty line inside a record; under CRLF it is a lone CR, which kseq drops only because the ACCUMULATED sequence is longer than one byte (found by scripts/cpu_fuzz.py)
        fa = str(tmp_path / ("c%d.fa" % case)) + (".gz" if case == 4 else "")
        with (gzip.open(fa, "wt", newline="") if case == 4 else open(fa, "w", newline="")) as f:
            f.write(txt)
        ref_prefix = str(tmp_path / ("ref%d" % case)); shutil.copy(fa, ref_prefix)
        RefBwa().build_index(ref_prefix)                      # the reference: FASTA at <prefix>, files next to it
        mine = str(tmp_path / ("mine%d" % case))
        l_pac = ArachneCuda().pack_fasta(fa, mine)
        assert l_pac == sum(len(s) for _, _, s in recs)
        for ext in ("pac", "ann", "amb"):
            assert open(mine + "." + ext, "rb").read() == open(ref_prefix + "." + ext, "rb").read(), (case, ext)
