"""GPU parity tier (pytest -m gpu): every stage of the CUDA path through the C ABI against the oracle
(oracle/restate_*.c) and, when present, against the reference's own code (oracle/_ref)."""
import ctypes as C
import os

import numpy as np
import pytest

from tests import util
from tests.util import S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def A():
    from arachne_b200 import ArachneCuda
    a = ArachneCuda()
    assert a.device_count() > 0, "no CUDA device"
    return a


@pytest.fixture(scope="module")
def O():
    return util.oracle_lib()


@pytest.fixture(scope="module")
def R():
    from oracle.refbwa import RefBwa, available
    if not available():
        pytest.skip("oracle/_ref not built")
    return RefBwa()


@pytest.mark.parametrize("codec", ["byte", "wide"])
def test_ksw_extend_parity(A, O, codec):
    """k_ext_dp against the oracle's ksw_extend2 (itself pinned to oracle/_ref).  "byte": every problem has h0 + qlen <= 255, the
    column codec the pipeline runs (E<<24 | H<<16 | PRMT selector); "wide": unrestricted h0, the 13-bit codec of the stage entry."""
    qs, ts, h0, w = util.extension_problems(101, 4000)
    if codec == "byte":
        h0 = np.minimum(h0, 255 - np.array([len(x) for x in qs])).astype(np.int32)
        keep = [i for i in range(len(qs)) if h0[i] > 0]
        qs = [qs[i] for i in keep]; ts = [ts[i] for i in keep]; h0 = h0[keep]; w = w[keep]
        assert len(qs) > 3000
    else:
        assert any(int(h0[i]) + len(qs[i]) > 255 for i in range(len(qs)))
    q, qo = util.pack(qs); t, to = util.pack(ts)
    for zdrop, eb in ((100, 5), (0, 0), (20, 5)):
        out, cells, _ = A.ksw_extend(q, qo, t, to, h0, w, end_bonus=eb, zdrop=zdrop)
        tot = 0
        for i in range(len(qs)):
            exp, c = util.orc_extend(O, qs[i], ts[i], int(w[i]), eb, zdrop, int(h0[i]))
            assert tuple(out[i]) == exp, (i, len(qs[i]), len(ts[i]), int(w[i]), int(h0[i]), tuple(out[i]), exp)
            tot += c
        assert cells == tot  # in-band cell count is the GCUPS numerator: must be the reference's own count


def test_ksw_global_parity(A, O):
    g = np.random.default_rng(5)
    qs, ts, ws = [], [], []
    for _ in range(1500):
        ql = int(g.integers(1, 251)); q = g.integers(0, 4, ql).astype(np.uint8)
        t = util.mutate(g, q, g.choice([0.0, 0.02, 0.1]), g.choice([0.0, 0.01, 0.05, 0.1]))
        if len(t) == 0:
            continue
        qs.append(q); ts.append(t); ws.append(int(max(abs(len(t) - ql) + 3, g.choice([3, 5, 10, 50, 100, 400]))))
    q, qo = util.pack(qs); t, to = util.pack(ts)
    sc, coff, cig = A.ksw_global(q, qo, t, to, ws)
    sc2, _, _ = A.ksw_global(q, qo, t, to, ws, want_cigar=False)
    for i in range(len(qs)):
        es, ec = util.orc_global(O, qs[i], ts[i], ws[i])
        assert sc[i] == es and sc2[i] == es, (i, sc[i], sc2[i], es)
        assert list(cig[coff[i]:coff[i + 1]]) == ec, (i, list(cig[coff[i]:coff[i + 1]]), ec)


def test_index_build_parity(A, R, tmp_path):
    """arachne_index_build (device suffix sort) writes .bwt/.sa byte-identical to the reference's bwa_idx_build."""
    cases = util.index_build_cases() + [("rand2m-cap", S.make_reference(2_000_000, 3, seed=21), 300_000),
                                         ("rep1m", S.make_reference(1_000_000, 2, seed=22, repeat_frac=0.3), 0)]
    for name, ctg, cap in cases:
        fa = str(tmp_path / (name + ".fa"))
        S.write_fasta(fa, ctg)
        R.build_index(fa)
        want = {e: open(fa + e, "rb").read() for e in (".bwt", ".sa")}
        mine = str(tmp_path / (name + ".mine"))
        S.write_bns(mine, ctg)
        info = A.build_index(mine, 0, cap)
        assert info["seq_len"] == 2 * sum(len(s) for _, s in ctg)
        for e in (".bwt", ".sa"):
            assert open(mine + e, "rb").read() == want[e], (name, e, info)
    # and the files load: same occ / SA answers as the reference on the last case
    A.load_index(mine); R.load(fa)
    g = np.random.default_rng(3)
    k = g.integers(0, 2 * A.l_pac, 2000).astype(np.uint64)
    assert np.array_equal(A.sa(k), np.array([R.sa(int(x)) for x in k], dtype=np.uint64))
    A.close()


def test_index_build_errors(A, tmp_path):
    from arachne_b200.lib import ArachneError
    with pytest.raises(ArachneError):
        A.build_index(str(tmp_path / "missing"))
    p = str(tmp_path / "bad")
    S.write_bns(p, S.make_reference(1000, 1, seed=1))
    open(p + ".pac", "ab").write(b"\x07")  # trailing byte no longer l_pac % 4
    with pytest.raises(ArachneError):
        A.build_index(p)


@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
def test_index_and_regions_parity(A, R, workdir, name):
    fa, ctg, rs = util.make_dataset(name, workdir)
    R.load(fa)
    A.load_index(fa)
    assert A.l_pac == R.l_pac
    prim, *L2, sl = R.bwt_meta()
    g = np.random.default_rng(3)
    ks = np.array([0, 1, prim - 1, prim, prim + 1, sl - 1, sl, 127, 128, 129, 255, 256] + [int(x) for x in g.integers(0, sl + 1, 3000)], dtype=np.uint64)
    ls = np.minimum(ks + g.integers(0, 300, len(ks)).astype(np.uint64), np.uint64(sl))
    got = A.occ(ks, ls)
    km1 = (ks - np.uint64(1))  # k-1 wraps to (bwtint_t)-1 for k == 0, as bwt_extend passes it
    got2 = A.occ(km1, ls)
    for i in range(len(ks)):
        a, b = R.occ4(int(ks[i]), int(ls[i]))
        assert list(got[i]) == a + b
        a, b = R.occ4(int(km1[i]), int(ls[i]))
        assert list(got2[i]) == a + b
    sa = A.sa(ks)
    for i in range(len(ks)):
        assert int(sa[i]) == R.sa(int(ks[i]))
    # mem_align1_core regions for every read (no rescue): byte-identical mem_alnreg_t records
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=1)
    ref = R.batch(rs.seq, rs.off, rescue=False, cigar=False, threads=8)
    assert np.array_equal(res["reg_off"], ref["reg_off"])
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    assert res["counters"]["n_ext_cells"] > 0 and res["counters"]["n_lf_steps"] >= 0  # 0 LF steps when HBM holds every SA row (index.cuh)


@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
def test_rescue_and_cigar_parity(A, R, workdir, name):
    """GoBwaMemMateSW (with mate rescue) + GoBwaSmithWaterman for every candidate, against the reference's own code."""
    fa, ctg, rs = util.make_dataset(name, workdir, damaged=True)
    R.load(fa); A.load_index(fa)
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=7)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    norescue = R.batch(rs.seq, rs.off, rescue=False, cigar=False, threads=8)
    assert ref["regs"].tobytes() != norescue["regs"].tobytes(), "dataset does not exercise rescue"
    assert np.array_equal(res["reg_off"], ref["reg_off"])
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    a, e = res["alns"], ref["aln"]  # e: pos rid is_rev NM n_cigar cigar_off score sub
    assert np.array_equal(a["pos"], e[:, 0]) and np.array_equal(a["rid"], e[:, 1]) and np.array_equal(a["is_rev"], e[:, 2])
    assert np.array_equal(a["nm"], e[:, 3]) and np.array_equal(a["n_cigar"], e[:, 4])
    assert np.array_equal(a["score"], e[:, 6]) and np.array_equal(a["sub"], e[:, 7])
    for i in range(len(a)):
        got = res["cigar"][a["cigar_off"][i]:a["cigar_off"][i] + a["n_cigar"][i]]
        exp = ref["cigar"][e[i, 5]:e[i, 5] + e[i, 4]]
        assert np.array_equal(got, exp), i


# ---- rows a15-a20: GetAlignments walk, tagBestAlignments, molecules, optimizer, MAPQ ----
def _place_inputs(A, rs, rfa):
    n_pairs = rs.n_pairs
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    run_rfa = np.full(len(rs.bc_off) - 1, rfa, np.uint8)
    return seeds, run_rfa


def _check_place(res, oc, ocands, olmp, what):
    bad = util.place_mismatches(res, oc, ocands, olmp)  # tolerances stated there
    assert not bad, (what, bad)


@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
@pytest.mark.parametrize("rfa", [0, 1])
def test_place_parity(A, R, workdir, name, rfa):
    """Stage PLACE against the CPU restatement of the Go code, fed with the same candidate regions / CIGARs."""
    fa, ctg, rs = util.make_dataset(name, workdir, damaged=True)
    R.load(fa); A.load_index(fa)
    seeds, run_rfa = _place_inputs(A, rs, rfa)
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=15, run_rfa=run_rfa, rand_seed=seeds)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    P = util.PlaceOracle(fa)
    # a15: the CIGAR / reference walk statistics of every candidate
    O = P.O
    import ctypes as C
    a = res["alns"]
    for g in range(0, len(a), 3):
        r = int(np.searchsorted(res["reg_off"], g, side="right") - 1)
        q = np.ascontiguousarray(rs.read(r)); out5 = (C.c_int * 5)(); mm = (C.c_int32 * (2 * len(q) + 8))()
        a8 = np.ascontiguousarray(ref["aln"][g]); reg = np.ascontiguousarray(ref["regs"][g:g + 1])
        n = O.orc_walk(C.byref(P.ref), len(q), q.ctypes.data_as(C.c_void_p), reg.ctypes.data_as(C.c_void_p), a8.ctypes.data_as(C.c_void_p),
                       np.ascontiguousarray(ref["cigar"]).ctypes.data_as(C.c_void_p), out5, mm)
        assert (a["matches"][g], a["mismatches"][g], a["indels"][g], a["soft_clipped"][g], a["soft_clipped_length"][g]) == tuple(out5), g
        assert a["n_mm"][g] == n and list(res["mm"][a["mm_off"][g]:a["mm_off"][g] + 2 * n]) == list(mm[:2 * n]), g
    oc, ocands, olmp = P.place(rs.seq, rs.off, rs.bc_off, run_rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    _check_place(res, oc, ocands, olmp, (name, rfa))
    act = ocands[ocands["active"] == 1]
    assert len(act) == 2 * rs.n_pairs
    if rfa:
        assert ocands["molecule_id"].max() >= 1 and act["active_molecule"].sum() > 0


def test_place_moves_and_ties(A, R, workdir):
    """A repeat-rich reference with few reads per molecule: the optimizer has moves to make, ties to break, and some pairs
    need more than 273 PRNG outputs (the stateful generator path)."""
    ctg = util.exact_repeat_reference(total_len=2_000_000, n_contigs=4, seed=9, families=12, copies=24)
    fa = os.path.join(workdir, "rep_exact.fa")
    if not os.path.exists(fa + ".bwt"):
        S.write_fasta(fa, ctg); R.build_index(fa)
    R.load(fa); A.load_index(fa)
    rs = S.simulate_linked_reads(ctg, n_barcodes=40, seed=77, sub=0.015, mol_per_bc=6.0, pairs_per_mol=12.0)
    seeds, run_rfa = _place_inputs(A, rs, 1)
    run_rfa[::5] = 0
    cs = np.full(len(ctg), -1, np.int64); ce = np.full(len(ctg), -1, np.int64); cs[0], ce[0] = 10_000, 30_000
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=15, run_rfa=run_rfa, rand_seed=seeds, improper_penalty=-4.0, centromeres=(cs, ce))
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    P = util.PlaceOracle(fa, centromeres=(cs, ce))
    oc, ocands, olmp = P.place(rs.seq, rs.off, rs.bc_off, run_rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    _check_place(res, oc, ocands, olmp, "rep_exact")
    act = ocands["active"] == 1
    assert (ocands["bwa_pick"][act] == 0).sum() > 0, "the optimizer never moved a read: dataset too easy"
    assert (ocands["mapq"][act] < 60).sum() > 10
    ncomb = np.diff(oc)[0::2] * np.diff(oc)[1::2]
    assert ncomb.max() > 273, "no pair exercises the stateful PRNG path"


def _edge_readset(ctg, seed):
    """Ragged / degenerate reads on top of a simulated set: 1-base reads, reads shorter than min_seed_len, all-N reads,
    N-rich reads, 250-base reads (the i16 mate-rescue kernel, l_ms >= 250 is never reached at 250 but 249/250 straddle the
    u8/i16 switch), homopolymers, reads hanging over a contig end, and single-pair barcodes."""
    g = np.random.default_rng(seed)
    base = S.simulate_linked_reads(ctg, n_barcodes=12, seed=seed, read_len=250, sub=0.02, ins=0.003, dele=0.003)
    seqs = [base.read(i).copy() for i in range(2 * base.n_pairs)]
    ref0 = ctg[0][1]
    def rc(s):
        r = s[::-1].copy(); m = r < 4; r[m] = 3 - r[m]; return r
    for p in range(0, base.n_pairs, 5):
        kind = (p // 5) % 9
        i = 2 * p + int(g.integers(0, 2))
        if kind == 0: seqs[i] = seqs[i][:1]
        elif kind == 1: seqs[i] = seqs[i][:int(g.integers(2, 19))]
        elif kind == 2: seqs[i] = np.full(int(g.integers(20, 120)), 4, np.uint8)
        elif kind == 3:
            s = seqs[i].copy(); s[g.random(len(s)) < 0.15] = 4; seqs[i] = s
        elif kind == 4: seqs[i] = np.full(150, int(g.integers(0, 4)), np.uint8)
        elif kind == 5: seqs[i] = seqs[i][:249]
        elif kind == 6: seqs[i] = np.concatenate([ref0[-90:], g.integers(0, 4, 60).astype(np.uint8)])      # runs off the contig end
        elif kind == 7: seqs[i] = rc(np.concatenate([g.integers(0, 4, 40).astype(np.uint8), ref0[:110]]))  # ... and off its start
        else: seqs[i] = seqs[i][:int(g.integers(19, 40))]
    off = np.zeros(len(seqs) + 1, np.int64); np.cumsum([len(s) for s in seqs], out=off[1:])
    bc = list(base.bc_off)
    extra = sorted(set(int(x) for x in g.integers(1, base.n_pairs, 6)) - set(bc))  # cut a few barcodes into 1..few-pair units
    bc_off = np.array(sorted(set(bc) | set(extra)), dtype=np.int64)
    return np.concatenate(seqs).astype(np.uint8), off, bc_off, base.names


@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
def test_edge_inputs_all_stages(A, R, workdir, name):
    fa, ctg, _ = util.make_dataset(name, workdir)
    R.load(fa); A.load_index(fa)
    seq, off, bc_off, names = _edge_readset(ctg, 41)
    seeds = np.array([A.name_seed(n) for n in names], dtype=np.int64)
    run_rfa = (np.arange(len(bc_off) - 1) % 3 != 0).astype(np.uint8)
    res = A.align_batch(seq, off, bc_off, stages=15, run_rfa=run_rfa, rand_seed=seeds)
    ref = R.batch(seq, off, rescue=True, cigar=True, threads=8)
    assert np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes()
    a, e = res["alns"], ref["aln"]
    for k, f in enumerate(("pos", "rid", "is_rev", "nm", "n_cigar")):
        assert np.array_equal(a[f], e[:, k]), f
    for i in range(len(a)):
        assert np.array_equal(res["cigar"][a["cigar_off"][i]:a["cigar_off"][i] + a["n_cigar"][i]], ref["cigar"][e[i, 5]:e[i, 5] + e[i, 4]]), i
    P = util.PlaceOracle(fa)
    oc, ocands, olmp = P.place(seq, off, bc_off, run_rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    _check_place(res, oc, ocands, olmp, ("edge", name))
    assert (ocands["reg"] == -1).sum() > 5, "no unmapped placeholders exercised"


@pytest.mark.parametrize("cfg", ["stlfr", "tellseq", "len100", "len250"])
def test_config_flavours_all_stages(A, R, workdir, cfg):
    """BASELINE.json configs[2..4] as parity cases: stLFR-style barcodes (hundreds of pairs per barcode: long candidate lists,
    many molecules per barcode in the optimizer), TELLseq-style with 30 % invalid barcodes (unbarcoded path: run_rfa = 0 for
    those, aligner.go:1018-1030), and the microbench read lengths 100 / 250."""
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    R.load(fa); A.load_index(fa)
    if cfg == "stlfr":
        rs = S.simulate_linked_reads(ctg, n_barcodes=4, seed=31, style="stlfr", dash_suffix=True, mol_per_bc=14.0, pairs_per_mol=22.0, sub=0.015)
        assert np.diff(rs.bc_off).max() > 250
        run_rfa = np.ones(len(rs.bc_off) - 1, np.uint8)
    elif cfg == "tellseq":
        rs = S.simulate_linked_reads(ctg, n_barcodes=60, seed=32, style="tellseq", invalid_frac=0.3, sub=0.015)
        g = np.random.default_rng(32)
        run_rfa = (g.random(len(rs.bc_off) - 1) >= 0.3).astype(np.uint8)  # VX:i:0 barcodes are not molecule-grouped
        assert 0 < run_rfa.sum() < len(run_rfa)
    else:
        rs = S.simulate_linked_reads(ctg, n_barcodes=25, seed=33, read_len=int(cfg[3:]), sub=0.02, ins=0.002, dele=0.002)
        run_rfa = np.ones(len(rs.bc_off) - 1, np.uint8)
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=15, run_rfa=run_rfa, rand_seed=seeds)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    assert np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes(), cfg
    a, e = res["alns"], ref["aln"]
    for f, jx in (("pos", 0), ("rid", 1), ("is_rev", 2), ("nm", 3), ("n_cigar", 4), ("score", 6), ("sub", 7)):
        assert np.array_equal(a[f], e[:, jx]), (cfg, f)
    for i in range(len(a)):
        assert np.array_equal(res["cigar"][a["cigar_off"][i]:a["cigar_off"][i] + a["n_cigar"][i]], ref["cigar"][e[i, 5]:e[i, 5] + e[i, 4]]), (cfg, i)
    oc, ocands, olmp = util.PlaceOracle(fa).place(rs.seq, rs.off, rs.bc_off, run_rfa, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    _check_place(res, oc, ocands, olmp, cfg)


def test_fetch_seq_matches_reference_text(A, workdir):
    """arachne_index_fetch_seq = bns_fetch_seq as GetSeq calls it: forward ranges read the contig, reverse ranges its
    reverse complement, both clamped to the contig that holds `mid`."""
    from arachne_b200.lib import ArachneError
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    A.load_index(fa)
    ann = util.read_ann(fa)
    l_pac = A.l_pac
    genome = np.concatenate([s for _, s in ctg])
    nmask = genome > 3  # N became a random base in the .pac (bntseq.c:261): compare elsewhere
    g = np.random.default_rng(8)
    offs = np.cumsum([0] + [len(s) for _, s in ctg])
    for _ in range(300):
        ci = int(g.integers(0, len(ctg))); L = len(ctg[ci][1])
        a = int(g.integers(-50, L - 10)); b = a + int(g.integers(1, 400))
        beg, end = offs[ci] + a, offs[ci] + b          # may hang over either contig end
        mid = int(np.clip((beg + end) >> 1, offs[ci], offs[ci] + L - 1))
        seq, cb, ce, rid = A.fetch_seq(int(beg), mid, int(end))
        assert rid == ci and cb == max(beg, offs[ci]) and ce == min(end, offs[ci] + L) and len(seq) == ce - cb
        ok = ~nmask[cb:ce]
        assert np.array_equal(seq[ok], genome[cb:ce][ok])
        rb, re_, rmid = 2 * l_pac - ce, 2 * l_pac - cb, 2 * l_pac - 1 - mid  # the same bases on the reverse strand
        rseq, rcb, rce, rrid = A.fetch_seq(int(rb), int(rmid), int(re_))
        assert rrid == ci and (rcb, rce) == (rb, re_)
        assert np.array_equal((3 - rseq[::-1])[ok], genome[cb:ce][ok])
    with pytest.raises(ArachneError):
        A.fetch_seq(10, 500, 20)  # mid outside the range


def test_concurrent_batches_on_one_handle(A, R, workdir):
    """Two host threads in flight on one handle (the reference's worker goroutines share one index, aligner.go:335-337):
    each call gets its own workspace; results equal the sequential ones and the reference's."""
    from concurrent.futures import ThreadPoolExecutor
    fa, ctg, rs = util.make_dataset("rand1m", workdir, damaged=True)
    R.load(fa); A.load_index(fa)
    nb = len(rs.bc_off) - 1
    parts = [rs.subset(nb * k // 6, nb * (k + 1) // 6) for k in range(6)]
    seq = [A.align_batch(p.seq, p.off, p.bc_off, stages=7) for p in parts]
    with ThreadPoolExecutor(3) as ex:
        par = list(ex.map(lambda p: A.align_batch(p.seq, p.off, p.bc_off, stages=7), parts * 2))
    for k, p in enumerate(parts):
        ref = R.batch(p.seq, p.off, rescue=True, cigar=True, threads=4)
        for res in (seq[k], par[k], par[k + 6]):
            assert np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes(), k
            a, b = res["alns"], seq[k]["alns"]  # the CIGAR pool is filled in completion order: compare through the offsets
            for f in a.dtype.names:
                if f != "cigar_off" and f != "mm_off":
                    assert np.array_equal(a[f], b[f]), (k, f)
            for i in range(0, len(a), 7):
                assert np.array_equal(res["cigar"][a["cigar_off"][i]:a["cigar_off"][i] + a["n_cigar"][i]],
                                      seq[k]["cigar"][b["cigar_off"][i]:b["cigar_off"][i] + b["n_cigar"][i]]), (k, i)


def test_empty_and_tiny_batches(A, R, workdir):
    fa, ctg, rs = util.make_dataset("rand1m", workdir)
    A.load_index(fa); R.load(fa)
    res = A.align_batch(np.zeros(0, np.uint8), np.zeros(1, np.int64), None, stages=15)
    assert len(res["reg_off"]) == 1 and len(res["regs"]) == 0
    one = rs.subset(0, 1)
    seq, off = one.seq[:one.off[2]], one.off[:3]
    res = A.align_batch(seq, off, None, stages=15)   # one pair, no barcode table, no seeds
    ref = R.batch(seq, off, rescue=True, cigar=True, threads=1)
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    assert res["cands"]["active"].sum() == 2 and len(res["cand_off"]) == 3
    from arachne_b200 import ArachneError
    with pytest.raises(ArachneError):  # reads longer than ARACHNE_MAX_READ_LEN are refused, not truncated
        A.align_batch(np.zeros(600, np.uint8), np.array([0, 300, 600], np.int64), None, stages=7)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["rand1m", "rep600k"])
def test_seed_intervals_every_shortcut(A, R, workdir, name):
    """k_seed2's shortcuts -- k-mer tables (LAST-like pass, forward windows, short backward extensions), direct text comparison
    of unique forward matches (SA / inverse SA / both-strand text), the forward window in a trip of its own -- against the
    reference's own bwt_smem1 / bwt_seed_strategy1 composed into mem_collect_intv, and against plain stepping: identical
    interval arrays (x0, x1, x2, info) for every read, including exact reference substrings of both strands (direct comparison
    running into the read's end), reads with an N inside an exact match, all-N and too-short reads."""
    fa, ctg, rs = util.make_dataset(name, workdir)
    R.load(fa); A.load_index(fa)
    g = np.random.default_rng(11)
    ref = np.concatenate([np.asarray(c[1], dtype=np.uint8) for c in ctg])
    reads = [rs.seq[rs.off[r]:rs.off[r + 1]] for r in range(0, min(len(rs.off) - 1, 240))]
    for k in range(24):
        p = int(g.integers(0, len(ref) - 260)); L = int(g.choice([30, 100, 150, 250]))
        s = ref[p:p + L].copy()
        if s.max() > 3:
            continue
        if k % 3 == 1:
            s = (3 - s)[::-1].copy()
        if k % 3 == 2:
            s[int(g.integers(20, L - 5))] = 4
        reads.append(s)
    reads += [np.full(40, 4, dtype=np.uint8), np.zeros(12, dtype=np.uint8), np.array([0, 1, 2, 3] * 30, dtype=np.uint8)]
    seq, off = util.pack(reads)
    base = None
    try:
        for kmer, direct in [(0, 0), (1, 0), (0, 1), (1, 1), (5, 1)]:
            A.set_option("seed_kmer", kmer); A.set_option("seed_direct", direct)
            ioff, iv = A.seed(seq, off)
            if base is None:
                base = (ioff.copy(), iv.copy())
                for r, q in enumerate(reads):
                    got = [tuple(int(v) for v in a) for a in iv[ioff[r]:ioff[r + 1]]]
                    assert got == util.ref_collect_intv(R, q), (name, r)
            else:
                assert np.array_equal(ioff, base[0]) and np.array_equal(iv, base[1]), (name, kmer, direct)
    finally:
        A.set_option("seed_kmer", 1); A.set_option("seed_direct", 1)


# ---- row b: the gobwa-shaped layer (include/arachne_compat.h), what integration/gobwa/gobwa_cuda.go binds ----
def test_compat_layer_per_pair_calls_batched(A, R, workdir):
    """One barcode's worth of per-pair GoBwaMemMateSW calls + per-candidate GoBwaSmithWaterman calls, made from many threads the
    way the reference's worker goroutines make them, against the reference's own C code driven in gobwa.go's call order
    (oracle/ref_shim.c).  Also: the calls really were batched, region arrays stay valid until released, GetSeq = bns_fetch_seq."""
    from concurrent.futures import ThreadPoolExecutor
    from arachne_b200 import Compat
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    R.load(fa); A.load_index(fa)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    n = min(rs.n_pairs, 1500)
    K = Compat(A, max_batch_pairs=256, max_wait_us=2000)
    try:
        def one(p):
            r1 = np.ascontiguousarray(rs.seq[rs.off[2 * p]:rs.off[2 * p + 1]]); r2 = np.ascontiguousarray(rs.seq[rs.off[2 * p + 1]:rs.off[2 * p + 2]])
            g1, g2, a1, a2, owner = K.mem_mate_sw(r1, r2, 25)
            alns = [K.reg2aln(base + 88 * k) for base, g in ((a1, g1), (a2, g2)) for k in range(len(g))]
            return p, g1, g2, alns, owner
        with ThreadPoolExecutor(64) as ex:
            out = list(ex.map(one, range(n)))
        st = K.stats()
        assert st["n_calls"] == n and st["n_batches"] < n // 4 and st["max_batch"] > 8, st  # dynamic batching happened
        assert st["n_live_blocks"] == n
        e = ref["aln"]
        for p, g1, g2, alns, owner in out:
            lo, mid, hi = (int(ref["reg_off"][2 * p + k]) for k in range(3))
            assert g1.tobytes() == ref["regs"][lo:mid].tobytes() and g2.tobytes() == ref["regs"][mid:hi].tobytes(), p
            assert len(alns) == hi - lo
            for k, a in enumerate(alns):
                x = e[lo + k]
                assert (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) == (x[0], x[1], x[2], x[3], x[4], x[6], x[7]), (p, k)
                assert np.array_equal(a["cigar"], ref["cigar"][x[5]:x[5] + x[4]]), (p, k)
        # a region handed out earlier is still addressable after every other call returned (gobwa's Arena lifetime) ...
        first = next((o, a1 or a2) for o in out for a1, a2 in [(len(o[1]) > 0, len(o[2]) > 0)] if a1 or a2)[0]
        r1 = np.ascontiguousarray(rs.seq[rs.off[2 * first[0]]:rs.off[2 * first[0] + 1]]); r2 = np.ascontiguousarray(rs.seq[rs.off[2 * first[0] + 1]:rs.off[2 * first[0] + 2]])
        g1, g2, a1, a2, ow2 = K.mem_mate_sw(r1, r2, 25)
        addr = a1 or a2
        again = K.reg2aln(addr)
        assert again["pos"] == first[3][0]["pos"] and np.array_equal(again["cigar"], first[3][0]["cigar"])
        K.release(ow2)
        # ... and not after its block is released
        from arachne_b200 import ArachneError
        with pytest.raises(ArachneError):
            K.reg2aln(addr)
        for _, _, _, _, ow in out:
            K.release(ow)
        assert K.stats()["n_live_blocks"] == 0
        # GetSeq (gobwa.go:50-80): bns_fetch_seq with the clamp to the contig, both strands
        l_pac = A.l_pac
        for beg, end in ((100, 400), (l_pac - 50, l_pac + 70), (2 * l_pac - 300, 2 * l_pac - 20)):
            got, b, en, rid = K.fetch_seq(beg, (beg + end) >> 1, end)
            exp, eb, ee, erid = A.fetch_seq(beg, (beg + end) >> 1, end)  # pinned to the reference text in test_fetch_seq_matches_reference_text
            assert (b, en, rid) == (eb, ee, erid) and np.array_equal(got, exp)
    finally:
        K.close()


# ---- row f1: record assembly (markDuplicates, CheckSplitReads, AppendBam's fields, SAM text) ----
def _records_dataset(ctg, seed):
    """Linked reads with what record assembly branches on: exact duplicate pairs inside a barcode, chimeric reads (two halves from
    different places: split alignments), random reads (unmapped records, weak-alignment demotion), mates far apart (improper)."""
    rs = S.simulate_linked_reads(ctg, n_barcodes=40, seed=seed, dash_suffix=True, sub=0.01)
    g = np.random.default_rng(seed)
    seq = rs.seq.copy(); off = rs.off
    genome = np.concatenate([s for _, s in ctg])
    for b in range(len(rs.bc_off) - 1):
        p0, p1 = int(rs.bc_off[b]), int(rs.bc_off[b + 1])
        for p in range(p0, p1 - 1):
            u = g.random()
            r1 = slice(off[2 * p], off[2 * p + 1]); r2 = slice(off[2 * p + 1], off[2 * p + 2])
            L = r1.stop - r1.start
            if u < 0.10:      # duplicate of the next pair
                seq[off[2 * p + 2]:off[2 * p + 3]] = seq[r1]; seq[off[2 * p + 3]:off[2 * p + 4]] = seq[r2]
            elif u < 0.30:    # chimera: the tail of read 1 comes from somewhere else
                cut = int(g.integers(50, L - 45)); at = int(g.integers(0, len(genome) - L))
                seq[r1.start + cut:r1.stop] = np.minimum(genome[at:at + L - cut], 3)
            elif u < 0.36:    # read 2 is noise
                seq[r2] = g.integers(0, 4, r2.stop - r2.start)
            elif u < 0.40:    # both noise
                seq[r1] = g.integers(0, 4, L); seq[r2] = g.integers(0, 4, r2.stop - r2.start)
            elif u < 0.46:    # mate from far away: improper pair
                at = int(g.integers(0, len(genome) - L)); seq[r2] = np.minimum(genome[at:at + (r2.stop - r2.start)], 3)
    rs.seq = seq
    return rs


@pytest.mark.parametrize("rfa", [1, 0])
def test_records_parity(A, R, workdir, rfa):
    """ARACHNE_STAGE_RECORDS against the oracle's restatement of markDuplicates / CheckSplitReads / AppendBam, field by field, and the
    library's SAM text against an independent formatter run on the oracle's records.  (Go side: parity unpinned, as rows a15-a20.)"""
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    R.load(fa); A.load_index(fa)
    rs = _records_dataset(ctg, 77)
    nb = len(rs.bc_off) - 1
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    run_rfa = np.full(nb, rfa, np.uint8)
    attach = (np.arange(nb) % 5 != 0).astype(np.uint8)  # work.unique_barcode
    if rfa:
        attach[:] = 1                                   # worthRunningRFA implies a unique barcode (aligner.go:1018-1030)
    quals = ["".join(chr(33 + (7 * i + k) % 40) for k in range(int(rs.off[i + 1] - rs.off[i]))) for i in range(2 * rs.n_pairs)]
    ubar = [rs.barcodes[b] for b in range(nb)]
    res = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=31, run_rfa=run_rfa, rand_seed=seeds, attach_bx=attach,
                        sam=dict(names=rs.names, quals=quals, barcodes=ubar, read_group="sample:lib:1"))
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    assert res["regs"].tobytes() == ref["regs"].tobytes()
    P = util.PlaceOracle(fa)
    oc, ocands, olmp, orecs = P.place_records(rs.seq, rs.off, rs.bc_off, run_rfa, attach, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    _check_place(res, oc, ocands, olmp, "records")
    g = res["records"]
    assert len(g) == len(orecs) and np.array_equal(res["rec_off"][1:] - res["rec_off"][:-1], np.bincount(orecs["read"], minlength=2 * rs.n_pairs))
    for f in g.dtype.names:
        if f in ("_pad", "dm"):
            continue
        d = np.nonzero(g[f] != orecs[f])[0]
        assert len(d) == 0, (f, len(d), int(d[0]), g[d[0]], orecs[d[0]])
    assert np.allclose(g["dm"], orecs["dm"], rtol=1e-9, atol=1e-12)
    # the dataset exercises the branches
    fl = g["flag"]
    assert (g["primary"] == 0).sum() > 20 and (fl & 0x400).sum() > 10 and (fl & 0x4).sum() > 20 and (fl & 0x8).sum() > 20 and ((fl & 0x2) == 0).sum() > 50
    assert (g["sa_reg"] >= 0).sum() > 40 and (g["has_xc"] == 1).sum() > 50 and (g["tlen"] != 0).sum() > 500
    # SAM text: the library's formatter on its records vs a pure-Python formatter on the oracle's records
    e = ref["aln"]
    O = util.oracle_lib()
    def mm_of_reg(gi):
        r = int(np.searchsorted(ref["reg_off"], gi, side="right") - 1)
        q = np.ascontiguousarray(rs.seq[rs.off[r]:rs.off[r + 1]])
        out5 = np.zeros(5, np.int32); mm = np.zeros(2 * len(q) + 8, np.int32)
        n = O.orc_walk(C.byref(P.ref), len(q), q.ctypes.data_as(C.c_void_p), ref["regs"][gi:gi + 1].ctypes.data_as(C.c_void_p), e[gi].ctypes.data_as(C.c_void_p),
                       ref["cigar"].ctypes.data_as(C.c_void_p), out5.ctypes.data_as(C.c_void_p), mm.ctypes.data_as(C.c_void_p))  # orc_walk indexes the pool with aln8[5]
        return mm[:n], mm[n:2 * n]
    names = [c[0] for c in util.read_ann(fa)[1]]
    exp = util.sam_lines(orecs, names, rs.seq, rs.off, rs.bc_off, rs.names, quals, ubar, "sample:lib:1", lambda gi: (int(e[gi, 4]), int(e[gi, 5])), ref["cigar"], mm_of_reg)
    got = res["sam"].rstrip("\n").split("\n")
    assert len(got) == len(exp)
    for i, (x, y) in enumerate(zip(got, exp)):
        assert x == y, (i, x, y)


def test_determinism_two_runs(A, R, workdir):
    """The same batch twice, and once more while a second host thread runs another batch on the handle's other workspace: every output
    array byte-identical.  Stands in for racecheck, which the GPU pool refuses (profiles/r2a_sanitizer_unavailable.txt): the kernels that
    write without atomics (interval halves, extension queues, copy-on-write rescue lists) would show up here as run-to-run differences."""
    from concurrent.futures import ThreadPoolExecutor
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    A.load_index(fa)
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    rfa = np.ones(len(rs.bc_off) - 1, np.uint8)
    def run(_=0):
        return A.align_batch(rs.seq, rs.off, rs.bc_off, stages=31, run_rfa=rfa, rand_seed=seeds)
    a = run()
    b = run()
    with ThreadPoolExecutor(2) as ex:
        c, d = list(ex.map(run, range(2)))
    def pools(x):  # CIGAR / mismatch words per region: the pool OFFSETS are handed out by atomics and may differ run to run, the words may not
        al = x["alns"]
        cg = np.concatenate([x["cigar"][o:o + n] for o, n in zip(al["cigar_off"].tolist(), al["n_cigar"].tolist())] + [np.zeros(0, np.uint32)])
        mm = np.concatenate([x["mm"][o:o + 2 * n] for o, n in zip(al["mm_off"].tolist(), al["n_mm"].tolist())] + [np.zeros(0, np.int32)])
        return cg, mm
    cga, mma = pools(a)
    for other in (b, c, d):
        for k in ("reg_off", "regs", "cand_off", "cands", "lmp", "rec_off", "records"):
            assert a[k].tobytes() == other[k].tobytes(), k
        for f in a["alns"].dtype.names:
            if f not in ("cigar_off", "mm_off"):
                assert np.array_equal(a["alns"][f], other["alns"][f]), f
        cgo, mmo = pools(other)
        assert np.array_equal(cga, cgo) and np.array_equal(mma, mmo)
    assert a["counters"]["n_ext_cells"] == b["counters"]["n_ext_cells"] and a["counters"]["n_occ_blocks"] == b["counters"]["n_occ_blocks"]


def test_out_no_regions_flag(A, R, workdir):
    """ARACHNE_OUT_NO_REGIONS: the region array is not copied back (regs is NULL) and every other array -- and the SAM text made from
    them -- is what the plain call returns."""
    from arachne_b200.lib import OUT_NO_REGIONS
    fa, ctg, rs = util.make_dataset("rep600k", workdir, damaged=True)
    A.load_index(fa)
    seeds = np.array([A.name_seed(n) for n in rs.names], dtype=np.int64)
    rfa = np.ones(len(rs.bc_off) - 1, np.uint8)
    sam = dict(names=list(rs.names), quals=None, barcodes=None, read_group=None)
    a = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=31, run_rfa=rfa, rand_seed=seeds, sam=sam)
    b = A.align_batch(rs.seq, rs.off, rs.bc_off, stages=31, run_rfa=rfa, rand_seed=seeds, sam=sam, flags=OUT_NO_REGIONS)
    assert len(a["regs"]) > 0 and b["regs"] is None
    for k in ("reg_off", "cand_off", "cands", "lmp", "rec_off", "records"):
        assert a[k].tobytes() == b[k].tobytes(), k
    for f in a["alns"].dtype.names:
        if f not in ("cigar_off", "mm_off"):
            assert np.array_equal(a["alns"][f], b["alns"][f]), f
    assert a["sam"] == b["sam"] and a["sam"].count("\n") == len(a["records"])


def test_fastq_to_sam_end_to_end(A, R, workdir, tmp_path):
    """scripts/align_fastq.py's path: FASTQ files -> native ingest (work units, md5 seeds, flags) -> all stages -> SAM text, against the same
    reads fed directly + the oracle's records formatted by the independent Python formatter.  The pieces are pinned separately; this pins
    that they fit together (offsets, unit flags, per-pair read groups, names with mate suffixes stripped)."""
    import subprocess, sys as _sys
    fa, ctg, _ = util.make_dataset("rep600k", workdir)
    rs = _records_dataset(ctg, 91)
    nb = len(rs.bc_off) - 1
    quals = ["".join(chr(35 + (3 * i + k) % 38) for k in range(int(rs.off[i + 1] - rs.off[i]))) for i in range(2 * rs.n_pairs)]
    dec = np.frombuffer(b"ACGTN", np.uint8)
    r1, r2 = str(tmp_path / "r1.fq"), str(tmp_path / "r2.fq")
    with open(r1, "w") as f1, open(r2, "w") as f2:
        for b in range(nb):
            for p in range(int(rs.bc_off[b]), int(rs.bc_off[b + 1])):
                for f, i, m in ((f1, 2 * p, 1), (f2, 2 * p + 1, 2)):
                    s = dec[np.minimum(rs.seq[rs.off[i]:rs.off[i + 1]], 4)].tobytes().decode()
                    f.write("@%s/%d\tBX:Z:%s\tVX:i:1\n%s\n+\n%s\n" % (rs.names[p], m, rs.barcodes[b], s, quals[i]))
    out = str(tmp_path / "out.sam")
    subprocess.run([_sys.executable, os.path.join(util.ROOT, "scripts", "align_fastq.py"), fa, r1, r2, "-o", out, "--batch-pairs", "700", "--threads", "3"], check=True)
    got = [l for l in open(out).read().split("\n") if l and not l.startswith("@")]
    # expectation: the reference's C code + the oracle's records on the same reads; RG is the header's last field (reader.go:143-152)
    R.load(fa)
    ref = R.batch(rs.seq, rs.off, rescue=True, cigar=True, threads=8)
    P = util.PlaceOracle(fa)
    from arachne_b200 import workunits as W
    uo, uq, ub = W.work_units(rs.bc_off, rs.barcodes)
    rfa = W.worth_running_rfa(uo, uq, ub, rs.barcodes)
    seeds = np.array([P.name_seed(n) for n in rs.names], dtype=np.int64)
    oc, ocands, olmp, orecs = P.place_records(rs.seq, rs.off, uo, rfa, uq, seeds, ref["reg_off"], ref["regs"], ref["aln"], ref["cigar"])
    e = ref["aln"]; O = util.oracle_lib()
    def mm_of_reg(gi):
        r = int(np.searchsorted(ref["reg_off"], gi, side="right") - 1)
        q = np.ascontiguousarray(rs.seq[rs.off[r]:rs.off[r + 1]])
        out5 = np.zeros(5, np.int32); mm = np.zeros(2 * len(q) + 8, np.int32)
        n = O.orc_walk(C.byref(P.ref), len(q), q.ctypes.data_as(C.c_void_p), ref["regs"][gi:gi + 1].ctypes.data_as(C.c_void_p), e[gi].ctypes.data_as(C.c_void_p),
                       ref["cigar"].ctypes.data_as(C.c_void_p), out5.ctypes.data_as(C.c_void_p), mm.ctypes.data_as(C.c_void_p))
        return mm[:n], mm[n:2 * n]
    names = [c[0] for c in util.read_ann(fa)[1]]
    ubar = [rs.barcodes[b] for b in ub.tolist()]
    exp = util.sam_lines(orecs, names, rs.seq, rs.off, uo, rs.names, quals, ubar, "VX:i:1", lambda gi: (int(e[gi, 4]), int(e[gi, 5])), ref["cigar"], mm_of_reg)
    assert len(got) == len(exp) and len(got) >= 2 * rs.n_pairs
    for i, (x, y) in enumerate(zip(got, exp)):
        assert x == y, (i, x, y)
