"""GPU parity tier (pytest -m gpu): every stage of the CUDA path through the C ABI against the oracle
(oracle/restate_*.c) and, when present, against the reference's own code (oracle/_ref)."""
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


def test_ksw_extend_parity(A, O):
    qs, ts, h0, w = util.extension_problems(101, 4000)
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
    assert res["counters"]["n_ext_cells"] > 0 and res["counters"]["n_lf_steps"] > 0


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
    g = res["cands"]
    assert np.array_equal(res["cand_off"], oc), what
    assert len(g) == len(ocands)
    for f in ("reg", "read", "id", "pos", "aend", "active", "is_proper", "bwa_pick", "molecule_id", "active_molecule"):
        bad = np.nonzero(g[f] != ocands[f])[0]
        assert len(bad) == 0, (what, f, bad[:5], g[f][bad[:5]], ocands[f][bad[:5]])
    assert np.array_equal(g["lap"], ocands["lap"]), what
    act = ocands["active"] == 1
    for f in ("second_best", "mate"):
        assert np.array_equal(g[f][act], ocands[f][act]), (what, f)
    # floating point: pair scores are sums of dyadic rationals (exact); the molecule penalty goes through log10 and MAPQ
    # through pow/log10, which differ from the host libm in the last bits.  Tolerance: 1e-9 relative on scores,
    # MAPQ equal unless the un-truncated value sits within 1e-6 of an integer (then +-1).
    for f in ("as_score", "xs_score", "molecule_difference", "molecule_confidence"):
        assert np.allclose(g[f][act], ocands[f][act], rtol=1e-9, atol=1e-9), (what, f)
    assert np.allclose(g["sum_move"][act], ocands["sum_move"][act], rtol=1e-9), what
    assert np.allclose(res["lmp"], olmp, rtol=1e-12, atol=1e-12), what
    dm = g["mapq"][act] != ocands["mapq"][act]
    if dm.any():
        raw = ocands["raw_mapq"][act][dm]
        assert np.all(np.abs(raw - np.round(raw)) < 1e-6) and np.all(np.abs(g["mapq"][act][dm] - ocands["mapq"][act][dm]) <= 1), (what, raw[:5])
    assert np.allclose(g["raw_mapq"][act], ocands["raw_mapq"][act], rtol=1e-9, atol=1e-9, equal_nan=True), what


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
