"""GPU parity tier (pytest -m gpu): every stage of the CUDA path through the C ABI against the oracle
(oracle/restate_*.c) and, when present, against the reference's own code (oracle/_ref)."""
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
