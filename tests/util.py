"""Shared helpers for the test tiers (test infrastructure; may use oracle/)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from arachne_b200 import simulate as S  # noqa: E402


class OrcSc(C.Structure):
    _fields_ = [(n, C.c_int) for n in "a b o_del e_del o_ins e_ins".split()]


DEFAULT_SC = OrcSc(1, 4, 6, 1, 6, 1)


def oracle_lib():
    """liboracle.so (our CPU restatement); built on demand with plain gcc/g++."""
    path = os.path.join(ROOT, "oracle", "liboracle.so")
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "restate"], check=True, stdout=subprocess.DEVNULL)
    return C.CDLL(os.environ.get("ARACHNE_ORACLE_LIB") or path)  # override: the sanitizer build (scripts/host_sanitize.sh)


def u8p(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, a.ctypes.data_as(C.POINTER(C.c_uint8))


def orc_extend(O, q, t, w, end_bonus, zdrop, h0):
    q, qp = u8p(q); t, tp = u8p(t); out = (C.c_int * 5)(); cells = C.c_uint64(0)
    s = O.orc_extend(C.byref(DEFAULT_SC), len(q), qp, len(t), tp, w, end_bonus, zdrop, h0, out, C.byref(cells))
    return (s, *out), cells.value


def orc_global(O, q, t, w):
    q, qp = u8p(q); t, tp = u8p(t); cap = len(q) + len(t) + 8; cig = (C.c_uint32 * cap)(); nc = C.c_int(0)
    s = O.orc_global(C.byref(DEFAULT_SC), len(q), qp, len(t), tp, w, cig, cap, C.byref(nc))
    return s, list(cig[:nc.value])


def mutate(g, s, sub, indel):
    out = []
    for x in s:
        r = g.random()
        if r < sub:
            out.append((int(x) + int(g.integers(1, 4))) & 3)
        elif r < sub + indel / 2:
            continue
        elif r < sub + indel:
            out.append(int(x)); out.append(int(g.integers(0, 4)))
        else:
            out.append(int(x))
    return np.array(out, dtype=np.uint8)


def extension_problems(seed, n, qmax=250):
    """Random + adversarial ksw_extend2 problems: N bases, truncated targets, tiny/huge bands, long seeds (stale eh[] reads)."""
    g = np.random.default_rng(seed)
    qs, ts, h0s, ws = [], [], [], []
    for _ in range(n):
        ql = int(g.integers(1, qmax + 1))
        q = g.integers(0, 4, ql).astype(np.uint8)
        if g.random() < 0.1:
            q[g.integers(0, ql)] = 4
        t = mutate(g, q, g.choice([0.0, 0.02, 0.1, 0.3]), g.choice([0.0, 0.01, 0.05]))
        t = np.concatenate([t, g.integers(0, 4, int(g.integers(0, 120))).astype(np.uint8)])
        if g.random() < 0.2:
            t = t[:int(g.integers(0, len(t) + 1))]
        if g.random() < 0.05:
            t = g.integers(0, 4, int(g.integers(0, 300))).astype(np.uint8)
        qs.append(q); ts.append(t)
        h0s.append(int(g.integers(1, 250))); ws.append(int(g.choice([1, 3, 8, 16, 32, 64, 100, 128, 200, 256])))
    return qs, ts, np.array(h0s, dtype=np.int32), np.array(ws, dtype=np.int32)


def pack(seqs):
    off = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=off[1:])
    return (np.concatenate(seqs).astype(np.uint8) if len(seqs) and off[-1] else np.zeros(0, np.uint8)), off


def damage(rs, frac, sub, seed):
    """Wreck one mate of a fraction of pairs so that only mate rescue can place it."""
    g = np.random.default_rng(seed); seq = rs.seq.copy()
    for p in range(rs.n_pairs):
        if g.random() < frac:
            i = 2 * p + int(g.integers(0, 2)); a, b = rs.off[i], rs.off[i + 1]
            m = g.random(b - a) < sub
            seq[a:b][m] = (seq[a:b][m] + g.integers(1, 4, int(m.sum()))) & 3
    rs.seq = seq
    return rs


DATASETS = {
    # name: (reference kwargs, read kwargs)
    "rand1m": (dict(total_len=1_000_000, n_contigs=1, seed=1), dict(n_barcodes=60, seed=11)),
    "rep600k": (dict(total_len=600_000, n_contigs=3, seed=2, repeat_frac=0.3, n_frac=0.002),
                dict(n_barcodes=60, seed=12, sub=0.02, ins=0.002, dele=0.002)),
}


def make_dataset(name, workdir, damaged=False):
    """Reference FASTA + BWA index (built with the reference's own `bwa index` code from oracle/_ref) + simulated reads."""
    from oracle.refbwa import RefBwa
    rk, qk = DATASETS[name]
    ctg = S.make_reference(**rk)
    fa = os.path.join(workdir, name + ".fa")
    if not os.path.exists(fa + ".bwt"):
        S.write_fasta(fa, ctg)
        RefBwa().build_index(fa)
    rs = S.simulate_linked_reads(ctg, **qk)
    if damaged:
        rs = damage(rs, 0.3, 0.11, 5)
    return fa, ctg, rs


# ------------------------------------------------------------------------------------------------ golden fixtures
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name, workdir=None):
    """tests/golden/<name>.npz (made by scripts/make_golden.py from the reference's own code).  With workdir, the BWA
    index files the fixture carries are written to <workdir>/<name>.fa.{bwt,sa,pac,ann,amb} and the prefix returned."""
    d = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    if workdir is None:
        return d
    prefix = os.path.join(workdir, name + ".fa")
    for ext in ("bwt", "sa", "pac", "ann", "amb"):
        p = prefix + "." + ext
        if not os.path.exists(p):
            with open(p, "wb") as f:
                f.write(d["idx_" + ext].tobytes())
    return d, prefix


# ------------------------------------------------------------------------------------------------ host simulation
class HostSim:
    """tests/hostsim/libhostsim.so: the product's thread-sequential device logic (csrc/*.cuh) compiled for the host."""

    def __init__(self):
        d = os.path.join(ROOT, "tests", "hostsim")
        subprocess.run(["make", "-s", "-C", d], check=True, stdout=subprocess.DEVNULL)
        L = self.lib = C.CDLL(os.environ.get("ARACHNE_HOSTSIM_LIB") or os.path.join(d, "libhostsim.so"))  # override: the sanitizer build (scripts/hostsim_sanitize.sh)
        L.hs_load.restype = C.c_void_p; L.hs_load.argtypes = [C.c_char_p]
        L.hs_free.argtypes = [C.c_void_p]
        L.hs_occ.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.hs_sa.restype = C.c_uint64; L.hs_sa.argtypes = [C.c_void_p, C.c_uint64]
        L.hs_seed.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.c_int]
        L.hs_align1.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.hs_pair.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_void_p, C.POINTER(C.c_int),
                              C.c_void_p, C.POINTER(C.c_int), C.c_int]
        L.hs_reg2aln.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.c_int, C.POINTER(C.c_int32)]
        L.hs_introsort_regs.argtypes = [C.c_int, C.c_int, C.c_void_p]
        L.hs_place.restype = C.c_long
        L.hs_place.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_long)]
        L.hs_reg2aln_warp.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.c_int, C.POINTER(C.c_int32)]
        L.hs_pair_rescue_warp.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_void_p, C.POINTER(C.c_int),
                                          C.c_void_p, C.POINTER(C.c_int), C.c_int]
        L.hs_ext_dp.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.c_void_p, C.c_void_p,
                                C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]
        L.hs_cigar_tiers.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.c_void_p, C.POINTER(C.c_uint64)]
        L.hs_seed_index.restype = C.c_void_p; L.hs_seed_index.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int]
        L.hs_seed_index_free.argtypes = [C.c_void_p]
        L.hs_seed2.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.c_int, C.c_int, C.c_int, C.c_int,
                               C.POINTER(C.c_uint64), C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        L.hs_patch_dp_calls.restype = C.c_long
        L.hs_gorand_float64.argtypes = [C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.hs_chain.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_float)]
        self.h = None

    def chain(self, q, do_flt=False, cap=4096):
        """csrc/chain.cuh on the host: chain rows {pos, rid, n, w, kept, first_seed_row}, seed rows {rbeg, qbeg, len, score}, frac_rep."""
        q, qp = u8p(q)
        ch = np.zeros((cap, 6), dtype=np.int64); sd = np.zeros((cap * 4, 4), dtype=np.int64); ns = C.c_int(0); fr = C.c_float(0)
        n = self.lib.hs_chain(self.h, len(q), qp, int(do_flt), ch.ctypes.data_as(C.c_void_p), cap, sd.ctypes.data_as(C.c_void_p), cap * 4, C.byref(ns), C.byref(fr))
        assert 0 <= n <= cap and ns.value <= cap * 4, (n, ns.value)
        return ch[:n], sd[:ns.value], fr.value

    def gorand_float64(self, seed, n, switch_at):
        out = (C.c_double * n)()
        assert self.lib.hs_gorand_float64(seed, n, switch_at, out) == 0
        return list(out)

    def seed_index(self, prefix, kmer_k=-1, direct=True):
        """The HBM-only index structures of the loaded index (k-mer tables of depth kmer_k, -1 = the loader's own rule min(14, log4(n) - 1);
        dense SA + inverse SA + both-strand text when `direct`), built by the product's own builder kernels run on the host."""
        pac = np.fromfile(prefix + ".pac", dtype=np.uint8)
        l_pac = int(open(prefix + ".ann").readline().split()[0])
        if kmer_k < 0:
            n, kmer_k = 2 * l_pac, 0
            while n > 3:
                n >>= 2; kmer_k += 1
            kmer_k = min(kmer_k - 1, 14)
        self._seed_keep = pac
        return self.lib.hs_seed_index(self.h, pac.ctypes.data, l_pac, max(kmer_k, 0), int(direct)), kmer_k

    def seed_index_free(self, si):
        self.lib.hs_seed_index_free(si)

    def seed2(self, si, opt, seq, off, use_kmer=1, use_direct=1, slist_want=-1, lane=0):
        """k_pack_reads -> k_seed2 -> k_sort_intv (csrc/seed_kernel.cuh, the shipped kernels) run one lane at a time on the host:
        -> (intv_off [n + 1], intervals [m, 4] = x0 x1 x2 info, (bwt_extend calls, reference occ blocks))."""
        seq, sp = u8p(seq); off = np.ascontiguousarray(off, dtype=np.int64); n = len(off) - 1
        cap = 64 * n + 1024
        iv = np.zeros((cap, 4), np.uint64); io = np.zeros(n + 1, np.int64); st = (C.c_uint64 * 2)()
        rc = self.lib.hs_seed2(si, C.byref(opt), n, sp, off.ctypes.data_as(C.POINTER(C.c_int64)), use_kmer, use_direct, slist_want, lane,
                               iv.ctypes.data_as(C.POINTER(C.c_uint64)), cap, io.ctypes.data_as(C.POINTER(C.c_int64)), st)
        assert rc == 0, rc
        return io, iv[:io[-1]], (int(st[0]), int(st[1]))

    def stages(self, opt, seq, off, intv_off, intv, rescue=True, score_delta=25, warp=0, wide=False, block=False):
        """SA intervals -> final region lists through the shipped stage kernels in batch.inl's launch order (hostsim_stages.cpp): k_expand, k_sa,
        k_chain, the extension rounds (k_ext_drive / k_ext_binbase / k_ext_scatter / k_ext_dp on the packed reference / k_ext_roll),
        k_ext_finish, k_init_loc, k_rescue, k_gather_regs2.  block: the warp-cooperative kernels as whole blocks (all warps as fibers sharing the queue counters) instead of one warp.
        -> (reg_off [n + 1], regs, info dict)."""
        from arachne_b200.lib import ALNREG
        seq, sp = u8p(seq); off = np.ascontiguousarray(off, dtype=np.int64); n = len(off) - 1
        intv = np.ascontiguousarray(intv, dtype=np.uint64); intv_off = np.ascontiguousarray(intv_off, dtype=np.int64)
        cap = 600 * n + 4096
        regs = np.zeros(cap, dtype=ALNREG); ro = np.zeros(n + 1, np.int64); info = np.zeros(6, np.int64)
        self.lib.hs_stages.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        rc = self.lib.hs_stages(self.h, C.addressof(opt), n, seq.ctypes.data, off.ctypes.data, intv.ctypes.data, intv_off.ctypes.data, 3 if rescue else 1,
                                score_delta, (warp & 3) | (4 if wide else 0) | (8 if block else 0), regs.ctypes.data, cap, ro.ctypes.data, info.ctypes.data)
        assert rc == 0, (rc, info)
        return ro, regs[:ro[-1]].copy(), dict(rounds=int(info[0]), total_sa=int(info[1]), problems=int(info[2]), rescued=int(info[3]))

    def k_cigar(self, opt, seq, off, reg_off, regs, list3=None, warp=0, block=False):
        """The shipped wide-band CIGAR kernel (k_cigar, csrc/pair_kernels.cuh) as a warp of fibers over the regions in list3 (None: all).
        -> (alns [n_regs] valid at the listed regions, cigar pool, mismatch pool)."""
        from arachne_b200.lib import ALN
        seq, sp = u8p(seq); off = np.ascontiguousarray(off, dtype=np.int64); n_reads = len(off) - 1
        regs = np.ascontiguousarray(regs); n = len(regs)
        reg_read = np.repeat(np.arange(n_reads, dtype=np.int32), np.diff(reg_off).astype(np.int64))
        alns = np.zeros(n + 1, dtype=ALN); cig_cap = 64 * n + 4096; mm_cap = 600 * n + 4096
        cig = np.zeros(cig_cap, np.uint32); mm = np.zeros(mm_cap, np.int32); used = (C.c_uint64 * 2)()
        lst = None if list3 is None else np.ascontiguousarray(list3, dtype=np.uint32)
        self.lib.hs_k_cigar.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]
        rc = self.lib.hs_k_cigar(self.h, C.addressof(opt), n_reads, seq.ctypes.data, off.ctypes.data, n, regs.ctypes.data, reg_read.ctypes.data,
                                 None if lst is None else lst.ctypes.data, 0 if lst is None else len(lst), (warp & 3) | (8 if block else 0), alns.ctypes.data, cig.ctypes.data, cig_cap,
                                 mm.ctypes.data, mm_cap, used)
        assert rc == 0, rc
        return alns[:n], cig[:used[0]], mm[:used[1]]

    def cigar_tiers(self, opt, seq, off, reg_off, regs):
        """The shipped lane-per-region CIGAR tiers (csrc/cigar_kernel.cuh) run one lane at a time over every region of the batch.
        -> (alns [n_regs] ALN records, cigar pool, mismatch pool, list3 = regions left to the warp tier, dict of counts)."""
        from arachne_b200.lib import ALN
        seq, sp = u8p(seq); off = np.ascontiguousarray(off, dtype=np.int64); n_reads = len(off) - 1
        regs = np.ascontiguousarray(regs); n = len(regs)
        reg_read = np.repeat(np.arange(n_reads, dtype=np.int32), np.diff(reg_off).astype(np.int64))
        alns = np.zeros(n + 1, dtype=ALN); cig_cap = 64 * n + 4096; mm_cap = 600 * n + 4096
        cig = np.zeros(cig_cap, np.uint32); mm = np.zeros(mm_cap, np.int32); used = (C.c_uint64 * 2)(); counts = (C.c_uint64 * 3)()
        list3 = np.zeros(n + 1, np.uint32)
        rc = self.lib.hs_cigar_tiers(self.h, C.byref(opt), n_reads, sp, off.ctypes.data_as(C.POINTER(C.c_int64)), n, regs.ctypes.data, reg_read.ctypes.data,
                                     alns.ctypes.data, cig.ctypes.data, cig_cap, mm.ctypes.data, mm_cap, used, list3.ctypes.data, counts)
        assert rc == 0, rc
        return alns[:n], cig[:used[0]], mm[:used[1]], list3[:counts[1]].copy(), dict(tier2=int(counts[0]), tier3=int(counts[1]), band_cells=int(counts[2]))

    def ext_dp(self, opt, qs, ts, h0, w, end_bonus=5, zdrop=100, codec=0, warp=0):
        """The shipped extension DP kernel k_ext_dp (csrc/ext_stage.cuh) as one warp of 32 fibers on the host (tests/hostsim/hostsim_extdp.cpp).
        -> (n x 6 array {score, qle, tle, gtle, gscore, max_off}, in-band cells)."""
        q, qo = pack(qs); t, to = pack(ts)
        q, qp = u8p(np.concatenate([q, np.zeros(16, np.uint8)])); t, tp = u8p(np.concatenate([t, np.zeros(16, np.uint8)]))
        h0 = np.ascontiguousarray(h0, dtype=np.int32); w = np.ascontiguousarray(w, dtype=np.int32)
        out = np.zeros((len(qs), 6), np.int32); cells = C.c_uint64(0)
        rc = self.lib.hs_ext_dp(C.byref(opt), len(qs), qp, qo.ctypes.data_as(C.POINTER(C.c_int64)), tp, to.ctypes.data_as(C.POINTER(C.c_int64)), h0.ctypes.data, w.ctypes.data,
                                end_bonus, zdrop, codec, warp, out.ctypes.data, C.byref(cells))
        assert rc == 0, rc
        return out, cells.value

    def xs_static(self, opt, qs, ts, h0, w, end_bonus=5, zdrop=100, max_qlen=None, warp=0, block=False):
        """arachne_ksw_extend_static's launch sequence on the host (csrc/xs_kernels.cuh around k_ext_dp; hostsim_stages.cpp::hs_xs_static): the
        scratch persists between calls as the static device array does.  -> (n x 6, status, in-band cells)."""
        q, qo = pack(qs); t, to = pack(ts)
        q, _ = u8p(np.concatenate([q, np.zeros(16, np.uint8)])); t, _ = u8p(np.concatenate([t, np.zeros(16, np.uint8)]))
        h0 = np.ascontiguousarray(h0, dtype=np.int32); w = np.ascontiguousarray(w, dtype=np.int32)
        n = len(qs); out = np.full((n + 1, 6), 12345, np.int32); cells = C.c_uint64(0); status = C.c_int32(77)
        mq = max_qlen if max_qlen is not None else max([len(x) for x in qs] + [1])
        self.lib.hs_xs_static.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 6 + [C.c_int] * 4 + [C.c_void_p] * 3
        rc = self.lib.hs_xs_static(C.addressof(opt), n, q.ctypes.data, qo.ctypes.data, t.ctypes.data, to.ctypes.data, h0.ctypes.data, w.ctypes.data, mq, end_bonus, zdrop, (warp & 3) | (8 if block else 0),
                                   out.ctypes.data, C.addressof(status), C.addressof(cells))
        assert rc == 0, rc
        assert tuple(out[n]) == (12345,) * 6
        return out[:n], status.value, cells.value

    def k_ksw_global(self, opt, qs, ts, w, warp=0, block=False):
        """arachne_ksw_global_batch's kernel (k_ksw_global: ksw_global2 + CIGAR on the warp context) as fibers.  -> (scores, list of CIGAR lists)."""
        q, qo = pack(qs); t, to = pack(ts)
        q, _ = u8p(np.concatenate([q, np.zeros(16, np.uint8)])); t, _ = u8p(np.concatenate([t, np.zeros(16, np.uint8)]))
        w = np.ascontiguousarray(w, dtype=np.int32); n = len(qs)
        cap = max([len(a) + len(b) + 2 for a, b in zip(qs, ts)] + [4])
        sc = np.zeros(n + 1, np.int32); cig = np.zeros((n + 1, cap), np.uint32); nc = np.zeros(n + 1, np.int32)
        self.lib.hs_k_ksw_global.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        rc = self.lib.hs_k_ksw_global(C.addressof(opt), n, q.ctypes.data, qo.ctypes.data, t.ctypes.data, to.ctypes.data, w.ctypes.data, (warp & 3) | (8 if block else 0),
                                      sc.ctypes.data, cig.ctypes.data, cap, nc.ctypes.data)
        assert rc == 0, rc
        return sc[:n], [list(int(x) for x in cig[i, :nc[i]]) for i in range(n)]

    def k_occ(self, k, l):
        """arachne_occ_batch's kernel (k_occ) on the host image.  -> [n, 8] counts."""
        k = np.ascontiguousarray(k, dtype=np.uint64); l = np.ascontiguousarray(l, dtype=np.uint64); out = np.zeros((len(k), 8), np.uint64)
        self.lib.hs_k_occ.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]; self.lib.hs_k_occ.restype = None
        self.lib.hs_k_occ(self.h, len(k), k.ctypes.data, l.ctypes.data, out.ctypes.data)
        return out

    def pair_warp(self, opt, r1, r2, regs1, regs2, score_delta=25, cap=512):
        """GoBwaMemMateSW's rescue on the device's own warp-cooperative context (csrc/warp_ctx.cuh) as one warp of fibers: the region lists
        of both reads after mem_align1_core in, the lists after rescue out."""
        from arachne_b200.lib import ALNREG
        r1, p1 = u8p(r1); r2, p2 = u8p(r2)
        o1 = np.zeros(cap, dtype=ALNREG); o2 = np.zeros(cap, dtype=ALNREG); o1[:len(regs1)] = regs1; o2[:len(regs2)] = regs2
        n1 = C.c_int(len(regs1)); n2 = C.c_int(len(regs2))
        rc = self.lib.hs_pair_rescue_warp(self.h, C.byref(opt), len(r1), p1, len(r2), p2, score_delta, o1.ctypes.data, C.byref(n1), o2.ctypes.data, C.byref(n2), cap)
        assert rc == 0, rc
        return o1[:n1.value], o2[:n2.value]

    def reg2aln_warp(self, opt, q, reg):
        """reg2aln on the device's warp tier (csrc/cigar.cuh with the WarpCtx of csrc/warp_ctx.cuh) as one warp of fibers."""
        from arachne_b200.lib import ALN
        q, qp = u8p(q); out = np.zeros(1, dtype=ALN); cap = 2 * len(q) + 64; cig = (C.c_uint32 * cap)(); mm = (C.c_int32 * (2 * len(q) + 8))()
        r = np.ascontiguousarray(reg)
        rc = self.lib.hs_reg2aln_warp(self.h, C.byref(opt), len(q), qp, r.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), cig, cap, mm)
        assert rc == 0, rc
        a = out[0]
        return a, np.array(cig[:a["n_cigar"]], dtype=np.uint32), np.array(mm[:2 * a["n_mm"]], dtype=np.int32)

    def all_alns(self, opt, reads, seq, off, reg_off, regs):
        """The CIGAR stage's per-region records for a batch, as the device stages them: the shipped lane tiers (cigar_tiers) + the warp tier's
        logic (reg2aln) for what they leave.  -> ALN array [n_regs] (pool offsets are not meaningful across the two sources)."""
        alns, cig, mm, list3, cnt = self.cigar_tiers(opt, seq, off, reg_off, regs)
        alns = alns.copy()
        reg_read = np.repeat(np.arange(len(off) - 1), np.diff(reg_off).astype(np.int64))
        for g in list3:
            a, _c, _m = self.reg2aln(reads[reg_read[g]], regs[g:g + 1])
            alns[g] = a
        return alns

    def place(self, seq_off, reg_off, regs, alns, bc_off, run_rfa, seeds, improper=-4.0, delta=17, records=None):
        """ARACHNE_STAGE_PLACE by the shipped kernels on the host (csrc/place.cuh, csrc/records.cuh's markDuplicates; k_place_rfa as a warp of
        fibers), staged as csrc/place_host.inl stages them.  -> dict(cand_off, cands, lmp) shaped like the library's result."""
        from arachne_b200.lib import CAND
        seq_off = np.ascontiguousarray(seq_off, np.int64); n_reads = len(seq_off) - 1
        reg_off = np.ascontiguousarray(reg_off, np.uint64); regs = np.ascontiguousarray(regs); alns = np.ascontiguousarray(alns)
        bc_off = np.ascontiguousarray(bc_off, np.int64); run_rfa = np.ascontiguousarray(run_rfa, np.uint8); seeds = np.ascontiguousarray(seeds, np.int64)
        cap = len(regs) + n_reads + 8
        cand_off = np.zeros(n_reads + 1, np.int64); cands = np.zeros(cap, CAND); lmp = np.zeros(max(1, len(bc_off) - 1))
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        if records is None:
            n = self.lib.hs_place(self.h, n_reads, vp(seq_off), vp(reg_off), vp(regs), vp(alns), len(bc_off) - 1, vp(bc_off), vp(run_rfa), vp(seeds), improper, delta,
                                  vp(cand_off), vp(cands), cap, vp(lmp), None, None, None, None, None)
            assert n >= 0, n
            return dict(cand_off=cand_off, cands=cands[:n], lmp=lmp)
        # records = (attach_bx per barcode, the CIGAR pool alns["cigar_off"] points into): also ARACHNE_STAGE_RECORDS
        from arachne_b200.lib import RECORD
        attach = np.ascontiguousarray(records[0], np.uint8); pool = np.ascontiguousarray(records[1], np.uint32)
        rec_off = np.zeros(n_reads + 1, np.int64); recs = np.zeros(2 * n_reads + 1, RECORD); nrec = C.c_long(0)
        n = self.lib.hs_place(self.h, n_reads, vp(seq_off), vp(reg_off), vp(regs), vp(alns), len(bc_off) - 1, vp(bc_off), vp(run_rfa), vp(seeds), improper, delta,
                              vp(cand_off), vp(cands), cap, vp(lmp), vp(attach), vp(pool), vp(rec_off), vp(recs), C.byref(nrec))
        assert n >= 0, n
        return dict(cand_off=cand_off, cands=cands[:n], lmp=lmp, rec_off=rec_off, records=recs[:nrec.value])

    def introsort_regs(self, mode, regs):
        out = regs.copy()
        self.lib.hs_introsort_regs(mode, len(out), out.ctypes.data_as(C.c_void_p))
        return out

    def load(self, prefix):
        self.h = L = self.lib.hs_load(prefix.encode())
        if not L:
            raise RuntimeError("hs_load failed: " + prefix)
        return self

    def occ(self, k, l):
        a = (C.c_uint64 * 4)(); b = (C.c_uint64 * 4)()
        self.lib.hs_occ(self.h, C.c_uint64(int(k) & (2**64 - 1)), C.c_uint64(int(l) & (2**64 - 1)), a, b)
        return list(a) + list(b)

    def sa(self, k):
        return self.lib.hs_sa(self.h, int(k))

    def seed(self, q, cap=1024):
        q, qp = u8p(q); out = (C.c_uint64 * (4 * cap))()
        n = self.lib.hs_seed(self.h, len(q), qp, out, cap)
        return np.array(out[:4 * max(n, 0)], dtype=np.uint64).reshape(-1, 4)

    def pair(self, r1, r2, score_delta=25, cap=512):
        from arachne_b200.lib import ALNREG
        r1, p1 = u8p(r1); r2, p2 = u8p(r2)
        o1 = np.zeros(cap, dtype=ALNREG); o2 = np.zeros(cap, dtype=ALNREG); n1 = C.c_int(0); n2 = C.c_int(0)
        rc = self.lib.hs_pair(self.h, len(r1), p1, len(r2), p2, score_delta, o1.ctypes.data_as(C.c_void_p), C.byref(n1),
                              o2.ctypes.data_as(C.c_void_p), C.byref(n2), cap)
        assert rc == 0, rc
        return o1[:n1.value], o2[:n2.value]

    def align1(self, q, cap=512):
        from arachne_b200.lib import ALNREG
        q, qp = u8p(q); out = np.zeros(cap, dtype=ALNREG); cnt = (C.c_uint64 * 3)()
        n = self.lib.hs_align1(self.h, len(q), qp, out.ctypes.data_as(C.c_void_p), cap, cnt)
        assert n >= 0, n
        return out[:n]

    def reg2aln(self, q, reg):
        from arachne_b200.lib import ALN
        q, qp = u8p(q); out = np.zeros(1, dtype=ALN); cap = 2 * len(q) + 64; cig = (C.c_uint32 * cap)(); mm = (C.c_int32 * (2 * len(q) + 8))()
        r = np.ascontiguousarray(reg)
        rc = self.lib.hs_reg2aln(self.h, len(q), qp, r.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), cig, cap, mm)
        assert rc == 0, rc
        a = out[0]
        return a, np.array(cig[:a["n_cigar"]], dtype=np.uint32), np.array(mm[:2 * a["n_mm"]], dtype=np.int32)


class CompatMockBackend:
    """tests/hostsim/libcompat_mock.so: csrc/compat.cu (the gobwa-shaped layer, host code) over batch entry points served by the reference's
    own C code.  Quacks like ArachneCuda as far as arachne_b200.lib.Compat needs (lib, idx, opt, _ck)."""

    def __init__(self, prefix):
        from arachne_b200.lib import MemOpt, ArachneError
        d = os.path.join(ROOT, "tests", "hostsim")
        if not os.environ.get("ARACHNE_COMPAT_MOCK_LIB"):
            subprocess.run(["make", "-s", "-C", d, "libcompat_mock.so"], check=True, stdout=subprocess.DEVNULL)
        L = self.lib = C.CDLL(os.environ.get("ARACHNE_COMPAT_MOCK_LIB") or os.path.join(d, "libcompat_mock.so"))
        L.arachne_last_error.restype = C.c_char_p
        L.mock_index_open.restype = C.c_void_p; L.mock_index_open.argtypes = [C.c_char_p]
        L.mock_index_close.argtypes = [C.c_void_p]
        L.arachne_opts_default.argtypes = [C.POINTER(MemOpt)]
        L.mock_fail_next_batches.argtypes = [C.c_int]
        self._err = ArachneError
        self.idx = C.c_void_p(L.mock_index_open(prefix.encode()))
        assert self.idx
        self.opt = MemOpt(); assert L.arachne_opts_default(C.byref(self.opt)) == 0

    def _ck(self, rc):
        if rc != 0:
            raise self._err("%d: %s" % (rc, self.lib.arachne_last_error().decode()))

    def close(self):
        self.lib.mock_index_close(self.idx)


def orc_local(O, q, t, xtra):
    q, qp = u8p(q); t, tp = u8p(t); out = (C.c_int * 7)()
    O.orc_local(C.byref(DEFAULT_SC), len(q), qp, len(t), tp, xtra, out)
    return tuple(out)


def ALNREG_DT():
    from arachne_b200.lib import ALNREG
    return ALNREG


# ------------------------------------------------------------------------------------------------ placement oracle
class OrcRef(C.Structure):
    _fields_ = [("l_pac", C.c_int64), ("n_seqs", C.c_int32), ("ctg_off", C.POINTER(C.c_int64)), ("ctg_len", C.POINTER(C.c_int32)),
                ("pac", C.POINTER(C.c_uint8)), ("centro_start", C.POINTER(C.c_int64)), ("centro_end", C.POINTER(C.c_int64))]


CAND = np.dtype([("reg", "<i4"), ("read", "<i4"), ("id", "<i4"), ("active", "<i4"), ("is_proper", "<i4"), ("bwa_pick", "<i4"), ("mapq", "<i4"),
                 ("molecule_id", "<i4"), ("active_molecule", "<i4"), ("second_best", "<i4"), ("mate", "<i4"), ("duplicate", "<i4"),
                 ("pos", "<i8"), ("aend", "<i8"), ("lap", "<f8"), ("as_score", "<f8"), ("xs_score", "<f8"), ("molecule_difference", "<f8"),
                 ("sum_move", "<f8"), ("molecule_confidence", "<f8"), ("raw_mapq", "<f8")])
assert CAND.itemsize == 120


def read_ann(prefix):
    """(l_pac, [(name, offset, len)]) from a bwa .ann file (bntseq.c:93-130)."""
    with open(prefix + ".ann") as f:
        l_pac, n, _ = f.readline().split()
        ctg = []
        for _ in range(int(n)):
            name = f.readline().split()[1]
            off, ln, _na = f.readline().split()
            ctg.append((name, int(off), int(ln)))
    return int(l_pac), ctg


class PlaceOracle:
    def __init__(self, prefix, centromeres=None):
        self.O = oracle_lib()
        self.l_pac, self.ctg = read_ann(prefix)
        self.pac = np.fromfile(prefix + ".pac", dtype=np.uint8)
        self.coff = np.array([c[1] for c in self.ctg], dtype=np.int64); self.clen = np.array([c[2] for c in self.ctg], dtype=np.int32)
        self.ref = OrcRef(self.l_pac, len(self.ctg), self.coff.ctypes.data_as(C.POINTER(C.c_int64)), self.clen.ctypes.data_as(C.POINTER(C.c_int32)),
                          self.pac.ctypes.data_as(C.POINTER(C.c_uint8)), None, None)
        if centromeres is not None:
            self.cs = np.ascontiguousarray(centromeres[0], dtype=np.int64); self.ce = np.ascontiguousarray(centromeres[1], dtype=np.int64)
            self.ref.centro_start = self.cs.ctypes.data_as(C.POINTER(C.c_int64)); self.ref.centro_end = self.ce.ctypes.data_as(C.POINTER(C.c_int64))
        self.O.orc_place.argtypes = [C.POINTER(OrcRef), C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        self.O.orc_name_seed.restype = C.c_int64; self.O.orc_name_seed.argtypes = [C.c_char_p, C.c_int]

    def name_seed(self, name: str) -> int:
        b = name.encode()
        return self.O.orc_name_seed(b, len(b))

    def place(self, seq, off, bc_off, run_rfa, seeds, reg_off, regs, aln8, cigar, improper=-4.0, delta=17):
        seq = np.ascontiguousarray(seq, np.uint8); off = np.ascontiguousarray(off, np.int64); bc_off = np.ascontiguousarray(bc_off, np.int64)
        run_rfa = np.ascontiguousarray(run_rfa, np.uint8); seeds = np.ascontiguousarray(seeds, np.int64)
        reg_off = np.ascontiguousarray(reg_off, np.int64); regs = np.ascontiguousarray(regs); aln8 = np.ascontiguousarray(aln8, np.int64)
        cigar = np.ascontiguousarray(cigar, np.uint32)
        n_pairs = (len(off) - 1) // 2; cap = len(regs) + 2 * n_pairs + 8
        cand_off = np.zeros(2 * n_pairs + 1, np.int64); cands = np.zeros(cap, CAND); lmp = np.zeros(max(1, len(bc_off) - 1))
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        rc = self.O.orc_place(C.byref(self.ref), n_pairs, vp(seq), vp(off), len(bc_off) - 1, vp(bc_off), vp(run_rfa), vp(seeds), vp(reg_off), vp(regs),
                              vp(aln8), vp(cigar), improper, delta, vp(cand_off), vp(cands), cap, vp(lmp))
        assert rc == 0, rc
        return cand_off, cands[:cand_off[-1]], lmp


def _oracle_records(self, seq, off, bc_off, run_rfa, attach_bx, seeds, reg_off, regs, aln8, cigar, improper=-4.0, delta=17):
    """orc_place_records: placement + markDuplicates + CheckSplitReads + AppendBam's fields (oracle/restate_place.cpp)."""
    from arachne_b200.lib import RECORD
    seq = np.ascontiguousarray(seq, np.uint8); off = np.ascontiguousarray(off, np.int64); bc_off = np.ascontiguousarray(bc_off, np.int64)
    run_rfa = np.ascontiguousarray(run_rfa, np.uint8); seeds = np.ascontiguousarray(seeds, np.int64)
    attach_bx = None if attach_bx is None else np.ascontiguousarray(attach_bx, np.uint8)
    reg_off = np.ascontiguousarray(reg_off, np.int64); regs = np.ascontiguousarray(regs); aln8 = np.ascontiguousarray(aln8, np.int64)
    cigar = np.ascontiguousarray(cigar, np.uint32)
    n_pairs = (len(off) - 1) // 2; cap = len(regs) + 2 * n_pairs + 8
    cand_off = np.zeros(2 * n_pairs + 1, np.int64); cands = np.zeros(cap, CAND); lmp = np.zeros(max(1, len(bc_off) - 1))
    recs = np.zeros(4 * n_pairs + 8, RECORD); n_recs = C.c_int64(0)
    vp = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    self.O.orc_place_records.argtypes = [C.POINTER(OrcRef), C.c_int64] + [C.c_void_p] * 2 + [C.c_int64] + [C.c_void_p] * 8 + [C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
    rc = self.O.orc_place_records(C.byref(self.ref), n_pairs, vp(seq), vp(off), len(bc_off) - 1, vp(bc_off), vp(run_rfa), vp(attach_bx), vp(seeds), vp(reg_off),
                                  vp(regs), vp(aln8), vp(cigar), improper, delta, vp(cand_off), vp(cands), cap, vp(lmp), vp(recs), len(recs), C.byref(n_recs))
    assert rc == 0, rc
    return cand_off, cands[:cand_off[-1]], lmp, recs[:n_recs.value]


PlaceOracle.place_records = _oracle_records


def sam_lines(recs, contig_names, seq, off, bc_off, names, quals, barcodes, read_group, aln_of_reg, cigar, mm_of_reg):
    """Independent (pure Python) SAM formatter for record fields, used on the ORACLE's records: field and tag order of AppendBam
    (bamwriter.go:283-568).  aln_of_reg(g) -> (n_cigar, cigar_off); mm_of_reg(g) -> (ref loci, read offsets)."""
    dec, cmpl = "ACGTN", "TGCAN"
    out = []
    def cig(g, hard, rev):
        n, o = aln_of_reg(g)
        ws = [int(cigar[o + k]) for k in range(n)]
        if rev:
            ws = ws[::-1]
        return "".join("%d%s" % (w >> 4, "MIDSH"[4 if (w & 15) == 3 and hard else (w & 15)]) for w in ws)
    def mml(g):
        if g < 0:
            return ""
        a, b = mm_of_reg(g)
        return "".join("%d,%d,1;" % (int(x), int(y)) for x, y in zip(a, b))
    for R in recs:
        r = int(R["read"]); pair = r >> 1
        b = int(np.searchsorted(bc_off, pair, side="right") - 1)
        q = seq[off[r]:off[r + 1]]; L = len(q)
        rng = range(int(R["clip_start"]), int(R["clip_end"]))
        rev = bool(R["reversed"])
        s = "".join((cmpl if rev else dec)[min(int(q[L - 1 - k] if rev else q[k]), 4)] for k in rng) or "*"
        ql = quals[r] if quals is not None else None
        qs = ("".join((ql[L - 1 - k] if rev else ql[k]) for k in rng) if ql is not None and len(rng) else "*")
        g = int(R["reg"])
        f = [names[pair].rstrip("\n"), str(int(R["flag"])), contig_names[R["rid"]] if R["rid"] >= 0 else "*", str(int(R["pos"]) + 1), str(int(R["mapq"])),
             (cig(g, bool(R["hard"]), False) if g >= 0 and aln_of_reg(g)[0] > 0 else "*"),
             "*" if R["mate_rid"] < 0 else ("=" if R["mate_rid"] == R["rid"] else contig_names[R["mate_rid"]]), str(int(R["mate_pos"]) + 1), str(int(R["tlen"])), s, qs]
        if read_group:
            f.append("RG:Z:" + read_group)
        f += ["XS:i:%d" % R["xs"], "XC:Z:" + (mml(int(R["xc_reg"])) if R["has_xc"] else ""), "AC:Z:" + mml(g), "AS:i:%d" % R["as"], "XM:Z:%d" % R["xm"], "AM:Z:%d" % R["am"],
              "XT:i:%d" % R["xt"]]
        if R["sa_reg"] >= 0:
            f.append("SA:Z:%s,%d,%s,%s,%d,%d;" % (contig_names[R["sa_rid"]], R["sa_pos"], "-" if R["sa_rev"] else "+", cig(int(R["sa_reg"]), bool(R["sa_hard"]), bool(R["sa_rev"])),
                                                 R["sa_mapq"], R["sa_nm"]))
        bc = barcodes[b] if barcodes is not None else None
        if R["attach_bx"] and bc is not None and "-" in bc:
            f += ["BX:Z:" + bc, "VX:i:1"]
            if R["has_dm"]:
                f.append("DM:Z:%.6f" % R["dm"])
        out.append("\t".join(f))
    return out


def place_mismatches(res, oc, ocands, olmp):
    """Compare the library's placement outputs with the oracle's.  Integers exactly; floating point: pair scores are sums of
    dyadic rationals (exact), the molecule penalty goes through log10 and MAPQ through pow/log10, which differ from the host
    libm in the last bits -> 1e-9 relative on scores, MAPQ equal unless the un-truncated value sits within 1e-6 of an
    integer (then +-1).  Returns a list of human-readable mismatches (empty = parity)."""
    bad = []
    g = res["cands"]
    if not np.array_equal(res["cand_off"], oc) or len(g) != len(ocands):
        return ["candidate lists differ (%d vs %d)" % (len(g), len(ocands))]
    for f in ("reg", "read", "id", "pos", "aend", "active", "is_proper", "bwa_pick", "molecule_id", "active_molecule", "duplicate"):
        d = np.nonzero(g[f] != ocands[f])[0]
        if len(d):
            bad.append("%s: %d differ, first at %d (%s vs %s)" % (f, len(d), d[0], g[f][d[0]], ocands[f][d[0]]))
    if not np.array_equal(g["lap"], ocands["lap"]):
        bad.append("log_alignment_probability differs")
    act = ocands["active"] == 1
    for f in ("second_best", "mate"):
        if not np.array_equal(g[f][act], ocands[f][act]):
            bad.append(f + " differs")
    for f in ("as_score", "xs_score", "molecule_difference", "molecule_confidence"):
        if not np.allclose(g[f][act], ocands[f][act], rtol=1e-9, atol=1e-9):
            bad.append(f + " beyond 1e-9")
    if not np.allclose(g["sum_move"][act], ocands["sum_move"][act], rtol=1e-9):
        bad.append("sum_move beyond 1e-9")
    if not np.allclose(res["lmp"], olmp, rtol=1e-12, atol=1e-12):
        bad.append("log_molecule_penalty beyond 1e-12")
    dm = g["mapq"][act] != ocands["mapq"][act]
    if dm.any():
        raw = ocands["raw_mapq"][act][dm]
        if not (np.all(np.abs(raw - np.round(raw)) < 1e-6) and np.all(np.abs(g["mapq"][act][dm] - ocands["mapq"][act][dm]) <= 1)):
            bad.append("mapq: %d differ beyond the integer-boundary tolerance" % int(dm.sum()))
    if not np.allclose(g["raw_mapq"][act], ocands["raw_mapq"][act], rtol=1e-9, atol=1e-9, equal_nan=True):
        bad.append("raw_mapq beyond 1e-9")
    return bad


def exact_repeat_reference(total_len=300_000, n_contigs=4, seed=9, families=6, copies=24, rep_len=700):
    """Random contigs with families of EXACT repeat copies: reads inside a copy get many equal-score candidates, so
    tagBestAlignments' PRNG tie-break decides, pairs need hundreds of PRNG outputs, and the optimizer has moves to make."""
    ctg = S.make_reference(total_len=total_len, n_contigs=n_contigs, seed=seed)
    g = np.random.default_rng(seed + 1000)
    for _ in range(families):
        ci = int(g.integers(0, len(ctg))); s = ctg[ci][1]
        a = int(g.integers(0, len(s) - rep_len)); seg = s[a:a + rep_len].copy()
        for _ in range(copies):
            cj = int(g.integers(0, len(ctg))); t = ctg[cj][1]
            b = int(g.integers(0, len(t) - rep_len)); t[b:b + rep_len] = seg
    return ctg


# ------------------------------------------------------------------------------------------------ index construction
def bwt_payload_words(n):
    return (n + 15) // 16 + ((n + 127) // 128 + 1) * 8


def hostsim_index_build(pac_bytes, l_pac, batch_cap=0):
    """The product's index-construction driver (csrc/index_build.cuh) on the TEST-ONLY host backend -> the bytes of
    <prefix>.bwt and <prefix>.sa as bwt_dump_bwt / bwt_dump_sa lay them out (bwa/bwt.c:385-409) + driver statistics."""
    HostSim()  # builds libhostsim.so
    H = C.CDLL(os.environ.get("ARACHNE_HOSTSIM_LIB") or os.path.join(ROOT, "tests", "hostsim", "libhostsim.so"))
    H.hostsim_index_build.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    n = 2 * l_pac
    buf = np.zeros(l_pac // 4 + 16, np.uint8)
    k = min(len(pac_bytes), l_pac // 4 + 1)
    buf[:k] = np.frombuffer(bytes(pac_bytes[:k]), np.uint8)
    out = np.zeros(bwt_payload_words(n), np.uint32); sa = np.zeros((n + 32) // 32, np.uint64); meta = np.zeros(8, np.uint64)
    rc = H.hostsim_index_build(buf.ctypes.data, l_pac, batch_cap, out.ctypes.data, sa.ctypes.data, meta.ctypes.data)
    assert rc == 0
    bwt = meta[:5].tobytes() + out.tobytes()
    saf = meta[:5].tobytes() + np.array([32, n], np.uint64).tobytes() + sa[1:].tobytes()
    return bwt, saf, dict(n_batches=int(meta[5]), tied=int(meta[6]), rounds=int(meta[7]))


def index_build_cases():
    """(name, contigs, batch_cap): references that stress the suffix sorter -- iid text, batches forced small, diverged and
    exact repeats, homopolymer runs (deep tie refinement, the '$'-is-smallest rule at the text end), tandem repeats,
    lengths around every packing boundary (4 / 16 / 32 / 128 bases), single-letter texts."""
    g = np.random.default_rng(5)
    cases = [("rand10k", S.make_reference(10007, 1, seed=1), 0), ("rand10k-cap500", S.make_reference(10007, 1, seed=1), 500),
             ("rep60k", S.make_reference(60000, 1, seed=2, repeat_frac=0.3), 3000), ("multi100k", S.make_reference(100000, 3, seed=3), 0),
             ("exactrep", exact_repeat_reference(120_000, 2, 9, 3, 8, 700), 20000)]
    s = g.integers(0, 4, 5000, dtype=np.uint8); s[1000:1400] = 0; s[-50:] = 0; s[:37] = 3
    cases.append(("polyA", [("c", s)], 0))
    cases.append(("tandem", [("c", np.tile(g.integers(0, 4, 7, dtype=np.uint8), 300))], 0))
    for L in (1, 2, 3, 4, 5, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 129, 255, 256, 257):
        cases.append((f"L{L}", [("c", g.integers(0, 4, L, dtype=np.uint8))], 0))
    cases.append(("allA", [("c", np.zeros(300, np.uint8))], 0))
    cases.append(("allT", [("c", np.full(300, 3, np.uint8))], 0))
    return cases


def ref_collect_intv(R, q, min_seed_len=19, split_len=28, split_width=10, max_mem_intv=20):
    """mem_collect_intv (bwamem.c:114-162) composed from the reference's own bwt_smem1 / bwt_seed_strategy1 (oracle/_ref),
    sorted by info: the array k_seed2 + k_sort_intv must produce for the read."""
    L = len(q); exp = []
    x = 0
    while x < L:
        if q[x] < 4:
            x, iv = R.smem1(q, x)
            exp += [tuple(int(v) for v in a) for a in iv if (int(a[3]) & 0xffffffff) - (int(a[3]) >> 32) >= min_seed_len]
        else:
            x += 1
    for a in list(exp):
        s, e = a[3] >> 32, a[3] & 0xffffffff
        if e - s < split_len or a[2] > split_width:
            continue
        _, iv = R.smem1(q, (s + e) >> 1, a[2] + 1)
        exp += [tuple(int(v) for v in b) for b in iv if (int(b[3]) & 0xffffffff) - (int(b[3]) >> 32) >= min_seed_len]
    x = 0
    while x < L:
        if q[x] < 4:
            x, m = R.seed_strategy1(q, x, min_seed_len, max_mem_intv)
            if m[2] > 0:
                exp.append(tuple(int(v) for v in m))
        else:
            x += 1
    exp.sort(key=lambda a: a[3])
    return exp
