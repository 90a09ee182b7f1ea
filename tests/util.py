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
    return C.CDLL(path)


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
