#!/usr/bin/env python
"""Config 5 of BASELINE.json: ksw_extend2 banded-SW microbench, band 16-256 x read length 100-250, GCUPS on the B200 next to
the reference's own ksw_extend2 (oracle/_ref, one host thread).

A problem = one right extension as mem_chain2aln issues it (bwamem.c:746-771): query = the read flank beyond a seed,
target = the reference window under it (1 % substitutions, 14 % of problems carry a 1-3 base indel) plus the slack the caller's
max-gap estimate adds, h0 = seed length, end_bonus 5, zdrop 100.  GCUPS counts in-band cells exactly as ksw.c:424-448 visits
them (the library's count is asserted equal to the oracle's in tests/test_gpu_parity.py).

  python scripts/ksw_microbench.py [--n 131072] [--cpu-n 4000] [--out profiles/r1_ksw_microbench.json]
"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import ArachneCuda


def problems(n, L, seed):
    g = np.random.default_rng(seed)
    seedlen = g.integers(19, max(20, L // 3), n)
    qlen = L - seedlen                                  # flank beyond the seed
    Q = g.integers(0, 4, (n, L), dtype=np.uint8)
    T = Q.copy()
    m = g.random((n, L)) < 0.01
    T[m] = (T[m] + g.integers(1, 4, int(m.sum()), dtype=np.uint8)) & 3
    qs, ts = [], []
    ind = g.random(n) < 0.14
    for i in range(n):
        q = Q[i, :qlen[i]]; t = T[i, :qlen[i]]
        if ind[i] and qlen[i] > 30:
            p = int(g.integers(10, qlen[i] - 10)); l = int(g.integers(1, 4))
            t = np.concatenate([t[:p], t[p + l:]]) if g.random() < 0.5 else np.concatenate([t[:p], g.integers(0, 4, l).astype(np.uint8), t[p:]])
        slack = int(min(qlen[i] - 5, 100)) if qlen[i] > 5 else 0  # cal_max_gap-like extra target
        ts.append(np.concatenate([t, g.integers(0, 4, max(slack, 0)).astype(np.uint8)])); qs.append(q)
    return qs, ts, seedlen.astype(np.int32)


def pack(seqs):
    off = np.zeros(len(seqs) + 1, np.int64); np.cumsum([len(s) for s in seqs], out=off[1:])
    return np.concatenate(seqs).astype(np.uint8), off


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 17); ap.add_argument("--cpu-n", type=int, default=4000)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r1_ksw_microbench.json"))
    args = ap.parse_args()
    A = ArachneCuda()
    assert A.device_count() > 0, "needs a CUDA device"
    R = None
    try:
        from oracle.refbwa import RefBwa, available
        R = RefBwa() if available() else None
    except Exception:
        pass
    rows = []
    for L in (100, 150, 200, 250):
        qs, ts, h0 = problems(args.n, L, 15 + L)
        q, qo = pack(qs); t, to = pack(ts)
        for w in (16, 32, 64, 100, 128, 200, 256):
            wv = np.full(args.n, w, np.int32)
            out, cells, ms = A.ksw_extend(q, qo, t, to, h0, wv, end_bonus=5, zdrop=100, repeat=5)
            row = {"read_len": L, "w": w, "problems": args.n, "cells": int(cells), "gpu_ms": ms, "gpu_gcups": cells / ms / 1e6}
            if R is not None:
                k = min(args.cpu_n, args.n); c0 = time.perf_counter()
                for i in range(k):
                    r = R.ksw_extend2(qs[i], ts[i], w, 5, 100, int(h0[i]))
                    assert tuple(out[i]) == tuple(r), (L, w, i, tuple(out[i]), r)   # parity on every timed CPU problem
                dt = time.perf_counter() - c0
                # cells of the CPU subset: same per-problem distribution, scale by count
                row.update({"cpu_problems": k, "cpu_s": dt, "cpu_gcups_1thread": cells * (k / args.n) / dt / 1e9, "cpu_note": "ctypes call overhead included"})
            rows.append(row)
            print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
