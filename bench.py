#!/usr/bin/env python
"""bench.py -- read pairs/s of the candidate-alignment hot path (seeding -> chaining -> extension -> mate rescue ->
CIGAR) through libarachne_cuda, with the reference's own CPU code (oracle/_ref) timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, barcode-sharded, no collective)

A step = one pass of the hot path over one batch of `--pairs` simulated 2x150 read pairs grouped by barcode.
`value`   : pairs/s with the batch already resident in HBM and the results left in HBM (CUDA events, max over ranks).
`e2e`     : pairs/s through the C-ABI call a host makes: pinned host buffers in, H2D + kernels + D2H of every result.
`roofline`: the seeding kernel's algorithmic occ-block bytes / its CUDA-event time against the measured HBM copy peak.
`cpu_baseline`: the reference's C code (GoBwaMemMateSW + GoBwaSmithWaterman per candidate, oracle/_ref) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from arachne_b200 import simulate as S  # noqa: E402

STAGES = 15  # regions + rescue + cigar + placement (tagBestAlignments .. optimizer .. MAPQ)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="arachne_b200", choices=["arachne_b200", "reference"])
    ap.add_argument("--config", default="haplotag", choices=["haplotag", "stlfr", "tellseq", "ksw"],
                    help="workload (BASELINE.json configs): haplotag = configs[1] (default, what BENCH/SCALE record), stlfr = configs[2] "
                         "(lognormal pairs per barcode, mean 200, one over-cap barcode per batch), tellseq = configs[3] (30 %% invalid barcodes), "
                         "ksw = configs[4] (ksw_extend2 microbench grid, GCUPS)")
    ap.add_argument("--ref-mb", type=float, default=float(os.environ.get("ARACHNE_BENCH_REF_MB", "3100")), help="synthetic reference size, Mbp (configs[1]: 3100)")
    ap.add_argument("--pairs", type=int, default=int(os.environ.get("ARACHNE_BENCH_PAIRS", "1000000")), help="read pairs per step per GPU")
    ap.add_argument("--batches", type=int, default=2, help="distinct input batches cycled over the steps")
    ap.add_argument("--cpu-sample", type=int, default=300000, help="pairs in the bounded CPU-baseline sample")
    ap.add_argument("--index-builder", default="device", choices=["device", "reference"],
                    help="data preparation: arachne_index_build on the GPU (byte-identical files, tests/test_gpu_parity.py::test_index_build_parity) or the reference's single-threaded `bwa index`")
    ap.add_argument("--workdir", default=os.environ.get("ARACHNE_BENCH_DIR", "/tmp/arachne_bench"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--plain-barcodes", action="store_true", help="barcodes without the '-1' suffix: worthRunningRFA is false, the optimizer is skipped (what the reference does with haplotagging barcodes as shipped)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- dataset
def dataset(args, rank: int, world: int, need_index_files: bool = True):
    """Reference (seed 2, human-like contig proportions) + its index + per-rank read batches (seed 12 + rank).
    Data preparation, outside every timed region.  Rank 0 writes the index files once per box (a marker file lets the
    reference arm and the N = 1, 2, 4, 8 runs of one round share them); the other ranks wait on the marker."""
    os.makedirs(args.workdir, exist_ok=True)
    L = int(args.ref_mb * 1e6)
    n_ctg = 24 if L >= 24_000_000 else 1
    t = time.time()
    ctg = S.make_reference(L, n_ctg, seed=2)
    log(f"rank {rank}: reference of {L} bases generated in {time.time() - t:.1f}s")
    prefix = os.path.join(args.workdir, f"ref{L}")
    done = prefix + ".done"
    if rank == 0 and not os.path.exists(done):
        t = time.time()
        build_index(prefix, ctg, args)
        open(done, "w").write("ok")
        log(f"index files written in {time.time() - t:.1f}s")
    while not os.path.exists(done):
        time.sleep(0.5)
    return prefix, ctg


def build_index(prefix: str, ctg, args):
    """Data preparation, not the measured path.  `device`: the packed reference is written as the reference's pack step
    writes it (simulate.write_bns, pinned to bns_fasta2bntseq in the CPU tests) and the BWT + sampled SA come from
    arachne_index_build (bwtindex.c:61-173 + bwt.c:62-84 on the GPU; byte-identical files).  `reference`: FASTA through the
    reference's own `bwa index` (oracle/_ref), single-threaded -- minutes per 100 Mb, hours at 3.1 Gb."""
    if args.index_builder == "reference":
        from oracle.refbwa import RefBwa
        S.write_fasta(prefix, ctg)
        RefBwa().build_index(prefix)
        return
    from arachne_b200 import ArachneCuda
    S.write_bns(prefix, ctg)
    info = ArachneCuda().build_index(prefix, int(os.environ.get("LOCAL_RANK", "0")), 0)
    log("arachne_index_build: " + json.dumps(info))


WORKLOADS = {
    # --config: (simulate_workload style, seed base (12 / 13 / 14), extra kwargs, description)
    "haplotag": ("haplotagging", 12, {}, "simulated haplotagging 2x150 bp pairs: molecules/barcode ~ Poisson(3)+1, pairs/molecule ~ Poisson(8), "
                                         "1% subs, 0.05%/base ins, 0.05%/base del (1-3 bp), 0.2% N"),
    "stlfr": ("stlfr", 13, {"oversize_pairs": 30500}, "simulated stLFR 2x150 bp pairs: pairs/barcode ~ lognormal(mean 200, sigma 1), one molecule per ~20 pairs, "
                                                      "one 30500-pair barcode per batch (reader cap 30000 + 201-record continuation units), same error model"),
    "tellseq": ("tellseq", 14, {"invalid_frac": 0.3}, "simulated TELLseq 2x150 bp pairs: 18-mer barcodes, 30% of barcodes invalid (N, VX:i:0: random positions, no RFA), "
                                                      "same molecule and error model as haplotagging"),
}


def make_batch(ctg, args, seed, n_pairs):
    """One batch of the --config workload, cut into the reference reader's work units (reader.go:208-291): ReadSet.bc_off
    becomes the work-unit offsets, .unique / .run_rfa the per-unit flags (aligner.go:1018-1030)."""
    from arachne_b200 import workunits as W
    style, seed0, kw, _ = WORKLOADS[args.config]
    rs = S.simulate_workload(ctg, n_pairs, seed=seed0 + seed, style=style, dash_suffix=not args.plain_barcodes, **kw)
    uo, uq, ub = W.work_units(rs.bc_off, rs.barcodes)
    rs.run_rfa = W.worth_running_rfa(uo, uq, ub, rs.barcodes)
    rs.barcodes = [rs.barcodes[b] for b in ub.tolist()]
    rs.bc_off = uo
    return rs


def make_batches(ctg, args, rank):
    return [make_batch(ctg, args, 1000 * rank + b, args.pairs) for b in range(args.batches)]


def subset_units(rs, lo, hi):
    sub = rs.subset(lo, hi)
    sub.run_rfa = rs.run_rfa[lo:hi].copy()
    return sub


def place_inputs(b, args):
    """Per-batch inputs of the placement stage: PRNG seeds = md5(read name)[0:8] (aligner.go:1413-1416) and the
    worthRunningRFA flag per work unit (aligner.go:1018-1030: unique barcode with a '-', >= 5 pairs)."""
    import hashlib
    seeds = np.fromiter((int.from_bytes(hashlib.md5(n.encode()).digest()[:8], "little", signed=True) for n in b.names), dtype=np.int64, count=b.n_pairs)
    return seeds, b.run_rfa


def log(msg):
    print(f"[bench] {msg}", file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev: int):
        self.dev, self.rows, self.p = dev, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.dev), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        sm = [int(r[0]) for r in self.rows if r and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None, "reasons": reasons, "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------- reference arm / cpu baseline
_REF_CACHE = {}


def cpu_reference(fa, rs, n_sample, threads, args):
    """The reference's CPU implementation of the path on `threads` host threads: the C part is the reference's own code
    (oracle/_ref: unmodified bwa sources driven as gobwa.go drives them, pthreads over pairs); the Go part (GetAlignments
    walk, tagBestAlignments .. optimizer .. MAPQ) cannot be built here (no Go toolchain) and is timed as the oracle's C++
    restatement, one worker thread per chunk of barcodes like aligner.WorkerThread (aligner.go:405-419)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.refbwa import RefBwa
    from tests import util
    if fa not in _REF_CACHE:
        _REF_CACHE[fa] = (RefBwa().load(fa), util.PlaceOracle(fa))
    R, P = _REF_CACHE[fa]
    n_bc = int(np.searchsorted(rs.bc_off, min(n_sample, rs.n_pairs), side="left"))
    sub = subset_units(rs, 0, max(1, n_bc))
    seeds, rfa = place_inputs(sub, args)
    t = time.perf_counter()
    ref = R.batch(sub.seq, sub.off, rescue=True, cigar=True, threads=threads)
    t1 = time.perf_counter()
    nb = len(sub.bc_off) - 1
    # chunks of work units balanced by pair count (a 30000-pair unit must not drag a whole thread's share behind it)
    tgt = [sub.n_pairs * k // threads for k in range(threads + 1)]
    cuts = sorted(set(int(np.searchsorted(sub.bc_off, x, side="left")) for x in tgt) | {0, nb})

    def work(k):
        lo, hi = cuts[k], cuts[k + 1]
        p0, p1 = int(sub.bc_off[lo]), int(sub.bc_off[hi])
        o = sub.off[2 * p0:2 * p1 + 1]; ro = ref["reg_off"][2 * p0:2 * p1 + 1]
        aln = ref["aln"][ro[0]:ro[-1]].copy(); c0 = int(aln[0, 5]) if len(aln) else 0
        aln[:, 5] -= c0
        return P.place(sub.seq[o[0]:o[-1]], o - o[0], sub.bc_off[lo:hi + 1] - p0, rfa[lo:hi], seeds[p0:p1], ro - ro[0], ref["regs"][ro[0]:ro[-1]], aln, ref["cigar"][c0:])
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, range(len(cuts) - 1)))
    dt = time.perf_counter() - t
    log(f"cpu reference: {sub.n_pairs} pairs, C part {t1 - t:.2f}s + placement port {dt - (t1 - t):.2f}s on {threads} threads")
    _REF_CACHE["last"] = (sub, ref, parts, seeds, rfa)
    return sub.n_pairs, dt


def merge_place_parts(sub, ref, parts, cuts_pairs):
    """Concatenate the placement oracle's per-chunk outputs into batch-global arrays (indices re-based)."""
    offs, cands, lmps = [np.zeros(1, np.int64)], [], []
    cb = 0
    for (oc, ca, lmp), p0 in zip(parts, cuts_pairs):
        ca = ca.copy()
        rb = int(ref["reg_off"][2 * p0])
        ca["reg"][ca["reg"] >= 0] += rb
        ca["read"] += 2 * p0
        for f in ("second_best", "mate"):
            ca[f][ca[f] >= 0] += cb
        offs.append(oc[1:] + cb); cands.append(ca); lmps.append(lmp)
        cb += int(oc[-1])
    return np.concatenate(offs), np.concatenate(cands), np.concatenate(lmps)


def mapping_truth(sub, cand_off, cands, rid, is_rev, tol=20):
    """Fraction of the sample's reads whose selected (active) alignment sits on the simulated contig / strand within `tol`
    bases of the simulated position (ReadSet.truth): pins the index and the whole path independently of either arm."""
    act = np.nonzero(cands["active"] == 1)[0]
    read = cands["read"][act]
    assert len(np.unique(read)) == len(read) == 2 * sub.n_pairs, "not exactly one active alignment per read"
    order = np.argsort(read); act = act[order]
    reg = cands["reg"][act]
    L = np.diff(sub.off)
    t = sub.truth
    pair = np.arange(2 * sub.n_pairs) // 2; mate = np.arange(2 * sub.n_pairs) & 1
    # R1 is the reverse-strand read of the fragment when flip (simulate.py): read 2p is reverse iff flip, read 2p+1 iff not flip
    rev = (t[pair, 3] == 1) ^ (mate == 1)
    exp_pos = np.where(rev, t[pair, 1] + t[pair, 2] - L, t[pair, 1])
    ok = reg >= 0
    r = np.where(ok, reg, 0)
    good = ok & (rid[r] == t[pair, 0]) & ((is_rev[r] != 0) == rev) & (np.abs(cands["pos"][act] - exp_pos) <= tol)
    return float(good.mean())


def parity_sample(A):
    """The CPU-baseline sample (whole work units of batch 0) through ALL stages of the library against the reference arm on
    the same pairs: regions / alignments / CIGARs from the reference's own C code (oracle/_ref), placement (candidates,
    active picks, molecules, MAPQ) from the oracle's restatement of the Go code with the tolerances of
    tests/test_gpu_parity.py, and both arms against the simulation truth.  Any false flag makes bench.py exit non-zero.
    Reference coordinates beyond 2^32 only exist at the configs[1] index size, which no pytest case reaches."""
    from tests import util
    sub, ref, parts, seeds, rfa = _REF_CACHE["last"]
    res = A.align_batch(sub.seq, sub.off, sub.bc_off, stages=15, run_rfa=rfa, rand_seed=seeds)
    regs = bool(np.array_equal(res["reg_off"], ref["reg_off"]) and res["regs"].tobytes() == ref["regs"].tobytes())
    a, e = res["alns"], ref["aln"]
    alns = regs and all(np.array_equal(a[f], e[:, j]) for f, j in (("pos", 0), ("rid", 1), ("is_rev", 2), ("nm", 3), ("n_cigar", 4), ("score", 6), ("sub", 7)))
    if alns:  # CIGAR pools: same words in the same order once offsets are rebased
        idx_g = np.concatenate([np.arange(o, o + n) for o, n in zip(a["cigar_off"].tolist(), a["n_cigar"].tolist())]) if len(a) else np.zeros(0, np.int64)
        idx_r = np.concatenate([np.arange(o, o + n) for o, n in zip(e[:, 5].tolist(), e[:, 4].tolist())]) if len(a) else np.zeros(0, np.int64)
        alns = bool(np.array_equal(res["cigar"][idx_g], ref["cigar"][idx_r]))
    nb = len(sub.bc_off) - 1
    # the chunk boundaries cpu_reference used (same formula)
    threads = len(parts)
    tgt = [sub.n_pairs * k // (os.cpu_count() or 1) for k in range((os.cpu_count() or 1) + 1)]
    cuts = sorted(set(int(np.searchsorted(sub.bc_off, x, side="left")) for x in tgt) | {0, nb})
    assert len(cuts) - 1 == threads
    oc, ocands, olmp = merge_place_parts(sub, ref, parts, [int(sub.bc_off[c]) for c in cuts[:-1]])
    bad = util.place_mismatches(res, oc, ocands, olmp) if regs else ["regions differ"]
    out = {"pairs": int(sub.n_pairs), "work_units": nb, "regions_identical": regs, "alignments_and_cigars_identical": bool(alns),
           "placement_identical": not bad, "placement_mismatches": bad[:6], "n_regions": int(ref["reg_off"][-1]), "n_candidates": int(oc[-1]),
           "placement_tolerance": "integers exact; scores 1e-9 rel; MAPQ equal or +-1 where the untruncated value is within 1e-6 of an integer"}
    out["mapping_truth"] = {"library": mapping_truth(sub, res["cand_off"], res["cands"], a["rid"], a["is_rev"]) if regs else None,
                            "reference": mapping_truth(sub, oc, ocands, e[:, 1], e[:, 2]),
                            "rule": "active alignment on the simulated contig and strand, |pos - simulated pos| <= 20"}
    out["ok"] = bool(regs and alns and not bad and out["mapping_truth"]["library"] is not None and out["mapping_truth"]["library"] > 0.97
                     and out["mapping_truth"]["reference"] > 0.97)
    return out


# ---------------------------------------------------------------------------------------------- configs[4]: ksw_extend2 microbench
def ksw_problems(n, L, seed):
    """Extension problems with the shape mem_chain2aln's call sites give them (bwamem.c:715-771) for single-seed chains of simulated
    reads: a read of length L copied from a random reference with 1 % substitutions and 0.05 %/base insertions and deletions; the seed is
    its longest error-free stretch (>= 19); the LEFT flank (read and reference reversed, h0 = seed length, target = flank +
    cal_max_gap(flank) bases of reference) and the RIGHT flank (h0 = seed length + left flank length as an upper bound of the left
    score) each give one problem.  Half of the n problems are of this kind, half are pure-random flanks against unrelated targets (the
    early-termination path).  Returns lists of query / target arrays and h0."""
    g = np.random.default_rng(seed)
    qs, ts, h0 = [], [], []
    def max_gap(q):  # cal_max_gap (bwamem.c:621-628) with a=1, o=6, e=1, w=100
        l = max(int((q * 1 - 6) / 1 + 1.), 1)
        return min(l, 200)
    n_real = n // 2
    while len(qs) < n_real:
        m = 4096
        ref = g.integers(0, 4, (m, L + 420), dtype=np.uint8)
        start = 210
        for k in range(m):
            r = ref[k]
            read = r[start:start + L].copy()
            sub = np.nonzero(g.random(L) < 0.01)[0]
            read[sub] = (read[sub] + g.integers(1, 4, len(sub))) & 3
            brk = set(sub.tolist())
            tgt = r
            if g.random() < 0.14:  # one 1-3 base indel (0.05 %/base each way over 2 x L bases)
                p = int(g.integers(10, L - 10)); l = int(g.integers(1, 4))
                if g.random() < 0.5:
                    read = np.concatenate([read[:p], g.integers(0, 4, l).astype(np.uint8), read[p:L - l]])
                else:
                    read = np.concatenate([read[:p], read[p + l:], r[start + L:start + L + l]])
                brk.add(p)
            edges = [-1] + sorted(brk) + [L]
            best = max(range(len(edges) - 1), key=lambda i: edges[i + 1] - edges[i])
            sb, se = edges[best] + 1, edges[best + 1]
            if se - sb < 19:
                continue
            if sb > 0:       # left flank, reversed
                q = read[:sb][::-1]; tl = sb + max_gap(sb)
                t = tgt[max(start + sb - tl, 0):start + sb][::-1]
                qs.append(q); ts.append(t); h0.append(se - sb)
            if se < L:       # right flank
                q = read[se:]; tl = (L - se) + max_gap(L - se)
                t = tgt[start + se:start + se + tl]
                qs.append(q); ts.append(t); h0.append(min(se, 255 - (L - se)))
            if len(qs) >= n_real:
                break
    qs, ts, h0 = qs[:n_real], ts[:n_real], h0[:n_real]
    for _ in range(n - n_real):
        ql = int(g.integers(5, max(6, L - 19))); tl = ql + max_gap(ql)
        qs.append(g.integers(0, 4, ql).astype(np.uint8)); ts.append(g.integers(0, 4, tl).astype(np.uint8)); h0.append(int(g.integers(19, max(20, min(L - ql, 255 - ql) + 1))))
    return qs, ts, np.array(h0, np.int32)


def run_ksw(args):
    """--config ksw: BASELINE.json configs[4], ksw_extend2 (bwa/ksw.c:380-479) on the grid band 16-256 x read length 100-250, GCUPS counted
    in in-band cells exactly as ksw.c:424-448 visits them.  value = the L=150, w=100 point (the pipeline's shape)."""
    import torch
    from arachne_b200 import ArachneCuda
    from tests import util
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libarachne_cuda has no CPU path")
    A = ArachneCuda()
    n = args.pairs if args.pairs != 1000000 else 1 << 20
    R = None
    try:
        from oracle.refbwa import RefBwa, available
        R = RefBwa() if available() else None
    except Exception:
        R = None
    clk = ClockSampler(0); clk.start()
    grid, head, e2e_head = [], None, None
    for L in (100, 150, 200, 250):
        qs, ts, h0 = ksw_problems(n, L, 15 + L)
        q, qo = util.pack(qs); t, to = util.pack(ts)
        for w in (16, 32, 64, 100, 128, 200, 256):
            wv = np.full(n, w, np.int32)
            for _ in range(max(0, args.warmup - 1) if (L, w) == (150, 100) else 0):
                A.ksw_extend(q, qo, t, to, h0, wv, end_bonus=5, zdrop=100, repeat=1)
            t0 = time.perf_counter()
            out, cells, ms = A.ksw_extend(q, qo, t, to, h0, wv, end_bonus=5, zdrop=100, repeat=args.steps)
            wall = time.perf_counter() - t0
            row = {"read_len": L, "w": w, "problems": n, "cells": int(cells), "kernel_ms": ms, "gcups": cells / ms / 1e6}
            if R is not None and not args.no_cpu_baseline:
                k = min(400, n); c0 = time.perf_counter(); ccells = 0
                for i in range(k):
                    r = R.ksw_extend2(qs[i], ts[i], w, 5, 100, int(h0[i]))
                    if tuple(out[i]) != tuple(r):
                        row["parity_failure"] = [i, [int(x) for x in out[i]], [int(x) for x in r]]
                dt = time.perf_counter() - c0
                row.update({"cpu_problems": k, "cpu_gcups_1thread": cells * (k / n) / dt / 1e9})
            grid.append(row)
            log(json.dumps(row))
            if (L, w) == (150, 100):
                head = row
                # end to end through the C ABI: host arrays in, results out, one launch (the call above minus its warm-up launch and repeats)
                t0 = time.perf_counter(); A.ksw_extend(q, qo, t, to, h0, wv, end_bonus=5, zdrop=100, repeat=1); e2e_head = time.perf_counter() - t0
    clocks = clk.stop()
    dpx = A.dpx_peak()
    ok = not any("parity_failure" in r for r in grid)
    line = {"metric": "ksw_extend2 banded-SW GCUPS", "value": head["gcups"], "unit": "GCUPS", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["kernel_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32 (scores as bytes, DPX max-plus)",
            "data": "synthetic",
            "config": {"workload": f"configs[4]: ksw_extend2 microbench, {n} problems per point: half with the shape of mem_chain2aln's call sites for simulated 1 %-error reads "
                                   "(left + right flank of the longest exact seed, target = flank + cal_max_gap), half pure-random; grid band {16,32,64,100,128,200,256} x read length "
                                   "{100,150,200,250}; value = the L=150, w=100 point; h0 = seed length, end_bonus 5, zdrop 100", "name": "ksw",
                       "l2_policy": "each point's queries + targets + results (>= 150 MB) exceed nothing but are streamed once per launch; launches of different points alternate"},
            "grid": grid,
            "roofline": {"kernel": "k_ext_dp (ksw_extend2, one problem per lane)", "bound": "dpx issue", "unit": "G DPX-class ops/s", "dpx_ops_per_cell": 4.75,
                         "achieved": head["gcups"] * 4.75, "peak": dpx, "frac": head["gcups"] * 4.75 / dpx if dpx else None, "traffic": None,
                         "peak_source": "measured in this run: independent VIADDMNMX streams (arachne_dpx_peak); the integer ALU pipe they share with the cell's 3 PRMT + 1 SHF "
                                        "+ key max is the binding unit: 9.25 ALU-pipe instructions per cell (SASS of the 4-cell group)",
                         "alu_ops_per_cell": 9.25, "alu_frac": head["gcups"] * 9.25 / dpx if dpx else None},
            "e2e": {"value": head["cells"] / e2e_head / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": int(sum(len(x) for x in qs) + sum(len(x) for x in ts) + 16 * n + 8 * n),
                    "d2h_bytes_per_step": 24 * n, "timing": "host wall clock around one arachne_ksw_extend_batch call with host arrays (L=250 set sizes for the byte counts)"},
            "gpu_launches": (args.steps + 1) * 28, "clocks": clocks, "parity_ok": ok}
    if head.get("cpu_gcups_1thread"):
        line["cpu_baseline"] = {"value": head["cpu_gcups_1thread"], "unit": "GCUPS", "cores": 1, "kind": "reference",
                                "sample": f"{head['cpu_problems']} problems of the L=150, w=100 point through the reference's own ksw_extend2 (oracle/_ref), one thread, ctypes call overhead included"}
    emit(line)
    return 0 if ok else 3


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation on this box's host cores, same workload/metric.  Rank 0 only."""
    if rank != 0:
        return
    fa, ctg = dataset(args, 0, 1)
    threads = os.cpu_count() or 1
    n = max(1000, min(args.cpu_sample, args.pairs))
    batches = [make_batch(ctg, args, b, n) for b in range(min(args.batches, 2))]
    times = []
    for it in range(args.warmup + args.steps):
        npairs, dt = cpu_reference(fa, batches[it % len(batches)], n, threads, args)
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    val = npairs / (ms / 1e3)
    line = base_line(args, world, val, ms)  # `config` is the arm's own, word for word; what the CPU step really was is said beside it
    line["reference_arm_step"] = f"{npairs} pairs per step (a bounded sample of the arm's {args.pairs}-pair step: same generator, same work-unit rule); rank 0's host cores only"
    line.update({"impl": "reference", "dtype": "int32", "n_gpus": args.gpus,
                 "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": threads, "kind": "reference",
                                  "sample": f"{npairs} pairs per step of the same workload (oracle/_ref: the reference's BWA C sources, gobwa.go call sequence, {threads} pthreads over pairs)"},
                 "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0, "roofline": None})
    emit(line)


def base_line(args, world, value, ms):
    big = args.ref_mb >= 3000
    cfgno = {"haplotag": 1, "stlfr": 2, "tellseq": 3}[args.config]
    what = (f"configs[{cfgno}]: synthetic 3.1 Gb human-scale reference" if big else f"configs[{cfgno}] scaled down to a synthetic {args.ref_mb:g} Mb reference") + \
           " (24 human-like contigs, iid, seed 2; index built by arachne_index_build, byte-identical to bwa index)"
    return {"metric": "read pairs/sec aligned", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"{what}, {WORKLOADS[args.config][3]}; work units cut by the reference reader's rule (reader.go:208-291); "
                                   f"{args.pairs} pairs per step per GPU ({args.steps} timed steps = {args.pairs * args.steps * world} pairs); "
                                   f"stages: seeding+SA+chaining+extension+mate rescue+CIGAR/NM+placement (tagBestAlignments, molecules, optimizer, MAPQ; " + ("plain barcodes: no RFA" if args.plain_barcodes else "'-1' barcodes: RFA on") + ")",
                       "name": args.config, "ref_mb": args.ref_mb, "pairs_per_step_per_gpu": args.pairs, "read_len": 150, "parallelism": f"barcode-sharded x{world}, index replicated, no collective",
                       "l2_policy": "index (3.1 GB BWT + 1.55 GB SA + 0.78 GB pac at 3.1 Gb) and per-batch arenas exceed the 126 MB L2; distinct batches are cycled"}}


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line goes to the process's real stdout; everything else that writes to fd 1 during the run (NCCL
    prints its version banner there) has been pointed at stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.config == "ksw":
        if rank != 0:
            return 0
        if args.impl == "reference":
            args.no_cpu_baseline = False
        return run_ksw(args)
    if args.impl == "reference":
        return run_reference(args, rank, world)
    import torch
    import torch.distributed as dist
    from arachne_b200 import ArachneCuda
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libarachne_cuda has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)

    rc = 0
    fa, ctg = dataset(args, rank, world)
    barrier()
    batches = make_batches(ctg, args, rank)
    A = ArachneCuda().load_index(fa, device=local)
    stream = torch.cuda.Stream()
    A.set_stream(stream.cuda_stream)
    # resident inputs
    pl = [place_inputs(b, args) for b in batches]
    dev = [(torch.from_numpy(b.seq).cuda(), torch.from_numpy(b.off).cuda(), b, torch.from_numpy(b.bc_off).cuda(), torch.from_numpy(rfa).cuda(),
            torch.from_numpy(sd).cuda()) for b, (sd, rfa) in zip(batches, pl)]
    pinned = [tuple(torch.from_numpy(x).pin_memory() for x in (b.seq, b.off, b.bc_off, rfa, sd)) for b, (sd, rfa) in zip(batches, pl)]
    torch.cuda.synchronize()

    def step_resident(i):
        s, o, b, bc, rfa, sd = dev[i % len(dev)]
        return A.align_batch_device(s.data_ptr(), o.data_ptr(), b.n_pairs, int(b.off[-1]), 150, stages=STAGES, d_bc_off_ptr=bc.data_ptr(),
                                    n_barcodes=len(b.bc_off) - 1, d_run_rfa_ptr=rfa.data_ptr(), d_rand_seed_ptr=sd.data_ptr())

    def step_host(i, flags=4):
        # flags=4 (ARACHNE_OUT_NO_REGIONS): everything a consumer of the placed alignments reads comes back -- alignments, CIGARs,
        # mismatch loci, candidates with MAPQ, molecule penalties -- but not the intermediate mem_alnreg_t array; flags=0 copies that too.
        s, o, bc, rfa, sd = pinned[i % len(pinned)]
        return A.align_batch_host(s.numpy(), o.numpy(), stages=STAGES, bc_off=bc.numpy(), run_rfa=rfa.numpy(), rand_seed=sd.numpy(), flags=flags)

    # ---- device-resident timing (value) ----
    with torch.cuda.stream(stream):
        # algorithmic occ-block touches of each distinct batch = what the reference's bwt_2occ4 calls touch: counted once,
        # untimed, with every bwt_extend call made (the k-mer tables and the direct extension of unique matches skip some of them in the timed steps)
        A.set_option("seed_kmer", 0); A.set_option("seed_direct", 0)
        algo_blocks = [step_resident(i)["counters"]["n_occ_blocks"] for i in range(len(dev))]
        A.set_option("seed_kmer", int(os.environ.get("ARACHNE_BENCH_SEED_KMER", "1")))
        A.set_option("seed_direct", int(os.environ.get("ARACHNE_BENCH_SEED_DIRECT", "1")))
        if os.environ.get("ARACHNE_BENCH_SEED_FWD_LEVELS"):
            A.set_option("seed_fwd_levels", int(os.environ["ARACHNE_BENCH_SEED_FWD_LEVELS"]))
        A.set_option("seed_blocks_per_sm", int(os.environ.get("ARACHNE_BENCH_SEED_BLOCKS", "3")))
        A.set_option("seed_shared_list", int(os.environ.get("ARACHNE_BENCH_SEED_SLIST", "-1")))
        for i in range(args.warmup):
            step_resident(i)
        torch.cuda.synchronize(); barrier()
        clk = ClockSampler(local); clk.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        acc = {}
        e0.record(stream)
        for i in range(args.steps):
            r = step_resident(i)
            for k, v in {**r["ms"], **r["counters"]}.items():
                acc[k] = acc.get(k, 0) + v
        e1.record(stream)
        torch.cuda.synchronize(); barrier()
        ms_total = e0.elapsed_time(e1)
        # ---- end to end through the host C-ABI (e2e) ----
        # Host buffers in, every result copied back to pinned host memory, all inside the timed region.  The calls come from
        # E2E_THREADS host threads (the reference's host is a pool of worker goroutines calling into the library
        # concurrently, aligner.go:335-337); the handle has two workspaces, so one batch's copies overlap the other's kernels.
        from concurrent.futures import ThreadPoolExecutor
        A.set_stream(None)  # each workspace runs on its own stream
        n_thr = max(1, int(os.environ.get("ARACHNE_BENCH_E2E_THREADS", "2")))
        with ThreadPoolExecutor(n_thr) as ex:
            list(ex.map(step_host, range(min(2, args.warmup))))
            torch.cuda.synchronize(); barrier()
            t0 = time.perf_counter()
            res = list(ex.map(step_host, range(args.steps)))
            torch.cuda.synchronize(); ms_e2e = (time.perf_counter() - t0) * 1e3; barrier()
            # the same with the region array copied back as well (every array the library can return)
            list(ex.map(lambda i: step_host(i, 0), range(min(2, args.warmup))))
            torch.cuda.synchronize(); barrier()
            t0 = time.perf_counter()
            res_all = list(ex.map(lambda i: step_host(i, 0), range(args.steps)))
            torch.cuda.synchronize(); ms_e2e_all = (time.perf_counter() - t0) * 1e3; barrier()
        h2d = sum(r["h2d_bytes"] for r in res); d2h = sum(r["d2h_bytes"] for r in res)
        d2h_all = sum(r["d2h_bytes"] for r in res_all)
        A.set_stream(stream.cuda_stream)
        clocks = clk.stop()
    t = torch.tensor([ms_total, ms_e2e, ms_e2e_all], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, ms_e2e_all = float(t[0]), float(t[1]), float(t[2])
    if rank == 0:
        pairs = args.pairs * args.steps * world
        value = pairs / (ms_total / 1e3)
        line = base_line(args, world, value, ms_total / args.steps)
        peak, how = peaks()
        seed_ms = acc["seed"] / args.steps
        blocks = sum(algo_blocks[i % len(algo_blocks)] for i in range(args.steps)) / args.steps
        achieved = blocks * 64 / (seed_ms / 1e3) / 1e9
        traffic, traffic_note = None, "no ncu capture committed for this workload"
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tp):  # dram bytes of ONE k_seed2 launch from `ncu --set full` on this same command (scripts/gpu_round.sh)
            tj = json.load(open(tp))
            if tj.get("pairs_per_step") == args.pairs and tj.get("ref_mb") == args.ref_mb and tj.get("config", "haplotag") == args.config:
                traffic, traffic_note = tj["dram_bytes_per_launch"], tj.get("note", "")
        dpx = A.dpx_peak()
        ext_ops = 3 * acc["n_ext_cells"] / args.steps  # DPX-class ops per in-band cell in k_extend2's SASS: VIMNMX3 + 2 x VIADDMNMX.RELU
        line["roofline"] = {"kernel": "k_seed2 (FM-index SMEM seeding: bwt_2occ4 block reads)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                            "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note, "peak_source": how,
                            # what the DRAM really delivered: ncu dram bytes of one launch / this run's kernel time (null without a matching capture)
                            "delivered": traffic / (seed_ms / 1e3) / 1e9 if traffic else None, "frac_delivered": traffic / (seed_ms / 1e3) / 1e9 / peak if traffic else None,
                            "frac_definition": "frac = ALGORITHMIC bytes (64 B per occ block the reference's bwt_2occ4 calls touch, SURVEY 8d) / kernel time / peak; frac_delivered = ncu dram bytes / kernel time / peak",
                            "extension": {"kernel": "k_extend2 (ksw_extend2)", "bound": "dpx issue", "gcups": acc["n_ext_cells"] / (acc["extend"] / 1e3) / 1e9 if acc["extend"] else None,
                                          "dpx_ops_per_cell": 3, "achieved_gops": ext_ops / (acc["extend"] / args.steps / 1e3) / 1e9 if acc["extend"] else None,
                                          "peak_gops": dpx, "frac": (ext_ops / (acc["extend"] / args.steps / 1e3) / 1e9) / dpx if acc["extend"] and dpx else None,
                                          "peak_source": "measured in this run: independent VIADDMNMX streams (arachne_dpx_peak)"},
                            "algorithmic_bytes_per_launch": blocks * 64, "kernel_ms": seed_ms,
                            "stage_ms_per_step": {k: acc[k] / args.steps for k in ("seed", "sa", "chain", "extend", "rescue", "cigar", "place", "total")},
                            "ext_gcups": acc["n_ext_cells"] / (acc["extend"] / 1e3) / 1e9 if acc["extend"] else None}
        line["e2e"] = {"value": pairs / (ms_e2e / 1e3), "unit": "pairs/s", "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
                       "host_threads": n_thr, "timing": "host wall clock around all steps, device synchronised on both sides, max over ranks",
                       "result": "ARACHNE_OUT_NO_REGIONS: reg_off, alignments (pos, CIGAR, NM, score, mismatch loci), candidates with MAPQ and "
                                 "molecule penalties in pinned host memory; the intermediate mem_alnreg_t array stays on the device",
                       "all_arrays": {"value": pairs / (ms_e2e_all / 1e3), "unit": "pairs/s", "d2h_bytes_per_step": d2h_all // args.steps,
                                      "note": "the same call with the mem_alnreg_t array copied back too (flags = 0)"}}
        line["gpu_launches"] = int(acc["n_kernel_launches"])
        line["clocks"] = clocks
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            n, dt = cpu_reference(fa, batches[0], args.cpu_sample, threads, args)
            line["cpu_baseline"] = {"value": n / dt, "unit": "pairs/s", "cores": threads, "kind": "reference",
                                    "sample": f"first {n} pairs of batch 0 (whole barcodes), oracle/_ref = the reference's BWA C sources driven as gobwa.go drives them, {threads} pthreads"}
            line["parity_sample"] = parity_sample(A)
        emit(line)
        if "parity_sample" in line and not line["parity_sample"]["ok"]:
            log("PARITY FAILURE at the bench size: " + json.dumps(line["parity_sample"]))
            rc = 3
    if world > 1:
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main() or 0)
