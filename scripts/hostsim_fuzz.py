#!/usr/bin/env python
"""Differential fuzzing of the thread-sequential device logic (tests/hostsim: chaining, extension driver, de-duplication / patching, mate
rescue, mem_reg2aln of the warp tier, the shipped seeding kernel k_seed2 and the shipped CIGAR tiers k_cigar_fast / k_cigar_band run one lane at a time -- the same csrc/*.cuh the kernels
compile) against the reference's own C code (oracle/_ref), on
references and reads drawn to reach the rare branches: diverged and exact repeats, tandem arrays, short contigs (window clamps), read
lengths 30..250, heavy substitutions / indels, N runs, chimeric reads (region patching and colinear merges), wrecked mates (rescue).
CPU only.   python scripts/hostsim_fuzz.py [seconds=300] [first_seed=1]   -> one JSON line; non-zero exit on the first divergence."""
import json, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import simulate as S
from arachne_b200.lib import ALNREG
from oracle.refbwa import RefBwa
from tests import util

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 300.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
HS = util.HostSim()
from arachne_b200 import ArachneCuda
ARACHNE = ArachneCuda(); OPT = ARACHNE.opt; PO_CLS = util.PlaceOracle
t0 = time.time(); stats = dict(staged_reads=0, staged_dp_problems=0, staged_rescued_pairs=0, rounds=0, pairs=0, regions=0, rescued_reads=0, cigars=0, gapped_cigars=0, chains_checked=0, patched_or_merged=0, occ_sa_checked=0, seed_reads_checked=0, seed_intervals=0, tier_regions=0, tier2_regions=0, tier3_regions=0, warp_tier_regions=0, warp_rescues=0, placed_candidates=0, molecules=0)
rc = lambda s: np.where(s[::-1] < 4, 3 - s[::-1], 4).astype(np.uint8)


def make_reference(g, kind):
    if kind == 0:   # diverged repeats + N
        return S.make_reference(int(g.integers(80_000, 300_000)), int(g.integers(1, 5)), seed=int(g.integers(1 << 30)), repeat_frac=float(g.uniform(0.1, 0.5)), n_frac=0.003)
    if kind == 1:   # exact repeat families
        return util.exact_repeat_reference(total_len=int(g.integers(100_000, 300_000)), n_contigs=int(g.integers(1, 5)), seed=int(g.integers(1 << 30)),
                                           families=int(g.integers(2, 8)), copies=int(g.integers(4, 40)), rep_len=int(g.integers(200, 1500)))
    if kind == 2:   # tandem arrays and homopolymer runs inside random sequence, many short contigs
        ctg = []
        for c in range(int(g.integers(3, 12))):
            parts = []
            while sum(len(p) for p in parts) < int(g.integers(2_000, 30_000)):
                u = g.random()
                if u < 0.15: parts.append(np.tile(g.integers(0, 4, int(g.integers(1, 7))), int(g.integers(10, 120))))
                elif u < 0.2: parts.append(np.full(int(g.integers(20, 200)), int(g.integers(0, 4))))
                else: parts.append(g.integers(0, 4, int(g.integers(200, 3000))))
            ctg.append(("t%d" % c, np.concatenate(parts).astype(np.uint8)))
        return ctg
    return S.make_reference(int(g.integers(50_000, 150_000)), int(g.integers(1, 3)), seed=int(g.integers(1 << 30)))


def make_reads(g, ctg):
    L = int(g.choice([30, 50, 75, 100, 150, 151, 200, 249, 250]))
    rs = S.simulate_linked_reads(ctg, n_barcodes=int(g.integers(3, 9)), seed=int(g.integers(1 << 30)), read_len=L, sub=float(g.choice([0.0, 0.01, 0.03, 0.08])),
                                 ins=float(g.choice([0.0, 0.001, 0.01, 0.03])), dele=float(g.choice([0.0, 0.001, 0.01, 0.03])), nrate=float(g.choice([0.0, 0.002, 0.02])))
    reads = [rs.seq[rs.off[i]:rs.off[i + 1]].copy() for i in range(2 * rs.n_pairs)]
    genome = np.concatenate([s for _, s in ctg])
    for p in range(rs.n_pairs):
        u = g.random(); a, b = 2 * p, 2 * p + 1
        if u < 0.12:   # wrecked mate: only rescue places it
            k = a if g.random() < 0.5 else b; m = g.random(len(reads[k])) < 0.11
            reads[k][m] = (reads[k][m] + g.integers(1, 4, int(m.sum()))) & 3
        elif u < 0.24 and len(reads[a]) > 80:   # chimera: tail from elsewhere (two regions per read, sometimes colinear: patching)
            cut = int(g.integers(30, len(reads[a]) - 30)); at = int(g.integers(0, max(1, len(genome) - len(reads[a]))))
            reads[a][cut:] = np.minimum(genome[at:at + len(reads[a]) - cut], 3)[:len(reads[a]) - cut]
        elif u < 0.32 and len(reads[a]) > 100:  # a deletion of 20..120 reference bases inside the read: two colinear regions that may merge
            k = a; src = rs.truth[p]
            ref = ctg[int(src[0])][1]; fs = int(src[1]); gap = int(g.integers(20, 120)); half = len(reads[k]) // 2
            if fs + len(reads[k]) + gap < len(ref):
                reads[k] = np.minimum(np.concatenate([ref[fs:fs + half], ref[fs + half + gap:fs + len(reads[k]) + gap]]), 3).astype(np.uint8)
        elif u < 0.36: reads[b] = rc(reads[b])
        elif u < 0.40: reads[a], reads[b] = reads[b], reads[a]
        elif u < 0.43: reads[a] = g.integers(0, 4, len(reads[a])).astype(np.uint8)
        elif u < 0.45: reads[a] = reads[a][:int(g.integers(0, 25))]   # shorter than a seed, or empty
    return reads, rs


seed = seed0
with tempfile.TemporaryDirectory() as wd:
    while time.time() - t0 < budget:
        g = np.random.default_rng(seed)
        ctg = make_reference(g, int(g.integers(0, 4)))
        fa = os.path.join(wd, "f%d.fa" % seed)
        S.write_fasta(fa, ctg); R = RefBwa(); R.build_index(fa); R.load(fa); HS.load(fa)
        prim, *L2, sl = R.bwt_meta()
        ks = [0, 1, prim - 1, prim, prim + 1, sl - 1, sl, 63, 64, 65, 127, 128, 129] + [int(x) for x in g.integers(0, sl + 1, 400)]
        for k in ks:   # rows a1 / a2 / a6: the re-cut 32-byte-sector image, occ counts at block edges / primary / the k - 1 wrap, SA rows
            l = min(sl, k + int(g.integers(0, 300)))
            if k < 0 or k > sl: continue
            if HS.occ(k, l) != sum(R.occ4(k, l), []) or HS.occ(k - 1, l) != sum(R.occ4(k - 1, l), []) or HS.sa(k) != R.sa(k):
                print(json.dumps(dict(divergence="occ/sa", seed=seed, k=k, l=l, **stats))); sys.exit(1)
            stats["occ_sa_checked"] += 1
        reads, rs = make_reads(g, ctg)
        seq, off = util.pack(reads)
        # rows a3-a5 on the SHIPPED seeding kernel (csrc/seed_kernel.cuh run one lane at a time, tests/hostsim/hostsim_seed.cpp) with a random
        # choice of its shortcuts, against mem_collect_intv composed from the reference's bwt_smem1 / bwt_seed_strategy1
        si, K = HS.seed_index(fa, kmer_k=int(g.choice([-1, -1, 0, 4, 6])), direct=bool(g.random() < 0.7))
        use_kmer = int(g.choice([0, 1, 5, 1 | (int(g.integers(1, 15)) << 8)])); use_direct = int(g.random() < 0.7)
        io, iv, _st = HS.seed2(si, OPT, seq, off, use_kmer, use_direct, int(g.choice([-1, 0, 3, 16])), int(g.integers(0, 256)))
        HS.seed_index_free(si)
        for r in range(0, len(reads), 3):
            if [tuple(int(v) for v in a) for a in iv[io[r]:io[r + 1]]] != util.ref_collect_intv(R, reads[r]):
                print(json.dumps(dict(divergence="seed intervals", seed=seed, read=r, use_kmer=use_kmer, use_direct=use_direct, K=K, **stats))); sys.exit(1)
            stats["seed_reads_checked"] += 1; stats["seed_intervals"] += int(io[r + 1] - io[r])
        ref = R.batch(seq, off, rescue=True, cigar=True, threads=8); plain = R.batch(seq, off, rescue=False, cigar=False, threads=8)
        exp = ref["regs"].view(ALNREG); ro = ref["reg_off"]; exp1 = plain["regs"].view(ALNREG); ro1 = plain["reg_off"]
        # rows a6-a12 on the SHIPPED stage kernels in batch.inl's launch order (k_expand .. k_ext_dp on the packed reference .. k_rescue ..
        # k_gather_regs2 as blocks of fibers, tests/hostsim/hostsim_stages.cpp), fed with the intervals the shipped seeding kernel just produced
        st_rescue = bool(g.random() < 0.5); st_warp = int(g.integers(0, 4)); st_wide = bool(g.random() < 0.3)
        st_block = bool(g.random() < 0.5)  # the warp-cooperative kernels as whole blocks (their warps interleaving on the shared queue counters) or as one warp
        sro, sregs, sinfo = HS.stages(OPT, seq, off, io, iv, rescue=st_rescue, warp=st_warp, wide=st_wide, block=st_block)
        want = ref if st_rescue else plain
        if not np.array_equal(sro, want["reg_off"]) or sregs.tobytes() != want["regs"].view(ALNREG).tobytes():
            print(json.dumps(dict(divergence="staged kernels", seed=seed, rescue=st_rescue, warp=st_warp, wide=st_wide, block=st_block, **stats))); sys.exit(1)
        stats["staged_reads"] += len(reads); stats["staged_dp_problems"] += sinfo["problems"]; stats["staged_rescued_pairs"] += sinfo["rescued"]
        # row a13 on the SHIPPED lane-per-region CIGAR tiers (csrc/cigar_kernel.cuh run one lane at a time, tests/hostsim/hostsim_cigar.cpp)
        t_alns, t_cig, t_mm, t_list3, t_cnt = HS.cigar_tiers(OPT, seq, off, ro, exp)
        t_left = set(int(x) for x in t_list3)
        for j in range(len(exp)):
            if j in t_left: continue
            a = t_alns[j]; e = ref["aln"][j]
            if (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) != (e[0], e[1], e[2], e[3], e[4], e[6], e[7]) or \
                    not np.array_equal(t_cig[a["cigar_off"]:a["cigar_off"] + a["n_cigar"]], ref["cigar"][e[5]:e[5] + e[4]]):
                print(json.dumps(dict(divergence="cigar tiers", seed=seed, region=j, **stats))); sys.exit(1)
        # what the lane tiers leave goes to the warp tier: csrc/cigar.cuh on the device's WarpCtx as a warp of 32 fibers (hostsim_rescue.cpp)
        reg_read = np.repeat(np.arange(len(reads)), np.diff(ro))
        for j in sorted(t_left)[:40]:
            wa, wcig, _wmm = HS.reg2aln_warp(OPT, reads[reg_read[j]], exp[j:j + 1]); e = ref["aln"][j]
            if (wa["pos"], wa["rid"], wa["is_rev"], wa["nm"], wa["n_cigar"], wa["score"], wa["sub"]) != (e[0], e[1], e[2], e[3], e[4], e[6], e[7]) or \
                    not np.array_equal(wcig, ref["cigar"][e[5]:e[5] + e[4]]):
                print(json.dumps(dict(divergence="warp cigar tier", seed=seed, region=int(j), **stats))); sys.exit(1)
            stats["warp_tier_regions"] += 1
        stats["tier_regions"] += len(exp) - len(t_left); stats["tier2_regions"] += t_cnt["tier2"]; stats["tier3_regions"] += t_cnt["tier3"]
        # rows a15-a20 on the SHIPPED placement kernels (csrc/place.cuh + markDuplicates; k_place_rfa as a warp of fibers, hostsim_place.cpp)
        # against the restatement of the Go code, on this round's barcodes with a random choice of which of them run the optimizer
        t_all = t_alns.copy()
        for j in t_left:
            t_all[j] = HS.reg2aln(reads[reg_read[j]], exp[j:j + 1])[0]
        seeds = np.array([ARACHNE.name_seed(nm) for nm in rs.names], dtype=np.int64)
        run_rfa = (g.random(len(rs.bc_off) - 1) < 0.7).astype(np.uint8)
        pres = HS.place(off, ro, exp, t_all, rs.bc_off, run_rfa, seeds)
        oc, ocands, olmp = PO_CLS(fa).place(seq, off, rs.bc_off, run_rfa, seeds, ro, ref["regs"], ref["aln"], ref["cigar"])
        bad = util.place_mismatches(pres, oc, ocands, olmp)
        if bad:
            print(json.dumps(dict(divergence="placement", seed=seed, what=bad[:3], **stats))); sys.exit(1)
        stats["placed_candidates"] += len(ocands); stats["molecules"] += int(max(0, ocands["molecule_id"].max() + 1)) if len(ocands) else 0
        for p in range(len(reads) // 2):
            for k in (2 * p, 2 * p + 1):
                got = HS.align1(reads[k])
                if got.tobytes() != exp1[ro1[k]:ro1[k + 1]].tobytes():
                    print(json.dumps(dict(divergence="align1", seed=seed, read=k, **stats))); sys.exit(1)
                if len(reads[k]) >= 19 and k % 4 == 0:
                    for flt in (False, True):
                        ech, esd, efr = R.mem_chain(reads[k], do_flt=flt); gch, gsd, gfr = HS.chain(reads[k], do_flt=flt)
                        if not (np.array_equal(gch, ech) and np.array_equal(gsd, esd)):
                            print(json.dumps(dict(divergence="chain", seed=seed, read=k, flt=flt, **stats))); sys.exit(1)
                        stats["chains_checked"] += len(ech)
            if (ro[2 * p + 2] - ro[2 * p]) != (ro1[2 * p + 2] - ro1[2 * p]) and stats["warp_rescues"] < 40 * (stats["rounds"] + 1):
                # a pair that rescue changed: the same on the device's own warp context (striped local SW, leader writes between barriers)
                w1, w2 = HS.pair_warp(OPT, reads[2 * p], reads[2 * p + 1], exp1[ro1[2 * p]:ro1[2 * p + 1]], exp1[ro1[2 * p + 1]:ro1[2 * p + 2]])
                if w1.tobytes() != exp[ro[2 * p]:ro[2 * p + 1]].tobytes() or w2.tobytes() != exp[ro[2 * p + 1]:ro[2 * p + 2]].tobytes():
                    print(json.dumps(dict(divergence="warp rescue", seed=seed, pair=p, **stats))); sys.exit(1)
                stats["warp_rescues"] += 1
            o1, o2 = HS.pair(reads[2 * p], reads[2 * p + 1])
            for k, got in ((2 * p, o1), (2 * p + 1, o2)):
                if got.tobytes() != exp[ro[k]:ro[k + 1]].tobytes():
                    print(json.dumps(dict(divergence="pair", seed=seed, read=k, **stats))); sys.exit(1)
                stats["rescued_reads"] += int(ro[k + 1] - ro[k] != ro1[k + 1] - ro1[k]); stats["regions"] += int(ro[k + 1] - ro[k])
                stats["patched_or_merged"] += int(np.any((got["n_comp_is_alt"] & 0x3fffffff) > 1)) if "n_comp_is_alt" in got.dtype.names else 0
                for j in range(ro[k], ro[k + 1]):
                    a, cig, _mm = HS.reg2aln(reads[k], exp[j:j + 1]); e = ref["aln"][j]
                    if (a["pos"], a["rid"], a["is_rev"], a["nm"], a["n_cigar"], a["score"], a["sub"]) != (e[0], e[1], e[2], e[3], e[4], e[6], e[7]) or \
                            not np.array_equal(cig, ref["cigar"][e[5]:e[5] + e[4]]):
                        print(json.dumps(dict(divergence="reg2aln", seed=seed, read=k, region=int(j - ro[k]), **stats))); sys.exit(1)
                    stats["cigars"] += 1; stats["gapped_cigars"] += int(np.any(((cig & 0xf) == 1) | ((cig & 0xf) == 2)))
        stats["rounds"] += 1; stats["pairs"] += len(reads) // 2
        for ext in ("", ".bwt", ".sa", ".pac", ".ann", ".amb"):
            try: os.remove(fa + ext)
            except OSError: pass
        seed += 1
stats["patch_realignments"] = HS.lib.hs_patch_dp_calls()
print(json.dumps(dict(divergence=None, seconds=round(time.time() - t0, 1), first_seed=seed0, next_seed=seed, **stats)))
