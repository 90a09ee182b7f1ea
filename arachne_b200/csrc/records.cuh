// Record assembly on the batch outputs: what the reference does between estimateMapQualities and the BAM writer.
//   markDuplicates    src/aligner/aligner.go:611-641
//   CheckSplitReads   src/aligner/split.go:31-160 (GetSplitAlignment)
//   AppendBam         src/aligner/bamwriter.go:283-568 -- the FIELDS of every record (flag, RNAME/POS/MAPQ, RNEXT/PNEXT/TLEN, SEQ/QUAL
//                     orientation and hard clipping, AS XS XC AC XM AM XT SA BX VX DM); strings are formatted on the host (records_host.inl)
// Organisation: duplicates are found by one batch-wide stable radix sort of a 64-bit hash of (work unit, read1, strand, contig, pos,
// mate contig, mate pos) followed by an exact comparison inside equal-hash runs (the first read of a run in read order stays
// unmarked, as the reference's map insertion order decides); split selection is one thread per read over ALL its regions (the
// reference's un-filtered `full` list); records are written by one thread per PAIR because read 2's mate fields see the "too weak:
// report as unmapped" edit that writing read 1's record made (bamwriter.go:287-290).
#pragma once
#include "place.cuh"
#include "../../include/arachne_cuda.h"

namespace arachne {

struct Record {  // == arachne_record_t
    int32_t read, reg, primary, flag, rid, mapq, mate_rid, duplicate;
    int64_t pos, mate_pos, tlen;
    int32_t reversed, clip_start, clip_end, hard, as, xs, has_xc, xc_reg, xm, am, xt;
    int32_t sa_reg, sa_rid, sa_rev, sa_hard, sa_mapq, sa_nm;
    int64_t sa_pos;
    int32_t attach_bx, has_dm;
    double dm;
};
static_assert(sizeof(Record) == 152 && sizeof(Record) == sizeof(arachne_record_t), "Record must match arachne_record_t");

struct RecArgs {
    PlaceArgs P;
    const uint8_t* attach_bx;   // per work unit; nullptr: run_rfa
    int* act;                   // per read: its active candidate
    // split selection per read
    int* split_reg; int* split_mapq; int* split_proper; double* split_as; double* split_xs;
    unsigned long long* rec_cnt; const unsigned long long* rec_off; Record* recs;
};

// an alignment of the reference's `full` list, rebuilt from a region (k_place_build's arithmetic)
struct FullAln { int64_t pos, aend; int rid, reversed, score, qb, qe; double base; bool mapped; };

__device__ __forceinline__ FullAln rec_full_aln(const PlaceArgs& A, long g) {
    FullAln f;
    const AlnReg rg = A.regs[g]; const AlnRec a = A.alns[g];
    const int64_t coff = A.contigs[rg.rid].offset;
    const int64_t cpos = rg.rb < A.l_pac ? rg.rb - coff : A.l_pac * 2 - 1 - rg.rb - coff;
    const int64_t caend = rg.re < A.l_pac ? rg.re - coff : A.l_pac * 2 - 1 - rg.re - coff;
    f.pos = cpos; f.aend = caend;
    if (cpos != -1 && a.is_rev) { f.pos = caend + 1; f.aend = cpos + 1; }
    f.rid = a.rid; f.reversed = a.is_rev != 0; f.score = rg.score; f.qb = rg.qb; f.qe = rg.qe;
    double base = (double)(a.mismatches * -2 + a.indels * -3);
    if (a.soft_clipped > 0) { base -= 5.0 * (double)a.soft_clipped; base -= (double)a.soft_clipped_length * 0.5; }
    f.base = base; f.mapped = true;
    return f;
}
__device__ __forceinline__ bool rec_is_pair(int rev_a, int rid_a, int64_t pos_a, const Hot& b) {
    if ((rev_a != 0) == (b.reversed != 0) || rid_a != b.rid) return false;
    const int64_t dist = rev_a ? pos_a - b.pos : b.pos - pos_a;
    return dist >= -35 && dist < 750;
}

__global__ void k_rec_active(PlaceArgs A, int* act) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < A.n_reads; r += gridDim.x * blockDim.x) {
        int a = -1;
        for (long c = (long)A.cand_off[r], e = (long)A.cand_off[r + 1]; c < e; ++c) if (A.hot[c].active) { a = (int)c; break; }
        act[r] = a;
    }
}

// ---- markDuplicates ----
struct DupTuple { int bc, read1, reversed, rid, mrid; int64_t pos, mpos; };
__device__ __forceinline__ DupTuple rec_dup_tuple(const PlaceArgs& A, const int* act, int r) {
    DupTuple t;
    const Hot h = A.hot[act[r]], m = A.hot[act[r ^ 1]];  // the mate_alignment of an active alignment is the mate's active alignment
    t.bc = place_bc_of_read(A, r); t.read1 = !(r & 1); t.reversed = h.reversed; t.rid = h.rid; t.pos = h.pos; t.mrid = m.rid; t.mpos = m.pos;
    return t;
}
__device__ __forceinline__ unsigned long long rec_mix(unsigned long long x) {  // splitmix64 finaliser
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x;
}
__global__ void k_rec_dup_keys(PlaceArgs A, const int* __restrict__ act, unsigned long long* keys, unsigned* vals) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < A.n_reads; r += gridDim.x * blockDim.x) {
        const DupTuple t = rec_dup_tuple(A, act, r);
        unsigned long long h = rec_mix((unsigned long long)t.bc * 4 + t.read1 * 2 + t.reversed);
        h = rec_mix(h ^ (unsigned long long)(t.rid + 1)); h = rec_mix(h ^ (unsigned long long)(t.pos + 1));
        h = rec_mix(h ^ (unsigned long long)(t.mrid + 1)); h = rec_mix(h ^ (unsigned long long)(t.mpos + 1));
        keys[r] = h; vals[r] = (unsigned)r;
    }
}
__global__ void k_rec_dup_mark(PlaceArgs A, const int* __restrict__ act, const unsigned long long* __restrict__ keys, const unsigned* __restrict__ vals) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.n_reads; i += gridDim.x * blockDim.x) {
        const int r = (int)vals[i];
        const DupTuple t = rec_dup_tuple(A, act, r);
        bool dup = false;
        for (int j = i - 1; j >= 0 && keys[j] == keys[i] && !dup; --j) {  // the sort is stable: earlier in the run == earlier read of the batch
            const DupTuple u = rec_dup_tuple(A, act, (int)vals[j]);
            dup = u.bc == t.bc && u.read1 == t.read1 && u.reversed == t.reversed && u.rid == t.rid && u.pos == t.pos && u.mrid == t.mrid && u.mpos == t.mpos;
        }
        if (dup) A.cands[act[r]].duplicate = 1;
    }
}

// ---- CheckSplitReads ----
__global__ void k_rec_split(RecArgs R) {
    const PlaceArgs& A = R.P;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < A.n_reads; r += gridDim.x * blockDim.x) {
        int sreg = -1, smapq = 0, sproper = 0; double sas = 0.0, sxs = 0.0;
        const int ca = R.act[r];
        const Cand pc = A.cands[ca];
        const Hot pmate = A.hot[R.act[r ^ 1]];
        const double pmate_base = A.aux[R.act[r ^ 1]].base;
        const int read_len = (int)(A.seq_off[r + 1] - A.seq_off[r]);
        if (pc.reg >= 0 && pc.pos != -1) {
            const AlnReg prg = A.regs[pc.reg];
            int Ps = prg.qb, Pe = prg.qe;
            if (Ps > Pe) { const int t = Ps; Ps = Pe; Pe = t; }
            if (!(Pe - Ps > read_len - 15)) {
                long best = -1, second = -1; int best_s = -1, second_s = -1; int n_c = 0; int best_proper = 0;
                for (long g = (long)A.reg_off[r], e = (long)A.reg_off[r + 1]; g < e; ++g) {
                    if (g == pc.reg) continue;  // the active alignment (the only one of the read)
                    const FullAln f = rec_full_aln(A, g);
                    if (f.pos == -1) continue;
                    int Ss = f.qb, Se = f.qe, overlap;
                    if (Ss > Se) { const int t = Ss; Ss = Se; Se = t; }
                    if ((Ps < Ss && Pe > Se) || (Ss < Ps && Se > Pe)) continue;
                    else if (Ps < Ss) overlap = Pe - Ss;
                    else overlap = Se - Ps;
                    if (!(overlap < (Se - Ss) / 2)) continue;
                    const int proper = rec_is_pair(f.reversed, f.rid, f.pos, pmate);
                    if (!(f.score >= 36 || proper)) continue;
                    ++n_c;
                    // sort.Sort by score descending; ties keep list order here (insertion sort up to 12 candidates does; above: unpinned)
                    if (f.score > best_s) { second = best; second_s = best_s; best = g; best_s = f.score; best_proper = proper; }
                    else if (f.score > second_s) { second = g; second_s = f.score; }
                }
                if (n_c > 0) {
                    const FullAln c = rec_full_aln(A, best);
                    const double p_base = A.aux[ca].base;
                    double mapq;
                    // scoreAlignment(primary, nil, 0) + psuedoCountAlignmentScore(c0, 0)
                    double second_best = (p_base + A.improper + (pc.active_molecule ? 0.0 : 0.0)) + (0.0 - 10.0 - ((double)read_len - 25.0) * 0.5);
                    if (n_c > 1) {
                        mapq = (double)(best_s - second_s);
                        const FullAln c1 = rec_full_aln(A, second);
                        Hot h1; h1.reversed = (uint8_t)c1.reversed; h1.rid = c1.rid; h1.pos = c1.pos;
                        const bool paired = rec_is_pair(A.hot[ca].reversed, A.hot[ca].rid, A.hot[ca].pos, h1);
                        second_best = p_base + c1.base + (paired ? 0.0 : A.improper);
                    } else mapq = (double)best_s;
                    long long cs = -1, ce = -1;
                    if (A.centro_start && c.rid >= 0) { cs = A.centro_start[c.rid]; ce = A.centro_end[c.rid]; }
                    if (c.pos > cs && c.pos <= ce) mapq = 0.0;
                    if (mapq > 60) mapq = 60;
                    sreg = (int)best; smapq = (int)mapq; sproper = best_proper; sxs = second_best;
                    sas = c.base + pmate_base + (rec_is_pair(c.reversed, c.rid, c.pos, pmate) ? 0.0 : A.improper);  // scoreAlignment(split, active.mate_alignment, 0)
                }
            }
        }
        R.split_reg[r] = sreg; R.split_mapq[r] = smapq; R.split_proper[r] = sproper; R.split_as[r] = sas; R.split_xs[r] = sxs;
        R.rec_cnt[r] = sreg >= 0 ? 2ull : 1ull;
        if (r == A.n_reads - 1) R.rec_cnt[A.n_reads] = 0;
    }
}

// ---- AppendBam fields ----
struct RecAln {  // the fields AppendBam reads from an *Alignment
    int reg, rid, reversed, score, mapq, is_proper, active_molecule, molecule_id, n_mm, indel_len, read_len;
    int64_t pos, aend;
    int s_lead, s_trail;  // leading / trailing soft clip of the forward-strand CIGAR (HardClip)
};

__global__ void k_rec_emit(RecArgs R, const uint32_t* __restrict__ cig_pool) {
    const PlaceArgs& A = R.P;
    const int n_pairs = A.n_reads >> 1;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_pairs; p += gridDim.x * blockDim.x) {
        const int b = place_bc_of_read(A, 2 * p);
        const int attach = R.attach_bx ? R.attach_bx[b] != 0 : (A.run_rfa && A.run_rfa[b]);
        int64_t eff_pos[2]; int eff_mapq[2];  // pos / mapq of the two active alignments as AppendBam has left them so far
        for (int k = 0; k < 2; ++k) { const Cand c = A.cands[R.act[2 * p + k]]; eff_pos[k] = c.pos; eff_mapq[k] = c.mapq; }
        for (int k = 0; k < 2; ++k) {
            const int r = 2 * p + k;
            const int ca = R.act[r], cm = R.act[r ^ 1];
            const Cand pc = A.cands[ca], mc = A.cands[cm];
            const Hot ph = A.hot[ca], mh = A.hot[cm];
            const int p_score = pc.reg >= 0 ? A.regs[pc.reg].score : 0, m_score = mc.reg >= 0 ? A.regs[mc.reg].score : 0;
            const int read_len = (int)(A.seq_off[r + 1] - A.seq_off[r]);
            const int sreg = R.split_reg[r];
            Record* out = R.recs + R.rec_off[r];
            for (int which = 0; which < (sreg >= 0 ? 2 : 1); ++which) {
                const bool is_primary = which == 0;
                // ---- the alignment being written ----
                int a_reg, a_rid, a_rev, a_score, a_proper, a_mapq, a_am, a_mol; int64_t a_pos, a_aend;
                if (is_primary) {
                    a_reg = pc.reg; a_rid = ph.rid; a_rev = ph.reversed; a_score = p_score; a_proper = pc.is_proper; a_am = pc.active_molecule; a_mol = pc.molecule_id;
                    a_pos = pc.pos; a_aend = pc.aend; a_mapq = pc.mapq;
                } else {
                    const FullAln f = rec_full_aln(A, sreg);
                    a_reg = sreg; a_rid = f.rid; a_rev = f.reversed; a_score = f.score; a_proper = R.split_proper[r]; a_pos = f.pos; a_aend = f.aend; a_mapq = R.split_mapq[r];
                    a_am = 0; a_mol = -1;
                    for (long c = (long)A.cand_off[r], e = (long)A.cand_off[r + 1]; c < e; ++c)  // a split alignment that is also a candidate carries its placement state
                        if (A.cands[c].reg == sreg) { a_am = A.cands[c].active_molecule; a_mol = A.cands[c].molecule_id; }
                }
                if (!a_proper && a_score - 17 < 19) { a_pos = -1; a_mapq = 0; if (is_primary) { eff_pos[k] = -1; eff_mapq[k] = 0; } }
                // ---- primary.mate_alignment as it stands now ----
                const int64_t m_pos = eff_pos[k ^ 1];
                Record rec;
                rec.read = r; rec.reg = a_reg; rec.primary = is_primary; rec.mate_rid = -1; rec.mate_pos = -1; rec.tlen = 0;
                int flags = 1;
                Hot mnow = mh; mnow.pos = m_pos;
                if (a_proper) {
                    if (is_primary) flags |= 0x2;
                    else if (rec_is_pair(a_rev, a_rid, a_pos, mnow)) flags |= 0x2;
                }
                if (m_pos == -1 || (!pc.is_proper && m_score - 17 < 19)) flags |= 0x8;
                else { if (mh.reversed) flags |= 0x20; rec.mate_rid = mh.rid; rec.mate_pos = m_pos; }
                flags |= (r & 1) ? 0x80 : 0x40;
                const int dup = is_primary ? pc.duplicate : 0;
                if (!is_primary) for (long c = (long)A.cand_off[r], e = (long)A.cand_off[r + 1]; c < e; ++c) if (A.cands[c].reg == sreg && A.cands[c].duplicate) flags |= 0x400;
                if (dup) flags |= 0x400;
                if (m_pos == -1) rec.mate_rid = -1;
                else if (is_primary && a_rid == mh.rid && (pc.is_proper || m_score - 17 >= 19))
                    rec.tlen = a_rev ? -(a_aend - m_pos) : mc.aend - a_pos;
                if (!is_primary) flags |= 256;
                rec.rid = a_rid; rec.mapq = a_mapq;
                if (a_pos == -1) { flags |= 0x4; rec.mapq = 0; rec.rid = -1; }
                if (a_rev) flags |= 0x10;
                rec.flag = flags; rec.duplicate = (flags & 0x400) != 0; rec.reversed = a_rev; rec.pos = a_pos;
                rec.clip_start = 0; rec.clip_end = read_len; rec.hard = 0;
                if (!is_primary) {
                    rec.hard = 1;
                    const AlnRec ar = A.alns[a_reg];
                    if (ar.n_cigar >= 1 && (cig_pool[ar.cigar_off] & 0xf) == 3) rec.clip_start = (int)(cig_pool[ar.cigar_off] >> 4);
                    if (ar.n_cigar >= 2 && (cig_pool[ar.cigar_off + ar.n_cigar - 1] & 0xf) == 3) rec.clip_end -= (int)(cig_pool[ar.cigar_off + ar.n_cigar - 1] >> 4);
                }
                // ---- tags ----
                const int sb = is_primary ? pc.second_best : -1;
                rec.as = (int32_t)(is_primary ? pc.as_score : R.split_as[r]);
                rec.xs = (int32_t)(is_primary ? pc.xs_score : R.split_xs[r]);
                rec.has_xc = sb >= 0; rec.xc_reg = sb >= 0 ? A.cands[sb].reg : -1;
                rec.xm = sb >= 0 && A.cands[sb].active_molecule; rec.am = a_am; rec.xt = sb >= 0 && a_mol == A.cands[sb].molecule_id;
                // SA: the other alignment of the (primary, split) couple, if it is still mapped
                rec.sa_reg = -1; rec.sa_rid = rec.sa_rev = rec.sa_hard = rec.sa_mapq = rec.sa_nm = 0; rec.sa_pos = 0;
                if (sreg >= 0) {
                    int o_reg, o_rid, o_rev, o_mapq; int64_t o_pos;
                    if (is_primary) { const FullAln f = rec_full_aln(A, sreg); o_reg = sreg; o_rid = f.rid; o_rev = f.reversed; o_pos = f.pos; o_mapq = R.split_mapq[r]; }
                    else { o_reg = pc.reg; o_rid = ph.rid; o_rev = ph.reversed; o_pos = eff_pos[k]; o_mapq = eff_mapq[k]; }
                    if (o_pos > -1) {
                        rec.sa_reg = o_reg; rec.sa_rid = o_rid; rec.sa_rev = o_rev; rec.sa_pos = o_pos; rec.sa_mapq = o_mapq; rec.sa_hard = is_primary;
                        const AlnRec oa = A.alns[o_reg];
                        int indel = 0;
                        for (int x = 0; x < oa.n_cigar; ++x) { const uint32_t w = cig_pool[oa.cigar_off + x]; if ((w & 0xf) == 1 || (w & 0xf) == 2) indel += (int)(w >> 4); }
                        rec.sa_nm = oa.n_mm + indel;
                    }
                }
                rec.attach_bx = attach; rec.has_dm = attach && a_am; rec.dm = 0.0;
                if (is_primary) rec.dm = pc.molecule_difference;
                else for (long c = (long)A.cand_off[r], e = (long)A.cand_off[r + 1]; c < e; ++c) if (A.cands[c].reg == sreg) rec.dm = A.cands[c].molecule_difference;
                out[which] = rec;
            }
        }
    }
}

}  // namespace arachne
