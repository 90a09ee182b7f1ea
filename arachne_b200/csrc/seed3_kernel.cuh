// k_seed3: mem_collect_intv (bwa/bwamem.c:114-162) with the automaton's states UNBOUND from lanes.
//
// k_seed2 binds one read to one lane; its ncu capture in the timed configuration (profiles/r2l_*) shows what that costs: every
// trip of a warp executes the code of every kind of step its 32 lanes happen to be in (forward-window setup, occ popcounts,
// table steps, list packing, direct extension ...): 38.9 G warp instructions at 12.2 lanes per instruction, issue-bound at
// 0.36 of the HBM copy peak.  Here a block keeps ~800 read states in shared memory (structure of arrays) and one queue per KIND
// of step; a warp pops up to 32 states of ONE kind, executes that kind's code convergently, and pushes each state to the queue
// of its next kind.  The step semantics (what is looked up, what is pushed, the shortcuts and their counters) are k_seed2's,
// statement for statement; only the scheduling differs.  Kinds:
//   SETUP  bookkeeping that ends or starts passes (work fetch, scans over ambiguous bases, re-seed starts, pass ends)
//   WIN    table window of a fresh forward pass (cp.async of levels 1..K into the state's list head, replay)
//   OCCF   one bwt_extend of a forward pass or of the LAST-like pass        OCCB  one bwt_extend of a backward pass
//   TAB    a backward extension answered by the k-mer table
//   SA     SA / inverse-SA entry of the direct extension                     TXT   its text comparison
#pragma once
#include "seed_kernel.cuh"

namespace arachne {

enum { Q_SETUP = 0, Q_WIN, Q_OCCF, Q_OCCB, Q_TAB, Q_SA, Q_TXT, Q_N, Q_DEAD = 100 };
constexpr int S3_LH = 8, S3_WORDS = 16;
constexpr unsigned S3_RING = 1024;  // slots per queue ring (power of two, >= states per block)

struct S3Plan { int ns, nwq; size_t smem; };
inline S3Plan seed3_plan(int maxlen) {
    S3Plan p;
    p.nwq = (maxlen > 0 ? (maxlen - 1) / 32 : 0) + 2;
    const int per_state = S3_WORDS * 4 + p.nwq * 12 + S3_LH * 16;
    p.ns = ((227 * 1024 - 1024 - (int)(Q_N * S3_RING * 2)) / per_state) / 32 * 32;
    if (p.ns > (int)S3_RING) p.ns = (int)S3_RING;
    p.smem = (size_t)p.ns * per_state + Q_N * S3_RING * 2 + 256;
    return p;
}

struct S3 {  // a state in registers
    uint64_t kx0, kx1, kx2, lc;  // ik; lastcur_x2 of the backward pass / text position of the direct extension
    int r, min_intv, phase, fresh, len, x, ret, i, n, lo, hi, j, ncur, last_start, n_rec, rec_i, emit;
};

template <int NT>
__global__ void __launch_bounds__(NT, 1) k_seed3(FMIndex fm, MemOpt opt, const int64_t* __restrict__ off, int n_reads, int maxlen, uint4* lists,
                                                         int rows, SeedOut out, unsigned int* next, Counters* cnt, int use_kmer, int ns,
                                                         const ulonglong2* __restrict__ qpk, int qstride, int nwq, int use_direct) {
    extern __shared__ uint4 s3_smem[];
    uint4* sl = s3_smem;                                                   // [S3_LH][ns]
    uint64_t* qv = reinterpret_cast<uint64_t*>(sl + (size_t)S3_LH * ns);   // [nwq][ns]
    uint32_t* stw = reinterpret_cast<uint32_t*>(qv + (size_t)nwq * ns);    // [S3_WORDS][ns]
    uint32_t* qm = stw + (size_t)S3_WORDS * ns;                            // [nwq][ns]
    unsigned short* qslot = reinterpret_cast<unsigned short*>(qm + (size_t)nwq * ns);  // [Q_N][S3_RING]: state id + 1, 0 = not written yet
    __shared__ unsigned q_head[Q_N], q_tail[Q_N];
    __shared__ unsigned n_dead;
    const unsigned lane = threadIdx.x & 31u;
    const size_t T = (size_t)gridDim.x * ns;
    for (int k = threadIdx.x; k < (int)(Q_N * S3_RING); k += blockDim.x) qslot[k] = 0;
    if (threadIdx.x < Q_N) { q_head[threadIdx.x] = 0; q_tail[threadIdx.x] = threadIdx.x == Q_SETUP ? (unsigned)ns : 0u; }
    if (threadIdx.x == 0) n_dead = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < ns; k += blockDim.x) {  // every state starts idle, queued for SETUP
        qslot[Q_SETUP * S3_RING + k] = (unsigned short)(k + 1);
        stw[9 * ns + k] = (uint32_t)SP_IDLE << 8;
    }
    __syncthreads();

    const int split_len = (int)(opt.min_seed_len * opt.split_factor + .499f);
    const int kmer_k = ((use_kmer & 1) && fm.kmer && fm.kmer_k <= opt.min_seed_len) ? fm.kmer_k : 0;
    const int kf_cap = (use_kmer >> 8) & 15;
    const int kf = min(min(kmer_k, S3_LH), kf_cap ? kf_cap : 8);
    const bool kmer_fwd = kmer_k && fm.kmer_all && kf >= 4;
    const bool kmer_bwd = kmer_k && fm.kmer_all;
    const bool direct = use_direct && fm.isa != nullptr && fm.text != nullptr && fm.sa_intv == 1;
    unsigned st_calls = 0, st_blocks = 0;

    for (;;) {
        // ------------------------------------------------------------------ pick a kind and pop up to 32 of its states
        int K = -1; unsigned base = 0, take = 0;
        for (;;) {
            // lanes 0..Q_N-1 look at one queue each; a full batch of real work beats bookkeeping, otherwise the fullest queue
            unsigned a = 0, h = 0;
            if (lane < Q_N) { h = *(volatile unsigned*)&q_head[lane]; a = *(volatile unsigned*)&q_tail[lane] - h; if (a > (unsigned)ns) a = 0; }
            const unsigned full = __ballot_sync(0xffffffffu, a >= 32u) & ~1u;  // bit 0 = SETUP
            const unsigned score = a >= 32u && lane != Q_SETUP ? 0x40000000u + a : a;
            const unsigned best = __reduce_max_sync(0xffffffffu, score);
            (void)full;
            if (best == 0) {
                if (*(volatile unsigned*)&n_dead >= (unsigned)ns) break;
                __nanosleep(100);
                continue;
            }
            const int bk = __ffs(__ballot_sync(0xffffffffu, score == best)) - 1;
            const unsigned ba = __shfl_sync(0xffffffffu, a, bk), bh = __shfl_sync(0xffffffffu, h, bk);
            const unsigned want = ba < 32u ? ba : 32u;
            unsigned ok = 0;
            if (lane == 0) ok = atomicCAS(&q_head[bk], bh, bh + want) == bh;
            ok = __shfl_sync(0xffffffffu, ok, 0);
            if (ok) { K = bk; base = bh; take = want; break; }
        }
        if (K < 0) break;
        const bool act = lane < take;
        int sid = 0;
        if (act) {
            volatile unsigned short* slot = qslot + (unsigned)K * S3_RING + ((base + lane) & (S3_RING - 1u));
            unsigned v;
            while ((v = *slot) == 0) {}
            *slot = 0;
            sid = (int)v - 1;
        }
        __threadfence_block();
        // ------------------------------------------------------------------ the state, its read and its list
        S3 s;
        s.kx0 = s.kx1 = s.kx2 = s.lc = 0; s.r = 0; s.min_intv = 1; s.phase = SP_IDLE; s.fresh = 0; s.len = s.x = s.ret = s.i = s.n = s.lo = s.hi = s.j = s.ncur = 0;
        s.last_start = 0xffff; s.n_rec = s.rec_i = s.emit = 0;
        uint32_t* w = stw + sid;
        if (act) {
            s.kx0 = (uint64_t)w[1 * ns] << 32 | w[0]; s.kx1 = (uint64_t)w[3 * ns] << 32 | w[2 * ns]; s.kx2 = (uint64_t)w[5 * ns] << 32 | w[4 * ns];
            s.lc = (uint64_t)w[7 * ns] << 32 | w[6 * ns]; s.r = (int)w[8 * ns];
            uint32_t v = w[9 * ns]; s.min_intv = v & 0xff; s.phase = (v >> 8) & 0xff; s.fresh = (v >> 16) & 0xff; s.len = v >> 24;
            v = w[10 * ns]; s.x = v & 0xffff; s.ret = v >> 16;
            v = w[11 * ns]; s.i = (int)(v & 0xffff) - 1; s.n = v >> 16;
            v = w[12 * ns]; s.lo = v & 0xffff; s.hi = v >> 16;
            v = w[13 * ns]; s.j = v & 0xffff; s.ncur = v >> 16;
            v = w[14 * ns]; s.last_start = v & 0xffff; s.n_rec = v >> 16;
            v = w[15 * ns]; s.rec_i = v & 0xffff; s.emit = v >> 16;
        }
        const size_t gsid = (size_t)blockIdx.x * ns + sid;
        uint4* list = lists + gsid;                                // entry e >= S3_LH at list[e * T]
        uint4* recs = lists + (size_t)(rows >> 1) * T + gsid;      // re-seed records in the upper half
        uint4* slh = sl + sid;
        auto lput = [&](int e, uint4 v) { if (e < S3_LH) slh[(unsigned)e * (unsigned)ns] = v; else list[(size_t)e * T] = v; };
        auto lget = [&](int e) -> uint4 { return e < S3_LH ? slh[(unsigned)e * (unsigned)ns] : list[(size_t)e * T]; };
        const uint64_t* qvs = qv + sid; const uint32_t* qms = qm + sid;
        auto qb = [&](int p) -> int {
            const int wd = p >> 5, sft = p & 31;
            return ((qms[(unsigned)wd * (unsigned)ns] >> (31 - sft)) & 1u) ? 4 : (int)((qvs[(unsigned)wd * (unsigned)ns] >> (62 - 2 * sft)) & 3);
        };
        auto qwin = [&](int p) -> uint64_t { const int wd = p >> 5; return fsh64(qvs[(unsigned)wd * (unsigned)ns], qvs[(unsigned)(wd + 1) * (unsigned)ns], (p & 31) * 2); };
        auto mwin = [&](int p) -> uint32_t { const int wd = p >> 5; return __funnelshift_l(qms[(unsigned)(wd + 1) * (unsigned)ns], qms[(unsigned)wd * (unsigned)ns], p & 31); };
        auto stat = [&](unsigned calls, unsigned blocks) { st_calls += calls; st_blocks += blocks; };
        auto emit = [&](uint64_t e0, uint64_t e1, uint64_t e2, int start, int end, bool filter) {
            const int slen = end - start;
            if (filter && slen < opt.min_seed_len) return;
            const int kind = filter ? 0 : 1, slot = s.emit;
            if (slot >= (out.cap >> 1)) { raise(cnt, ARACHNE_ERR_OVERFLOW, s.r); return; }
            s.emit = slot + 1;
            out.cnt2[2 * s.r + kind] = slot + 1;
            Intv v; v.x0 = e0; v.x1 = e1; v.x2 = e2; v.info = (uint64_t)start << 32 | (unsigned)end;
            out.intv[(size_t)s.r * out.cap + kind * (out.cap >> 1) + slot] = v;
            if (s.rec_i == 0 && filter && slen >= split_len && e2 <= (uint64_t)opt.split_width)
                recs[(size_t)(s.n_rec++) * T] = make_uint4((unsigned)start, (unsigned)end, (unsigned)e2, 0u);
        };
        auto start_bwd = [&]() { s.hi = s.n; s.lo = 0; s.j = s.hi - 1; s.ncur = 0; s.i = s.x - 1; s.phase = SP_BWD; };
        auto end_pass = [&]() { if (s.rec_i == 0) { s.x = s.ret; s.phase = SP_P1_SCAN; } else s.phase = SP_P2_NEXT; };
        // what the state needs next once it sits in one of the stepping phases (k_seed2's PRE for SP_FWD / SP_BWD / SP_P3 / direct);
        // every other phase goes through SETUP
        auto next_kind = [&]() -> int {
            for (;;) {
                if (s.phase == SP_FWD) {
                    if (s.fresh) return Q_SETUP;
                    if (qb(s.i) < 4) return Q_OCCF;
                    lput(s.n++, pk_pack(s.kx0, s.kx1, s.kx2, (unsigned)s.i));  // ambiguous base or end of read closes the forward pass (bwt.c:311-316)
                    s.ret = s.i; start_bwd();
                    continue;
                }
                if (s.phase == SP_BWD) {
                    const int b = s.i < 0 ? 4 : qb(s.i);
                    if (b > 3) {  // no extension possible: only the longest survivor can be a new SMEM (bwt.c:331-341)
                        const uint4 praw = lget(s.j);
                        if (s.ncur == 0 && s.j == s.hi - 1 && s.i + 1 < s.last_start) {
                            uint64_t p0, p1, p2; unsigned pe;
                            pk_unpack(praw, p0, p1, p2, pe);
                            emit(p0, p1, p2, s.i + 1, (int)(praw.w >> 24), true);
                        }
                        end_pass();
                        return Q_SETUP;
                    }
                    const uint4 praw = lget(s.j);
                    const int L = (int)(praw.w >> 24) - s.i;
                    return (kmer_bwd && L <= kmer_k) ? Q_TAB : Q_OCCB;
                }
                if (s.phase == SP_P3) {
                    const int b = s.i < s.len ? qb(s.i) : 5;
                    if (b < 4) return Q_OCCF;
                    if (b == 4) { s.x = s.i + 1; s.phase = SP_P3_SCAN; } else s.phase = SP_IDLE;
                    return Q_SETUP;
                }
                if (s.phase == SP_FSA || s.phase == SP_FISA) return Q_SA;
                if (s.phase == SP_FTXT) return Q_TXT;
                if (s.phase == SP_DONE) return Q_DEAD;
                return Q_SETUP;
            }
        };
        int nk = Q_DEAD;

        if (K == Q_SETUP) {
            // ---- k_seed2's PRE for the phases that start or end things; runs until the state needs a step ----
            if (act) {
                for (;;) {
                    if (s.phase == SP_IDLE) {
                        // one queue per task kind (next[0]: SMEM passes, next[1]: LAST-like pass); which one a state tries first alternates 1 : 2 / 1 : 5
                        int kind = (opt.max_mem_intv > 0 && (gsid % (kmer_k ? 6 : 3)) == 2) ? 1 : 0;
                        unsigned t = atomicAdd(next + kind, 1u);
                        if (t >= (unsigned)n_reads) {
                            kind ^= 1;
                            t = atomicAdd(next + kind, 1u);
                            if (t >= (unsigned)n_reads) { s.phase = SP_DONE; break; }
                        }
                        s.r = (int)t;
                        s.len = (int)(off[s.r + 1] - off[s.r]);
                        if (s.len > maxlen) { if (kind == 0) raise(cnt, ARACHNE_ERR_ARG, s.r); s.len = 0; continue; }
                        if (s.len < opt.min_seed_len) continue;
                        for (int wd = 0; wd < nwq; ++wd) {
                            const ulonglong2 pw = __ldg(qpk + (size_t)s.r * qstride + wd);
                            qv[(size_t)wd * ns + sid] = pw.x; qm[(size_t)wd * ns + sid] = (uint32_t)pw.y;
                        }
                        s.x = 0; s.n_rec = 0; s.rec_i = 0; s.emit = 0;
                        s.phase = kind ? ((opt.max_mem_intv > 0) ? SP_P3_SCAN : SP_IDLE) : SP_P1_SCAN;
                        continue;
                    }
                    if (s.phase == SP_P2_NEXT) {
                        if (s.rec_i < s.n_rec) {
                            const uint4 rc = recs[(size_t)(s.rec_i++) * T];
                            s.x = (int)((rc.x + rc.y) >> 1);
                            s.min_intv = (int)rc.z + 1;
                            const int b = qb(s.x);
                            if (b > 3) continue;  // bwt_smem1a returns at once on an ambiguous base
                            s.kx0 = fm_L2(fm, b) + 1; s.kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); s.kx1 = fm_L2(fm, 3 - b) + 1;
                            s.i = s.x + 1; s.n = 0; s.last_start = 0xffff;
                            s.phase = SP_FWD; s.fresh = 1;
                        } else { s.phase = SP_IDLE; continue; }
                    }
                    if (s.phase == SP_P1_SCAN) {
                        while (s.x < s.len && qb(s.x) > 3) ++s.x;
                        if (s.x >= s.len) { s.phase = SP_P2_NEXT; continue; }
                        const int b = qb(s.x);
                        s.kx0 = fm_L2(fm, b) + 1; s.kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); s.kx1 = fm_L2(fm, 3 - b) + 1;
                        s.min_intv = 1;
                        s.i = s.x + 1; s.n = 0; s.last_start = 0xffff;
                        s.phase = SP_FWD; s.fresh = 1;
                    }
                    if (s.phase == SP_P3_SCAN) {
                        bool idle = false;
                        for (;;) {
                            while (s.x < s.len && qb(s.x) > 3) ++s.x;
                            if (s.len - s.x <= opt.min_seed_len) { s.phase = SP_IDLE; idle = true; break; }
                            const int b = qb(s.x);
                            if (kmer_k) {
                                const int zn = __clz((int)mwin(s.x));
                                if (zn < kmer_k) { s.x += zn + 1; continue; }
                                const uint32_t km = (uint32_t)(qwin(s.x) >> (64 - 2 * kmer_k));
                                unsigned spare;
                                pk_unpack(__ldg(reinterpret_cast<const uint4*>(fm.kmer) + km), s.kx0, s.kx1, s.kx2, spare);
                                s.i = s.x + kmer_k;
                                stat(kmer_k - 1, 2 * (kmer_k - 1));
                            } else {
                                s.kx0 = fm_L2(fm, b) + 1; s.kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); s.kx1 = fm_L2(fm, 3 - b) + 1;
                                s.i = s.x + 1;
                            }
                            s.phase = SP_P3;
                            break;
                        }
                        if (idle) continue;
                    }
                    if (s.phase == SP_FWD && s.fresh) {
                        // the table window of a fresh forward pass, when it pays (k_seed2: m >= 2 unambiguous bases ahead)
                        if (kmer_fwd) {
                            int m = __clz((int)mwin(s.x));
                            m = m < kf ? m : kf;
                            if (m >= 2) { nk = Q_WIN; break; }
                        }
                        s.fresh = 0;
                    }
                    nk = next_kind();
                    if (nk != Q_SETUP) break;
                }
                if (s.phase == SP_DONE) nk = Q_DEAD;
            }
        } else if (K == Q_WIN) {
            // ---- bwt_smem1a's forward loop (bwt.c:300-316) over the first m <= K bases from the table ----
            int m = 0;
            if (act) {
                m = __clz((int)mwin(s.x));
                m = m < kf ? m : kf;
                const uint32_t km = (uint32_t)(qwin(s.x) >> (64 - 2 * m));
                const uint4* all = reinterpret_cast<const uint4*>(fm.kmer_all);
                for (int L = 1; L <= m; ++L) cp_async16(&slh[(unsigned)(L - 1) * (unsigned)ns], all + klev(L) + (km >> (2 * (m - L))));
            }
            cp_async_wait_all();
            __syncwarp();
            if (act) {
                unsigned spare;
                pk_unpack(slh[0], s.kx0, s.kx1, s.kx2, spare);
                bool stop = false;
                int L = 1;
                for (; L < m; ++L) {
                    uint64_t a0, a1, a2;
                    pk_unpack(slh[(unsigned)L * (unsigned)ns], a0, a1, a2, spare);
                    if (a2 != s.kx2) {
                        lput(s.n++, pk_pack(s.kx0, s.kx1, s.kx2, (unsigned)(s.x + L)));
                        if (a2 < (uint64_t)s.min_intv) { stop = true; break; }
                    }
                    s.kx0 = a0; s.kx1 = a1; s.kx2 = a2;
                }
                stat((unsigned)(stop ? L : L - 1), 2u * (unsigned)(stop ? L : L - 1));
                s.fresh = 0;
                if (stop) { s.ret = s.x + L; start_bwd(); } else { s.i = s.x + m; s.phase = SP_FWD; }
                nk = next_kind();
            }
        } else if (K == Q_OCCF || K == Q_OCCB) {
            // ---- one bwt_extend (bwt.c:262-275) ----
            uint32_t R[16];
            uint64_t xb = 0; int kbit = 0, lbit = 0, c = 0; bool kneg = false, same = true, hasp = false;
            uint4 praw = make_uint4(0, 0, 0, 0);
            if (act) {
                uint64_t xa, x2;
                if (K == Q_OCCB) { unsigned pe; praw = lget(s.j); pk_unpack(praw, xa, xb, x2, pe); c = qb(s.i); }
                else { xa = s.kx1; xb = s.kx0; x2 = s.kx2; c = 3 - qb(s.i); }
                const uint64_t k = xa - 1, l = xa - 1 + x2;
                stat(1u, fm_ref_blocks(fm, k, l));
                kneg = k == (uint64_t)-1;
                hasp = xa <= fm.primary && xa + x2 - 1 >= fm.primary;
                const uint64_t l_ = l - (l >= fm.primary), k_ = kneg ? l_ : k - (k >= fm.primary);
                const uint64_t kb = k_ >> 6, lb = l_ >> 6;
                kbit = (int)(k_ & 63); lbit = (int)(l_ & 63);
                same = kb == lb;
                ld_sector(R, fm.blocks + kb);
                if (!same) ld_sector(R + 8, fm.blocks + lb);
            }
            __syncwarp();
            if (act) {
                uint64_t tk[4], tl[4];
                Blk32 bk, bl;
#pragma unroll
                for (int u = 0; u < 4; ++u) { bk.cnt[u] = R[u]; bk.w[u] = R[4 + u]; bl.cnt[u] = same ? R[u] : R[8 + u]; bl.w[u] = same ? R[4 + u] : R[12 + u]; }
                fm_occ4_blk(bk, (uint64_t)kbit, tk);
                fm_occ4_blk(bl, (uint64_t)lbit, tl);
                if (kneg) tk[0] = tk[1] = tk[2] = tk[3] = 0;
                const uint64_t ns3 = tl[3] - tk[3], ns2 = tl[2] - tk[2], ns1 = tl[1] - tk[1], ns0 = tl[0] - tk[0];
                uint64_t nb = xb + (hasp ? 1 : 0);
                if (c < 3) nb += ns3;
                if (c < 2) nb += ns2;
                if (c < 1) nb += ns1;
                const uint64_t ox_b = nb;
                const uint64_t ox_a = fm_L2(fm, c) + 1 + (c == 0 ? tk[0] : c == 1 ? tk[1] : c == 2 ? tk[2] : tk[3]);
                const uint64_t ox2 = c == 0 ? ns0 : c == 1 ? ns1 : c == 2 ? ns2 : ns3;
                if (s.phase == SP_FWD) {
                    bool stop = false;
                    if (ox2 != s.kx2) {
                        lput(s.n++, pk_pack(s.kx0, s.kx1, s.kx2, (unsigned)s.i));
                        if (ox2 < (uint64_t)s.min_intv) stop = true;
                    }
                    if (stop) { s.ret = s.i; start_bwd(); }
                    else {
                        s.kx0 = ox_b; s.kx1 = ox_a; s.kx2 = ox2;
                        ++s.i;
                        if (direct && s.kx2 == 1 && s.min_intv <= 1 && s.len - s.i >= DIRECT_MIN_LEFT) s.phase = SP_FSA;
                    }
                } else if (s.phase == SP_BWD) {
                    if (ox2 < (uint64_t)s.min_intv) {
                        if (s.ncur == 0 && s.i + 1 < s.last_start) {
                            uint64_t p0, p1, p2; unsigned pe;
                            pk_unpack(praw, p0, p1, p2, pe);
                            emit(p0, p1, p2, s.i + 1, (int)(praw.w >> 24), true); s.last_start = s.i + 1;
                        }
                    } else if (s.ncur == 0 || ox2 != s.lc) {
                        lput(s.hi - 1 - s.ncur, pk_pack(ox_a, ox_b, ox2, praw.w >> 24));
                        ++s.ncur; s.lc = ox2;
                    }
                    if (--s.j < s.lo) {
                        if (s.ncur == 0) end_pass();
                        else { s.lo = s.hi - s.ncur; s.ncur = 0; s.j = s.hi - 1; --s.i; }
                    }
                } else {  // SP_P3: bwt_seed_strategy1 (bwt.c:358-379)
                    if (ox2 < opt.max_mem_intv && s.i - s.x >= opt.min_seed_len) {
                        if (ox2 > 0) emit(ox_b, ox_a, ox2, s.x, s.i + 1, false);
                        s.x = s.i + 1; s.phase = SP_P3_SCAN;
                    } else { s.kx0 = ox_b; s.kx1 = ox_a; s.kx2 = ox2; ++s.i; }
                }
                nk = next_kind();
            }
        } else if (K == Q_TAB) {
            // ---- a backward extension whose result is at most kmer_k long: one table entry ----
            uint32_t R[4]; uint4 praw = make_uint4(0, 0, 0, 0);
            if (act) {
                praw = lget(s.j);
                const int L = (int)(praw.w >> 24) - s.i;
                const uint64_t code = qwin(s.i) >> (64 - 2 * L);
                const uint4* tp = reinterpret_cast<const uint4*>(fm.kmer_all) + klev(L) + code;
                stat(1u, 2u);
                asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(R[0]), "=r"(R[1]), "=r"(R[2]), "=r"(R[3]) : "l"(tp));
            }
            __syncwarp();
            if (act) {
                const uint64_t ox2 = ((uint64_t)R[3] << 32 | R[2]) >> 16 & 0xffffffffffull;
                if (ox2 < (uint64_t)s.min_intv) {
                    if (s.ncur == 0 && s.i + 1 < s.last_start) {
                        uint64_t p0, p1, p2; unsigned pe;
                        pk_unpack(praw, p0, p1, p2, pe);
                        emit(p0, p1, p2, s.i + 1, (int)(praw.w >> 24), true); s.last_start = s.i + 1;
                    }
                } else if (s.ncur == 0 || ox2 != s.lc) {
                    lput(s.hi - 1 - s.ncur, make_uint4(R[0], R[1], R[2], (R[3] & 0xffffffu) | (praw.w & 0xff000000u)));
                    ++s.ncur; s.lc = ox2;
                }
                if (--s.j < s.lo) {
                    if (s.ncur == 0) end_pass();
                    else { s.lo = s.hi - s.ncur; s.ncur = 0; s.j = s.hi - 1; --s.i; }
                }
                nk = next_kind();
            }
        } else if (K == Q_SA) {
            uint32_t R[2];
            if (act) {
                const uint64_t* a = s.phase == SP_FSA ? fm.sa + s.kx0 : fm.isa + (fm.seq_len - s.lc - (uint64_t)(s.i - s.x));
                asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(R[0]), "=r"(R[1]) : "l"(a));
            }
            __syncwarp();
            if (act) {
                const uint64_t got = (uint64_t)R[1] << 32 | R[0];
                if (s.phase == SP_FSA) { s.lc = got; s.phase = SP_FTXT; }
                else {
                    s.kx1 = got;
                    lput(s.n++, pk_pack(s.kx0, s.kx1, s.kx2, (unsigned)s.i));
                    s.ret = s.i; start_bwd();
                }
                nk = next_kind();
            }
        } else {  // Q_TXT
            uint32_t R[2 * DIRECT_WORDS + 2];
            if (act) {
                const uint64_t* tp = fm.text + ((s.lc + (uint64_t)(s.i - s.x)) >> 5);
#pragma unroll
                for (int d = 0; d <= DIRECT_WORDS; ++d)
                    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(R[2 * d]), "=r"(R[2 * d + 1]) : "l"(tp + d));
            }
            __syncwarp();
            if (act) {
                int dt = 32 * DIRECT_WORDS; bool dstop = false, dnstop = false;
                if (s.len - s.i < dt) { dt = s.len - s.i; dnstop = true; }
                const uint64_t ts = s.lc + (uint64_t)(s.i - s.x), left = fm.seq_len - ts;
                const int tsh = (int)(ts & 31) * 2;
#pragma unroll
                for (int d = DIRECT_WORDS - 1; d >= 0; --d) {
                    if (s.i + 32 * d >= s.len) continue;
                    const uint64_t td = fsh64((uint64_t)R[2 * d + 1] << 32 | R[2 * d], (uint64_t)R[2 * d + 3] << 32 | R[2 * d + 2], tsh);
                    const int zx = __clzll((long long)(td ^ qwin(s.i + 32 * d))) >> 1, zn = __clz((int)mwin(s.i + 32 * d));
                    if (zx < 32 || zn < 32) { dt = 32 * d + (zx < zn ? zx : zn); dnstop = zn <= zx; }
                }
                dstop = dt < 32 * DIRECT_WORDS;
                if (left < (uint64_t)dt || (left == (uint64_t)dt && !dstop)) { dt = (int)left; dnstop = false; dstop = true; }
                const unsigned calls = (unsigned)dt + ((dstop && !dnstop) ? 1u : 0u);
                stat(calls, calls);
                s.i += dt;
                if (dstop) s.phase = SP_FISA;
                nk = next_kind();
            }
        }
        // ------------------------------------------------------------------ write the state back, queue it for its next kind
        if (act) {
            w[0] = (uint32_t)s.kx0; w[1 * ns] = (uint32_t)(s.kx0 >> 32); w[2 * ns] = (uint32_t)s.kx1; w[3 * ns] = (uint32_t)(s.kx1 >> 32);
            w[4 * ns] = (uint32_t)s.kx2; w[5 * ns] = (uint32_t)(s.kx2 >> 32); w[6 * ns] = (uint32_t)s.lc; w[7 * ns] = (uint32_t)(s.lc >> 32);
            w[8 * ns] = (uint32_t)s.r;
            w[9 * ns] = (uint32_t)(s.min_intv & 0xff) | (uint32_t)s.phase << 8 | (uint32_t)s.fresh << 16 | (uint32_t)s.len << 24;
            w[10 * ns] = (uint32_t)s.x | (uint32_t)s.ret << 16;
            w[11 * ns] = (uint32_t)(s.i + 1) | (uint32_t)s.n << 16;
            w[12 * ns] = (uint32_t)s.lo | (uint32_t)s.hi << 16;
            w[13 * ns] = (uint32_t)s.j | (uint32_t)s.ncur << 16;
            w[14 * ns] = (uint32_t)(s.last_start & 0xffff) | (uint32_t)s.n_rec << 16;
            w[15 * ns] = (uint32_t)s.rec_i | (uint32_t)s.emit << 16;
        }
        __threadfence_block();
        if (!act) nk = -1;
        {
            const unsigned grp = __match_any_sync(0xffffffffu, nk);   // the lanes that go to the same queue
            const int leader = __ffs(grp) - 1;
            const bool queued = nk >= 0 && nk < Q_N;
            unsigned b = 0;
            if (queued && (int)lane == leader) b = atomicAdd(&q_tail[nk], (unsigned)__popc(grp));
            b = __shfl_sync(0xffffffffu, b, leader);
            if (queued) {
                volatile unsigned short* slot = qslot + (unsigned)nk * S3_RING + ((b + __popc(grp & ((1u << lane) - 1u))) & (S3_RING - 1u));
                while (*slot != 0) {}  // (a slot is free again once its previous occupant was popped: at most ns <= S3_RING states exist)
                *slot = (unsigned short)(sid + 1);
            }
        }
        const unsigned dm = __ballot_sync(0xffffffffu, nk == Q_DEAD);
        if (dm && lane == 0) atomicAdd(&n_dead, (unsigned)__popc(dm));
    }
    if (st_calls) { atomicAdd(&cnt->occ, (unsigned long long)st_calls); atomicAdd(&cnt->occ_blocks, (unsigned long long)st_blocks); }
}

}  // namespace arachne
