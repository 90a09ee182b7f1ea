// Host orchestration of one batch (included inside pipeline.cu's extern "C" block).
// All device arenas live in the index handle's Workspace and are reused by later batches; results are copied into
// pinned host buffers recycled through arachne_batch_free.

static Workspace* get_ws(arachne_index_t* idx, int slot) {
    if (!idx->ws[slot]) {
        auto* w = new Workspace();
        if (w->init()) { w->destroy(); delete w; return nullptr; }
        idx->ws[slot] = w;
    }
    return reinterpret_cast<Workspace*>(idx->ws[slot]);
}

// Take a free workspace slot: the first idle one; when all are busy, sleep until a batch hands its slot back.
struct WsLease {
    arachne_index_t* idx = nullptr; int slot = -1;
    WsLease() = default;
    WsLease(WsLease&& o) noexcept : idx(o.idx), slot(o.slot) { o.idx = nullptr; }
    WsLease(const WsLease&) = delete;
    ~WsLease() {
        if (!idx) return;
        { std::lock_guard<std::mutex> g(idx->ws_gate); idx->ws_busy[slot] = false; }
        idx->ws_cv.notify_one();
    }
};
static WsLease take_ws(arachne_index_t* idx, int* slot) {
    std::unique_lock<std::mutex> g(idx->ws_gate);
    for (;;) {
        for (int k = 0; k < arachne_index::N_WS; ++k)
            if (!idx->ws_busy[k]) { idx->ws_busy[k] = true; *slot = k; WsLease l; l.idx = idx; l.slot = k; return l; }
        idx->ws_cv.wait(g);
    }
}

int arachne_index_set_option(arachne_index_t* idx, const char* name, int64_t value) {
    if (!idx || !name) { set_error("arachne_index_set_option: null argument"); return ARACHNE_ERR_ARG; }
    std::lock_guard<std::mutex> lk(idx->mu);  // takes effect with the next batch that starts
    if (!strcmp(name, "seed_kmer")) idx->opt_seed_kmer = (idx->opt_seed_kmer & ~0xff) | (int)(value & 7);  // bit 0: tables on; bit 2: the forward window's table copies complete in a trip of their own
    else if (!strcmp(name, "seed_blocks_per_sm") && value >= 1 && value <= 4) idx->opt_seed_blocks = (int)value;
    else if (!strcmp(name, "seed_fwd_levels") && value >= 0 && value <= 15) idx->opt_seed_kmer = (idx->opt_seed_kmer & 0xff) | (int)value << 8;
    else if (!strcmp(name, "seed_direct")) idx->opt_seed_direct = value != 0;
    else if (!strcmp(name, "seed_impl") && (value == 2 || value == 3)) idx->opt_seed_impl = (int)value;
    else if (!strcmp(name, "seed3_threads") && (value == 512 || value == 768 || value == 1024)) idx->opt_seed3_threads = (int)value;
    else if (!strcmp(name, "seed_shared_list") && value >= -1 && value <= 16) idx->opt_seed_slist = (int)value;
    else { set_error("arachne_index_set_option: unknown option or value: %s=%lld", name, (long long)value); return ARACHNE_ERR_ARG; }
    return ARACHNE_OK;
}

int arachne_index_set_stream(arachne_index_t* idx, void* cuda_stream, int use) {
    if (!idx) { set_error("arachne_index_set_stream: null index"); return ARACHNE_ERR_ARG; }
    std::lock_guard<std::mutex> g(idx->mu);
    idx->user_stream = cuda_stream;
    idx->has_user_stream = use != 0;
    return ARACHNE_OK;
}

int arachne_align_batch(arachne_index_t* idx, const arachne_mem_opt_t* opt_, const arachne_batch_in_t* in, arachne_batch_out_t** out_) {
    if (!idx || !in || !out_ || !in->seq_off || (!in->seq && in->n_pairs > 0)) { set_error("arachne_align_batch: null argument"); return ARACHNE_ERR_ARG; }
    int rc = validate_opts(opt_);
    if (rc) return rc;
    MemOpt opt;
    memcpy(&opt, opt_, sizeof(opt));
    const bool in_dev = (in->flags & ARACHNE_IN_DEVICE) != 0, out_dev = (in->flags & ARACHNE_OUT_DEVICE) != 0;
    const int64_t n_reads = in->n_pairs * 2;
    if (n_reads < 0 || n_reads > 0x7fffffff / 2) { set_error("batch too large"); return ARACHNE_ERR_ARG; }
    int64_t n_bases = 0;
    int maxlen = 0;
    if (in_dev) {  // device-resident input: lengths cannot be checked on the host; the caller states the bounds
        if (in->max_read_len <= 0 || in->max_read_len > ARACHNE_MAX_READ_LEN || in->n_bases < 0) {
            set_error("ARACHNE_IN_DEVICE needs max_read_len in 1..%d and n_bases", ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG;
        }
        if (reinterpret_cast<uintptr_t>(in->seq) & 7) { set_error("ARACHNE_IN_DEVICE: seq must be 8-byte aligned"); return ARACHNE_ERR_ARG; }
        maxlen = in->max_read_len; n_bases = in->n_bases;
    } else {
        n_bases = n_reads ? in->seq_off[n_reads] : 0;
        for (int64_t i = 0; i < n_reads; ++i) {
            int64_t l = in->seq_off[i + 1] - in->seq_off[i];
            if (l < 0 || l > ARACHNE_MAX_READ_LEN) { set_error("read %lld has length %lld; supported 0..%d", (long long)i, (long long)l, ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG; }
            if (l > maxlen) maxlen = (int)l;
        }
    }
    if ((long)opt.a * maxlen > 8191) { set_error("a * read length = %ld exceeds the extension kernel's 13-bit score fields", (long)opt.a * maxlen); return ARACHNE_ERR_OPTS; }
    int ws_slot = 0;
    WsLease guard = take_ws(idx, &ws_slot);
    AR_CUDA(cudaSetDevice(idx->device));
    Workspace* wsp = get_ws(idx, ws_slot);
    if (!wsp) { set_error("cannot create the batch workspace (stream/events/pinned memory)"); return ARACHNE_ERR_CUDA; }
    Workspace& W = *wsp;
    const int sms = sm_count(idx->device);
    cudaStream_t st;
    int use_seed_kmer, seed_bpsm_opt, seed_slist, use_seed_direct, seed_impl;
    BatchPriv* priv = nullptr;
    {
        std::lock_guard<std::mutex> g(idx->mu);
        st = idx->has_user_stream ? reinterpret_cast<cudaStream_t>(idx->user_stream) : W.own_stream;
        use_seed_kmer = idx->opt_seed_kmer; seed_bpsm_opt = idx->opt_seed_blocks; seed_slist = idx->opt_seed_slist; use_seed_direct = idx->opt_seed_direct; seed_impl = idx->opt_seed_impl;
        if (!idx->priv_pool.empty()) { priv = reinterpret_cast<BatchPriv*>(idx->priv_pool.back()); idx->priv_pool.pop_back(); }
    }
    cudaEvent_t* ev = W.ev;  // seed | sa | chain | extend | rescue | gather | cigar | end

    auto* out = new arachne_batch_out_t();
    memset(out, 0, sizeof(*out));
    if (!priv) priv = new BatchPriv();
    priv->owner = idx;
    out->priv = priv;
    out->n_reads = n_reads;
    auto fail = [&](int code) { { std::lock_guard<std::mutex> g(idx->mu); idx->priv_pool.push_back(priv); } delete out; return code; };
#define AR_TRY(x) do { const cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); return fail(ARACHNE_ERR_CUDA); } } while (0)
#define AR_ALLOC(buf, bytes) do { if ((buf).ensure(bytes)) { set_error("out of device memory allocating %zu bytes (%s)", (size_t)(bytes), #buf); return fail(ARACHNE_ERR_CUDA); } } while (0)
#define AR_PIN(buf, bytes) do { if ((buf).ensure(bytes)) { set_error("out of pinned host memory allocating %zu bytes (%s)", (size_t)(bytes), #buf); return fail(ARACHNE_ERR_CUDA); } } while (0)

    AR_PIN(priv->reg_off, (n_reads + 1) * 8);
    if (n_reads == 0) { priv->reg_off.as<int64_t>()[0] = 0; out->reg_off = priv->reg_off.as<int64_t>(); *out_ = out; return ARACHNE_OK; }

    const int L1 = maxlen + 1;
    const uint8_t* d_seq; const int64_t* d_off;
    if (in_dev) { d_seq = in->seq; d_off = in->seq_off; }
    else {
        AR_ALLOC(W.d_seq, n_bases + 16);
        AR_ALLOC(W.d_off, (n_reads + 1) * 8);
        AR_TRY(cudaMemcpyAsync(W.d_seq.p, in->seq, n_bases, cudaMemcpyHostToDevice, st));
        AR_TRY(cudaMemcpyAsync(W.d_off.p, in->seq_off, (n_reads + 1) * 8, cudaMemcpyHostToDevice, st));
        d_seq = W.d_seq.as<uint8_t>(); d_off = W.d_off.as<int64_t>();
    }
    AR_ALLOC(W.d_misc, 64);
    AR_ALLOC(W.d_intv_off, n_reads * 8);
    AR_ALLOC(W.d_intv_n, n_reads * 4);
    AR_ALLOC(W.d_intv_cnt2, n_reads * 8);
    AR_ALLOC(W.d_nsa, (n_reads + 1) * 4);
    AR_ALLOC(W.d_sa_off, (n_reads + 1) * 8);
    AR_ALLOC(W.d_cnt, sizeof(Counters));
    AR_ALLOC(W.d_nchain, n_reads * 4);
    AR_ALLOC(W.d_nkseed, n_reads * 4);
    AR_ALLOC(W.d_nregs, (n_reads + 1) * 4);
    AR_ALLOC(W.d_reg_off, (n_reads + 1) * 8);
    AR_TRY(cudaMemsetAsync(W.d_cnt.p, 0, sizeof(Counters), st));
    AR_TRY(cudaMemsetAsync(W.d_nsa.p, 0, (n_reads + 1) * 4, st));
    AR_TRY(cudaMemsetAsync(W.d_nregs.p, 0, (n_reads + 1) * 4, st));

    int nl = 0;  // kernels launched by this batch
    Counters& hc = *W.h_cnt;
    unsigned long long* hs = W.h_scalars;
    // ---- seeding (retry with a larger per-read interval capacity if a read overflows it) ----
    if (idx->fm.seq_len >= (1ull << 40) - 2) { set_error("reference too long for the packed interval lists (2*l_pac must stay below 2^40)"); return fail(ARACHNE_ERR_ARG); }
    const int seed_bpsm = seed_bpsm_opt >= 1 && seed_bpsm_opt <= 4 ? seed_bpsm_opt : 3, seed_blocks = sms * seed_bpsm, list_rows = 2 * L1;
    const S3Plan s3 = seed3_plan(maxlen);
    const size_t seed_T = seed_impl == 3 ? (size_t)sms * s3.ns : (size_t)seed_blocks * 256;  // lanes (k_seed2) / states (k_seed3) that own a list
    AR_ALLOC(W.d_scratch, seed_T * list_rows * sizeof(uint4));
    int intv_cap = 32;
    cudaEventRecord(ev[0], st);
    // the reads 2-bit packed with their masks: what the seeding lanes keep in shared memory (seed_kernel.cuh)
    const ulonglong2* d_qpk = nullptr;
    const int qstride = seed_qstride(maxlen);
    const int seed_direct = use_seed_direct && idx->fm.isa && idx->fm.text;
    {
        AR_ALLOC(W.d_qpk, (size_t)n_reads * qstride * sizeof(ulonglong2));
        k_pack_reads<<<sms * 8, 256, 0, st>>>(d_seq, d_off, (int)n_reads, qstride, W.d_qpk.as<ulonglong2>());
        AR_TRY(cudaGetLastError());
        d_qpk = W.d_qpk.as<ulonglong2>();
        ++nl;
    }
    for (int attempt = 0;; ++attempt) {
        AR_ALLOC(W.d_arena, (size_t)n_reads * intv_cap * sizeof(Intv));
        AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
        AR_TRY(cudaMemsetAsync(W.d_intv_cnt2.p, 0, n_reads * 8, st));
        AR_TRY(cudaMemsetAsync(W.d_nsa.p, 0, (n_reads + 1) * 4, st));
        SeedOut so; so.intv = W.d_arena.as<Intv>(); so.cnt2 = W.d_intv_cnt2.as<int>(); so.cap = intv_cap;
        ++nl;
        // shared-memory head of the interval lists: as many entries as fit next to the resident blocks (16 at 3 blocks/SM)
        int slist_n = 0, nwq = 0;
        const size_t seed_smem = seed_smem_plan(maxlen, seed_bpsm, seed_slist, &slist_n, &nwq);
        if (seed_impl == 3) {
            const int nt = idx->opt_seed3_threads;
            auto kfn = nt == 512 ? k_seed3<512> : nt == 1024 ? k_seed3<1024> : k_seed3<768>;
            AR_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s3.smem));
            kfn<<<sms, nt, s3.smem, st>>>(idx->fm, opt, d_off, (int)n_reads, maxlen, W.d_scratch.as<uint4>(), list_rows, so,
                                          W.d_misc.as<unsigned int>(), W.d_cnt.as<Counters>(), use_seed_kmer, s3.ns, d_qpk, qstride, s3.nwq, seed_direct);
        } else if (seed_bpsm == 4) {
            AR_TRY(cudaFuncSetAttribute(k_seed2<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)seed_smem));
            k_seed2<4><<<seed_blocks, 256, seed_smem, st>>>(idx->fm, opt, d_off, (int)n_reads, maxlen, W.d_scratch.as<uint4>(), list_rows, so,
                                                             W.d_misc.as<unsigned int>(), W.d_cnt.as<Counters>(), use_seed_kmer, slist_n, d_qpk, qstride, nwq, seed_direct);
        } else {
            AR_TRY(cudaFuncSetAttribute(k_seed2<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)seed_smem));
            k_seed2<3><<<seed_blocks, 256, seed_smem, st>>>(idx->fm, opt, d_off, (int)n_reads, maxlen, W.d_scratch.as<uint4>(), list_rows, so,
                                                             W.d_misc.as<unsigned int>(), W.d_cnt.as<Counters>(), use_seed_kmer, slist_n, d_qpk, qstride, nwq, seed_direct);
        }
        AR_TRY(cudaGetLastError());
        AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
        AR_TRY(W.sync(st));
        if (hc.err == 0) break;
        if (hc.err == ARACHNE_ERR_ARG) { set_error("read %d is longer than the stated max_read_len %d", hc.err_read, maxlen); return fail(ARACHNE_ERR_ARG); }
        if (attempt >= 3) { set_error("seeding: more than %d intervals at read %d", intv_cap, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
        intv_cap *= 4;
        AR_TRY(cudaMemsetAsync(W.d_cnt.p, 0, sizeof(Counters), st));
    }
    ++nl;
    k_sort_intv<<<sms * 8, 128, 0, st>>>((int)n_reads, W.d_arena.as<Intv>(), W.d_intv_cnt2.as<int>(), intv_cap, opt.max_occ, W.d_intv_n.as<int>(), W.d_nsa.as<unsigned int>());
    AR_TRY(cudaGetLastError());
    cudaEventRecord(ev[1], st);
    // ---- seeds per read -> offsets ----
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, W.d_nsa.as<unsigned int>(), W.d_sa_off.as<unsigned long long>(), (int)n_reads + 1, st);
    AR_ALLOC(W.d_tmp, tmp_bytes + 16);
    nl += 2;  // cub scan = init + scan kernels
    AR_TRY(cub::DeviceScan::ExclusiveSum(W.d_tmp.p, tmp_bytes, W.d_nsa.as<unsigned int>(), W.d_sa_off.as<unsigned long long>(), (int)n_reads + 1, st));
    AR_TRY(cudaMemcpyAsync(&hs[0], W.d_sa_off.as<unsigned long long>() + n_reads, 8, cudaMemcpyDeviceToHost, st));
    AR_TRY(W.sync(st));
    const unsigned long long total_sa = hs[0];
    const size_t tsa = (size_t)total_sa + 8;
    AR_ALLOC(W.d_rows, tsa * 8);
    AR_ALLOC(W.d_seeds, tsa * sizeof(Seed));
    AR_ALLOC(W.d_kseeds, tsa * sizeof(Seed));
    AR_ALLOC(W.d_chains, tsa * sizeof(Chain));
    AR_ALLOC(W.d_nodes, (tsa / 4 + 4 * (size_t)n_reads + 8) * sizeof(BNode));
    AR_ALLOC(W.d_order, tsa * 4);
    AR_ALLOC(W.d_kept, tsa * 4);
    AR_ALLOC(W.d_srt, tsa * 8);
    AR_ALLOC(W.d_regs, tsa * sizeof(AlnReg));
    ++nl;
    k_expand<<<sms * 4, 256, 0, st>>>(opt, (int)n_reads, W.d_arena.as<Intv>(), W.d_intv_off.as<unsigned long long>(), intv_cap, W.d_intv_n.as<int>(),
                                       W.d_sa_off.as<unsigned long long>(), W.d_rows.as<uint64_t>());
    AR_TRY(cudaGetLastError());
    ++nl;
    AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
    k_sa<<<sms * 8, 256, 0, st>>>(idx->fm, total_sa, W.d_rows.as<uint64_t>(), W.d_misc.as<unsigned long long>(), W.d_cnt.as<Counters>());
    AR_TRY(cudaGetLastError());
    cudaEventRecord(ev[2], st);
    // ---- chaining ----
    ChainBufs cb;
    cb.seeds = W.d_seeds.as<Seed>(); cb.kseeds = W.d_kseeds.as<Seed>(); cb.chains = W.d_chains.as<Chain>(); cb.nodes = W.d_nodes.as<BNode>();
    cb.order = W.d_order.as<int32_t>(); cb.kept = W.d_kept.as<int32_t>(); cb.n_chain = W.d_nchain.as<int>(); cb.n_kseed = W.d_nkseed.as<int>();
    ++nl;
    k_chain<<<sms * 8, 128, 0, st>>>(opt, idx->ref, d_off, (int)n_reads, W.d_arena.as<Intv>(), W.d_intv_off.as<unsigned long long>(),
                                      W.d_intv_n.as<int>(), W.d_sa_off.as<unsigned long long>(), W.d_rows.as<uint64_t>(), cb, W.d_cnt.as<Counters>());
    AR_TRY(cudaGetLastError());
    cudaEventRecord(ev[3], st);
    // ---- extension + dedup: rounds of (driver -> sorted problem queue -> DP), ext_stage.cuh ----
    const int ext_blocks = sms * 8;
    const int win_cap = 2 * maxlen + 1024 + 4 * opt.w;
    const long cg_warps = (long)sms * 4 * EXT_WARPS;
    AR_ALLOC(W.d_win, (size_t)std::max<long>(cg_warps, (long)ext_blocks * EXT_WARPS) * win_cap);
    {
        const int fin_blocks = sms * 4, drv_T = fin_blocks * 128, drv_blocks = (int)std::min<int64_t>((n_reads + 127) / 128, (int64_t)sms * 16);
        AR_ALLOC(W.d_hscr, (size_t)drv_T * 2 * (maxlen + 2) * sizeof(int));
        AR_ALLOC(W.d_rscr, (size_t)drv_T * win_cap);
        AR_ALLOC(W.d_xstate, (size_t)n_reads * sizeof(ExtState));
        AR_ALLOC(W.d_xprob, (size_t)n_reads * sizeof(ExtProb));
        AR_ALLOC(W.d_xres, (size_t)n_reads * sizeof(ExtRes));
        AR_ALLOC(W.d_xrank, (size_t)n_reads * 4);
        AR_ALLOC(W.d_xperm, (size_t)n_reads * sizeof(ExtProb) * 2);  // the sorted problem queues of two consecutive rounds
        AR_ALLOC(W.d_xctl, sizeof(ExtCtl));
        AR_TRY(cudaMemsetAsync(W.d_xctl.p, 0, sizeof(ExtCtl), st));
        ExtBufs XB;
        XB.seq = d_seq; XB.off = d_off; XB.n_reads = (int)n_reads; XB.sa_off = W.d_sa_off.as<unsigned long long>(); XB.b = cb;
        XB.srt = W.d_srt.as<uint64_t>(); XB.regs = W.d_regs.as<AlnReg>(); XB.n_regs = W.d_nregs.as<int>();
        XB.hscr = W.d_hscr.as<int>(); XB.rscr = W.d_rscr.as<uint8_t>(); XB.rcap = win_cap; XB.maxlen = maxlen;
        XB.state = W.d_xstate.as<ExtState>(); XB.prob = W.d_xprob.as<ExtProb>(); XB.res = W.d_xres.as<ExtRes>(); XB.rank = W.d_xrank.as<unsigned>();
        XB.queue_cur = W.d_xperm.as<ExtProb>(); XB.queue_next = XB.queue_cur + n_reads;
        XB.ctl = W.d_xctl.as<ExtCtl>(); XB.cnt = W.d_cnt.as<Counters>();
        ExtDpArgs DA;
        DA.seq = d_seq; DA.tbytes = nullptr; DA.res = XB.res; DA.n_items = &XB.ctl->n_next; DA.head = &XB.ctl->head;
        DA.ncols = std::max(1, maxlen - opt.min_seed_len) + 2;  // a flank is never longer than len - min_seed_len (every seed is at least that long)
        DA.zdrop = opt.zdrop; DA.one = 1; DA.cnt = XB.cnt;
        const size_t smem = (size_t)DA.ncols * XDP_BLOCK * sizeof(uint32_t);
        if (smem > 200 * 1024) { set_error("read length %d does not fit the extension kernel's shared-memory columns", maxlen); return fail(ARACHNE_ERR_ARG); }
        // byte-wide score fields whenever no score can exceed 255 (a = 1: always); the 13-bit codec otherwise
        auto dp_kernel = (long)opt.a * maxlen <= 255 ? k_ext_dp<true, false> : k_ext_dp<true, true>;
        AR_TRY(cudaFuncSetAttribute(dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        AR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_kernel, XDP_BLOCK, smem));
        if (per_sm < 1) per_sm = 1;
        const int dp_blocks = sms * per_sm;
        // A read needs one round per ksw_extend2 call it makes (two for most reads).  Rounds are launched four at a time without
        // looking; the host reads the number of problems the last round emitted and stops at zero.
        int round = 0;
        for (;;) {
            for (int k = 0; k < 4; ++k, ++round) {
                k_ext_drive<<<drv_blocks, 128, 0, st>>>(opt, idx->ref, XB, round);
                k_ext_binbase<<<1, 256, 0, st>>>(XB.ctl);
                k_ext_scatter<<<sms * 4, 256, 0, st>>>(XB, round);
                DA.queue = XB.queue_next;
                dp_kernel<<<dp_blocks, XDP_BLOCK, smem, st>>>(opt, idx->ref, DA);
                k_ext_roll<<<1, 1, 0, st>>>(XB.ctl);
                nl += 5;
                std::swap(XB.queue_cur, XB.queue_next);
            }
            AR_TRY(cudaGetLastError());
            AR_TRY(cudaMemcpyAsync(&hs[5], &XB.ctl->n_next, 4, cudaMemcpyDeviceToHost, st));
            AR_TRY(W.sync(st));
            if ((unsigned)hs[5] == 0) break;
            if (round > 4096) { set_error("extension rounds do not terminate"); return fail(ARACHNE_ERR_OVERFLOW); }
        }
        k_ext_finish<<<fin_blocks, 128, 0, st>>>(opt, idx->ref, XB);
        ++nl;
    }
    cudaEventRecord(ev[4], st);
    int stages = in->stages ? in->stages : (ARACHNE_STAGE_REGIONS | ARACHNE_STAGE_RESCUE | ARACHNE_STAGE_CIGAR);
    if (stages & ARACHNE_STAGE_RECORDS) stages |= ARACHNE_STAGE_PLACE;
    if (stages & ARACHNE_STAGE_PLACE) stages |= ARACHNE_STAGE_CIGAR;  // placement scores the candidates' CIGAR statistics
    // ---- mate rescue ----
    AR_ALLOC(W.d_reg_loc, (n_reads + 1) * 8);
    ++nl;
    k_init_loc<<<sms * 2, 256, 0, st>>>((int)n_reads, W.d_sa_off.as<unsigned long long>(), W.d_reg_loc.as<unsigned long long>());
    AR_TRY(cudaGetLastError());
    if (stages & ARACHNE_STAGE_RESCUE) {
        const int bcap = win_cap;  // one row-maximum entry per target row at most
        AR_ALLOC(W.d_blist, (size_t)ext_blocks * EXT_WARPS * bcap * 8);
        unsigned long long resc_cap = (unsigned long long)n_reads * (unsigned)idx->env_resc_mult + 4096;
        AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
        AR_TRY(W.sync(st));
        if (hc.err != 0) { set_error("internal scratch overflow (code %d) at read %d", hc.err, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
        AR_ALLOC(W.d_nregs_bak, (n_reads + 1) * 4);
        AR_TRY(cudaMemcpyAsync(W.d_nregs_bak.p, W.d_nregs.p, (n_reads + 1) * 4, cudaMemcpyDeviceToDevice, st));
        for (int attempt = 0;; ++attempt) {
            AR_ALLOC(W.d_resc, resc_cap * sizeof(AlnReg));
            AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
            ++nl;
            k_rescue<<<ext_blocks, EXT_WARPS * 32, 0, st>>>(opt, idx->ref, in->score_delta, d_seq, d_off, (int)in->n_pairs,
                                                            W.d_sa_off.as<unsigned long long>(), W.d_regs.as<AlnReg>(), W.d_nregs.as<int>(),
                                                            W.d_reg_loc.as<unsigned long long>(), W.d_resc.as<AlnReg>(), resc_cap,
                                                            W.d_misc.as<unsigned long long>() + 1, W.d_win.as<uint8_t>(), win_cap, W.d_blist.as<uint64_t>(), bcap,
                                                            W.d_misc.as<unsigned int>(), W.d_cnt.as<Counters>());
            AR_TRY(cudaGetLastError());
            AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
            AR_TRY(W.sync(st));
            if (hc.err == 0) break;
            if (hc.err != -100 || attempt >= 4) { set_error("rescue: scratch overflow (code %d) at pair %d", hc.err, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
            // the rescue arena was too small: regions in place are untouched (copy-on-write), so the stage can simply be re-run
            resc_cap *= 8;
            AR_TRY(cudaMemcpyAsync(W.d_nregs.p, W.d_nregs_bak.p, (n_reads + 1) * 4, cudaMemcpyDeviceToDevice, st));
            AR_TRY(cudaMemsetAsync(W.d_cnt.as<char>() + offsetof(Counters, err), 0, 8, st));
            ++nl;
            k_init_loc<<<sms * 2, 256, 0, st>>>((int)n_reads, W.d_sa_off.as<unsigned long long>(), W.d_reg_loc.as<unsigned long long>());
        }
    }
    cudaEventRecord(ev[5], st);
    // ---- compact regions ----
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, W.d_nregs.as<int>(), W.d_reg_off.as<unsigned long long>(), (int)n_reads + 1, st);
    AR_ALLOC(W.d_tmp, tmp_bytes + 16);
    nl += 2;  // cub scan = init + scan kernels
    AR_TRY(cub::DeviceScan::ExclusiveSum(W.d_tmp.p, tmp_bytes, W.d_nregs.as<int>(), W.d_reg_off.as<unsigned long long>(), (int)n_reads + 1, st));
    AR_TRY(cudaMemcpyAsync(&hs[1], W.d_reg_off.as<unsigned long long>() + n_reads, 8, cudaMemcpyDeviceToHost, st));
    AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    AR_TRY(W.sync(st));
    const unsigned long long total_regs = hs[1];
    if (hc.err != 0) { set_error("internal scratch overflow (code %d) at read %d", hc.err, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
    AR_ALLOC(W.d_out, (total_regs + 1) * sizeof(AlnReg));
    AR_ALLOC(W.d_reg_read, (total_regs + 1) * 4);
    ++nl;
    k_gather_regs2<<<sms * 4, 256, 0, st>>>((int)n_reads, W.d_reg_loc.as<unsigned long long>(), W.d_reg_off.as<unsigned long long>(), W.d_regs.as<AlnReg>(),
                                             W.d_resc.as<AlnReg>(), W.d_out.as<AlnReg>(), W.d_reg_read.as<int>());
    AR_TRY(cudaGetLastError());
    static_assert(sizeof(unsigned long long) == sizeof(int64_t), "");
    const bool out_regs = !out_dev && !(in->flags & ARACHNE_OUT_NO_REGIONS);
    if (!out_dev) {
        if (out_regs) AR_PIN(priv->regs, (total_regs + 1) * sizeof(AlnReg));
        AR_TRY(cudaMemcpyAsync(priv->reg_off.p, W.d_reg_off.p, (n_reads + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (total_regs && out_regs) AR_TRY(cudaMemcpyAsync(priv->regs.p, W.d_out.p, total_regs * sizeof(AlnReg), cudaMemcpyDeviceToHost, st));
    }
    cudaEventRecord(ev[6], st);
    // ---- CIGAR + candidate statistics ----
    if ((stages & ARACHNE_STAGE_CIGAR) && total_regs) {
        const int cg_blocks = sms * 4;
        const long nw = (long)cg_blocks * EXT_WARPS;
        const long z_cap = (long)maxlen * (maxlen + 1024);
        const int cig_scr_cap = 2 * maxlen + 64, mm_scr_cap = 2 * maxlen + 8;
        AR_ALLOC(W.d_z, (size_t)nw * z_cap);
        AR_ALLOC(W.d_cig_scr, (size_t)nw * cig_scr_cap * 4);
        AR_ALLOC(W.d_mm_scr, (size_t)nw * mm_scr_cap * 4);
        AR_ALLOC(W.d_alns, (total_regs + 1) * sizeof(AlnRec));
        unsigned long long cig_cap = total_regs * 4 + 4096, mm_cap = total_regs * 12 + 4096;
        // thread-per-region tiers (cigar_kernel.cuh); ARACHNE_CIGAR_TIERS=0 sends every region to the warp tier
        const bool cigar_tiers = idx->env_cigar_tiers;
        const size_t band_smem = (size_t)BAND_RING * CIG_BLOCK * sizeof(int2);
        int band_per_sm = 1;
        if (cigar_tiers) {
            AR_TRY(cudaFuncSetAttribute(k_cigar_band, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)band_smem));
            AR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&band_per_sm, k_cigar_band, CIG_BLOCK, band_smem));
            if (band_per_sm < 1) band_per_sm = 1;
        }
        const int band_blocks = sms * band_per_sm;
        const int zband_rows = maxlen + 2 * BAND_WMAX + 4;  // rlen <= l_query + band
        if (cigar_tiers) {
            AR_ALLOC(W.d_list2, (total_regs + 1) * 4 * 4);  // list, sorted list, keys, sorted keys
            AR_ALLOC(W.d_list3, (total_regs + 1) * 4);
            AR_ALLOC(W.d_zband, (size_t)band_blocks * CIG_BLOCK * zband_rows * BAND_ZCHUNK * 8);
        }
        for (int attempt = 0;; ++attempt) {
            AR_ALLOC(W.d_cig, cig_cap * 4);
            AR_ALLOC(W.d_mm, mm_cap * 4);
            AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
            // misc words: u32[0] queue head of the warp tier, u64[1..2] pool use, u32[8] / u32[9] lengths of the tier-2 / tier-3 lists
            CigArgs CA;
            CA.seq = d_seq; CA.off = d_off; CA.n_regs = (long)total_regs; CA.regs = W.d_out.as<AlnReg>(); CA.reg_read = W.d_reg_read.as<int>();
            CA.alns = W.d_alns.as<AlnRec>(); CA.cig_pool = W.d_cig.as<uint32_t>(); CA.cig_cap = cig_cap; CA.mm_pool = W.d_mm.as<int32_t>(); CA.mm_cap = mm_cap;
            CA.used = W.d_misc.as<unsigned long long>() + 1; CA.list2 = W.d_list2.as<unsigned>(); CA.n2 = W.d_misc.as<unsigned>() + 8;
            CA.list3 = W.d_list3.as<unsigned>(); CA.n3 = W.d_misc.as<unsigned>() + 9; CA.z = W.d_zband.as<unsigned long long>(); CA.z_rows = zband_rows;
            CA.cnt = W.d_cnt.as<Counters>();
            const unsigned* l3 = nullptr; const unsigned* n3 = nullptr;
            if (cigar_tiers) {
                nl += 2;
                unsigned* l2 = W.d_list2.as<unsigned>(); const size_t nr1 = (size_t)total_regs + 1;
                CA.list2 = l2; CA.key2 = l2 + 2 * nr1;
                AR_TRY(cudaMemsetAsync(CA.key2, 0xff, nr1 * 4, st));
                k_cigar_fast<<<sms * 8, CIG_BLOCK, 0, st>>>(opt, idx->ref, CA);
                size_t sbytes = 0;
                cub::DeviceRadixSort::SortPairs(nullptr, sbytes, CA.key2, l2 + 3 * nr1, l2, l2 + nr1, (int)total_regs, 0, 16, st);
                AR_ALLOC(W.d_tmp, sbytes + 16);
                sbytes = W.d_tmp.cap;
                AR_TRY(cub::DeviceRadixSort::SortPairs(W.d_tmp.p, sbytes, CA.key2, l2 + 3 * nr1, l2, l2 + nr1, (int)total_regs, 0, 16, st));
                nl += 3;
                CA.list2 = l2 + nr1;
                k_cigar_band<<<band_blocks, CIG_BLOCK, band_smem, st>>>(opt, idx->ref, CA);
                l3 = CA.list3; n3 = CA.n3;
            }
            ++nl;
            k_cigar<<<cg_blocks, EXT_WARPS * 32, 0, st>>>(opt, idx->ref, d_seq, d_off, (long)total_regs, W.d_out.as<AlnReg>(),
                                                          W.d_reg_read.as<int>(), W.d_alns.as<AlnRec>(), W.d_cig.as<uint32_t>(), cig_cap, W.d_mm.as<int32_t>(), mm_cap,
                                                          W.d_misc.as<unsigned long long>() + 1, W.d_win.as<uint8_t>(), win_cap, W.d_z.as<uint8_t>(), z_cap,
                                                          W.d_cig_scr.as<uint32_t>(), cig_scr_cap, W.d_mm_scr.as<int32_t>(), mm_scr_cap,
                                                          W.d_misc.as<unsigned int>(), W.d_cnt.as<Counters>(), l3, n3);
            AR_TRY(cudaGetLastError());
            AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
            AR_TRY(cudaMemcpyAsync(&hs[2], W.d_misc.as<unsigned long long>() + 1, 16, cudaMemcpyDeviceToHost, st));
            AR_TRY(W.sync(st));
            if (hc.err == 0) break;
            if (hc.err != -101 || attempt >= 4) { set_error("cigar: scratch overflow (code %d) at read %d", hc.err, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
            cig_cap *= 4; mm_cap *= 4;
            AR_TRY(cudaMemsetAsync(W.d_cnt.as<char>() + offsetof(Counters, err), 0, 8, st));
        }
        const unsigned long long used0 = hs[2], used1 = hs[3];
        out->n_cigar = (int64_t)used0; out->n_mm = (int64_t)used1;
        if (!out_dev) {
            AR_PIN(priv->alns, (total_regs + 1) * sizeof(AlnRec)); AR_PIN(priv->cigar, (used0 + 1) * 4); AR_PIN(priv->mm, (used1 + 1) * 4);
            AR_TRY(cudaMemcpyAsync(priv->alns.p, W.d_alns.p, total_regs * sizeof(AlnRec), cudaMemcpyDeviceToHost, st));
            if (used0) AR_TRY(cudaMemcpyAsync(priv->cigar.p, W.d_cig.p, used0 * 4, cudaMemcpyDeviceToHost, st));
            if (used1) AR_TRY(cudaMemcpyAsync(priv->mm.p, W.d_mm.p, used1 * 4, cudaMemcpyDeviceToHost, st));
            out->alns = priv->alns.as<arachne_aln_t>(); out->cigar = priv->cigar.as<uint32_t>(); out->mm = priv->mm.as<int32_t>();
        }
    }
    cudaEventRecord(ev[7], st);
    // ---- placement: tagBestAlignments .. optimizer .. MAPQ ----
    if (stages & ARACHNE_STAGE_PLACE) {
#include "place_host.inl"
    }
    if (!(stages & ARACHNE_STAGE_PLACE)) { cudaEventRecord(ev[8], st); cudaEventRecord(ev[9], st); }
    AR_TRY(W.sync(st));

    out->n_regs = (int64_t)total_regs;
    if (!out_dev) { out->reg_off = priv->reg_off.as<int64_t>(); out->regs = out_regs ? priv->regs.as<arachne_alnreg_t>() : nullptr; }
    out->n_occ_lookups = hc.occ;
    out->n_occ_blocks = hc.occ_blocks;
    out->n_sa_lookups = total_sa;
    out->n_lf_steps = hc.lf_steps;
    out->n_ext_calls = hc.ext_calls;
    out->n_ext_cells = hc.ext_cells;
    out->n_glb_cells = hc.glb_cells;
    out->n_kernel_launches = nl;
    cudaEventElapsedTime(&out->ms_seed, ev[0], ev[1]);
    cudaEventElapsedTime(&out->ms_sa, ev[1], ev[2]);
    cudaEventElapsedTime(&out->ms_chain, ev[2], ev[3]);
    cudaEventElapsedTime(&out->ms_extend, ev[3], ev[4]);
    cudaEventElapsedTime(&out->ms_rescue, ev[4], ev[5]);
    cudaEventElapsedTime(&out->ms_cigar, ev[6], ev[7]);
    cudaEventElapsedTime(&out->ms_place, ev[7], ev[8]);
    cudaEventElapsedTime(&out->ms_records, ev[8], ev[9]);
    cudaEventElapsedTime(&out->ms_total, ev[0], ev[9]);
    *out_ = out;
    return ARACHNE_OK;
#undef AR_TRY
#undef AR_ALLOC
#undef AR_PIN
}

void arachne_batch_free(arachne_batch_out_t* out) {
    if (!out) return;
    auto* priv = reinterpret_cast<BatchPriv*>(out->priv);
    if (priv) {
        arachne_index_t* idx = priv->owner;
        bool pooled = false;
        if (idx) {
            std::lock_guard<std::mutex> g(idx->mu);
            if (idx->priv_pool.size() < 8) { idx->priv_pool.push_back(priv); pooled = true; }
        }
        if (!pooled) { priv->release(); delete priv; }
    }
    delete out;
}
