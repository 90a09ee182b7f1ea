// Host orchestration of ARACHNE_STAGE_PLACE (included inside arachne_align_batch, after the CIGAR stage).
// Uses the AR_TRY / AR_ALLOC / AR_PIN macros and the locals of batch.inl: W, st, sms, idx, in, out, priv, nl, hc, hs,
// n_reads, total_regs, d_off, in_dev, out_dev.
{
    const int64_t n_pairs = in->n_pairs;
    int n_bc = (int)in->n_barcodes;
    std::vector<int64_t> h_bc;                       // host view of barcode_off when it is needed here
    std::vector<uint8_t> h_rfa;
    const int64_t* d_bc = nullptr; const uint8_t* d_rfa = nullptr; const int64_t* d_seed = nullptr;
    if (n_bc <= 0 || !in->barcode_off) {             // no grouping given: the whole batch is one work unit
        n_bc = 1; h_bc = {0, n_pairs};
        AR_ALLOC(W.d_bc_off, 16);
        AR_TRY(cudaMemcpyAsync(W.d_bc_off.p, h_bc.data(), 16, cudaMemcpyHostToDevice, st));
        d_bc = W.d_bc_off.as<int64_t>();
    } else if (in_dev) {
        d_bc = in->barcode_off;
    } else {
        AR_ALLOC(W.d_bc_off, (size_t)(n_bc + 1) * 8);
        AR_TRY(cudaMemcpyAsync(W.d_bc_off.p, in->barcode_off, (size_t)(n_bc + 1) * 8, cudaMemcpyHostToDevice, st));
        d_bc = W.d_bc_off.as<int64_t>();
    }
    h_rfa.assign((size_t)n_bc, 0);
    if (in->run_rfa && in->n_barcodes > 0) {
        if (in_dev) { AR_TRY(cudaMemcpyAsync(h_rfa.data(), in->run_rfa, (size_t)n_bc, cudaMemcpyDeviceToHost, st)); AR_TRY(W.sync(st)); d_rfa = in->run_rfa; }
        else {
            memcpy(h_rfa.data(), in->run_rfa, (size_t)n_bc);
            AR_ALLOC(W.d_run_rfa, (size_t)n_bc + 8);
            AR_TRY(cudaMemcpyAsync(W.d_run_rfa.p, in->run_rfa, (size_t)n_bc, cudaMemcpyHostToDevice, st));
            d_rfa = W.d_run_rfa.as<uint8_t>();
        }
    }
    if (in->rand_seed) {
        if (in_dev) d_seed = in->rand_seed;
        else {
            AR_ALLOC(W.d_rand_seed, (size_t)n_pairs * 8 + 8);
            AR_TRY(cudaMemcpyAsync(W.d_rand_seed.p, in->rand_seed, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, st));
            d_seed = W.d_rand_seed.as<int64_t>();
        }
    }
    const int n_ctg = (int)idx->contigs.size();
    const int64_t *d_cs = nullptr, *d_ce = nullptr;
    if (in->centro_start && in->centro_end) {
        AR_ALLOC(W.d_centro, (size_t)n_ctg * 16 + 16);
        AR_TRY(cudaMemcpyAsync(W.d_centro.p, in->centro_start, (size_t)n_ctg * 8, cudaMemcpyHostToDevice, st));
        AR_TRY(cudaMemcpyAsync(W.d_centro.as<int64_t>() + n_ctg, in->centro_end, (size_t)n_ctg * 8, cudaMemcpyHostToDevice, st));
        d_cs = W.d_centro.as<int64_t>(); d_ce = d_cs + n_ctg;
    }
    // ---- candidates: count, scan, build ----
    AR_ALLOC(W.d_pl_cnt, (size_t)(n_reads + 1) * 8);
    AR_ALLOC(W.d_pl_scan, (size_t)(n_reads + 2) * 8);
    AR_ALLOC(W.d_cand_off, (size_t)(n_reads + 1) * 8);
    AR_ALLOC(W.d_lmp, (size_t)n_bc * 8 + 8);
    AR_TRY(cudaMemsetAsync(W.d_lmp.p, 0, (size_t)n_bc * 8 + 8, st));
    const int delta = in->delta > 0 ? in->delta : 17;
    const double improper = (in->place_flags & ARACHNE_PLACE_DEFAULTS) ? -4.0 : in->improper_penalty;
    ++nl;
    k_place_count<<<sms * 4, 256, 0, st>>>((int)n_reads, W.d_reg_off.as<unsigned long long>(), W.d_out.as<AlnReg>(), delta, W.d_pl_cnt.as<unsigned long long>());
    AR_TRY(cudaGetLastError());
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, W.d_pl_cnt.as<unsigned long long>(), W.d_pl_scan.as<unsigned long long>(), (int)n_reads + 1, st);
    AR_ALLOC(W.d_tmp, tb + 16);
    nl += 2;
    AR_TRY(cub::DeviceScan::ExclusiveSum(W.d_tmp.p, tb, W.d_pl_cnt.as<unsigned long long>(), W.d_pl_scan.as<unsigned long long>(), (int)n_reads + 1, st));
    AR_TRY(cudaMemcpyAsync(&hs[4], W.d_pl_scan.as<unsigned long long>() + n_reads, 8, cudaMemcpyDeviceToHost, st));
    AR_TRY(W.sync(st));
    const long n_cands = (long)(hs[4] & 0xffffffffull);
    AR_ALLOC(W.d_cands, (size_t)(n_cands + 1) * sizeof(Cand));
    AR_ALLOC(W.d_hot, (size_t)(n_cands + 1) * sizeof(Hot));
    AR_ALLOC(W.d_aux, (size_t)(n_cands + 1) * sizeof(Aux));
    PlaceArgs PA;
    PA.regs = W.d_out.as<AlnReg>(); PA.alns = W.d_alns.as<AlnRec>(); PA.reg_off = W.d_reg_off.as<unsigned long long>(); PA.seq_off = d_off;
    PA.contigs = idx->ref.contigs; PA.l_pac = idx->l_pac; PA.bc_off = d_bc; PA.n_bc = n_bc; PA.n_reads = (int)n_reads;
    PA.run_rfa = d_rfa; PA.rand_seed = d_seed; PA.centro_start = d_cs; PA.centro_end = d_ce; PA.improper = improper; PA.delta = delta;
    PA.scan = W.d_pl_scan.as<unsigned long long>(); PA.cand_off = W.d_cand_off.as<unsigned long long>();
    PA.cands = W.d_cands.as<Cand>(); PA.hot = W.d_hot.as<Hot>(); PA.aux = W.d_aux.as<Aux>(); PA.lmp = W.d_lmp.as<double>(); PA.cnt = W.d_cnt.as<Counters>();
    ++nl;
    k_place_build<<<sms * 4, 256, 0, st>>>(PA);
    AR_TRY(cudaGetLastError());
    // ---- tagBestAlignments ----
    unsigned slow_cap = 256;
    for (int attempt = 0;; ++attempt) {
        AR_ALLOC(W.d_slow, (size_t)slow_cap * GORAND_LEN * 8);
        AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
        ++nl;
        k_place_tag<<<sms * 4, 128, 0, st>>>(PA, (int)n_pairs, W.d_slow.as<uint64_t>(), slow_cap, W.d_misc.as<unsigned>());
        AR_TRY(cudaGetLastError());
        AR_TRY(cudaMemcpyAsync(&hc, W.d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
        AR_TRY(W.sync(st));
        if (hc.err == 0) break;
        if (hc.err != -102 || attempt >= 3) { set_error("placement: scratch overflow (code %d) at read %d", hc.err, hc.err_read); return fail(ARACHNE_ERR_OVERFLOW); }
        slow_cap *= 16;  // pairs with very many candidate combinations: the kernel is idempotent, re-run it with a larger pool
        AR_TRY(cudaMemsetAsync(W.d_cnt.as<char>() + offsetof(Counters, err), 0, 8, st));
    }
    // ---- molecules + optimizer for the barcodes that run RFA ----
    std::vector<int> slot_of_bc((size_t)n_bc, -1), bc_of_slot;
    for (int b = 0; b < n_bc; ++b) if (h_rfa[b]) { slot_of_bc[b] = (int)bc_of_slot.size(); bc_of_slot.push_back(b); }
    const int n_slots = (int)bc_of_slot.size();
    if (n_slots > 0 && n_cands > 0) {
        const int n_ctg1 = n_ctg + 1;
        if ((size_t)n_slots * n_ctg1 > (size_t)1 << 31) { set_error("placement: %d RFA barcodes x %d contigs exceeds the first-appearance table; split the batch", n_slots, n_ctg); return fail(ARACHNE_ERR_ARG); }
        AR_ALLOC(W.d_slot_of_bc, (size_t)n_bc * 4 + 8);
        AR_ALLOC(W.d_bc_of_slot, (size_t)n_slots * 4 + 8);
        AR_TRY(cudaMemcpyAsync(W.d_slot_of_bc.p, slot_of_bc.data(), (size_t)n_bc * 4, cudaMemcpyHostToDevice, st));
        AR_TRY(cudaMemcpyAsync(W.d_bc_of_slot.p, bc_of_slot.data(), (size_t)n_slots * 4, cudaMemcpyHostToDevice, st));
        AR_ALLOC(W.d_first, (size_t)n_slots * n_ctg1 * 4);
        AR_TRY(cudaMemsetAsync(W.d_first.p, 0xff, (size_t)n_slots * n_ctg1 * 4, st));
        AR_ALLOC(W.d_keys_a, (size_t)(n_cands + 1) * 8); AR_ALLOC(W.d_keys_b, (size_t)(n_cands + 1) * 8);
        AR_ALLOC(W.d_vals_a, (size_t)(n_cands + 1) * 4); AR_ALLOC(W.d_vals_b, (size_t)(n_cands + 1) * 4);
        AR_ALLOC(W.d_seg_off, (size_t)(n_slots + 2) * 8 * 2);
        nl += 3;
        k_place_first<<<sms * 4, 256, 0, st>>>(PA, n_cands, W.d_slot_of_bc.as<int>(), n_ctg1, W.d_first.as<unsigned>());
        k_place_keys1<<<sms * 4, 256, 0, st>>>(PA, n_cands, W.d_keys_a.as<unsigned long long>(), W.d_vals_a.as<unsigned>());
        AR_TRY(cudaGetLastError());
        size_t sb1 = 0, sb2 = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, sb1, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_a.as<unsigned>(),
                                        W.d_vals_b.as<unsigned>(), (int)n_cands, 0, 40, st);
        cub::DeviceRadixSort::SortPairs(nullptr, sb2, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_b.as<unsigned>(),
                                        W.d_vals_a.as<unsigned>(), (int)n_cands, 0, 64, st);
        AR_ALLOC(W.d_tmp, std::max(sb1, sb2) + 16);
        sb1 = sb2 = W.d_tmp.cap;
        AR_TRY(cub::DeviceRadixSort::SortPairs(W.d_tmp.p, sb1, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_a.as<unsigned>(),
                                               W.d_vals_b.as<unsigned>(), (int)n_cands, 0, 40, st));
        k_place_keys2<<<sms * 4, 256, 0, st>>>(PA, n_cands, W.d_vals_b.as<unsigned>(), W.d_slot_of_bc.as<int>(), n_ctg1, W.d_first.as<unsigned>(), W.d_keys_a.as<unsigned long long>());
        AR_TRY(cudaGetLastError());
        AR_TRY(cub::DeviceRadixSort::SortPairs(W.d_tmp.p, sb2, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_b.as<unsigned>(),
                                               W.d_vals_a.as<unsigned>(), (int)n_cands, 0, 64, st));
        nl += 8;  // two radix sorts (histogram + onesweep passes, counted as 4 launches each)
        // ranges of the sort order per slot
        unsigned long long* d_segc = W.d_seg_off.as<unsigned long long>() + (n_slots + 2);
        ++nl;
        k_place_seg<<<(n_slots + 256) / 256, 256, 0, st>>>(PA, n_slots, W.d_bc_of_slot.as<int>(), d_segc);
        AR_TRY(cudaGetLastError());
        cub::DeviceScan::ExclusiveSum(nullptr, tb, d_segc, W.d_seg_off.as<unsigned long long>(), n_slots + 1, st);
        AR_ALLOC(W.d_tmp, tb + 16);
        nl += 2;
        AR_TRY(cub::DeviceScan::ExclusiveSum(W.d_tmp.p, tb, d_segc, W.d_seg_off.as<unsigned long long>(), n_slots + 1, st));
        AR_ALLOC(W.d_mol, (size_t)(n_cands + 1) * (10 * 4 + 8));
        RfaArgs RA;
        RA.bc_of_slot = W.d_bc_of_slot.as<int>(); RA.n_slots = n_slots; RA.ord = W.d_vals_a.as<unsigned>(); RA.seg_off = W.d_seg_off.as<unsigned long long>();
        int* mi = W.d_mol.as<int>(); const size_t nc1 = (size_t)n_cands + 1;
        RA.k_p = reinterpret_cast<double*>(mi); mi += 2 * nc1;
        RA.m_beg = mi; RA.m_end = mi + nc1; RA.m_nact = mi + 2 * nc1; RA.m_nreads = mi + 3 * nc1; RA.m_new = mi + 4 * nc1;
        RA.k_beg = mi + 5 * nc1; RA.k_end = mi + 6 * nc1; RA.k_nact = mi + 7 * nc1; RA.k_nreads = mi + 8 * nc1; RA.k_isact = mi + 9 * nc1;
        AR_TRY(cudaMemsetAsync(W.d_misc.p, 0, 64, st));
        RA.next = W.d_misc.as<unsigned>();
        ++nl;
        k_place_rfa<<<std::min(n_slots, sms * (1024 / PLACE_T)), PLACE_T, 0, st>>>(PA, RA);
        AR_TRY(cudaGetLastError());
    }
    // ---- MAPQ ----
    ++nl;
    k_place_mapq<<<sms * 4, 128, 0, st>>>(PA);
    AR_TRY(cudaGetLastError());
    cudaEventRecord(ev[8], st);
    // ---- record assembly: markDuplicates, CheckSplitReads, AppendBam's fields (records.cuh) ----
    if (n_cands > 0) {
        const bool want_records = (stages & ARACHNE_STAGE_RECORDS) != 0;
        const uint8_t* d_attach = nullptr;
        if (in->attach_bx && in->n_barcodes > 0) {
            if (in_dev) d_attach = in->attach_bx;
            else {
                AR_ALLOC(W.d_attach, (size_t)n_bc + 8);
                AR_TRY(cudaMemcpyAsync(W.d_attach.p, in->attach_bx, (size_t)n_bc, cudaMemcpyHostToDevice, st));
                d_attach = W.d_attach.as<uint8_t>();
            }
        }
        AR_ALLOC(W.d_rec_i, (size_t)(n_reads + 1) * 4 * 4);                  // act, split_reg, split_mapq, split_proper
        AR_ALLOC(W.d_rec_d, (size_t)(n_reads + 1) * 8 * 2);                  // split_as, split_xs
        AR_ALLOC(W.d_rec_cnt, (size_t)(n_reads + 2) * 8 * 2);                // counts, offsets
        AR_ALLOC(W.d_keys_a, (size_t)(n_reads + 1) * 8); AR_ALLOC(W.d_keys_b, (size_t)(n_reads + 1) * 8);
        AR_ALLOC(W.d_vals_a, (size_t)(n_reads + 1) * 4); AR_ALLOC(W.d_vals_b, (size_t)(n_reads + 1) * 4);
        RecArgs RA2;
        RA2.P = PA; RA2.attach_bx = d_attach;
        int* ri = W.d_rec_i.as<int>(); const size_t nr1 = (size_t)n_reads + 1;
        RA2.act = ri; RA2.split_reg = ri + nr1; RA2.split_mapq = ri + 2 * nr1; RA2.split_proper = ri + 3 * nr1;
        RA2.split_as = W.d_rec_d.as<double>(); RA2.split_xs = RA2.split_as + nr1;
        RA2.rec_cnt = W.d_rec_cnt.as<unsigned long long>(); RA2.rec_off = RA2.rec_cnt + (n_reads + 2); RA2.recs = nullptr;
        k_rec_active<<<sms * 4, 256, 0, st>>>(PA, RA2.act);
        k_rec_dup_keys<<<sms * 4, 256, 0, st>>>(PA, RA2.act, W.d_keys_a.as<unsigned long long>(), W.d_vals_a.as<unsigned>());
        AR_TRY(cudaGetLastError());
        size_t sbr = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, sbr, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_a.as<unsigned>(),
                                        W.d_vals_b.as<unsigned>(), (int)n_reads, 0, 64, st);
        AR_ALLOC(W.d_tmp, sbr + 16);
        sbr = W.d_tmp.cap;
        AR_TRY(cub::DeviceRadixSort::SortPairs(W.d_tmp.p, sbr, W.d_keys_a.as<unsigned long long>(), W.d_keys_b.as<unsigned long long>(), W.d_vals_a.as<unsigned>(),
                                               W.d_vals_b.as<unsigned>(), (int)n_reads, 0, 64, st));
        k_rec_dup_mark<<<sms * 4, 256, 0, st>>>(PA, RA2.act, W.d_keys_b.as<unsigned long long>(), W.d_vals_b.as<unsigned>());  // markDuplicates closes DoRFAForOneBarcode (aligner.go:473,497)
        AR_TRY(cudaGetLastError());
        nl += 4 + 3;
      if (want_records) {
        k_rec_split<<<sms * 4, 256, 0, st>>>(RA2);
        AR_TRY(cudaGetLastError());
        cub::DeviceScan::ExclusiveSum(nullptr, tb, RA2.rec_cnt, const_cast<unsigned long long*>(RA2.rec_off), (int)n_reads + 1, st);
        AR_ALLOC(W.d_tmp, tb + 16);
        AR_TRY(cub::DeviceScan::ExclusiveSum(W.d_tmp.p, tb, RA2.rec_cnt, const_cast<unsigned long long*>(RA2.rec_off), (int)n_reads + 1, st));
        nl += 1 + 2;
        // at most two records per read: no size read-back before the last kernel
        AR_ALLOC(W.d_recs, (size_t)(2 * n_reads + 1) * sizeof(Record));
        RA2.recs = W.d_recs.as<Record>();
        k_rec_emit<<<sms * 4, 128, 0, st>>>(RA2, W.d_cig.as<uint32_t>());
        AR_TRY(cudaGetLastError());
        ++nl;
        AR_TRY(cudaMemcpyAsync(&hs[6], RA2.rec_off + n_reads, 8, cudaMemcpyDeviceToHost, st));
        AR_TRY(W.sync(st));
        out->n_records = (int64_t)hs[6];
        if (!out_dev) {
            AR_PIN(priv->rec_off, (size_t)(n_reads + 1) * 8); AR_PIN(priv->recs, (size_t)(out->n_records + 1) * sizeof(Record));
            AR_TRY(cudaMemcpyAsync(priv->rec_off.p, RA2.rec_off, (size_t)(n_reads + 1) * 8, cudaMemcpyDeviceToHost, st));
            AR_TRY(cudaMemcpyAsync(priv->recs.p, W.d_recs.p, (size_t)out->n_records * sizeof(Record), cudaMemcpyDeviceToHost, st));
            out->rec_off = priv->rec_off.as<int64_t>(); out->records = priv->recs.as<arachne_record_t>();
        }
      }
    }
    cudaEventRecord(ev[9], st);
    out->n_cands = n_cands;
    if (!out_dev) {
        AR_PIN(priv->cand_off, (size_t)(n_reads + 1) * 8); AR_PIN(priv->cands, (size_t)(n_cands + 1) * sizeof(Cand)); AR_PIN(priv->lmp, (size_t)n_bc * 8 + 8);
        AR_TRY(cudaMemcpyAsync(priv->cand_off.p, W.d_cand_off.p, (size_t)(n_reads + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (n_cands) AR_TRY(cudaMemcpyAsync(priv->cands.p, W.d_cands.p, (size_t)n_cands * sizeof(Cand), cudaMemcpyDeviceToHost, st));
        AR_TRY(cudaMemcpyAsync(priv->lmp.p, W.d_lmp.p, (size_t)n_bc * 8, cudaMemcpyDeviceToHost, st));
        out->cand_off = priv->cand_off.as<int64_t>(); out->cands = priv->cands.as<arachne_cand_t>(); out->log_molecule_penalty = priv->lmp.as<double>();
    }
}
