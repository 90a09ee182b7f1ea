// arachne_records_sam: SAM text from the record fields of ARACHNE_STAGE_RECORDS (host only; included inside pipeline.cu's extern "C").
// Field and tag order follow AppendBam (src/aligner/bamwriter.go:283-568); the tag bodies are the strings it builds there.

namespace {
struct SamBuf {
    std::string s;
    void num(long long v) { char b[24]; int n = snprintf(b, sizeof(b), "%lld", v); s.append(b, (size_t)n); }
    void tab() { s.push_back('\t'); }
};
// CIGAR string of a region as the record shows it: ops MIDSH = 01234; S -> H when `hard`; op order reversed when `rev_order`
void sam_cigar(SamBuf& o, const arachne_batch_out_t* out, int reg, bool hard, bool rev_order) {
    const arachne_aln_t& a = out->alns[reg];
    for (int k = 0; k < a.n_cigar; ++k) {
        const uint32_t w = out->cigar[a.cigar_off + (rev_order ? a.n_cigar - 1 - k : k)];
        int op = (int)(w & 0xf);
        if (op == 3 && hard) op = 4;
        o.num((long long)(w >> 4));
        o.s.push_back("MIDSH"[op > 4 ? 0 : op]);
    }
}
// "ref,read,1;" for every mismatch locus of a region (XC / AC, bamwriter.go:420-441)
void sam_mm_list(SamBuf& o, const arachne_batch_out_t* out, int reg) {
    if (reg < 0) return;
    const arachne_aln_t& a = out->alns[reg];
    for (int i = 0; i < a.n_mm; ++i) { o.num(out->mm[a.mm_off + i]); o.s.push_back(','); o.num(out->mm[a.mm_off + a.n_mm + i]); o.s.append(",1;"); }
}
}  // namespace

int arachne_records_sam(const arachne_index_t* idx, const arachne_batch_in_t* in, const arachne_batch_out_t* out, const char* const* names,
                        const char* const* quals, const char* const* barcodes, const char* const* read_groups, char** sam, size_t* sam_len) {
    if (!idx || !in || !out || !names || !sam || !sam_len) { set_error("arachne_records_sam: null argument"); return ARACHNE_ERR_ARG; }
    if (!out->records || !out->rec_off || (out->n_regs && (!out->alns || !out->cigar))) { set_error("arachne_records_sam: the batch has no host-side records (run ARACHNE_STAGE_RECORDS without ARACHNE_OUT_DEVICE)"); return ARACHNE_ERR_ARG; }
    if (in->flags & ARACHNE_IN_DEVICE) { set_error("arachne_records_sam: needs the host copy of the reads"); return ARACHNE_ERR_ARG; }
    static const char dec[5] = {'A', 'C', 'G', 'T', 'N'}, cmp[5] = {'T', 'G', 'C', 'A', 'N'};
    SamBuf o;
    o.s.reserve((size_t)out->n_records * 420);
    int64_t bc = 0;
    for (int64_t i = 0; i < out->n_records; ++i) {
        const arachne_record_t& R = out->records[i];
        const int64_t pair = R.read >> 1;
        while (in->n_barcodes > 0 && bc + 1 < in->n_barcodes && in->barcode_off[bc + 1] <= pair) ++bc;
        const char* nm = names[pair];
        size_t nl = strlen(nm);
        while (nl && nm[nl - 1] == '\n') --nl;  // strings.TrimRight(name, "\n")
        o.s.append(nm, nl); o.tab();
        o.num(R.flag); o.tab();
        if (R.rid >= 0) o.s += idx->names[R.rid]; else o.s.push_back('*');
        o.tab(); o.num(R.pos + 1); o.tab(); o.num(R.mapq); o.tab();
        if (R.reg >= 0 && out->alns[R.reg].n_cigar > 0) sam_cigar(o, out, R.reg, R.hard != 0, false); else o.s.push_back('*');
        o.tab();
        if (R.mate_rid < 0) o.s.push_back('*'); else if (R.mate_rid == R.rid) o.s.push_back('='); else o.s += idx->names[R.mate_rid];
        o.tab(); o.num(R.mate_pos + 1); o.tab(); o.num(R.tlen); o.tab();
        const uint8_t* q = in->seq + in->seq_off[R.read];
        const int len = (int)(in->seq_off[R.read + 1] - in->seq_off[R.read]);
        if (R.clip_end <= R.clip_start) o.s.push_back('*');
        for (int k = R.clip_start; k < R.clip_end; ++k) { const int c = R.reversed ? q[len - 1 - k] : q[k]; o.s.push_back(R.reversed ? cmp[c > 4 ? 4 : c] : dec[c > 4 ? 4 : c]); }
        o.tab();
        const char* ql = quals ? quals[R.read] : nullptr;
        if (!ql || R.clip_end <= R.clip_start) o.s.push_back('*');
        else for (int k = R.clip_start; k < R.clip_end; ++k) o.s.push_back(R.reversed ? ql[len - 1 - k] : ql[k]);
        if (read_groups && read_groups[pair] && *read_groups[pair]) { o.s.append("\tRG:Z:"); o.s += read_groups[pair]; }
        o.s.append("\tXS:i:"); o.num(R.xs);
        o.s.append("\tXC:Z:"); if (R.has_xc) sam_mm_list(o, out, R.xc_reg);
        o.s.append("\tAC:Z:"); sam_mm_list(o, out, R.reg);
        o.s.append("\tAS:i:"); o.num(R.as);
        o.s.append("\tXM:Z:"); o.num(R.xm);
        o.s.append("\tAM:Z:"); o.num(R.am);
        o.s.append("\tXT:i:"); o.num(R.xt);
        if (R.sa_reg >= 0) {
            o.s.append("\tSA:Z:"); o.s += idx->names[R.sa_rid]; o.s.push_back(','); o.num(R.sa_pos); o.s.push_back(','); o.s.push_back(R.sa_rev ? '-' : '+'); o.s.push_back(',');
            sam_cigar(o, out, R.sa_reg, R.sa_hard != 0, R.sa_rev != 0);
            o.s.push_back(','); o.num(R.sa_mapq); o.s.push_back(','); o.num(R.sa_nm); o.s.push_back(';');
        }
        const char* bcs = barcodes && in->n_barcodes > 0 ? barcodes[bc] : nullptr;
        if (R.attach_bx && bcs && strchr(bcs, '-')) {  // len(strings.Split(barcode, "-")) > 1 && attach_bx
            o.s.append("\tBX:Z:"); o.s += bcs; o.s.append("\tVX:i:1");
            if (R.has_dm) { char b[64]; snprintf(b, sizeof(b), "%.6f", R.dm); o.s.append("\tDM:Z:"); o.s += b; }
        }
        o.s.push_back('\n');
    }
    char* buf = static_cast<char*>(malloc(o.s.size() + 1));
    if (!buf) { set_error("arachne_records_sam: out of host memory"); return ARACHNE_ERR_CUDA; }
    memcpy(buf, o.s.data(), o.s.size()); buf[o.s.size()] = 0;
    *sam = buf; *sam_len = o.s.size();
    return ARACHNE_OK;
}
