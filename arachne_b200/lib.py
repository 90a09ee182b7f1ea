"""ctypes view of libarachne_cuda.so (include/arachne_cuda.h).  No fallback: a missing library or device raises."""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

ALNREG = np.dtype([("rb", "<i8"), ("re", "<i8"), ("qb", "<i4"), ("qe", "<i4"), ("rid", "<i4"), ("score", "<i4"),
                   ("truesc", "<i4"), ("sub", "<i4"), ("alt_sc", "<i4"), ("csub", "<i4"), ("sub_n", "<i4"),
                   ("w", "<i4"), ("seedcov", "<i4"), ("secondary", "<i4"), ("secondary_all", "<i4"),
                   ("seedlen0", "<i4"), ("n_comp_is_alt", "<i4"), ("frac_rep", "<f4"), ("hash", "<u8")])
ALN = np.dtype([("pos", "<i8"), ("rid", "<i4"), ("is_rev", "<i4"), ("nm", "<i4"), ("n_cigar", "<i4"), ("cigar_off", "<i8"),
                ("score", "<i4"), ("sub", "<i4"), ("matches", "<i4"), ("mismatches", "<i4"), ("indels", "<i4"),
                ("soft_clipped", "<i4"), ("soft_clipped_length", "<i4"), ("n_mm", "<i4"), ("mm_off", "<i8")])
CAND = np.dtype([("reg", "<i4"), ("read", "<i4"), ("id", "<i4"), ("active", "<i4"), ("is_proper", "<i4"), ("bwa_pick", "<i4"), ("mapq", "<i4"),
                 ("molecule_id", "<i4"), ("active_molecule", "<i4"), ("second_best", "<i4"), ("mate", "<i4"), ("duplicate", "<i4"),
                 ("pos", "<i8"), ("aend", "<i8"), ("lap", "<f8"), ("as_score", "<f8"), ("xs_score", "<f8"), ("molecule_difference", "<f8"),
                 ("sum_move", "<f8"), ("molecule_confidence", "<f8"), ("raw_mapq", "<f8")])  # arachne_cand_t
assert ALNREG.itemsize == 88 and ALN.itemsize == 72 and CAND.itemsize == 120

STAGE_REGIONS, STAGE_RESCUE, STAGE_CIGAR, STAGE_PLACE = 1, 2, 4, 8
IN_DEVICE, OUT_DEVICE, OUT_NO_REGIONS = 1, 2, 4
COUNTERS = ("n_occ_lookups", "n_occ_blocks", "n_sa_lookups", "n_lf_steps", "n_ext_calls", "n_ext_cells", "n_glb_cells", "n_kernel_launches")
STAGE_MS = ("seed", "sa", "chain", "extend", "rescue", "cigar", "place", "total")


class ArachneError(RuntimeError):
    pass


def lib_path() -> str:
    if os.environ.get("ARACHNE_LIB"):  # tuning builds (scripts/): same ABI, different compile-time constants
        return os.environ["ARACHNE_LIB"]
    return os.path.join(HERE, "libarachne_cuda.so")


class MemOpt(C.Structure):  # arachne_mem_opt_t == mem_opt_t (bwa/bwamem.h:27-66)
    _fields_ = [("a", C.c_int), ("b", C.c_int), ("o_del", C.c_int), ("e_del", C.c_int), ("o_ins", C.c_int), ("e_ins", C.c_int),
                ("pen_unpaired", C.c_int), ("pen_clip5", C.c_int), ("pen_clip3", C.c_int), ("w", C.c_int), ("zdrop", C.c_int),
                ("max_mem_intv", C.c_uint64), ("T", C.c_int), ("flag", C.c_int), ("min_seed_len", C.c_int),
                ("min_chain_weight", C.c_int), ("max_chain_extend", C.c_int), ("split_factor", C.c_float), ("split_width", C.c_int),
                ("max_occ", C.c_int), ("max_chain_gap", C.c_int), ("n_threads", C.c_int), ("chunk_size", C.c_int),
                ("mask_level", C.c_float), ("drop_ratio", C.c_float), ("XA_drop_ratio", C.c_float), ("mask_level_redun", C.c_float),
                ("mapQ_coef_len", C.c_float), ("mapQ_coef_fac", C.c_int), ("max_ins", C.c_int), ("max_matesw", C.c_int),
                ("max_XA_hits", C.c_int), ("max_XA_hits_alt", C.c_int), ("mat", C.c_int8 * 25)]


class IndexBuildInfo(C.Structure):  # arachne_index_build_info_t
    _fields_ = [(n, C.c_int64) for n in ("seq_len", "primary", "n_batches", "max_batch", "tied_suffixes", "max_rounds", "device_peak_bytes")] + \
               [(n, C.c_double) for n in ("seconds_read", "seconds_build", "seconds_write")]


class BatchIn(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("seq", C.POINTER(C.c_uint8)), ("seq_off", C.POINTER(C.c_int64)), ("n_barcodes", C.c_int64),
                ("barcode_off", C.POINTER(C.c_int64)), ("score_delta", C.c_int32), ("stages", C.c_int32),
                ("flags", C.c_int32), ("max_read_len", C.c_int32), ("n_bases", C.c_int64),
                ("run_rfa", C.POINTER(C.c_uint8)), ("rand_seed", C.POINTER(C.c_int64)), ("improper_penalty", C.c_double), ("delta", C.c_int32),
                ("place_flags", C.c_int32), ("centro_start", C.POINTER(C.c_int64)), ("centro_end", C.POINTER(C.c_int64)),
                ("attach_bx", C.POINTER(C.c_uint8))]


class BatchOut(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("n_regs", C.c_int64), ("n_cigar", C.c_int64), ("n_mm", C.c_int64),
                ("reg_off", C.POINTER(C.c_int64)), ("regs", C.c_void_p), ("alns", C.c_void_p), ("cigar", C.POINTER(C.c_uint32)),
                ("mm", C.POINTER(C.c_int32)),
                ("n_occ_lookups", C.c_uint64), ("n_occ_blocks", C.c_uint64), ("n_sa_lookups", C.c_uint64), ("n_lf_steps", C.c_uint64), ("n_ext_calls", C.c_uint64),
                ("n_ext_cells", C.c_uint64), ("n_glb_cells", C.c_uint64), ("n_kernel_launches", C.c_uint64),
                ("ms_seed", C.c_float), ("ms_sa", C.c_float), ("ms_chain", C.c_float), ("ms_extend", C.c_float), ("ms_rescue", C.c_float),
                ("ms_cigar", C.c_float), ("ms_place", C.c_float), ("ms_total", C.c_float),
                ("n_cands", C.c_int64), ("cand_off", C.POINTER(C.c_int64)), ("cands", C.c_void_p), ("log_molecule_penalty", C.POINTER(C.c_double)),
                ("n_records", C.c_int64), ("rec_off", C.POINTER(C.c_int64)), ("records", C.c_void_p), ("ms_records", C.c_float),
                ("priv", C.c_void_p)]


RECORD = np.dtype([("read", "<i4"), ("reg", "<i4"), ("primary", "<i4"), ("flag", "<i4"), ("rid", "<i4"), ("mapq", "<i4"), ("mate_rid", "<i4"),
                   ("duplicate", "<i4"), ("pos", "<i8"), ("mate_pos", "<i8"), ("tlen", "<i8"), ("reversed", "<i4"), ("clip_start", "<i4"),
                   ("clip_end", "<i4"), ("hard", "<i4"), ("as", "<i4"), ("xs", "<i4"), ("has_xc", "<i4"), ("xc_reg", "<i4"), ("xm", "<i4"),
                   ("am", "<i4"), ("xt", "<i4"), ("sa_reg", "<i4"), ("sa_rid", "<i4"), ("sa_rev", "<i4"), ("sa_hard", "<i4"), ("sa_mapq", "<i4"),
                   ("sa_nm", "<i4"), ("_pad", "<i4"), ("sa_pos", "<i8"), ("attach_bx", "<i4"), ("has_dm", "<i4"), ("dm", "<f8")])
assert RECORD.itemsize == 152


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class IngestBatch(C.Structure):  # arachne_ingest_batch_t (include/arachne_ingest.h)
    _fields_ = [("n_pairs", C.c_int64), ("n_units", C.c_int64), ("seq", C.POINTER(C.c_uint8)), ("seq_off", C.POINTER(C.c_int64)), ("qual", C.c_void_p),
                ("unit_off", C.POINTER(C.c_int64)), ("unique", C.POINTER(C.c_uint8)), ("run_rfa", C.POINTER(C.c_uint8)), ("valid", C.POINTER(C.c_uint8)),
                ("rand_seed", C.POINTER(C.c_int64)), ("names", C.c_void_p), ("name_off", C.POINTER(C.c_int64)), ("barcodes", C.c_void_p),
                ("barcode_char_off", C.POINTER(C.c_int64)), ("read_groups", C.c_void_p), ("read_group_off", C.POINTER(C.c_int64)), ("priv", C.c_void_p)]


class Ingest:
    """ctypes mirror of include/arachne_ingest.h: barcode-grouped FASTQ batches (reader.go:208-291) ready for align_batch.  Host only."""

    def __init__(self, r1: str, r2: str, n_threads: int = 4, path: str | None = None):
        L = self.lib = C.CDLL(path or lib_path())
        L.arachne_last_error.restype = C.c_char_p
        L.arachne_ingest_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.arachne_ingest_next.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.POINTER(IngestBatch))]
        L.arachne_ingest_batch_free.argtypes = [C.POINTER(IngestBatch)]
        L.arachne_ingest_close.argtypes = [C.c_void_p]
        self.h = C.c_void_p(None)
        rc = L.arachne_ingest_open(r1.encode(), r2.encode(), n_threads, C.byref(self.h))
        if rc:
            raise ArachneError(f"arachne_ingest_open {rc}: {L.arachne_last_error().decode()}")

    def next(self, target_pairs: int, copy: bool = True):
        """-> dict of numpy arrays / string lists for the next batch, or None at the end of the input."""
        pb = C.POINTER(IngestBatch)()
        rc = self.lib.arachne_ingest_next(self.h, target_pairs, C.byref(pb))
        if rc == 1:
            return None
        if rc:
            raise ArachneError(f"arachne_ingest_next {rc}: {self.lib.arachne_last_error().decode()}")
        try:
            b = pb.contents; n, u = b.n_pairs, b.n_units
            off = np.ctypeslib.as_array(b.seq_off, shape=(2 * n + 1,)).copy()
            nb = int(off[-1])
            out = dict(n_pairs=n, n_units=u, off=off, seq=np.ctypeslib.as_array(b.seq, shape=(max(nb, 1),))[:nb].copy(),
                       unit_off=np.ctypeslib.as_array(b.unit_off, shape=(u + 1,)).copy(), unique=np.ctypeslib.as_array(b.unique, shape=(u,)).copy(),
                       run_rfa=np.ctypeslib.as_array(b.run_rfa, shape=(u,)).copy(), valid=np.ctypeslib.as_array(b.valid, shape=(u,)).copy(),
                       rand_seed=np.ctypeslib.as_array(b.rand_seed, shape=(n,)).copy())
            if copy:
                no = np.ctypeslib.as_array(b.name_off, shape=(n + 1,)); bo = np.ctypeslib.as_array(b.barcode_char_off, shape=(u + 1,)); go = np.ctypeslib.as_array(b.read_group_off, shape=(n + 1,))
                names = C.string_at(b.names, int(no[-1])); bcs = C.string_at(b.barcodes, int(bo[-1])); rgs = C.string_at(b.read_groups, int(go[-1])); ql = C.string_at(b.qual, nb)
                out["names"] = [names[no[i]:no[i + 1]].decode() for i in range(n)]
                out["barcodes"] = [bcs[bo[i]:bo[i + 1]].decode() for i in range(u)]
                out["read_groups"] = [rgs[go[i]:go[i + 1]].decode() for i in range(n)]
                out["quals"] = [ql[off[i]:off[i + 1]].decode() for i in range(2 * n)]
            return out
        finally:
            self.lib.arachne_ingest_batch_free(pb)

    def close(self):
        if self.h:
            self.lib.arachne_ingest_close(self.h)
            self.h = C.c_void_p(None)


class AlnRegV(C.Structure):  # mem_alnreg_v / arachne_alnreg_v
    _fields_ = [("n", C.c_size_t), ("m", C.c_size_t), ("a", C.c_void_p)]


class GoAln(C.Structure):  # GO_mem_aln_t / arachne_go_aln_t (bwa_bridge.h:19-29)
    _fields_ = [("pos", C.c_int64), ("rid", C.c_int), ("flag", C.c_int), ("flag2", C.c_uint32), ("n_cigar", C.c_int),
                ("cigar", C.POINTER(C.c_uint32)), ("XA", C.c_char_p), ("score", C.c_int), ("sub", C.c_int), ("alt_sc", C.c_int)]


class CompatStats(C.Structure):
    _fields_ = [("n_calls", C.c_uint64), ("n_batches", C.c_uint64), ("max_batch", C.c_uint64), ("n_live_blocks", C.c_uint64)]


class Compat:
    """ctypes mirror of include/arachne_compat.h: the per-pair / per-candidate calls gobwa makes, served from GPU batches formed
    from the calls in flight (what integration/gobwa/gobwa_cuda.go binds).  Thread-safe: call it from many threads."""

    def __init__(self, A: "ArachneCuda", max_batch_pairs: int = 0, max_wait_us: int = 0):
        self.A, L = A, A.lib
        L.arachne_compat_open.argtypes = [C.c_void_p, C.POINTER(MemOpt), C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.arachne_compat_close.argtypes = [C.c_void_p]
        L.arachne_compat_mem_mate_sw.argtypes = [C.c_void_p, C.POINTER(C.c_uint8), C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_int,
                                                 C.POINTER(AlnRegV), C.POINTER(AlnRegV), C.POINTER(C.c_void_p)]
        L.arachne_compat_reg2aln.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(GoAln)]
        L.arachne_compat_release.argtypes = [C.c_void_p, C.c_void_p]
        L.arachne_compat_fetch_seq.restype = C.c_void_p
        L.arachne_compat_fetch_seq.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        L.arachne_compat_stats.argtypes = [C.c_void_p, C.POINTER(CompatStats)]
        self.h = C.c_void_p(None)
        A._ck(L.arachne_compat_open(A.idx, C.byref(A.opt), max_batch_pairs, max_wait_us, C.byref(self.h)))

    def mem_mate_sw(self, r1: np.ndarray, r2: np.ndarray, score_delta: int = 25):
        """GoBwaMemMateSW for one pair -> (regions of read 1, regions of read 2, first-region addresses, owner).  Regions are
        numpy copies of the 88-byte records; addr1/addr2 are the addresses GoBwaSmithWaterman would pass back."""
        o1, o2, owner = AlnRegV(), AlnRegV(), C.c_void_p(None)
        self.A._ck(self.A.lib.arachne_compat_mem_mate_sw(self.h, _p(r1, C.c_uint8), len(r1), _p(r2, C.c_uint8), len(r2), score_delta,
                                                          C.byref(o1), C.byref(o2), C.byref(owner)))
        def view(v):
            return np.ctypeslib.as_array((C.c_uint8 * (88 * v.n)).from_address(v.a)).copy().view(ALNREG) if v.n else np.zeros(0, ALNREG)
        return view(o1), view(o2), (o1.a or 0), (o2.a or 0), owner

    def reg2aln(self, reg_addr: int):
        g = GoAln()
        self.A._ck(self.A.lib.arachne_compat_reg2aln(self.h, C.c_void_p(reg_addr), C.byref(g)))
        cig = np.array([g.cigar[i] for i in range(g.n_cigar)], dtype=np.uint32)
        return {"pos": g.pos, "rid": g.rid, "is_rev": g.flag2 & 1, "nm": g.flag2 >> 10, "n_cigar": g.n_cigar, "cigar": cig, "score": g.score, "sub": g.sub}

    def release(self, owner):
        self.A.lib.arachne_compat_release(self.h, owner)

    def fetch_seq(self, beg: int, mid: int, end: int):
        b, e, rid = C.c_int64(beg), C.c_int64(end), C.c_int(-1)
        p = self.A.lib.arachne_compat_fetch_seq(self.A.idx, C.byref(b), mid, C.byref(e), C.byref(rid))
        if not p:
            raise ArachneError(self.A.lib.arachne_last_error().decode())
        out = np.ctypeslib.as_array((C.c_uint8 * max(e.value - b.value, 0)).from_address(p)).copy()
        C.CDLL(None).free(C.c_void_p(p))
        return out, b.value, e.value, rid.value

    def stats(self) -> dict:
        s = CompatStats()
        self.A.lib.arachne_compat_stats(self.h, C.byref(s))
        return {n: getattr(s, n) for n, _ in CompatStats._fields_}

    def close(self):
        if self.h:
            self.A.lib.arachne_compat_close(self.h)
            self.h = C.c_void_p(None)


class ArachneCuda:
    """One loaded index on one device + the kernel entry points."""

    def __init__(self, path: str | None = None):
        path = path or lib_path()
        if not os.path.exists(path):
            raise ArachneError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = self.lib = C.CDLL(path)
        L.arachne_last_error.restype = C.c_char_p
        L.arachne_index_load.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.arachne_index_free.argtypes = [C.c_void_p]
        L.arachne_index_build.argtypes = [C.c_char_p, C.c_int, C.c_int64, C.POINTER(IndexBuildInfo)]
        L.arachne_index_l_pac.restype = C.c_int64; L.arachne_index_l_pac.argtypes = [C.c_void_p]
        L.arachne_index_n_seqs.argtypes = [C.c_void_p]
        L.arachne_index_device_bytes.restype = C.c_size_t; L.arachne_index_device_bytes.argtypes = [C.c_void_p]
        L.arachne_index_contig.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
        L.arachne_index_set_stream.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.arachne_index_fetch_seq.restype = C.c_int64
        L.arachne_index_fetch_seq.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)]
        L.arachne_index_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
        L.arachne_opts_default.argtypes = [C.POINTER(MemOpt)]
        L.arachne_align_batch.argtypes = [C.c_void_p, C.POINTER(MemOpt), C.POINTER(BatchIn), C.POINTER(C.POINTER(BatchOut))]
        L.arachne_batch_free.argtypes = [C.POINTER(BatchOut)]
        L.arachne_occ_batch.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.arachne_sa_batch.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.arachne_seed_batch.argtypes = [C.c_void_p, C.POINTER(MemOpt), C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64),
                                         C.POINTER(C.c_int64), C.POINTER(C.c_uint64), C.c_int64]
        L.arachne_ksw_extend_batch.argtypes = [C.c_int, C.POINTER(MemOpt), C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64),
                                               C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                               C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_float), C.c_int]
        L.arachne_ksw_global_batch.argtypes = [C.c_int, C.POINTER(MemOpt), C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64),
                                               C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                               C.POINTER(C.c_int64), C.POINTER(C.c_uint32), C.c_int64]
        L.arachne_dpx_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]
        L.arachne_name_seed.restype = C.c_int64; L.arachne_name_seed.argtypes = [C.c_char_p, C.c_int32]
        self.opt = MemOpt()
        self._ck(L.arachne_opts_default(C.byref(self.opt)))
        self.idx = C.c_void_p(None)
        self.device = 0

    def _ck(self, rc: int):
        if rc != 0:
            raise ArachneError(f"libarachne_cuda error {rc}: {self.lib.arachne_last_error().decode()}")

    def device_count(self) -> int:
        return self.lib.arachne_device_count()

    # ---- index construction (the BWT/SA half of `bwa index`, bwtindex.c:244-300) ----
    def build_index(self, prefix: str, device: int = 0, batch_cap: int = 0) -> dict:
        """<prefix>.pac/.ann -> <prefix>.bwt/.sa on the device, byte-identical to `bwa index`."""
        info = IndexBuildInfo()
        self._ck(self.lib.arachne_index_build(prefix.encode(), device, batch_cap, C.byref(info)))
        return {n: getattr(info, n) for n, _ in IndexBuildInfo._fields_}

    def pack_fasta(self, fasta: str, prefix: str) -> int:
        """FASTA -> <prefix>.pac/.ann/.amb as the reference's pack step writes them (bns_fasta2bntseq, bntseq.c:275-328). Host only."""
        lp = C.c_int64(0)
        self.lib.arachne_index_pack.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int64)]
        self._ck(self.lib.arachne_index_pack(fasta.encode(), prefix.encode(), C.byref(lp)))
        return lp.value

    # ---- index (GoBwaLoadReference, gobwa.go:128-147) ----
    def load_index(self, prefix: str, device: int = 0):
        self.close()
        self.device = device
        self._ck(self.lib.arachne_index_load(prefix.encode(), device, C.byref(self.idx)))
        return self

    def close(self):
        if self.idx:
            self.lib.arachne_index_free(self.idx)
            self.idx = C.c_void_p(None)

    @property
    def l_pac(self) -> int:
        return self.lib.arachne_index_l_pac(self.idx)

    def contigs(self):
        out = []
        for i in range(self.lib.arachne_index_n_seqs(self.idx)):
            name = C.c_char_p(); off = C.c_int64(); ln = C.c_int32()
            self._ck(self.lib.arachne_index_contig(self.idx, i, C.byref(name), C.byref(off), C.byref(ln)))
            out.append((name.value.decode(), off.value, ln.value))
        return out

    # ---- stage entry points ----
    def occ(self, k, l):
        k = np.ascontiguousarray(k, dtype=np.uint64); l = np.ascontiguousarray(l, dtype=np.uint64)
        out = np.zeros((len(k), 8), dtype=np.uint64)
        self._ck(self.lib.arachne_occ_batch(self.idx, len(k), _p(k, C.c_uint64), _p(l, C.c_uint64), _p(out, C.c_uint64)))
        return out

    def sa(self, k):
        k = np.ascontiguousarray(k, dtype=np.uint64); out = np.zeros(len(k), dtype=np.uint64)
        self._ck(self.lib.arachne_sa_batch(self.idx, len(k), _p(k, C.c_uint64), _p(out, C.c_uint64)))
        return out

    def seed(self, seq, off, cap_per_read=64):
        seq = np.ascontiguousarray(seq, dtype=np.uint8); off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        ioff = np.zeros(n + 1, dtype=np.int64); cap = n * cap_per_read + 1024
        intv = np.zeros((cap, 4), dtype=np.uint64)
        self._ck(self.lib.arachne_seed_batch(self.idx, C.byref(self.opt), n, _p(seq, C.c_uint8), _p(off, C.c_int64), _p(ioff, C.c_int64),
                                             _p(intv, C.c_uint64), cap))
        return ioff, intv[:ioff[-1]]

    def ksw_extend(self, q, q_off, t, t_off, h0, w, end_bonus=5, zdrop=100, repeat=1, device=None):
        q = np.ascontiguousarray(q, dtype=np.uint8); t = np.ascontiguousarray(t, dtype=np.uint8)
        q_off = np.ascontiguousarray(q_off, dtype=np.int64); t_off = np.ascontiguousarray(t_off, dtype=np.int64)
        h0 = np.ascontiguousarray(h0, dtype=np.int32); w = np.ascontiguousarray(w, dtype=np.int32)
        n = len(h0); out = np.zeros((n, 6), dtype=np.int32); cells = C.c_uint64(0); ms = C.c_float(0)
        self._ck(self.lib.arachne_ksw_extend_batch(self.device if device is None else device, C.byref(self.opt), n, _p(q, C.c_uint8),
                                                   _p(q_off, C.c_int64), _p(t, C.c_uint8), _p(t_off, C.c_int64), _p(h0, C.c_int32),
                                                   _p(w, C.c_int32), end_bonus, zdrop, _p(out, C.c_int32), C.byref(cells), C.byref(ms), repeat))
        return out, cells.value, ms.value

    def ksw_global(self, q, q_off, t, t_off, w, want_cigar=True, device=None):
        q = np.ascontiguousarray(q, dtype=np.uint8); t = np.ascontiguousarray(t, dtype=np.uint8)
        q_off = np.ascontiguousarray(q_off, dtype=np.int64); t_off = np.ascontiguousarray(t_off, dtype=np.int64)
        w = np.ascontiguousarray(w, dtype=np.int32); n = len(w)
        sc = np.zeros(n, dtype=np.int32); coff = np.zeros(n + 1, dtype=np.int64); cap = int(q_off[-1] + t_off[-1] + 8 * n + 8)
        cig = np.zeros(cap, dtype=np.uint32)
        self._ck(self.lib.arachne_ksw_global_batch(self.device if device is None else device, C.byref(self.opt), n, _p(q, C.c_uint8),
                                                   _p(q_off, C.c_int64), _p(t, C.c_uint8), _p(t_off, C.c_int64), _p(w, C.c_int32),
                                                   _p(sc, C.c_int32), _p(coff, C.c_int64) if want_cigar else None, _p(cig, C.c_uint32), cap))
        return sc, coff, cig[:coff[-1]]

    def fetch_seq(self, beg: int, mid: int, end: int):
        """bns_fetch_seq as GetSeq uses it (gobwa.go:59): (bases, clamped beg, clamped end, rid)."""
        b, e, rid = C.c_int64(beg), C.c_int64(end), C.c_int32(-1)
        out = np.zeros(max(1, abs(end - beg)), np.uint8)
        n = self.lib.arachne_index_fetch_seq(self.idx, C.byref(b), mid, C.byref(e), C.byref(rid), _p(out, C.c_uint8))
        if n < 0:
            self._ck(int(n))
        return out[:n].copy(), b.value, e.value, rid.value

    def set_option(self, name: str, value: int):
        self._ck(self.lib.arachne_index_set_option(self.idx, name.encode(), int(value)))

    def set_stream(self, cuda_stream: int | None):
        """Run batches on the caller's stream (e.g. torch.cuda.current_stream().cuda_stream); None = the handle's own."""
        self._ck(self.lib.arachne_index_set_stream(self.idx, C.c_void_p(cuda_stream or 0), 0 if cuda_stream is None else 1))

    def align_batch_device(self, d_seq_ptr: int, d_off_ptr: int, n_pairs: int, n_bases: int, max_read_len: int, score_delta=25, stages=0,
                           d_bc_off_ptr: int = 0, n_barcodes: int = 0, d_run_rfa_ptr: int = 0, d_rand_seed_ptr: int = 0, improper_penalty=-4.0):
        """Inputs already resident in HBM, results left in the device workspace: returns counts, counters and stage times."""
        bi = BatchIn(); bi.n_pairs = n_pairs
        bi.seq = C.cast(C.c_void_p(d_seq_ptr), C.POINTER(C.c_uint8)); bi.seq_off = C.cast(C.c_void_p(d_off_ptr), C.POINTER(C.c_int64))
        bi.score_delta = score_delta; bi.stages = stages; bi.flags = IN_DEVICE | OUT_DEVICE
        bi.max_read_len = max_read_len; bi.n_bases = n_bases; bi.improper_penalty = improper_penalty; bi.delta = 17
        if d_bc_off_ptr:
            bi.n_barcodes = n_barcodes; bi.barcode_off = C.cast(C.c_void_p(d_bc_off_ptr), C.POINTER(C.c_int64))
        if d_run_rfa_ptr:
            bi.run_rfa = C.cast(C.c_void_p(d_run_rfa_ptr), C.POINTER(C.c_uint8))
        if d_rand_seed_ptr:
            bi.rand_seed = C.cast(C.c_void_p(d_rand_seed_ptr), C.POINTER(C.c_int64))
        po = C.POINTER(BatchOut)()
        self._ck(self.lib.arachne_align_batch(self.idx, C.byref(self.opt), C.byref(bi), C.byref(po)))
        try:
            o = po.contents
            return dict(n_regs=o.n_regs, n_cigar=o.n_cigar, n_mm=o.n_mm, n_cands=o.n_cands, counters={k: getattr(o, k) for k in COUNTERS},
                        ms={k: getattr(o, "ms_" + k) for k in STAGE_MS})
        finally:
            self.lib.arachne_batch_free(po)

    def align_batch_host(self, seq, off, score_delta=25, stages=0, bc_off=None, run_rfa=None, rand_seed=None, improper_penalty=-4.0, flags=0):
        """The C-ABI call exactly as a host makes it (host buffers in, pinned host results out), without converting the
        results to numpy: returns counts, the bytes that crossed PCIe in each direction, counters and stage times."""
        n_pairs = (len(off) - 1) // 2
        bi = BatchIn(); bi.n_pairs = n_pairs; bi.seq = _p(seq, C.c_uint8); bi.seq_off = _p(off, C.c_int64)
        bi.score_delta = score_delta; bi.stages = stages; bi.improper_penalty = improper_penalty; bi.delta = 17; bi.flags = flags
        h2d = seq.nbytes + off.nbytes
        if bc_off is not None:
            bi.n_barcodes = len(bc_off) - 1; bi.barcode_off = _p(bc_off, C.c_int64); h2d += bc_off.nbytes
        if run_rfa is not None:
            bi.run_rfa = _p(run_rfa, C.c_uint8); h2d += run_rfa.nbytes
        if rand_seed is not None:
            bi.rand_seed = _p(rand_seed, C.c_int64); h2d += rand_seed.nbytes
        po = C.POINTER(BatchOut)()
        self._ck(self.lib.arachne_align_batch(self.idx, C.byref(self.opt), C.byref(bi), C.byref(po)))
        try:
            o = po.contents
            d2h = (o.n_reads + 1) * 8 + (o.n_regs * 88 if o.regs else 0) + (o.n_regs * 72 + o.n_cigar * 4 + o.n_mm * 4 if o.alns else 0)
            if o.cands:
                d2h += (o.n_reads + 1) * 8 + o.n_cands * 120 + max(1, bi.n_barcodes) * 8
            mapq_sum = int(np.frombuffer(C.string_at(o.cands, min(o.n_cands, 64) * 120), dtype=CAND)["mapq"].sum()) if o.cands and o.n_cands else -1
            return dict(n_regs=o.n_regs, n_cigar=o.n_cigar, n_mm=o.n_mm, n_cands=o.n_cands, h2d_bytes=h2d, d2h_bytes=d2h,
                        first_score=(int(np.frombuffer(C.string_at(o.regs, 88), dtype=ALNREG)["score"][0]) if o.n_regs and o.regs else
                                     int(np.frombuffer(C.string_at(o.alns, 72), dtype=ALN)["score"][0]) if o.n_regs and o.alns else -1), mapq_probe=mapq_sum,
                        counters={k: getattr(o, k) for k in COUNTERS}, ms={k: getattr(o, "ms_" + k) for k in STAGE_MS})
        finally:
            self.lib.arachne_batch_free(po)

    def dpx_peak(self, device=None) -> float:
        """Measured DPX (VIADDMNMX) issue peak in giga-ops/s."""
        g = C.c_double(0); ms = C.c_float(0)
        self._ck(self.lib.arachne_dpx_peak(self.device if device is None else device, C.byref(g), C.byref(ms)))
        return g.value

    def name_seed(self, name: str) -> int:
        b = name.encode()
        return self.lib.arachne_name_seed(b, len(b))

    # ---- batch hot path ----
    def align_batch(self, seq, off, bc_off=None, score_delta=25, stages=0, run_rfa=None, rand_seed=None, improper_penalty=-4.0, centromeres=None,
                    attach_bx=None, sam=None, flags=0):
        """sam: dict(names=[...n_pairs], quals=[...2*n_pairs] or None, barcodes=[...n_barcodes] or None, read_groups=[...n_pairs] or
        read_group=str (the same for every pair) or None) -> res["sam"] (needs stages & 16)."""
        seq = np.ascontiguousarray(seq, dtype=np.uint8); off = np.ascontiguousarray(off, dtype=np.int64)
        n_pairs = (len(off) - 1) // 2
        bi = BatchIn(); bi.n_pairs = n_pairs; bi.seq = _p(seq, C.c_uint8); bi.seq_off = _p(off, C.c_int64)
        if bc_off is not None:
            bc_off = np.ascontiguousarray(bc_off, dtype=np.int64)
            bi.n_barcodes = len(bc_off) - 1; bi.barcode_off = _p(bc_off, C.c_int64)
        bi.score_delta = score_delta; bi.stages = stages; bi.improper_penalty = improper_penalty; bi.delta = 17; bi.flags = flags
        if run_rfa is not None:
            run_rfa = np.ascontiguousarray(run_rfa, dtype=np.uint8); bi.run_rfa = _p(run_rfa, C.c_uint8)
        if rand_seed is not None:
            rand_seed = np.ascontiguousarray(rand_seed, dtype=np.int64); bi.rand_seed = _p(rand_seed, C.c_int64)
        if centromeres is not None:
            cs = np.ascontiguousarray(centromeres[0], dtype=np.int64); ce = np.ascontiguousarray(centromeres[1], dtype=np.int64)
            bi.centro_start = _p(cs, C.c_int64); bi.centro_end = _p(ce, C.c_int64)
        if attach_bx is not None:
            attach_bx = np.ascontiguousarray(attach_bx, dtype=np.uint8); bi.attach_bx = _p(attach_bx, C.c_uint8)
        po = C.POINTER(BatchOut)()
        self._ck(self.lib.arachne_align_batch(self.idx, C.byref(self.opt), C.byref(bi), C.byref(po)))
        try:
            o = po.contents
            res = dict(reg_off=np.ctypeslib.as_array(o.reg_off, shape=(o.n_reads + 1,)).copy(),
                       regs=(np.frombuffer(C.string_at(o.regs, o.n_regs * 88), dtype=ALNREG).copy() if o.n_regs and o.regs else
                             None if o.n_regs else np.zeros(0, ALNREG)),
                       counters={k: getattr(o, k) for k in COUNTERS}, ms={k: getattr(o, "ms_" + k) for k in STAGE_MS})
            if o.alns and o.n_regs:
                res["alns"] = np.frombuffer(C.string_at(o.alns, o.n_regs * 72), dtype=ALN).copy()
                res["cigar"] = np.ctypeslib.as_array(o.cigar, shape=(max(1, o.n_cigar),))[:o.n_cigar].copy()
                res["mm"] = np.ctypeslib.as_array(o.mm, shape=(max(1, o.n_mm),))[:o.n_mm].copy()
            if o.cand_off:
                res["cand_off"] = np.ctypeslib.as_array(o.cand_off, shape=(o.n_reads + 1,)).copy()
                res["cands"] = np.frombuffer(C.string_at(o.cands, o.n_cands * 120), dtype=CAND).copy() if o.n_cands else np.zeros(0, CAND)
                nb = bi.n_barcodes if bi.n_barcodes > 0 else 1
                res["lmp"] = np.ctypeslib.as_array(o.log_molecule_penalty, shape=(nb,)).copy()
            if o.rec_off:
                res["rec_off"] = np.ctypeslib.as_array(o.rec_off, shape=(o.n_reads + 1,)).copy()
                res["records"] = np.frombuffer(C.string_at(o.records, o.n_records * 152), dtype=RECORD).copy() if o.n_records else np.zeros(0, RECORD)
                if sam is not None:
                    arr = lambda xs: (C.c_char_p * len(xs))(*[x.encode() if isinstance(x, str) else x for x in xs]) if xs is not None else None
                    txt, ln = C.c_void_p(None), C.c_size_t(0)
                    self.lib.arachne_records_sam.argtypes = [C.c_void_p, C.POINTER(BatchIn), C.POINTER(BatchOut), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                             C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
                    nm, ql, bcs = arr(sam["names"]), arr(sam.get("quals")), arr(sam.get("barcodes"))
                    rgl = sam.get("read_groups") or ([sam["read_group"]] * n_pairs if sam.get("read_group") else None)
                    rgs = arr(rgl)
                    self._ck(self.lib.arachne_records_sam(self.idx, C.byref(bi), po, nm, ql, bcs, rgs, C.byref(txt), C.byref(ln)))
                    res["sam"] = C.string_at(txt.value, ln.value).decode()
                    C.CDLL(None).free(txt)
        finally:
            self.lib.arachne_batch_free(po)
        return res


FQ_FORMATS = ("unknown", "standard", "haplotagging", "stlfr", "tellseq")  # ARACHNE_FQ_* (include/arachne_preprocess.h)


class PreprocessOpt(C.Structure):  # arachne_preprocess_opt_t
    _fields_ = [("tag", C.c_char_p), ("standardize", C.c_int32), ("sort", C.c_int32), ("device", C.c_int32), ("compress_level", C.c_int32)]


class PreprocessStats(C.Structure):  # arachne_preprocess_stats_t
    _fields_ = [("n_pairs", C.c_int64), ("n_barcodes", C.c_int64), ("n_no_tag", C.c_int64), ("n_invalid", C.c_int64), ("format", C.c_int32),
                ("ms_read", C.c_float), ("ms_sort", C.c_float), ("ms_write", C.c_float)]


def _pp_lib(path=None):
    L = C.CDLL(path or lib_path())
    L.arachne_last_error.restype = C.c_char_p
    L.arachne_fastq_format.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int)]
    L.arachne_fastq_header_format.argtypes = [C.c_char_p]
    L.arachne_preprocess.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(PreprocessOpt), C.POINTER(PreprocessStats)]
    return L


def fastq_header_format(header: str) -> str:
    """The format patterns of src/preprocess/standardize.go:16-25 on one header line (without '@')."""
    return FQ_FORMATS[_pp_lib().arachne_fastq_header_format(header.encode())]


def fastq_format(r1: str, r2: str) -> str:
    """findFastqFormat (standardize.go:98-125). Host only."""
    L = _pp_lib(); f = C.c_int(0)
    rc = L.arachne_fastq_format(r1.encode(), r2.encode(), C.byref(f))
    if rc:
        raise ArachneError(f"arachne_fastq_format {rc}: {L.arachne_last_error().decode()}")
    return FQ_FORMATS[f.value]


def preprocess(r1: str, r2: str, out_r1: str, out_r2: str, tag: str = "BX", standardize: bool = True, sort: bool = True, device: int = 0,
               compress_level: int = 4) -> dict:
    """`arachne preprocess` (src/preprocess/preprocess.go:43-110 + standardize.go) without samtools: standard headers, pairs in barcode
    order (stable device radix sort).  sort=True needs a CUDA device."""
    L = _pp_lib()
    o = PreprocessOpt(tag.encode(), int(standardize), int(sort), device, compress_level); st = PreprocessStats()
    rc = L.arachne_preprocess(r1.encode(), r2.encode(), out_r1.encode(), out_r2.encode(), C.byref(o), C.byref(st))
    if rc:
        raise ArachneError(f"arachne_preprocess {rc}: {L.arachne_last_error().decode()}")
    d = {k: getattr(st, k) for k, _ in PreprocessStats._fields_}
    d["format"] = FQ_FORMATS[st.format]
    return d
