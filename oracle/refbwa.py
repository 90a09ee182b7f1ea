"""TEST INFRASTRUCTURE ONLY: ctypes view of oracle/_ref/libbwa_ref.so (the reference's own BWA C sources
compiled by oracle/Makefile + the gobwa.go-mirroring driver in oracle/ref_shim.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libbwa_ref.so")

# byte-for-byte the layout of mem_alnreg_t (bwa/bwamem.h:86-104); 88 bytes
ALNREG = np.dtype([("rb", "<i8"), ("re", "<i8"), ("qb", "<i4"), ("qe", "<i4"), ("rid", "<i4"), ("score", "<i4"),
                   ("truesc", "<i4"), ("sub", "<i4"), ("alt_sc", "<i4"), ("csub", "<i4"), ("sub_n", "<i4"),
                   ("w", "<i4"), ("seedcov", "<i4"), ("secondary", "<i4"), ("secondary_all", "<i4"),
                   ("seedlen0", "<i4"), ("n_comp_is_alt", "<i4"), ("frac_rep", "<f4"), ("hash", "<u8")])
assert ALNREG.itemsize == 88


def available() -> bool:
    return os.path.exists(LIB_PATH)


class _Batch(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("n_regs", C.c_int64), ("n_cigar", C.c_int64), ("reg_off", C.POINTER(C.c_int64)),
                ("regs", C.c_void_p), ("aln", C.POINTER(C.c_int64)), ("cigar", C.POINTER(C.c_uint32))]


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, a.ctypes.data_as(C.POINTER(C.c_uint8))


class RefBwa:
    def __init__(self):
        if not available():
            raise RuntimeError("oracle/_ref/libbwa_ref.so missing: run `make -C oracle ref` where /root/reference exists")
        L = self.lib = C.CDLL(LIB_PATH)
        L.ref_idx_load.restype = C.c_void_p; L.ref_idx_load.argtypes = [C.c_char_p]
        L.ref_idx_free.argtypes = [C.c_void_p]
        L.ref_idx_build.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_idx_lpac.restype = C.c_int64; L.ref_idx_lpac.argtypes = [C.c_void_p]
        L.ref_idx_nseqs.argtypes = [C.c_void_p]
        L.ref_batch.restype = C.POINTER(_Batch)
        L.ref_batch.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_batch_free.argtypes = [C.POINTER(_Batch)]
        L.ref_bwt_sa.restype = C.c_uint64; L.ref_bwt_sa.argtypes = [C.c_void_p, C.c_uint64]
        L.ref_bwt_2occ4.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.ref_bwt_meta.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        # every entry that takes the index handle declares its arguments: an undeclared Python int is passed as a 32-bit C int, which
        # only worked while the handle happened to lie in the low heap (found by running the suite under the ASan allocator)
        u8, i64, u64, i32 = C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_uint64), C.POINTER(C.c_int)
        L.ref_smem1.argtypes = [C.c_void_p, C.c_int, u8, C.c_int, C.c_int, u64, C.c_int, i32]
        L.ref_seed_strategy1.argtypes = [C.c_void_p, C.c_int, u8, C.c_int, C.c_int, C.c_int, u64]
        L.ref_mem_chain.argtypes = [C.c_void_p, C.c_int, u8, C.c_int, i64, C.c_int, i64, C.c_int, i32, C.POINTER(C.c_float)]
        L.ref_align1.argtypes = [C.c_void_p, C.c_int, u8, C.c_void_p, C.c_int]
        L.ref_pair.argtypes = [C.c_void_p, C.c_int, u8, C.c_int, u8, C.c_int, C.c_int, C.c_void_p, C.c_int, i32, C.c_void_p, C.c_int, i32]
        L.ref_reg2aln.argtypes = [C.c_void_p, C.c_int, u8, C.c_void_p, i64, C.POINTER(C.c_uint32), C.c_int]
        L.ref_ksw_extend2.argtypes = [C.c_int, u8, C.c_int, u8, C.c_int, C.c_int, C.c_int, C.c_int, i32]
        L.ref_ksw_global2.argtypes = [C.c_int, u8, C.c_int, u8, C.c_int, i32, C.POINTER(C.c_uint32), C.c_int]
        L.ref_ksw_align2.argtypes = [C.c_int, u8, C.c_int, u8, C.c_int, i32]
        assert L.ref_sizeof_alnreg() == 88
        self.idx = None

    # ---- index ----
    def build_index(self, fasta: str, prefix: str | None = None) -> str:
        prefix = prefix or fasta
        rc = self.lib.ref_idx_build(fasta.encode(), prefix.encode())
        if rc != 0:
            raise RuntimeError("bwa_idx_build failed")
        return prefix

    def load(self, prefix: str):
        self.idx = self.lib.ref_idx_load(prefix.encode())
        if not self.idx:
            raise RuntimeError(f"bwa_idx_load({prefix}) failed")
        return self

    @property
    def l_pac(self) -> int:
        return self.lib.ref_idx_lpac(self.idx)

    def opt_default(self) -> bytes:
        buf = C.create_string_buffer(self.lib.ref_sizeof_opt())
        self.lib.ref_opt_default(buf)
        return buf.raw

    # ---- primitives ----
    def ksw_extend2(self, q, t, w, end_bonus, zdrop, h0):
        q, qp = _u8(q); t, tp = _u8(t)
        out = (C.c_int * 5)()
        sc = self.lib.ref_ksw_extend2(len(q), qp, len(t), tp, w, end_bonus, zdrop, h0, out)
        return (sc, *list(out))  # score qle tle gtle gscore max_off

    def ksw_global2(self, q, t, w, want_cigar=True):
        q, qp = _u8(q); t, tp = _u8(t)
        n = C.c_int(0); cap = len(q) + len(t) + 8
        cig = (C.c_uint32 * cap)()
        sc = self.lib.ref_ksw_global2(len(q), qp, len(t), tp, w, C.byref(n) if want_cigar else None, cig, cap)
        return sc, list(cig[:n.value])

    def ksw_align2(self, q, t, xtra):
        q, qp = _u8(q); t, tp = _u8(t)
        out = (C.c_int * 7)()
        self.lib.ref_ksw_align2(len(q), qp, len(t), tp, xtra, out)
        return tuple(out)  # score te qe score2 te2 tb qb

    def occ4(self, k, l):
        a = (C.c_uint64 * 4)(); b = (C.c_uint64 * 4)()
        self.lib.ref_bwt_2occ4(self.idx, C.c_uint64(k & (2**64 - 1)), C.c_uint64(l & (2**64 - 1)), a, b)
        return list(a), list(b)

    def sa(self, k):
        return self.lib.ref_bwt_sa(self.idx, k)

    def bwt_meta(self):
        m = (C.c_uint64 * 7)(); self.lib.ref_bwt_meta(self.idx, m)
        return list(m)  # primary L2[0..4] seq_len

    def smem1(self, q, x, min_intv=1):
        q, qp = _u8(q); cap = len(q) + 4
        out = (C.c_uint64 * (4 * cap))(); n = C.c_int(0)
        ret = self.lib.ref_smem1(self.idx, len(q), qp, x, min_intv, out, cap, C.byref(n))
        return ret, np.array(out[:4 * n.value], dtype=np.uint64).reshape(-1, 4)

    def seed_strategy1(self, q, x, min_len=19, max_intv=20):
        q, qp = _u8(q); out = (C.c_uint64 * 4)()
        ret = self.lib.ref_seed_strategy1(self.idx, len(q), qp, x, min_len, max_intv, out)
        return ret, list(out)

    def mem_chain(self, q, do_flt=False, cap=4096):
        q, qp = _u8(q)
        ch = np.zeros((cap, 6), dtype=np.int64); sd = np.zeros((cap * 4, 4), dtype=np.int64)
        ns = C.c_int(0); fr = C.c_float(0)
        n = self.lib.ref_mem_chain(self.idx, len(q), qp, int(do_flt), ch.ctypes.data_as(C.POINTER(C.c_int64)), cap,
                                   sd.ctypes.data_as(C.POINTER(C.c_int64)), cap * 4, C.byref(ns), C.byref(fr))
        return ch[:n], sd[:ns.value], fr.value

    def align1(self, q, cap=512):
        q, qp = _u8(q); out = np.zeros(cap, dtype=ALNREG)
        n = self.lib.ref_align1(self.idx, len(q), qp, out.ctypes.data_as(C.c_void_p), cap)
        return out[:n]

    def pair(self, r1, r2, score_delta=25, rescue=True, cap=512):
        r1, p1 = _u8(r1); r2, p2 = _u8(r2)
        o1 = np.zeros(cap, dtype=ALNREG); o2 = np.zeros(cap, dtype=ALNREG); n1 = C.c_int(0); n2 = C.c_int(0)
        self.lib.ref_pair(self.idx, len(r1), p1, len(r2), p2, score_delta, int(rescue), o1.ctypes.data_as(C.c_void_p), cap,
                          C.byref(n1), o2.ctypes.data_as(C.c_void_p), cap, C.byref(n2))
        return o1[:n1.value], o2[:n2.value]

    def reg2aln(self, q, reg):
        q, qp = _u8(q); out = (C.c_int64 * 7)(); cap = 2 * len(q) + 8; cig = (C.c_uint32 * cap)()
        r = np.ascontiguousarray(reg)
        n = self.lib.ref_reg2aln(self.idx, len(q), qp, r.ctypes.data_as(C.c_void_p), out, cig, cap)
        return dict(pos=out[0], rid=out[1], is_rev=out[2], NM=out[3], n_cigar=out[4], score=out[5], sub=out[6],
                    cigar=list(cig[:max(0, n)]))

    # ---- batch (threads over pairs) ----
    def batch(self, seq, off, score_delta=25, rescue=True, cigar=True, threads=1):
        seq, sp = _u8(seq)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n_pairs = (len(off) - 1) // 2
        b = self.lib.ref_batch(self.idx, n_pairs, sp, off.ctypes.data_as(C.POINTER(C.c_int64)), score_delta, int(rescue), int(cigar), threads)
        try:
            bb = b.contents
            reg_off = np.ctypeslib.as_array(bb.reg_off, shape=(bb.n_reads + 1,)).copy()
            regs = np.frombuffer(C.string_at(bb.regs, bb.n_regs * 88), dtype=ALNREG).copy() if bb.n_regs else np.zeros(0, ALNREG)
            aln = np.ctypeslib.as_array(bb.aln, shape=(bb.n_regs + 1, 8))[:bb.n_regs].copy()
            cig = np.ctypeslib.as_array(bb.cigar, shape=(bb.n_cigar + 1,))[:bb.n_cigar].copy()
        finally:
            self.lib.ref_batch_free(b)
        return dict(reg_off=reg_off, regs=regs, aln=aln, cigar=cig)
