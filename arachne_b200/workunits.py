"""Work units as the reference's FASTQ reader cuts them (src/fastqreader/reader.go:208-291 ReadBarcodeSet) and the
worthRunningRFA gate (src/aligner/aligner.go:1018-1030).  Host-side mirror for tests and bench.py; the native ingest
path (csrc/ingest.cpp) applies the same rule while parsing.

A record is one read PAIR (ReadOneLine consumes one R1 and one R2 record, reader.go:129-198).  ReadBarcodeSet returns
  * all consecutive records that share a barcode, at most 30000 (reader.go:236): a run that hits the cap comes back with
    full_barcode = false;
  * after a capped unit, continuation units of the SAME barcode break at record index 200, i.e. carry 201 records
    (reader.go:266-270), again with full_barcode = false;
  * the last continuation unit, ended by a barcode change before index 200 or by EOF, comes back with full_barcode = true
    (the reference's quirk: the tail of an over-cap barcode is treated as a barcode of its own);
  * a unit ended by a barcode change or EOF is full_barcode = true.
"""
from __future__ import annotations

import numpy as np

CAP = 30000
CONT = 200


def work_units(bc_off, barcodes=None):
    """bc_off: n_barcodes+1 pair offsets of a barcode-sorted input (equal adjacent barcode strings are merged first, as the
    reader only sees strings).  Returns (unit_off [n_units+1], unique [n_units] uint8, unit_bc [n_units] index of the barcode)."""
    bc_off = np.asarray(bc_off, dtype=np.int64)
    n_bc = len(bc_off) - 1
    runs = []  # (start, end, barcode index) of maximal runs of one barcode string
    b = 0
    while b < n_bc:
        e = b + 1
        if barcodes is not None:
            while e < n_bc and barcodes[e] == barcodes[b]:
                e += 1
        if bc_off[e] > bc_off[b]:
            runs.append((int(bc_off[b]), int(bc_off[e]), b))
        b = e
    uo, uq, ub = [0], [], []
    for s, e, b in runs:
        pos = s
        first = True
        while pos < e:
            left = e - pos
            if first:
                if left > CAP:          # cap reached without a barcode change: index ran to 30000, nothing dropped, not unique
                    n, uniq = CAP, 0
                elif left == CAP:       # the 30000th record fills the array before the next barcode's record is read: capped, not unique
                    n, uniq = CAP, 0
                else:
                    n, uniq = left, 1
                first = False
            else:                       # LastBarcode == this barcode: break at index 200 with 201 records unless the barcode changes first
                if left > CONT:
                    n, uniq = CONT + 1, 0
                else:
                    n, uniq = left, 1
            uo.append(pos + n); uq.append(uniq); ub.append(b)
            pos += n
    if not uq:
        return np.zeros(1, np.int64), np.zeros(0, np.uint8), np.zeros(0, np.int64)
    uo = np.array(uo, dtype=np.int64)
    uo = uo - uo[0] + (runs[0][0] if runs else 0)
    return uo, np.array(uq, dtype=np.uint8), np.array(ub, dtype=np.int64)


def worth_running_rfa(unit_off, unique, unit_bc, barcodes):
    """aligner.go:1018-1030: unique barcode, the barcode string contains a '-', at least 5 records."""
    n = np.diff(np.asarray(unit_off, dtype=np.int64))
    dash = np.array(["-" in barcodes[b] for b in unit_bc], dtype=bool) if len(unit_bc) else np.zeros(0, bool)
    return ((np.asarray(unique) != 0) & dash & (n >= 5)).astype(np.uint8)
