"""Barcode sharding across GPUs: the only multi-GPU strategy of this path.

A barcode's read pairs are an independent work unit all the way to its records (DoRFAForOneBarcode touches only its own
WorkUnit, the read-only index and the output channel; src/aligner/aligner.go:440-501), so the barcode-sorted input is cut
into `world` contiguous barcode ranges balanced by pair count, every rank runs the whole hot path on its range against
its own replica of the index, and the per-rank results are concatenated in rank order on the host.  No data-path
collective exists; torch.distributed is used only for the barrier / timing reduction and (optionally) to gather results.
"""
from __future__ import annotations

import numpy as np


def shard_barcodes(bc_off, world: int):
    """Cut n_barcodes work units into `world` contiguous ranges with near-equal pair counts.
    bc_off: n_barcodes+1 pair offsets.  Returns [(bc_lo, bc_hi)] * world (ranges may be empty); a barcode never
    straddles two ranks (the reader's over-cap continuation chunks of one barcode, reader.go:236,266-270, are separate
    work units in the reference as well)."""
    bc_off = np.asarray(bc_off, dtype=np.int64)
    n_bc = len(bc_off) - 1
    total = int(bc_off[-1] - bc_off[0]) if n_bc > 0 else 0
    cuts = [0]
    for r in range(1, world):
        target = bc_off[0] + (total * r) // world
        c = int(np.searchsorted(bc_off, target, side="left"))
        # choose the boundary closest to the target
        if c > 0 and c <= n_bc and abs(int(bc_off[c - 1]) - target) <= abs(int(bc_off[min(c, n_bc)]) - target):
            c -= 1
        cuts.append(min(max(c, cuts[-1]), n_bc))
    cuts.append(n_bc)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def shard_view(seq, off, bc_off, bc_lo: int, bc_hi: int):
    """The (seq, off, bc_off) triple of barcodes [bc_lo, bc_hi) re-based to zero (no copy of the bases beyond the slice)."""
    off = np.asarray(off, dtype=np.int64); bc_off = np.asarray(bc_off, dtype=np.int64)
    p0, p1 = int(bc_off[bc_lo]), int(bc_off[bc_hi])
    o = off[2 * p0:2 * p1 + 1]
    return np.ascontiguousarray(seq[o[0]:o[-1]]), (o - o[0]).copy(), (bc_off[bc_lo:bc_hi + 1] - p0).copy()


def merge_results(parts):
    """Concatenate per-rank align_batch results (rank order == barcode order): CSR offsets are re-based."""
    parts = [p for p in parts if p is not None]
    out = {}
    reg_off = [np.zeros(1, dtype=np.int64)]
    base = 0
    for p in parts:
        reg_off.append(np.asarray(p["reg_off"][1:], dtype=np.int64) + base)
        base += int(p["reg_off"][-1])
    out["reg_off"] = np.concatenate(reg_off)
    out["regs"] = np.concatenate([p["regs"] for p in parts]) if parts else None
    if parts and all("alns" in p for p in parts):
        alns, cb, mb = [], 0, 0
        for p in parts:
            a = p["alns"].copy(); a["cigar_off"] += cb; a["mm_off"] += mb
            cb += len(p["cigar"]); mb += len(p["mm"]); alns.append(a)
        out["alns"] = np.concatenate(alns)
        out["cigar"] = np.concatenate([p["cigar"] for p in parts]); out["mm"] = np.concatenate([p["mm"] for p in parts])
    return out
