"""Synthetic references and simulated linked reads.

Host-side input generation for tests and bench.py only; nothing here is on the hot path.
All draws come from numpy.random.Generator(PCG64(seed)), so a (config, seed) pair names a dataset.

Sequences are handled as uint8 codes A=0 C=1 G=2 T=3 N=4, the encoding the reference's
`SequenceConvert` produces (src/gobwa/gobwa.go:159-167) before anything reaches the C side.
"""
from __future__ import annotations

import dataclasses
import numpy as np

BASES = np.frombuffer(b"ACGTN", dtype=np.uint8)
HUMAN_LIKE = [248, 242, 198, 190, 181, 171, 159, 145, 138, 133, 135, 133, 114, 107, 102, 90, 83, 80, 58, 64, 46, 50,
              156, 57]  # Mb, chr1..22,X,Y-ish; rescaled to the requested total


def rng(seed: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(seed))


def make_reference(total_len: int, n_contigs: int = 1, seed: int = 1, repeat_frac: float = 0.0, n_frac: float = 0.0):
    """iid-uniform ACGT contigs. Returns (list of (name, uint8 codes))."""
    g = rng(seed)
    if n_contigs == 1:
        lens = [total_len]
    else:
        w = np.array(HUMAN_LIKE[:n_contigs], dtype=np.float64)
        lens = np.maximum(1000, (w / w.sum() * total_len).astype(np.int64)).tolist()
    contigs = []
    for i, L in enumerate(lens):
        s = g.integers(0, 4, size=L, dtype=np.uint8)
        if repeat_frac > 0:  # 2-5 kb repeats copied with 1-3 % divergence (exercises rescue / optimizer)
            n_rep = int(L * repeat_frac / 3500)
            for _ in range(n_rep):
                rl = int(g.integers(2000, 5000))
                if L < 3 * rl:
                    break
                a = int(g.integers(0, L - rl)); b = int(g.integers(0, L - rl))
                seg = s[a:a + rl].copy()
                div = g.uniform(0.01, 0.03)
                m = g.random(rl) < div
                seg[m] = (seg[m] + g.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) & 3
                s[b:b + rl] = seg
        if n_frac > 0:
            n_run = max(1, int(L * n_frac / 100))
            for _ in range(n_run):
                a = int(g.integers(0, max(1, L - 100)))
                s[a:a + 100] = 4
        contigs.append((f"chr{i + 1}", s))
    return contigs


def write_fasta(path: str, contigs, width: int = 60) -> None:
    with open(path, "wb") as f:
        for name, s in contigs:
            f.write(b">" + name.encode() + b"\n")
            a = BASES[s]
            n = len(a) // width * width
            if n:
                block = np.empty((n // width, width + 1), dtype=np.uint8)
                block[:, :width] = a[:n].reshape(-1, width)
                block[:, width] = 10
                f.write(block.tobytes())
            if n < len(a):
                f.write(a[n:].tobytes() + b"\n")


def write_bns(prefix: str, contigs) -> int:
    """<prefix>.pac/.ann/.amb exactly as the reference's pack step leaves them after `bwa index`
    (bns_fasta2bntseq(for_only=1) + bns_dump, bwa/bntseq.c:66-96,275-328) for contigs without N: forward strand, 4 bases
    per byte with the first base in the top bits, a zero byte when l_pac % 4 == 0, then the byte l_pac % 4; .ann carries
    seed 11 and "(null)" annotations, .amb no holes.  Skips FASTA text for references too large to round-trip through it."""
    l_pac = int(sum(len(s) for _, s in contigs))
    with open(prefix + ".pac", "wb") as f:
        carry = np.zeros(0, np.uint8)
        for _, s in contigs:
            if (s > 3).any():
                raise ValueError("write_bns: N bases need the reference's own pack step (lrand48 replacement, bntseq.c:261)")
            a = np.concatenate([carry, s]) if len(carry) else s
            k = len(a) // 4 * 4
            q = a[:k].reshape(-1, 4)
            f.write((q[:, 0] << 6 | q[:, 1] << 4 | q[:, 2] << 2 | q[:, 3]).astype(np.uint8).tobytes())
            carry = a[k:].copy()
        if len(carry):
            v = 0
            for i, c in enumerate(carry):
                v |= int(c) << (6 - 2 * i)
            f.write(bytes([v]))
        else:
            f.write(b"\0")
        f.write(bytes([l_pac % 4]))
    with open(prefix + ".ann", "w") as f:
        f.write(f"{l_pac} {len(contigs)} 11\n")
        off = 0
        for name, s in contigs:
            f.write(f"0 {name} (null)\n{off} {len(s)} 0\n")
            off += len(s)
    with open(prefix + ".amb", "w") as f:
        f.write(f"{l_pac} {len(contigs)} 0\n")
    return l_pac


@dataclasses.dataclass
class ReadSet:
    """Barcode-sorted read pairs. seq is the concatenation of all reads (R1 of pair 0, R2 of pair 0, ...),
    off has 2*n_pairs+1 entries, bc_off has n_barcodes+1 entries (pair indices)."""
    seq: np.ndarray
    off: np.ndarray
    bc_off: np.ndarray
    barcodes: list
    names: list
    truth: np.ndarray  # per pair: contig index, fragment start, fragment length, r1 on reverse strand

    @property
    def n_pairs(self) -> int:
        return (len(self.off) - 1) // 2

    def read(self, i: int) -> np.ndarray:
        return self.seq[self.off[i]:self.off[i + 1]]

    def subset(self, bc_lo: int, bc_hi: int) -> "ReadSet":
        p0, p1 = int(self.bc_off[bc_lo]), int(self.bc_off[bc_hi])
        o = self.off[2 * p0:2 * p1 + 1]
        return ReadSet(self.seq[o[0]:o[-1]].copy(), (o - o[0]).copy(), (self.bc_off[bc_lo:bc_hi + 1] - p0).copy(),
                       self.barcodes[bc_lo:bc_hi], self.names[p0:p1], self.truth[p0:p1].copy())


def _revcomp(s: np.ndarray) -> np.ndarray:
    r = s[::-1].copy()
    m = r < 4
    r[m] = 3 - r[m]
    return r


def _mutate(g, s: np.ndarray, sub: float, ins: float, dele: float, nrate: float, L: int) -> np.ndarray:
    """Apply substitutions / indels / N to a template slice (len >= L + slack) and cut to L."""
    s = s.copy()
    n = len(s)
    m = g.random(n) < sub
    k = int(m.sum())
    if k:
        s[m] = (s[m] + g.integers(1, 4, size=k, dtype=np.uint8)) & 3
    if ins > 0 or dele > 0:
        r = g.random(n)
        idx = np.nonzero(r < ins + dele)[0]
        if len(idx):
            out = []
            last = 0
            for p in idx:
                p = int(p)
                if p < last:
                    continue
                out.append(s[last:p])
                if r[p] < ins:
                    out.append(g.integers(0, 4, size=int(g.integers(1, 4)), dtype=np.uint8))
                    last = p
                else:
                    last = p + int(g.integers(1, 4))
            out.append(s[last:])
            s = np.concatenate(out)
    s = s[:L]
    if nrate > 0:
        m = g.random(len(s)) < nrate
        s[m] = 4
    return s


def haplotag_name(g) -> str:
    a, c, b, d = g.integers(1, 97, size=4)
    return f"A{a:02d}C{c:02d}B{b:02d}D{d:02d}"


def simulate_linked_reads(contigs, n_barcodes: int, seed: int, *, read_len: int = 150, style: str = "haplotagging",
                          dash_suffix: bool = False, sub: float = 0.01, ins: float = 0.0005, dele: float = 0.0005,
                          nrate: float = 0.002, mol_per_bc: float = 3.0, pairs_per_mol: float = 8.0,
                          invalid_frac: float = 0.0, max_pairs: int | None = None) -> ReadSet:
    """Molecule-structured 2xL FR pairs, insert ~N(300,50) clipped to [max(160,L+10),500]."""
    g = rng(seed)
    clen = np.array([len(s) for _, s in contigs], dtype=np.int64)
    cprob = clen / clen.sum()
    seqs, lens, names, truth, bcs, bc_off = [], [], [], [], [], [0]
    n_pairs = 0
    slack = 16
    for b in range(n_barcodes):
        if style == "haplotagging":
            bc = haplotag_name(g)
        elif style == "stlfr":
            i, j, k = g.integers(1, 1537, size=3)
            bc = f"{i}_{j}_{k}"
        else:  # tellseq
            bc = "".join("ACGT"[x] for x in g.integers(0, 4, size=18))
        invalid = style == "tellseq" and g.random() < invalid_frac
        if dash_suffix:
            bc += "-1"
        n_mol = int(g.poisson(mol_per_bc)) + 1
        for _ in range(n_mol):
            ci = int(g.choice(len(contigs), p=cprob))
            ref = contigs[ci][1]
            mlen = int(np.clip(g.exponential(40000), 5000, 150000))
            mlen = min(mlen, len(ref) - 1)
            mstart = int(g.integers(0, len(ref) - mlen))
            npair = max(1, int(g.poisson(pairs_per_mol)))
            for _ in range(npair):
                isz = int(np.clip(g.normal(300, 50), max(160, read_len + 10), 500))
                if invalid:  # unbarcoded: random position, no molecule structure
                    ci = int(g.choice(len(contigs), p=cprob)); ref = contigs[ci][1]
                    fs = int(g.integers(0, len(ref) - isz - slack))
                else:
                    fs = mstart + int(g.integers(0, max(1, mlen - isz - slack)))
                    fs = min(fs, len(ref) - isz - slack - 1)
                frag = ref[fs:fs + isz + slack]
                fwd = _mutate(g, frag[:read_len + slack], sub, ins, dele, nrate, read_len)
                rv = _mutate(g, _revcomp(frag[:isz])[:read_len + slack] if isz >= read_len + slack else _revcomp(frag[:isz]), sub, ins, dele, nrate, read_len)
                flip = bool(g.random() < 0.5)
                r1, r2 = (rv, fwd) if flip else (fwd, rv)
                seqs.append(r1); seqs.append(r2); lens.append(len(r1)); lens.append(len(r2))
                names.append(f"sim{seed}_{n_pairs}")
                truth.append((ci, fs, isz, int(flip)))
                n_pairs += 1
        bcs.append(bc + ("" if not invalid else ""))
        bc_off.append(n_pairs)
        if max_pairs is not None and n_pairs >= max_pairs:
            break
    off = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(np.array(lens, dtype=np.int64), out=off[1:])
    return ReadSet(np.concatenate(seqs).astype(np.uint8), off, np.array(bc_off, dtype=np.int64), bcs, names,
                   np.array(truth, dtype=np.int64).reshape(-1, 4))


def simulate_pairs_fast(contigs, n_pairs: int, seed: int, *, read_len: int = 150, pairs_per_barcode: int = 30,
                        sub: float = 0.01, indel_read_frac: float = 0.14, nrate: float = 0.002,
                        mol_per_bc: int = 4) -> ReadSet:
    """Vectorised generator for bench-scale inputs (millions of pairs): same error model as
    simulate_linked_reads but one 1-3 bp indel per affected read and fixed pairs per barcode."""
    g = rng(seed)
    L = read_len
    slack = 8
    clen = np.array([len(s) for _, s in contigs], dtype=np.int64)
    cstart = np.concatenate([[0], np.cumsum(clen)])
    genome = np.concatenate([s for _, s in contigs]) if len(contigs) > 1 else contigs[0][1]
    n_bc = (n_pairs + pairs_per_barcode - 1) // pairs_per_barcode
    n_mol = n_bc * mol_per_bc
    mol_c = g.choice(len(contigs), size=n_mol, p=clen / clen.sum())
    mol_len = np.clip(g.exponential(40000, size=n_mol), 5000, 150000).astype(np.int64)
    mol_len = np.minimum(mol_len, clen[mol_c] - 600)
    mol_start = (g.random(n_mol) * (clen[mol_c] - mol_len - 600)).astype(np.int64)
    pair_bc = np.arange(n_pairs) // pairs_per_barcode
    pair_mol = pair_bc * mol_per_bc + g.integers(0, mol_per_bc, size=n_pairs)
    isz = np.clip(g.normal(300, 50, size=n_pairs), max(160, L + 10), 500).astype(np.int64)
    fs = mol_start[pair_mol] + (g.random(n_pairs) * np.maximum(1, mol_len[pair_mol] - isz)).astype(np.int64)
    ci = mol_c[pair_mol]
    gpos = cstart[ci] + fs
    ar = np.arange(L + slack, dtype=np.int64)
    fwd = genome[gpos[:, None] + ar[None, :]]
    rvt = genome[(gpos + isz - 1)[:, None] - ar[None, :]]
    rv = np.where(rvt < 4, 3 - rvt, rvt).astype(np.uint8)

    def mutate(a):
        m = g.random(a.shape) < sub
        a = np.where(m, (a + g.integers(1, 4, size=a.shape, dtype=np.uint8)) & 3, a).astype(np.uint8)
        rows = np.nonzero(g.random(a.shape[0]) < indel_read_frac)[0]
        pos = g.integers(10, L - 10, size=len(rows)); ln = g.integers(1, 4, size=len(rows)); isdel = g.random(len(rows)) < 0.5
        for r, p, l, d in zip(rows.tolist(), pos.tolist(), ln.tolist(), isdel.tolist()):
            row = a[r]
            if d:
                a[r, p:L + slack - l] = row[p + l:L + slack].copy()
            else:
                tmp = row[p:L + slack - l].copy()
                a[r, p:p + l] = g.integers(0, 4, size=l, dtype=np.uint8)
                a[r, p + l:L + slack] = tmp
        a = a[:, :L]
        if nrate > 0:
            a = np.where(g.random(a.shape) < nrate, 4, a).astype(np.uint8)
        return np.ascontiguousarray(a)

    fwd = mutate(fwd); rv = mutate(rv)
    flip = g.random(n_pairs) < 0.5
    seq = np.empty((n_pairs, 2, L), dtype=np.uint8)
    seq[:, 0] = np.where(flip[:, None], rv, fwd)
    seq[:, 1] = np.where(flip[:, None], fwd, rv)
    off = np.arange(2 * n_pairs + 1, dtype=np.int64) * L
    bc_off = np.minimum(np.arange(n_bc + 1, dtype=np.int64) * pairs_per_barcode, n_pairs)
    bcs = [f"A{(i % 96) + 1:02d}C{(i // 96 % 96) + 1:02d}B{(i // 9216 % 96) + 1:02d}D{(i // 884736 % 96) + 1:02d}" for i in range(n_bc)]
    names = [f"f{seed}_{i}" for i in range(n_pairs)] if n_pairs <= 2_000_000 else []
    truth = np.stack([ci, fs, isz, flip.astype(np.int64)], axis=1)
    return ReadSet(seq.reshape(-1), off, bc_off, bcs, names, truth)


# ---------------------------------------------------------------------------------------------- bench-scale workloads
def _barcode_plan(g, n_pairs: int, style: str, oversize_pairs: int = 0):
    """Pairs per molecule for every barcode until n_pairs are covered.  Returns (bc_sizes, mol_sizes, mol_bc).
    haplotagging / tellseq: molecules per barcode ~ Poisson(3) + 1, pairs per molecule ~ max(1, Poisson(8)).
    stlfr: pairs per barcode ~ lognormal(mean 200, sigma 1), one molecule per ~20 pairs; `oversize_pairs` > 0 puts one
    barcode of that many pairs first, so that the reader's 30000-record cap and 201-record continuation chunks
    (src/fastqreader/reader.go:236,266-270) are exercised."""
    bc_sizes, mol_sizes, mol_bc = [], [], []
    tot = 0
    est = 36 if style != "stlfr" else 200
    while tot < n_pairs:
        nb = max(16, int((n_pairs - tot) / est * 1.1) + 1)
        if style == "stlfr":
            sz = np.maximum(1, g.lognormal(np.log(200.0) - 0.5, 1.0, size=nb)).astype(np.int64)
            if oversize_pairs and not bc_sizes:
                sz[0] = oversize_pairs
            nm = np.maximum(1, np.rint(sz / 20.0)).astype(np.int64)
            for b in range(nb):
                cuts = np.sort(g.integers(0, sz[b] + 1, size=nm[b] - 1)) if nm[b] > 1 else np.zeros(0, np.int64)
                ms = np.diff(np.concatenate([[0], cuts, [sz[b]]]))
                ms = ms[ms > 0]
                mol_sizes.append(ms); mol_bc.append(np.full(len(ms), len(bc_sizes), np.int64)); bc_sizes.append(int(sz[b]))
                tot += int(sz[b])
                if tot >= n_pairs:
                    break
        else:
            nm = g.poisson(3.0, size=nb) + 1
            ms_all = np.maximum(1, g.poisson(8.0, size=int(nm.sum())))
            o = 0
            for b in range(nb):
                ms = ms_all[o:o + nm[b]]; o += nm[b]
                mol_sizes.append(ms); mol_bc.append(np.full(len(ms), len(bc_sizes), np.int64)); bc_sizes.append(int(ms.sum()))
                tot += int(ms.sum())
                if tot >= n_pairs:
                    break
    return np.array(bc_sizes, np.int64), np.concatenate(mol_sizes).astype(np.int64), np.concatenate(mol_bc)


def simulate_workload(contigs, n_pairs: int, seed: int, *, style: str = "haplotagging", read_len: int = 150, dash_suffix: bool = True,
                      sub: float = 0.01, ins: float = 0.0005, dele: float = 0.0005, nrate: float = 0.002, invalid_frac: float = 0.0,
                      oversize_pairs: int = 0, want_names: bool = True) -> ReadSet:
    """Vectorised molecule-structured generator for bench-scale inputs: per-base substitution / insertion /
    deletion rates (indels 1-3 bp), insert ~ N(300, 50) clipped to [max(160, L + 10), 500], FR pairs, R1 strand random.
    style: haplotagging (AxxCxxBxxDxx), stlfr (i_j_k, lognormal pairs per barcode), tellseq (18-mer; `invalid_frac` of the
    barcodes carry an N, are not molecule-grouped and never run RFA).  The last barcode is cut so that exactly n_pairs come out.
    `barcodes` holds one string per barcode with the '-1' suffix when dash_suffix (worthRunningRFA needs a '-', aligner.go:1022)."""
    g = rng(seed)
    L = read_len
    slack = 16
    clen = np.array([len(s) for _, s in contigs], dtype=np.int64)
    cstart = np.concatenate([[0], np.cumsum(clen)])
    genome = np.concatenate([s for _, s in contigs]) if len(contigs) > 1 else contigs[0][1]
    bc_sizes, mol_sizes, mol_bc = _barcode_plan(g, n_pairs, style, oversize_pairs)
    n_mol = len(mol_sizes)
    pair_mol = np.repeat(np.arange(n_mol), mol_sizes)[:n_pairs]
    pair_bc = mol_bc[pair_mol]
    n_bc = int(pair_bc[-1]) + 1
    bc_off = np.concatenate([[0], np.cumsum(bc_sizes[:n_bc])]).astype(np.int64)
    bc_off[-1] = n_pairs
    mol_c = g.choice(len(contigs), size=n_mol, p=clen / clen.sum())
    mol_len = np.clip(g.exponential(40000, size=n_mol), 5000, 150000).astype(np.int64)
    mol_len = np.minimum(mol_len, clen[mol_c] - 600)
    mol_start = (g.random(n_mol) * (clen[mol_c] - mol_len - 600)).astype(np.int64)
    isz = np.clip(g.normal(300, 50, size=n_pairs), max(160, L + 10), 500).astype(np.int64)
    ci = mol_c[pair_mol]
    fs = mol_start[pair_mol] + (g.random(n_pairs) * np.maximum(1, mol_len[pair_mol] - isz)).astype(np.int64)
    bc_invalid = np.zeros(n_bc, bool)
    if style == "tellseq" and invalid_frac > 0:  # unbarcoded pairs: random positions, no molecule structure
        bc_invalid = g.random(n_bc) < invalid_frac
        inv = bc_invalid[pair_bc]
        k = int(inv.sum())
        ci2 = g.choice(len(contigs), size=k, p=clen / clen.sum())
        ci = ci.copy(); fs = fs.copy()
        ci[inv] = ci2
        fs[inv] = (g.random(k) * (clen[ci2] - 600)).astype(np.int64)
    gpos = cstart[ci] + fs
    ar = np.arange(L + slack, dtype=np.int64)
    fwd = genome[gpos[:, None] + ar[None, :]]
    rvt = genome[(gpos + isz - 1)[:, None] - ar[None, :]]
    rv = np.where(rvt < 4, 3 - rvt, rvt).astype(np.uint8)

    def mutate(a):
        m = g.random(a.shape) < sub
        a = np.where(m, (a + g.integers(1, 4, size=a.shape, dtype=np.uint8)) & 3, a).astype(np.uint8)
        n_ev = g.poisson((ins + dele) * L, size=a.shape[0])
        rows = np.nonzero(n_ev)[0]
        W = L + slack
        for r in rows.tolist():
            row = a[r]
            for _ in range(int(n_ev[r])):
                p = int(g.integers(5, L - 5)); ln = int(g.integers(1, 4))
                if g.random() < dele / (ins + dele):
                    row[p:W - ln] = row[p + ln:W].copy()
                else:
                    tmp = row[p:W - ln].copy()
                    row[p:p + ln] = g.integers(0, 4, size=ln, dtype=np.uint8)
                    row[p + ln:W] = tmp
        a = a[:, :L]
        if nrate > 0:
            a = np.where(g.random(a.shape) < nrate, 4, a).astype(np.uint8)
        return np.ascontiguousarray(a)

    fwd = mutate(fwd); rv = mutate(rv)
    flip = g.random(n_pairs) < 0.5
    seq = np.empty((n_pairs, 2, L), dtype=np.uint8)
    seq[:, 0] = np.where(flip[:, None], rv, fwd)
    seq[:, 1] = np.where(flip[:, None], fwd, rv)
    off = np.arange(2 * n_pairs + 1, dtype=np.int64) * L
    sfx = "-1" if dash_suffix else ""
    if style == "stlfr":
        ijk = g.integers(1, 1537, size=(n_bc, 3))
        bcs = [f"{a}_{b}_{c}{sfx}" for a, b, c in ijk.tolist()]
    elif style == "tellseq":
        codes = g.integers(0, 4, size=(n_bc, 18))
        bcs = ["".join("ACGT"[x] for x in row) for row in codes.tolist()]
        bcs = [(b[:7] + "N" + b[8:]) if bad else b + sfx for b, bad in zip(bcs, bc_invalid.tolist())]
    else:
        bcs = [f"A{(i % 96) + 1:02d}C{(i // 96 % 96) + 1:02d}B{(i // 9216 % 96) + 1:02d}D{(i // 884736 % 96) + 1:02d}{sfx}" for i in range(n_bc)]
    names = [f"w{seed}_{i}" for i in range(n_pairs)] if want_names else []
    truth = np.stack([ci, fs, isz, flip.astype(np.int64)], axis=1)
    return ReadSet(seq.reshape(-1), off, bc_off, bcs, names, truth)
