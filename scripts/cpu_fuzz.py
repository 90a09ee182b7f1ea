#!/usr/bin/env python
"""Time-boxed differential fuzzing of what runs without a GPU, each against the reference's own compiled C code (oracle/_ref):
  ksw    the oracle's restatement of ksw_extend2 / ksw_global2 / ksw_align2 (the checker of the GPU DP tests) on random + adversarial problems,
         and the SHIPPED extension DP kernel k_ext_dp run as one warp of fibers on the host
  index  the index-construction driver (csrc/index_build.cuh on the test-only host backend) -> .bwt / .sa bytes vs `bwa index`
  pack   arachne_index_pack -> .pac / .ann / .amb vs `bwa index` on FASTA with N runs, IUPAC codes, lower case, CRLF, odd line widths
  ingest the FASTQ ingest (csrc/ingest.cu) vs the line-by-line Python restatement of reader.go in tests/test_cpu_tier.py (the one part
         without a compiled reference: Go) on headers with odd white space, repeated / empty / trailing BX and VX tags, short names,
         barcode runs around the 200- and 30000-record rules, plain / gzip / BGZF files
  headers the preprocess row's hand-written header scanners (the reference's five patterns, standardize.go:16-25) vs Python `re` on the same
         patterns (oracle/preprocess_oracle.py), and the header rewriting of whole files in the four formats
python scripts/cpu_fuzz.py [seconds per part=120] [first_seed=1] [parts=ksw,index,pack,ingest,headers]   -> one JSON line; non-zero exit on the first divergence."""
import json, os, shutil, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import ArachneCuda, simulate as S
from oracle.refbwa import RefBwa
from tests import util

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
parts = (sys.argv[3] if len(sys.argv) > 3 else "ksw,index,pack,ingest,headers").split(",")
out = dict(first_seed=seed0)


def fail(what, **kw):
    print(json.dumps(dict(divergence=what, **kw, **out))); sys.exit(1)


# ---- ksw ----
O = util.oracle_lib(); R = RefBwa(); t0 = time.time(); n = 0; seed = seed0; n_dp = 0
HS = util.HostSim(); OPT = ArachneCuda().opt
while "ksw" in parts and time.time() - t0 < budget:
    g = np.random.default_rng(seed)
    qs, ts, h0, w = util.extension_problems(seed, 300)
    # the SHIPPED DP kernel k_ext_dp as one warp of fibers on the host (tests/hostsim/hostsim_extdp.cpp), against the reference itself
    zd, eb = [(100, 5), (0, 0), (20, 5), (10, 0), (1, 0)][seed % 5]
    got, _ = HS.ext_dp(OPT, qs, ts, h0, w, end_bonus=eb, zdrop=zd, codec=seed % 2, warp=seed % 4)
    for i in range(len(qs)):
        if tuple(int(x) for x in got[i]) != R.ksw_extend2(qs[i], ts[i], int(w[i]), eb, zd, int(h0[i])):
            fail("k_ext_dp", seed=seed, i=i)
    n_dp += len(qs)
    # the allocation-free entry's launch sequence around the same kernel (csrc/xs_kernels.cuh; scratch persisting between calls), with refusals
    hb = h0.copy(); bad = sorted(set(int(x) for x in g.integers(0, len(qs), 3))); hb[bad] = 0
    got2, status, _ = HS.xs_static(OPT, qs, ts, hb, w, end_bonus=eb, zdrop=zd, warp=(seed + 1) % 4)
    if status != bad[0] + 1 or any(tuple(got2[i]) != ((-1,) * 6 if i in bad else tuple(got[i])) for i in range(len(qs))):
        fail("arachne_ksw_extend_static kernels", seed=seed)
    for i in range(len(qs)):
        zd, eb = [(100, 5), (0, 0), (20, 5), (10, 0), (1, 0)][i % 5]
        if util.orc_extend(O, qs[i], ts[i], int(w[i]), eb, zd, int(h0[i]))[0] != R.ksw_extend2(qs[i], ts[i], int(w[i]), eb, zd, int(h0[i])):
            fail("ksw_extend2", seed=seed, i=i)
        if len(ts[i]) and len(qs[i]):
            ww = max(abs(len(ts[i]) - len(qs[i])) + 3, int(w[i]))
            if util.orc_global(O, qs[i], ts[i], ww) != R.ksw_global2(qs[i], ts[i], ww):
                fail("ksw_global2", seed=seed, i=i)
        if i % 3 == 0 and len(qs[i]) >= 20 and len(ts[i]):
            L = len(qs[i]); xtra = 0x40000 | 0x80000 | (0x10000 if L < 250 else 0) | 19
            t = np.concatenate([g.integers(0, 4, int(g.integers(0, 200))).astype(np.uint8), ts[i], g.integers(0, 4, int(g.integers(0, 200))).astype(np.uint8)])
            if util.orc_local(O, qs[i], t, xtra) != R.ksw_align2(qs[i], t, xtra):
                fail("ksw_align2", seed=seed, i=i)
        n += 1
    seed += 1
out["ksw"] = dict(problems=n, shipped_dp_kernel_problems=n_dp, seconds=round(time.time() - t0, 1))

# ---- index construction ----
t0 = time.time(); n = 0; seed = seed0; total = 0
with tempfile.TemporaryDirectory() as wd:
    while "index" in parts and time.time() - t0 < budget:
        g = np.random.default_rng(seed)
        kind = int(g.integers(0, 5)); L = int(g.integers(1, 6000)) if kind else int(g.integers(1, 40))
        if kind <= 1: s = g.integers(0, 4, L)
        elif kind == 2: s = np.tile(g.integers(0, 4, int(g.integers(1, 9))), L // 4 + 1)[:L]            # tandem
        elif kind == 3: s = np.concatenate([g.integers(0, 4, L), np.full(int(g.integers(1, 300)), int(g.integers(0, 4)))])  # homopolymer at the end
        else:
            unit = g.integers(0, 4, int(g.integers(20, 400)))                                            # exact repeats with unique spacers
            s = np.concatenate([np.concatenate([unit, g.integers(0, 4, int(g.integers(0, 50)))]) for _ in range(int(g.integers(2, 12)))])
        cuts = sorted(set(int(x) for x in g.integers(1, max(2, len(s)), int(g.integers(0, 4))))) if len(s) > 2 else []
        ctg = [("c%d" % k, np.asarray(p, np.uint8)) for k, p in enumerate(np.split(np.asarray(s, np.uint8), cuts)) if len(p)]
        fa = os.path.join(wd, "i%d.fa" % seed)
        S.write_fasta(fa, ctg); RefBwa().build_index(fa)
        l_pac = sum(len(p) for _, p in ctg)
        bwt, sa, st = util.hostsim_index_build(open(fa + ".pac", "rb").read(), l_pac, batch_cap=int(g.choice([0, 0, 64, 1000])))
        if bwt != open(fa + ".bwt", "rb").read() or sa != open(fa + ".sa", "rb").read():
            fail("index_build", seed=seed, l_pac=l_pac)
        for ext in ("", ".bwt", ".sa", ".pac", ".ann", ".amb"):
            os.remove(fa + ext)
        n += 1; total += l_pac; seed += 1
out["index"] = dict(references=n, bases=total, seconds=round(time.time() - t0, 1))

# ---- pack step ----
A = ArachneCuda(); t0 = time.time(); n = 0; seed = seed0
with tempfile.TemporaryDirectory() as wd:
    while "pack" in parts and time.time() - t0 < budget:
        # kseq's corner cases on purpose: CR-only lines, CRLF, text before the first header, '@' headers, '>' / '@' / '+' inside lines, lines
        # that START with '+' (the record turns into FASTQ: accepted when the quality length matches, reading stops when it does not),
        # comments that are a lone CR, names run into the end of the file, blank lines, no final newline
        g = np.random.default_rng(seed); txt = ""
        eol = "\r\n" if seed % 5 == 3 else "\n"
        if seed % 7 == 2: txt += "leading junk without a header" + eol + eol
        for c in range(int(g.integers(1, 7))):
            L = int(g.integers(1, 3000))
            s = np.array(list("ACGT"))[g.integers(0, 4, L)]
            for _ in range(int(g.integers(0, 6))):
                a = int(g.integers(0, L)); k = int(g.integers(1, 60)); s[a:a + k] = g.choice(list("NnRYKMSWBDHVryu-.*>@+ "))
            low = g.random(L) < g.choice([0.0, 0.3]); s[low] = np.char.lower(s[low])
            s = "".join(s)
            head = (">" if g.random() < 0.8 else "@") + "ctg%d_%d" % (seed, c)
            u = g.random()
            if u < 0.4: head += (" " if c % 3 else "\t") + "note %d x" % c
            elif u < 0.5: head += " "            # an empty comment; with CRLF the comment is a lone CR
            txt += head + eol
            w = int(g.integers(1, 120)); fastq = g.random() < 0.15; seq_chars = 0
            for i in range(0, L, w):
                line = s[i:i + w]
                if line[0] in ">@+" and g.random() < 0.7: line = "A" + line[1:]   # mostly keep the record going; sometimes let kseq see a header / '+' here
                txt += line + eol
                if g.random() < 0.03: txt += eol
            if fastq:
                txt += "+" + ("anything" if g.random() < 0.5 else "") + eol
                ql = L if g.random() < 0.7 else int(g.integers(0, L + 5))
                q = "".join(chr(33 + int(x)) for x in g.integers(0, 40, ql))
                for i in range(0, ql, w): txt += q[i:i + w] + eol
        if seed % 11 == 4: txt = txt.rstrip("\r\n")
        if seed % 13 == 5: txt += ">lastname_without_newline"
        fa = os.path.join(wd, "p%d.fa" % seed); open(fa, "w", newline="").write(txt)
        ref = os.path.join(wd, "ref%d" % seed); shutil.copy(fa, ref); RefBwa().build_index(ref)
        mine = os.path.join(wd, "mine%d" % seed)
        ref_empty = open(ref + ".ann").read().split()[1] == "0"  # no record survived kseq: the reference leaves degenerate files, we refuse
        try:
            A.pack_fasta(fa, mine)
            if ref_empty: fail("pack accepted an input without records", seed=seed)
        except Exception as e:
            if not ref_empty: fail("pack error: %s" % e, seed=seed)
            for p in os.listdir(wd): os.remove(os.path.join(wd, p))
            n += 1; seed += 1
            continue
        for ext in ("pac", "ann", "amb"):
            if open(mine + "." + ext, "rb").read() != open(ref + "." + ext, "rb").read(): fail("pack ." + ext, seed=seed)
        for p in os.listdir(wd): os.remove(os.path.join(wd, p))
        n += 1; seed += 1
out["pack"] = dict(fasta_files=n, seconds=round(time.time() - t0, 1))
# ---- FASTQ ingest ----
import gzip
sys.path.insert(0, os.path.join(ROOT, "tests"))
from tests.test_cpu_tier import _reader_go, _bgzf_bytes
from arachne_b200 import Ingest
t0 = time.time(); n = 0; seed = seed0; npairs = 0
with tempfile.TemporaryDirectory() as wd:
    while "ingest" in parts and time.time() - t0 < budget:
        g = np.random.default_rng(seed)
        ws = [" ", "\t", "  ", "\t ", "\v", "\f", "\r"]
        def header(k, mate, bc, vx):
            name = g.choice(["read%d" % k, "r", "", "x%d" % k, "a b"]) + "/%d" % mate
            tags = []
            u = g.random()
            if bc is not None:
                if u < 0.85: tags.append("BX:Z:" + bc)
                elif u < 0.9: tags += ["BX:Z:", "BX:Z:" + bc]          # an empty BX first: the regex moves on to the next one
                elif u < 0.95: tags += ["xBX:Z:" + bc]                   # matches inside a longer token too
                else: tags += ["BX:Z:" + bc + "\vmore"]                 # a vertical tab is \S for the regex
            if vx is not None: tags.append("VX:i:%s" % vx)
            if g.random() < 0.3: tags.append("RG:%d" % (k % 3))
            g.shuffle(tags)
            h = "@" + name
            for t in tags: h += str(g.choice(ws)) + t
            if g.random() < 0.1: h += str(g.choice(ws))
            return h + "\n"
        r1, r2, k = [], [], 0
        plan = []
        for _ in range(int(g.integers(1, 12))):
            bc = g.choice(["A01C02B03D04-1", "B%02d-1" % g.integers(0, 99), "plain%d" % g.integers(0, 9), None, "N-N"])
            cnt = int(g.choice([1, 2, 4, 5, 6, 30, 199, 200, 201, 202, 403])) if g.random() < 0.97 else 30000 + int(g.integers(0, 500))
            plan.append((None if bc is None else str(bc), cnt))
        if g.random() < 0.3: plan.append(plan[-1])                       # the same barcode again right after: the continuation rule
        for bc, cnt in plan:
            for _ in range(cnt):
                L1 = int(g.integers(0, 152)) if g.random() < 0.05 else 40; L2 = 40
                s1 = "".join("ACGTNacgt-"[int(x)] for x in g.integers(0, 10 if k % 40 == 0 else 4, L1)); s2 = "ACGT" * 10
                q1 = "I" * (L1 if g.random() < 0.97 else max(0, L1 - 3)); q2 = "I" * L2
                vx = g.choice(["0", "1", None, "2"])
                r1 += [header(k, 1, bc, None if vx is None else str(vx)), s1 + "\n", "+\n", q1 + "\n"]; r2 += [header(k, 2, bc, None), s2 + "\n", "+\n", q2 + "\n"]
                k += 1
        if g.random() < 0.2: r2 = r2[:len(r2) - 4 * int(g.integers(1, 3))]   # R2 ends early
        if g.random() < 0.2: r1[-1] = r1[-1][:-1]                            # no final newline
        kind = int(g.integers(0, 4))
        p1, p2 = os.path.join(wd, "a%d" % seed), os.path.join(wd, "b%d" % seed)
        for p, lines in ((p1, r1), (p2, r2)):
            text = "".join(lines).encode("latin-1")
            blob = text if kind == 0 else gzip.compress(text, 1) if kind == 1 else _bgzf_bytes(text, int(g.choice([0xff00, 500, 4096])))
            open(p, "wb").write(blob)
        if not r1[-1].endswith("\n"): r1[-1] += "\n"
        recs, units = _reader_go(r1, r2)
        # quality shorter than the sequence is padded with '!' by the ingest (the reference would index past it)
        I = Ingest(p1, p2, n_threads=int(g.integers(1, 5)))
        try:
            got_units, pair = [], 0
            while True:
                b = I.next(int(g.choice([1, 7, 300, 10**9])))
                if b is None: break
                for u in range(b["n_units"]):
                    lo, hi = int(b["unit_off"][u]), int(b["unit_off"][u + 1])
                    got_units.append((pair + lo, pair + hi, bool(b["unique"][u])))
                    first = recs[pair + lo]
                    if b["barcodes"][u] != first[1] or bool(b["valid"][u]) != first[2] or bool(b["run_rfa"][u]) != (bool(b["unique"][u]) and "-" in first[1] and hi - lo >= 5):
                        fail("ingest unit fields", seed=seed, unit=len(got_units) - 1)
                for i in range(b["n_pairs"]):
                    name, bc, valid, rg, s1, q1, s2, q2 = recs[pair + i]
                    o = b["off"][2 * i:2 * i + 3]
                    if b["names"][i] != name or b["read_groups"][i] != rg or o[1] - o[0] != len(s1) or o[2] - o[1] != len(s2) or b["quals"][2 * i] != (q1 + "!" * len(s1))[:len(s1)]:
                        fail("ingest record fields", seed=seed, pair=pair + i, got=(b["names"][i], b["read_groups"][i]), want=(name, rg))
                pair += b["n_pairs"]
            if pair != len(recs) or got_units != units: fail("ingest units", seed=seed, pairs=(pair, len(recs)))
        finally:
            I.close()
        for p in os.listdir(wd): os.remove(os.path.join(wd, p))
        n += 1; npairs += len(recs); seed += 1
out["ingest"] = dict(file_pairs=n, pairs=npairs, seconds=round(time.time() - t0, 1))
# ---- preprocess: header patterns and rewriting ----
from arachne_b200 import lib as L
from oracle import preprocess_oracle as PO
from tests.test_preprocess import make_fastq
t0 = time.time(); n = 0; nf = 0; seed = seed0
pieces = ["r1", "/1", "/2", "\t", " ", "BX:Z:", "VX:i:", "0", "1", "2", "A01C02B03D04", "A01C02B03D0", "A99C00B12D77", "#", "12_7_1444", "1_2", "_", "x", ":", "ACGTNACGTACGTACGTA",
          "ACGTACGTACGTACGTAC", "1:N:0:ATCACG", "2:N:0:", "atc", "N", "RG:Z:l", "\t\t", "  ", "", "A01", "C02", "B03D04", "D04\t", "Y", ",", "+", "0_0_0", "1536_1536_1536"]
with tempfile.TemporaryDirectory() as wd:
    while "headers" in parts and time.time() - t0 < budget:
        g = np.random.default_rng(seed)
        for _ in range(2000):
            h = "".join(str(g.choice(pieces)) for _ in range(int(g.integers(0, 9))))
            if L.fastq_header_format(h) != PO.header_format(h): fail("header_format", seed=seed, header=h, got=L.fastq_header_format(h), want=PO.header_format(h))
            n += 1
        fmt = str(g.choice(["standard", "haplotagging", "stlfr", "tellseq"])); gz = bool(g.random() < 0.3)
        d = os.path.join(wd, "f%d" % seed); os.makedirs(d)
        r1, r2 = make_fastq(d, fmt, n=int(g.integers(1, 600)), seed=int(g.integers(1 << 30)), gz=gz)
        o1, o2 = os.path.join(d, "o1"), os.path.join(d, "o2")
        try:
            st = L.preprocess(r1, r2, o1, o2, sort=False); e1, e2, est = PO.preprocess(r1, r2, sort=False)
            if open(o1, "rb").read() != e1 or open(o2, "rb").read() != e2 or any(st[k] != est[k] for k in ("n_pairs", "format", "n_barcodes", "n_no_tag", "n_invalid")):
                fail("standardize", seed=seed, fmt=fmt)
        except L.ArachneError as e:   # the first 200 headers may not decide a format for a tiny file: both sides must then refuse
            try:
                PO.preprocess(r1, r2, sort=False); fail("standardize refused: %s" % e, seed=seed, fmt=fmt)
            except Exception:
                pass
        shutil.rmtree(d); nf += 1; seed += 1
out["headers"] = dict(headers=n, files=nf, seconds=round(time.time() - t0, 1))
print(json.dumps(dict(divergence=None, **out)))
