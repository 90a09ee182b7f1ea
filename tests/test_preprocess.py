"""FASTQ preprocessing (include/arachne_preprocess.h) against oracle/preprocess_oracle.py.

Parity for this row is UNPINNED (no compilable reference source, no samtools in the image: see the oracle's header); the tests pin the
library to the restatement and to properties the reference's pipeline must have: pairs stay together, the order is the stable
strcmp order of the tag with untagged pairs first, nothing but the header changes, and the reader groups the output into one unit per
barcode."""
import gzip
import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import lib as L  # noqa: E402
from oracle import preprocess_oracle as O  # noqa: E402


def _seq(rng, n):
    return "".join(rng.choice("ACGT") for _ in range(n))


def make_fastq(d, fmt, n=400, seed=5, gz=False):
    """Unsorted paired FASTQ in one of the four formats, with the awkward cases of each."""
    rng = random.Random(seed)
    if fmt == "haplotagging":
        pool = ["A%02dC%02dB%02dD%02d" % tuple(rng.randrange(0, 5) for _ in range(4)) for _ in range(40)]
    elif fmt == "stlfr":
        pool = ["%d_%d_%d" % tuple(rng.randrange(0, 30) for _ in range(3)) for _ in range(40)] + ["0_0_0", "10_0_3", "7_12_0", "0_4_4"]
    elif fmt == "tellseq":
        pool = ["".join(rng.choice("ACGTN" if k % 7 == 0 else "ACGT") for _ in range(18)) for k in range(40)]
    else:  # standard: free-form barcodes of very different lengths (1..32 bytes: all four key words)
        pool = ["b" * k + "%d" % k for k in range(0, 30, 3)] + ["A01C02B03D04", "A01C02B03D0", "A01C02B03D044", "Z", "a", "x" * 32]
    p1, p2 = os.path.join(d, fmt + ".R1.fq" + (".gz" if gz else "")), os.path.join(d, fmt + ".R2.fq" + (".gz" if gz else ""))
    o1, o2 = [], []
    for i in range(n):
        bc = rng.choice(pool)
        name = "r%05d:%d:x" % (i, rng.randrange(100))
        l1, l2 = rng.randrange(30, 151), rng.randrange(30, 151)
        s1, s2 = _seq(rng, l1), _seq(rng, l2)
        q1, q2 = "".join(chr(33 + rng.randrange(40)) for _ in range(l1)), "".join(chr(33 + rng.randrange(40)) for _ in range(l2))
        if fmt == "haplotagging":
            h1, h2 = "%s/1\tBX:Z:%s\tRX:Z:ACGT" % (name, bc), "%s/2\tBX:Z:%s" % (name, bc)
            if i % 37 == 5:  # a record that lost its tag
                h1, h2 = name + "/1", name + "/2"
        elif fmt == "stlfr":
            h1, h2 = "%s#%s/1" % (name, bc), "%s#%s/2 extra words" % (name, bc)
        elif fmt == "tellseq":
            h1, h2 = "%s:%s 1:N:0:ATCACG" % (name, bc), "%s:%s 2:N:0:ATCACG" % (name, bc)
        else:
            vx = rng.randrange(2)
            h1, h2 = "%s/1 BX:Z:%s\tVX:i:%d\tRG:Z:lane%d" % (name, bc, vx, i % 3), "%s/2\tBX:Z:%s VX:i:%d casava:stuff" % (name, bc, vx)
            if i % 41 == 7:
                h1, h2 = "%s/1\tVX:i:1" % name, "%s/2\tVX:i:1" % name
        o1.append("@%s\n%s\n+\n%s\n" % (h1, s1, q1)); o2.append("@%s\n%s\n+\n%s\n" % (h2, s2, q2))
    op = gzip.open if gz else open
    with op(p1, "wb") as f:
        f.write("".join(o1).encode())
    with op(p2, "wb") as f:
        f.write("".join(o2).encode())
    return p1, p2


HEADERS = [
    ("r1/1\tBX:Z:A01C02B03D04\tVX:i:1", "standard"),
    ("r1/1 BX:Z:whatever VX:i:0 more", "standard"),
    ("r1/1\tBX:Z:A01C02B03D04", "haplotagging"),
    ("r1/1\tBX:Z:A01C02B03D04\tVX:i:2", "haplotagging"),       # VX must be 0 or 1
    ("r1/1\tBX:Z:A01C02B03D0", "unknown"),                       # 11 characters
    ("r1/1\tBX:Z:A01C02B03D045", "unknown"),                     # 13: no white space after the 12th
    ("V300#12_7_1444/1", "stlfr"),
    ("V300#12_7_1444/2 extra", "stlfr"),
    ("V300#12_7/1", "unknown"),
    ("V300#12_7_x/1", "unknown"),
    ("A00:1:HX:1:1101:1000:2000:ACGTNACGTACGTACGTA 1:N:0:ATCACG", "tellseq"),
    ("plain_name 1:N:0:ATCACG", "tellseq"),                      # the casava index also fits the pattern, as in the reference
    ("plain_name 1:N:0:atcacg", "unknown"),
    ("plain_name", "unknown"),
    ("", "unknown"),
    ("x#1_2_3/1\tBX:Z:q\tVX:i:1", "standard"),                  # the chain tests `standard` first
]


@pytest.mark.parametrize("header,expect", HEADERS)
def test_header_patterns(header, expect):
    assert O.header_format(header) == expect
    assert L.fastq_header_format(header) == expect


@pytest.mark.parametrize("fmt", ["standard", "haplotagging", "stlfr", "tellseq"])
@pytest.mark.parametrize("gz", [False, True])
def test_standardize_matches_restatement(tmp_path, fmt, gz):
    """sort=0 needs no device: format sniffing, header rewriting, tag passthrough, plain and gzip in and out."""
    d = str(tmp_path)
    r1, r2 = make_fastq(d, fmt, gz=gz)
    assert L.fastq_format(r1, r2) == fmt
    o1, o2 = os.path.join(d, "o1.fq" + (".gz" if gz else "")), os.path.join(d, "o2.fq" + (".gz" if gz else ""))
    st = L.preprocess(r1, r2, o1, o2, sort=False)
    e1, e2, est = O.preprocess(r1, r2, sort=False)
    rd = (lambda p: gzip.open(p, "rb").read()) if gz else (lambda p: open(p, "rb").read())
    assert rd(o1) == e1 and rd(o2) == e2
    for k in ("n_pairs", "format", "n_barcodes", "n_no_tag", "n_invalid"):
        assert st[k] == est[k], k
    # the output is in the standard format and the reader's ParseHeader finds the same barcode in it
    assert L.fastq_format(o1, o2) == "standard"


def test_unknown_format_and_bad_input_fail(tmp_path):
    d = str(tmp_path)
    a, b = os.path.join(d, "a.fq"), os.path.join(d, "b.fq")
    open(a, "w").write("@n1/1\nACGT\n+\nIIII\n@n2/1\nACGT\n+\nIIII\n"); open(b, "w").write("@n1/2\nACGT\n+\nIIII\n@n2/2\nACGT\n+\nIIII\n")
    assert L.fastq_format(a, b) == "unknown"
    with pytest.raises(L.ArachneError):
        L.preprocess(a, b, a + ".o", b + ".o", sort=False)
    st = L.preprocess(a, b, a + ".o", b + ".o", sort=False, standardize=False)  # kept as it is
    assert st["n_pairs"] == 2 and st["n_no_tag"] == 2 and open(a + ".o").read() == open(a).read()
    open(b, "w").write("@n1/2\nACGT\n+\nIIII\n")
    with pytest.raises(L.ArachneError):
        L.preprocess(a, b, a + ".o", b + ".o", sort=False, standardize=False)  # 2 records against 1
    with pytest.raises(L.ArachneError):
        L.fastq_format(os.path.join(d, "missing.fq"), b)
    open(a, "w").write("n1/1\nACGT\n+\nIIII\n")
    with pytest.raises(L.ArachneError):
        L.fastq_format(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["standard", "haplotagging", "stlfr", "tellseq"])
def test_preprocess_sorted_matches_restatement(tmp_path, fmt):
    d = str(tmp_path)
    r1, r2 = make_fastq(d, fmt, n=3000, seed=11)
    o1, o2 = os.path.join(d, "BX_sort.R1.fq"), os.path.join(d, "BX_sort.R2.fq")
    st = L.preprocess(r1, r2, o1, o2)
    e1, e2, est = O.preprocess(r1, r2)
    assert open(o1, "rb").read() == e1 and open(o2, "rb").read() == e2
    for k in ("n_pairs", "format", "n_barcodes", "n_no_tag", "n_invalid"):
        assert st[k] == est[k], k
    # properties of `samtools sort -t BX`: same multiset of records, tag order non-decreasing with untagged first, ties in input order
    inp = {O.split_header(h)[0]: (s, q) for h, s, q in O.read_fastq(r1)}
    keys, idx = [], []
    for h, s, q in O.read_fastq(o1):
        name = O.split_header(h)[0]
        assert inp[name] == (s, q)
        keys.append(O.passthrough(h, "BX")[2].encode()); idx.append(int(name[1:6]))
    assert len(keys) == 3000 and keys == sorted(keys)
    assert all(idx[i] < idx[i + 1] for i in range(len(idx) - 1) if keys[i] == keys[i + 1])
    # the reader cuts the result into one work unit per barcode
    from arachne_b200 import Ingest
    ing = Ingest(o1, o2)
    units = []
    while True:
        b = ing.next(1 << 20)
        if b is None:
            break
        units += b["barcodes"]
    tagged = [k.decode() for k in keys if k]
    assert [u for u in units if u] == sorted(set(tagged), key=lambda x: x.encode())


@pytest.mark.gpu
def test_preprocess_other_tag_and_long_key(tmp_path):
    d = str(tmp_path)
    r1, r2 = make_fastq(d, "standard", n=500, seed=3)
    o1, o2 = os.path.join(d, "o1.fq"), os.path.join(d, "o2.fq")
    L.preprocess(r1, r2, o1, o2, tag="RG")
    e1, e2, _ = O.preprocess(r1, r2, tag="RG")
    assert open(o1, "rb").read() == e1 and open(o2, "rb").read() == e2
    a = os.path.join(d, "long.R1.fq"); b = os.path.join(d, "long.R2.fq")
    open(a, "w").write("@n/1\tBX:Z:%s\tVX:i:1\nAC\n+\nII\n" % ("q" * 33)); open(b, "w").write("@n/2\tBX:Z:%s\tVX:i:1\nAC\n+\nII\n" % ("q" * 33))
    with pytest.raises(L.ArachneError):
        L.preprocess(a, b, o1, o2)
