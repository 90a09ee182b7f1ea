"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's `arachne preprocess` for the parity tests.

PARITY UNPINNED: the reference's src/preprocess/standardize.go does not compile and samtools, which
src/preprocess/preprocess.go:43-110 shells out to, is not in this image; there is no golden vector, fixture or runnable reference for
this row.  What is restated:
  * the five patterns, verbatim                                   standardize.go:16-25
  * findFastqFormat's if-chain over the first 200 records         standardize.go:98-125
  * the standard header `<name>/1\\tBX:Z:<bc>\\tVX:i:<v>`            standardize.go:73-95
  * the converters' barcode extraction                            standardize.go:128-186
  * samtools import -T '*' | sort -t TAG | fastq -T '*'           preprocess.go:47-55 (documented samtools behaviour: mate suffixes
    cut on import and written back by fastq, SAM-tag-shaped comment fields kept tab-separated, records ordered by the tag's value with
    strcmp, records without the tag first, ties in input order)
Choices where the broken source contradicts itself (also in DESIGN.md): VX:i:1 means "valid" for all three converters (the source
negates only the stLFR test); the patterns are applied to the header with the mate suffix cut from the name (the source applies them
to ParseHeader's name, which never ends in white space, so they could not match).
"""
import gzip
import re

WS = r"[\t\n\f\r ]"  # Go's \s
BX_RE = re.compile(r"BX:Z:([^\t\n\f\r ]+)" + WS)
VX_RE = re.compile(r"VX:i:([01])" + WS)
HAPLOTAGGING_RE = re.compile(r"BX:Z:(A[0-9]{2}C[0-9]{2}B[0-9]{2}D[0-9]{2})" + WS)
STLFR_RE = re.compile(r"#([0-9]+_[0-9]+_[0-9]+)" + WS)
STLFR_INVALID_RE = re.compile(r"^0_|_0_|_0$")
TELLSEQ_RE = re.compile(r":([ATCGN]+)" + WS)
SAM_TAG_RE = re.compile(r"^[A-Za-z][A-Za-z0-9]:[AifZHB]:")


def split_header(h):
    """header line without '@' / newline -> (name without mate suffix, rest of the line)"""
    m = re.match(r"[^ \t\n\r\v\f]*", h)
    name, rest = m.group(0), h[m.end():]
    if len(name) >= 2 and name[-2] == "/" and name[-1] in "12":
        name = name[:-2]
    return name, rest


def pattern_text(h):
    name, rest = split_header(h)
    return name + rest + "\n"


def header_format(h):
    t = pattern_text(h.rstrip("\r\n"))
    if BX_RE.search(t) and VX_RE.search(t):
        return "standard"
    if HAPLOTAGGING_RE.search(t):
        return "haplotagging"
    if STLFR_RE.search(t):
        return "stlfr"
    if TELLSEQ_RE.search(t):
        return "tellseq"
    return "unknown"


def read_fastq(path):
    op = gzip.open if open(path, "rb").read(2) == b"\x1f\x8b" else open
    with op(path, "rb") as f:
        lines = f.read().decode().split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    assert len(lines) % 4 == 0
    return [(lines[i][1:].rstrip("\r"), lines[i + 1].rstrip("\r"), lines[i + 3].rstrip("\r")) for i in range(0, len(lines), 4)]


def file_format(recs1):
    for h, _, _ in recs1[:200]:
        f = header_format(h)
        if f != "unknown":
            return f
    return "unknown"


def convert(h, fmt):
    """-> (name, tag string beginning with a tab, barcode, valid)"""
    name, _ = split_header(h)
    t = pattern_text(h)
    bc = ""
    if fmt == "haplotagging":
        m = HAPLOTAGGING_RE.search(t) or BX_RE.search(t)
        bc = m.group(1) if m else ""
        valid = bool(bc) and "00" not in bc
    elif fmt == "stlfr":
        m = STLFR_RE.search(t)
        bc = m.group(1) if m else ""
        valid = bool(bc) and not STLFR_INVALID_RE.search(bc)
    else:
        m = TELLSEQ_RE.search(t)
        bc = m.group(1) if m else ""
        valid = bool(bc) and "N" not in bc
    return name, "\tBX:Z:%s\tVX:i:%d" % (bc, 1 if valid else 0), bc, valid


def passthrough(h, tag):
    name, rest = split_header(h)
    fields = [x for x in re.split(r"[ \t\n\r\v\f]+", rest) if x and SAM_TAG_RE.match(x)]
    key = ""
    for x in fields:
        if x.startswith(tag + ":") and not key:
            key = x[5:]
    return name, "".join("\t" + x for x in fields), key


def preprocess(r1, r2, tag="BX", standardize=True, sort=True):
    """-> (bytes of the R1 output, bytes of the R2 output, stats dict)"""
    a, b = read_fastq(r1), read_fastq(r2)
    assert len(a) == len(b)
    fmt = file_format(a)
    conv = standardize and fmt != "standard"
    if conv and fmt == "unknown":
        raise ValueError("unknown format")
    rows, keys, n_invalid = [], [], 0
    for (h1, s1, q1), (h2, s2, q2) in zip(a, b):
        out = []
        for k, (h, s, q) in enumerate(((h1, s1, q1), (h2, s2, q2))):
            if conv:
                name, tags, bc, valid = convert(h, fmt)
                key = bc if tag == "BX" else ("1" if valid else "0") if tag == "VX" else ""
                if k == 0:
                    n_invalid += not valid
            else:
                name, tags, key = passthrough(h, tag)
                if k == 0:
                    n_invalid += any(x == "VX:i:0" for x in tags.split("\t"))
            if k == 0:
                keys.append(key.encode())
            out.append("@%s/%d%s\n%s\n+\n%s\n" % (name, k + 1, tags, s, q))
        rows.append(out)
    order = sorted(range(len(rows)), key=lambda i: keys[i]) if sort else list(range(len(rows)))
    o1 = "".join(rows[i][0] for i in order).encode()
    o2 = "".join(rows[i][1] for i in order).encode()
    return o1, o2, dict(n_pairs=len(rows), format=fmt, n_barcodes=len({k for k in keys if k}), n_no_tag=sum(1 for k in keys if not k), n_invalid=int(n_invalid))
