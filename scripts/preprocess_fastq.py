#!/usr/bin/env python
"""`arachne preprocess <options> sample.R1.fq sample.R2.fq` (src/preprocess/preprocess.go:112-162) on libarachne_cuda: brings the
pair of FASTQ files to the standard header form and sorts the pairs by barcode on the GPU.  Writes <TAG>_sort.<basename> next to the
current directory, as the reference does (preprocess.go:44-45).  No samtools needed."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arachne_b200 import fastq_format, preprocess  # noqa: E402

ap = argparse.ArgumentParser(description=__doc__)
ap.add_argument("--tag", default="BX", help="which SAM tag has the barcode to sort by (default: BX)")
ap.add_argument("--threads", type=int, default=4, help="accepted for compatibility; the sort runs on the GPU")
ap.add_argument("--no-standardize", action="store_true", help="keep the headers as they are (the reference's compilable behaviour)")
ap.add_argument("--device", type=int, default=0)
ap.add_argument("r1"); ap.add_argument("r2")
a = ap.parse_args()
for p in (a.r1, a.r2):
    if not os.access(p, os.R_OK):
        sys.exit("Error: FASTQ file %s does not exist or does not have read permissions." % p)
o1, o2 = a.tag + "_sort." + os.path.basename(a.r1), a.tag + "_sort." + os.path.basename(a.r2)
print("format:", fastq_format(a.r1, a.r2), file=sys.stderr)
st = preprocess(a.r1, a.r2, o1, o2, tag=a.tag, standardize=not a.no_standardize, device=a.device)
print("%d pairs, %d barcodes, %d without %s, %d invalid -> %s %s (read %.0f ms, sort %.0f ms, write %.0f ms)" %
      (st["n_pairs"], st["n_barcodes"], st["n_no_tag"], a.tag, st["n_invalid"], o1, o2, st["ms_read"], st["ms_sort"], st["ms_write"]), file=sys.stderr)
