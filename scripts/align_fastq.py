#!/usr/bin/env python
"""FASTQ pair -> SAM through the whole path: native ingest (work units, name seeds) -> arachne_align_batch with every stage incl. record
assembly -> arachne_records_sam.  What `arachne -reads ... -genome ...` does, minus BAM encoding (the reference hands its records to
biogo/hts; here they are SAM text).

   python scripts/align_fastq.py <index prefix> <R1.fq[.gz]> <R2.fq[.gz]> [-o out.sam] [--batch-pairs 1000000] [--device 0] [--threads N]
   <index prefix>: FASTA path whose .bwt/.sa/.pac/.ann/.amb exist (make them with --index: arachne_index_pack + arachne_index_build)."""
import argparse, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arachne_b200 import ArachneCuda, Ingest


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("prefix"); ap.add_argument("r1"); ap.add_argument("r2")
    ap.add_argument("-o", "--out", default="-"); ap.add_argument("--batch-pairs", type=int, default=1_000_000)
    ap.add_argument("--device", type=int, default=0); ap.add_argument("--threads", type=int, default=os.cpu_count() or 4)
    ap.add_argument("--index", action="store_true", help="build the index files first (pack step + device BWT/SA)")
    ap.add_argument("--improper-penalty", type=float, default=-4.0)
    a = ap.parse_args()
    A = ArachneCuda()
    if a.index:
        A.pack_fasta(a.prefix, a.prefix)
        A.build_index(a.prefix, a.device, 0)
    A.load_index(a.prefix, device=a.device)
    out = sys.stdout if a.out == "-" else open(a.out, "w")
    out.write("@HD\tVN:1.3\tSO:unsorted\n")
    for name, off, ln in A.contigs():
        out.write("@SQ\tSN:%s\tLN:%d\n" % (name, ln))
    I = Ingest(a.r1, a.r2, n_threads=a.threads)
    n = 0; t0 = time.time()
    while True:
        b = I.next(a.batch_pairs)
        if b is None:
            break
        res = A.align_batch(b["seq"], b["off"], b["unit_off"], stages=31, run_rfa=b["run_rfa"], rand_seed=b["rand_seed"], attach_bx=b["unique"],
                            improper_penalty=a.improper_penalty,
                            sam=dict(names=b["names"], quals=b["quals"], barcodes=b["barcodes"], read_groups=b["read_groups"]))
        out.write(res["sam"])
        n += b["n_pairs"]
    I.close(); A.close()
    print("%d pairs in %.1f s" % (n, time.time() - t0), file=sys.stderr)


if __name__ == "__main__":
    main()
