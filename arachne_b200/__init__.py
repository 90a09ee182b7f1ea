"""arachne_b200: B200-native (sm_100a) candidate-alignment / read-placement hot path of Arachne, behind the
reference's Go<->C seam (src/gobwa).  The product is csrc/ -> libarachne_cuda.so (C ABI, include/arachne_cuda.h);
this package is the host-side mirror used by tests and bench.py."""
from .lib import ArachneCuda, ArachneError, ALNREG, Compat, Ingest, fastq_format, lib_path, preprocess  # noqa: F401
