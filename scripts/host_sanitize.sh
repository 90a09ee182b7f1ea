#!/usr/bin/env bash
# compute-sanitizer is refused by the GPU pool (profiles/r2a_sanitizer_unavailable.txt).  What CAN be sanitised is everything that compiles for
# the host: the thread-sequential device logic that tests/hostsim builds from csrc/*.cuh (chaining, B-tree, extension driver, de-duplication /
# patching, mate rescue, CIGAR / NM walk, introsort, the index-construction driver) and the multi-threaded FASTQ ingest.  This script builds both
# with AddressSanitizer + UndefinedBehaviorSanitizer (any finding aborts), the ingest also with ThreadSanitizer, and runs the CPU tests on them.
#   bash scripts/host_sanitize.sh > profiles/r2z_host_sanitizers.txt 2>&1
set -u
cd "$(dirname "$0")/.."
ROOT=$PWD; W=${TMPDIR:-/tmp}/arachne_sanitize; mkdir -p "$W"
ASAN=$(gcc -print-file-name=libasan.so); TSAN=$(gcc -print-file-name=libtsan.so)
make -s -C oracle restate
echo "== build: hostsim with -fsanitize=address,undefined -fno-sanitize-recover=undefined"
( cd tests/hostsim && g++ -O1 -g -std=c++17 -fPIC -fno-extern-tls-init -shared -fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer -x c++ \
    hostsim.cpp hostsim_index.cpp hostsim_seed.cpp hostsim_cigar.cpp hostsim_extdp.cpp hostsim_rescue.cpp hostsim_place.cpp hostsim_stages.cpp -o "$W/libhostsim_asan.so" -L../../oracle -loracle -Wl,-rpath,"$ROOT/oracle" -lm ) || exit 1
echo "== run: hostsim tests (goldens, live against the reference's C code, chain level, introsort permutations, index build)"
ASAN_OPTIONS=detect_leaks=0 UBSAN_OPTIONS=print_stacktrace=1 LD_PRELOAD=$ASAN ARACHNE_HOSTSIM_LIB="$W/libhostsim_asan.so" \
    python -m pytest tests/test_cpu_tier.py -q -k "hostsim" -p no:cacheprovider 2>&1 | grep -v "site-packages\|<frozen" | tail -15
cat > "$W/stub.cpp" <<'EOS'
#include <cstdarg>
#include <cstdio>
namespace arachne { static thread_local char g_err[512]; void set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap); } }
extern "C" const char* arachne_last_error() { return arachne::g_err; }
EOS
cat > "$W/run_ingest_tests.py" <<EOS
import os, sys, pathlib, tempfile
sys.path.insert(0, "$ROOT")
import tests.test_cpu_tier as T
from arachne_b200 import lib
orig = lib.Ingest.__init__
def patched(self, r1, r2, n_threads=4, path=None): orig(self, r1, r2, n_threads, os.environ["INGLIB"])
lib.Ingest.__init__ = patched
for gz in (False, True, "bgzf", "bgzf_small", "members"):
    T.test_ingest_matches_reader_go(pathlib.Path(tempfile.mkdtemp()), gz); print("ingest ok:", gz, flush=True)
T.test_ingest_bad_inputs_fail_loudly(pathlib.Path(tempfile.mkdtemp())); print("ingest bad inputs ok", flush=True)
EOS
for san in address,undefined thread; do
    echo "== build + run: FASTQ ingest (csrc/ingest.cu) with -fsanitize=$san"
    extra=""; [ "$san" = "address,undefined" ] && extra="-fno-sanitize-recover=undefined"
    g++ -O1 -g -std=c++17 -shared -fPIC -fsanitize=$san $extra -x c++ arachne_b200/csrc/ingest.cu "$W/stub.cpp" -o "$W/libingest_san.so" -lz -lpthread || exit 1
    if [ "$san" = "thread" ]; then pre=$TSAN; else pre=$ASAN; fi
    INGLIB="$W/libingest_san.so" ASAN_OPTIONS=detect_leaks=0 TSAN_OPTIONS="report_signal_unsafe=0 exitcode=0" LD_PRELOAD=$pre python "$W/run_ingest_tests.py" > "$W/ingest_$san.log" 2>&1
    echo "exit code $?"; grep "ingest ok\|ingest bad" "$W/ingest_$san.log"
    echo "sanitizer reports: $(grep -c 'WARNING: ThreadSanitizer\|ERROR: AddressSanitizer\|runtime error' "$W/ingest_$san.log")"
done
echo "== build + run: the whole library with its HOST code under ASan + UBSan (nvcc -Xcompiler -fsanitize=...), CPU tier against it"
nvcc -gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -shared -Xcompiler -fPIC,-fsanitize=address,-fsanitize=undefined,-fno-sanitize-recover=undefined,-fno-omit-frame-pointer \
    -Iinclude -o "$W/libarachne_asan.so" arachne_b200/csrc/*.cu -lz -lasan -lubsan || exit 1
ARACHNE_LIB="$W/libarachne_asan.so" ASAN_OPTIONS=detect_leaks=0:protect_shadow_gap=0 UBSAN_OPTIONS=print_stacktrace=1 LD_PRELOAD=$ASAN \
    python -m pytest tests/ -q -m "not gpu" -p no:cacheprovider -k "not gloo" 2>&1 | grep -v "site-packages\|<frozen" | tail -8
echo "== build + run: the gobwa-shaped layer (csrc/compat.cu) over the mock batch backend with -fsanitize=thread, 64 caller threads"
( cd tests/hostsim && g++ -O1 -g -std=c++17 -fPIC -shared -fsanitize=thread -Wno-subobject-linkage -I/usr/local/cuda/include -x c++ compat_mock.cpp \
    -o "$W/libcompat_mock_tsan.so" -L../../oracle/_ref -lbwa_ref -Wl,-rpath,"$ROOT/oracle/_ref" -lpthread ) || exit 1
ARACHNE_COMPAT_MOCK_LIB="$W/libcompat_mock_tsan.so" TSAN_OPTIONS="report_signal_unsafe=0 exitcode=0" LD_PRELOAD=$TSAN \
    python -m pytest tests/test_cpu_tier.py -q -k "compat_layer" -p no:cacheprovider > "$W/compat_tsan.log" 2>&1
tail -1 "$W/compat_tsan.log"; echo "sanitizer reports: $(grep -c 'WARNING: ThreadSanitizer' "$W/compat_tsan.log")"
echo "== build + run: the whole library with its HOST code under ThreadSanitizer; preprocess / ingest / pack / ABI tests of the CPU tier against it"
nvcc -gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -shared -Xcompiler -fPIC,-fsanitize=thread -Iinclude -o "$W/libarachne_tsan.so" arachne_b200/csrc/*.cu -lz -ltsan || exit 1
ARACHNE_LIB="$W/libarachne_tsan.so" TSAN_OPTIONS="report_signal_unsafe=0 exitcode=0" LD_PRELOAD=$TSAN python -m pytest tests/test_preprocess.py tests/test_cpu_tier.py -q -m "not gpu" \
    -p no:cacheprovider -k "preprocess or standardize or ingest or pack or abi or header" > "$W/lib_tsan.log" 2>&1
tail -1 "$W/lib_tsan.log"; echo "sanitizer reports: $(grep -c 'WARNING: ThreadSanitizer' "$W/lib_tsan.log")"
echo "== build + run: the ORACLE itself (oracle/restate_ksw.c, restate_place.cpp: the checker of the GPU tests) under ASan + UBSan"
( cd oracle && g++ -O1 -g -fPIC -shared -fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer -x c restate_ksw.c -x c++ -std=c++17 restate_place.cpp \
    -o "$W/liboracle_asan.so" -lm -lpthread ) || exit 1
ARACHNE_ORACLE_LIB="$W/liboracle_asan.so" ASAN_OPTIONS=detect_leaks=0 UBSAN_OPTIONS=print_stacktrace=1 LD_PRELOAD=$ASAN \
    python -m pytest tests/test_cpu_tier.py -q -p no:cacheprovider -k "oracle or place or go_math" 2>&1 | grep -v "site-packages\|<frozen" | tail -3
