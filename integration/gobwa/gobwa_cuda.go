// gobwa over libarachne_cuda: a drop-in replacement for src/gobwa/gobwa.go of pdimens/arachne.
//
// Install: copy this file to src/gobwa/gobwa_cuda.go and delete (or build-tag away) src/gobwa/gobwa.go, bwa_bridge.c,
// bwa_bridge.h and the vendored bwa/ directory.  Nothing outside src/gobwa changes: the exported API below is the one
// src/aligner uses today (aligner.go:305,307,452,475,500,1500,1519,1638; bamwriter.go:56,141), with the same argument
// meaning, the same result structs and the same lifetime rules (results of a barcode live until arena.Free()).
//
// What is different underneath: the per-pair GoBwaMemMateSW calls that the worker goroutines make concurrently
// (aligner.go:335-337, -threads) are gathered by the library into GPU batches (include/arachne_compat.h), so run the host
// with MANY workers (-threads 2048 or more: a goroutine blocked in a call is cheap) to give the GPU large batches.
// GoBwaSmithWaterman is a lookup: the pair's batch already produced the CIGAR of every region it returned.
//
// NOT COMPILED in the image this was written in (no Go toolchain there, `go version`: not found).  The C functions it
// binds are exercised with the same arguments by tests/test_gpu_parity.py::test_compat_layer_per_pair_calls_batched
// through ctypes (arachne_b200/lib.py::Compat).

package gobwa

// #cgo CFLAGS: -I${SRCDIR}/../../third_party/arachne_b200/include
// #cgo LDFLAGS: -L${SRCDIR}/../../third_party/arachne_b200/arachne_b200 -larachne_cuda -lcudart -lstdc++
// #include <stdlib.h>
// #include "arachne_compat.h"
import "C"

import (
	"log"
	"os"
	"strconv"
	"unsafe"
)

// GoBwaReference holds a loaded reference (gobwa.go:19-23).  BWTData keeps its name and its opaqueness; it now points at
// the library's index handle instead of a bwaidx_t.
type GoBwaReference struct {
	BWTData    unsafe.Pointer // *C.arachne_index_t
	contigTids map[string]int32
	names      []string
	offsets    []int64
	lengths    []int64
	lPac       int64
	batcher    *C.arachne_compat_t
}

// GoBwaSettings wraps the aligner options (gobwa.go:44-46): a mem_opt_t-shaped struct with mem_opt_init's defaults.
type GoBwaSettings struct {
	Settings unsafe.Pointer // *C.arachne_mem_opt_t
}

// EasyAlignment is a candidate region (gobwa.go:88-98).  ChainedHit points into the pair's region array, which stays valid
// and unmoved until arena.Free(); aligner.go hands it back to GoBwaSmithWaterman as an unsafe.Pointer (aligner.go:1653,1500).
type EasyAlignment struct {
	Offset        int64
	Alignment_end int64
	Contig        string
	Reversed      bool
	ChainedHit    *C.arachne_alnreg_t
	Score         int
	Secondary     bool
	ReadS         int
	ReadE         int
}

// Arena collects what a barcode's calls allocated (gobwa.go:107-125).  Here an entry is a pair's result block.
type Arena struct {
	Pointers []uintptr
	batcher  *C.arachne_compat_t
}

func (a *Arena) Push(p uintptr) { a.Pointers = append(a.Pointers, p) }

func (a *Arena) Free() {
	for i := 0; i < len(a.Pointers); i++ {
		C.arachne_compat_release(a.batcher, unsafe.Pointer(a.Pointers[i]))
	}
	a.Pointers = a.Pointers[0:0]
}

func NewArena() *Arena {
	a := new(Arena)
	a.Pointers = make([]uintptr, 0, 0)
	return a
}

func envInt(name string, def int) int {
	if v, err := strconv.Atoi(os.Getenv(name)); err == nil {
		return v
	}
	return def
}

// GoBwaLoadReference loads <path>.bwt/.sa/.pac/.ann/.amb into the HBM of one GPU (gobwa.go:128-147).  Returns nil and logs
// on failure, like the original.  ARACHNE_DEVICE picks the GPU (default 0); ARACHNE_BATCH_PAIRS / ARACHNE_BATCH_WAIT_US tune
// the batcher (defaults 4096 pairs / 200 us).
func GoBwaLoadReference(path string) *GoBwaReference {
	cs := C.CString(path)
	defer C.free(unsafe.Pointer(cs))
	var idx *C.arachne_index_t
	if C.arachne_index_load(cs, C.int(envInt("ARACHNE_DEVICE", 0)), &idx) != C.ARACHNE_OK {
		log.Printf("THAT DIDN'T WORK!!! FILE PROBLEMS? %s", C.GoString(C.arachne_last_error()))
		return nil
	}
	r := &GoBwaReference{BWTData: unsafe.Pointer(idx), contigTids: map[string]int32{}}
	r.lPac = int64(C.arachne_index_l_pac(idx))
	n := int(C.arachne_index_n_seqs(idx))
	for i := 0; i < n; i++ {
		var name *C.char
		var off C.int64_t
		var ln C.int32_t
		C.arachne_index_contig(idx, C.int32_t(i), &name, &off, &ln)
		r.names = append(r.names, C.GoString(name))
		r.offsets = append(r.offsets, int64(off))
		r.lengths = append(r.lengths, int64(ln))
		r.contigTids[r.names[i]] = int32(i)
	}
	if C.arachne_compat_open(idx, nil, C.int(envInt("ARACHNE_BATCH_PAIRS", 0)), C.int(envInt("ARACHNE_BATCH_WAIT_US", 0)), &r.batcher) != C.ARACHNE_OK {
		log.Printf("arachne_compat_open: %s", C.GoString(C.arachne_last_error()))
		return nil
	}
	return r
}

// GetReferenceContigsInfo: contig names and lengths (gobwa.go:26-39).
func (r GoBwaReference) GetReferenceContigsInfo() ([]string, []int64) {
	return append([]string{}, r.names...), append([]int64{}, r.lengths...)
}

// EnumerateContigs calls back once per contig in index order (gobwa.go:436-446).
func EnumerateContigs(ref *GoBwaReference, callback func(name string, length int)) {
	for i := range ref.names {
		callback(ref.names[i], int(ref.lengths[i]))
	}
}

func GoBwaAllocSettings() *GoBwaSettings {
	opt := (*C.arachne_mem_opt_t)(C.malloc(C.size_t(unsafe.Sizeof(C.arachne_mem_opt_t{}))))
	C.arachne_opts_default(opt)
	return &GoBwaSettings{Settings: unsafe.Pointer(opt)}
}

var nt4 = func() (t [256]byte) { // nst_nt4_table (bwa/bntseq.c:47-64): A/a C/c G/g T/t -> 0..3, '-' -> 5, everything else 4
	for i := range t {
		t[i] = 4
	}
	t['-'] = 5
	for i, c := range "ACGT" {
		t[byte(c)] = byte(i)
		t[byte(c)+32] = byte(i)
	}
	return
}()

// SequenceConvert: ASCII bases -> codes 0..4 (gobwa.go:159-167).
func SequenceConvert(seq string) []byte {
	buffer := make([]byte, len(seq))
	for i := 0; i < len(seq); i++ {
		buffer[i] = nt4[seq[i]]
	}
	return buffer
}

var twoBitToSeq = [4]byte{'A', 'C', 'G', 'T'}
var twoBitToSeqComp = [4]byte{'T', 'G', 'C', 'A'}

// GetSeq: reference bases of [start, end) on a contig, reverse-complemented when reversed (gobwa.go:50-80).
func (r GoBwaReference) GetSeq(chrom string, start, end int64, reversed bool) []byte {
	tid := r.contigTids[chrom]
	offstart := C.int64_t(start + r.offsets[tid])
	offend := C.int64_t(end + r.offsets[tid])
	var rid C.int
	res := C.arachne_compat_fetch_seq((*C.arachne_index_t)(r.BWTData), &offstart, (offstart+offend)>>1, &offend, &rid)
	str := make([]byte, int(end-start))
	if res == nil {
		log.Printf("GetSeq: %s", C.GoString(C.arachne_last_error()))
		return str
	}
	n := int(offend - offstart)
	bases := unsafe.Slice((*byte)(unsafe.Pointer(res)), n)
	for i := 0; i < n; i++ {
		if reversed {
			str[n-i-1] = twoBitToSeqComp[bases[i]]
		} else {
			str[i] = twoBitToSeq[bases[i]]
		}
	}
	C.free(unsafe.Pointer(res))
	return str
}

// InterpretAlign: region -> contig coordinates (gobwa.go:339-371).
func InterpretAlign(ref *GoBwaReference, caln *C.arachne_alnreg_t) EasyAlignment {
	var res EasyAlignment
	off := ref.offsets[int(caln.rid)]
	if int64(caln.rb) < ref.lPac {
		res.Offset = int64(caln.rb) - off
		res.Reversed = false
	} else {
		res.Offset = ref.lPac*2 - 1 - int64(caln.rb) - off
		res.Reversed = true
	}
	if int64(caln.re) < ref.lPac {
		res.Alignment_end = int64(caln.re) - off
	} else {
		res.Alignment_end = ref.lPac*2 - 1 - int64(caln.re) - off
	}
	res.Contig = ref.names[int(caln.rid)]
	res.Secondary = int(caln.secondary) >= 0 || int(caln.secondary_all) > 0
	res.ChainedHit = caln
	res.Score = int(caln.score)
	res.ReadS = int(caln.qb)
	res.ReadE = int(caln.qe)
	return res
}

func interpretList(ref *GoBwaReference, v C.arachne_alnreg_v) []EasyAlignment {
	out := make([]EasyAlignment, int(v.n))
	for i := uintptr(0); i < uintptr(v.n); i++ {
		a := (*C.arachne_alnreg_t)(unsafe.Pointer(uintptr(unsafe.Pointer(v.a)) + i*unsafe.Sizeof(*v.a)))
		out[i] = InterpretAlign(ref, a)
	}
	return out
}

// GoBwaMemMateSW: candidate regions of both reads of a pair after mate rescue (gobwa.go:226-337: mem_align1_core x2, then
// mem_matesw from every hit within score_delta of the mate's best, read 1 first).  Blocks until the GPU batch that this call
// joined has run.
func GoBwaMemMateSW(ref *GoBwaReference, settings *GoBwaSettings, read1 *[]byte, read2 *[]byte, arena *Arena, score_delta int) ([]EasyAlignment, []EasyAlignment) {
	c1 := SequenceConvert(string(*read1))
	c2 := SequenceConvert(string(*read2))
	var p1, p2 *C.uint8_t
	if len(c1) > 0 {
		p1 = (*C.uint8_t)(unsafe.Pointer(&c1[0]))
	}
	if len(c2) > 0 {
		p2 = (*C.uint8_t)(unsafe.Pointer(&c2[0]))
	}
	var v1, v2 C.arachne_alnreg_v
	var owner unsafe.Pointer
	if rc := C.arachne_compat_mem_mate_sw(ref.batcher, p1, C.int(len(c1)), p2, C.int(len(c2)), C.int(score_delta), &v1, &v2, &owner); rc != C.ARACHNE_OK {
		panic("arachne_compat_mem_mate_sw: " + C.GoString(C.arachne_last_error())) // the reference aborts inside BWA here
	}
	arena.batcher = ref.batcher
	arena.Push(uintptr(owner))
	return interpretList(ref, v1), interpretList(ref, v2)
}

// GoBwaAlign: single-end candidates of one read (gobwa.go:174-204; unused by src/aligner).
func GoBwaAlign(ref *GoBwaReference, settings *GoBwaSettings, seq string, arena *Arena) []EasyAlignment {
	r1 := []byte(seq)
	r2 := []byte{}
	a, _ := GoBwaMemMateSW(ref, settings, &r1, &r2, arena, 0)
	return a
}

type SingleReadAlignment struct {
	Pos                 int64
	Chrom               string
	Flag                int
	Reversed            bool
	Alt                 int
	Mapq                int
	EditDistance        int
	Cigar               []uint32
	AlternativeMappings string
	Score               int
	Sub                 int
	AltSC               int
	ReadS               int
	ReadE               int
}

// GoBwaSmithWaterman: base-level alignment of a read under one of its regions (gobwa.go:400-415 -> mem_reg2aln).  `alignment`
// is the ChainedHit of an EasyAlignment whose arena has not been freed.
func GoBwaSmithWaterman(ref *GoBwaReference, settings *GoBwaSettings, seq string, alignment unsafe.Pointer, arena *Arena) SingleReadAlignment {
	var g C.arachne_go_aln_t
	if rc := C.arachne_compat_reg2aln(ref.batcher, (*C.arachne_alnreg_t)(alignment), &g); rc != C.ARACHNE_OK {
		panic("arachne_compat_reg2aln: " + C.GoString(C.arachne_last_error()))
	}
	return InterpretSingleReadAlignment(ref, &g)
}

// InterpretSingleReadAlignment: gobwa.go:449-488, including its bit arithmetic on the flag word.
func InterpretSingleReadAlignment(ref *GoBwaReference, alignment *C.arachne_go_aln_t) SingleReadAlignment {
	var result SingleReadAlignment
	fixed_flags := alignment.flag2
	result.Pos = int64(alignment.pos)
	result.Chrom = ref.names[int(alignment.rid)]
	cigar_ops := uintptr(alignment.n_cigar)
	result.Cigar = make([]uint32, cigar_ops*2)
	for i := uintptr(0); i < cigar_ops; i++ {
		raw := *(*uint32)(unsafe.Pointer(uintptr(unsafe.Pointer(alignment.cigar)) + i*unsafe.Sizeof(C.uint32_t(0))))
		result.Cigar[i*2] = raw & 0xf
		result.Cigar[i*2+1] = raw >> 4
	}
	result.Alt = int(fixed_flags & 0x2)
	result.Mapq = int(fixed_flags&0x2C) >> 2
	result.Reversed = (uint32(0) != uint32(uint32(fixed_flags)&uint32(0x1)))
	result.Score = int(alignment.score)
	result.Sub = int(alignment.sub)
	result.EditDistance = int(uint(fixed_flags) >> 10)
	result.AltSC = int(alignment.alt_sc)
	return result
}
