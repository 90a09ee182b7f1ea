// Host helper: re-cut the payload of a `bwa index` .bwt file (64-byte blocks of 128 symbols, bwa/bwtindex.c:151-173)
// into the 32-byte sector blocks the kernels read (fmindex.cuh).  Used by the index loader and by the test-only host
// simulation.
#pragma once
#include <cstring>
#include <vector>
#include "fmindex.cuh"

namespace arachne {

// payload = file bytes after the 40-byte header; returns false when it is shorter than an occ-interleaved BWT of
// seq_len symbols must be.  image receives 2*ceil(seq_len/128)+2 sector blocks (zero padded).
inline bool bwt_recut_image(const uint8_t* payload, size_t payload_bytes, uint64_t seq_len, std::vector<uint8_t>& image) {
    // 32 bytes of counts per started 128-symbol block, 4 bytes per started 16-symbol word, one trailing 32-byte count record
    const size_t n_ref_blocks = (seq_len + 127) / 128;
    const size_t want = n_ref_blocks * 32 + ((seq_len + 15) / 16) * 4 + 32;
    if (payload_bytes < want) return false;
    const size_t body = want - 32;
    image.assign((2 * n_ref_blocks + 2) * sizeof(Blk32), 0);
    Blk32* dst = reinterpret_cast<Blk32*>(image.data());
    uint8_t last[64];
    for (size_t b = 0; b < n_ref_blocks; ++b) {
        const uint8_t* src = payload + b * 64;
        if ((b + 1) * 64 > body) {  // partial last block: the words that were never written read as zero
            memset(last, 0, 64);
            memcpy(last, src, body - b * 64);
            src = last;
        }
        uint32_t raw[16];
        memcpy(raw, src, 64);
        fm_recut_block(raw, dst + 2 * b);
    }
    return true;
}

}  // namespace arachne
