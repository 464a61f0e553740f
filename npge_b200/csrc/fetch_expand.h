// fetch_expand.h -- host side of the compact fetch (api.cu: fetch_compact): the
// 4-byte payloads the emit kernel leaves (position in the packed genome << 1 |
// strand) -> the (sequence, position, strand) arrays of npge_anchor_result.
// Plain C++ so that tests/fetch_expand_host.cpp can run it without a device.
#pragma once

#include <stdint.h>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace npge_b200 {

// Payloads [lo, hi) -> seq / pos / ori; the host twin of locate() in
// bucket_kernels.cu.
//   lut[q]    scan-order sequence holding position q << shift
//   start[s]  first position of scan-order sequence s, [S + 1] entries, the last
//             one above every position (pads belong to the sequence before them)
//   caller[s] index of that sequence as the caller passed it
// The arrays are written once and not read by this thread again: streaming
// stores keep them out of the cache and skip the read-for-ownership of every
// line (the expansion is bound by memory bandwidth: 13 bytes per fragment over
// all cores).  lo must be a multiple of 8 and ori + lo 8-byte aligned.
inline void expand_payloads(const uint32_t* pay, uint64_t lo, uint64_t hi, const uint32_t* lut,
                            int shift, const uint32_t* start, const uint32_t* caller,
                            uint32_t* seq, uint32_t* pos, int8_t* ori) {
    uint64_t f = lo;
#if defined(__x86_64__)
    for (; f + 8 <= hi; f += 8) {
        uint64_t oris = 0;
        for (int j = 0; j < 8; j++) {
            const uint32_t p = pay[f + j], g = p >> 1;
            uint32_t s = lut[g >> shift];
            while (start[s + 1] <= g) s++;
            _mm_stream_si32(reinterpret_cast<int*>(seq + f + j), int(caller[s]));
            _mm_stream_si32(reinterpret_cast<int*>(pos + f + j), int(g - start[s]));
            oris |= uint64_t((p & 1) ? 0x01u : 0xffu) << (8 * j);
        }
        _mm_stream_si64(reinterpret_cast<long long*>(ori + f), static_cast<long long>(oris));
    }
#endif
    for (; f < hi; f++) {
        const uint32_t p = pay[f], g = p >> 1;
        uint32_t s = lut[g >> shift];
        while (start[s + 1] <= g) s++;
        seq[f] = caller[s];
        pos[f] = g - start[s];
        ori[f] = (p & 1) ? 1 : -1;
    }
#if defined(__x86_64__)
    _mm_sfence();  // the streaming stores are visible before the piece counts as done
#endif
}

}  // namespace npge_b200
