// block_hash_core.h -- the reference's block_hash (src/model/block_hash.cpp:29-54)
// over blocks given as flat fragment arrays, written once for the device
// kernel (block_hash_kernels.cu) and for the host harness of the CPU tests
// (tests/block_hash_host.cpp compiles this header with g++ and checks it against
// the oracle; the product library only ever runs it on the device).
//
// block_hash(block): ids of the fragments ("name_begin_last",
// Fragment::id, src/model/Fragment.cpp:173-183), once as they are and once with
// every fragment inverted; both lists sorted as strings; the smaller list
// joined with ' ', padded with ' ' to a multiple of 16 bytes and folded
// a = 1; a *= word[2i]; a ^= word[2i+1] over little-endian 64-bit words.
//
// No strings are built.  An id is ordered by a 128-bit key
//   (rank of name + "_" among the names, begin as nibbles, last as nibbles)
// whose unsigned order is the string order of the id (see make_key), and the
// bytes of the joint string are regenerated from the sorted keys on the fly.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define NPGE_HD __host__ __device__ __forceinline__
#else
#define NPGE_HD inline
#endif

namespace npge_b200 {

struct IdKey {
    uint64_t hi, lo;
};

NPGE_HD bool key_less(const IdKey& a, const IdKey& b) {
    return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}

// Decimal digits of v, most significant first, one per nibble, left-aligned in
// 44 bits (11 nibbles: 10 digits of a 32-bit value + one terminator).
// digit d -> nibble d + add; the nibble after the last digit = term.
NPGE_HD uint64_t decimal_nibbles(uint32_t v, uint32_t add, uint32_t term) {
    uint64_t digits = 0;  // least significant digit in the low nibble
    int n = 0;
    do {
        digits |= uint64_t(v % 10u + add) << (4 * n);
        v /= 10u;
        n++;
    } while (v);
    // reverse into most-significant-first, left-aligned
    uint64_t out = 0;
    for (int i = 0; i < n; i++) {
        out |= ((digits >> (4 * (n - 1 - i))) & 0xf) << (40 - 4 * i);
    }
    if (n < 11) out |= uint64_t(term) << (40 - 4 * n);
    return out;
}

// Key of the id "name_begin_last".
//  * names: `rank` orders name + "_" as a string (host: name_ranks()).
//  * begin is followed by '_', which is greater than any digit: digits 0..9,
//    terminator 0xF, so "12_" > "123".
//  * last is followed by the end of the string (or the ' ' of the join), which
//    is smaller than anything: '-' 1, digits 2..11, terminator 0; last = -1
//    ("a == b && ori == -1", a one-letter reverse fragment) is the string "-1".
NPGE_HD IdKey make_key(uint32_t rank, uint32_t begin, uint32_t last, bool last_is_minus1) {
    const uint64_t b = decimal_nibbles(begin, 0, 0xf);
    const uint64_t l = last_is_minus1 ? ((uint64_t(1) << 40) | (uint64_t(3) << 36))
                                      : decimal_nibbles(last, 2, 0);
    IdKey k;
    k.hi = (uint64_t(rank) << 32) | (b >> 12);
    k.lo = ((b & 0xfff) << 52) | (l << 8);
    return k;
}

// Keys of one fragment as it is (dir) and inverted (inv).
NPGE_HD void fragment_keys(uint32_t rank, uint32_t min_pos, uint32_t max_pos, int ori,
                           IdKey& dir, IdKey& inv) {
    const bool one = min_pos == max_pos;
    const uint32_t a = ori == 1 ? min_pos : max_pos, b = ori == 1 ? max_pos : min_pos;
    dir = make_key(rank, a, b, one && ori == -1);
    inv = make_key(rank, b, a, one && ori != -1);
}

// In-place ascending sort; ids of one block are distinct, so stability is moot.
NPGE_HD void sort_keys(IdKey* a, uint64_t n) {
    if (n <= 24) {
        for (uint64_t i = 1; i < n; i++) {
            const IdKey x = a[i];
            uint64_t j = i;
            while (j > 0 && key_less(x, a[j - 1])) {
                a[j] = a[j - 1];
                j--;
            }
            a[j] = x;
        }
        return;
    }
    // heap sort: O(n log n) whatever the block looks like, no extra memory
    for (uint64_t start = n / 2; start-- > 0;) {
        uint64_t root = start;
        const IdKey x = a[root];
        while (true) {
            uint64_t c = 2 * root + 1;
            if (c >= n) break;
            if (c + 1 < n && key_less(a[c], a[c + 1])) c++;
            if (!key_less(x, a[c])) break;
            a[root] = a[c];
            root = c;
        }
        a[root] = x;
    }
    for (uint64_t end = n - 1; end > 0; end--) {
        const IdKey x = a[end];
        a[end] = a[0];
        uint64_t root = 0;
        while (true) {
            uint64_t c = 2 * root + 1;
            if (c >= end) break;
            if (c + 1 < end && key_less(a[c], a[c + 1])) c++;
            if (!key_less(x, a[c])) break;
            a[root] = a[c];
            root = c;
        }
        a[root] = x;
    }
}

// ids_dir < ids_inv of two sorted lists of the same length (std::vector<
// std::string>::operator<): decided by the first position that differs.
NPGE_HD bool list_less(const IdKey* a, const IdKey* b, uint64_t n) {
    for (uint64_t i = 0; i < n; i++) {
        if (key_less(a[i], b[i])) return true;
        if (key_less(b[i], a[i])) return false;
    }
    return false;
}

// The fold of block_hash.cpp:41-53 fed one byte at a time.
struct HashFold {
    uint64_t a = 1, w0 = 0, w1 = 0;
    uint32_t n = 0;  // bytes in the current 16-byte group
    NPGE_HD void put(uint32_t c) {
        if (n < 8) {
            w0 |= uint64_t(c) << (8 * n);
        } else {
            w1 |= uint64_t(c) << (8 * (n - 8));
        }
        if (++n == 16) {
            a *= w0;
            a ^= w1;
            w0 = w1 = 0;
            n = 0;
        }
    }
    NPGE_HD uint64_t finish() {
        while (n) put(' ');  // joint.resize(new_size, ' ')
        return a;
    }
};

// names of the ranks: bytes name_bytes[name_off[r] .. name_off[r + 1])
NPGE_HD uint64_t fold_sorted(const IdKey* a, uint64_t n, const uint8_t* name_bytes,
                             const uint32_t* name_off) {
    HashFold h;
    for (uint64_t i = 0; i < n; i++) {
        if (i) h.put(' ');
        const uint32_t rank = uint32_t(a[i].hi >> 32);
        for (uint32_t p = name_off[rank]; p < name_off[rank + 1]; p++) h.put(name_bytes[p]);
        h.put('_');
        const uint64_t b = ((a[i].hi & 0xffffffffull) << 12) | (a[i].lo >> 52);
        for (int s = 40; s >= 0; s -= 4) {
            const uint32_t d = uint32_t(b >> s) & 0xf;
            if (d == 0xf) break;
            h.put('0' + d);
        }
        h.put('_');
        const uint64_t l = (a[i].lo >> 8) & 0xfffffffffffull;
        for (int s = 40; s >= 0; s -= 4) {
            const uint32_t d = uint32_t(l >> s) & 0xf;
            if (d == 0) break;
            h.put(d == 1 ? uint32_t('-') : '0' + d - 2);
        }
    }
    return h.finish();
}

// block_hash of the block whose fragments are [f0, f1) of the flat arrays.
// max_pos == nullptr: every fragment is `frag_len` long (anchors).
// scratch: 2 * (f1 - f0) keys.
NPGE_HD uint64_t block_hash_flat(uint64_t f0, uint64_t f1, const uint32_t* frag_seq,
                                 const uint32_t* min_pos, const uint32_t* max_pos,
                                 uint32_t frag_len, const int8_t* frag_ori,
                                 const uint32_t* seq_rank, const uint8_t* name_bytes,
                                 const uint32_t* name_off, IdKey* scratch) {
    const uint64_t n = f1 - f0;
    IdKey* dir = scratch;
    IdKey* inv = scratch + n;
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t f = f0 + i;
        const uint32_t lo = min_pos[f];
        const uint32_t hi = max_pos ? max_pos[f] : lo + frag_len - 1;
        fragment_keys(seq_rank[frag_seq[f]], lo, hi, frag_ori[f], dir[i], inv[i]);
    }
    sort_keys(dir, n);
    sort_keys(inv, n);
    const IdKey* ids = list_less(dir, inv, n) ? dir : inv;
    return fold_sorted(ids, n, name_bytes, name_off);
}

}  // namespace npge_b200
