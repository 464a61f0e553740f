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
// bytes of the joint string are regenerated from the sorted keys on the fly,
// eight at a time where they come from a name.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define NPGE_HD __host__ __device__ __forceinline__
#else
#define NPGE_HD inline
#endif

namespace npge_b200 {

struct alignas(16) IdKey {
    uint64_t hi, lo;
};

NPGE_HD bool key_less(const IdKey& a, const IdKey& b) {
    return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}

NPGE_HD int leading_zeros(uint64_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
    return __clzll(static_cast<long long>(x));
#else
    return __builtin_clzll(x);
#endif
}

NPGE_HD uint64_t byte_swap(uint64_t x) {
#if defined(__CUDA_ARCH__)
    const uint32_t lo = uint32_t(x), hi = uint32_t(x >> 32);
    return (uint64_t(__byte_perm(lo, 0, 0x0123)) << 32) | __byte_perm(hi, 0, 0x0123);
#else
    return __builtin_bswap64(x);
#endif
}

// bcd4[v] = the four decimal digits of v < 10000, one per nibble (host:
// make_bcd4 in block_hash_names.h; 20 KB, lives in L1).  Ten digits of a 32-bit
// value = three lookups instead of ten divisions.
constexpr uint32_t kBcdEntries = 10000;
NPGE_HD uint64_t bcd10(uint32_t v, const uint16_t* bcd4) {
    const uint32_t hi = v / 100000000u, r = v - hi * 100000000u;
    const uint32_t mid = r / 10000u, lo = r - mid * 10000u;
    return (uint64_t(bcd4[hi]) << 32) | (uint64_t(bcd4[mid]) << 16) | bcd4[lo];
}

// A position printed in decimal, n digits, in the two encodings the key needs
// (digits most significant first, one per nibble, left-aligned in 44 bits =
// 10 digits + one terminator; nibbles after the terminator 0):
//  * begin is followed by '_', which is greater than any digit: digits 0..9,
//    terminator 0xF, so "12_" > "123".
//  * last is followed by the end of the string (or the ' ' of the join), which
//    is smaller than anything: '-' 1, digits 2..11, terminator 0; last = -1
//    ("a == b && ori == -1", a one-letter reverse fragment) is the string "-1".
struct DecimalField {
    uint64_t begin, last;
    uint32_t n;
};
NPGE_HD uint64_t last_digit_bias(uint32_t n) {  // 2 in each of the n digit nibbles
    return 0x22222222222ull & ~((uint64_t(1) << (44 - 4 * n)) - 1);
}
NPGE_HD DecimalField decimal_field(uint32_t v, const uint16_t* bcd4) {
    const uint64_t bcd = bcd10(v, bcd4);
    DecimalField f;
    f.n = bcd ? uint32_t(63 - leading_zeros(bcd)) / 4u + 1u : 1u;
    const uint64_t d = bcd << (44 - 4 * f.n);
    f.begin = d | (uint64_t(0xf) << (40 - 4 * f.n));
    f.last = d + last_digit_bias(f.n);
    return f;
}
constexpr uint64_t kLastMinus1 = (uint64_t(1) << 40) | (uint64_t(3) << 36);  // "-1"

// Key of the id "name_begin_last": `rank` orders name + "_" as a string (host:
// build_name_table), then the begin field, then the last field.  The low byte
// carries the digit counts (begin, last; 0 for "-1") for the fold; it cannot
// decide a comparison (equal fields above it = equal strings).
NPGE_HD IdKey make_key(uint32_t rank, uint64_t begin_field, uint32_t begin_n,
                       uint64_t last_field, uint32_t last_n) {
    IdKey k;
    k.hi = (uint64_t(rank) << 32) | (begin_field >> 12);
    k.lo = ((begin_field & 0xfff) << 52) | (last_field << 8) | (last_n << 4) | begin_n;
    return k;
}

// Keys of one fragment as it is (dir) and inverted (inv).
NPGE_HD void fragment_keys(uint32_t rank, uint32_t min_pos, uint32_t max_pos, int ori,
                           const uint16_t* bcd4, IdKey& dir, IdKey& inv) {
    const bool one = min_pos == max_pos;
    const DecimalField lo = decimal_field(min_pos, bcd4);
    const DecimalField hi = one ? lo : decimal_field(max_pos, bcd4);
    const DecimalField& a = ori == 1 ? lo : hi;  // begin of the fragment as it is
    const DecimalField& b = ori == 1 ? hi : lo;  // its last position
    const bool dir_m1 = one && ori == -1, inv_m1 = one && ori != -1;
    dir = make_key(rank, a.begin, a.n, dir_m1 ? kLastMinus1 : b.last, dir_m1 ? 0 : b.n);
    inv = make_key(rank, b.begin, b.n, inv_m1 ? kLastMinus1 : a.last, inv_m1 ? 0 : a.n);
}

// n digits held most significant first in the top nibbles of a 44-bit field ->
// ASCII, first digit in the low byte: chunk[0] the first 8, chunk[1] the rest.
NPGE_HD uint64_t nibbles_to_ascii(uint32_t x) {  // 8 nibbles, the top one first
    uint64_t t = x;
    t = (t | (t << 16)) & 0x0000ffff0000ffffull;
    t = (t | (t << 8)) & 0x00ff00ff00ff00ffull;
    t = (t | (t << 4)) & 0x0f0f0f0f0f0f0f0full;
    return byte_swap(t) + 0x3030303030303030ull;
}
NPGE_HD uint64_t low_bytes(uint64_t v, uint32_t len) {  // len in 1..8
    return len >= 8 ? v : v & ((uint64_t(1) << (8 * len)) - 1);
}

// "_" + begin + "_" + last of a key as ASCII bytes, at most 23; returns the count.
NPGE_HD uint32_t id_tail(const IdKey& k, unsigned char* out) {
    const uint64_t lo = k.lo;
    const uint32_t nb = uint32_t(lo) & 0xf, nl = uint32_t(lo >> 4) & 0xf;
    uint32_t n = 0;
    out[n++] = '_';
    const uint64_t b = ((k.hi & 0xffffffffull) << 12) | (lo >> 52);  // 44 bits, digits on top
    for (uint32_t i = 0; i < nb; i++) out[n++] = (unsigned char)('0' + ((b >> (40 - 4 * i)) & 0xf));
    out[n++] = '_';
    if (nl == 0) {
        out[n++] = '-';
        out[n++] = '1';
    } else {
        const uint64_t l = ((lo >> 8) & 0xfffffffffffull) - last_digit_bias(nl);
        for (uint32_t i = 0; i < nl; i++) out[n++] = (unsigned char)('0' + ((l >> (40 - 4 * i)) & 0xf));
    }
    return n;
}

// How the ids of a block are ordered.  Two ids with different names are ordered by
// the rank of name + "_" (the first difference lies inside both prefixes) unless one
// name + "_" is a prefix of the other ("s1" next to "s1_2"): then the digits of one id
// meet the letters of the other, and the order is the byte-wise order of the two id
// strings, regenerated here (slow[rank] marks the names that have such a partner;
// nullptr: none has).
struct KeyOrder {
    const uint64_t* name_words;
    const uint32_t* name_pos;
    const uint32_t* name_len;
    const unsigned char* slow;

    NPGE_HD unsigned char name_byte(uint32_t rank, uint32_t i) const {
        return (unsigned char)(name_words[name_pos[rank] + i / 8] >> (8 * (i % 8)));
    }
    NPGE_HD bool string_less(const IdKey& a, const IdKey& b) const {
        const uint32_t ra = uint32_t(a.hi >> 32), rb = uint32_t(b.hi >> 32);
        const uint32_t la = name_len[ra], lb = name_len[rb];
        unsigned char ta[24], tb[24];
        const uint32_t na = la + id_tail(a, ta), nb = lb + id_tail(b, tb);
        const uint32_t n = na < nb ? na : nb;
        for (uint32_t i = 0; i < n; i++) {
            const unsigned char ca = i < la ? name_byte(ra, i) : ta[i - la];
            const unsigned char cb = i < lb ? name_byte(rb, i) : tb[i - lb];
            if (ca != cb) return ca < cb;
        }
        return na < nb;
    }
    NPGE_HD bool less(const IdKey& a, const IdKey& b) const {
        if (slow) {
            const uint32_t ra = uint32_t(a.hi >> 32), rb = uint32_t(b.hi >> 32);
            if (ra != rb && slow[ra] && slow[rb]) return string_less(a, b);
        }
        return key_less(a, b);
    }
};

// Keys of one thread interleaved with those of its neighbours (shared memory of
// the kernel: element i of thread t at base[i * stride + t]); stride 1 = an array.
struct KeySlots {
    IdKey* base;
    uint32_t stride;
    NPGE_HD IdKey& operator[](uint64_t i) const { return base[i * stride]; }
};

// In-place ascending sort; ids of one block are distinct, so stability is moot.
template <typename Keys>
NPGE_HD void sort_keys(Keys a, uint64_t n, const KeyOrder& ord) {
    if (n <= 24) {
        for (uint64_t i = 1; i < n; i++) {
            const IdKey x = a[i];
            uint64_t j = i;
            while (j > 0 && ord.less(x, a[j - 1])) {
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
            if (c + 1 < n && ord.less(a[c], a[c + 1])) c++;
            if (!ord.less(x, a[c])) break;
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
            if (c + 1 < end && ord.less(a[c], a[c + 1])) c++;
            if (!ord.less(x, a[c])) break;
            a[root] = a[c];
            root = c;
        }
        a[root] = x;
    }
}

// ids_dir < ids_inv of two sorted lists of the same length (std::vector<
// std::string>::operator<): decided by the first position that differs.
template <typename Keys>
NPGE_HD bool list_less(Keys a, Keys b, uint64_t n, const KeyOrder& ord) {
    for (uint64_t i = 0; i < n; i++) {
        if (ord.less(a[i], b[i])) return true;
        if (ord.less(b[i], a[i])) return false;
    }
    return false;
}

// The fold of block_hash.cpp:41-53 fed with chunks of up to 8 bytes (little-
// endian in a 64-bit word, unused high bytes 0).
struct HashFold {
    uint64_t a = 1, w0 = 0, w1 = 0, w2 = 0;
    uint32_t n = 0;  // bytes in the current 16-byte group
    NPGE_HD void put(uint64_t v, uint32_t len) {
        const uint32_t sh = (n & 7u) * 8u;
        const uint64_t lo = v << sh, hi = sh ? v >> (64u - sh) : 0;
        if (n < 8) {
            w0 |= lo;
            w1 |= hi;
        } else {
            w1 |= lo;
            w2 |= hi;
        }
        n += len;
        if (n >= 16) {
            a *= w0;
            a ^= w1;
            w0 = w2;
            w1 = w2 = 0;
            n -= 16;
        }
    }
    NPGE_HD uint64_t finish() {
        if (n) {  // joint.resize(new_size, ' ')
            const uint64_t spaces = 0x2020202020202020ull;
            if (n < 8) {
                w0 |= spaces << (8 * n);
                w1 = spaces;
            } else if (n < 16 && n > 8) {
                w1 |= spaces << (8 * (n - 8));
            } else if (n == 8) {
                w1 = spaces;
            }
            a *= w0;
            a ^= w1;
        }
        return a;
    }
};

// Names of the ranks: name_len[r] bytes packed little-endian into the 64-bit
// words name_words[name_pos[r] ...] (unused high bytes of the last word 0).
template <typename Keys>
NPGE_HD uint64_t fold_sorted(Keys a, uint64_t n, const uint64_t* name_words,
                             const uint32_t* name_pos, const uint32_t* name_len) {
    HashFold h;
    for (uint64_t i = 0; i < n; i++) {
        if (i) h.put(' ', 1);
        const uint32_t rank = uint32_t(a[i].hi >> 32);
        const uint64_t* w = name_words + name_pos[rank];
        for (uint32_t left = name_len[rank]; left > 0; w++) {
            const uint32_t len = left < 8 ? left : 8;
            h.put(*w, len);
            left -= len;
        }
        // "_" + begin + "_" + last
        const uint64_t lo = a[i].lo;
        const uint32_t nb = uint32_t(lo) & 0xf, nl = uint32_t(lo >> 4) & 0xf;
        const uint64_t b = (((a[i].hi & 0xffffffffull) << 12) | (lo >> 52)) << 20;  // digits on top
        const uint64_t b0 = nibbles_to_ascii(uint32_t(b >> 32));
        if (nb <= 7) {
            h.put('_' | (low_bytes(b0, nb) << 8), nb + 1);
        } else {
            h.put('_', 1);
            h.put(b0, 8);
            if (nb > 8) h.put(low_bytes(nibbles_to_ascii(uint32_t(b)), nb - 8), nb - 8);
        }
        if (nl == 0) {
            h.put('_' | (uint64_t('-') << 8) | (uint64_t('1') << 16), 3);
        } else {
            const uint64_t l = ((((lo >> 8) & 0xfffffffffffull) - last_digit_bias(nl))) << 20;
            const uint64_t l0 = nibbles_to_ascii(uint32_t(l >> 32));
            if (nl <= 7) {
                h.put('_' | (low_bytes(l0, nl) << 8), nl + 1);
            } else {
                h.put('_', 1);
                h.put(l0, 8);
                if (nl > 8) h.put(low_bytes(nibbles_to_ascii(uint32_t(l)), nl - 8), nl - 8);
            }
        }
    }
    return h.finish();
}

// The flat block list and the name table (see block_hash_names.h).
// max_pos == nullptr: every fragment is `frag_len` long (anchors).
struct FlatBlocks {
    const uint32_t* frag_seq;
    const uint32_t* min_pos;
    const uint32_t* max_pos;
    uint32_t frag_len;
    const int8_t* frag_ori;
    const uint32_t* seq_rank;
    const uint64_t* name_words;
    const uint32_t* name_pos;
    const uint32_t* name_len;
    const uint16_t* bcd4;  // [kBcdEntries]
    const unsigned char* name_slow;  // [ranks] or nullptr (KeyOrder::slow)
    NPGE_HD KeyOrder order() const {
        const KeyOrder o = {name_words, name_pos, name_len, name_slow};
        return o;
    }
    NPGE_HD void keys(uint64_t f, IdKey& dir, IdKey& inv) const {
        const uint32_t lo = min_pos[f];
        const uint32_t hi = max_pos ? max_pos[f] : lo + frag_len - 1;
        fragment_keys(seq_rank[frag_seq[f]], lo, hi, frag_ori[f], bcd4, dir, inv);
    }
};

// block_hash of the block whose fragments are [f0, f1), both id lists built and
// sorted.  scratch: 2 * (f1 - f0) keys.
NPGE_HD uint64_t block_hash_two_lists(const FlatBlocks& fb, uint64_t f0, uint64_t f1,
                                      IdKey* scratch) {
    const uint64_t n = f1 - f0;
    IdKey* dir = scratch;
    IdKey* inv = scratch + n;
    const KeyOrder ord = fb.order();
    for (uint64_t i = 0; i < n; i++) fb.keys(f0 + i, dir[i], inv[i]);
    sort_keys(dir, n, ord);
    sort_keys(inv, n, ord);
    const IdKey* ids = list_less(dir, inv, n, ord) ? dir : inv;
    return fold_sorted(ids, n, fb.name_words, fb.name_pos, fb.name_len);
}

// The same with one list: sorted lists compare by their first elements, i.e. by
// the smallest id of either list, unless those are equal (an id of one
// fragment equals the inverted id of another one: then *tie is set and nothing
// is computed).  Only the winning list is built and sorted.  slots: f1 - f0 keys.
template <typename Keys>
NPGE_HD uint64_t block_hash_one_list(const FlatBlocks& fb, uint64_t f0, uint64_t f1,
                                     Keys slots, bool* tie) {
    const uint64_t n = f1 - f0;
    const KeyOrder ord = fb.order();
    IdKey min_dir = {0, 0}, min_inv = min_dir;
    for (uint64_t i = 0; i < n; i++) {
        IdKey d, v;
        fb.keys(f0 + i, d, v);
        if (i == 0 || ord.less(d, min_dir)) min_dir = d;
        if (i == 0 || ord.less(v, min_inv)) min_inv = v;
    }
    const bool use_dir = ord.less(min_dir, min_inv);
    *tie = !use_dir && !ord.less(min_inv, min_dir);
    if (*tie) return 0;
    for (uint64_t i = 0; i < n; i++) {
        IdKey d, v;
        fb.keys(f0 + i, d, v);
        slots[i] = use_dir ? d : v;
    }
    sort_keys(slots, n, ord);
    return fold_sorted(slots, n, fb.name_words, fb.name_pos, fb.name_len);
}

// What the kernel runs per block: up to `fast_cap` fragments through the one-list
// path in `slots`, larger blocks and ties through `scratch` (2 keys per fragment).
template <typename Keys>
NPGE_HD uint64_t block_hash_flat(const FlatBlocks& fb, uint64_t f0, uint64_t f1, Keys slots,
                                 uint32_t fast_cap, IdKey* scratch) {
    if (f1 - f0 <= fast_cap) {
        bool tie;
        const uint64_t h = block_hash_one_list(fb, f0, f1, slots, &tie);
        if (!tie) return h;
    }
    return block_hash_two_lists(fb, f0, f1, scratch);
}

}  // namespace npge_b200
