// bucket_kernels.cu -- last stage of the anchor sort, fused with the grouping:
// every CTA takes one "unit" (consecutive (A, B) bins of one A bucket packed greedily
// up to kBucketCap records, unit_build_kernel; an oversized bin is a unit of its own), sorts
// it by the remaining key bits with an LSD radix sort in shared memory (byte
// digits, stable), finds the runs of equal canonical k-mers and writes the
// anchor fragments and blocks.  Three launches, no CTA ever waits on another:
// sort (+ count) -> scan of the per-unit totals -> emit from the sorted units, so
// the result comes out in ascending (k-mer, position) order.
//
// Replaces fragmenttg_postprocess (src/algo/AnchorFinder.cpp:356-391): the
// linear pass that opens a Block the first time a hash repeats,
// FoundFragment::make_fragment (:240-246), the max-anchor-fragments cut
// (:361-363), and -- in the two anchor-similar launches -- the neighbour rule of
// BloomTask::test_and_add (:170-182) with an exact filter.
//
// A unit larger than kBucketCap (one (A, B) bin with thousands of records: a
// k-mer repeated thousands of times, low-complexity sequence) takes the slow
// path of the same CTA: the same passes, chunk by chunk, through global memory.
#include <type_traits>

#include "kernels.h"
#include "rank.cuh"

#ifndef NPGE_EMIT_CTAS
#define NPGE_EMIT_CTAS 2
#endif
#ifndef NPGE_EMIT_DIRECT
#define NPGE_EMIT_DIRECT 1
#endif

namespace npge_b200 {

namespace {

constexpr int kCap = kBucketCap;
constexpr int kPartials = 64;
// unit totals and prefixes: blocks << 31 | fragments (both < 2^31)

__device__ __forceinline__ uint64_t rem_key(uint64_t rec, uint32_t hi) {
    return (uint64_t(hi) << 32) | (rec >> 32);
}

// SortedVector::has_elem (src/util/SortedVector.hpp:108-111) on the device
__device__ __forceinline__ bool is_used(const BucketArgs& a, uint64_t key) {
    uint64_t lo = 0, hi = a.n_used;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (a.used[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo < a.n_used && a.used[lo] == key;
}

__device__ __forceinline__ bool found_bit(const BucketArgs& a, uint32_t g) {
    return (a.found_bits[g >> 5] >> (g & 31)) & 1u;
}

template <bool WIDE>
struct __align__(16) BucketSmem {
    uint32_t grun[256];       // slow path: running digit bases of the unit
    uint64_t warp_tot[kRW];
    int32_t warp_head[kRW];
    uint64_t red[3][kRW];
    uint32_t carry;           // slow path: head of the group entering a chunk
    uint32_t cut;             // set by the thread that wrote the record of rank M-1
    // the unit: half[0] is enough for the kernels that only read a sorted unit
    // (count, emit), which are launched with the smaller size (kSingle)
    struct __align__(16) Half {
        uint64_t buf[kCap + 2];
        uint32_t hib[WIDE ? kCap + 2 : 4];
    } half[2];
    RankSmem<8> rk;           // ranking scratch of the sort kernel
    static constexpr size_t kSingle() { return sizeof(BucketSmem) - sizeof(Half) - sizeof(RankSmem<8>); }
};

// One sorted chunk of a unit in shared memory: records at K[1 .. cn], halo
// records at K[0] / K[cn + 1] when has_prev / has_next.
template <bool WIDE>
struct Chunk {
    const uint64_t* rec;   // shared: rec[0] = halo before, rec[1..cn], rec[cn+1] halo
    const uint32_t* hi;
    uint32_t cn;
    bool has_prev, has_next;
    uint32_t g0;           // global record index of rec[1]
    uint32_t carry_head;   // global index of the head of rec[1]'s group
    __device__ __forceinline__ uint64_t key(uint32_t j) const {  // j in 0 .. cn+1
        return rem_key(rec[j], WIDE ? hi[j] : 0u);
    }
};


// Item i of lane l of warp w is chunk element j = w * 32 * kRI + i * 32 + l.
__device__ __forceinline__ uint32_t item_index(int i) {
    return (threadIdx.x >> 5) * (32 * kRI) + i * 32 + (threadIdx.x & 31);
}

// Per-item flags of a sorted chunk, one bit per item round i:
//   valid    the item exists
//   preveq   same key as the record before it (= not the head of its group)
//   inb      its group has at least two members
template <bool WIDE>
__device__ __forceinline__ void unit_flags(const Chunk<WIDE>& ch, uint32_t& valid,
                                           uint32_t& preveq, uint32_t& inb) {
    valid = preveq = inb = 0;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t j = item_index(i);
        if (j < ch.cn) {
            const uint64_t key = ch.key(j + 1);
            const bool pe = (j > 0 || ch.has_prev) && ch.key(j) == key;
            const bool ne = (j + 1 < ch.cn || ch.has_next) && ch.key(j + 2) == key;
            valid |= 1u << i;
            preveq |= (pe ? 1u : 0u) << i;
            inb |= ((pe || ne) ? 1u : 0u) << i;
        }
    }
}

// Global record index of the head of every item's group.  Contains barriers.
template <bool WIDE>
__device__ __forceinline__ void unit_heads(BucketSmem<WIDE>& sm, const Chunk<WIDE>& ch,
                                           uint32_t valid, uint32_t preveq,
                                           uint32_t (&hidx)[kRI]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t le = (2u << lane) - 1;  // lanes <= lane
    int32_t last = -1;
    uint32_t hb[kRI];
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const bool head = ((valid >> i) & 1) && !((preveq >> i) & 1);
        hb[i] = __ballot_sync(0xffffffffu, head);
        if (hb[i]) last = int32_t(warp * (32 * kRI) + i * 32 + (31 - __clz(hb[i])));
    }
    __syncthreads();  // warp_head may still be read by a previous call
    if (lane == 0) sm.warp_head[warp] = last;
    __syncthreads();
    uint32_t carry = ch.carry_head;
    for (int w = 0; w < warp; w++) {
        int32_t h = sm.warp_head[w];
        if (h >= 0) carry = ch.g0 + uint32_t(h);
    }
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t m = hb[i] & le;
        const uint32_t base = warp * (32 * kRI) + i * 32;
        hidx[i] = m ? ch.g0 + base + (31 - __clz(m)) : carry;
        if (hb[i]) carry = ch.g0 + base + (31 - __clz(hb[i]));
    }
}

// Selection of the chunk's items and their count.  MODE 0: parity / exact
// (every group of >= 2 whose hash is not used); MODE 2: anchor-similar (groups
// holding a candidate: found(p) and not found(p-1), AnchorFinder.cpp:176).
// Returns the CTA total (blocks << 31 | fragments) in every thread; leaves the
// per-warp totals in sm.warp_tot for unit_write.  Contains barriers.
template <bool WIDE, int MODE>
__device__ __forceinline__ uint64_t unit_select(BucketSmem<WIDE>& sm,
                                                const Chunk<WIDE>& ch,
                                                const BucketArgs& a, uint64_t key_hi,
                                                uint32_t valid, uint32_t preveq,
                                                uint32_t inb, uint32_t& sel) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    sel = 0;
    if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            if ((inb >> i) & 1) {
                bool u = a.n_used &&
                         is_used(a, key_hi | ch.key(item_index(i) + 1));
                sel |= (u ? 0u : 1u) << i;
            }
        }
    } else {
        uint32_t hidx[kRI];
        unit_heads<WIDE>(sm, ch, valid, preveq, hidx);
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            if ((preveq >> i) & 1) {
                const uint32_t g = uint32_t(ch.rec[item_index(i) + 1]) >> 1;
                if (found_bit(a, g) && !(g > 0 && found_bit(a, g - 1))) {
                    a.sel[hidx[i]] = 1;
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            if (((inb >> i) & 1) && a.sel[hidx[i]]) sel |= 1u << i;
        }
    }
    uint32_t nf = 0, nb = 0;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const bool s = (sel >> i) & 1;
        nf += __popc(__ballot_sync(0xffffffffu, s));
        nb += __popc(__ballot_sync(0xffffffffu, s && !((preveq >> i) & 1)));
    }
    __syncthreads();  // warp_tot may still be read by a previous unit_write
    if (lane == 0) sm.warp_tot[warp] = (uint64_t(nb) << 31) | nf;
    __syncthreads();
    uint64_t tot = 0;
#pragma unroll
    for (int w = 0; w < kRW; w++) tot += sm.warp_tot[w];
    return tot;
}

struct Digest {
    uint64_t d0 = 0, d1 = 0, nb = 0;
};

// (scan-order sequence, position in it) of global position g
__device__ __forceinline__ void locate(const BucketArgs& a, uint32_t g, uint32_t& seq,
                                       uint32_t& pos) {
    const uint32_t q = g >> a.lut_shift;
    uint32_t lo = a.seq_lut[q], hi = a.seq_lut[q + 1] + 1;  // answer in [lo, hi)
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (a.seq_start[mid] <= g) lo = mid; else hi = mid;
    }
    seq = a.seq_caller[lo];
    pos = g - uint32_t(a.seq_start[lo]);
}

// The same from the packed (start, caller) table: two dependent loads per position.
__device__ __forceinline__ void locate_pair(const uint32_t* __restrict__ lut,
                                            const uint2* __restrict__ pair, int lut_shift,
                                            uint32_t g, uint32_t& seq, uint32_t& pos) {
    const uint32_t q = g >> lut_shift;
    uint32_t lo = lut[q];
    const uint32_t hi = lut[q + 1];
    uint2 p = pair[lo];
    while (lo < hi) {  // rare: a sequence starts inside this chunk of positions
        const uint2 nx = pair[lo + 1];
        if (nx.x > g) break;
        p = nx;
        lo++;
    }
    seq = p.y;
    pos = g - p.x;
}

// Writes the selected items of the chunk; `base` = global ranks before it.
template <bool WIDE>
__device__ __forceinline__ void unit_write(BucketSmem<WIDE>& sm, const Chunk<WIDE>& ch,
                                           const BucketArgs& a, uint64_t key_hi,
                                           uint32_t preveq, uint32_t sel,
                                           uint64_t base, Digest& dg) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1;
    for (int w = 0; w < warp; w++) base += sm.warp_tot[w];
    uint64_t fr = base & 0x7fffffffull, br = base >> 31;
    const uint64_t limit = a.max_fragments;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const bool s = (sel >> i) & 1;
        const bool head = s && !((preveq >> i) & 1);
        const uint32_t fb = __ballot_sync(0xffffffffu, s);
        const uint32_t bb = __ballot_sync(0xffffffffu, head);
        if (s) {
            const uint64_t f = fr + __popc(fb & lt);
            if (f < limit) {
                const uint32_t j = item_index(i);
                const uint32_t pay = uint32_t(ch.rec[j + 1]);
                const uint64_t key = key_hi | ch.key(j + 1);
                uint32_t seq, pos;
                locate(a, pay >> 1, seq, pos);
                if (a.r_seq) {  // (nullptr: compact result layout, payloads only)
                    a.r_seq[f] = seq;
                    a.r_pos[f] = pos;
                    a.r_ori[f] = (pay & 1) ? 1 : -1;
                }
                if (a.r_pay) a.r_pay[f] = pay;
                const uint64_t t = (uint64_t(seq) << 33) | (uint64_t(pos) << 1) | (pay & 1);
                uint64_t x, y;
                digest_terms(t, key, x, y);
                dg.d0 += x;
                dg.d1 += y;
                if (head) {
                    const uint64_t b = br + __popc(bb & lt);
                    a.r_block_off[b] = f;
                    a.r_block_hash[b] = key;
                    dg.nb++;
                }
                if (f + 1 == limit) {
                    a.counters[kCntCutIndex] = ch.g0 + j;
                    sm.cut = 1;
                }
            }
        }
        fr += __popc(fb);
        br += __popc(bb);
    }
}


// head of the group of sorted record `i` (key == key of record i) inside
// [lo_bound, i]: gallop back, then binary search (keys ascending)
template <bool WIDE>
__device__ uint32_t find_head_global(const uint64_t* rec, const uint32_t* hi,
                                     uint32_t lo_bound, uint32_t i, uint64_t key) {
    auto K = [&](uint32_t x) { return rem_key(rec[x], WIDE ? hi[x] : 0u); };
    uint32_t hi_i = i, step = 1, lo;
    while (true) {
        if (hi_i - lo_bound < step) { lo = lo_bound; break; }
        if (K(hi_i - step) != key) { lo = hi_i - step + 1; break; }
        hi_i -= step;
        step <<= 1;
    }
    while (lo < hi_i) {
        uint32_t mid = (lo + hi_i) >> 1;
        if (K(mid) == key) hi_i = mid; else lo = mid + 1;
    }
    return lo;
}

template <bool WIDE>
__device__ __forceinline__ void mark_found(const Chunk<WIDE>& ch, const BucketArgs& a,
                                           uint64_t key_hi, uint32_t preveq) {
    // found(p): the window's hash is not used and was seen at an earlier scan
    // position (BloomTask::test_and_add, AnchorFinder.cpp:170-182, exact filter):
    // every member of a sorted group but the first
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        if ((preveq >> i) & 1) {
            const uint32_t j = item_index(i);
            if (a.n_used && is_used(a, key_hi | ch.key(j + 1))) continue;
            const uint32_t g = uint32_t(ch.rec[j + 1]) >> 1;
            atomicOr(a.found_bits + (g >> 5), 1u << (g & 31));
        }
    }
}

template <bool WIDE>
__device__ __forceinline__ void flush_digest(BucketSmem<WIDE>& sm, const BucketArgs& a,
                                             uint32_t w, Digest dg) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dg.d0 += __shfl_xor_sync(0xffffffffu, dg.d0, o);
        dg.d1 += __shfl_xor_sync(0xffffffffu, dg.d1, o);
        dg.nb += __shfl_xor_sync(0xffffffffu, dg.nb, o);
    }
    if (lane == 0) {
        sm.red[0][warp] = dg.d0;
        sm.red[1][warp] = dg.d1;
        sm.red[2][warp] = dg.nb;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        uint64_t v = 0;
        for (int x = 0; x < kRW; x++) v += sm.red[threadIdx.x][x];
        if (v) {
            atomicAdd(reinterpret_cast<unsigned long long*>(
                          a.summary + kSumWords + (w % kPartials) * 4 + threadIdx.x),
                      (unsigned long long)v);
        }
    }
}

// Greedy packing of the (A, B) bins of one A bucket into units of at most kCap
// records (a bin larger than that is a unit of its own: slow path).  One warp per
// bucket: the lanes probe the next 32 bin ends at once.  WRITE = false counts the
// units of every bucket; WRITE = true writes them at the bucket's prefix.
template <bool WRITE>
__global__ void __launch_bounds__(256)
unit_build_kernel(BucketArgs a, uint32_t* __restrict__ unit_cnt) {
    const int lane = threadIdx.x & 31;
    const uint32_t ab = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    const uint32_t bins_a = 1u << a.plan.capa, nb = 1u << a.plan.wb;
    if (ab >= bins_a) return;
    // the bucket's row of bin starts (2^wb + 1 <= 1025 words) in shared memory: the
    // greedy loop below is a chain of dependent reads
    __shared__ uint32_t rows_sm[8][1025];
    uint32_t* fs = rows_sm[threadIdx.x >> 5];
    {
        const uint32_t* src = a.fine_start + (size_t(ab) << a.plan.wb);
        for (uint32_t x = lane; x <= nb; x += 32) fs[x] = src[x];
        __syncwarp();
    }
    uint32_t out = 0;
    if (WRITE) {  // units of the buckets before this one
        for (uint32_t x = lane; x < ab; x += 32) out += unit_cnt[x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) out += __shfl_xor_sync(0xffffffffu, out, o);
    }
    const uint32_t end_all = fs[nb];
    uint32_t cur = 0, count = 0;
    while (cur < nb) {
        const uint32_t s = fs[cur];
        if (s == end_all) break;  // only empty bins are left
        {   // skip empty bins: a unit starts with a record
            const uint32_t probe = cur + lane;
            const uint32_t m = __ballot_sync(0xffffffffu, probe < nb && fs[probe + 1] == s);
            const int run = m == 0xffffffffu ? 32 : __ffs(~m) - 1;
            if (run) {
                cur += run;
                continue;
            }
        }
        uint32_t e = cur + 1;     // the unit is [cur, e); grows while it fits
        while (true) {
            const uint32_t probe = e + lane;  // candidate bin to add: [cur, probe]
            const bool fits = probe < nb && fs[probe + 1] - s <= uint32_t(kCap);
            const uint32_t m = __ballot_sync(0xffffffffu, fits);
            const int run = m == 0xffffffffu ? 32 : __ffs(~m) - 1;
            e += run;
            if (run < 32) break;
        }
        if (WRITE && lane == 0) {
            a.units[out + count] = make_uint4(s, fs[e] - s, ab, 0u);
        }
        count++;
        cur = e;
    }
    if (lane == 0) {
        if (!WRITE) unit_cnt[ab] = count;
        if (WRITE && ab == bins_a - 1) a.counters[kCntWindows] = out + count;
    }
}

// The unit w of the CTA: (first record, size, A bucket).  False past the last
// unit.  Contains a barrier when true.
template <bool WIDE>
__device__ __forceinline__ bool unit_of_window(BucketSmem<WIDE>& sm, const BucketArgs& a,
                                               uint32_t w, uint32_t& s, uint32_t& n,
                                               uint64_t& key_hi) {
    if (w >= (a.unit_end ? a.unit_end : a.counters[kCntWindows])) return false;
    const uint4 u = a.units[w];
    if (threadIdx.x == 0) sm.cut = 0;
    __syncthreads();
    s = u.x;
    n = u.y;
    key_hi = uint64_t(u.z) << a.plan.rem;
    return true;
}

__device__ __forceinline__ uint32_t digit_at(uint64_t r, uint32_t h, int shift) {
    return uint32_t(rem_key(r, h) >> shift) & 255u;
}

// Loads the unit (n <= kCap) into sm.half[0].buf[1..n].  Contains a barrier.
template <bool WIDE>
__device__ __forceinline__ void load_unit(BucketSmem<WIDE>& sm, const BucketArgs& a,
                                          uint32_t s, uint32_t n) {
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t j = item_index(i);
        if (j < n) {
            sm.half[0].buf[1 + j] = a.rec_w[s + j];
            if (WIDE) sm.half[0].hib[1 + j] = a.hi_w[s + j];
        }
    }
    __syncthreads();
}

template <bool WIDE>
__device__ __forceinline__ void store_unit(BucketSmem<WIDE>& sm, const BucketArgs& a,
                                           uint32_t s, uint32_t n, int cur) {
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t j = item_index(i);
        if (j < n) {
            a.rec_w[s + j] = sm.half[cur].buf[1 + j];
            if (WIDE) a.hi_w[s + j] = sm.half[cur].hib[1 + j];
        }
    }
}

// LSD radix sort of sm.half[0].buf[1..n] by the low `rem` key bits, byte digits,
// stable.  Returns the buffer holding the result.  Contains barriers.
template <bool WIDE>
__device__ __forceinline__ int sort_unit_smem(BucketSmem<WIDE>& sm, const MsdPlan& pl,
                                              uint32_t n) {
    const int warp = threadIdx.x >> 5;
    const uint32_t first = warp * (32 * kRI);
    const int rounds = first >= n ? 0 : int(min(uint32_t(kRI), (n - first + 31) / 32));
    int cur = 0;
#pragma unroll 1
    for (int pass = 0; pass < pl.smem_passes; pass++) {
        uint64_t rec[kRI];
        uint32_t hi[kRI], digit[kRI], slot[kRI];
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            const uint32_t j = item_index(i);
            const bool in = j < n;
            rec[i] = in ? sm.half[cur].buf[1 + j] : 0ull;
            hi[i] = (WIDE && in) ? sm.half[cur].hib[1 + j] : 0u;
            digit[i] = in ? digit_at(rec[i], hi[i], 8 * pass) : 256u;
        }
        uint32_t c0, c1;
        tile_rank_abs<8>(sm.rk, digit, rounds, slot, c0, c1);
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            if (item_index(i) < n) {
                sm.half[cur ^ 1].buf[1 + slot[i]] = rec[i];
                if (WIDE) sm.half[cur ^ 1].hib[1 + slot[i]] = hi[i];
            }
        }
        __syncthreads();
        cur ^= 1;
    }
    return cur;
}

// Oversized unit: the same passes chunk by chunk through global memory, ping-pong
// between the unit's own range and the same range of the scratch buffer.
template <bool WIDE>
__device__ void sort_unit_global(BucketSmem<WIDE>& sm, const BucketArgs& a, uint32_t s,
                                 uint32_t n) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const MsdPlan& pl = a.plan;
    uint64_t* src = a.rec_w;
    uint32_t* srch = a.hi_w;
    uint64_t* dst = a.scratch_rec;
    uint32_t* dsth = a.scratch_hi;
#pragma unroll 1
    for (int pass = 0; pass < pl.slow_passes; pass++) {  // one bin: only the low bits differ
        const int shift = 8 * pass;
        __syncthreads();
        if (tid < 256) sm.grun[tid] = 0;
        __syncthreads();
        for (uint32_t x = tid; x < n; x += kRT) {
            atomicAdd(&sm.grun[digit_at(src[s + x], WIDE ? srch[s + x] : 0u, shift)], 1u);
        }
        __syncthreads();
        {   // exclusive scan of the 256 digit counts (warps 0..7 hold them)
            uint32_t v = tid < 256 ? sm.grun[tid] : 0u;
            uint32_t incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += u;
            }
            if (lane == 31) sm.rk.scan_tmp[warp] = incl;
            __syncthreads();
            uint32_t base = 0;
            for (int x = 0; x < warp; x++) base += sm.rk.scan_tmp[x];
            __syncthreads();
            if (tid < 256) sm.grun[tid] = base + incl - v;
            __syncthreads();
        }
#pragma unroll 1
        for (uint32_t c0 = 0; c0 < n; c0 += kCap) {
            const uint32_t cn = min(uint32_t(kCap), n - c0);
            uint64_t rec[kRI];
            uint32_t hi[kRI], digit[kRI], slot[kRI];
#pragma unroll
            for (int i = 0; i < kRI; i++) {
                const uint32_t j = item_index(i);
                const bool in = j < cn;
                rec[i] = in ? src[s + c0 + j] : 0ull;
                hi[i] = (WIDE && in) ? srch[s + c0 + j] : 0u;
                digit[i] = in ? digit_at(rec[i], hi[i], shift) : 256u;
            }
            const uint32_t first = warp * (32 * kRI);
            const int rounds =
                first >= cn ? 0 : int(min(uint32_t(kRI), (cn - first + 31) / 32));
            uint32_t cnt0, cnt1;
            tile_rank<8>(sm.rk, digit, rounds, slot, cnt0, cnt1);
#pragma unroll
            for (int i = 0; i < kRI; i++) {
                if (digit[i] < 256u) {
                    sm.half[0].buf[slot[i]] = rec[i];
                    if (WIDE) sm.half[0].hib[slot[i]] = hi[i];
                }
            }
            __syncthreads();
#pragma unroll
            for (int i = 0; i < kRI; i++) {
                const uint32_t j = tid + i * kRT;
                if (j < cn) {
                    const uint64_t r = sm.half[0].buf[j];
                    const uint32_t h = WIDE ? sm.half[0].hib[j] : 0u;
                    const uint32_t d = digit_at(r, h, shift);
                    const uint32_t at = s + sm.grun[d] + j - sm.rk.tile_off[d];
                    dst[at] = r;
                    if (WIDE) dsth[at] = h;
                }
            }
            __syncthreads();
            if (tid < 256) sm.grun[tid] += cnt0;
            __syncthreads();
        }
        uint64_t* t = src; src = dst; dst = t;
        uint32_t* th = srch; srch = dsth; dsth = th;
    }
    if (src != a.rec_w) {  // odd number of passes: bring the unit home
        for (uint32_t x = tid; x < n; x += kRT) {
            a.rec_w[s + x] = src[s + x];
            if (WIDE) a.hi_w[s + x] = srch[s + x];
        }
    }
    __syncthreads();
}

// Chunk c0 of a sorted oversized unit -> sm.half[0].buf with halos.  Contains barriers.
template <bool WIDE>
__device__ __forceinline__ void load_chunk(BucketSmem<WIDE>& sm, const BucketArgs& a,
                                           uint32_t s, uint32_t n, uint32_t c0,
                                           Chunk<WIDE>& ch) {
    const uint64_t* X = a.rec_w;
    const uint32_t* XH = a.hi_w;
    const uint32_t cn = min(uint32_t(kCap), n - c0);
    __syncthreads();  // the previous chunk is no longer read
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t j = item_index(i);
        if (j < cn) {
            sm.half[0].buf[1 + j] = X[s + c0 + j];
            if (WIDE) sm.half[0].hib[1 + j] = XH[s + c0 + j];
        }
    }
    if (threadIdx.x == 0) {
        sm.carry = s + c0;
        if (c0 > 0) {
            sm.half[0].buf[0] = X[s + c0 - 1];
            if (WIDE) sm.half[0].hib[0] = XH[s + c0 - 1];
            const uint64_t k0 = rem_key(X[s + c0 - 1], WIDE ? XH[s + c0 - 1] : 0u);
            const uint64_t k1 = rem_key(X[s + c0], WIDE ? XH[s + c0] : 0u);
            if (k0 == k1) sm.carry = find_head_global<WIDE>(X, XH, s, s + c0 - 1, k0);
        }
        if (c0 + cn < n) {
            sm.half[0].buf[cn + 1] = X[s + c0 + cn];
            if (WIDE) sm.half[0].hib[cn + 1] = XH[s + c0 + cn];
        }
    }
    __syncthreads();
    ch.rec = sm.half[0].buf;
    ch.hi = sm.half[0].hib;
    ch.cn = cn;
    ch.has_prev = c0 > 0;
    ch.has_next = c0 + cn < n;
    ch.g0 = s + c0;
    ch.carry_head = sm.carry;
}

template <bool WIDE>
__device__ __forceinline__ void whole_unit_chunk(BucketSmem<WIDE>& sm, uint32_t s,
                                                 uint32_t n, int cur, Chunk<WIDE>& ch) {
    ch.rec = sm.half[cur].buf;
    ch.hi = sm.half[cur].hib;
    ch.cn = n;
    ch.has_prev = ch.has_next = false;
    ch.g0 = s;
    ch.carry_head = s;
}

// ---- kernel 1: sort every unit; MODE 0 also counts its anchors, MODE 1
// (anchor-similar) sets the found bits instead.  The sorted unit is written back.
template <bool WIDE, int MODE>
__global__ void __launch_bounds__(kRT, 2)
bucket_sort_kernel(BucketArgs a) {
    extern __shared__ __align__(16) unsigned char raw[];
    BucketSmem<WIDE>& sm = *reinterpret_cast<BucketSmem<WIDE>*>(raw);
    const uint32_t w = blockIdx.x + a.unit_first;
    uint32_t s, n;
    uint64_t key_hi;
    if (!unit_of_window<WIDE>(sm, a, w, s, n, key_hi)) return;
    Chunk<WIDE> ch;
    uint32_t valid, preveq, inb, sel;
    uint64_t tot = 0;
    if (n <= uint32_t(kCap)) {
        load_unit<WIDE>(sm, a, s, n);
        const int cur = sort_unit_smem<WIDE>(sm, a.plan, n);
        whole_unit_chunk<WIDE>(sm, s, n, cur, ch);
        unit_flags<WIDE>(ch, valid, preveq, inb);
        if (MODE == 1) {
            mark_found<WIDE>(ch, a, key_hi, preveq);
        } else {
            tot = unit_select<WIDE, 0>(sm, ch, a, key_hi, valid, preveq, inb, sel);
            // the emit kernel ranks its fragments from these, without a count of its own
            if ((threadIdx.x & 31) == 0) {
                const uint64_t wt = sm.warp_tot[threadIdx.x >> 5];
                a.unit_wtot[size_t(w) * kRW + (threadIdx.x >> 5)] =
                    uint32_t(wt & 0xffff) | (uint32_t(wt >> 31) << 16);
            }
        }
        store_unit<WIDE>(sm, a, s, n, cur);
    } else {
        sort_unit_global<WIDE>(sm, a, s, n);
        for (uint32_t c0 = 0; c0 < n; c0 += kCap) {
            load_chunk<WIDE>(sm, a, s, n, c0, ch);
            unit_flags<WIDE>(ch, valid, preveq, inb);
            if (MODE == 1) {
                mark_found<WIDE>(ch, a, key_hi, preveq);
            } else {
                tot += unit_select<WIDE, 0>(sm, ch, a, key_hi, valid, preveq, inb, sel);
            }
        }
    }
    if (MODE == 0 && threadIdx.x == 0) a.unit_tot[w] = tot;
}

// ---- kernel 1b (anchor-similar): candidate rule and count on the sorted units -----
template <bool WIDE>
__global__ void __launch_bounds__(kRT, WIDE ? 2 : 3)
bucket_count_kernel(BucketArgs a) {
    extern __shared__ __align__(16) unsigned char raw[];
    BucketSmem<WIDE>& sm = *reinterpret_cast<BucketSmem<WIDE>*>(raw);
    const uint32_t w = blockIdx.x + a.unit_first;
    uint32_t s, n;
    uint64_t key_hi;
    if (!unit_of_window<WIDE>(sm, a, w, s, n, key_hi)) return;
    Chunk<WIDE> ch;
    uint32_t valid, preveq, inb, sel;
    uint64_t tot = 0;
    if (n <= uint32_t(kCap)) {
        load_unit<WIDE>(sm, a, s, n);
        whole_unit_chunk<WIDE>(sm, s, n, 0, ch);
        unit_flags<WIDE>(ch, valid, preveq, inb);
        tot = unit_select<WIDE, 2>(sm, ch, a, key_hi, valid, preveq, inb, sel);
    } else {
        // sweep 0: every candidate marks its group, also across chunk borders
        for (int sweep = 0; sweep < 2; sweep++) {
            tot = 0;
            for (uint32_t c0 = 0; c0 < n; c0 += kCap) {
                load_chunk<WIDE>(sm, a, s, n, c0, ch);
                unit_flags<WIDE>(ch, valid, preveq, inb);
                tot += unit_select<WIDE, 2>(sm, ch, a, key_hi, valid, preveq, inb, sel);
            }
        }
    }
    if (threadIdx.x == 0) a.unit_tot[w] = tot;
}

// ---- kernel 2: exclusive scan of the unit totals (one CTA) -------------------------
__global__ void __launch_bounds__(1024)
unit_scan_kernel(BucketArgs a) {
    __shared__ uint64_t part[1024];
    // units [u0, u1) (all of them by default), continuing at the carry of the ranges
    // scanned before (0 after the memset of the summary)
    const uint32_t u0 = a.unit_first;
    const uint32_t u1 = a.unit_end ? a.unit_end : a.counters[kCntWindows];
    const uint32_t n_win = u1 > u0 ? u1 - u0 : 0;
    const uint32_t per = (n_win + 1023) / 1024;
    const uint32_t b = min(u0 + threadIdx.x * per, u1), e = min(b + per, u1);
    const uint64_t carry = a.summary[kSumCarry];
    __syncthreads();  // everybody has read the carry before lane 31 of warp 0 replaces it
    uint64_t sum = 0;
    for (uint32_t i = b; i < e; i++) sum += a.unit_tot[i];
    // block-wide exclusive scan of the 1024 partial sums
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint64_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) part[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint64_t v = part[lane];
        uint64_t wi = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t u = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += u;
        }
        part[32 + lane] = wi - v;  // sum of the warps before this one
        if (lane == 31) {
            const uint64_t all = carry + wi;
            a.summary[kSumCarry] = all;
            a.summary[kSumFragments] = all & 0x7fffffffull;  // before truncation
            a.summary[kSumBlocks] = all >> 31;
        }
    }
    __syncthreads();
    uint64_t run = carry + part[32 + warp] + incl - sum;
    for (uint32_t i = b; i < e; i++) {
        uint64_t v = a.unit_tot[i];
        a.unit_tot[i] = run;
        run += v;
    }
}

// ---- kernel 3: write the anchors of every unit --------------------------------------
template <bool WIDE, int MODE>
__global__ void __launch_bounds__(kRT, WIDE ? 2 : 3)
bucket_emit_kernel(BucketArgs a) {
    extern __shared__ __align__(16) unsigned char raw[];
    BucketSmem<WIDE>& sm = *reinterpret_cast<BucketSmem<WIDE>*>(raw);
    const uint32_t w = blockIdx.x + a.unit_first;
    uint32_t s, n;
    uint64_t key_hi;
    if (!unit_of_window<WIDE>(sm, a, w, s, n, key_hi)) return;
    uint64_t base = a.unit_tot[w];
    if ((base & 0x7fffffffull) >= a.max_fragments) return;  // past the cut
    Chunk<WIDE> ch;
    uint32_t valid, preveq, inb, sel;
    Digest dg;
    if (n <= uint32_t(kCap)) {
        load_unit<WIDE>(sm, a, s, n);
        whole_unit_chunk<WIDE>(sm, s, n, 0, ch);
        unit_flags<WIDE>(ch, valid, preveq, inb);
        unit_select<WIDE, MODE>(sm, ch, a, key_hi, valid, preveq, inb, sel);
        unit_write<WIDE>(sm, ch, a, key_hi, preveq, sel, base, dg);
    } else {
        for (uint32_t c0 = 0; c0 < n; c0 += kCap) {
            load_chunk<WIDE>(sm, a, s, n, c0, ch);
            unit_flags<WIDE>(ch, valid, preveq, inb);
            const uint64_t t =
                unit_select<WIDE, MODE>(sm, ch, a, key_hi, valid, preveq, inb, sel);
            unit_write<WIDE>(sm, ch, a, key_hi, preveq, sel, base, dg);
            base += t;
        }
    }
    __syncthreads();
    if (sm.cut && threadIdx.x == 0) {
        a.counters[kCntCutStart] = s;
        a.counters[kCntCutSize] = n;
    }
    flush_digest<WIDE>(sm, a, w, dg);
}

// ---- kernel 3, parity / exact mode: write the anchors straight from the sorted
// records.  Every thread reads its kRI records with coalesced loads (item i of lane l
// of warp w = record w * 32 * kRI + i * 32 + l of the chunk, as in the sort kernel),
// the neighbours' keys come by warp shuffles, and the first fragment of every warp
// comes from the per-warp totals the sort kernel left (unit_wtot): no shared-memory
// copy of the unit, no count pass, no barrier before the writes.  An oversized unit
// goes chunk by chunk with a count pass per chunk.
struct __align__(16) DirectSmem {
    uint64_t warp_tot[kRW];
    uint64_t red[3][kRW];
    uint32_t cut;
};

template <bool WIDE>
__global__ void __launch_bounds__(kRT, NPGE_EMIT_CTAS)
bucket_emit_direct_kernel(BucketArgs a) {
    __shared__ DirectSmem sm;
    typedef typename std::conditional<WIDE, uint64_t, uint32_t>::type key_t;
    const uint32_t w = blockIdx.x + a.unit_first;
    if (w >= (a.unit_end ? a.unit_end : a.counters[kCntWindows])) return;
    const uint4 u = a.units[w];
    const uint32_t s = u.x, n = u.y;
    const uint64_t key_hi = uint64_t(u.z) << a.plan.rem;
    uint64_t base = a.unit_tot[w];
    if ((base & 0x7fffffffull) >= a.max_fragments) return;  // past the cut
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1;
    if (tid == 0) sm.cut = 0;
    const uint64_t* R = a.rec_w + s;
    const uint32_t* H = a.hi_w + s;
    auto key_at = [&](uint32_t j) -> key_t {
        return key_t(rem_key(R[j], WIDE ? H[j] : 0u));
    };
    // fragment and block indices fit 31 bits (unit_tot packs them so)
    const uint32_t limit = uint32_t(a.max_fragments < 0x80000000ull ? a.max_fragments : 0x80000000ull);
    uint32_t* __restrict__ r_seq = a.r_seq;
    uint32_t* __restrict__ r_pos = a.r_pos;
    int8_t* __restrict__ r_ori = a.r_ori;
    uint32_t* __restrict__ r_pay = a.r_pay;
    uint64_t* __restrict__ r_block_off = a.r_block_off;
    uint64_t* __restrict__ r_block_hash = a.r_block_hash;
    Digest dg;
#pragma unroll 1
    for (uint32_t c0 = 0; c0 < n; c0 += kCap) {
        const uint32_t cn = min(uint32_t(kCap), n - c0);
        const uint32_t j0 = c0 + warp * (32 * kRI);  // first record of the warp
        key_t key[kRI];
        uint32_t pay[kRI];
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            const uint32_t j = j0 + i * 32 + lane;
            key[i] = 0;
            pay[i] = 0;
            if (j < c0 + cn) {
                const uint64_t r = R[j];
                key[i] = key_t(rem_key(r, WIDE ? H[j] : 0u));
                pay[i] = uint32_t(r);
            }
        }
        // keys just outside the warp's range (lane 0 / lane 31 use them)
        key_t edge_prev = 0, edge_next = 0;
        if (lane == 0 && j0 > 0 && j0 - 1 < n) edge_prev = key_at(j0 - 1);
        if (lane == 31 && j0 + 32 * kRI < n) edge_next = key_at(j0 + 32 * kRI);
        // (sequence, position) of every record first: independent loads in flight
        // together, before any store the compiler would have to order them against
        uint32_t seq[kRI], pos[kRI];
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            seq[i] = pos[i] = 0;
            if (j0 + i * 32 + lane < c0 + cn) {
                locate_pair(a.seq_lut, a.seq_pair, a.lut_shift, pay[i] >> 1, seq[i], pos[i]);
            }
            pos[i] = (pos[i] << 1) | (pay[i] & 1);  // position << 1 | (ori == +1)
        }
        uint32_t preveq = 0, sel = 0;
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            const uint32_t j = j0 + i * 32 + lane;
            key_t kp = __shfl_up_sync(0xffffffffu, key[i], 1);
            key_t kn = __shfl_down_sync(0xffffffffu, key[i], 1);
            const key_t from31 = __shfl_sync(0xffffffffu, key[i > 0 ? i - 1 : 0], 31);
            const key_t from0 = __shfl_sync(0xffffffffu, key[i < kRI - 1 ? i + 1 : kRI - 1], 0);
            if (lane == 0) kp = i > 0 ? from31 : edge_prev;
            if (lane == 31) kn = i < kRI - 1 ? from0 : edge_next;
            const bool in = j < c0 + cn;
            const bool pe = in && j > 0 && kp == key[i];
            const bool ne = in && j + 1 < n && kn == key[i];
            bool s_ = pe || ne;
            if (s_ && a.n_used) s_ = !is_used(a, key_hi | uint64_t(key[i]));
            preveq |= (pe ? 1u : 0u) << i;
            sel |= (s_ ? 1u : 0u) << i;
        }
        // fragments and blocks before this warp's first item
        uint64_t before = 0, chunk_tot = 0;
        if (n <= uint32_t(kCap)) {
            uint32_t x = lane < warp ? a.unit_wtot[size_t(w) * kRW + lane] : 0u;
            uint32_t f = x & 0xffff, b = x >> 16;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                f += __shfl_xor_sync(0xffffffffu, f, o);
                b += __shfl_xor_sync(0xffffffffu, b, o);
            }
            before = (uint64_t(b) << 31) | f;
        } else {
            uint32_t nf = 0, nb = 0;
#pragma unroll
            for (int i = 0; i < kRI; i++) {
                const bool s_ = (sel >> i) & 1;
                nf += __popc(__ballot_sync(0xffffffffu, s_));
                nb += __popc(__ballot_sync(0xffffffffu, s_ && !((preveq >> i) & 1)));
            }
            __syncthreads();  // warp_tot may still be read for the previous chunk
            if (lane == 0) sm.warp_tot[warp] = (uint64_t(nb) << 31) | nf;
            __syncthreads();
            for (int x = 0; x < kRW; x++) {
                const uint64_t t = sm.warp_tot[x];
                if (x < warp) before += t;
                chunk_tot += t;
            }
        }
        uint32_t fr = uint32_t((base + before) & 0x7fffffffull), br = uint32_t((base + before) >> 31);
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            const bool s_ = (sel >> i) & 1;
            const bool head = s_ && !((preveq >> i) & 1);
            const uint32_t fb = __ballot_sync(0xffffffffu, s_);
            const uint32_t bb = __ballot_sync(0xffffffffu, head);
            if (s_) {
                const uint32_t f = fr + __popc(fb & lt);
                if (f < limit) {
                    const uint64_t k64 = key_hi | uint64_t(key[i]);
                    if (r_seq) {  // (nullptr: compact result layout, payloads only)
                        r_seq[f] = seq[i];
                        r_pos[f] = pos[i] >> 1;
                        r_ori[f] = int8_t(2 * int(pos[i] & 1) - 1);
                    }
                    if (r_pay) r_pay[f] = pay[i];
                    const uint64_t t = (uint64_t(seq[i]) << 33) | pos[i];
                    uint64_t x, y;
                    digest_terms(t, k64, x, y);
                    dg.d0 += x;
                    dg.d1 += y;
                    if (head) {
                        const uint32_t b = br + __popc(bb & lt);
                        r_block_off[b] = f;
                        r_block_hash[b] = k64;
                        dg.nb++;
                    }
                    if (f + 1 == limit) {
                        a.counters[kCntCutIndex] = s + j0 + i * 32 + lane;
                        sm.cut = 1;
                    }
                }
            }
            fr += __popc(fb);
            br += __popc(bb);
        }
        base += chunk_tot;
    }
    __syncthreads();
    if (sm.cut && tid == 0) {
        a.counters[kCntCutStart] = s;
        a.counters[kCntCutSize] = n;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dg.d0 += __shfl_xor_sync(0xffffffffu, dg.d0, o);
        dg.d1 += __shfl_xor_sync(0xffffffffu, dg.d1, o);
        dg.nb += __shfl_xor_sync(0xffffffffu, dg.nb, o);
    }
    if (lane == 0) {
        sm.red[0][warp] = dg.d0;
        sm.red[1][warp] = dg.d1;
        sm.red[2][warp] = dg.nb;
    }
    __syncthreads();
    if (tid < 3) {
        uint64_t v = 0;
        for (int x = 0; x < kRW; x++) v += sm.red[tid][x];
        if (v) {
            atomicAdd(reinterpret_cast<unsigned long long*>(
                          a.summary + kSumWords + (w % kPartials) * 4 + tid),
                      (unsigned long long)v);
        }
    }
}

__global__ void bucket_finalize_kernel(BucketArgs a) {
    if (threadIdx.x >= 32 || blockIdx.x != 0) return;
    uint64_t d0 = 0, d1 = 0, nb = 0;
    for (int s = threadIdx.x; s < kPartials; s += 32) {
        d0 += a.summary[kSumWords + s * 4 + 0];
        d1 += a.summary[kSumWords + s * 4 + 1];
        nb += a.summary[kSumWords + s * 4 + 2];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        d0 += __shfl_xor_sync(0xffffffffu, d0, o);
        d1 += __shfl_xor_sync(0xffffffffu, d1, o);
        nb += __shfl_xor_sync(0xffffffffu, nb, o);
    }
    if (threadIdx.x != 0) return;
    uint64_t total = a.summary[kSumFragments];
    uint64_t emitted = total < a.max_fragments ? total : a.max_fragments;
    a.summary[kSumDigest0] = d0;
    a.summary[kSumDigest1] = d1;
    a.summary[kSumEmitted] = emitted;
    a.summary[kSumBlocksEmitted] = nb;
    a.r_block_off[nb] = emitted;
}

// ---- found bitmap of a multi-GPU job inside one process: dst = OR of the ranks'
// bitmaps, read through peer pointers (every bit is set by exactly one rank)
struct FoundSources {
    const uint32_t* src[8];
    int n;
};

__global__ void __launch_bounds__(256)
found_or_kernel(FoundSources fs, uint32_t* __restrict__ dst, size_t n_vec) {
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n_vec; i += stride) {
        uint4 v = make_uint4(0, 0, 0, 0);
        for (int r = 0; r < fs.n; r++) {
            const uint4 x = reinterpret_cast<const uint4*>(fs.src[r])[i];
            v.x |= x.x;
            v.y |= x.y;
            v.z |= x.z;
            v.w |= x.w;
        }
        reinterpret_cast<uint4*>(dst)[i] = v;
    }
}

template <typename K>
void set_smem(K kernel, size_t bytes) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         int(bytes));
}

}  // namespace

size_t bucket_summary_words() { return kSumWords + kPartials * 4; }

// upper bound on the units of the bucket sort (two neighbours never fit in one):
// the grid of its kernels and the size of the unit tables
uint32_t bucket_max_windows(const MsdPlan& plan, uint64_t max_records) {
    return uint32_t((max_records + kWindow - 1) / kWindow) + (2u << plan.capa) + 1;
}

// The stages of the bucket sort, one launcher each so that the multi-GPU driver
// can put its collectives between them (found bitmap, totals).
//   kBucketUnits   pack the bins into units
//   kBucketSort    sort every unit + count its anchors (parity / exact)
//   kBucketMark    sort every unit + set the found bits (anchor-similar)
//   kBucketCount   candidate rule + count on the sorted units (anchor-similar)
//   kBucketScan    exclusive scan of the unit totals (summary: totals before the cut)
//   kBucketEmit / kBucketEmitSimilar   write the anchors
//   kBucketFinal   digest, emitted counts, closing block offset
void launch_bucket_stage(const BucketArgs& a, int stage, uint64_t max_records,
                         cudaStream_t s, uint32_t* launches) {
    const uint32_t grid = a.unit_end ? (a.unit_end > a.unit_first ? a.unit_end - a.unit_first : 0u)
                                     : bucket_max_windows(a.plan, max_records);
#define NPGE_LAUNCH(KERNEL, BYTES)                                              \
    do {                                                                        \
        set_smem(KERNEL, BYTES);                                                \
        KERNEL<<<grid, kRT, BYTES, s>>>(a);                                     \
        *launches += 1;                                                         \
    } while (0)
// FULL: both halves of the unit buffer (the sort ping-pongs between them)
#define NPGE_LAUNCH_W(FULL, NAME, ...)                                          \
    do {                                                                        \
        if (a.plan.wide) NPGE_LAUNCH((NAME<true, ##__VA_ARGS__>),               \
            FULL ? sizeof(BucketSmem<true>) : BucketSmem<true>::kSingle());     \
        else NPGE_LAUNCH((NAME<false, ##__VA_ARGS__>),                          \
            FULL ? sizeof(BucketSmem<false>) : BucketSmem<false>::kSingle());   \
    } while (0)
    if (stage == kBucketFinal) {
        bucket_finalize_kernel<<<1, 32, 0, s>>>(a);
        *launches += 1;
        return;
    }
    if (max_records == 0) return;
    if (a.unit_end && grid == 0 && stage != kBucketScan && stage != kBucketUnits) return;
    switch (stage) {
    case kBucketUnits: {
        const uint32_t bins_a = 1u << a.plan.capa;
        uint32_t* unit_cnt = reinterpret_cast<uint32_t*>(a.unit_tot);  // free until the sort
        unit_build_kernel<false><<<(bins_a + 7) / 8, 256, 0, s>>>(a, unit_cnt);
        unit_build_kernel<true><<<(bins_a + 7) / 8, 256, 0, s>>>(a, unit_cnt);
        *launches += 2;
        break;
    }
    case kBucketSort: NPGE_LAUNCH_W(true, bucket_sort_kernel, 0); break;
    case kBucketMark: NPGE_LAUNCH_W(true, bucket_sort_kernel, 1); break;
    case kBucketCount: NPGE_LAUNCH_W(false, bucket_count_kernel); break;
    case kBucketScan:
        unit_scan_kernel<<<1, 1024, 0, s>>>(a);
        *launches += 1;
        break;
    case kBucketEmit:
#if NPGE_EMIT_DIRECT
        if (a.plan.wide) bucket_emit_direct_kernel<true><<<grid, kRT, 0, s>>>(a);
        else bucket_emit_direct_kernel<false><<<grid, kRT, 0, s>>>(a);
        *launches += 1;
#else
        NPGE_LAUNCH_W(false, bucket_emit_kernel, 0);
#endif
        break;
    case kBucketEmitSimilar: NPGE_LAUNCH_W(false, bucket_emit_kernel, 2); break;
    }
#undef NPGE_LAUNCH_W
#undef NPGE_LAUNCH
}

void launch_found_or(const uint32_t* const* srcs, int n_src, uint32_t* dst, size_t bytes,
                     cudaStream_t s, uint32_t* launches) {
    FoundSources fs = {};
    fs.n = n_src < 8 ? n_src : 8;
    for (int r = 0; r < fs.n; r++) fs.src[r] = srcs[r];
    const size_t n_vec = bytes / 16;
    if (n_vec == 0) return;
    const size_t want = (n_vec + 255) / 256;
    found_or_kernel<<<uint32_t(want < 148 * 8 ? want : 148 * 8), 256, 0, s>>>(fs, dst, n_vec);
    *launches += 1;
}

void launch_bucket(const BucketArgs& a, bool similar, uint64_t max_records,
                   cudaStream_t s, cudaEvent_t* ev, uint32_t* launches) {
    launch_bucket_stage(a, kBucketUnits, max_records, s, launches);
    if (similar) {
        launch_bucket_stage(a, kBucketMark, max_records, s, launches);
        launch_bucket_stage(a, kBucketCount, max_records, s, launches);
    } else {
        launch_bucket_stage(a, kBucketSort, max_records, s, launches);
    }
    if (ev) cudaEventRecord(ev[0], s);
    launch_bucket_stage(a, kBucketScan, max_records, s, launches);
    launch_bucket_stage(a, similar ? kBucketEmitSimilar : kBucketEmit, max_records, s,
                        launches);
    launch_bucket_stage(a, kBucketFinal, max_records, s, launches);
}

}  // namespace npge_b200
