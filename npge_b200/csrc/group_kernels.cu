// group_kernels.cu -- run-length grouping of the sorted records and compaction
// into the flat anchor result; the anchor-similar rule.  sm_100a.
//
// Replaces fragmenttg_postprocess (src/algo/AnchorFinder.cpp:356-391): the
// linear pass that opens a Block the first time a hash repeats, and
// FoundFragment::make_fragment (:240-246).  Because the records are sorted by
// (key, position), "this record belongs to a block" is a purely local test
// (equal to the previous or the next key), so the pass is one device-wide scan
// with decoupled look-back over tiles of sorted records.
#include "kernels.h"

namespace npge_b200 {

namespace {

constexpr int kGThreads = 256;
constexpr int kGItems = 8;
constexpr int kGTile = kGThreads * kGItems;  // 2048 sorted records per tile
constexpr int kPartials = 64;

constexpr uint64_t kSFlagAggregate = 1ull << 62;
constexpr uint64_t kSFlagInclusive = 2ull << 62;
constexpr uint64_t kSValueMask = (1ull << 62) - 1;
// scan value: blocks << 31 | fragments (both < 2^31)

__device__ __forceinline__ uint64_t ld_relaxed64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_relaxed64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint64_t key_at(const GroupArgs& a, uint64_t i) {
    uint64_t k = a.klo[i];
    if (a.khi) k |= uint64_t(a.khi[i]) << 32;
    return k;
}

// SortedVector::has_elem (src/util/SortedVector.hpp:108-111) on the device
__device__ __forceinline__ bool is_used(const GroupArgs& a, uint64_t key) {
    uint64_t lo = 0, hi = a.n_used;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (a.used[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo < a.n_used && a.used[lo] == key;
}

// index of the first record of the group record i belongs to (gallop back)
__device__ uint64_t find_head(const GroupArgs& a, uint64_t i, uint64_t key) {
    uint64_t hi = i;  // key_at(hi) == key
    uint64_t step = 1;
    uint64_t lo;
    while (true) {
        if (hi < step) { lo = 0; break; }
        if (key_at(a, hi - step) != key) { lo = hi - step + 1; break; }
        hi -= step;
        step <<= 1;
    }
    // first index in [lo, hi] whose key == key (keys ascending)
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (key_at(a, mid) == key) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// ---- anchor-similar ----------------------------------------------------------
// found(p): the window's hash is not used and was seen at an earlier scan
// position (BloomTask::test_and_add, AnchorFinder.cpp:170-182, exact filter).
// Device positions are laid out in scan order, so "earlier" = smaller global
// position = every member of a sorted group but the first.
__global__ void __launch_bounds__(256)
mark_found_kernel(GroupArgs a) {
    const uint64_t n = a.counters[kCntValid];
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n;
            i += stride) {
        if (i == 0) continue;
        uint64_t key = key_at(a, i);
        if (key_at(a, i - 1) != key) continue;
        if (a.n_used && is_used(a, key)) continue;
        uint32_t g = a.pay[i] >> 1;
        atomicOr(a.found_bits + (g >> 5), 1u << (g & 31));
    }
}

// candidate iff found(p) and not found(p-1) (`hash_found && !prev_`, :176);
// prev_ resets per sequence (:191): position p-1 of a sequence start is the
// pad, which never holds a record.  A candidate selects its whole group.
__global__ void __launch_bounds__(256)
mark_selected_kernel(GroupArgs a) {
    const uint64_t n = a.counters[kCntValid];
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n;
            i += stride) {
        if (i == 0) continue;
        uint64_t key = key_at(a, i);
        if (key_at(a, i - 1) != key) continue;
        uint32_t g = a.pay[i] >> 1;
        if (!((a.found_bits[g >> 5] >> (g & 31)) & 1)) continue;  // used hash
        if (g > 0) {
            uint32_t q = g - 1;
            if ((a.found_bits[q >> 5] >> (q & 31)) & 1) continue;
        }
        a.sel[find_head(a, i - 1, key)] = 1;
    }
}

// ---- grouping + compaction -----------------------------------------------------
struct GroupSmem {
    uint32_t klo[kGTile + 2];
    uint32_t khi[kGTile + 2];
    uint64_t warp_tot[kGThreads / 32];
    uint32_t warp_head[kGThreads / 32];
    uint64_t red[3][kGThreads / 32];
    uint64_t tile_prefix;
    uint64_t carry_head;
    uint32_t ticket;
};

template <bool SIMILAR>
__global__ void __launch_bounds__(kGThreads)
group_emit_kernel(GroupArgs a) {
    __shared__ GroupSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) sm.ticket = atomicAdd(a.tickets, 1u);
    __syncthreads();
    const uint64_t tile = sm.ticket;
    const uint64_t n = a.counters[kCntValid];
    const uint64_t base = tile * kGTile;
    if (base >= n) return;
    const bool has_hi = a.khi != nullptr;
    for (int j = tid; j < kGTile + 2; j += kGThreads) {
        int64_t idx = int64_t(base) + j - 1;
        bool in = idx >= 0 && uint64_t(idx) < n;
        sm.klo[j] = in ? a.klo[idx] : 0;
        sm.khi[j] = (in && has_hi) ? a.khi[idx] : 0;
    }
    if (SIMILAR && tid == 0) {
        // head of the group that contains the tile's first record
        uint64_t key0 = key_at(a, base);
        sm.carry_head = (base > 0 && key_at(a, base - 1) == key0)
                        ? find_head(a, base - 1, key0) : base;
    }
    __syncthreads();
    auto K = [&](int j) { return (uint64_t(sm.khi[j]) << 32) | sm.klo[j]; };
    // per item flags
    uint32_t in_block = 0, is_head = 0;  // bit e = item e
    uint64_t keys[kGItems];
    uint32_t last_head = 0;  // local index + 1 of the thread's last head
#pragma unroll
    for (int e = 0; e < kGItems; e++) {
        int li = tid * kGItems + e;
        uint64_t i = base + li;
        keys[e] = K(li + 1);
        if (i < n) {
            bool prev_eq = i > 0 && K(li) == keys[e];
            bool next_eq = i + 1 < n && K(li + 2) == keys[e];
            if (!prev_eq) {
                is_head |= 1u << e;
                last_head = li + 1;
            }
            if (prev_eq || next_eq) in_block |= 1u << e;
        }
    }
    uint32_t selected = 0;
    if (SIMILAR) {
        // exclusive max-scan of last_head over the threads of the tile
        uint32_t incl = last_head;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl = max(incl, u);
        }
        if (lane == 31) sm.warp_head[warp] = incl;
        uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) excl = 0;
        __syncthreads();
        for (int w = 0; w < warp; w++) excl = max(excl, sm.warp_head[w]);
        uint64_t head = excl ? base + excl - 1 : sm.carry_head;
#pragma unroll
        for (int e = 0; e < kGItems; e++) {
            if (is_head & (1u << e)) head = base + tid * kGItems + e;
            if ((in_block >> e) & 1) {
                if (a.sel[head]) selected |= 1u << e;
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < kGItems; e++) {
            if ((in_block >> e) & 1) {
                if (!(a.n_used && is_used(a, keys[e]))) selected |= 1u << e;
            }
        }
    }
    // block scan of (blocks << 31 | fragments)
    uint64_t mine = uint64_t(__popc(selected)) |
                    (uint64_t(__popc(selected & is_head)) << 31);
    uint64_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) sm.warp_tot[warp] = incl;
    __syncthreads();
    uint64_t excl = incl - mine;
    uint64_t tile_total = 0;
#pragma unroll
    for (int w = 0; w < kGThreads / 32; w++) {
        if (w < warp) excl += sm.warp_tot[w];
        tile_total += sm.warp_tot[w];
    }
    if (tid == 0) {
        uint64_t* my = a.scan_status + tile;
        uint64_t prefix = 0;
        if (tile == 0) {
            st_relaxed64(my, kSFlagInclusive | tile_total);
        } else {
            st_relaxed64(my, kSFlagAggregate | tile_total);
            const uint64_t* p = my - 1;
            while (true) {
                uint64_t v = ld_relaxed64(p);
                if ((v >> 62) == 0) continue;
                prefix += v & kSValueMask;
                if (v & kSFlagInclusive) break;
                p -= 1;
            }
            st_relaxed64(my, kSFlagInclusive | (prefix + tile_total));
        }
        sm.tile_prefix = prefix;
        if (base + kGTile >= n) {  // last tile: totals before truncation
            uint64_t tot = prefix + tile_total;
            a.summary[kSumFragments] = tot & 0x7fffffffull;
            a.summary[kSumBlocks] = tot >> 31;
        }
    }
    __syncthreads();
    excl += sm.tile_prefix;
    uint64_t frag_rank = excl & 0x7fffffffull;
    uint64_t block_rank = excl >> 31;
    uint64_t d0 = 0, d1 = 0, nb = 0;
    const uint64_t limit = a.max_fragments;
#pragma unroll
    for (int e = 0; e < kGItems; e++) {
        if (!((selected >> e) & 1)) continue;
        uint64_t i = base + tid * kGItems + e;
        bool head = (is_head >> e) & 1;
        if (frag_rank < limit) {
            uint32_t pay = a.pay[i];
            uint64_t g = pay >> 1;
            // scan-order sequence holding g (binary search over the starts)
            uint32_t lo = 0, hi = a.n_seqs;
            while (hi - lo > 1) {
                uint32_t mid = (lo + hi) >> 1;
                if (a.seq_start[mid] <= g) lo = mid; else hi = mid;
            }
            uint32_t seq = a.seq_caller[lo];
            uint32_t pos = uint32_t(g - a.seq_start[lo]);
            a.r_seq[frag_rank] = seq;
            a.r_pos[frag_rank] = pos;
            a.r_ori[frag_rank] = (pay & 1) ? 1 : -1;
            uint64_t t = (uint64_t(seq) << 33) | (uint64_t(pos) << 1) | (pay & 1);
            uint64_t x, y;
            digest_terms(t, keys[e], x, y);
            d0 += x;
            d1 += y;
            if (head) {
                a.r_block_off[block_rank] = frag_rank;
                a.r_block_hash[block_rank] = keys[e];
                nb++;
            }
            if (frag_rank + 1 == limit) *a.cut_index = uint32_t(i);
        }
        frag_rank++;
        block_rank += head ? 1 : 0;
    }
    // tile reduction of digest and emitted block count -> 64 partial slots
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        d0 += __shfl_xor_sync(0xffffffffu, d0, o);
        d1 += __shfl_xor_sync(0xffffffffu, d1, o);
        nb += __shfl_xor_sync(0xffffffffu, nb, o);
    }
    if (lane == 0) {
        sm.red[0][warp] = d0;
        sm.red[1][warp] = d1;
        sm.red[2][warp] = nb;
    }
    __syncthreads();
    if (tid < 3) {
        uint64_t v = 0;
        for (int w = 0; w < kGThreads / 32; w++) v += sm.red[tid][w];
        if (v) {
            atomicAdd(reinterpret_cast<unsigned long long*>(
                          a.summary + kSumWords + (tile % kPartials) * 4 + tid),
                      (unsigned long long)v);
        }
    }
}

__global__ void finalize_kernel(GroupArgs a) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint64_t d0 = 0, d1 = 0, nb = 0;
    for (int s = 0; s < kPartials; s++) {
        d0 += a.summary[kSumWords + s * 4 + 0];
        d1 += a.summary[kSumWords + s * 4 + 1];
        nb += a.summary[kSumWords + s * 4 + 2];
    }
    uint64_t total = a.summary[kSumFragments];
    uint64_t emitted = total < a.max_fragments ? total : a.max_fragments;
    a.summary[kSumDigest0] = d0;
    a.summary[kSumDigest1] = d1;
    a.summary[kSumEmitted] = emitted;
    a.summary[kSumBlocksEmitted] = nb;
    a.r_block_off[nb] = emitted;
}

}  // namespace

void launch_similar(const GroupArgs& a, uint64_t n_padded, cudaStream_t s,
                    uint32_t* launches) {
    if (a.max_records == 0) return;
    uint64_t want = (a.max_records + 255) / 256;
    uint32_t blocks = uint32_t(want < 148 * 16 ? want : 148 * 16);
    mark_found_kernel<<<blocks, 256, 0, s>>>(a);
    mark_selected_kernel<<<blocks, 256, 0, s>>>(a);
    *launches += 2;
}

void launch_group_emit(const GroupArgs& a, bool similar, cudaStream_t s) {
    uint32_t tiles = uint32_t((a.max_records + kGTile - 1) / kGTile);
    if (tiles) {
        if (similar) {
            group_emit_kernel<true><<<tiles, kGThreads, 0, s>>>(a);
        } else {
            group_emit_kernel<false><<<tiles, kGThreads, 0, s>>>(a);
        }
    }
    finalize_kernel<<<1, 32, 0, s>>>(a);
}

// layout helpers for the host side
size_t group_summary_words() { return kSumWords + kPartials * 4; }
size_t group_scan_tiles(uint64_t max_records) {
    return size_t((max_records + kGTile - 1) / kGTile) + 1;
}

}  // namespace npge_b200
