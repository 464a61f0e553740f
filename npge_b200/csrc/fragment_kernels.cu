// fragment_kernels.cu -- FragmentFinder: every exact
// occurrence of a pattern, on either strand, in the uploaded sequences.
//
// Reference: src/algo/FragmentFinder.cpp:42-171.  It rolls the same two hashes
// as AnchorFinder over every sequence and compares them with the pattern's
// hash, verifying by string compare when the hash cannot be trusted (pattern
// longer than 32, N in the pattern or in the window).  Net effect: window == P
// gives a fragment with ori = +1, revcomp(window) == P one with ori = -1, N
// matching only N.  Here the compare is done directly on the packed genome:
// 32 bases (one u64) + 32 N-mask bits per step, against the packed pattern and
// the packed reverse complement of the pattern.
//
// Two launches over the genome (count per tile, then ordered write) and one
// single-CTA scan between them; the output is in scan order (sequence by size
// descending, position ascending, forward before reverse), cut to max-matches,
// which is what a one-worker run of the reference produces.
#include "kernels.h"

namespace npge_b200 {

namespace {

constexpr int kFT = 512;  // threads per CTA; 8 consecutive positions per thread

__device__ __forceinline__ uint32_t mask_bits(const uint32_t* __restrict__ M, uint32_t rel,
                                              int len) {
    const uint32_t w = rel >> 5, s = rel & 31;
    uint32_t m = __funnelshift_r(M[w], M[w + 1], s);
    if (len < 32) m &= (1u << len) - 1;
    return m;
}

// does the text at position g equal the pattern of strand `strand`?
__device__ __forceinline__ bool equals_at(const FragmentArgs& a, uint32_t g, int strand) {
    const uint64_t* pk = a.pat_key + size_t(strand) * a.n_chunks;
    const uint32_t* pm = a.pat_mask + size_t(strand) * a.n_chunks;
    for (uint32_t c = 0; c < a.n_chunks; c++) {
        const int len = int(min(32u, a.pattern_len - 32u * c));
        const uint32_t rel = g + 32u * c;
        if (forward_key(a.packed, rel, len) != pk[c]) return false;
        if (mask_bits(a.nmask, rel, len) != pm[c]) return false;
    }
    return true;
}

// sequence of position g (scan-order index), via the chunk look-up table
__device__ __forceinline__ uint32_t seq_of(const FragmentArgs& a, uint32_t g) {
    const uint32_t q = g >> a.lut_shift;
    uint32_t lo = a.seq_lut[q], hi = a.seq_lut[q + 1] + 1;  // answer in [lo, hi)
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a.seq_start[mid] <= g) lo = mid; else hi = mid;
    }
    return lo;
}

// bit 0: forward match at g, bit 1: reverse match
__device__ __forceinline__ uint32_t matches_at(const FragmentArgs& a, uint64_t g) {
    if (g + a.pattern_len > a.n_positions) return 0;
    uint32_t m = 0;
    if (equals_at(a, uint32_t(g), 0)) m |= 1;
    if (equals_at(a, uint32_t(g), 1)) m |= 2;
    if (m) {  // the window must lie inside one sequence (a pattern with N could
              // otherwise match across the pad between two sequences)
        const uint32_t s = seq_of(a, uint32_t(g));
        if (g + a.pattern_len > a.seq_start[s + 1] - 1) m = 0;
    }
    return m;
}

__device__ __forceinline__ uint32_t block_sum_excl(uint32_t v, uint32_t* tmp, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) tmp[warp] = incl;
    __syncthreads();
    uint32_t base = 0, all = 0;
    for (int w = 0; w < kFT / 32; w++) {
        const uint32_t t = tmp[w];
        if (w < warp) base += t;
        all += t;
    }
    if (total) *total = all;
    return base + incl - v;
}

// MODE 0: tile_count[tile] = matches in the tile.  MODE 1: write them in order.
template <int MODE>
__global__ void __launch_bounds__(kFT)
fragment_kernel(FragmentArgs a) {
    __shared__ uint32_t tmp[kFT / 32];
    const uint64_t g0 = (uint64_t(blockIdx.x) * kFT + threadIdx.x) * 8;
    uint64_t first = 0;
    if (MODE == 1) {
        first = a.tile_offset[blockIdx.x];
        if (first >= a.max_matches) return;  // whole tile past the cut
    }
    uint32_t bits = 0, count = 0;  // 2 bits per position
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t m = matches_at(a, g0 + i);
        bits |= m << (2 * i);
        count += __popc(m);
    }
    if (MODE == 0) {
        uint32_t total;
        block_sum_excl(count, tmp, &total);
        if (threadIdx.x == 0) a.tile_count[blockIdx.x] = total;
        return;
    }
    uint64_t at = first + block_sum_excl(count, tmp, nullptr);
    for (int i = 0; i < 8 && bits; i++) {
        for (int strand = 0; strand < 2; strand++) {
            if (!((bits >> (2 * i + strand)) & 1)) continue;
            if (at < a.max_matches) {
                const uint32_t g = uint32_t(g0 + i);
                const uint32_t s = seq_of(a, g);
                a.r_seq[at] = a.seq_caller[s];
                a.r_pos[at] = g - uint32_t(a.seq_start[s]);
                a.r_ori[at] = strand == 0 ? 1 : -1;
            }
            at++;
        }
    }
}

// exclusive scan of the tile counts by one CTA; summary[0] = total
__global__ void __launch_bounds__(1024)
fragment_scan_kernel(FragmentArgs a, uint32_t n_tiles) {
    __shared__ uint64_t part[1024];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t b = min(threadIdx.x * per, n_tiles), e = min(b + per, n_tiles);
    uint64_t sum = 0;
    for (uint32_t i = b; i < e; i++) sum += a.tile_count[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t run = 0;
        for (int i = 0; i < 1024; i++) {
            const uint64_t v = part[i];
            part[i] = run;
            run += v;
        }
        a.summary[0] = run;
    }
    __syncthreads();
    uint64_t run = part[threadIdx.x];
    for (uint32_t i = b; i < e; i++) {
        a.tile_offset[i] = run;
        run += a.tile_count[i];
    }
}

}  // namespace

uint32_t fragment_tiles(uint64_t n_padded) {
    return uint32_t((n_padded + kFT * 8 - 1) / (kFT * 8));
}

// stage 0: count + scan (summary[0] = matches before the cut); stage 1: write
void launch_fragment_find(const FragmentArgs& a, int stage, cudaStream_t s,
                          uint32_t* launches) {
    const uint32_t tiles = fragment_tiles(a.n_padded);
    if (tiles == 0) return;
    if (stage == 0) {
        fragment_kernel<0><<<tiles, kFT, 0, s>>>(a);
        fragment_scan_kernel<<<1, 1024, 0, s>>>(a, tiles);
        *launches += 2;
    } else {
        fragment_kernel<1><<<tiles, kFT, 0, s>>>(a);
        *launches += 1;
    }
}

}  // namespace npge_b200
