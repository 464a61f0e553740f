// rank.cuh -- device building blocks shared by the partition passes and the
// bucket sort: TMA bulk loads, and the stable in-tile ranking of kSortItems
// items per thread by a radix digit.  sm_100a.
#pragma once

#include "common.cuh"

namespace npge_b200 {

constexpr int kRT = kSortThreads;       // threads per CTA of every ranking kernel
constexpr int kRW = kSortThreads / 32;  // warps per CTA
constexpr int kRI = kSortItems;         // items per thread

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_load(void* dst, const void* src,
                                          uint32_t bytes, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ void mbar_init(unsigned long long* bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}

// The packed-genome tile (kTile positions + 16 B of halo) of the scan kernels.
struct __align__(16) GenomeStage {
    uint32_t w[kTile / 16 + 4];
    uint32_t m[kTile / 32 + 4];
    unsigned long long bar;
};

// Thread 0 issues the copies; everybody waits.  One use per kernel (parity 0).
__device__ __forceinline__ void stage_genome_tile(GenomeStage& st,
                                                  const uint32_t* __restrict__ packed,
                                                  const uint32_t* __restrict__ nmask,
                                                  uint32_t tile) {
    constexpr uint32_t kBytesW = kTile / 4 + 16, kBytesM = kTile / 8 + 16;
    if (threadIdx.x == 0) {
        mbar_init(&st.bar);
        mbar_expect(&st.bar, kBytesW + kBytesM);
        bulk_load(st.w, packed + size_t(tile) * (kTile / 16), kBytesW, &st.bar);
        bulk_load(st.m, nmask + size_t(tile) * (kTile / 32), kBytesM, &st.bar);
    }
    __syncthreads();  // the init must be visible before anybody polls
    mbar_wait(&st.bar, 0);
}

template <int CAPBITS>
struct __align__(16) RankSmem {
    static constexpr int BINS = 1 << CAPBITS;
    uint16_t warp_hist[kRW][BINS];  // per-warp digit counts, then the first slot of (warp, digit)
    uint32_t tile_off[BINS];        // first slot of the digit in tile-sorted order
    uint32_t scan_tmp[kRW];
    uint32_t total;                 // items in the tile
};

// Rank of every item among the items of its own warp with the same digit, and
// the warp's digit counts in sm.warp_hist[warp].  FULL: all kRI rounds hold
// items (straight-line code); otherwise rounds >= `rounds` are skipped.
template <int CAPBITS, bool FULL>
__device__ __forceinline__ void warp_rank(RankSmem<CAPBITS>& sm,
                                          const uint32_t (&digit)[kRI], int rounds,
                                          uint32_t (&rank)[kRI]) {
    constexpr int BINS = 1 << CAPBITS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1;
    uint16_t* wh = sm.warp_hist[warp];
    uint32_t peers[kRI];
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        peers[i] = (FULL || i < rounds) ? __match_any_sync(0xffffffffu, digit[i]) : 0u;
    }
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        rank[i] = 0;
        if (FULL || i < rounds) {
            const uint32_t d = digit[i];
            const int leader = __ffs(peers[i]) - 1;
            uint32_t old = 0;
            if (lane == leader && d < BINS) {
                old = wh[d];
                wh[d] = uint16_t(old + __popc(peers[i]));
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            rank[i] = old + __popc(peers[i] & lt);
            __syncwarp();
        }
    }
}

// Stable rank of the tile's items by digit.  Item i of lane l of warp w is
// element w * 32 * kRI + i * 32 + l of the tile (ascending input order);
// digit == BINS means "no item".  `rounds` (warp-uniform) = item rounds of this
// warp that hold at least one item.  All threads of the CTA must call it.
// On return: slot[i] = index of the item in tile-sorted order, sm.tile_off and
// sm.total are valid, (c0, c1) = counts of the digits the thread owns
// (digits 2 tid, 2 tid + 1 when BINS = 1024; digit tid < BINS otherwise).
//
// All match.any of a thread are issued back to back: the instruction has
// ~200 cycles of latency on sm_100a (profiles/r01b_*), so the rounds must
// overlap; only the counter updates are ordered.
// tile_rank_abs: the same contract, but the per-warp counters end up holding the
// ABSOLUTE first slot of (warp, digit), so that an item needs one lookup instead of
// two.  Costs eight more registers between the barriers: used where registers are
// not the occupancy limit (the bucket sort; measured -0.04 ms there, +0.04 ms in
// the partition passes).
template <int CAPBITS>
__device__ __forceinline__ void tile_rank_abs(RankSmem<CAPBITS>& sm,
                                          const uint32_t (&digit)[kRI], int rounds,
                                          uint32_t (&slot)[kRI], uint32_t& c0,
                                          uint32_t& c1) {
    constexpr int BINS = 1 << CAPBITS;
    constexpr int PER = BINS / kRT > 0 ? BINS / kRT : 1;
    static_assert(PER == 1 || PER == 2, "one or two digits per thread");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        uint4* wh = reinterpret_cast<uint4*>(sm.warp_hist[warp]);
#pragma unroll
        for (int i = 0; i < BINS / 256; i++) wh[lane + 32 * i] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    uint32_t rank[kRI];
    if (rounds == kRI) {
        warp_rank<CAPBITS, true>(sm, digit, rounds, rank);
    } else {
        warp_rank<CAPBITS, false>(sm, digit, rounds, rank);
    }
    // digit and rank share a register from here on (digit <= 1024, rank <= 256):
    // a caller that does not read digit[] after the call frees its registers
    uint32_t dr[kRI];
#pragma unroll
    for (int i = 0; i < kRI; i++) dr[i] = digit[i] | (rank[i] << 16);
    __syncthreads();
    // column scan: the thread owning a digit walks the kRW per-warp counts and
    // keeps their exclusive prefixes in registers (two 16-bit values per word);
    // once the digit's first slot is known it writes back, per warp, the ABSOLUTE
    // first slot of (warp, digit): one lookup per item afterwards
    c0 = 0;
    c1 = 0;
    uint32_t pre[PER == 2 ? kRW : kRW / 2];
    if (PER == 2) {
        uint32_t run = 0;  // two packed 16-bit running sums (each <= kTile)
#pragma unroll
        for (int w = 0; w < kRW; w++) {
            pre[w] = run;
            run += reinterpret_cast<const uint32_t*>(sm.warp_hist[w])[tid];
        }
        c0 = run & 0xffff;
        c1 = run >> 16;
    } else if (tid < BINS) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kRW; w += 2) {
            const uint32_t v0 = sm.warp_hist[w][tid], v1 = sm.warp_hist[w + 1][tid];
            pre[w / 2] = run | ((run + v0) << 16);
            run += v0 + v1;
        }
        c0 = run;
    }
    const uint32_t mine = c0 + c1;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) sm.scan_tmp[warp] = incl;
    __syncthreads();
    // sum of the totals of the warps before this one: one load + a butterfly
    uint32_t wbase = (lane < kRW && lane < warp) ? sm.scan_tmp[lane] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wbase += __shfl_xor_sync(0xffffffffu, wbase, o);
    const uint32_t toff0 = wbase + incl - mine;
    if (PER == 2) {
        sm.tile_off[2 * tid] = toff0;
        sm.tile_off[2 * tid + 1] = toff0 + c0;
        const uint32_t add = toff0 | ((toff0 + c0) << 16);  // no carry: slots <= kTile
#pragma unroll
        for (int w = 0; w < kRW; w++) {
            reinterpret_cast<uint32_t*>(sm.warp_hist[w])[tid] = pre[w] + add;
        }
    } else if (tid < BINS) {
        sm.tile_off[tid] = toff0;
#pragma unroll
        for (int w = 0; w < kRW; w += 2) {
            sm.warp_hist[w][tid] = uint16_t(toff0 + (pre[w / 2] & 0xffff));
            sm.warp_hist[w + 1][tid] = uint16_t(toff0 + (pre[w / 2] >> 16));
        }
    }
    if (tid == kRT - 1) sm.total = toff0 + mine;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t d = dr[i] & 0xffffu;
        slot[i] = d < BINS ? sm.warp_hist[warp][d] + (dr[i] >> 16) : 0u;
    }
}


template <int CAPBITS>
__device__ __forceinline__ void tile_rank(RankSmem<CAPBITS>& sm,
                                          const uint32_t (&digit)[kRI], int rounds,
                                          uint32_t (&slot)[kRI], uint32_t& c0,
                                          uint32_t& c1) {
    constexpr int BINS = 1 << CAPBITS;
    constexpr int PER = BINS / kRT > 0 ? BINS / kRT : 1;
    static_assert(PER == 1 || PER == 2, "one or two digits per thread");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        uint4* wh = reinterpret_cast<uint4*>(sm.warp_hist[warp]);
#pragma unroll
        for (int i = 0; i < BINS / 256; i++) wh[lane + 32 * i] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    uint32_t rank[kRI];
    if (rounds == kRI) {
        warp_rank<CAPBITS, true>(sm, digit, rounds, rank);
    } else {
        warp_rank<CAPBITS, false>(sm, digit, rounds, rank);
    }
    __syncthreads();
    c0 = 0;
    c1 = 0;
    if (PER == 2) {
        uint32_t run = 0;  // two packed 16-bit running sums (each <= kTile)
#pragma unroll
        for (int w = 0; w < kRW; w++) {
            uint32_t* p = reinterpret_cast<uint32_t*>(sm.warp_hist[w]) + tid;
            uint32_t v = *p;
            *p = run;
            run += v;
        }
        c0 = run & 0xffff;
        c1 = run >> 16;
    } else if (tid < BINS) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kRW; w++) {
            uint32_t v = sm.warp_hist[w][tid];
            sm.warp_hist[w][tid] = uint16_t(run);
            run += v;
        }
        c0 = run;
    }
    const uint32_t mine = c0 + c1;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) sm.scan_tmp[warp] = incl;
    __syncthreads();
    // sum of the totals of the warps before this one: one load + a butterfly
    uint32_t wbase = (lane < kRW && lane < warp) ? sm.scan_tmp[lane] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wbase += __shfl_xor_sync(0xffffffffu, wbase, o);
    const uint32_t toff0 = wbase + incl - mine;
    if (PER == 2) {
        sm.tile_off[2 * tid] = toff0;
        sm.tile_off[2 * tid + 1] = toff0 + c0;
    } else if (tid < BINS) {
        sm.tile_off[tid] = toff0;
    }
    if (tid == kRT - 1) sm.total = toff0 + mine;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        const uint32_t d = digit[i];
        slot[i] = d < BINS ? sm.tile_off[d] + sm.warp_hist[warp][d] + rank[i] : 0u;
    }
}


}  // namespace npge_b200
