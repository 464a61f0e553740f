// sort_kernels.cu -- hand-written LSD radix sort of (canonical k-mer, position,
// strand) records, one-sweep style: every pass reads its input once and
// scatters it once; tile prefixes are resolved by decoupled look-back.
//
// Pass 0 never reads records: it regenerates the keys from the 2-bit packed
// genome (staged into shared memory with a TMA bulk copy) and scatters them by
// the lowest digit, so the (k-mer, position) list of the reference's pass 2
// (FragmentTask::test_and_push, src/algo/AnchorFinder.cpp:303-312) is never
// written in unsorted form.  Replaces std::sort of FoundFragments
// (AnchorFinder.cpp:234-238,358; src/util/SortedVector.hpp:47-49).
//
// Stability: tiles are taken in ticket order, items inside a tile are ranked in
// (warp, item, lane) order which is ascending input order, so after the last
// pass records are ordered by (key, global position).
#include "kernels.h"

namespace npge_b200 {

namespace {

struct __align__(16) TileSmem {
    uint32_t stage_w[kTile / 16 + 4];  // packed genome tile + 16 B halo (pass 0)
    uint32_t stage_m[kTile / 32 + 4];  // N-mask tile + 16 B halo       (pass 0)
    uint32_t warp_hist[kWarps][kRadix];
    uint32_t tile_off[kRadix];
    uint32_t bin_base[kRadix];
    uint32_t scan_tmp[kWarps];
    uint32_t klo[kTile];
    uint32_t khi[kTile];
    uint32_t pay[kTile];
    unsigned long long bar;
    uint32_t ticket;
    uint32_t total;
};

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// TMA 1-D bulk copy global -> shared, completion on an mbarrier (UBLKCP).
__device__ __forceinline__ void bulk_load(void* dst, const void* src,
                                          uint32_t bytes, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ uint32_t digit_of(uint32_t lo, uint32_t hi, int pass) {
    return pass < 4 ? ((lo >> (8 * pass)) & 0xff) : ((hi >> (8 * (pass - 4))) & 0xff);
}

// Ranks the tile's items by `digit` (256 = no item), resolves the tile's global
// digit bases by look-back and writes the items out.  All threads call it.
template <bool HAS_HI>
__device__ __forceinline__ void rank_and_scatter(
        TileSmem& sm, const uint32_t (&lo)[kItemsPerThread],
        const uint32_t (&hi)[kItemsPerThread],
        const uint32_t (&pay)[kItemsPerThread],
        const uint32_t (&digit)[kItemsPerThread], uint32_t tile, int pass,
        const uint32_t* __restrict__ hist_base, uint32_t* __restrict__ status,
        uint32_t* __restrict__ klo_out, uint32_t* __restrict__ khi_out,
        uint32_t* __restrict__ pay_out) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1;
#pragma unroll
    for (int i = 0; i < kRadix / 32; i++) sm.warp_hist[warp][lane + 32 * i] = 0;
    __syncwarp();
    uint32_t rank[kItemsPerThread];
#pragma unroll
    for (int i = 0; i < kItemsPerThread; i++) {
        uint32_t d = digit[i];
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && d < kRadix) old = sm.warp_hist[warp][d];
        old = __shfl_sync(0xffffffffu, old, leader);
        if (lane == leader && d < kRadix) sm.warp_hist[warp][d] = old + __popc(peers);
        rank[i] = old + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    // thread d owns digit d: warp counts -> warp offsets, tile count
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) {
        uint32_t c = sm.warp_hist[w][tid];
        sm.warp_hist[w][tid] = count;
        count += c;
    }
    // exclusive scan of the 256 counts
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) sm.scan_tmp[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) wbase += (w < warp) ? sm.scan_tmp[w] : 0;
    uint32_t toff = wbase + incl - count;
    sm.tile_off[tid] = toff;
    if (tid == kRadix - 1) sm.total = toff + count;
    // decoupled look-back over earlier tiles of this pass, digit = tid
    uint32_t* my = status + size_t(tile) * kRadix + tid;
    uint32_t excl = 0;
    if (tile == 0) {
        st_relaxed(my, kFlagInclusive | count);
    } else {
        st_relaxed(my, kFlagAggregate | count);
        const uint32_t* p = my - kRadix;
        while (true) {
            uint32_t v = ld_relaxed(p);
            if ((v >> 30) == 0) continue;
            excl += v & kValueMask;
            if (v & kFlagInclusive) break;
            p -= kRadix;
        }
        st_relaxed(my, kFlagInclusive | (excl + count));
    }
    sm.bin_base[tid] = hist_base[pass * kRadix + tid] + excl - toff;
    __syncthreads();
    // place the items in tile-sorted order in shared memory
#pragma unroll
    for (int i = 0; i < kItemsPerThread; i++) {
        uint32_t d = digit[i];
        if (d < kRadix) {
            uint32_t slot = sm.tile_off[d] + sm.warp_hist[warp][d] + rank[i];
            sm.klo[slot] = lo[i];
            if (HAS_HI) sm.khi[slot] = hi[i];
            sm.pay[slot] = pay[i];
        }
    }
    __syncthreads();
    const uint32_t total = sm.total;
    for (uint32_t j = tid; j < total; j += kTileThreads) {
        uint32_t l = sm.klo[j];
        uint32_t h = HAS_HI ? sm.khi[j] : 0;
        uint32_t idx = sm.bin_base[digit_of(l, h, pass)] + j;
        klo_out[idx] = l;
        if (HAS_HI) khi_out[idx] = h;
        pay_out[idx] = sm.pay[j];
    }
}

__device__ __forceinline__ uint32_t take_ticket(TileSmem& sm, uint32_t* counter) {
    if (threadIdx.x == 0) sm.ticket = atomicAdd(counter, 1u);
    __syncthreads();
    return sm.ticket;
}

// ---- pass 0: generate + scatter -------------------------------------------------
template <bool HAS_HI>
__global__ void __launch_bounds__(kTileThreads)
sort_pass0_kernel(const uint32_t* __restrict__ packed,
                  const uint32_t* __restrict__ nmask, int k,
                  const uint32_t* __restrict__ hist_base,
                  uint32_t* __restrict__ status, uint32_t* __restrict__ counters,
                  uint32_t* __restrict__ klo_out, uint32_t* __restrict__ khi_out,
                  uint32_t* __restrict__ pay_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TileSmem& sm = *reinterpret_cast<TileSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&sm.bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const uint32_t tile = take_ticket(sm, counters + kCntTicket0);
    constexpr uint32_t kBytesW = kTile / 4 + 16, kBytesM = kTile / 8 + 16;
    if (tid == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                         smem_u32(&sm.bar)),
                     "r"(kBytesW + kBytesM)
                     : "memory");
        bulk_load(sm.stage_w, packed + size_t(tile) * (kTile / 16), kBytesW, &sm.bar);
        bulk_load(sm.stage_m, nmask + size_t(tile) * (kTile / 32), kBytesM, &sm.bar);
    }
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(smem_u32(&sm.bar))
                : "memory");
        }
    }
    uint32_t lo[kItemsPerThread], hi[kItemsPerThread], pay[kItemsPerThread],
             digit[kItemsPerThread];
    const uint32_t gbase = tile * uint32_t(kTile);
#pragma unroll
    for (int i = 0; i < kItemsPerThread; i++) {
        uint32_t rel = warp * (32 * kItemsPerThread) + i * 32 + lane;
        bool bad = window_has_n(sm.stage_m, rel, k);
        uint64_t dir = forward_key(sm.stage_w, rel, k);
        uint64_t rev = revcomp_key(dir, k);
        uint64_t h = dir < rev ? dir : rev;
        lo[i] = uint32_t(h);
        hi[i] = uint32_t(h >> 32);
        // payload: global position << 1 | (ori == +1), ori = +1 iff h == dir
        pay[i] = ((gbase + rel) << 1) | (h == dir ? 1u : 0u);
        digit[i] = bad ? uint32_t(kRadix) : (lo[i] & 0xff);
    }
    rank_and_scatter<HAS_HI>(sm, lo, hi, pay, digit, tile, 0, hist_base, status,
                             klo_out, khi_out, pay_out);
}

// ---- pass p >= 1: read + scatter ---------------------------------------------------
template <bool HAS_HI>
__global__ void __launch_bounds__(kTileThreads)
sort_pass_kernel(int pass, const uint32_t* __restrict__ klo_in,
                 const uint32_t* __restrict__ khi_in,
                 const uint32_t* __restrict__ pay_in,
                 const uint32_t* __restrict__ hist_base,
                 uint32_t* __restrict__ status, uint32_t* __restrict__ counters,
                 uint32_t* __restrict__ klo_out, uint32_t* __restrict__ khi_out,
                 uint32_t* __restrict__ pay_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TileSmem& sm = *reinterpret_cast<TileSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = take_ticket(sm, counters + kCntTicket0 + pass);
    const uint32_t n = counters[kCntValid];
    const uint64_t tbase = uint64_t(tile) * kTile;
    if (tbase >= n) return;
    uint32_t lo[kItemsPerThread], hi[kItemsPerThread], pay[kItemsPerThread],
             digit[kItemsPerThread];
#pragma unroll
    for (int i = 0; i < kItemsPerThread; i++) {
        uint64_t idx = tbase + warp * (32 * kItemsPerThread) + i * 32 + lane;
        bool in = idx < n;
        lo[i] = in ? klo_in[idx] : 0;
        hi[i] = (HAS_HI && in) ? khi_in[idx] : 0;
        pay[i] = in ? pay_in[idx] : 0;
        digit[i] = in ? digit_of(lo[i], hi[i], pass) : uint32_t(kRadix);
    }
    rank_and_scatter<HAS_HI>(sm, lo, hi, pay, digit, tile, pass, hist_base,
                             status, klo_out, khi_out, pay_out);
}

template <typename K>
void set_smem(K kernel) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         int(sizeof(TileSmem)));
}

}  // namespace

void launch_sort_pass0(const uint32_t* packed, const uint32_t* nmask,
                       uint64_t n_padded, int k, const uint32_t* hist_base,
                       uint32_t* status, uint32_t* counters, uint32_t* klo_out,
                       uint32_t* khi_out, uint32_t* pay_out, cudaStream_t s) {
    uint32_t tiles = uint32_t(n_padded / kTile);
    if (tiles == 0) return;
    if (k > 16) {
        set_smem(sort_pass0_kernel<true>);
        sort_pass0_kernel<true><<<tiles, kTileThreads, sizeof(TileSmem), s>>>(
            packed, nmask, k, hist_base, status, counters, klo_out, khi_out, pay_out);
    } else {
        set_smem(sort_pass0_kernel<false>);
        sort_pass0_kernel<false><<<tiles, kTileThreads, sizeof(TileSmem), s>>>(
            packed, nmask, k, hist_base, status, counters, klo_out, khi_out, pay_out);
    }
}

void launch_sort_pass(int pass, uint64_t max_records, const uint32_t* klo_in,
                      const uint32_t* khi_in, const uint32_t* pay_in,
                      const uint32_t* hist_base, uint32_t* status,
                      uint32_t* counters, uint32_t* klo_out, uint32_t* khi_out,
                      uint32_t* pay_out, bool has_hi, cudaStream_t s) {
    uint32_t tiles = uint32_t((max_records + kTile - 1) / kTile);
    if (tiles == 0) return;
    if (has_hi) {
        set_smem(sort_pass_kernel<true>);
        sort_pass_kernel<true><<<tiles, kTileThreads, sizeof(TileSmem), s>>>(
            pass, klo_in, khi_in, pay_in, hist_base, status, counters, klo_out,
            khi_out, pay_out);
    } else {
        set_smem(sort_pass_kernel<false>);
        sort_pass_kernel<false><<<tiles, kTileThreads, sizeof(TileSmem), s>>>(
            pass, klo_in, khi_in, pay_in, hist_base, status, counters, klo_out,
            khi_out, pay_out);
    }
}

}  // namespace npge_b200
