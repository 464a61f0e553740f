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
// Digits are up to 10 bits wide (SortPlan): 4 passes over HBM at k = 20.
// Ranking inside a tile is warp-synchronous: all match.any of a thread's items
// are issued back to back (the instruction has ~200 cycles of latency on
// sm_100a, ncu capture profiles/r01b), then the per-warp digit counters are
// updated in item order.
//
// Stability: tiles are taken in ticket order, items inside a tile are ranked in
// (warp, item, lane) order which is ascending input order, so after the last
// pass records are ordered by (key, global position).
#include "kernels.h"

namespace npge_b200 {

namespace {

__device__ int g_exp_nochain = 0;  // experiment: timing without the look-back chain

constexpr int kST = kSortThreads;
constexpr int kSI = kSortItems;
constexpr int kSW = kSortThreads / 32;

template <int CAPBITS>
struct __align__(16) TileSmem {
    static constexpr int BINS = 1 << CAPBITS;
    uint32_t stage_w[kTile / 16 + 4];  // packed genome tile + 16 B halo (pass 0)
    uint32_t stage_m[kTile / 32 + 4];  // N-mask tile + 16 B halo       (pass 0)
    uint16_t warp_hist[kSW][BINS];     // per-warp digit counts -> warp offsets
    uint32_t tile_off[BINS];           // first slot of the digit in the tile
    uint32_t bin_base[BINS];           // global index of slot 0 of the digit
    uint32_t scan_tmp[kSW];
    uint32_t klo[kTile];
    uint32_t khi[kTile];
    uint32_t pay[kTile];
    unsigned long long bar;
    uint32_t ticket;
    uint32_t total;
};

__device__ __forceinline__ uint2 ld_relaxed2(const uint32_t* p) {
    uint2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];"
                 : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_relaxed2(uint32_t* p, uint32_t a, uint32_t b) {
    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a),
                 "r"(b) : "memory");
}

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

__device__ __forceinline__ uint32_t digit_of(uint32_t lo, uint32_t hi, int shift,
                                             uint32_t mask) {
    // bits [shift, shift + width) of hi:lo; shift < 64
    uint32_t v = shift >= 32 ? (hi >> (shift - 32))
                             : __funnelshift_r(lo, hi, shift);
    return v & mask;
}

// Ranks the tile's items by `digit` (BINS = no item), resolves the tile's global
// digit bases by look-back and writes the items out.  All threads call it.
template <int CAPBITS, bool HAS_HI>
__device__ __forceinline__ void rank_and_scatter(
        TileSmem<CAPBITS>& sm, const uint32_t (&lo)[kSI], const uint32_t (&hi)[kSI],
        const uint32_t (&pay)[kSI], const uint32_t (&digit)[kSI], uint32_t tile,
        int shift, uint32_t mask, const uint32_t* __restrict__ hist_base,
        uint32_t* __restrict__ status, uint32_t* __restrict__ klo_out,
        uint32_t* __restrict__ khi_out, uint32_t* __restrict__ pay_out) {
    constexpr int BINS = 1 << CAPBITS;
    constexpr int PER = BINS / kST > 0 ? BINS / kST : 1;  // digits per thread
    static_assert(PER == 1 || PER == 2, "one or two digits per thread");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1;
    {
        uint32_t* wh = reinterpret_cast<uint32_t*>(sm.warp_hist[warp]);
#pragma unroll
        for (int i = 0; i < BINS / 64; i++) wh[lane + 32 * i] = 0;
    }
    __syncwarp();
    // all matches first: independent, so their latencies overlap
    uint32_t peers[kSI];
#pragma unroll
    for (int i = 0; i < kSI; i++) peers[i] = __match_any_sync(0xffffffffu, digit[i]);
    uint32_t rank[kSI];
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        const uint32_t d = digit[i];
        const int leader = __ffs(peers[i]) - 1;
        uint32_t old = 0;
        if (lane == leader && d < BINS) {
            old = sm.warp_hist[warp][d];
            sm.warp_hist[warp][d] = uint16_t(old + __popc(peers[i]));
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers[i] & lt);
        __syncwarp();
    }
    __syncthreads();
    // thread t owns digits PER*t .. PER*t+PER-1: warp counts -> warp offsets
    uint32_t c0 = 0, c1 = 0;
    if (PER == 2) {
        uint32_t run = 0;  // two packed 16-bit running sums (each <= kTile)
#pragma unroll
        for (int w = 0; w < kSW; w++) {
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
        for (int w = 0; w < kSW; w++) {
            uint32_t v = sm.warp_hist[w][tid];
            sm.warp_hist[w][tid] = uint16_t(run);
            run += v;
        }
        c0 = run;
    }
    // exclusive scan of the digit counts over the CTA
    const uint32_t mine = c0 + c1;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) sm.scan_tmp[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < kSW; w++) wbase += (w < warp) ? sm.scan_tmp[w] : 0;
    const uint32_t toff0 = wbase + incl - mine, toff1 = toff0 + c0;
    if (tid == kST - 1) sm.total = toff0 + mine;
    // decoupled look-back over earlier tiles of this pass
    if (PER == 2) {
        sm.tile_off[2 * tid] = toff0;
        sm.tile_off[2 * tid + 1] = toff1;
        uint32_t* my = status + size_t(tile) * BINS + 2 * tid;
        uint32_t e0 = 0, e1 = 0;
        if (tile == 0) {
            st_relaxed2(my, kFlagInclusive | c0, kFlagInclusive | c1);
        } else {
            st_relaxed2(my, kFlagAggregate | c0, kFlagAggregate | c1);
            const uint32_t* p = my - BINS;
            bool done0 = false, done1 = false;
            while (true) {
                uint2 v = ld_relaxed2(p);
                if ((!done0 && (v.x >> 30) == 0) || (!done1 && (v.y >> 30) == 0)) {
                    continue;
                }
                if (!done0) {
                    e0 += v.x & kValueMask;
                    done0 = (v.x & kFlagInclusive) != 0;
                }
                if (!done1) {
                    e1 += v.y & kValueMask;
                    done1 = (v.y & kFlagInclusive) != 0;
                }
                if (done0 && done1) break;
                p -= BINS;
            }
            st_relaxed2(my, kFlagInclusive | (e0 + c0), kFlagInclusive | (e1 + c1));
        }
        sm.bin_base[2 * tid] = hist_base[2 * tid] + e0 - toff0;
        sm.bin_base[2 * tid + 1] = hist_base[2 * tid + 1] + e1 - toff1;
    } else if (tid < BINS) {
        sm.tile_off[tid] = toff0;
        uint32_t* my = status + size_t(tile) * BINS + tid;
        uint32_t e0 = 0;
        if (g_exp_nochain) {
            e0 = tile * (kTile / BINS);
        } else if (tile == 0) {
            st_relaxed(my, kFlagInclusive | c0);
        } else {
            st_relaxed(my, kFlagAggregate | c0);
            const uint32_t* p = my - BINS;
            while (true) {
                uint32_t v = ld_relaxed(p);
                if ((v >> 30) == 0) continue;
                e0 += v & kValueMask;
                if (v & kFlagInclusive) break;
                p -= BINS;
            }
            st_relaxed(my, kFlagInclusive | (e0 + c0));
        }
        sm.bin_base[tid] = hist_base[tid] + e0 - toff0;
    }
    __syncthreads();
    // place the items in tile-sorted order in shared memory
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        uint32_t d = digit[i];
        if (d < BINS) {
            uint32_t slot = sm.tile_off[d] + sm.warp_hist[warp][d] + rank[i];
            sm.klo[slot] = lo[i];
            if (HAS_HI) sm.khi[slot] = hi[i];
            sm.pay[slot] = pay[i];
        }
    }
    __syncthreads();
    const uint32_t total = sm.total;
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        uint32_t j = tid + i * kST;
        if (j < total) {
            uint32_t l = sm.klo[j];
            uint32_t h = HAS_HI ? sm.khi[j] : 0;
            uint32_t idx = sm.bin_base[digit_of(l, h, shift, mask)] + j;
            klo_out[idx] = l;
            if (HAS_HI) khi_out[idx] = h;
            pay_out[idx] = sm.pay[j];
        }
    }
}

// Zeroes this tile's look-back row of the NEXT pass (its buffer was last used two
// passes ago), so no memset runs between the passes.
template <int CAPBITS>
__device__ __forceinline__ void clear_next_row(uint32_t* __restrict__ status_next,
                                               uint32_t tile) {
    if (!status_next) return;
    constexpr int BINS = 1 << CAPBITS;
    uint32_t* row = status_next + size_t(tile) * BINS;
    for (int i = threadIdx.x; i < BINS; i += kST) row[i] = 0;
}

template <int CAPBITS>
__device__ __forceinline__ uint32_t take_ticket(TileSmem<CAPBITS>& sm,
                                                uint32_t* counter) {
    if (threadIdx.x == 0) sm.ticket = atomicAdd(counter, 1u);
    __syncthreads();
    return sm.ticket;
}

// ---- pass 0: generate + scatter -------------------------------------------------
template <int CAPBITS, bool HAS_HI>
__global__ void __launch_bounds__(kST, 2)
sort_pass0_kernel(const uint32_t* __restrict__ packed,
                  const uint32_t* __restrict__ nmask, int k, uint32_t mask,
                  const uint32_t* __restrict__ hist_base,
                  uint32_t* __restrict__ status, uint32_t* __restrict__ status_next,
                  uint32_t* __restrict__ counters,
                  uint32_t* __restrict__ klo_out, uint32_t* __restrict__ khi_out,
                  uint32_t* __restrict__ pay_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TileSmem<CAPBITS>& sm = *reinterpret_cast<TileSmem<CAPBITS>*>(smem_raw);
    constexpr int BINS = 1 << CAPBITS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&sm.bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const uint32_t tile = take_ticket(sm, counters + kCntTicket0);
    clear_next_row<CAPBITS>(status_next, tile);
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
    uint32_t lo[kSI], hi[kSI], pay[kSI], digit[kSI];
    const uint32_t gbase = tile * uint32_t(kTile);
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        uint32_t rel = warp * (32 * kSI) + i * 32 + lane;
        bool bad = window_has_n(sm.stage_m, rel, k);
        uint64_t dir = forward_key(sm.stage_w, rel, k);
        uint64_t rev = revcomp_key(dir, k);
        uint64_t h = dir < rev ? dir : rev;
        lo[i] = uint32_t(h);
        hi[i] = uint32_t(h >> 32);
        // payload: global position << 1 | (ori == +1), ori = +1 iff h == dir
        pay[i] = ((gbase + rel) << 1) | (h == dir ? 1u : 0u);
        digit[i] = bad ? uint32_t(BINS) : (lo[i] & mask);
    }
    rank_and_scatter<CAPBITS, HAS_HI>(sm, lo, hi, pay, digit, tile, 0, mask,
                                      hist_base, status, klo_out, khi_out, pay_out);
}

// ---- pass p >= 1: read + scatter ---------------------------------------------------
template <int CAPBITS, bool HAS_HI>
__global__ void __launch_bounds__(kST, 2)
sort_pass_kernel(int pass, int shift, uint32_t mask,
                 const uint32_t* __restrict__ klo_in,
                 const uint32_t* __restrict__ khi_in,
                 const uint32_t* __restrict__ pay_in,
                 const uint32_t* __restrict__ hist_base,
                 uint32_t* __restrict__ status, uint32_t* __restrict__ status_next,
                 uint32_t* __restrict__ counters,
                 uint32_t* __restrict__ klo_out, uint32_t* __restrict__ khi_out,
                 uint32_t* __restrict__ pay_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TileSmem<CAPBITS>& sm = *reinterpret_cast<TileSmem<CAPBITS>*>(smem_raw);
    constexpr int BINS = 1 << CAPBITS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = take_ticket(sm, counters + kCntTicket0 + pass);
    clear_next_row<CAPBITS>(status_next, tile);
    const uint32_t n = counters[kCntValid];
    const uint64_t tbase = uint64_t(tile) * kTile;
    if (tbase >= n) return;
    uint32_t lo[kSI], hi[kSI], pay[kSI], digit[kSI];
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        uint64_t idx = tbase + warp * (32 * kSI) + i * 32 + lane;
        bool in = idx < n;
        lo[i] = in ? klo_in[idx] : 0;
        hi[i] = (HAS_HI && in) ? khi_in[idx] : 0;
        pay[i] = in ? pay_in[idx] : 0;
    }
#pragma unroll
    for (int i = 0; i < kSI; i++) {
        uint64_t idx = tbase + warp * (32 * kSI) + i * 32 + lane;
        digit[i] = idx < n ? digit_of(lo[i], hi[i], shift, mask) : uint32_t(BINS);
    }
    rank_and_scatter<CAPBITS, HAS_HI>(sm, lo, hi, pay, digit, tile, shift, mask,
                                      hist_base, status, klo_out, khi_out, pay_out);
}

template <typename K>
void set_smem(K kernel, size_t bytes) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         int(bytes));
}

}  // namespace

void exp_set_nochain(int v) {
    cudaMemcpyToSymbol(g_exp_nochain, &v, sizeof(int));
}

SortPlan make_sort_plan(int k, int max_digit_bits) {
    SortPlan p;
    const int total = 2 * k;
    if (max_digit_bits < 4) max_digit_bits = 4;
    if (max_digit_bits > kMaxDigitBits) max_digit_bits = kMaxDigitBits;
    p.passes = (total + max_digit_bits - 1) / max_digit_bits;
    p.width = (total + p.passes - 1) / p.passes;
    p.capbits = p.width <= 8 ? 8 : 10;
    return p;
}

void launch_sort_pass0(const SortPlan& plan, const uint32_t* packed,
                       const uint32_t* nmask, uint64_t n_padded, int k,
                       const uint32_t* hist_base, uint32_t* status,
                       uint32_t* status_next, uint32_t* counters,
                       uint32_t* klo_out, uint32_t* khi_out, uint32_t* pay_out,
                       cudaStream_t s) {
    uint32_t tiles = uint32_t(n_padded / kTile);
    if (tiles == 0) return;
    const uint32_t mask = plan.mask(0, k);
#define NPGE_PASS0(CB, HI)                                                        \
    do {                                                                          \
        set_smem(sort_pass0_kernel<CB, HI>, sizeof(TileSmem<CB>));                \
        sort_pass0_kernel<CB, HI><<<tiles, kST, sizeof(TileSmem<CB>), s>>>(       \
            packed, nmask, k, mask, hist_base, status, status_next, counters,     \
            klo_out, khi_out, pay_out);                                           \
    } while (0)
    const bool hi = k > 16;
    if (plan.capbits == 8) {
        if (hi) NPGE_PASS0(8, true); else NPGE_PASS0(8, false);
    } else {
        if (hi) NPGE_PASS0(10, true); else NPGE_PASS0(10, false);
    }
#undef NPGE_PASS0
}

void launch_sort_pass(const SortPlan& plan, int pass, int k, uint64_t max_records,
                      const uint32_t* klo_in, const uint32_t* khi_in,
                      const uint32_t* pay_in, const uint32_t* hist_base,
                      uint32_t* status, uint32_t* status_next, uint32_t* counters,
                      uint32_t* klo_out, uint32_t* khi_out, uint32_t* pay_out,
                      cudaStream_t s) {
    uint32_t tiles = uint32_t((max_records + kTile - 1) / kTile);
    if (tiles == 0) return;
    const int shift = pass * plan.width;
    const uint32_t mask = plan.mask(pass, k);
    const uint32_t* hb = hist_base + size_t(pass) * plan.bins();
#define NPGE_PASS(CB, HI)                                                         \
    do {                                                                          \
        set_smem(sort_pass_kernel<CB, HI>, sizeof(TileSmem<CB>));                 \
        sort_pass_kernel<CB, HI><<<tiles, kST, sizeof(TileSmem<CB>), s>>>(        \
            pass, shift, mask, klo_in, khi_in, pay_in, hb, status, status_next,   \
            counters, klo_out, khi_out, pay_out);                                 \
    } while (0)
    const bool hi = k > 16;
    if (plan.capbits == 8) {
        if (hi) NPGE_PASS(8, true); else NPGE_PASS(8, false);
    } else {
        if (hi) NPGE_PASS(10, true); else NPGE_PASS(10, false);
    }
#undef NPGE_PASS
}

}  // namespace npge_b200
