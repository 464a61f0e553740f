// msd_kernels.cu -- the two partition passes of the anchor sort.
//
// (canonical k-mer, position, strand) records are ordered in three stages
// (DESIGN.md "Sort"):
//   pass A   regenerate the keys from the packed genome and partition them,
//            stably, by the top `wa` bits of the canonical k-mer
//   pass B   partition every A bucket, stably, by the next `wb` bits
//   bucket   (bucket_kernels.cu) LSD radix sort of the remaining bits of each
//            small run of (A, B) bins entirely in shared memory, fused with the
//            run-length grouping and the anchor output
// Both partition passes are chainless: a counting kernel writes one digit
// histogram row per tile, a scan turns the rows into scatter bases, and the
// scatter kernel never waits on another CTA.  Counting pass A costs one more
// read of the 2-bit genome (0.375 B/base); counting pass B one read of the
// records.  Ranking inside a tile is the stable warp-synchronous scheme of
// rank.cuh, so records with equal keys stay in ascending position order, which
// is the scan order of the reference (SeqI.hpp:45-57, AnchorFinder.cpp:184-196).
//
// Replaces std::sort of FoundFragments (src/algo/AnchorFinder.cpp:234-238,358;
// src/util/SortedVector.hpp:47-49) together with bucket_kernels.cu.
#include <algorithm>

#include "kernels.h"
#include "rank.cuh"

#ifndef NPGE_EXP_PASSB
#define NPGE_EXP_PASSB 0
#endif
#ifndef NPGE_EXP_SORTPASSES
#define NPGE_EXP_SORTPASSES -1
#endif

namespace npge_b200 {

MsdPlan make_msd_plan(int k, uint64_t n_positions, int force_f, int force_wa) {
    MsdPlan p;
    p.k = k;
    p.total = 2 * k;
    // Fine bits.  The bucket sort costs a fixed time per unit of up to 4096 records, and
    // units are whole (A, B) bins packed greedily: bins of ~256 records fill them to
    // 97 %, bins of ~1024 to 84 % (brucella17: sort 0.98 against 1.21 ms, emit 0.45
    // against 0.50).  Finer bins need wider rows in the two partition passes, which
    // cost more than they save beyond 512 bins (9 + 9 bits): so ~256 records per bin up
    // to 18 bits, and above that what keeps a bin below 512 records (the canonical
    // minimum skews the key density up to 2x: bins of 1000-2000 records leave units a
    // third empty, and on several GPUs the rank that owns them sets the step).
    int f_fine = 2, f_mid = 2;
    while (f_fine < 20 && (uint64_t(256) << f_fine) < n_positions) f_fine++;
    while (f_mid < 20 && (uint64_t(512) << f_mid) < n_positions) f_mid++;
    int f = std::min(f_fine, std::max(f_mid, 18));
    if (force_f >= 2 && force_f <= 20) f = force_f;  // tests: the wider digit plans
    if (f > p.total) f = p.total;
    const int need = p.total - 32;  // bits a narrow (8 B) record cannot hold
    // the halves of f go to pass A and pass B; of an odd f the larger half goes to pass A
    // while its rows stay at 512 bins (measured: 9 + 8 beats 8 + 9, 9 + 10 beats 10 + 9:
    // 1024-bin rows cost pass A 0.15 ms per 56 M records, pass B 0.10)
    int wa = (f + 1) / 2 <= 9 ? (f + 1) / 2 : f / 2;
    if (need > 0 && need <= kMaxDigitBits && wa < need) wa = need;
    if (wa > kMaxDigitBits) wa = kMaxDigitBits;
    if (force_wa >= 1 && force_wa <= kMaxDigitBits &&
        (need <= 0 || need > kMaxDigitBits || force_wa >= need)) wa = force_wa;
    if (wa > p.total) wa = p.total;
    if (wa < 1) wa = 1;
    int wb = f - wa;
    if (wb < 1) wb = 1;
    if (wb > kMaxDigitBits) wb = kMaxDigitBits;
    if (wb > p.total - wa) wb = p.total - wa;
    p.wa = wa;
    p.wb = wb;
    p.rem = p.total - wa;
    p.wide = p.rem > 32;
    p.capa = wa <= 8 ? 8 : (wa == 9 ? 9 : 10);  // histogram rows of 256 / 512 / 1024 bins
    p.capb = wb <= 8 ? 8 : (wb == 9 ? 9 : 10);
    p.smem_passes = (p.rem + 7) / 8;
    if (NPGE_EXP_SORTPASSES >= 0) p.smem_passes = NPGE_EXP_SORTPASSES;  // timing experiment
    p.slow_passes = (p.rem - wb + 7) / 8;
    return p;
}

namespace {

// canonical key and validity of the window at tile-relative position rel
__device__ __forceinline__ bool window_key(const GenomeStage& st, uint32_t rel, int k,
                                           uint64_t& h, uint32_t& ori_bit) {
    const bool bad = window_has_n(st.m, rel, k);
    const uint64_t dir = forward_key(st.w, rel, k);
    const uint64_t rev = revcomp_key(dir, k);
    h = dir < rev ? dir : rev;
    ori_bit = h == dir ? 1u : 0u;  // ori = +1 iff h == dir (AnchorFinder.cpp:303-312)
    return !bad;
}

// ---- count A: one histogram row per genome tile ---------------------------------
__global__ void __launch_bounds__(kRT)
count_a_kernel(const uint32_t* __restrict__ packed, const uint32_t* __restrict__ nmask,
               MsdPlan plan, uint32_t win_limit, uint32_t* __restrict__ rows) {
    extern __shared__ __align__(16) unsigned char raw[];
    GenomeStage& st = *reinterpret_cast<GenomeStage*>(raw);
    uint32_t* hist = reinterpret_cast<uint32_t*>(raw + sizeof(GenomeStage));
    const int bins = 1 << plan.capa;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < bins; i += kRT) hist[i] = 0;
    stage_genome_tile(st, packed, nmask, blockIdx.x);
    const int sa = plan.total - plan.wa;
    // a histogram does not care which thread sees which window: every thread walks
    // kRI CONSECUTIVE windows and rolls the two hashes (two bits in, two bits out)
    // instead of extracting each one from scratch
    {
        (void)lane;
        (void)warp;
        const int k = plan.k;
        const uint32_t rel0 = tid * kRI;  // kRI = 8 positions = 16 bits = half a word
        const uint32_t g_first = blockIdx.x * uint32_t(kTile) + rel0;
        const uint32_t* W = st.w + (rel0 >> 4);
        const uint32_t sh = (rel0 & 15) * 2;
        // 128 bits of bases from position rel0 on (the tile carries 16 B of halo)
        uint64_t lo = (uint64_t(__funnelshift_r(W[1], W[2], sh)) << 32) |
                      __funnelshift_r(W[0], W[1], sh);
        uint64_t hi = (uint64_t(__funnelshift_r(W[3], 0u, sh)) << 32) |
                      __funnelshift_r(W[2], W[3], sh);
        const uint32_t* M = st.m + (rel0 >> 5);
        const uint32_t ms = rel0 & 31;
        uint64_t mm = (uint64_t(__funnelshift_r(M[1], M[2], ms)) << 32) |
                      __funnelshift_r(M[0], M[1], ms);
        const uint64_t kmask = k == 32 ? ~0ull : ((1ull << (2 * k)) - 1);
        const uint32_t nm = k == 32 ? ~0u : ((1u << k) - 1);
        uint64_t dir = lo & kmask;
        uint64_t rev = revcomp_key(dir, k);
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            if ((uint32_t(mm) & nm) == 0 && g_first + i < win_limit) {
                const uint64_t h = dir < rev ? dir : rev;
                atomicAdd(&hist[uint32_t(h >> sa)], 1u);
            }
            // the base entering the window: position rel0 + i + k
            const uint32_t in = uint32_t(k == 32 ? hi : (lo >> (2 * k))) & 3u;
            lo = (lo >> 2) | (hi << 62);
            hi >>= 2;
            mm >>= 1;
            dir = lo & kmask;
            rev = ((rev << 2) | (in ^ 1u)) & kmask;
        }
    }
    __syncthreads();
    uint32_t* row = rows + size_t(blockIdx.x) * bins;
    for (int i = tid; i < bins; i += kRT) row[i] = hist[i];
}

// ---- column scan of the histogram rows ---------------------------------------------
// rows[t][d] (t < n_rows) -> exclusive prefix over t of column d, in place; the
// column totals go to rows[n_rows][d].  Three launches: block sums, scan of the
// block sums (one CTA), row prefix.
constexpr int kScanRows = 16;  // rows per block of the two-level scan

__global__ void __launch_bounds__(1024)
scan_rows_sum_kernel(const uint32_t* __restrict__ rows, uint32_t n_rows,
                     uint32_t* __restrict__ blk) {
    const int bins = blockDim.x, d = threadIdx.x;
    const uint32_t r0 = blockIdx.x * kScanRows;
    const uint32_t r1 = min(r0 + kScanRows, n_rows);
    uint32_t s = 0;
#pragma unroll 8
    for (uint32_t r = r0; r < r1; r++) s += rows[size_t(r) * bins + d];
    blk[size_t(blockIdx.x) * bins + d] = s;
}

// 32 columns x 32 groups of blocks per CTA (grid = bins / 32): every group sums
// its share of the block sums, a warp shuffle scans the 32 group sums of a column,
// the group writes its exclusive prefixes.  Loads are 128-byte rows of 32 columns.
__global__ void __launch_bounds__(1024)
scan_rows_blk_kernel(uint32_t* __restrict__ blk, uint32_t n_blk, int bins,
                     uint32_t* __restrict__ totals) {
    __shared__ uint32_t part[32][33];
    const int c = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int d = blockIdx.x * 32 + c;
    const uint32_t per = (n_blk + 31) / 32;
    const uint32_t b0 = min(uint32_t(g) * per, n_blk), b1 = min(b0 + per, n_blk);
    uint32_t sum = 0;
#pragma unroll 4
    for (uint32_t b = b0; b < b1; b++) sum += blk[size_t(b) * bins + d];
    part[g][c] = sum;
    __syncthreads();
    {   // warp g scans column g: lane = group
        const uint32_t v = part[c][g];
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (c >= o) incl += u;
        }
        part[c][g] = incl - v;
        if (c == 31) totals[blockIdx.x * 32 + g] = incl;
    }
    __syncthreads();
    uint32_t run = part[g][c];
    for (uint32_t b = b0; b < b1; b++) {
        const uint32_t v = blk[size_t(b) * bins + d];
        blk[size_t(b) * bins + d] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(1024)
scan_rows_apply_kernel(uint32_t* __restrict__ rows, uint32_t n_rows,
                       const uint32_t* __restrict__ blk,
                       const uint32_t* __restrict__ add) {
    const int bins = blockDim.x, d = threadIdx.x;
    const uint32_t r0 = blockIdx.x * kScanRows;
    const uint32_t r1 = min(r0 + kScanRows, n_rows);
    uint32_t run = blk[size_t(blockIdx.x) * bins + d] + (add ? add[d] : 0u);
    for (uint32_t r = r0; r < r1; r += 8) {
        uint32_t v[8];
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = r + j < r1 ? rows[size_t(r + j) * bins + d] : 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (r + j < r1) rows[size_t(r + j) * bins + d] = run;
            run += v[j];
        }
    }
}

// block-wide exclusive scan of one value per thread (blockDim.x <= 1024)
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* warp_tmp,
                                                    uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    __syncthreads();  // warp_tmp may still be read from a previous call
    if (lane == 31) warp_tmp[warp] = incl;
    __syncthreads();
    uint32_t base = 0, all = 0;
    const int nw = (blockDim.x + 31) / 32;
    for (int w = 0; w < nw; w++) {
        uint32_t t = warp_tmp[w];
        if (w < warp) base += t;
        all += t;
    }
    if (total) *total = all;
    return base + incl - v;
}

// After the A rows are scanned: the segment tables of pass B for one GPU, where
// segment a = A bucket a of this GPU's own pass A output.  One CTA of 1 << capa
// threads.  (With several GPUs the host builds the same tables from the
// all-gathered histograms: one segment per (A bucket, source rank).)
__global__ void __launch_bounds__(1024)
a_tables_kernel(const uint32_t* __restrict__ totals, SegTables t,
                uint32_t* __restrict__ counters) {
    __shared__ uint32_t tmp[32];
    const int bins = blockDim.x, d = threadIdx.x;
    const uint32_t c = totals[d];
    uint32_t all;
    uint32_t s = block_excl_scan(c, tmp, &all);
    t.a_start[d] = s;
    t.seg_begin[d] = s;
    t.seg_end[d] = s + c;
    t.seg_a[d] = d;
    if (d == bins - 1) {
        t.a_start[bins] = all;
        counters[kCntValid] = all;
    }
    uint32_t tl = block_excl_scan((c + kTile - 1) / kTile, tmp, &all);
    t.seg_tile[d] = tl;
    t.a_tile0[d] = tl;
    if (d == bins - 1) {
        t.seg_tile[bins] = all;
        t.a_tile0[bins] = all;
    }
}

// Several GPUs, pass A in chunks of tiles (so that the copies of a finished chunk
// overlap the ranking of the next): the output is chunk-major, bucket-major inside a
// chunk.  chunk_start[c][d] = first record of bucket d of chunk c (c's end in entry
// [bins]), from the per-chunk column totals.  One CTA of 1 << capa threads.
__global__ void __launch_bounds__(1024)
chunk_tables_kernel(const uint32_t* __restrict__ chunk_tot, int n_chunks,
                    uint32_t* __restrict__ chunk_start, uint32_t* __restrict__ counters) {
    __shared__ uint32_t tmp[32];
    const int bins = blockDim.x, d = threadIdx.x;
    uint32_t base = 0;
    for (int c = 0; c < n_chunks; c++) {
        const uint32_t v = chunk_tot[size_t(c) * bins + d];
        uint32_t all;
        const uint32_t s = block_excl_scan(v, tmp, &all);
        chunk_start[size_t(c) * (bins + 1) + d] = base + s;
        if (d == bins - 1) chunk_start[size_t(c) * (bins + 1) + bins] = base + all;
        base += all;
    }
    if (d == 0) counters[kCntValid] = base;
}

// key bits kept in a record after pass A (the low `rem` bits of the k-mer)
__device__ __forceinline__ uint64_t rem_key(uint64_t rec, uint32_t hi) {
    return (uint64_t(hi) << 32) | (rec >> 32);
}

// last a with table[a] <= t (table ascending, n + 1 entries, table[n] > t)
__device__ __forceinline__ uint32_t bucket_of(const uint32_t* __restrict__ table,
                                              uint32_t n, uint32_t t) {
    uint32_t lo = 0, hi = n;  // answer in [lo, hi)
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (table[mid] <= t) lo = mid; else hi = mid;
    }
    return lo;
}

template <int CAPBITS, bool WIDE>
struct __align__(16) PassSmem {
    static constexpr int BINS = 1 << CAPBITS;
    GenomeStage st;  // pass A only
    RankSmem<CAPBITS> rk;
    uint32_t bin_base[BINS];
    uint64_t rec[kTile];
    uint32_t hi[WIDE ? kTile : 4];
    uint16_t dig[kTile];  // pass A: the digit is not part of the record
};

// ---- pass A: generate + partition by the top wa bits -----------------------------
// The payload carries the JOB-WIDE position (pm.g0 + local position): one rank of a
// multi-GPU job scans a range of the job's position space; windows that start at or
// after pm.win_limit belong to the next rank (their bases are this rank's halo).
// PEER: every bucket's run goes straight into the owning rank's receive buffer.
// 8-bit digits with narrow records: 52 KB of shared memory and (at some spill) 32
// registers let four CTAs share an SM; the kernel is latency-bound, occupancy pays
// (measured on brucella17: 0.545 ms at 2 CTAs, 0.483 at 3, 0.467 at 4; pass B is
// best at 3: 0.608 / 0.495 / 0.522).
#ifndef NPGE_PASS_A_CTAS
#define NPGE_PASS_A_CTAS 4
#endif
template <int CAPBITS, bool WIDE, bool PEER>
__global__ void __launch_bounds__(kRT, WIDE ? 2 : (CAPBITS == 8 ? NPGE_PASS_A_CTAS : (CAPBITS == 9 ? 3 : 2)))
pass_a_kernel(const uint32_t* __restrict__ packed, const uint32_t* __restrict__ nmask,
              MsdPlan plan, const uint32_t* __restrict__ rows, PosMap pm, uint32_t tile0,
              uint64_t* __restrict__ out_rec, uint32_t* __restrict__ out_hi) {
    extern __shared__ __align__(16) unsigned char raw[];
    using Smem = PassSmem<CAPBITS, WIDE>;
    Smem& sm = *reinterpret_cast<Smem*>(raw);
    constexpr int BINS = Smem::BINS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = blockIdx.x + tile0;  // (a chunk of the tiles per launch)
    const uint32_t gbase = tile * uint32_t(kTile);
    stage_genome_tile(sm.st, packed, nmask, tile);  // (barrier inside)
    const int sa = plan.rem;
    const uint64_t rmask = (1ull << plan.rem) - 1;  // rem <= 63
    // only the key bits stay in registers across the ranking: the payload (position
    // and strand bit) is rebuilt when the record is written into the sorted tile
    uint32_t klo[kRI], hi[kRI], digit[kRI], slot[kRI];
    uint32_t obits = 0;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        uint32_t rel = warp * (32 * kRI) + i * 32 + lane;
        uint64_t h;
        uint32_t ob;
        bool ok = window_key(sm.st, rel, plan.k, h, ob) && gbase + rel < pm.win_limit;
        uint64_t kr = h & rmask;
        klo[i] = uint32_t(kr);
        hi[i] = uint32_t(kr >> 32);
        obits |= ob << i;
        digit[i] = ok ? uint32_t(h >> sa) : uint32_t(BINS);
    }
    // the scatter bases of the thread's digit(s): loaded before the ranking, whose
    // barriers hide the latency
    const uint32_t* row = rows + size_t(tile) * BINS;
    uint32_t base0 = 0, base1 = 0;
    if (BINS > kRT) {
        base0 = row[2 * tid];
        base1 = row[2 * tid + 1];
    } else if (tid < BINS) {
        base0 = row[tid];
    }
    uint32_t c0, c1;
    tile_rank<CAPBITS>(sm.rk, digit, kRI, slot, c0, c1);
    if (BINS > kRT) {
        sm.bin_base[2 * tid] = base0 - sm.rk.tile_off[2 * tid];
        sm.bin_base[2 * tid + 1] = base1 - sm.rk.tile_off[2 * tid + 1];
    } else if (tid < BINS) {
        sm.bin_base[tid] = base0 - sm.rk.tile_off[tid];
    }
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        if (digit[i] < BINS) {
            // payload: position << 1 | (ori == +1)
            const uint32_t g = pm.g0 + gbase + warp * (32 * kRI) + i * 32 + lane;
            sm.rec[slot[i]] = (uint64_t(klo[i]) << 32) | ((g << 1) | ((obits >> i) & 1u));
            if (WIDE) sm.hi[slot[i]] = hi[i];
            sm.dig[slot[i]] = uint16_t(digit[i]);
        }
    }
    __syncthreads();
    const uint32_t total = sm.rk.total;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        uint32_t j = tid + i * kRT;
        if (j < total) {
            const uint32_t d = sm.dig[j];
            uint32_t idx = sm.bin_base[d] + j;
            if (PEER) {  // straight into the owner's HBM
                reinterpret_cast<uint64_t*>(pm.peer_rec[d])[idx] = sm.rec[j];
                if (WIDE) reinterpret_cast<uint32_t*>(pm.peer_hi[d])[idx] = sm.hi[j];
            } else {
                out_rec[idx] = sm.rec[j];
                if (WIDE) out_hi[idx] = sm.hi[j];
            }
        }
    }
}

// ---- tile descriptors of pass B: the segment search once per tile, not once per CTA of
// the counting and the scatter kernel (a chain of dependent loads at their start)
__global__ void __launch_bounds__(256)
tile_desc_kernel(SegTables t, uint32_t n_tiles, uint4* __restrict__ desc) {
    const uint32_t tb = blockIdx.x * blockDim.x + threadIdx.x;
    if (tb >= n_tiles) return;
    uint4 d = make_uint4(0u, 0u, 0xffffffffu, 0u);
    if (tb < t.seg_tile[t.n_seg]) {
        const uint32_t sg = bucket_of(t.seg_tile, t.n_seg, tb);
        const uint32_t a = t.seg_a[sg];
        const uint32_t s = t.seg_begin[sg] + (tb - t.seg_tile[sg]) * uint32_t(kTile);
        d = make_uint4(s, min(s + uint32_t(kTile), t.seg_end[sg]), a, t.a_tile0[a]);
    }
    desc[tb] = d;
}

// ---- count B: one histogram row per bucket-aligned tile of A-sorted records ----------
template <bool WIDE>
__global__ void __launch_bounds__(kRT)
count_b_kernel(const uint64_t* __restrict__ rec, const uint32_t* __restrict__ hi,
               MsdPlan plan, const uint4* __restrict__ desc, uint32_t* __restrict__ rows) {
    extern __shared__ __align__(16) unsigned char raw[];
    uint32_t* hist = reinterpret_cast<uint32_t*>(raw);
    const int bins = 1 << plan.capb;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < bins; i += kRT) hist[i] = 0;
    __syncthreads();
    const uint32_t tb = blockIdx.x;
    const uint4 td = desc[tb];
    if (td.z != 0xffffffffu) {
        const uint32_t s = td.x, e = td.y;
        const int sb = plan.rem - plan.wb;
        const uint32_t mb = (1u << plan.wb) - 1;
#pragma unroll
        for (int i = 0; i < kRI; i++) {
            uint32_t idx = s + warp * (32 * kRI) + i * 32 + lane;
            if (idx < e) {
                uint64_t kr = rem_key(rec[idx], WIDE ? hi[idx] : 0u);
                atomicAdd(&hist[uint32_t(kr >> sb) & mb], 1u);
            }
        }
    }
    __syncthreads();
    uint32_t* row = rows + size_t(tb) * bins;
    for (int i = tid; i < bins; i += kRT) row[i] = hist[i];
}

// fine_start[(a << wb) + b] = first record of the (a, b) bin; one CTA per a.
// rows hold the exclusive column prefix P over ALL B tiles (totals in row n_rows).
__global__ void __launch_bounds__(1024)
fine_tables_kernel(const uint32_t* __restrict__ rows, MsdPlan plan, SegTables t,
                   uint32_t* __restrict__ fine_start) {
    __shared__ uint32_t tmp[32];
    const int bins = blockDim.x, b = threadIdx.x;
    const uint32_t a = blockIdx.x;
    const uint32_t* a_start = t.a_start;
    const uint32_t t0 = t.a_tile0[a], t1 = t.a_tile0[a + 1];
    const uint32_t c = rows[size_t(t1) * bins + b] - rows[size_t(t0) * bins + b];
    const uint32_t s = block_excl_scan(c, tmp, nullptr);
    if (b < (1 << plan.wb)) fine_start[(a << plan.wb) + b] = a_start[a] + s;
    if (a == gridDim.x - 1 && b == 0) {
        fine_start[gridDim.x << plan.wb] = a_start[gridDim.x];
    }
}

// ---- pass B: partition every A bucket by the next wb bits ------------------------------
template <int CAPBITS, bool WIDE>
__global__ void __launch_bounds__(kRT, (CAPBITS <= 9 && !WIDE) ? 3 : 2)
pass_b_kernel(const uint64_t* __restrict__ in_rec, const uint32_t* __restrict__ in_hi,
              MsdPlan plan, const uint4* __restrict__ desc, const uint32_t* __restrict__ rows,
              const uint32_t* __restrict__ fine_start,
              uint64_t* __restrict__ out_rec, uint32_t* __restrict__ out_hi) {
    extern __shared__ __align__(16) unsigned char raw[];
    using Smem = PassSmem<CAPBITS, WIDE>;
    Smem& sm = *reinterpret_cast<Smem*>(raw);
    constexpr int BINS = Smem::BINS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tb = blockIdx.x;
    const uint4 td = desc[tb];  // (first record, end, A bucket, first tile of the bucket)
    if (td.z == 0xffffffffu) return;
    const uint32_t s = td.x, e = td.y, a = td.z, t0 = td.w;
    const int sb = plan.rem - plan.wb;
    const uint32_t mb = (1u << plan.wb) - 1;
    uint64_t rec[kRI];
    uint32_t hi[kRI], digit[kRI], slot[kRI];
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        uint32_t idx = s + warp * (32 * kRI) + i * 32 + lane;
        bool in = idx < e;
        rec[i] = in ? in_rec[idx] : 0ull;
        hi[i] = (WIDE && in) ? in_hi[idx] : 0u;
    }
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        uint32_t idx = s + warp * (32 * kRI) + i * 32 + lane;
        digit[i] = idx < e ? (uint32_t(rem_key(rec[i], hi[i]) >> sb) & mb)
                           : uint32_t(BINS);
    }
    const uint32_t first = s + warp * (32 * kRI);
    const int rounds = first >= e ? 0 : int(min(uint32_t(kRI), (e - first + 31) / 32));
    // the scatter bases of the thread's digit(s): loaded before the ranking, whose
    // barriers hide the latency
    const uint32_t* row = rows + size_t(tb) * BINS;
    const uint32_t* row0 = rows + size_t(t0) * BINS;
    const uint32_t* fs = fine_start + (a << plan.wb);
    uint32_t base0 = 0, base1 = 0;
    if (BINS > kRT) {
        const int d = 2 * tid;
        if (d <= int(mb)) base0 = fs[d] + row[d] - row0[d];
        if (d + 1 <= int(mb)) base1 = fs[d + 1] + row[d + 1] - row0[d + 1];
    } else if (tid < BINS && tid <= int(mb)) {
        base0 = fs[tid] + row[tid] - row0[tid];
    }
    uint32_t c0, c1;
#if NPGE_EXP_PASSB == 1  // timing experiment: no ranking
    (void)rounds;
#pragma unroll
    for (int i = 0; i < kRI; i++) slot[i] = warp * (32 * kRI) + i * 32 + lane;
    if (tid == 0) sm.rk.total = e - s;
    c0 = c1 = 0;
#else
    tile_rank<CAPBITS>(sm.rk, digit, rounds, slot, c0, c1);
#endif
    if (BINS > kRT) {
        sm.bin_base[2 * tid] = base0 - sm.rk.tile_off[2 * tid];
        sm.bin_base[2 * tid + 1] = base1 - sm.rk.tile_off[2 * tid + 1];
    } else if (tid < BINS) {
        sm.bin_base[tid] = base0 - sm.rk.tile_off[tid];
    }
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        if (digit[i] < BINS) {
            sm.rec[slot[i]] = rec[i];
            if (WIDE) sm.hi[slot[i]] = hi[i];
        }
    }
    __syncthreads();
    const uint32_t total = sm.rk.total;
#pragma unroll
    for (int i = 0; i < kRI; i++) {
        uint32_t j = tid + i * kRT;
        if (j < total) {
            uint64_t r = sm.rec[j];
            uint32_t h = WIDE ? sm.hi[j] : 0u;
            uint32_t idx = sm.bin_base[uint32_t(rem_key(r, h) >> sb) & mb] + j;
#if NPGE_EXP_PASSB  // timing experiment: sequential instead of scattered writes
            idx = s + j;
#endif
            out_rec[idx] = r;
            if (WIDE) out_hi[idx] = h;
        }
    }
}

// totals: where the column totals go (default: the row after the last one)
void scan_rows(uint32_t* rows, uint32_t n_rows, int bins, uint32_t* blk,
               cudaStream_t s, uint32_t* totals = nullptr) {
    const uint32_t n_blk = (n_rows + kScanRows - 1) / kScanRows;
    scan_rows_sum_kernel<<<n_blk, bins, 0, s>>>(rows, n_rows, blk);
    scan_rows_blk_kernel<<<bins / 32, 1024, 0, s>>>(
        blk, n_blk, bins, totals ? totals : rows + size_t(n_rows) * bins);
}

void apply_rows(uint32_t* rows, uint32_t n_rows, int bins, const uint32_t* blk,
                const uint32_t* add, cudaStream_t s) {
    const uint32_t n_blk = (n_rows + kScanRows - 1) / kScanRows;
    scan_rows_apply_kernel<<<n_blk, bins, 0, s>>>(rows, n_rows, blk, add);
}

template <typename K>
void set_smem(K kernel, size_t bytes) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         int(bytes));
}

}  // namespace

size_t msd_tiles_b(uint64_t max_records, uint32_t max_seg) {
    return size_t((max_records + kTile - 1) / kTile) + max_seg;
}

size_t msd_rows_needed(const MsdPlan& plan, uint64_t n_padded, uint64_t max_records,
                       uint32_t max_seg) {
    const size_t ta = size_t(n_padded / kTile);
    const size_t tb = size_t((max_records + kTile - 1) / kTile) + max_seg;
    const size_t wa = (ta + 1) << plan.capa, wb = (tb + 1) << plan.capb;
    return wa > wb ? wa : wb;  // u32 words
}

size_t msd_blk_needed(const MsdPlan& plan, uint64_t n_padded, uint64_t max_records,
                      uint32_t max_seg) {
    const size_t ta = size_t(n_padded / kTile);
    const size_t tb = size_t((max_records + kTile - 1) / kTile) + max_seg;
    const size_t wa = ((ta + kScanRows - 1) / kScanRows + 1) << plan.capa;
    const size_t wb = ((tb + kScanRows - 1) / kScanRows + 1) << plan.capb;
    return wa > wb ? wa : wb;
}

SegTables carve_seg_tables(uint32_t* base, uint32_t max_seg, const MsdPlan& plan) {
    SegTables t;
    const uint32_t bins_a = 1u << plan.capa;
    t.n_seg = bins_a;
    t.seg_begin = base;
    t.seg_end = t.seg_begin + max_seg;
    t.seg_a = t.seg_end + max_seg;
    t.seg_tile = t.seg_a + max_seg;
    t.a_tile0 = t.seg_tile + max_seg + 1;
    t.a_start = t.a_tile0 + bins_a + 1;
    return t;
}

size_t seg_table_words(uint32_t max_seg, const MsdPlan& plan) {
    return size_t(4) * max_seg + 1 + 2 * ((size_t(1) << plan.capa) + 1);
}

// genome -> records partitioned by the top wa key bits (stable) in b.rec[0] /
// b.hi[0]; on one GPU also the segment tables of pass B (b.seg).
void launch_msd_pass_a(const MsdPlan& plan, const uint32_t* packed,
                       const uint32_t* nmask, uint64_t n_padded, const MsdBuffers& b,
                       const PosMap& pm, uint32_t* counters, cudaStream_t s,
                       cudaEvent_t* ev, uint32_t* launches, int part,
                       const PassAChunks* ch) {
    const uint32_t tiles_a = uint32_t(n_padded / kTile);
    if (tiles_a == 0) return;
    const int bins_a = 1 << plan.capa;
    const int n_chunks = ch ? ch->n_chunks : 1;
    // chunk c = tiles [c * per, (c + 1) * per)
    const uint32_t per = (tiles_a + n_chunks - 1) / n_chunks;
    if (part != 2) {
        count_a_kernel<<<tiles_a, kRT, sizeof(GenomeStage) + bins_a * 4, s>>>(
            packed, nmask, plan, pm.win_limit, b.rows);
        *launches += 1;
        if (!ch) {
            scan_rows(b.rows, tiles_a, bins_a, b.blk, s);
            a_tables_kernel<<<1, bins_a, 0, s>>>(b.rows + size_t(tiles_a) * bins_a, b.seg,
                                                 counters);
            apply_rows(b.rows, tiles_a, bins_a, b.blk, b.seg.a_start, s);
            *launches += 4;
        } else {
            for (int c = 0; c < n_chunks; c++) {
                const uint32_t t0 = min(uint32_t(c) * per, tiles_a), t1 = min(t0 + per, tiles_a);
                uint32_t* tot = ch->chunk_tot + size_t(c) * bins_a;
                if (t1 > t0) {
                    scan_rows(b.rows + size_t(t0) * bins_a, t1 - t0, bins_a,
                              b.blk + (size_t(t0) / kScanRows + c) * bins_a, s, tot);
                    *launches += 2;
                } else {
                    cudaMemsetAsync(tot, 0, size_t(bins_a) * 4, s);
                }
            }
            chunk_tables_kernel<<<1, bins_a, 0, s>>>(ch->chunk_tot, n_chunks, ch->chunk_start,
                                                     counters);
            *launches += 1;
            for (int c = 0; c < n_chunks; c++) {
                const uint32_t t0 = min(uint32_t(c) * per, tiles_a), t1 = min(t0 + per, tiles_a);
                if (t1 > t0) {
                    apply_rows(b.rows + size_t(t0) * bins_a, t1 - t0, bins_a,
                               b.blk + (size_t(t0) / kScanRows + c) * bins_a,
                               ch->chunk_start + size_t(c) * (bins_a + 1), s);
                    *launches += 1;
                }
            }
        }
        if (ev) cudaEventRecord(ev[0], s);
    }
    if (part == 1) return;
#define NPGE_PASS_A(CB, W, G)                                                      \
    do {                                                                           \
        set_smem(pass_a_kernel<CB, W, G>, sizeof(PassSmem<CB, W>));                \
        pass_a_kernel<CB, W, G><<<t1 - t0, kRT, sizeof(PassSmem<CB, W>), s>>>(     \
            packed, nmask, plan, b.rows, pm, t0, b.rec[0], b.hi[0]);               \
    } while (0)
#define NPGE_PASS_A2(CB, W)                                                        \
    do {                                                                           \
        if (pm.peer_rec) NPGE_PASS_A(CB, W, true); else NPGE_PASS_A(CB, W, false); \
    } while (0)
    for (int c = 0; c < n_chunks; c++) {
        const uint32_t t0 = min(uint32_t(c) * per, tiles_a), t1 = min(t0 + per, tiles_a);
        if (t1 > t0) {
            if (plan.capa == 8) {
                if (plan.wide) NPGE_PASS_A2(8, true); else NPGE_PASS_A2(8, false);
            } else if (plan.capa == 9) {
                if (plan.wide) NPGE_PASS_A2(9, true); else NPGE_PASS_A2(9, false);
            } else {
                if (plan.wide) NPGE_PASS_A2(10, true); else NPGE_PASS_A2(10, false);
            }
            *launches += 1;
        }
        if (ch && ch->done) cudaEventRecord(ch->done[c], s);
    }
#undef NPGE_PASS_A2
#undef NPGE_PASS_A
    if (ev) cudaEventRecord(ev[1], s);
}

// records of the segments b.seg describes (in in_rec / in_hi, each segment
// holding one A bucket) -> sorted by (A, B), stable in segment order, in
// b.rec[1] / b.hi[1]; b.fine_start gives the first record of every (A, B) bin.
void launch_msd_pass_b(const MsdPlan& plan, const uint64_t* in_rec, const uint32_t* in_hi,
                       uint64_t max_records, uint32_t max_seg, const MsdBuffers& b,
                       cudaStream_t s, cudaEvent_t* ev, uint32_t* launches) {
    const int bins_a = 1 << plan.capa, bins_b = 1 << plan.capb;
    const uint32_t tiles_b = uint32_t(msd_tiles_b(max_records, max_seg));
    tile_desc_kernel<<<(tiles_b + 255) / 256, 256, 0, s>>>(b.seg, tiles_b, b.tile_desc);
    *launches += 1;
    if (plan.wide) {
        count_b_kernel<true><<<tiles_b, kRT, bins_b * 4, s>>>(in_rec, in_hi, plan,
                                                            b.tile_desc, b.rows);
    } else {
        count_b_kernel<false><<<tiles_b, kRT, bins_b * 4, s>>>(in_rec, in_hi, plan,
                                                             b.tile_desc, b.rows);
    }
    scan_rows(b.rows, tiles_b, bins_b, b.blk, s);
    apply_rows(b.rows, tiles_b, bins_b, b.blk, nullptr, s);
    fine_tables_kernel<<<bins_a, bins_b, 0, s>>>(b.rows, plan, b.seg, b.fine_start);
    *launches += 5;
    if (ev) cudaEventRecord(ev[2], s);
#define NPGE_PASS_B(CB, W)                                                         \
    do {                                                                           \
        set_smem(pass_b_kernel<CB, W>, sizeof(PassSmem<CB, W>));                   \
        pass_b_kernel<CB, W><<<tiles_b, kRT, sizeof(PassSmem<CB, W>), s>>>(        \
            in_rec, in_hi, plan, b.tile_desc, b.rows, b.fine_start, b.rec[1], b.hi[1]);  \
    } while (0)
    if (plan.capb == 8) {
        if (plan.wide) NPGE_PASS_B(8, true); else NPGE_PASS_B(8, false);
    } else if (plan.capb == 9) {
        if (plan.wide) NPGE_PASS_B(9, true); else NPGE_PASS_B(9, false);
    } else {
        if (plan.wide) NPGE_PASS_B(10, true); else NPGE_PASS_B(10, false);
    }
#undef NPGE_PASS_B
    *launches += 1;
    if (ev) cudaEventRecord(ev[3], s);
}

}  // namespace npge_b200
