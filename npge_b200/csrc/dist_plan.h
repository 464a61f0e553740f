// dist_plan.h -- host tables of the multi-GPU exchange (pure C++, no CUDA): who
// owns which A buckets, where every rank's slice lands in the owner's receive buffer,
// and the (bucket, source rank) segments pass B walks there.  Every rank computes the
// same tables from the all-gathered per-bucket histograms (a few KB), in
// microseconds: the step has no Python or numpy between its kernels.
//
// The reference has no counterpart: AnchorFinder shares one hash -> positions map
// between the threads of one process (src/algo/AnchorFinder.cpp:252-292, 330-357).
#pragma once

#include <stdint.h>

#include <vector>

namespace npge_b200 {

constexpr uint32_t kPlanTile = 4096;  // records per pass-B tile (kTile)

// Every rank runs pass A in `chunks` chunks of its tiles (the copies of a finished
// chunk overlap the ranking of the next); a (rank, chunk) pair is one SOURCE of the
// exchange: n_src = world * chunks histograms, source rank * chunks + c.
struct DistPlan {
    uint32_t world = 0, chunks = 1, bins = 0, rank = 0;
    std::vector<uint32_t> cuts;       // [world + 1] rank r owns buckets cuts[r] .. cuts[r + 1] - 1
    std::vector<uint64_t> n_recv;     // [world] records every rank receives
    // this rank as a SENDER, for every one of its chunks c and every destination d:
    // records [send_first, send_first + send_n) of its pass-A output (chunk-major) go
    // to records [send_to, ...) of d's receive buffer
    std::vector<uint64_t> send_first, send_n, send_to;  // [chunks][world]
    // per bucket, chunk 0 (the fused scatter runs unchunked): element idx of the
    // A-sorted order lands at record dest_off[a] + idx of dest_rank[a]'s buffer
    std::vector<uint32_t> dest_rank;  // [bins]
    std::vector<int64_t> dest_off;    // [bins]
    // this rank as a RECEIVER: the receive buffer holds source 0's slice, then source
    // 1's, ...; inside a slice the buckets ascend.  Segments are bucket-major, sources
    // in order (= position order) inside a bucket, so that pass B stays stable.
    std::vector<uint32_t> seg_begin, seg_end, seg_a, seg_tile;  // seg_tile: [n_seg + 1]
    std::vector<uint32_t> a_tile0, a_start;                     // [bins + 1]
    uint32_t n_seg = 0;
};

// Contiguous bucket ranges balanced by records: the boundary whose prefix sum is
// nearest to r / world of the total (the canonical minimum skews the k-mer
// distribution; a fixed split of the hash space would overload rank 0).
inline void dist_bucket_ranges(const uint32_t* hist_all, uint32_t n_src, uint32_t world,
                               uint32_t bins, std::vector<uint32_t>& cuts) {
    std::vector<uint64_t> csum(bins + 1, 0);
    for (uint32_t a = 0; a < bins; a++) {
        uint64_t t = 0;
        for (uint32_t s = 0; s < n_src; s++) t += hist_all[size_t(s) * bins + a];
        csum[a + 1] = csum[a] + t;
    }
    const uint64_t total = csum[bins];
    cuts.assign(world + 1, 0);
    for (uint32_t r = 1; r < world; r++) {
        // target = total * r / world, compared exactly in units of 1 / world
        const unsigned __int128 target = (unsigned __int128)total * r;
        uint32_t j = 0;  // first boundary with csum[j] * world >= target
        {
            uint32_t lo = 0, hi = bins + 1;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) / 2;
                if ((unsigned __int128)csum[mid] * world < target) lo = mid + 1; else hi = mid;
            }
            j = lo;
        }
        if (j > bins) j = bins;
        if (j > 0) {
            const unsigned __int128 below = target - (unsigned __int128)csum[j - 1] * world;
            const unsigned __int128 above = (unsigned __int128)csum[j] * world - target;
            if (below < above) j--;
        }
        if (j < cuts[r - 1]) j = cuts[r - 1];
        cuts[r] = j;
    }
    cuts[world] = bins;
}

// a_first[c][bins + 1]: first record of every bucket of chunk c in this rank's pass-A
// output (chunk-major), as the device computed it.
inline void dist_make_plan(const uint32_t* hist_all, uint32_t world, uint32_t chunks,
                           uint32_t bins, uint32_t rank, const uint32_t* a_first,
                           DistPlan& p) {
    const uint32_t n_src = world * chunks;
    p.world = world;
    p.chunks = chunks;
    p.bins = bins;
    p.rank = rank;
    dist_bucket_ranges(hist_all, n_src, world, bins, p.cuts);
    auto H = [&](uint32_t s, uint32_t a) -> uint64_t { return hist_all[size_t(s) * bins + a]; };
    // sender side
    p.send_first.assign(size_t(chunks) * world, 0);
    p.send_n.assign(size_t(chunks) * world, 0);
    p.send_to.assign(size_t(chunks) * world, 0);
    p.dest_rank.assign(bins, 0);
    p.dest_off.assign(bins, 0);
    p.n_recv.assign(world, 0);
    for (uint32_t d = 0; d < world; d++) {
        const uint32_t lo = p.cuts[d], hi = p.cuts[d + 1];
        uint64_t before = 0;  // records of the sources before the current one
        for (uint32_t s = 0; s < n_src; s++) {
            uint64_t t = 0;
            for (uint32_t a = lo; a < hi; a++) t += H(s, a);
            if (s / chunks == rank) {
                const uint32_t c = s % chunks;
                // without a_first: a single chunk whose buckets start at the prefix sums
                uint64_t first = 0;
                if (a_first) {
                    first = a_first[size_t(c) * (bins + 1) + lo];
                } else {
                    for (uint32_t a = 0; a < lo; a++) first += H(s, a);
                }
                p.send_first[size_t(c) * world + d] = first;
                p.send_n[size_t(c) * world + d] = t;
                p.send_to[size_t(c) * world + d] = before;
                if (c == 0) {
                    for (uint32_t a = lo; a < hi; a++) {
                        p.dest_rank[a] = d;
                        p.dest_off[a] = int64_t(before) - int64_t(first);
                    }
                }
            }
            before += t;
        }
        p.n_recv[d] = before;
    }
    // receiver side
    const uint32_t lo = p.cuts[rank], hi = p.cuts[rank + 1];
    std::vector<uint64_t> src_base(n_src + 1, 0);
    for (uint32_t s = 0; s < n_src; s++) {
        uint64_t t = 0;
        for (uint32_t a = lo; a < hi; a++) t += H(s, a);
        src_base[s + 1] = src_base[s] + t;
    }
    std::vector<uint64_t> inside(n_src, 0);  // offset of the current bucket in every slice
    p.seg_begin.clear();
    p.seg_end.clear();
    p.seg_a.clear();
    p.seg_tile.assign(1, 0);
    p.a_tile0.assign(bins + 1, 0);
    p.a_start.assign(bins + 1, 0);
    uint32_t tiles = 0;
    uint64_t records = 0;
    for (uint32_t a = 0; a < bins; a++) {
        p.a_tile0[a] = tiles;
        p.a_start[a] = uint32_t(records);
        if (a < lo || a >= hi) continue;
        for (uint32_t s = 0; s < n_src; s++) {
            const uint64_t c = H(s, a);
            if (c) {
                const uint64_t b = src_base[s] + inside[s];
                p.seg_begin.push_back(uint32_t(b));
                p.seg_end.push_back(uint32_t(b + c));
                p.seg_a.push_back(a);
                tiles += uint32_t((c + kPlanTile - 1) / kPlanTile);
                p.seg_tile.push_back(tiles);
                records += c;
            }
            inside[s] += c;
        }
    }
    p.a_tile0[bins] = tiles;
    p.a_start[bins] = uint32_t(records);
    if (p.seg_begin.empty()) {  // nothing received: one empty segment keeps the tables valid
        p.seg_begin.push_back(0);
        p.seg_end.push_back(0);
        p.seg_a.push_back(lo < bins ? lo : 0);
        p.seg_tile.push_back(0);
    }
    p.n_seg = uint32_t(p.seg_begin.size());
}

// Share of max-anchor-fragments of every rank: the cut runs over the list ordered by
// hash, and the ranks own ascending hash ranges.
inline uint64_t dist_local_limit(const uint64_t* totals, uint32_t rank, uint64_t limit) {
    uint64_t before = 0;
    for (uint32_t r = 0; r < rank; r++) before += totals[r];
    if (before >= limit) return 0;
    const uint64_t room = limit - before;
    return room < totals[rank] ? room : totals[rank];
}

// Equal ranges of the job's position space: rank r owns the windows that start in
// [cut[r], cut[r + 1]) (multiples of `align` but the last).
inline void dist_position_cuts(uint64_t positions, uint32_t world, uint64_t align,
                               std::vector<uint64_t>& cuts) {
    cuts.assign(world + 1, 0);
    const uint64_t per = ((positions + world - 1) / world + align - 1) / align * align;
    for (uint32_t r = 1; r <= world; r++) {
        const uint64_t c = per * r;
        cuts[r] = c < positions ? c : positions;
    }
}

}  // namespace npge_b200
