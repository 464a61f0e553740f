// common.cuh -- shared constants, device helpers and the context of the
// B200 anchor finder.  sm_100a only.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/npge_anchor.h"
#include "dist_plan.h"

namespace npge_b200 {

// ---- device data layout ------------------------------------------------------
// Sequences with size >= k are concatenated in SCAN ORDER (size descending,
// ties by caller index: SeqBase::make_seqs, src/algo/SeqI.hpp:45-57) into one
// global position space; one pad position marked N follows every sequence, so
// a window that would cross a sequence end is invalid through the N mask alone
// (no circular wrap).  Base at global position g:
//   packed[g / 16] bits [2*(g%16), +2)   codes A0 T1 G2 C3 (char_to_size.hpp:24-36)
//   nmask[g / 32]  bit  g % 32           1 = N or pad
// which is the bit layout of CompactLowNSequence (Sequence.cpp:590-596) and of
// make_hash (make_hash.hpp:29-42): the forward hash of the window at g is bits
// [2g, 2g+2k) of the packed stream.

#ifndef NPGE_SORT_THREADS
#define NPGE_SORT_THREADS 512
#endif
constexpr int kSortThreads = NPGE_SORT_THREADS;
constexpr int kSortItems = 4096 / NPGE_SORT_THREADS;
constexpr int kTile = kSortThreads * kSortItems;  // 4096 records / positions
constexpr int kMaxDigitBits = 10;                 // widest partition digit
constexpr int kMaxPasses = 16;                    // event slots per find

// Plan of the MSD partition + shared-memory bucket sort for 2k-bit keys
// (msd_kernels.cu): pass A takes the top wa bits, pass B the next wb, the bucket
// sort the low `rem` = 2k - wa bits (B included) in `smem_passes` byte digits.
struct MsdPlan {
    int k = 0, total = 0, wa = 0, wb = 0, rem = 0;
    int capa = 8, capb = 8;  // histogram rows hold 1 << cap bins
    int wide = 0;            // rem > 32: records carry a second u32 of key bits
    int smem_passes = 0;     // byte passes of the shared-memory sort over the rem bits
    int slow_passes = 0;     // oversized unit (one bin): byte passes below the B digit
};

constexpr int kBucketCap = kTile;        // records one CTA sorts in shared memory
constexpr int kWindow = kBucketCap / 2;  // bins starting in one window = one unit

// positions are padded so every tile can bulk-copy tile + 16 B of halo
constexpr uint64_t kPosAlign = 4096;

constexpr uint64_t kMaxRecords = (1u << 30) - 1;      // per sort
constexpr uint64_t kMaxPositions = (1ull << 31) - 1; // payload = gpos << 1 | ori

__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// The two digest terms of one fragment (include/npge_anchor.h).
__host__ __device__ inline void digest_terms(uint64_t t, uint64_t h, uint64_t& a,
                                             uint64_t& b) {
    uint64_t u = t * 0x9E3779B97F4A7C15ull + h * 0xC2B2AE3D27D4EB4Full;
    u ^= u >> 32;
    a = u;
    uint64_t v = u * 0xD6E8FEB86659FD93ull;
    b = v ^ (v >> 29);
}

#ifdef __CUDACC__
// Canonical key of the window of k bases at global position p, read from
// packed words `W` (16 bases per u32) whose word 0 holds position `base`.
// dir: make_hash ori=+1; rev: ori=-1 (complement = code ^ 1, reversed), as
// SeqI::init_state / next_hash maintain them (SeqI.hpp:86-119).
__device__ __forceinline__ uint64_t forward_key(const uint32_t* W, uint32_t rel,
                                                int k) {
    uint32_t w = rel >> 4, s = (rel & 15) * 2;
    uint32_t w0 = W[w], w1 = W[w + 1], w2 = W[w + 2];
    uint32_t lo = __funnelshift_r(w0, w1, s);
    uint32_t hi = __funnelshift_r(w1, w2, s);
    uint64_t x = (uint64_t(hi) << 32) | lo;
    return k == 32 ? x : (x & ((1ull << (2 * k)) - 1));
}

__device__ __forceinline__ uint64_t revcomp_key(uint64_t dir, int k) {
    uint64_t x = __brevll(dir);
    const uint64_t m = 0x5555555555555555ull;
    x = ((x >> 1) & m) | ((x & m) << 1);  // undo the bit swap inside each base
    x ^= m;                               // complement_letter: code ^ 1
    return x >> (64 - 2 * k);
}

// 1 iff any of the k positions starting at rel is N / pad.
__device__ __forceinline__ bool window_has_n(const uint32_t* M, uint32_t rel,
                                             int k) {
    uint32_t w = rel >> 5, s = rel & 31;
    uint32_t m = __funnelshift_r(M[w], M[w + 1], s);
    if (k < 32) m &= (1u << k) - 1;
    return m != 0;
}
#endif

// ---- host-side context ---------------------------------------------------------
struct DeviceBuffer {
    void* ptr = nullptr;
    size_t bytes = 0;
};

struct SeqEntry {
    uint32_t caller_index;
    uint64_t size;
    uint64_t start;  // global position of base 0
};

}  // namespace npge_b200

struct npge_anchor_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_pass[npge_b200::kMaxPasses + 1] = {};

    // uploaded genome
    bool uploaded = false;
    int32_t upload_k = 0;
    std::vector<npge_b200::SeqEntry> seqs;  // scan order
    uint64_t n_positions = 0;               // sum(size + 1)
    uint64_t n_padded = 0;                  // rounded up to kPosAlign
    uint64_t n_bases = 0;                   // sum(size)
    uint32_t n_caller_seqs = 0;
    npge_b200::DeviceBuffer ascii, packed, nmask, seq_start, seq_caller, seq_pair;

    // sort / grouping scratch
    npge_b200::DeviceBuffer counters, used;
    npge_b200::DeviceBuffer found_bits, found_tmp, sel;
    // MSD partition + bucket sort (msd_kernels.cu, bucket_kernels.cu)
    npge_b200::DeviceBuffer rows, blk, segtab, fine_start, tile_desc, bstatus, unit_wtot, units;
    npge_b200::DeviceBuffer rec[2], rhi[2], seq_lut;
    int lut_shift = 12;
    npge_b200::MsdPlan plan;
    int force_f = 0, force_wa = 0;  // npge_anchor_debug_plan (tests)
    // several GPUs (npge_anchor_dist_*): job-wide sequence table and exchange
    bool dist = false;
    std::vector<npge_b200::SeqEntry> gseqs;  // ALL sequences, job-wide scan order
    uint64_t g_positions = 0;                // job-wide positions (with pads)
    uint64_t g_padded = 0;
    npge_b200::DeviceBuffer g_seq_start, g_seq_caller, g_seq_lut, g_seq_pair;
    npge_b200::DistPlan dplan;               // tables of the last npge_anchor_dist_exchange
    bool planned = false;
    std::vector<cudaStream_t> copy_streams;  // one per destination rank
    std::vector<cudaEvent_t> copy_events;
    std::vector<cudaEvent_t> chunk_events;   // pass A chunk c has been scattered
    int dist_chunks = 1;                     // chunks of pass A (npge_anchor_dist_set_chunks)
    int a_chunks = 1;                        // chunks of the last count A
    npge_b200::DeviceBuffer chunk_tab;       // per-chunk column totals and bucket starts
    uint64_t dist_g0 = 0, dist_n_own = 0;    // this rank's range of the job-wide positions
    npge_b200::DeviceBuffer recv_rec, recv_hi;
    uint64_t n_recv = 0;
    uint32_t dist_launches = 0;
    uint32_t dist_max_seg = 0;
    // peer exchange: receive buffers of all ranks mapped through CUDA IPC
    std::vector<void*> peer_opened;          // to close
    std::vector<void*> peer_rec, peer_hi;    // [world]; own rank: recv_rec / recv_hi
    uint32_t dist_rank = 0;                  // this rank's index in peer_rec
    uint64_t peer_capacity = 0;              // records every receive buffer can take
    npge_b200::DeviceBuffer frag_pat, frag_tiles, frag_out;  // FragmentFinder scratch
    npge_b200::DeviceBuffer raw, ingest;                     // upload_raw scratch
    npge_b200::DeviceBuffer bh_in, bh_names, bh_scratch, bh_out;  // block_hash
    float bh_ms = 0.f;                       // device time of the last block-hash call
    npge_b200::DeviceBuffer peer_tab;        // per-bucket destination addresses
    bool a_counted = false;
    std::vector<uint32_t> a_start_h;         // host copy of this rank's a_start
    bool a_scatter_async = false;            // pass A still in flight on the stream
    bool a_scattered = false;                // rec[0] holds the local pass A output
    uint64_t dist_n_used = 0;
    float dist_ms[8] = {};
    float dist_exchange_ms = 0.f;            // start of pass A .. last copy landed
    // device result
    npge_b200::DeviceBuffer r_seq, r_pos, r_ori, r_block_off, r_block_hash;
    npge_b200::DeviceBuffer summary;
    // compact fetch (api.cu: fetch_compact): the emit kernel also leaves the
    // 4-byte payload (position << 1 | strand) of every fragment; the host turns it
    // into (sequence, position, strand) while the next chunk crosses PCIe
    npge_b200::DeviceBuffer r_pay;
    bool have_pay = false;
    int layout = 0;            // npge_anchor_set_result_layout: 0 full, 1 compact
    // npge_anchor_find_compact, exact mode without a cut: the units are sorted and emitted
    // range by range and every finished range is copied to these HOST buffers while the
    // next one is sorted (api.cu: find_device_msd)
    struct FetchPipe {
        uint32_t* pos_ori;
        uint64_t* block_off;
        uint64_t* block_hash;
        bool used;
    };
    FetchPipe* pipe = nullptr;
    cudaStream_t fetch_stream = nullptr;
    uint64_t* pipe_carry = nullptr;  // pinned: blocks << 31 | fragments after every range
    bool result_compact = false;  // the resident result holds payloads only
    std::vector<uint32_t> h_start32, h_caller, h_lut;  // host copies of the position tables
    std::vector<cudaEvent_t> fetch_ev;

    bool have_result = false;
    npge_anchor_params last_params = {0, 0, 0};
    npge_anchor_stats stats = {};
    uint64_t total_selected = 0;  // before truncation
    bool lone_tail = false;       // the cut left one member of a group: dropped,
    uint64_t lone_hash = 0;       // but its hash is appended to new_used
};
