// kernels.h -- launchers of the hand-written sm_100a kernels (host-callable).
#pragma once

#include "common.cuh"

namespace npge_b200 {

// counters[] slots (u32 unless noted) in ctx->counters
enum {
    kCntValid = 0,      // number of N-free windows (records)
    kCntBadInput = 1,   // bytes outside ATGCN seen by pack
    kCntCutIndex = 22,  // sorted index of the record with selected-rank M-1
    kCntWindows = 23,   // bucket-sort windows of the last partition
    kCntCutStart = 24,  // unit holding the cut record: first record, size
    kCntCutSize = 25,
    kCntWords = 32
};

// summary[] slots (u64) in ctx->summary
enum {
    kSumFragments = 0,  // selected records (before truncation)
    kSumBlocks = 1,
    kSumDigest0 = 2,
    kSumDigest1 = 3,
    kSumEmitted = 4,    // fragments actually written (after truncation)
    kSumBlocksEmitted = 5,
    kSumCarry = 6,      // blocks << 31 | fragments of the unit ranges scanned so far
    kSumWords = 8
};

void launch_pack(const uint8_t* ascii, uint32_t* packed, uint32_t* nmask,
                 uint64_t n_padded, uint32_t* counters, cudaStream_t s);

// ---- MSD partition + bucket sort (msd_kernels.cu, bucket_kernels.cu) -----------
// Segments of pass B: segment s holds records [seg_begin[s], seg_end[s]) of the
// input buffer, all of A bucket seg_a[s]; segments are ordered by (bucket,
// source), tiles of kTile records never cross a segment.  One GPU: segment =
// bucket.  Several GPUs: one segment per (bucket, source rank).
struct SegTables {
    uint32_t n_seg;
    uint32_t* seg_begin;
    uint32_t* seg_end;
    uint32_t* seg_a;
    uint32_t* seg_tile;  // [n_seg + 1] first tile of every segment
    uint32_t* a_tile0;   // [(1 << capa) + 1] first tile of every bucket
    uint32_t* a_start;   // [(1 << capa) + 1] first OUTPUT record of every bucket
};

// One rank of a multi-GPU job scans a RANGE of the job-wide position space: local
// position p is job-wide position g0 + p, and only the windows that start before
// win_limit are this rank's (the k - 1 positions after them are the halo that
// completes its last windows).  One GPU: g0 = 0, win_limit = all.
struct PosMap {
    uint32_t g0 = 0;
    uint32_t win_limit = 0xffffffffu;
    // optional, per A bucket: where the bucket's records go, as addresses such
    // that element idx (the index in this rank's A-sorted order) lands in the
    // owning rank's receive buffer -- peer memory over NVLink.  nullptr: local.
    const unsigned long long* peer_rec = nullptr;
    const unsigned long long* peer_hi = nullptr;
};

struct MsdBuffers {
    uint32_t* rows;        // histogram rows / scatter bases of the current pass
    uint32_t* blk;         // block sums of the row scan
    SegTables seg;
    uint32_t* fine_start;  // [((1 << capa) << wb) + 1]
    uint4* tile_desc;      // per pass-B tile: (first record, end, A bucket, first tile of the bucket)
    uint64_t* rec[2];      // key bits [0, 32) << 32 | payload
    uint32_t* hi[2];       // key bits [32, rem) when plan.wide
};

// force_f / force_wa > 0 override the fine bits and the first split (tests)
MsdPlan make_msd_plan(int k, uint64_t n_positions, int force_f = 0, int force_wa = 0);
size_t msd_rows_needed(const MsdPlan& plan, uint64_t n_padded, uint64_t max_records,
                       uint32_t max_seg);
size_t msd_blk_needed(const MsdPlan& plan, uint64_t n_padded, uint64_t max_records,
                      uint32_t max_seg);
size_t seg_table_words(uint32_t max_seg, const MsdPlan& plan);
SegTables carve_seg_tables(uint32_t* base, uint32_t max_seg, const MsdPlan& plan);
size_t msd_tiles_b(uint64_t max_records, uint32_t max_seg);  // pass-B tiles (upper bound)
// ev (optional): [0] after count A, [1] after pass A, [2] after count B, [3] pass B
// part: 0 = count + scatter, 1 = count A and the tables only, 2 = the scatter only
// Several GPUs: pass A in chunks of tiles, output chunk-major (msd_kernels.cu)
struct PassAChunks {
    int n_chunks;
    uint32_t* chunk_tot;    // [n_chunks][bins] column totals of every chunk
    uint32_t* chunk_start;  // [n_chunks][bins + 1] first record of every (chunk, bucket)
    cudaEvent_t* done;      // [n_chunks] recorded after the scatter of every chunk (optional)
};
void launch_msd_pass_a(const MsdPlan& plan, const uint32_t* packed,
                       const uint32_t* nmask, uint64_t n_padded, const MsdBuffers& b,
                       const PosMap& pm, uint32_t* counters, cudaStream_t s,
                       cudaEvent_t* ev, uint32_t* launches, int part = 0,
                       const PassAChunks* ch = nullptr);
constexpr int kMaxChunks = 8;
void launch_msd_pass_b(const MsdPlan& plan, const uint64_t* in_rec, const uint32_t* in_hi,
                       uint64_t max_records, uint32_t max_seg, const MsdBuffers& b,
                       cudaStream_t s, cudaEvent_t* ev, uint32_t* launches);

struct BucketArgs {
    MsdPlan plan;
    uint64_t* rec_w;        // records sorted by (A, B): buffer 1 of MsdBuffers
    uint32_t* hi_w;
    uint64_t* scratch_rec;  // buffer 0: scratch of oversized units
    uint32_t* scratch_hi;
    const uint32_t* fine_start;
    uint4* units;                // (first record, size, A bucket, -) of every unit
    uint32_t* counters;
    const uint64_t* used;
    uint64_t n_used;
    const uint64_t* seq_start;   // [S+1] global start of each scan-order sequence
    const uint32_t* seq_caller;  // [S] caller index
    const uint32_t* seq_lut;     // sequence holding position q << lut_shift
    const uint2* seq_pair;       // [S + 1] (start, caller index) of every scan-order sequence
    int lut_shift;
    uint64_t max_fragments;      // truncation M
    uint32_t* found_bits;        // anchor-similar: 1 bit per global position
    uint8_t* sel;                // anchor-similar: per record index of a group head
    uint32_t* r_seq;             // nullptr (with r_pos, r_ori): compact layout, only r_pay
    uint32_t* r_pos;
    int8_t* r_ori;
    uint32_t* r_pay;             // optional: payload (position << 1 | strand) per fragment
    uint64_t* r_block_off;
    uint64_t* r_block_hash;
    // units [unit_first, unit_end) of this launch (unit_end 0: all of them): the
    // pipelined host-buffer call sorts and emits the units range by range, copying a
    // finished range to the host while the next one is sorted
    uint32_t unit_first, unit_end;
    uint64_t* unit_tot;          // per unit: blocks << 31 | fragments, then its prefix
    uint32_t* unit_wtot;         // per unit and warp of the sort CTA: blocks << 16 | fragments
    uint64_t* summary;
};

size_t bucket_summary_words();
uint32_t bucket_max_windows(const MsdPlan& plan, uint64_t max_records);
// sort (+ found bits / count), scan of the unit totals, emit
enum {
    kBucketUnits, kBucketSort, kBucketMark, kBucketCount, kBucketScan, kBucketEmit,
    kBucketEmitSimilar, kBucketFinal
};
void launch_bucket_stage(const BucketArgs& a, int stage, uint64_t max_records,
                         cudaStream_t s, uint32_t* launches);
// dst = OR of n_src (<= 8) bitmaps of `bytes` bytes (a multiple of 16), peer pointers
void launch_found_or(const uint32_t* const* srcs, int n_src, uint32_t* dst, size_t bytes,
                     cudaStream_t s, uint32_t* launches);
// ev (optional, 1 event): recorded after the sort (+ count) kernels
void launch_bucket(const BucketArgs& a, bool similar, uint64_t max_records,
                   cudaStream_t s, cudaEvent_t* ev, uint32_t* launches);

// ---- ingest (ingest_kernels.cu): raw text -> to_atgcn -> padded ASCII layout -----------
struct IngestArgs {
    const uint8_t* raw;
    uint64_t n_raw;
    uint32_t n_seqs;
    const uint64_t* raw_offsets;  // [n_seqs + 1], relative to raw
    uint32_t* tile_count;
    uint64_t* tile_offset;        // [tiles + 1] kept bytes before every tile
    uint64_t* seq_rank;           // [n_seqs + 1] kept bytes before every sequence
    const uint64_t* dst_base;     // [n_seqs] start in `ascii`, ~0 = dropped
    uint8_t* ascii;
};
uint32_t ingest_tiles(uint64_t n_raw);
void launch_ingest(const IngestArgs& a, int stage, cudaStream_t s, uint32_t* launches);

// ---- FragmentFinder (fragment_kernels.cu) --------------------------------------------
struct FragmentArgs {
    const uint32_t* packed;
    const uint32_t* nmask;
    uint64_t n_positions, n_padded;
    const uint64_t* pat_key;   // [2][n_chunks]: pattern, reverse complement of it
    const uint32_t* pat_mask;  // [2][n_chunks]: N positions
    uint32_t pattern_len, n_chunks;
    const uint64_t* seq_start;
    const uint32_t* seq_caller;
    const uint32_t* seq_lut;
    int lut_shift;
    uint64_t max_matches;
    uint32_t* tile_count;
    uint64_t* tile_offset;
    uint64_t* summary;         // [0] matches before the cut
    uint32_t* r_seq;
    uint32_t* r_pos;
    int8_t* r_ori;
};
uint32_t fragment_tiles(uint64_t n_padded);
void launch_fragment_find(const FragmentArgs& a, int stage, cudaStream_t s,
                          uint32_t* launches);

// ---- block_hash of flat block lists (block_hash_kernels.cu) ----------------------
struct BlockHashArgs {
    const uint64_t* block_off;   // [n_blocks + 1]
    uint64_t b0, b1;             // blocks of this launch
    uint64_t f_base, f_end;      // fragments of the launch (f_base = scratch origin)
    uint64_t n_fragments;
    const uint32_t* frag_seq;
    const uint32_t* min_pos;
    const uint32_t* max_pos;     // nullptr: min_pos + frag_len - 1
    uint32_t frag_len;
    const int8_t* frag_ori;
    uint32_t n_seqs;
    const uint32_t* seq_rank;    // [n_seqs] rank of the sequence's name
    const uint64_t* name_words;  // names in rank order, 8 bytes per word
    const uint32_t* name_pos;    // [ranks] first word of the name
    const uint32_t* name_len;    // [ranks] bytes
    const unsigned char* name_slow;  // [ranks] or nullptr: names ordered by id strings
    const uint16_t* bcd4;        // [10000] decimal digits of v, one per nibble
    void* scratch;               // 2 IdKey per fragment of the launch
    uint64_t* hashes;            // [n_blocks]
    uint64_t* xor_out;           // XOR of the hashes of blocks with > 1 fragment
    uint32_t* error;             // bit 0: bad offsets, bit 1: bad sequence index
};
void launch_block_hash_cuts(const uint64_t* block_off, uint64_t n_blocks, uint64_t cap,
                            uint32_t n_cuts, uint64_t* cut_block, uint64_t* cut_frag,
                            cudaStream_t s);
void launch_block_hash(const BlockHashArgs& a, cudaStream_t s);

// thread-local error text of the C ABI (api.cu), for the other translation units
int set_last_error(int code, const std::string& msg);

// pooled page-locked host memory of the result arrays (api.cu); take throws
// std::bad_alloc
void* pinned_take(size_t bytes);
void pinned_give(void* p, size_t bytes);
// the compact resident result of a context, copied into the caller's (pinned) arrays:
// n_fragments payloads, n_blocks offsets (no closing one) and hashes
int fetch_compact_into(npge_anchor_ctx* ctx, uint32_t* pos_ori, uint64_t* block_off,
                       uint64_t* block_hash);

}  // namespace npge_b200
