/*
 * npge_anchor.h -- C ABI of the B200-native anchor finder.
 *
 * This is the drop-in boundary for ONE hot path of NPG-explorer: the
 * AnchorFinder Processor.  Everything below is plain C99 (no STL, Boost or
 * torch types); the library behind it (npge_b200/libnpge_anchor.so) is
 * hand-written sm_100a CUDA and has NO CPU fallback: without a CUDA device
 * every compute entry point returns NPGE_ANCHOR_ERR_CUDA.
 *
 * What each entry point replaces in the reference (paths relative to the
 * reference tree):
 *   npge_anchor_find          AnchorFinder::run_impl, src/algo/AnchorFinder.cpp:393-406
 *                             (BloomTG/BloomTask :99-218, FragmentTG/FragmentTask
 *                             :222-326, fragmenttg_postprocess :356-391) with the
 *                             scanner src/algo/SeqI.hpp:45-119 and the hash
 *                             src/util/make_hash.hpp:25-108
 *   npge_anchor_params        AnchorFinderOptions, AnchorFinder.cpp:59-77
 *                             (options declared at :37-53)
 *   npge_anchor_result        the Blocks/Fragments run_impl inserts into the
 *                             target BlockSet (FoundFragment::make_fragment
 *                             :240-246, loop :366-387) and the hashes it appends
 *                             to Impl::used_hashes_ (:28-35, :385)
 *   npge_anchor_upload        Sequence storage feeding the scan: the 2-bit layout
 *                             of CompactLowNSequence, src/model/Sequence.cpp:525-596
 *   npge_anchor_last_error    npge::Exception / ASSERT_* text, src/util/Exception.hpp
 *
 * Semantics: codes A0 T1 G2 C3, first base in the low bits,
 * canonical h = min(dir, rev), ori = +1 iff h == dir, windows containing N
 * skipped, window starts 0..size-k per sequence (no circular wrap), sequences
 * shorter than k ignored.  The first pass is EXACT (no Bloom false positives,
 * one logical worker), i.e. the result equals the reference's with an ideal
 * filter; see DESIGN.md for what that means for parity.
 */
#ifndef NPGE_ANCHOR_H_
#define NPGE_ANCHOR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NPGE_ANCHOR_MAX_SIZE 32 /* MAX_ANCHOR_SIZE, src/util/make_hash.hpp:19 */

enum {
    NPGE_ANCHOR_OK = 0,
    NPGE_ANCHOR_ERR_PARAMS = 1,   /* option rules violated (AnchorFinder.cpp:49-51) */
    NPGE_ANCHOR_ERR_CUDA = 2,     /* no device / CUDA runtime error */
    NPGE_ANCHOR_ERR_MEMORY = 3,   /* host or device allocation failed */
    NPGE_ANCHOR_ERR_CAPACITY = 4, /* input exceeds a documented limit */
    NPGE_ANCHOR_ERR_INPUT = 5,    /* a byte outside ATGCN (input not to_atgcn-clean) */
    NPGE_ANCHOR_ERR_STATE = 6     /* call order (e.g. find_device before upload) */
};

typedef struct npge_anchor_ctx npge_anchor_ctx; /* opaque: device, stream, scratch */

typedef struct {
    int32_t anchor_size;          /* option anchor-size, 1..32 */
    int32_t anchor_similar;       /* option anchor-similar, 0/1 */
    int64_t max_anchor_fragments; /* option max-anchor-fragments; negative = no limit
                                     (compared as size_t, AnchorFinder.cpp:361) */
} npge_anchor_params;

/* Host result. All arrays are owned by the library until
 * npge_anchor_result_free.  Blocks are in ascending canonical-hash order;
 * fragments inside a block in scan order: sequences by size descending (ties by
 * caller index, SeqBase::make_seqs), then min_pos. */
typedef struct {
    uint64_t n_blocks, n_fragments;
    const uint64_t* block_offsets; /* [n_blocks+1] into the fragment arrays */
    const uint64_t* block_hash;    /* [n_blocks] canonical k-mer of each block */
    const uint32_t* frag_seq;      /* [n_fragments] index into caller's sequences */
    const uint32_t* frag_min_pos;  /* [n_fragments]; max_pos = min_pos + k - 1 */
    const int8_t* frag_ori;        /* [n_fragments] +1 / -1 */
    uint64_t n_new_used;
    const uint64_t* new_used;      /* hashes to append to used_hashes_ (F7) */
    uint64_t digest[2];            /* order-independent digest, see below */
    void* owner_;                  /* private */
} npge_anchor_result;

/* The compact result layout: 4 bytes per fragment instead of 9 on the device and over
 * PCIe (the bus is the slowest link of the host-buffer call at parity-mode sizes).  A
 * fragment is frag_pos_ori = (position in the scan-order LAYOUT << 1) | (ori == +1);
 * the layout is the sequences with size >= anchor_size in scan order (size descending,
 * ties by caller index, SeqBase::make_seqs), one pad position after each:
 * sequence layout_seq[i] (the caller's index) occupies positions
 * [layout_start[i], layout_start[i + 1] - 1).  npge_decode_compact() below maps a
 * fragment to (seq, min_pos, ori); a caller that builds Fragment objects anyway (the
 * Processor shim) does it while it builds them.  Blocks, hashes, new_used and digest
 * are those of npge_anchor_result. */
enum { NPGE_ANCHOR_LAYOUT_FULL = 0, NPGE_ANCHOR_LAYOUT_COMPACT = 1 };
typedef struct {
    uint64_t n_blocks, n_fragments;
    const uint64_t* block_offsets; /* [n_blocks+1] into frag_pos_ori */
    const uint64_t* block_hash;    /* [n_blocks] */
    const uint32_t* frag_pos_ori;  /* [n_fragments] */
    uint32_t n_layout;
    int32_t anchor_size;           /* max_pos = min_pos + anchor_size - 1 */
    const uint64_t* layout_start;  /* [n_layout+1] */
    const uint32_t* layout_seq;    /* [n_layout] */
    uint64_t n_new_used;
    const uint64_t* new_used;
    uint64_t digest[2];
    void* owner_;
} npge_anchor_compact_result;

/* (seq, min_pos, ori) of fragment f of a compact result */
static inline void npge_decode_compact(const npge_anchor_compact_result* r, uint64_t f,
                                       uint32_t* seq, uint32_t* min_pos, int* ori) {
    const uint64_t g = r->frag_pos_ori[f] >> 1;
    uint32_t lo = 0, hi = r->n_layout; /* last i with layout_start[i] <= g */
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) / 2;
        if (r->layout_start[mid] <= g) lo = mid; else hi = mid;
    }
    *seq = r->layout_seq[lo];
    *min_pos = (uint32_t)(g - r->layout_start[lo]);
    *ori = (r->frag_pos_ori[f] & 1) ? 1 : -1;
}

/* Counters and device timings of the last find; always available, also when
 * the result itself stays on the device. */
typedef struct {
    uint64_t n_positions;   /* bases uploaded (sequences with size >= k) */
    uint64_t n_windows;     /* N-free windows = records sorted */
    uint64_t n_fragments;   /* after selection and truncation */
    uint64_t n_blocks;
    uint64_t digest[2];
    uint32_t radix_passes;
    uint32_t gpu_launches;  /* kernels launched by the last find_device */
    float ms_hist, ms_sort, ms_group, ms_total; /* CUDA-event times */
    float ms_pass[16];      /* per radix pass (pass 0 generates the keys) */
    uint64_t bytes_algorithmic; /* SURVEY 8d model: N * A(k, f, b), byte digits */
    uint64_t bytes_moved;       /* compulsory HBM bytes of the passes really run */
} npge_anchor_stats;

/* digest: for every emitted fragment, t = (seq << 33) | (min_pos << 1) |
 * (ori == +1), h = its block's hash:
 *   u = t * 0x9E3779B97F4A7C15 + h * 0xC2B2AE3D27D4EB4F;  u ^= u >> 32;
 *   v = u * 0xD6E8FEB86659FD93;                           v ^= v >> 29;
 *   digest[0] = sum u,  digest[1] = sum v   (all mod 2^64; order-independent) */

int npge_anchor_create(npge_anchor_ctx** ctx, int device);
void npge_anchor_destroy(npge_anchor_ctx* ctx);

/* Test hook: force the digit plan a much larger job would get (fine_bits =
 * log2 of the (A, B) bins, first_split = bits of pass A; 0 = automatic), so that
 * the wide-row kernels are exercised on small inputs.  Results never depend on it. */
int npge_anchor_debug_plan(npge_anchor_ctx* ctx, int32_t fine_bits, int32_t first_split);
/* Test hook, no device needed: the plan a job of n_windows records gets (bits of the two
 * partition passes, byte passes of the shared-memory sort). */
int npge_anchor_debug_plan_of(int32_t anchor_size, uint64_t n_windows, int32_t* bits_a,
                              int32_t* bits_b, int32_t* sort_passes);

/* Test hook: every device buffer of the context ends in 256 guard bytes; this reads
 * them back.  *n_damaged = buffers whose guard was overwritten (a kernel stored past the
 * end of the buffer); npge_anchor_last_error() then names them.  (Stands in for
 * compute-sanitizer memcheck, which is closed on the development pool.) */
int npge_anchor_debug_check(npge_anchor_ctx* ctx, uint32_t* n_damaged);

/* Thread-local text of the last error returned to this thread. */
const char* npge_anchor_last_error(void);

/* The whole path with HOST buffers: bases = all sequences concatenated as
 * upper-case ATGCN bytes, seq_offsets[n_seqs+1].  used_sorted = the
 * Processor's used_hashes_ state (sorted ascending), may be NULL. */
int npge_anchor_find(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                     const uint8_t* bases, const uint64_t* seq_offsets,
                     uint32_t n_seqs, const uint64_t* used_sorted,
                     uint64_t n_used, npge_anchor_result* out);
void npge_anchor_result_free(npge_anchor_result* out);

/* The same path in three steps, so that the scan can be timed with inputs
 * resident in HBM and results left there:
 *   upload       host ATGCN -> device, packed to 2 bits/base + N mask
 *   find_device  hist -> radix sort -> grouping, everything on the device
 *   fetch        device result -> host arrays (+ truncation fix-up, new_used) */
int npge_anchor_upload(npge_anchor_ctx* ctx, int32_t anchor_size,
                       const uint8_t* bases, const uint64_t* seq_offsets,
                       uint32_t n_seqs);
/* upload with Sequence::to_atgcn done on the device (Sequence.cpp:151-179): `raw`
 * is sequence text as it sits in a FASTA file -- any case, line breaks, gaps,
 * IUPAC ambiguity codes (-> N); every other byte is dropped.  clean_sizes
 * (optional, [n_seqs]) receives the size of every sequence after cleaning;
 * positions in the results refer to the cleaned sequences, as in the reference. */
int npge_anchor_upload_raw(npge_anchor_ctx* ctx, int32_t anchor_size,
                           const uint8_t* raw, const uint64_t* raw_offsets,
                           uint32_t n_seqs, uint64_t* clean_sizes);
int npge_anchor_find_device(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                            const uint64_t* used_sorted, uint64_t n_used,
                            npge_anchor_stats* stats);
int npge_anchor_fetch(npge_anchor_ctx* ctx, npge_anchor_result* out);
/* The compact layout.  set_result_layout(COMPACT) makes the following find_device calls
 * write payloads only (npge_anchor_fetch still works: it expands them on the host;
 * npge_anchor_block_hashes on the resident result does not); fetch_compact returns
 * them as they are; find_compact = upload + find_device + fetch_compact with host
 * buffers in, whatever the context's layout.  On a rank of a multi-GPU job
 * (npge_anchor_dist_*) the positions and the layout table are the job's, and new_used
 * lists every hash of the rank's range. */
int npge_anchor_set_result_layout(npge_anchor_ctx* ctx, int32_t layout);
int npge_anchor_fetch_compact(npge_anchor_ctx* ctx, npge_anchor_compact_result* out);
int npge_anchor_find_compact(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                             const uint8_t* bases, const uint64_t* seq_offsets,
                             uint32_t n_seqs, const uint64_t* used_sorted, uint64_t n_used,
                             npge_anchor_compact_result* out);
void npge_anchor_compact_result_free(npge_anchor_compact_result* out);
int npge_anchor_get_stats(npge_anchor_ctx* ctx, npge_anchor_stats* stats);

/* ---- several GPUs, one context per GPU -------------------------------
 * The job's sequences, in scan order (size descending, ties by caller index) with one
 * pad position after each, form ONE position space.  It is cut into equal ranges:
 * every rank generates the windows that START in its range (it also holds the k - 1
 * bases after it, the halo), so a sequence may be split between GPUs -- the
 * reference's unit of parallel work is a whole sequence (AnchorFinder.hpp:27-29).
 * Canonical k-mers are then repartitioned by hash range between pass A and pass B.
 * The caller owns the rendezvous between the ranks (npge_b200/dist.py: processes and
 * torch.distributed; npge_anchor_multi_*: threads of one process); these entry points
 * are the per-rank stages between them:
 *   dist_upload_range  this rank's range of the position space
 *   dist_count_a       count A; returns the records per A bucket
 *   dist_scatter_local pass A into local memory (asynchronous)
 *   [all-gather of the histograms; also the step's barrier]
 *   dist_exchange      the job's tables (who owns which buckets, where every slice
 *                      lands, the (bucket, source) segments of pass B: dist_plan.h,
 *                      computed on the host in microseconds) and the copies of this
 *                      rank's slices into the owners' receive buffers, mapped through
 *                      CUDA IPC (dist_peer_alloc / dist_peer_open) or peer access:
 *                      one copy per destination, all in flight at once
 *   [barrier: every rank's copies have landed]
 *   dist_sort_planned  stage 0: pass B over the received segments, bucket sort, and
 *                      either the anchor count (exact mode: totals valid) or the found
 *                      bits (anchor-similar: totals are 0)
 *                      [anchor-similar: all-reduce of the found bitmap; every bit
 *                      is set by exactly one rank, so SUM of the words is OR]
 *                      stage 1 (anchor-similar only): candidate rule + count
 *                      returns the rank's totals {fragments, blocks} before the cut
 *   [all-gather of the totals -> every rank's share of max-anchor-fragments]
 *   dist_emit          writes the rank's anchors; npge_anchor_fetch then works as usual,
 *                      except that new_used holds EVERY hash of the rank's list (the
 *                      caller concatenates the ranks and drops the job's first one)
 * Positions in the records are job-wide, so every rank's result is literally the
 * slice of the one-GPU result for its hash range.
 * Variants kept for process groups without peer access and for comparison:
 * dist_pass_a + dist_buffers + an all-to-all by the caller + dist_sort with the
 * caller's tables ("collective"); dist_scatter_peer, where pass A itself writes every
 * bucket into the owner's memory over NVLink ("peer").
 */
typedef struct {
    void* send_rec;      /* u64 records after pass A, bucket a at a_start[a] */
    void* send_hi;       /* u32, NULL unless wide (anchor-size >= 22) */
    void* recv_rec;      /* room for n_recv records */
    void* recv_hi;
    void* found_bits;    /* u32 words, one bit per job-wide position */
    uint64_t found_words;
    int32_t wide;
    int32_t n_bins;      /* A buckets (1 << capa) */
} npge_anchor_dist_ptrs;

/* bases: host or device memory; piece i = bytes [src_off[i], src_off[i] + len[i]) placed
 * at LOCAL position dst_pos[i] (job-wide position g0 + dst_pos[i]); local positions no
 * piece covers are pads.  The rank owns the windows starting in [g0, g0 + n_own) and
 * holds [g0, g0 + n_have), n_own <= n_have < n_own + anchor_size.
 * global_start[n_global + 1]: first job-wide position of every sequence in scan order
 * (size + 1 apart: the pads); global_caller[n_global]: the caller's index of each. */
int npge_anchor_dist_upload_range(npge_anchor_ctx* ctx, int32_t anchor_size,
                                  const uint8_t* bases, uint32_t n_pieces,
                                  const uint64_t* src_off, const uint64_t* dst_pos,
                                  const uint64_t* len, uint64_t g0, uint64_t n_own,
                                  uint64_t n_have, const uint64_t* global_start,
                                  const uint32_t* global_caller, uint32_t n_global);
int npge_anchor_dist_pass_a(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                            uint32_t* hist_out, uint32_t hist_cap, uint32_t* n_bins);
/* Receive buffers shared between the ranks (CUDA IPC), and the "peer" variant of the
 * exchange: pass A scatters every bucket straight into the owning rank's receive
 * buffer over NVLink.
 *   dist_peer_alloc    (re)allocates this rank's receive buffers for `capacity`
 *                      records and returns their IPC handles
 *   [all-gather of the handles]
 *   dist_peer_open     maps the other ranks' buffers; dist_peer_close unmaps them
 *                      (before any rank reallocates)
 *   dist_count_a       count A only; returns the records per A bucket
 *   [all-gather of the histograms, as above]
 *   dist_scatter_peer  pass A; bucket a goes to rank dest_rank[a], at record
 *                      index dest_off[a] + (index in this rank's A-sorted order)
 *   [barrier: every rank's scatter has finished]  then dist_buffers / dist_sort */
#define NPGE_ANCHOR_IPC_BYTES 128
int npge_anchor_dist_peer_alloc(npge_anchor_ctx* ctx, uint64_t capacity,
                                uint8_t handles_out[NPGE_ANCHOR_IPC_BYTES]);
int npge_anchor_dist_peer_open(npge_anchor_ctx* ctx, uint32_t world, uint32_t rank,
                               const uint8_t* handles_all);
int npge_anchor_dist_peer_close(npge_anchor_ctx* ctx);
int npge_anchor_dist_count_a(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                             uint32_t* hist_out, uint32_t hist_cap, uint32_t* n_bins);
int npge_anchor_dist_scatter_peer(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                  const uint32_t* dest_rank, const int64_t* dest_off,
                                  uint32_t n_bins);
/* dist_count_a + dist_scatter_local is dist_pass_a in two steps; the second only
 * enqueues pass A, so that the all-gather of the histograms overlaps it. */
int npge_anchor_dist_scatter_local(npge_anchor_ctx* ctx, const npge_anchor_params* p);
/* Pass A of the calls that follow runs in `chunks` chunks of tiles (1..8; default 1):
 * dist_count_a then returns chunks * n_bins counts ([chunk][bucket]), and dist_exchange
 * copies every finished chunk while the next one is ranked.  The same on every rank. */
int npge_anchor_dist_set_chunks(npge_anchor_ctx* ctx, int32_t chunks);
/* hist_all[world][chunks][bins]: the all-gathered results of dist_count_a.  n_recv_out[world]
 * (optional): records every rank receives.  NPGE_ANCHOR_ERR_CAPACITY when a receive
 * buffer is too small (n_recv_out valid: regrow with dist_peer_alloc and call again). */
int npge_anchor_dist_exchange(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                              const uint32_t* hist_all, uint32_t world, uint32_t rank,
                              uint64_t* n_recv_out);
int npge_anchor_dist_sort_planned(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                  int32_t stage, const uint64_t* used_sorted,
                                  uint64_t n_used, uint64_t totals[2]);
/* The tables of dist_exchange alone, on the host (no device needed; tests).
 * hist_all[world * chunks][bins]: source rank * chunks + c is chunk c of a rank.
 * cuts [world + 1], n_recv [world], dest_rank / dest_off [bins] of `rank`'s chunk 0 as a
 * sender, and its segment tables as a receiver (seg_* sized world * chunks * bins,
 * seg_tile one more, a_tile0 / a_start [bins + 1]); any pointer may be NULL. */
int npge_anchor_dist_plan_tables(const uint32_t* hist_all, uint32_t world, uint32_t chunks,
                                 uint32_t bins, uint32_t rank, uint32_t* cuts,
                                 uint64_t* n_recv,
                                 uint32_t* dest_rank, int64_t* dest_off, uint32_t* n_seg,
                                 uint32_t* seg_begin, uint32_t* seg_end, uint32_t* seg_a,
                                 uint32_t* seg_tile, uint32_t* a_tile0, uint32_t* a_start);
int npge_anchor_dist_buffers(npge_anchor_ctx* ctx, uint64_t n_recv,
                             npge_anchor_dist_ptrs* out);
int npge_anchor_dist_sort(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                          int32_t stage, uint64_t n_recv, uint32_t n_seg,
                          const uint32_t* seg_begin, const uint32_t* seg_end,
                          const uint32_t* seg_a, const uint32_t* seg_tile,
                          const uint32_t* a_tile0, const uint32_t* a_start,
                          const uint64_t* used_sorted, uint64_t n_used,
                          uint64_t totals[2]);
int npge_anchor_dist_emit(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                          uint64_t local_limit, npge_anchor_stats* stats);

/* Contexts of ONE process (npge_anchor_multi_* below): the other ranks' receive buffers
 * are plain device pointers (rec[r], hi[r]; devices[r] = their CUDA device), reached
 * through peer access, which is enabled here.  After dist_peer_alloc on every rank. */
int npge_anchor_dist_peer_attach(npge_anchor_ctx* ctx, uint32_t world, uint32_t rank,
                                 void* const* rec, void* const* hi, const int* devices);
/* anchor-similar inside one process: this rank's found bitmap := OR of all ranks'
 * (found[r]: their current bitmaps), built in a second buffer; dist_found_swap on
 * every rank once all of them have finished reading. */
int npge_anchor_dist_found_or(npge_anchor_ctx* ctx, uint32_t world, const void* const* found);
int npge_anchor_dist_found_swap(npge_anchor_ctx* ctx);

/* ---- several GPUs of one process: the host NPGE's Processor calls -----------------------
 * One context and one host thread per GPU, the same per-rank stages as above with
 * in-process rendezvous (a barrier, a shared histogram table, peer access).  The
 * result is the one-GPU result, bit for bit: blocks in ascending hash order, the ranks'
 * slices joined.  device_ids may repeat a device (several ranks on one GPU: tests).
 *   multi_find          host buffers in, host arrays out (AnchorFinder::run_impl,
 *                       src/algo/AnchorFinder.cpp:393-406)
 *   multi_upload / multi_find_device / multi_fetch   the same in three steps; the packed
 *                       genome stays resident, so that a Processor re-run on the same
 *                       sequences (AnchorLoop, src/algo/MetaProcessor.cpp:32-38) skips
 *                       the upload
 *   multi_get_stats     rank >= 0: that rank's stats; rank < 0: the job's (counts and
 *                       digest summed, times = max over ranks) */
typedef struct npge_anchor_multi npge_anchor_multi;
int npge_anchor_multi_create(npge_anchor_multi** m, const int* device_ids, int n_devices);
void npge_anchor_multi_destroy(npge_anchor_multi* m);
int npge_anchor_multi_devices(const npge_anchor_multi* m);
int npge_anchor_multi_find(npge_anchor_multi* m, const npge_anchor_params* p,
                           const uint8_t* bases, const uint64_t* seq_offsets, uint32_t n_seqs,
                           const uint64_t* used_sorted, uint64_t n_used,
                           npge_anchor_result* out);
int npge_anchor_multi_upload(npge_anchor_multi* m, int32_t anchor_size, const uint8_t* bases,
                             const uint64_t* seq_offsets, uint32_t n_seqs);
int npge_anchor_multi_find_device(npge_anchor_multi* m, const npge_anchor_params* p,
                                  const uint64_t* used_sorted, uint64_t n_used,
                                  npge_anchor_stats* job_stats);
int npge_anchor_multi_fetch(npge_anchor_multi* m, npge_anchor_result* out);
void npge_anchor_multi_result_free(npge_anchor_result* out);
/* The compact layout over several GPUs: after multi_set_result_layout(COMPACT) every
 * rank writes payloads only, and multi_fetch_compact copies every rank's slice straight
 * into its place in the job's arrays (4 bytes per fragment over every GPU's PCIe link). */
int npge_anchor_multi_set_result_layout(npge_anchor_multi* m, int32_t layout);
int npge_anchor_multi_fetch_compact(npge_anchor_multi* m, npge_anchor_compact_result* out);
void npge_anchor_multi_compact_result_free(npge_anchor_compact_result* out);
int npge_anchor_multi_get_stats(npge_anchor_multi* m, int rank, npge_anchor_stats* stats);
int npge_anchor_multi_debug_check(npge_anchor_multi* m, uint32_t* n_damaged); /* all ranks */

/* ---- FragmentFinder (src/algo/FragmentFinder.cpp:42-171) ---------------------------
 * Every exact occurrence of `pattern` (ATGCN bytes, N matches only N) on either
 * strand of the sequences: ori = +1 where the text equals the pattern, -1 where
 * its reverse complement does (both at a palindrome).  Order: sequences by size
 * descending (ties by caller index), position ascending, +1 before -1 -- the
 * order of a one-worker run of the reference -- cut to max_matches (option
 * max-matches, >= 1).  One single-fragment Block per match on the host side. */
typedef struct {
    uint64_t n_matches;            /* min(total_matches, max_matches) */
    uint64_t total_matches;        /* before the cut */
    const uint32_t* frag_seq;      /* index of the sequence as passed */
    const uint32_t* frag_min_pos;  /* max_pos = min_pos + pattern_len - 1 */
    const int8_t* frag_ori;
    void* owner_;
} npge_fragment_result;
int npge_fragment_find(npge_anchor_ctx* ctx, const uint8_t* pattern, uint32_t pattern_len,
                       int64_t max_matches, const uint8_t* bases,
                       const uint64_t* seq_offsets, uint32_t n_seqs,
                       npge_fragment_result* out);
/* on the sequences of the last npge_anchor_upload (its anchor_size must not
 * exceed pattern_len: shorter sequences were dropped at upload) */
int npge_fragment_find_device(npge_anchor_ctx* ctx, const uint8_t* pattern,
                              uint32_t pattern_len, int64_t max_matches,
                              npge_fragment_result* out);
void npge_fragment_result_free(npge_fragment_result* out);

/* npge_fragment_find_device without any allocation.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The kernels' scratch is a file-scope device array of fixed size (about 193 KiB, held by
 * every GPU that loads the library until exit), hence the limits: at most
 * NPGE_FRAGMENT_NOALLOC_MAX_POSITIONS uploaded positions (bases + one pad per sequence,
 * rounded up to 4096) and patterns of at most NPGE_FRAGMENT_NOALLOC_MAX_PATTERN bases;
 * beyond either the call fails with NPGE_ANCHOR_ERR_CAPACITY.
 *   d_frag_seq / d_frag_min_pos / d_frag_ori   DEVICE arrays of `capacity` elements the
 *       caller allocated; at most min(max_matches, capacity) matches are written, in the
 *       order of npge_fragment_find_device
 *   n_matches, total_matches   host pointers (optional): matches written / found
 * Calls no cudaMalloc, cudaFree, cudaHostAlloc or other allocating routine.  All calls on
 * one GPU share the scratch array: calls that overlap on the GPU race, serialise them. */
#define NPGE_FRAGMENT_NOALLOC_MAX_POSITIONS (1u << 26)
#define NPGE_FRAGMENT_NOALLOC_MAX_PATTERN 1024
int npge_fragment_find_device_noalloc(npge_anchor_ctx* ctx, const uint8_t* pattern,
                                      uint32_t pattern_len, int64_t max_matches,
                                      uint32_t* d_frag_seq, uint32_t* d_frag_min_pos,
                                      int8_t* d_frag_ori, uint64_t capacity,
                                      uint64_t* n_matches, uint64_t* total_matches);

/* ---- block_hash / blockset_hash (src/model/block_hash.cpp:29-54,125-130) ------------
 * The reference's own fingerprint of a block: the ids "name_begin_last"
 * (Fragment::id, src/model/Fragment.cpp:173-183) of its fragments, as they are
 * and inverted, sorted; the smaller list joined with ' ' and folded with
 * multiply / xor.  MoveUnchanged (src/algo/MoveUnchanged.cpp:37-49) keeps a
 * sorted list of these values between the iterations of AnchorLoop
 * (src/algo/lua_lib.lua:719-721) and moves the blocks whose hash it has seen;
 * blockset_hash is their XOR over the blocks with more than one fragment.
 * Computed on the device for all blocks of a flat block list at once.
 *
 *  names[s]          name of sequence s (Sequence::name()), NUL-terminated; no
 *                    bytes <= ' ' (NPGE_ANCHOR_ERR_PARAMS otherwise).  Any other
 *                    name set works; where one name + "_" is a prefix of
 *                    another ("s1", "s1_2") the ids of those sequences are
 *                    ordered by their full strings on the device (slower)
 *  block_offsets     [n_blocks + 1] into the fragment arrays, ascending
 *  frag_max_pos      NULL: every fragment is frag_len long (anchors)
 *  hashes            [n_blocks] out (may be NULL)
 *  blockset_hash     out (may be NULL)
 * All pointers are host memory. */
int npge_block_hashes(npge_anchor_ctx* ctx, const char* const* names, uint32_t n_seqs,
                      uint64_t n_blocks, const uint64_t* block_offsets,
                      const uint32_t* frag_seq, const uint32_t* frag_min_pos,
                      const uint32_t* frag_max_pos, uint32_t frag_len,
                      const int8_t* frag_ori, uint64_t* hashes, uint64_t* blockset_hash);
/* The same over the device-resident result of the last npge_anchor_find_device
 * (nothing but the hashes crosses PCIe).  names[s]: the sequences as passed to
 * npge_anchor_upload.  NPGE_ANCHOR_ERR_STATE when max_anchor_fragments cut the
 * list inside a block: that block is completed on the host by
 * npge_anchor_fetch, so hash the fetched arrays with npge_block_hashes. */
int npge_anchor_block_hashes(npge_anchor_ctx* ctx, const char* const* names,
                             uint32_t n_seqs, uint64_t* hashes, uint64_t* blockset_hash);

/* device time (ms, CUDA events around the hash kernels) of the last call of
 * either function on this context */
float npge_block_hashes_last_ms(npge_anchor_ctx* ctx);

/* Library self-description: "npge_anchor <version> sm_100a" */
const char* npge_anchor_version(void);

#ifdef __cplusplus
}
#endif
#endif
