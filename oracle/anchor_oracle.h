/*
 * oracle/anchor_oracle.h -- C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a CPU restatement of NPGE's AnchorFinder
 * (reference: src/algo/AnchorFinder.cpp, src/algo/SeqI.hpp,
 * src/algo/BloomFilter.cpp, src/util/make_hash.hpp, src/util/complement.cpp,
 * src/model/block_hash.cpp).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (npge_b200/, include/npge_anchor.h) never links or calls anything here.
 *
 * Parity pinning: every golden vector the reference's own tests hold for this
 * path is replayed against this oracle in tests/test_oracle_golden.py, and the
 * hash arithmetic is cross-checked against the reference's own headers
 * compiled in place (oracle/_ref, see oracle/Makefile).
 */
#ifndef NPGE_ANCHOR_ORACLE_H_
#define NPGE_ANCHOR_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- hash arithmetic (src/util/make_hash.hpp, src/util/complement.cpp) ---- */

/* make_hash(start, length, ori): make_hash.hpp:72-83. For ori=-1 `start`
 * points at the LAST letter of the fragment and letters are read backwards. */
uint64_t oracle_make_hash(const char* start, int length, int ori);
/* reuse_hash: make_hash.hpp:88-108 */
uint64_t oracle_reuse_hash(uint64_t old_hash, int length, char remove_char,
                           char add_char, int forward);
/* complement_hash: complement.cpp:24-50 */
uint64_t oracle_complement_hash(uint64_t hash, int letters_number);
/* Sequence::to_atgcn: Sequence.cpp:151-179; in place, returns new length */
size_t oracle_to_atgcn(char* data, size_t size);
/* BloomFilter::optimal_bits / optimal_hashes: BloomFilter.cpp:147-164 */
size_t oracle_bloom_optimal_bits(size_t members, double error_prob);
size_t oracle_bloom_optimal_hashes(size_t members, size_t bits);

/* ---- AnchorFinder ---- */

typedef struct {
    int32_t anchor_size;           /* option anchor-size, 1..32 */
    int32_t anchor_similar;        /* option anchor-similar */
    int64_t max_anchor_fragments;  /* option max-anchor-fragments */
    double anchor_fp;              /* option anchor-fp (faithful only) */
    int32_t workers;               /* Processor workers (faithful only) */
    uint32_t bloom_seed;           /* replaces make_seed() (rand_name.cpp:16-23) */
    int32_t exact_bloom;           /* 1: replace the Bloom filter by an exact set */
    int32_t asserts;               /* 1: keep per-base range checks (NPGE_ASSERTS=ON) */
} oracle_params;

typedef struct {
    uint64_t n_blocks, n_fragments;
    uint64_t* block_offsets;       /* [n_blocks+1] */
    uint64_t* block_hash;          /* [n_blocks] canonical k-mer of each block */
    uint32_t* frag_seq;            /* [n_fragments] caller sequence index */
    uint32_t* frag_min_pos;        /* [n_fragments] */
    int8_t* frag_ori;              /* [n_fragments] +1 / -1 */
    uint64_t n_new_used;
    uint64_t* new_used;            /* hashes appended to used_hashes_ by this run */
    double seconds;                /* wall time of run_impl's restatement */
} oracle_result;

/* Sequences are given as one concatenated ATGCN byte array with offsets[S+1].
 * genome_ids[s] is an integer label for Sequence::genome() (Sequence.cpp:193-202),
 * -1 meaning "" ; it is used only for the Bloom sizing (estimate_length).
 * used_sorted: the Processor's used_hashes_ state before the run (sorted). */
int oracle_find_faithful(const oracle_params* p, const char* bases,
                         const uint64_t* offsets, uint32_t n_seqs,
                         const int32_t* genome_ids,
                         const uint64_t* used_sorted, uint64_t n_used,
                         oracle_result* out);

/* Sort-based exact-set semantics: the reference's
 * output with an ideal (FP=0) first pass and one worker. */
int oracle_find_ideal(const oracle_params* p, const char* bases,
                      const uint64_t* offsets, uint32_t n_seqs,
                      const uint64_t* used_sorted, uint64_t n_used,
                      oracle_result* out);

void oracle_result_free(oracle_result* r);

/* The same exact semantics in 2^slice_bits hash-range slices on `threads`
 * threads (<= 0: all cores): O(N / slices) memory per thread, counts and the
 * 128-bit digest only.  For BASELINE.json's large configurations.  With a
 * finite max-anchor-fragments that is reached before the last slice,
 * total_selected and n_windows are 0 (not counted). */
typedef struct {
    uint64_t n_windows, n_fragments, n_blocks, total_selected, n_new_used;
    uint64_t digest[2];
    double seconds;
} oracle_summary;
int oracle_find_ideal_sliced(const oracle_params* p, const char* bases,
                             const uint64_t* offsets, uint32_t n_seqs,
                             const uint64_t* used_sorted, uint64_t n_used,
                             int32_t slice_bits, int32_t threads,
                             oracle_summary* out);

/* blockset_hash (block_hash.cpp:29-54,125-130) over blocks given as flat
 * arrays; names[s] is the sequence name used by Fragment::id() (Fragment.cpp:173-183) */
uint64_t oracle_blockset_hash(const char* const* names, uint64_t n_blocks,
                              const uint64_t* block_offsets,
                              const uint32_t* frag_seq,
                              const uint32_t* frag_min_pos,
                              const uint32_t* frag_max_pos,
                              const int8_t* frag_ori);

/* block_hash (block_hash.cpp:29-54) of every block, hashes[n_blocks] */
void oracle_block_hashes(const char* const* names, uint64_t n_blocks,
                         const uint64_t* block_offsets,
                         const uint32_t* frag_seq,
                         const uint32_t* frag_min_pos,
                         const uint32_t* frag_max_pos,
                         const int8_t* frag_ori, uint64_t* hashes);

/* Order-independent 128-bit digest of (seq, min_pos, ori, h) tuples; the same
 * definition as npge_anchor_result.digest (include/npge_anchor.h). */
void oracle_digest(const oracle_result* r, uint64_t digest[2]);

#ifdef __cplusplus
}
#endif
#endif
