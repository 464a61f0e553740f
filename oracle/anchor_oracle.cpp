/*
 * oracle/anchor_oracle.cpp -- CPU restatement of NPGE's AnchorFinder.
 *
 * TEST INFRASTRUCTURE ONLY (see anchor_oracle.h).  Every function cites the
 * reference file:line it follows (paths relative to the reference tree).
 * Nothing here is copied from the reference; the algorithms are restated.
 *
 * Two oracles:
 *   oracle_find_faithful -- two passes, Bloom filter, one task per sequence on
 *       W threads, (hash, seq, pos') sort, truncation, block building, the
 *       used_hashes_ quirk.  Deliberate deviations, both documented in
 *       the sort tie-break uses the caller's sequence index instead
 *       of the Sequence* heap address, and make_seed() is injectable.
 *   oracle_find_ideal -- sort-based exact semantics (no Bloom false positives,
 *       one worker); the ground truth the GPU path is compared with.
 */
#include "anchor_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_set>
#include <vector>

namespace {

typedef uint64_t hash_t;

/* ---- src/util/char_to_size.hpp:24-36,53-56 ---- */
inline size_t letter_code(char c) {
    switch (c) {
    case 'A': return 0;
    case 'T': return 1;
    case 'G': return 2;
    case 'C': return 3;
    default: return 4;  /* 'N' */
    }
}

/* src/util/complement.hpp:19-32 (input is already upper-case ATGCN) */
inline char complement_char(char c) {
    switch (c) {
    case 'A': return 'T';
    case 'T': return 'A';
    case 'G': return 'C';
    case 'C': return 'G';
    default: return c;
    }
}

const int HASH_BITS = 64;
const int MAX_ANCHOR = 32;  /* make_hash.hpp:19 */

/* make_hash.hpp:25-27 */
inline int shift_in_hash(int pos) {
    return (pos * 2) % HASH_BITS;
}

/* make_hash.hpp:29-42: XOR of 2-bit codes, N -> 4&3 = 0, reverse strand
 * complements every letter except N. `at(j)` yields the j-th letter in
 * fragment order. */
template <typename At>
hash_t hash_of(At at, int length, int ori) {
    hash_t result = 0;
    for (int j = 0; j < length; j++) {
        char c = at(j);
        size_t s = letter_code(c) & 3;
        if (ori == -1 && c != 'N') {
            s ^= 1;
        }
        result ^= hash_t(s) << shift_in_hash(j);
    }
    return result;
}

/* make_hash.hpp:88-108 */
inline hash_t roll_hash(hash_t h, int length, char remove_char, char add_char,
                        bool forward) {
    hash_t remove = letter_code(remove_char) & 3;
    h ^= remove << shift_in_hash(forward ? 0 : length - 1);
    int occupied = std::min(2 * length, HASH_BITS);
    if (forward) {
        h = (h >> 2) | ((h & 3) << (occupied - 2));
    } else {
        h = (h << 2) | ((h >> (occupied - 2)) & 3);
    }
    hash_t add = letter_code(add_char) & 3;
    h ^= add << shift_in_hash(forward ? length - 1 : 0);
    return h;
}

/* complement.cpp:24-50 */
hash_t revcomp_hash(hash_t hash, int letters) {
    hash_t result = 0;
    int digits = std::min(letters, MAX_ANCHOR);
    size_t r = (letters / MAX_ANCHOR) % 2;
    size_t swap_pos = letters % MAX_ANCHOR;
    for (int i = 0; i < digits; i++) {
        int dest = (letters - 1 - i) % MAX_ANCHOR;
        hash_t two = hash & 3;
        two ^= (size_t(i) < swap_pos) ? (r ^ 1) : r;
        result ^= two << (2 * dest);
        hash >>= 2;
    }
    return result;
}

/* BloomFilter.cpp:147-156. NB the reference computes in `int`. */
size_t bloom_bits(size_t members, double error_prob) {
    const double ln2 = std::log(2.0);
    int result = members * (-std::log(error_prob) / (ln2 * ln2)) + 0.5;
    if (result % 2 == 0) {
        result += 1;
    }
    if (result < 1) {
        result = 1;
    }
    return result;
}

/* BloomFilter.cpp:158-164 */
size_t bloom_hashes(size_t members, size_t bits) {
    const double ln2 = std::log(2.0);
    int result = round(ln2 * bits / members);
    if (result < 1) {
        result = 1;
    }
    return result;
}

}  // namespace

extern "C" uint64_t oracle_make_hash(const char* start, int length, int ori) {
    if (ori == 1) {
        return hash_of([&](int j) { return start[j]; }, length, 1);
    }
    return hash_of([&](int j) { return start[-j]; }, length, -1);
}

extern "C" uint64_t oracle_reuse_hash(uint64_t old_hash, int length,
                                      char remove_char, char add_char,
                                      int forward) {
    return roll_hash(old_hash, length, remove_char, add_char, forward != 0);
}

extern "C" uint64_t oracle_complement_hash(uint64_t hash, int letters_number) {
    return revcomp_hash(hash, letters_number);
}

/* Sequence.cpp:151-179 */
extern "C" size_t oracle_to_atgcn(char* data, size_t size) {
    size_t d = 0;
    for (size_t f = 0; f < size; f++) {
        char c = toupper((unsigned char)data[f]);
        bool ok = false;
        if (c == 'A' || c == 'T' || c == 'G' || c == 'C' || c == 'N') {
            ok = true;
        } else if (strchr("RYMKWSBVHD", c) && c != 0) {
            c = 'N';
            ok = true;
        }
        if (ok) {
            data[d++] = c;
        }
    }
    return d;
}

extern "C" size_t oracle_bloom_optimal_bits(size_t members, double error_prob) {
    return bloom_bits(members, error_prob);
}

extern "C" size_t oracle_bloom_optimal_hashes(size_t members, size_t bits) {
    return bloom_hashes(members, bits);
}

/* ====================================================================== */
/* Shared model of the input and of the scanner state                      */
/* ====================================================================== */
namespace {

struct SeqView {
    const char* data;
    size_t size;
    uint32_t index;  /* caller index */
    bool asserts;

    /* Sequence::char_at, Sequence.cpp:230-237: range ASSERTs then the read */
    inline char char_at(long i) const {
        if (asserts) {
            if (i < 0 || size_t(i) >= size) {
                throw std::logic_error("Index out of sequence");
            }
        }
        return data[i];
    }
};

/* SeqI.hpp:70-119 */
struct Scanner {
    const SeqView* seq;
    long pos;
    int ns;
    int anchor;
    hash_t dir, rev;

    Scanner(const SeqView* s, int k) : seq(s), pos(0), ns(0), anchor(k),
        dir(0), rev(0) {
    }

    /* SeqI.hpp:86-96 (+ ns_in_fragment :60-68, Sequence::hash_impl
     * Sequence.cpp:288-299) */
    void init_state() {
        ns = 0;
        for (int i = 0; i < anchor; i++) {
            if (seq->char_at(i) == 'N') {
                ns += 1;
            }
        }
        const SeqView* s = seq;
        int k = anchor;
        dir = hash_of([&](int j) { return s->data[j]; }, k, 1);
        rev = hash_of([&](int j) { return s->data[k - 1 - j]; }, k, -1);
        if (ns == 0 && rev != revcomp_hash(dir, k)) {
            throw std::logic_error("rev != complement_hash(dir)");
        }
    }

    /* SeqI.hpp:105-119 */
    void next_hash() {
        char remove_char = seq->char_at(pos);
        char add_char = seq->char_at(pos + anchor);
        pos += 1;
        if (remove_char == 'N') {
            ns -= 1;
        }
        if (add_char == 'N') {
            ns += 1;
        }
        dir = roll_hash(dir, anchor, remove_char, add_char, true);
        remove_char = complement_char(remove_char);
        add_char = complement_char(add_char);
        rev = roll_hash(rev, anchor, remove_char, add_char, false);
    }
};

/* SeqBase::make_seqs, SeqI.hpp:45-57: sequences with size >= k, by size
 * descending.  The reference's std::sort is unstable over a pointer-ordered
 * set; ties are broken here by caller index (documented deviation). */
std::vector<SeqView> make_seqs(const char* bases, const uint64_t* offsets,
                               uint32_t n_seqs, int k, bool asserts) {
    std::vector<SeqView> seqs;
    for (uint32_t s = 0; s < n_seqs; s++) {
        size_t size = offsets[s + 1] - offsets[s];
        if (size >= size_t(k)) {
            SeqView v = {bases + offsets[s], size, s, asserts};
            seqs.push_back(v);
        }
    }
    std::stable_sort(seqs.begin(), seqs.end(),
    [](const SeqView& a, const SeqView& b) {
        return a.size > b.size;
    });
    return seqs;
}

/* FoundFragment, AnchorFinder.cpp:222-247 */
struct Found {
    hash_t hash;
    uint32_t seq;   /* caller index (reference: Sequence*) */
    uint64_t pos;   /* pos' : min_pos, + size if ori == -1 */

    bool operator<(const Found& o) const {
        if (hash != o.hash) return hash < o.hash;
        if (seq != o.seq) return seq < o.seq;
        return pos < o.pos;
    }
};

/* fragmenttg_postprocess, AnchorFinder.cpp:356-391, writing flat arrays
 * instead of Block/Fragment objects (make_fragment :240-246). */
void build_blocks(std::vector<Found>& ffs, const oracle_params* p,
                  const uint64_t* offsets, oracle_result* out) {
    std::sort(ffs.begin(), ffs.end());
    if (p->max_anchor_fragments >= 0 &&
            ffs.size() > size_t(p->max_anchor_fragments)) {
        ffs.resize(p->max_anchor_fragments);
    }
    std::vector<uint64_t> block_offsets, block_hash, new_used;
    std::vector<uint32_t> frag_seq, frag_min;
    std::vector<int8_t> frag_ori;
    auto emit = [&](const Found& f) {
        uint64_t size = offsets[f.seq + 1] - offsets[f.seq];
        bool direct = f.pos < size;
        frag_seq.push_back(f.seq);
        frag_min.push_back(uint32_t(direct ? f.pos : f.pos - size));
        frag_ori.push_back(direct ? 1 : -1);
    };
    const Found* prev = 0;
    bool in_block = false;
    for (size_t i = 0; i < ffs.size(); i++) {
        const Found& ff = ffs[i];
        if (prev == 0) {
            prev = &ff;
        } else if (prev->hash == ff.hash) {
            if (!in_block) {
                in_block = true;
                block_offsets.push_back(frag_seq.size());
                block_hash.push_back(ff.hash);
                emit(*prev);
            }
            emit(ff);
        } else {
            prev = &ff;
            in_block = false;
            new_used.push_back(ff.hash);  /* never the first hash (F7) */
        }
    }
    block_offsets.push_back(frag_seq.size());
    out->n_blocks = block_hash.size();
    out->n_fragments = frag_seq.size();
    auto dup64 = [](const std::vector<uint64_t>& v) {
        uint64_t* r = (uint64_t*)malloc(sizeof(uint64_t) * (v.size() + 1));
        if (!v.empty()) memcpy(r, v.data(), sizeof(uint64_t) * v.size());
        return r;
    };
    out->block_offsets = dup64(block_offsets);
    out->block_hash = dup64(block_hash);
    out->new_used = dup64(new_used);
    out->n_new_used = new_used.size();
    out->frag_seq = (uint32_t*)malloc(4 * (frag_seq.size() + 1));
    out->frag_min_pos = (uint32_t*)malloc(4 * (frag_seq.size() + 1));
    out->frag_ori = (int8_t*)malloc(frag_seq.size() + 1);
    if (!frag_seq.empty()) {
        memcpy(out->frag_seq, frag_seq.data(), 4 * frag_seq.size());
        memcpy(out->frag_min_pos, frag_min.data(), 4 * frag_min.size());
        memcpy(out->frag_ori, frag_ori.data(), frag_ori.size());
    }
}

inline bool has_elem(const uint64_t* v, uint64_t n, uint64_t e) {
    /* SortedVector::has_elem, SortedVector.hpp:108-111 */
    const uint64_t* it = std::lower_bound(v, v + n, e);
    return it != v + n && *it == e;
}

}  // namespace

/* ====================================================================== */
/* ref_faithful: the reference's two-pass algorithm                        */
/* ====================================================================== */
namespace {

/* BloomFilter, BloomFilter.cpp:29-76,166-170 */
struct Bloom {
    std::vector<bool> bits;
    std::vector<size_t> params;

    /* set_members + set_optimal_hashes (:33-35,45-47) + set_hashes (:55-63);
     * srand(make_seed()) becomes srand(seed). */
    void setup(size_t members, double fp, uint32_t seed) {
        bits.assign(bloom_bits(members, fp), false);
        params.assign(bloom_hashes(members, bits.size()), 0);
        std::srand(seed);
        for (size_t i = 0; i < params.size(); i++) {
            params[i] = std::rand();
        }
    }

    /* :65-76 with make_index :166-170 */
    bool test_and_add(hash_t hash) {
        bool result = true;
        hash_t m = bits.size();
        for (size_t i = 0; i < params.size(); i++) {
            size_t index = (hash ^ params[i]) % m;
            if (!bits[index]) {
                result = false;
            }
            bits[index] = true;
        }
        return result;
    }
};

/* estimate_length, AnchorFinder.cpp:81-97 */
size_t estimate_length(const uint64_t* offsets, uint32_t n_seqs,
                       const int32_t* genome_ids) {
    std::map<int32_t, size_t> gtl;
    size_t max_length = 0;
    for (uint32_t s = 0; s < n_seqs; s++) {
        size_t& gl = gtl[genome_ids ? genome_ids[s] : -1];
        gl += offsets[s + 1] - offsets[s];
        max_length = std::max(max_length, gl);
    }
    if (gtl.size() == 1) {
        return max_length;
    }
    return max_length / 2 * 3;
}

/* Runs `task(seq_view, worker_id)` over the sequences the way
 * ThreadGroup/ThreadWorker do (thread_group.cpp:84-94,128-139): W workers pull
 * the next sequence from a shared iterator under a mutex. */
template <typename Task>
void run_tasks(const std::vector<SeqView>& seqs, int workers, Task task) {
    if (workers <= 1) {
        for (size_t i = 0; i < seqs.size(); i++) {
            task(seqs[i], 0);
        }
        return;
    }
    std::mutex mutex;
    size_t next = 0;
    std::vector<std::string> errors(workers);
    std::vector<std::thread> threads;
    for (int w = 0; w < workers; w++) {
        threads.emplace_back([&, w]() {
            try {
                while (true) {
                    size_t i;
                    {
                        std::lock_guard<std::mutex> lock(mutex);
                        if (next >= seqs.size()) {
                            break;
                        }
                        i = next++;
                    }
                    task(seqs[i], w);
                }
            } catch (std::exception& e) {
                errors[w] = e.what();  /* thread_group.cpp:70-82 */
            }
        });
    }
    for (auto& t : threads) {
        t.join();
    }
    for (auto& e : errors) {
        if (!e.empty()) {
            throw std::runtime_error(e);
        }
    }
}

void find_faithful(const oracle_params* p, const char* bases,
                   const uint64_t* offsets, uint32_t n_seqs,
                   const int32_t* genome_ids, const uint64_t* used,
                   uint64_t n_used, oracle_result* out) {
    const int k = p->anchor_size;
    const bool similar = p->anchor_similar != 0;
    int workers = p->workers;
    if (workers < 0) {
        workers = std::max(1u, std::thread::hardware_concurrency());
    }
    /* AnchorFinderOptions ctor :59-77 */
    std::vector<SeqView> seqs = make_seqs(bases, offsets, n_seqs, k,
                                          p->asserts != 0);
    /* BloomTG::initialize_bloom :115-131 */
    size_t length_sum = estimate_length(offsets, n_seqs, genome_ids);
    if (k * 2 < 64) {
        size_t all_anchors = size_t(1) << (k * 2);
        if (all_anchors < length_sum) {
            length_sum = all_anchors;
        }
    }
    Bloom bloom;
    std::unordered_set<hash_t> exact;
    if (!p->exact_bloom) {
        bloom.setup(length_sum, p->anchor_fp, p->bloom_seed);
    }
    /* pass 1: BloomTask :153-196 */
    std::vector<std::vector<hash_t> > worker_hashes(std::max(workers, 1));
    run_tasks(seqs, workers, [&](const SeqView& seq, int w) {
        std::vector<hash_t>& hashes = worker_hashes[w];
        Scanner sc(&seq, k);
        bool prev = false;
        auto test_and_add = [&]() {
            bool hash_found = false;
            if (sc.ns == 0) {
                hash_t hash = std::min(sc.dir, sc.rev);
                if (!has_elem(used, n_used, hash)) {
                    if (p->exact_bloom) {
                        hash_found = !exact.insert(hash).second;
                    } else {
                        hash_found = bloom.test_and_add(hash);
                    }
                    if (hash_found && (!prev || !similar)) {
                        hashes.push_back(hash);
                    }
                }
            }
            prev = hash_found;
        };
        sc.init_state();
        test_and_add();
        size_t n = seq.size - k;
        for (size_t i = 0; i < n; i++) {
            sc.next_hash();
            test_and_add();
        }
    });
    /* ~BloomWorker :146-149 + bloomtg_postprocess :213-218 */
    std::vector<hash_t> H;
    for (auto& wh : worker_hashes) {
        H.insert(H.end(), wh.begin(), wh.end());
        std::vector<hash_t>().swap(wh);
    }
    std::vector<bool>().swap(bloom.bits);
    exact.clear();
    std::sort(H.begin(), H.end());
    H.erase(std::unique(H.begin(), H.end()), H.end());
    /* pass 2: FragmentTask :283-326 */
    std::vector<std::vector<Found> > worker_ffs(std::max(workers, 1));
    run_tasks(seqs, workers, [&](const SeqView& seq, int w) {
        std::vector<Found>& ffs = worker_ffs[w];
        Scanner sc(&seq, k);
        auto test_and_push = [&]() {
            if (sc.ns == 0) {
                hash_t hash = std::min(sc.dir, sc.rev);
                if (has_elem(H.data(), H.size(), hash)) {
                    bool direct = (hash == sc.dir);
                    Found f = {hash, seq.index,
                               uint64_t(sc.pos) + (direct ? 0 : seq.size)};
                    ffs.push_back(f);
                }
            }
        };
        sc.init_state();
        test_and_push();
        size_t n = seq.size - k;
        for (size_t i = 0; i < n; i++) {
            sc.next_hash();
            test_and_push();
        }
    });
    std::vector<Found> ffs;
    for (auto& wf : worker_ffs) {
        ffs.insert(ffs.end(), wf.begin(), wf.end());
        std::vector<Found>().swap(wf);
    }
    build_blocks(ffs, p, offsets, out);
}

}  // namespace

extern "C" int oracle_find_faithful(const oracle_params* p, const char* bases,
                                    const uint64_t* offsets, uint32_t n_seqs,
                                    const int32_t* genome_ids,
                                    const uint64_t* used_sorted,
                                    uint64_t n_used, oracle_result* out) {
    memset(out, 0, sizeof(*out));
    if (p->anchor_size <= 0 || p->anchor_size > MAX_ANCHOR) {
        return 1;  /* add_opt_rule, AnchorFinder.cpp:49-51 */
    }
    try {
        auto t0 = std::chrono::steady_clock::now();
        find_faithful(p, bases, offsets, n_seqs, genome_ids, used_sorted,
                      n_used, out);
        out->seconds = std::chrono::duration<double>(
                           std::chrono::steady_clock::now() - t0).count();
    } catch (std::exception&) {
        return 2;
    }
    return 0;
}

/* ====================================================================== */
/* ref_ideal: exact set semantics of the same algorithm                    */
/* ====================================================================== */
namespace {

struct Window {
    hash_t hash;
    uint64_t rank;  /* scan rank: sequences by size desc, then position */
    uint32_t seq;
    uint32_t pos;
    bool direct;
};

void find_ideal(const oracle_params* p, const char* bases,
                const uint64_t* offsets, uint32_t n_seqs,
                const uint64_t* used, uint64_t n_used, oracle_result* out) {
    const int k = p->anchor_size;
    std::vector<SeqView> seqs = make_seqs(bases, offsets, n_seqs, k, false);
    /* every N-free window of every sequence, hashes
     * computed directly, not rolled: an independent check of reuse_hash */
    std::vector<Window> ws;
    std::vector<uint64_t> scan_base(seqs.size() + 1, 0);
    const hash_t mask = k == 32 ? ~hash_t(0) : ((hash_t(1) << (2 * k)) - 1);
    for (size_t si = 0; si < seqs.size(); si++) {
        const SeqView& seq = seqs[si];
        scan_base[si + 1] = scan_base[si] + seq.size;
        hash_t dir = 0, rev = 0;
        int valid = 0;  /* number of trailing N-free letters */
        for (size_t i = 0; i < seq.size; i++) {
            char c = seq.data[i];
            if (c == 'N') {
                valid = 0;
                dir = rev = 0;
                continue;
            }
            hash_t code = letter_code(c);
            /* dir: first base in the low bits; rev: complement, reversed */
            dir = (dir >> 2) | (code << (2 * (k - 1)));
            rev = ((rev << 2) | (code ^ 1)) & mask;
            valid++;
            if (valid >= k) {
                size_t pos = i + 1 - k;
                hash_t h = std::min(dir, rev);
                Window w = {h, scan_base[si] + pos, seq.index,
                            uint32_t(pos), h == dir};
                ws.push_back(w);
            }
        }
    }
    /* group by hash; inside a group order by scan rank */
    std::sort(ws.begin(), ws.end(), [](const Window& a, const Window& b) {
        if (a.hash != b.hash) return a.hash < b.hash;
        return a.rank < b.rank;
    });
    const bool similar = p->anchor_similar != 0;
    /* found(rank): valid, hash not used, seen at an earlier scan position
     * (BloomTask::test_and_add :170-182 with an exact filter) */
    std::vector<bool> found;
    if (similar) {
        found.assign(scan_base.back() + 1, false);
    }
    std::vector<char> group_ok(ws.size(), 0);  /* per window: group is in H */
    for (size_t i = 0; i < ws.size();) {
        size_t j = i;
        while (j < ws.size() && ws[j].hash == ws[i].hash) {
            j++;
        }
        bool repeated = (j - i >= 2) && !has_elem(used, n_used, ws[i].hash);
        if (repeated) {
            if (similar) {
                for (size_t t = i + 1; t < j; t++) {
                    found[ws[t].rank] = true;
                }
            } else {
                for (size_t t = i; t < j; t++) {
                    group_ok[t] = 1;
                }
            }
        }
        i = j;
    }
    if (similar) {
        /* candidate iff found(p) and not found(p-1); prev_ resets at the
         * start of every sequence (:191) */
        std::vector<bool> seq_start(scan_base.back() + 1, false);
        for (size_t si = 0; si < seqs.size(); si++) {
            seq_start[scan_base[si]] = true;
        }
        for (size_t i = 0; i < ws.size();) {
            size_t j = i;
            bool in_H = false;
            while (j < ws.size() && ws[j].hash == ws[i].hash) {
                uint64_t r = ws[j].rank;
                if (found[r] && (seq_start[r] || !found[r - 1])) {
                    in_H = true;
                }
                j++;
            }
            if (in_H) {
                for (size_t t = i; t < j; t++) {
                    group_ok[t] = 1;
                }
            }
            i = j;
        }
    }
    std::vector<Found> ffs;
    for (size_t i = 0; i < ws.size(); i++) {
        if (group_ok[i]) {
            const Window& w = ws[i];
            uint64_t size = offsets[w.seq + 1] - offsets[w.seq];
            Found f = {w.hash, w.seq, uint64_t(w.pos) + (w.direct ? 0 : size)};
            ffs.push_back(f);
        }
    }
    std::vector<Window>().swap(ws);
    build_blocks(ffs, p, offsets, out);
}

}  // namespace

extern "C" int oracle_find_ideal(const oracle_params* p, const char* bases,
                                 const uint64_t* offsets, uint32_t n_seqs,
                                 const uint64_t* used_sorted, uint64_t n_used,
                                 oracle_result* out) {
    memset(out, 0, sizeof(*out));
    if (p->anchor_size <= 0 || p->anchor_size > MAX_ANCHOR) {
        return 1;
    }
    try {
        auto t0 = std::chrono::steady_clock::now();
        find_ideal(p, bases, offsets, n_seqs, used_sorted, n_used, out);
        out->seconds = std::chrono::duration<double>(
                           std::chrono::steady_clock::now() - t0).count();
    } catch (std::exception&) {
        return 2;
    }
    return 0;
}

extern "C" void oracle_result_free(oracle_result* r) {
    free(r->block_offsets);
    free(r->block_hash);
    free(r->frag_seq);
    free(r->frag_min_pos);
    free(r->frag_ori);
    free(r->new_used);
    memset(r, 0, sizeof(*r));
}

/* ====================================================================== */
/* blockset_hash and the 128-bit digest                                     */
/* ====================================================================== */
namespace {

/* Fragment::id, Fragment.cpp:173-183: name_begin_last with begin/last in
 * fragment direction; a one-letter reverse fragment prints last = -1 */
std::string fragment_id(const char* name, long min_pos, long max_pos, int ori) {
    long a = ori == 1 ? min_pos : max_pos;
    long b = ori == 1 ? max_pos : min_pos;
    if (a == b && ori == -1) {
        b = -1;
    }
    return std::string(name) + "_" + std::to_string(a) + "_" +
           std::to_string(b);
}

/* block_hash, block_hash.cpp:29-54 */
hash_t one_block_hash(const std::vector<std::string>& dir_in,
                      const std::vector<std::string>& inv_in) {
    std::vector<std::string> ids_dir(dir_in), ids_inv(inv_in);
    std::sort(ids_dir.begin(), ids_dir.end());
    std::sort(ids_inv.begin(), ids_inv.end());
    const std::vector<std::string>& ids = (ids_dir < ids_inv) ? ids_dir : ids_inv;
    std::string joint;
    for (size_t i = 0; i < ids.size(); i++) {
        if (i) {
            joint += " ";
        }
        joint += ids[i];
    }
    const size_t LOOP = sizeof(hash_t) * 2;
    joint.resize((joint.size() + LOOP - 1) / LOOP * LOOP, ' ');
    hash_t a = 1;
    for (size_t i = 0; i < joint.size() / LOOP; i++) {
        hash_t v0, v1;
        memcpy(&v0, joint.data() + i * LOOP, 8);
        memcpy(&v1, joint.data() + i * LOOP + 8, 8);
        a *= v0;
        a ^= v1;
    }
    return a;
}

// one fragment's two digest terms (three 64-bit multiplies)
inline void digest_terms(uint64_t t, uint64_t h, uint64_t& a, uint64_t& b) {
    uint64_t u = t * 0x9E3779B97F4A7C15ull + h * 0xC2B2AE3D27D4EB4Full;
    u ^= u >> 32;
    a = u;
    uint64_t v = u * 0xD6E8FEB86659FD93ull;
    b = v ^ (v >> 29);
}

}  // namespace

/* block_hash (block_hash.cpp:29-54) of every block of a flat block list, as
 * MoveUnchanged (MoveUnchanged.cpp:37-49) computes it block by block */
extern "C" void oracle_block_hashes(const char* const* names,
        uint64_t n_blocks, const uint64_t* block_offsets,
        const uint32_t* frag_seq, const uint32_t* frag_min_pos,
        const uint32_t* frag_max_pos, const int8_t* frag_ori,
        uint64_t* hashes) {
    for (uint64_t b = 0; b < n_blocks; b++) {
        std::vector<std::string> dir, inv;
        for (uint64_t f = block_offsets[b]; f < block_offsets[b + 1]; f++) {
            const char* name = names[frag_seq[f]];
            dir.push_back(fragment_id(name, frag_min_pos[f], frag_max_pos[f],
                                      frag_ori[f]));
            inv.push_back(fragment_id(name, frag_min_pos[f], frag_max_pos[f],
                                      -frag_ori[f]));
        }
        hashes[b] = one_block_hash(dir, inv);
    }
}

/* blockset_hash, block_hash.cpp:125-130 (+ HashGroup::create_task_impl
 * :101-113: blocks with <= 1 fragment are skipped) */
extern "C" uint64_t oracle_blockset_hash(const char* const* names,
        uint64_t n_blocks, const uint64_t* block_offsets,
        const uint32_t* frag_seq, const uint32_t* frag_min_pos,
        const uint32_t* frag_max_pos, const int8_t* frag_ori) {
    hash_t result = 0;
    for (uint64_t b = 0; b < n_blocks; b++) {
        uint64_t lo = block_offsets[b], hi = block_offsets[b + 1];
        if (hi - lo <= 1) {
            continue;
        }
        std::vector<std::string> dir, inv;
        for (uint64_t f = lo; f < hi; f++) {
            const char* name = names[frag_seq[f]];
            dir.push_back(fragment_id(name, frag_min_pos[f], frag_max_pos[f],
                                      frag_ori[f]));
            inv.push_back(fragment_id(name, frag_min_pos[f], frag_max_pos[f],
                                      -frag_ori[f]));
        }
        result ^= one_block_hash(dir, inv);
    }
    return result;
}

/* Same definition as npge_anchor_result.digest (include/npge_anchor.h):
 * per fragment t = (seq << 33) | (min_pos << 1) | (ori == +1), h = block hash;
 * u = t * 0x9E3779B97F4A7C15 + h * 0xC2B2AE3D27D4EB4F, u ^= u >> 32,
 * v = u * 0xD6E8FEB86659FD93, v ^= v >> 29;
 * digest[0] = sum of u, digest[1] = sum of v, all mod 2^64. */
extern "C" void oracle_digest(const oracle_result* r, uint64_t digest[2]) {
    uint64_t d0 = 0, d1 = 0;
    for (uint64_t b = 0; b < r->n_blocks; b++) {
        uint64_t h = r->block_hash[b];
        for (uint64_t f = r->block_offsets[b]; f < r->block_offsets[b + 1]; f++) {
            uint64_t t = (uint64_t(r->frag_seq[f]) << 33) |
                         (uint64_t(r->frag_min_pos[f]) << 1) |
                         (r->frag_ori[f] == 1 ? 1 : 0);
            uint64_t x, y;
            digest_terms(t, h, x, y);
            d0 += x;
            d1 += y;
        }
    }
    digest[0] = d0;
    digest[1] = d1;
}

/* ====================================================================== */
/* ref_ideal in hash-range slices: the same exact semantics as find_ideal  */
/* with O(N / S) memory, for the configurations whose window list does not */
/* fit (C3-C5 of BASELINE.json; 10^9 windows at C5).                       */
/* ====================================================================== */
/* Slice s holds the canonical hashes whose top `slice_bits` bits (of the
 * 2k significant ones) equal s, so slices ascend with the hash and a
 * group never spans two of them.  Every slice re-scans the sequences
 * (rolling the two hashes as find_ideal does) and keeps only its own
 * windows; the slices are independent, so they run on `threads` threads.
 *   parity mode      one sweep: groups of >= 2 unused windows -> blocks
 *   anchor-similar   sweep 1 sets found(rank) (every member of such a
 *                    group but the first in scan order), sweep 2 selects
 *                    the groups holding a candidate (found(p) and not
 *                    found(p-1), BloomTask::test_and_add :170-182)
 * With a finite max-anchor-fragments the selected records are gathered in
 * ascending slice order until the cut is covered and then go through the
 * same build_blocks as find_ideal (sort by (hash, seq, pos'), cut, group). */
namespace {

struct SliceRec {
    hash_t hash;
    uint64_t t;  /* scan index of the sequence << 33 | pos << 1 | direct */
    bool operator<(const SliceRec& o) const {
        return hash != o.hash ? hash < o.hash : t < o.t;
    }
};

struct SliceScan {
    std::vector<SeqView> seqs;          /* scan order */
    std::vector<uint64_t> scan_base;    /* rank of position 0 of every sequence */
    int k;
    int shift;                          /* hash >> shift = slice */
    hash_t mask;
};

/* every N-free window of slice `slice`, sorted by (hash, scan rank) */
void slice_windows(const SliceScan& sc, uint64_t slice, std::vector<SliceRec>& out) {
    out.clear();
    const int k = sc.k;
    for (size_t si = 0; si < sc.seqs.size(); si++) {
        const SeqView& seq = sc.seqs[si];
        hash_t dir = 0, rev = 0;
        int valid = 0;
        for (size_t i = 0; i < seq.size; i++) {
            const char c = seq.data[i];
            if (c == 'N') {
                valid = 0;
                dir = rev = 0;
                continue;
            }
            const hash_t code = letter_code(c);
            dir = (dir >> 2) | (code << (2 * (k - 1)));
            rev = ((rev << 2) | (code ^ 1)) & sc.mask;
            valid++;
            if (valid >= k) {
                const hash_t h = std::min(dir, rev);
                if ((sc.shift >= 64 ? 0 : (h >> sc.shift)) == slice) {
                    const uint64_t pos = i + 1 - k;
                    SliceRec r = {h, (uint64_t(si) << 33) | (pos << 1) | (h == dir ? 1u : 0u)};
                    out.push_back(r);
                }
            }
        }
    }
    std::sort(out.begin(), out.end());
}

inline uint64_t rank_of(const SliceScan& sc, uint64_t t) {
    return sc.scan_base[t >> 33] + ((t >> 1) & 0xffffffffull);
}

template <typename Fn>
void parallel_slices(uint64_t n_slices, int threads, Fn fn) {
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    std::vector<std::string> errors(threads);
    for (int w = 0; w < threads; w++) {
        pool.emplace_back([&, w]() {
            try {
                while (true) {
                    const uint64_t s = next.fetch_add(1);
                    if (s >= n_slices) break;
                    fn(s, w);
                }
            } catch (std::exception& e) {
                errors[w] = e.what();
            }
        });
    }
    for (auto& t : pool) t.join();
    for (auto& e : errors) {
        if (!e.empty()) throw std::runtime_error(e);
    }
}

}  // namespace

extern "C" int oracle_find_ideal_sliced(const oracle_params* p, const char* bases,
                                        const uint64_t* offsets, uint32_t n_seqs,
                                        const uint64_t* used, uint64_t n_used,
                                        int32_t slice_bits, int32_t threads,
                                        oracle_summary* out) {
    memset(out, 0, sizeof(*out));
    const int k = p->anchor_size;
    if (k <= 0 || k > MAX_ANCHOR || slice_bits < 0 || slice_bits > 16) {
        return 1;
    }
    try {
        auto t0 = std::chrono::steady_clock::now();
        if (slice_bits > 2 * k) slice_bits = 2 * k;
        if (threads < 1) threads = std::max(1u, std::thread::hardware_concurrency());
        SliceScan sc;
        sc.seqs = make_seqs(bases, offsets, n_seqs, k, false);
        sc.scan_base.assign(sc.seqs.size() + 1, 0);
        for (size_t si = 0; si < sc.seqs.size(); si++) {
            sc.scan_base[si + 1] = sc.scan_base[si] + sc.seqs[si].size;
        }
        sc.k = k;
        sc.shift = 2 * k - slice_bits;
        sc.mask = k == 32 ? ~hash_t(0) : ((hash_t(1) << (2 * k)) - 1);
        const uint64_t n_slices = uint64_t(1) << slice_bits;
        const bool similar = p->anchor_similar != 0;
        const bool limited = p->max_anchor_fragments >= 0 &&
                             uint64_t(p->max_anchor_fragments) < (uint64_t(1) << 62);
        std::vector<std::atomic<uint64_t> > found;
        std::atomic<uint64_t> n_windows(0);
        if (similar) {
            std::vector<std::atomic<uint64_t> > bits(sc.scan_base.back() / 64 + 2);
            found.swap(bits);
            for (auto& w : found) w.store(0, std::memory_order_relaxed);
            parallel_slices(n_slices, threads, [&](uint64_t s, int) {
                std::vector<SliceRec> ws;
                slice_windows(sc, s, ws);
                for (size_t i = 0; i < ws.size();) {
                    size_t j = i;
                    while (j < ws.size() && ws[j].hash == ws[i].hash) j++;
                    if (j - i >= 2 && !has_elem(used, n_used, ws[i].hash)) {
                        for (size_t t = i + 1; t < j; t++) {
                            const uint64_t r = rank_of(sc, ws[t].t);
                            found[r >> 6].fetch_or(uint64_t(1) << (r & 63),
                                                   std::memory_order_relaxed);
                        }
                    }
                    i = j;
                }
            });
        }
        auto is_found = [&](uint64_t r) {
            return (found[r >> 6].load(std::memory_order_relaxed) >> (r & 63)) & 1;
        };
        /* selected records of one slice as FoundFragments, in list order */
        auto select = [&](uint64_t s, std::vector<Found>& ffs) {
            std::vector<SliceRec> ws;
            slice_windows(sc, s, ws);
            n_windows.fetch_add(ws.size());
            ffs.clear();
            for (size_t i = 0; i < ws.size();) {
                size_t j = i;
                while (j < ws.size() && ws[j].hash == ws[i].hash) j++;
                bool ok = j - i >= 2 && !has_elem(used, n_used, ws[i].hash);
                if (ok && similar) {
                    ok = false;
                    for (size_t t = i; t < j && !ok; t++) {
                        const uint64_t r = rank_of(sc, ws[t].t);
                        const uint64_t pos = (ws[t].t >> 1) & 0xffffffffull;
                        ok = is_found(r) && (pos == 0 || !is_found(r - 1));
                    }
                }
                if (ok) {
                    for (size_t t = i; t < j; t++) {
                        const SeqView& seq = sc.seqs[ws[t].t >> 33];
                        const uint64_t pos = (ws[t].t >> 1) & 0xffffffffull;
                        Found f = {ws[t].hash, seq.index,
                                   pos + ((ws[t].t & 1) ? 0 : seq.size)};
                        ffs.push_back(f);
                    }
                }
                i = j;
            }
        };
        if (!limited) {
            struct Acc { uint64_t nf = 0, nb = 0, d0 = 0, d1 = 0; };
            std::vector<Acc> acc(threads);
            parallel_slices(n_slices, threads, [&](uint64_t s, int w) {
                std::vector<Found> ffs;
                select(s, ffs);
                Acc& a = acc[w];
                for (size_t i = 0; i < ffs.size(); i++) {
                    const Found& f = ffs[i];
                    const uint64_t size = offsets[f.seq + 1] - offsets[f.seq];
                    const bool direct = f.pos < size;
                    const uint64_t t = (uint64_t(f.seq) << 33) |
                                       ((direct ? f.pos : f.pos - size) << 1) |
                                       (direct ? 1 : 0);
                    uint64_t x, y;
                    digest_terms(t, f.hash, x, y);
                    a.d0 += x;
                    a.d1 += y;
                    a.nf++;
                    if (i == 0 || ffs[i - 1].hash != f.hash) a.nb++;
                }
            });
            for (const Acc& a : acc) {
                out->n_fragments += a.nf;
                out->n_blocks += a.nb;
                out->digest[0] += a.d0;
                out->digest[1] += a.d1;
            }
            out->total_selected = out->n_fragments;
            out->n_new_used = out->n_blocks ? out->n_blocks - 1 : 0;
            out->n_windows = n_windows.load();
        } else {
            /* batches of `threads` slices in ascending order until the cut is covered */
            const uint64_t M = uint64_t(p->max_anchor_fragments);
            std::vector<Found> all;
            bool complete = true;
            for (uint64_t s0 = 0; s0 < n_slices; s0 += threads) {
                const uint64_t n = std::min<uint64_t>(threads, n_slices - s0);
                std::vector<std::vector<Found> > part(n);
                parallel_slices(n, int(n), [&](uint64_t s, int) { select(s0 + s, part[s]); });
                for (auto& v : part) all.insert(all.end(), v.begin(), v.end());
                if (all.size() > M) {
                    complete = s0 + n >= n_slices;
                    break;
                }
            }
            out->total_selected = complete ? all.size() : 0;  /* 0: not counted */
            out->n_windows = complete ? n_windows.load() : 0;
            oracle_result r;
            memset(&r, 0, sizeof(r));
            build_blocks(all, p, offsets, &r);
            out->n_fragments = r.n_fragments;
            out->n_blocks = r.n_blocks;
            out->n_new_used = r.n_new_used;
            oracle_digest(&r, out->digest);
            oracle_result_free(&r);
        }
        out->seconds = std::chrono::duration<double>(
                           std::chrono::steady_clock::now() - t0).count();
    } catch (std::exception&) {
        return 2;
    }
    return 0;
}

/* ====================================================================== */
/* FragmentFinder: src/algo/FragmentFinder.cpp:42-171   */
/* ====================================================================== */
/* One worker (the reference's result order depends on thread timing with
 * more): sequences by size descending (make_seqs), positions ascending, the
 * forward match before the reverse one at a position (test(), :110-117);
 * the scan of a sequence stops once more than max_matches were collected
 * (:125-127) and the list is cut to max_matches (:155-164).
 * Output: one (seq, min_pos, ori) per match, arrays malloc'ed here.        */
extern "C" int oracle_fragment_find(const char* pattern, uint32_t pattern_len,
                                    int64_t max_matches, const char* bases,
                                    const uint64_t* offsets, uint32_t n_seqs,
                                    uint64_t* n_out, uint32_t** seq_out,
                                    uint32_t** pos_out, int8_t** ori_out) {
    if (pattern_len == 0 || max_matches < 1) {
        return 1;  /* the option checks of the processor (:18-28, :37) */
    }
    try {
        const std::string p(pattern, pattern_len);  /* already to_atgcn-clean */
        const int L = int(pattern_len);
        const hash_t ph = hash_of([&](int j) { return p[j]; }, L, 1);
        const bool has_n = p.find('N') != std::string::npos;
        std::vector<SeqView> seqs = make_seqs(bases, offsets, n_seqs, L, true);
        std::vector<uint32_t> fs, fp;
        std::vector<int8_t> fo;
        for (size_t si = 0; si < seqs.size(); si++) {
            const SeqView& seq = seqs[si];
            Scanner sc(&seq, L);
            /* add_match, :97-108: trust the hash only for short N-free input */
            auto add_match = [&](int ori) {
                bool ok = !has_n && sc.ns == 0 && L <= MAX_ANCHOR;
                if (!ok) {  /* seq_->substr(begin, anchor_, ori) == p_ */
                    ok = true;
                    for (int j = 0; j < L && ok; j++) {
                        char c = ori == 1 ? seq.data[sc.pos + j]
                                          : complement_char(seq.data[sc.pos + L - 1 - j]);
                        ok = c == p[j];
                    }
                }
                if (ok) {
                    fs.push_back(seq.index);
                    fp.push_back(uint32_t(sc.pos));
                    fo.push_back(int8_t(ori));
                }
            };
            auto test = [&]() {
                if (sc.dir == ph) add_match(1);
                if (sc.rev == ph) add_match(-1);
            };
            sc.init_state();
            test();
            const long n = long(seq.size) - L;
            for (long i = 0; i < n; i++) {
                sc.next_hash();
                test();
                if (int64_t(fs.size()) > max_matches) break;
            }
        }
        uint64_t keep = std::min<uint64_t>(fs.size(), uint64_t(max_matches));
        *n_out = keep;
        *seq_out = (uint32_t*)malloc(4 * (keep + 1));
        *pos_out = (uint32_t*)malloc(4 * (keep + 1));
        *ori_out = (int8_t*)malloc(keep + 1);
        memcpy(*seq_out, fs.data(), 4 * keep);
        memcpy(*pos_out, fp.data(), 4 * keep);
        memcpy(*ori_out, fo.data(), keep);
        return 0;
    } catch (std::exception&) {
        return 2;
    }
}

extern "C" void oracle_free(void* p) {
    free(p);
}
