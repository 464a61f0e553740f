// gpu_anchor_finder.cpp -- see gpu_anchor_finder.hpp.
//
// Mirrors, call for call, what the reference's processor does around its two
// passes (src/algo/AnchorFinder.cpp):
//   ctor        :37-53    options, rules, declare_bs
//   run_impl    :393-406  read options, collect sequences, find, build blocks
//   postprocess :356-391  one Block per hash group, Fragment(seq, min, max, ori),
//                         check_block, used_hashes_ update
// Linearising the sequences and rebuilding the object graph are the only host
// work; everything between is one call into the C ABI.
#include "gpu_anchor_finder.hpp"

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <exception>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/npge_anchor.h"

#ifdef NPGE_B200_REAL_HEADERS
#include <boost/bind.hpp>

#include "Sequence.hpp"
#include "Fragment.hpp"
#include "Block.hpp"
#include "BlockSet.hpp"
#include "Exception.hpp"
#endif

namespace npge {
namespace gpu {

struct AnchorFinder::Impl {
    npge_anchor_ctx* ctx = nullptr;      // one device
    npge_anchor_multi* multi = nullptr;  // several devices
    std::vector<int> devices = std::vector<int>(1, 0);
    std::vector<uint64_t> used_hashes;  // sorted; AnchorFinderImpl::used_hashes_ (:30-32)

    // threads of the object build: the processor's `workers` option (-1 = all cores,
    // Processor.cpp workers semantics)
    static int build_threads(int workers) {
        if (workers < 0) {
            const unsigned hc = std::thread::hardware_concurrency();
            return hc ? int(hc) : 1;
        }
        return workers < 1 ? 1 : workers;
    }

    // used_hashes_ stays sorted and unique (SortedVector::extend + sort, :385-389);
    // the library returns new_used ascending, so a merge does it
    void extend_used(const uint64_t* more, uint64_t n) {
        if (!n) return;
        const size_t old = used_hashes.size();
        used_hashes.insert(used_hashes.end(), more, more + n);
        if (!std::is_sorted(used_hashes.begin() + old, used_hashes.end())) {
            std::sort(used_hashes.begin() + old, used_hashes.end());
        }
        std::inplace_merge(used_hashes.begin(), used_hashes.begin() + old, used_hashes.end());
        used_hashes.erase(std::unique(used_hashes.begin(), used_hashes.end()), used_hashes.end());
    }
    // the genome resident on the device(s): AnchorLoop re-runs this instance on the same
    // sequences (src/algo/MetaProcessor.cpp:32-38, src/algo/Pipe.cpp:41-61)
    std::vector<const Sequence*> up_seqs;
    std::vector<uint64_t> up_sizes;
    uint64_t up_probe = 0;
    int up_k = 0;
    bool resident = false, reused = false;
    double ms[4] = {0, 0, 0, 0};  // [3]: the device search alone (part of [1])
    void drop() {
        if (ctx) npge_anchor_destroy(ctx);
        if (multi) npge_anchor_multi_destroy(multi);
        ctx = nullptr;
        multi = nullptr;
        resident = false;
    }
};

AnchorFinder::AnchorFinder() : impl_(new Impl) {
#ifdef NPGE_B200_REAL_HEADERS
    add_gopt("anchor-size", "anchor size", "ANCHOR_SIZE");
    add_gopt("anchor-fp",
             "Probability of false positive in Bloom filter "
             "(accepted for compatibility: the GPU method is exact)", "ANCHOR_FP");
    add_opt("anchor-similar", "If neighbour anchors are skipped", true);
    add_gopt("max-anchor-fragments", "Maximum number of anchors fragments to return",
             "MAX_ANCHOR_FRAGMENTS");
    add_opt_rule("anchor-size > 0");
    add_opt_rule("anchor-size <= 32");
#else
    // defaults of the globals: CMakeLists.txt:41-46 of the reference
    add_gopt("anchor-size", "anchor size", "ANCHOR_SIZE", 20);
    add_gopt("anchor-fp",
             "Probability of false positive in Bloom filter "
             "(accepted for compatibility: the GPU method is exact)", "ANCHOR_FP", 0.1);
    add_opt("anchor-similar", "If neighbour anchors are skipped", true);
    add_gopt("max-anchor-fragments", "Maximum number of anchors fragments to return",
             "MAX_ANCHOR_FRAGMENTS", 100000);
    add_opt_rule([this](std::string& m) {
        m = "anchor-size > 0";
        return opt_value("anchor-size").as<int>() > 0;
    });
    add_opt_rule([this](std::string& m) {
        m = "anchor-size <= 32";
        return opt_value("anchor-size").as<int>() <= 32;
    });
#endif
    declare_bs("target", "Blockset to search anchors in");
}

AnchorFinder::~AnchorFinder() {
    impl_->drop();
    delete impl_;
}

void AnchorFinder::set_device(int device) {
    set_devices(std::vector<int>(1, device));
}

void AnchorFinder::set_devices(const std::vector<int>& devices) {
    if (devices.empty() || devices.size() > 8) throw Exception("AnchorFinder (GPU): 1..8 devices");
    impl_->drop();
    impl_->devices = devices;
}

bool AnchorFinder::last_run_reused_upload() const {
    return impl_->reused;
}

const double* AnchorFinder::last_run_ms() const {
    return impl_->ms;
}

namespace {

void check(int rc, const char* what) {
    if (rc != NPGE_ANCHOR_OK) {
        throw Exception(std::string("AnchorFinder (GPU): ") + what + ": " +
                        npge_anchor_last_error());
    }
}

// The letters of the anchors of one block must agree up to the strand (check_block,
// AnchorFinder.cpp:342-354: letter by letter through raw_at, no strings): a cheap
// end-to-end self-check of the device result against the sequences it came from.
char letter_at(const Fragment* f, size_t pos) {  // Fragment::raw_at
    const Sequence* seq = f->seq();
    return f->ori() == 1 ? seq->char_at(f->min_pos() + pos)
                         : complement(seq->char_at(f->max_pos() - pos));
}

void check_block(const Block* block, size_t anchor) {
    if (block->size() < 2) throw Exception("AnchorFinder (GPU): block of one fragment");
    const Fragment* first = block->front();
    if (first->length() != anchor) throw Exception("AnchorFinder (GPU): wrong anchor length");
    for (size_t pos = 0; pos < anchor; pos++) {
        const char c = letter_at(first, pos);
        for (Block::const_iterator it = block->begin(); it != block->end(); ++it) {
            if (letter_at(*it, pos) != c) {
                throw Exception("AnchorFinder (GPU): fragments of a block differ: " +
                                first->str() + " vs " + (*it)->str());
            }
        }
    }
}

// One Block per hash group, one Fragment per occurrence (postprocess,
// AnchorFinder.cpp:356-391), built and self-checked by `threads` threads over
// contiguous block ranges (the objects of different blocks share nothing but the
// read-only sequences); the caller inserts them into the block set in order.
// frag(f, &seq, &min_pos, &ori) decodes fragment f of the result.
template <typename Frag>
void build_blocks(const uint64_t* block_offsets, uint64_t n_blocks, Frag frag,
                  const std::vector<SequencePtr>& seqs, size_t k, int threads,
                  std::vector<Block*>& built) {
    built.assign(n_blocks, static_cast<Block*>(0));
    auto range = [&](uint64_t b0, uint64_t b1) {
        for (uint64_t b = b0; b < b1; b++) {
            Block* block = new Block;
            built[b] = block;  // freed by the caller if anything below throws
            for (uint64_t f = block_offsets[b]; f < block_offsets[b + 1]; f++) {
                uint32_t seq, min_pos;
                int ori;
                frag(f, &seq, &min_pos, &ori);
                block->insert(new Fragment(seqs[seq].get(), min_pos, min_pos + k - 1, ori));
            }
            check_block(block, k);
        }
    };
    const uint64_t kMinPerThread = 1024;
    uint64_t T = threads < 1 ? 1 : uint64_t(threads);
    T = std::min<uint64_t>(T, std::max<uint64_t>(1, n_blocks / kMinPerThread));
    std::exception_ptr first_error;
    if (T == 1) {
        try {
            range(0, n_blocks);
        } catch (...) {
            first_error = std::current_exception();
        }
    } else {
        std::mutex mu;
        std::vector<std::thread> pool;
        for (uint64_t t = 0; t < T; t++) {
            pool.emplace_back([&, t] {
                try {
                    range(n_blocks * t / T, n_blocks * (t + 1) / T);
                } catch (...) {
                    std::lock_guard<std::mutex> lock(mu);
                    if (!first_error) first_error = std::current_exception();
                }
            });
        }
        for (size_t t = 0; t < pool.size(); t++) pool[t].join();
    }
    if (first_error) {
        for (uint64_t b = 0; b < n_blocks; b++) delete built[b];  // a Block owns its fragments
        built.clear();
        std::rethrow_exception(first_error);
    }
}

// A cheap fingerprint of the sequences' letters (the ends and 256 evenly spaced
// positions of each): the resident genome is reused when the same Sequence objects,
// with the same sizes and the same fingerprint, come again.
uint64_t probe_sequences(const std::vector<SequencePtr>& seqs) {
    uint64_t h = 0xcbf29ce484222325ull;
    auto mix = [&](uint64_t v) {
        h ^= v;
        h *= 0x100000001b3ull;
    };
    for (size_t i = 0; i < seqs.size(); i++) {
        const Sequence& s = *seqs[i];
        const size_t n = s.size();
        mix(n);
        const size_t edge = n < 32 ? n : 32;
        for (size_t j = 0; j < edge; j++) mix(uint8_t(s.char_at(j)));
        for (size_t j = 0; j < edge; j++) mix(uint8_t(s.char_at(n - 1 - j)));
        for (size_t j = 0; j < 256 && n > 64; j++) mix(uint8_t(s.char_at((n - 1) / 256 * j)));
    }
    return h;
}

}  // namespace

static double wall_ms() {
    return std::chrono::duration<double, std::milli>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
}

void AnchorFinder::run_impl() const {
    Impl& st = *impl_;
    const double t_start = wall_ms();
    npge_anchor_params params;
    params.anchor_size = opt_value("anchor-size").as<int>();
    params.anchor_similar = opt_value("anchor-similar").as<bool>() ? 1 : 0;
    // (a negative value never truncates: size_t comparison, AnchorFinder.cpp:361)
    params.max_anchor_fragments = opt_value("max-anchor-fragments").as<int>();
    BlockSet& bs = *block_set();
    // the caller index of a sequence is its index in `seqs` and comes back in frag_seq
    std::vector<SequencePtr> seqs = bs.seqs();
    const bool many = st.devices.size() > 1;
    if (many && !st.multi) {
        check(npge_anchor_multi_create(&st.multi, st.devices.data(), int(st.devices.size())),
              "no CUDA device");
    }
    if (!many && !st.ctx) check(npge_anchor_create(&st.ctx, st.devices[0]), "no CUDA device");
    // is the genome of the previous run still the one on the device(s)?
    std::vector<const Sequence*> ptrs(seqs.size());
    std::vector<uint64_t> sizes(seqs.size());
    for (size_t i = 0; i < seqs.size(); i++) {
        ptrs[i] = seqs[i].get();
        sizes[i] = seqs[i]->size();
    }
    const uint64_t probe = probe_sequences(seqs);
    st.reused = st.resident && st.up_k == params.anchor_size && st.up_seqs == ptrs &&
                st.up_sizes == sizes && st.up_probe == probe;
    if (!st.reused) {
        // one contiguous ASCII buffer + offsets
        std::vector<uint64_t> offsets(seqs.size() + 1, 0);
        for (size_t i = 0; i < seqs.size(); i++) offsets[i + 1] = offsets[i] + sizes[i];
        std::string bases;
        bases.reserve(offsets.back());
        for (size_t i = 0; i < seqs.size(); i++) bases += seqs[i]->contents();
        st.resident = false;
        const uint8_t* b = reinterpret_cast<const uint8_t*>(bases.data());
        if (many) {
            check(npge_anchor_multi_upload(st.multi, params.anchor_size, b, offsets.data(),
                                           uint32_t(seqs.size())), "npge_anchor_multi_upload");
        } else {
            check(npge_anchor_upload(st.ctx, params.anchor_size, b, offsets.data(),
                                     uint32_t(seqs.size())), "npge_anchor_upload");
        }
        st.up_seqs = ptrs;
        st.up_sizes = sizes;
        st.up_probe = probe;
        st.up_k = params.anchor_size;
        st.resident = true;
    }
    const size_t k = size_t(params.anchor_size);
    const double t_uploaded = wall_ms();
    st.ms[0] = t_uploaded - t_start;
    // the compact result (4 bytes per fragment over PCIe); every fragment is decoded
    // while its Fragment object is built
    npge_anchor_compact_result c;
    if (!many) {
        check(npge_anchor_set_result_layout(st.ctx, NPGE_ANCHOR_LAYOUT_COMPACT),
              "npge_anchor_set_result_layout");
        check(npge_anchor_find_device(st.ctx, &params, st.used_hashes.data(),
                                      st.used_hashes.size(), nullptr),
              "npge_anchor_find_device");
        st.ms[3] = wall_ms() - t_uploaded;
        check(npge_anchor_fetch_compact(st.ctx, &c), "npge_anchor_fetch_compact");
    } else {
        check(npge_anchor_multi_set_result_layout(st.multi, NPGE_ANCHOR_LAYOUT_COMPACT),
              "npge_anchor_multi_set_result_layout");
        check(npge_anchor_multi_find_device(st.multi, &params, st.used_hashes.data(),
                                            st.used_hashes.size(), nullptr),
              "npge_anchor_multi_find_device");
        st.ms[3] = wall_ms() - t_uploaded;
        check(npge_anchor_multi_fetch_compact(st.multi, &c), "npge_anchor_multi_fetch_compact");
    }
    const double t_found = wall_ms();
    st.ms[1] = t_found - t_uploaded;
    try {
        std::vector<Block*> built;
        build_blocks(c.block_offsets, c.n_blocks,
                     [&](uint64_t f, uint32_t* seq, uint32_t* min_pos, int* ori) {
                         npge_decode_compact(&c, f, seq, min_pos, ori);
                     }, seqs, k, st.build_threads(workers()), built);
        for (size_t b = 0; b < built.size(); b++) bs.insert(built[b]);  // owned by the blockset
        st.extend_used(c.new_used, c.n_new_used);
    } catch (...) {
        if (many) npge_anchor_multi_compact_result_free(&c); else npge_anchor_compact_result_free(&c);
        throw;
    }
    if (many) npge_anchor_multi_compact_result_free(&c); else npge_anchor_compact_result_free(&c);
    st.ms[2] = wall_ms() - t_found;
}

const char* AnchorFinder::name_impl() const {
    return "Find anchors";  // AnchorFinder.cpp:408-410
}

// ---- FragmentFinder (src/algo/FragmentFinder.cpp) ---------------------------------
struct FragmentFinder::Impl {
    npge_anchor_ctx* ctx = nullptr;
    int device = 0;
};

#ifdef NPGE_B200_REAL_HEADERS
// the option check of the reference's processor, FragmentFinder.cpp:18-28
static bool pattern_length(const Processor* f, std::string& message) {
    std::string p = f->opt_value("pattern").as<std::string>();
    Sequence::to_atgcn(p);
    if (p.empty()) {
        message = "'pattern' should not be empty";
        return false;
    }
    return true;
}
#endif

FragmentFinder::FragmentFinder() : impl_(new Impl) {
    add_opt("pattern", "Sequence searched for", std::string(""));
#ifdef NPGE_B200_REAL_HEADERS
    add_opt_check(boost::bind(pattern_length, this, _1));  // as FragmentFinder.cpp:18-35
    add_gopt("max-matches", "Maximum number of matches", "MAX_MATCHES");
    add_opt_rule("max-matches >= 1");
#else
    add_opt_rule([this](std::string& m) {  // pattern_length, FragmentFinder.cpp:18-28
        std::string p = opt_value("pattern").as<std::string>();
        Sequence::to_atgcn(p);
        m = "'pattern' should not be empty";
        return !p.empty();
    });
    add_gopt("max-matches", "Maximum number of matches", "MAX_MATCHES", 1000);
    add_opt_rule([this](std::string& m) {
        m = "max-matches >= 1";
        return opt_value("max-matches").as<int>() >= 1;
    });
#endif
    declare_bs("target", "search in sequences, add fragments");
}

FragmentFinder::~FragmentFinder() {
    if (impl_->ctx) npge_anchor_destroy(impl_->ctx);
    delete impl_;
}

void FragmentFinder::set_device(int device) {
    if (impl_->ctx) {
        npge_anchor_destroy(impl_->ctx);
        impl_->ctx = nullptr;
    }
    impl_->device = device;
}

void FragmentFinder::run_impl() const {
    Impl& st = *impl_;
    std::string pattern = opt_value("pattern").as<std::string>();
    Sequence::to_atgcn(pattern);  // FinderTG, :52-53
    const int max_matches = opt_value("max-matches").as<int>();
    BlockSet& bs = *block_set();
    std::vector<SequencePtr> seqs = bs.seqs();
    std::vector<uint64_t> offsets(seqs.size() + 1, 0);
    for (size_t i = 0; i < seqs.size(); i++) offsets[i + 1] = offsets[i] + seqs[i]->size();
    std::string bases;
    bases.reserve(offsets.back());
    for (size_t i = 0; i < seqs.size(); i++) bases += seqs[i]->contents();
    if (!st.ctx) check(npge_anchor_create(&st.ctx, st.device), "no CUDA device");
    npge_fragment_result r;
    check(npge_fragment_find(st.ctx, reinterpret_cast<const uint8_t*>(pattern.data()),
                             uint32_t(pattern.size()), max_matches,
                             reinterpret_cast<const uint8_t*>(bases.data()), offsets.data(),
                             uint32_t(seqs.size()), &r),
          "npge_fragment_find");
    try {
        for (uint64_t f = 0; f < r.n_matches; f++) {  // :155-164: one block per match
            Block* block = new Block;
            const size_t min_pos = r.frag_min_pos[f];
            block->insert(new Fragment(seqs[r.frag_seq[f]].get(), min_pos,
                                       min_pos + pattern.size() - 1, r.frag_ori[f]));
            bs.insert(block);
        }
    } catch (...) {
        npge_fragment_result_free(&r);
        throw;
    }
    npge_fragment_result_free(&r);
}

const char* FragmentFinder::name_impl() const {
    return "Locate fragment by its sequence";  // FragmentFinder.cpp:168-170
}

// ---- MoveUnchanged (src/algo/MoveUnchanged.cpp) -----------------------------------
struct MoveUnchanged::Impl {
    npge_anchor_ctx* ctx = nullptr;
    int device = 0;
    std::vector<uint64_t> hashes;  // sorted, unique: MoveUnchanged::hashes_
    std::vector<uint64_t> last;
};

MoveUnchanged::MoveUnchanged() : impl_(new Impl) {
    declare_bs("other", "Where unchanged blocks are looked for");
    declare_bs("target", "Where unchanged blocks are moved to");
}

MoveUnchanged::~MoveUnchanged() {
    if (impl_->ctx) npge_anchor_destroy(impl_->ctx);
    delete impl_;
}

void MoveUnchanged::set_device(int device) {
    if (impl_->ctx) {
        npge_anchor_destroy(impl_->ctx);
        impl_->ctx = nullptr;
    }
    impl_->device = device;
}

void MoveUnchanged::clear() {
    impl_->hashes.clear();
}

const std::vector<uint64_t>& MoveUnchanged::last_hashes() const {
    return impl_->last;
}

void MoveUnchanged::run_impl() const {
    Impl& st = *impl_;
    BlockSet& t = *block_set();
    BlockSet& o = *other();
    // flat copy of `other`: fragments as (sequence index, min, max, ori)
    std::vector<Block*> blocks(o.begin(), o.end());
    std::vector<Sequence*> seq_of;
    std::vector<const char*> names;
    std::vector<uint64_t> offsets(1, 0);
    std::vector<uint32_t> frag_seq, min_pos, max_pos;
    std::vector<int8_t> ori;
    for (size_t b = 0; b < blocks.size(); b++) {
        for (Block::const_iterator it = blocks[b]->begin(); it != blocks[b]->end(); ++it) {
            const Fragment* f = *it;
            size_t s = 0;
            while (s < seq_of.size() && seq_of[s] != f->seq()) s++;  // few sequences
            if (s == seq_of.size()) seq_of.push_back(f->seq());
            frag_seq.push_back(uint32_t(s));
            min_pos.push_back(uint32_t(f->min_pos()));
            max_pos.push_back(uint32_t(f->max_pos()));
            ori.push_back(int8_t(f->ori()));
        }
        offsets.push_back(frag_seq.size());
    }
    for (size_t s = 0; s < seq_of.size(); s++) names.push_back(seq_of[s]->name().c_str());
    st.last.assign(blocks.size(), 0);
    if (!blocks.empty()) {
        if (!st.ctx) check(npge_anchor_create(&st.ctx, st.device), "no CUDA device");
        check(npge_block_hashes(st.ctx, names.data(), uint32_t(names.size()), blocks.size(),
                                offsets.data(), frag_seq.data(), min_pos.data(),
                                max_pos.data(), 0, ori.data(), st.last.data(), nullptr),
              "npge_block_hashes");
    }
    // process_block_impl (:37-49): seen in an EARLIER run -> move, else memorize;
    // the hashes of this run join the set only afterwards (after_thread_impl)
    std::vector<uint64_t> fresh;
    for (size_t b = 0; b < blocks.size(); b++) {
        if (std::binary_search(st.hashes.begin(), st.hashes.end(), st.last[b])) {
            o.detach(blocks[b]);
            t.insert(blocks[b]);
        } else {
            fresh.push_back(st.last[b]);
        }
    }
    st.hashes.insert(st.hashes.end(), fresh.begin(), fresh.end());
    std::sort(st.hashes.begin(), st.hashes.end());  // finish_work_impl: sort_unique
    st.hashes.erase(std::unique(st.hashes.begin(), st.hashes.end()), st.hashes.end());
}

const char* MoveUnchanged::name_impl() const {
    return "Move unchanged blocks from other to target";  // MoveUnchanged.cpp:71-73
}

}  // namespace gpu
}  // namespace npge
