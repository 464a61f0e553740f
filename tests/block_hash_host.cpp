// Test harness: runs npge_b200/csrc/block_hash_core.h (the code of the device
// kernel, compiled here for the host) over a flat block list, one block after
// the other.  Built by tests/test_block_hash.py with g++; test infrastructure
// only -- the product library runs the same header on the device and has no
// host path.
#include <cstring>
#include <string>
#include <vector>

#include "../npge_b200/csrc/block_hash_core.h"
#include "../npge_b200/csrc/block_hash_names.h"

static std::string g_error;
static uint32_t fast_cap_override = 0;

extern "C" void bh_host_set_fast_cap(uint32_t cap) { fast_cap_override = cap; }

extern "C" const char* bh_host_error() { return g_error.c_str(); }

extern "C" int bh_host_block_hashes(const char* const* names, uint32_t n_seqs,
                                    uint64_t n_blocks, const uint64_t* block_off,
                                    const uint32_t* frag_seq, const uint32_t* min_pos,
                                    const uint32_t* max_pos, uint32_t frag_len,
                                    const int8_t* frag_ori, uint64_t* hashes) {
    npge_b200::NameTable t;
    g_error = npge_b200::build_name_table(names, n_seqs, t);
    if (!g_error.empty()) return 1;
    static const std::vector<uint16_t> bcd4 = npge_b200::make_bcd4();
    npge_b200::FlatBlocks fb = {frag_seq, min_pos, max_pos, frag_len, frag_ori,
                                t.seq_rank.data(), t.words.data(), t.pos.data(), t.len.data(),
                                bcd4.data(), t.slow.empty() ? nullptr : t.slow.data()};
    // the kernel's layout: the slots of a thread interleaved with its neighbours'
    const uint32_t stride = 7, fast_cap = fast_cap_override ? fast_cap_override : 20;
    std::vector<npge_b200::IdKey> slots(size_t(fast_cap) * stride + 1), scratch;
    for (uint64_t b = 0; b < n_blocks; b++) {
        scratch.resize(2 * (block_off[b + 1] - block_off[b]) + 1);
        npge_b200::KeySlots mine = {slots.data() + b % stride, stride};
        hashes[b] = npge_b200::block_hash_flat(fb, block_off[b], block_off[b + 1], mine,
                                               fast_cap, scratch.data());
        // both paths must agree on every block
        if (hashes[b] != npge_b200::block_hash_two_lists(fb, block_off[b], block_off[b + 1],
                                                         scratch.data())) {
            g_error = "one-list and two-list paths disagree";
            return 2;
        }
    }
    return 0;
}
