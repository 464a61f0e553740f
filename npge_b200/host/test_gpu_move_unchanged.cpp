// test_gpu_move_unchanged.cpp -- npge::gpu::MoveUnchanged against the
// reference's semantics (src/algo/MoveUnchanged.cpp:37-69) and its block_hash
// (src/model/block_hash.cpp:29-54), restated below on the facade's Fragment::id
// as the checker.  The way AnchorLoop uses it (src/algo/lua_lib.lua:718-721):
// AnchorFinder adds anchor blocks to `cons`, MoveUnchanged moves the ones seen
// in an earlier iteration away.
//
//   test_gpu_move_unchanged            all cases (needs a CUDA device)
//   test_gpu_move_unchanged --no-gpu   only the cases that never reach the device
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "gpu_anchor_finder.hpp"

using namespace npge;

static int failures = 0, checks = 0;

#define CHECK(cond)                                                             \
    do {                                                                        \
        checks++;                                                               \
        if (!(cond)) {                                                          \
            failures++;                                                         \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);         \
        }                                                                       \
    } while (0)

// block_hash.cpp:29-54 on strings (the checker)
static uint64_t cpu_block_hash(const Block* block) {
    std::vector<std::string> dir, inv;
    for (Block::const_iterator it = block->begin(); it != block->end(); ++it) {
        const Fragment* f = *it;
        dir.push_back(f->id());
        Fragment g(f->seq(), f->min_pos(), f->max_pos(), -f->ori());
        inv.push_back(g.id());
    }
    std::sort(dir.begin(), dir.end());
    std::sort(inv.begin(), inv.end());
    const std::vector<std::string>& ids = dir < inv ? dir : inv;
    std::string joint;
    for (size_t i = 0; i < ids.size(); i++) joint += (i ? " " : "") + ids[i];
    joint.resize((joint.size() + 15) / 16 * 16, ' ');
    uint64_t a = 1;
    for (size_t i = 0; i < joint.size(); i += 16) {
        uint64_t v0, v1;
        std::memcpy(&v0, joint.data() + i, 8);
        std::memcpy(&v1, joint.data() + i + 8, 8);
        a *= v0;
        a ^= v1;
    }
    return a;
}

static SequencePtr named(const char* name, const char* text) {
    SequencePtr s = std::make_shared<InMemorySequence>(text);
    s->set_name(name);
    return s;
}

static Block* block_of(std::initializer_list<Fragment*> fragments) {
    Block* b = new Block;
    for (Fragment* f : fragments) b->insert(f);
    return b;
}

static void names_and_options() {
    gpu::MoveUnchanged mu;
    CHECK(std::string(mu.name()) == "Move unchanged blocks from other to target");
    CHECK(mu.has_opt("workers"));
    Meta meta;
    gpu::register_processors(&meta);
    CHECK(meta.has("MoveUnchanged"));
}

static void empty_other_needs_no_device() {
    gpu::MoveUnchanged mu;
    mu.set_block_set(new_bs());
    mu.set_other(new_bs());
    mu.run();
    CHECK(mu.last_hashes().empty());
}

static void hashes_and_moves() {
    SequencePtr s1 = named("g1&chr1&c", "ATGCATGCATGCATGCATGCATGCATGCATGCATGCATGC");
    SequencePtr s2 = named("g2&chr1&c", "TTGCATGCATGCATGCATGCATGCATGCATGCATGCATTT");
    SequencePtr s10 = named("g10&chr1&c", "ATGCATGCATGCATGCATGCATGCATGCATGC");
    BlockSetPtr cons = new_bs(), null_bs = new_bs();
    for (SequencePtr s : {s1, s2, s10}) cons->add_sequence(s);
    Block* a = block_of({new Fragment(s1, 0, 9, 1), new Fragment(s2, 12, 21, -1),
                         new Fragment(s10, 3, 12, 1)});
    Block* b = block_of({new Fragment(s2, 5, 5, -1), new Fragment(s1, 5, 5, 1)});  // one letter
    Block* c = block_of({new Fragment(s10, 20, 29, -1), new Fragment(s1, 100, 123, 1),
                         new Fragment(s1, 10, 12, -1), new Fragment(s2, 9, 10, 1)});
    for (Block* x : {a, b, c}) cons->insert(x);

    gpu::MoveUnchanged mu;
    mu.set_block_set(null_bs);
    mu.set_other(cons);
    std::vector<Block*> order(cons->begin(), cons->end());
    mu.run();  // first run: everything is new, nothing moves
    CHECK(cons->size() == 3 && null_bs->size() == 0);
    CHECK(mu.last_hashes().size() == 3);
    for (size_t i = 0; i < order.size(); i++) {
        CHECK(mu.last_hashes()[i] == cpu_block_hash(order[i]));
    }
    // inverting every fragment of a block does not change its hash
    Block* a_inv = block_of({new Fragment(s1, 0, 9, -1), new Fragment(s2, 12, 21, 1),
                             new Fragment(s10, 3, 12, -1)});
    CHECK(cpu_block_hash(a_inv) == cpu_block_hash(a));
    // second iteration: `a` (inverted) and `c` come back, `d` is new
    Block* d = block_of({new Fragment(s1, 1, 10, 1), new Fragment(s2, 13, 22, -1)});
    cons->detach(b);
    delete b;
    cons->detach(a);
    delete a;
    cons->insert(a_inv);
    cons->insert(d);
    mu.run();
    CHECK(null_bs->size() == 2 && cons->size() == 1);
    CHECK(null_bs->has(a_inv) && null_bs->has(c) && cons->has(d));
    // third iteration: d is known now
    mu.run();
    CHECK(null_bs->size() == 3 && cons->empty());
    // clear(): nothing is known any more
    null_bs->detach(d);
    cons->insert(d);
    mu.clear();
    mu.run();
    CHECK(cons->size() == 1 && null_bs->size() == 2);
}

static void after_anchor_finder() {
    // AnchorFinder -> MoveUnchanged twice on the same sequences: the second run of
    // the pair finds nothing new to keep (every anchor block was seen before)
    BlockSetPtr cons = new_bs(), null_bs = new_bs();
    cons->add_sequence(named("a&chr1&c", "TGGTCCGAGATGCGGGCCCTTTCGGACAGTCATTGGTCCGAGATGCGGGCC"));
    cons->add_sequence(named("b&chr1&c", "CCTTGGTCCGAGATGCGGGCCAAACGTACGTACCAGT"));
    gpu::MoveUnchanged mu;
    mu.set_block_set(null_bs);
    mu.set_other(cons);
    size_t first = 0;
    for (int iteration = 0; iteration < 2; iteration++) {
        gpu::AnchorFinder af;  // a new instance: no used_hashes_ carried over
        af.set_block_set(cons);
        af.set_opt_value("anchor-size", 10);
        af.set_opt_value("anchor-similar", false);
        af.run();
        if (iteration == 0) {
            first = cons->size();
            CHECK(first > 0);
            for (BlockSet::const_iterator it = cons->begin(); it != cons->end(); ++it) {
                CHECK((*it)->size() >= 2);
            }
        } else {
            CHECK(cons->size() == 2 * first);
        }
        std::vector<Block*> order(cons->begin(), cons->end());
        mu.run();
        for (size_t i = 0; i < order.size(); i++) {
            CHECK(mu.last_hashes()[i] == cpu_block_hash(order[i]));
        }
        if (iteration == 0) {
            CHECK(null_bs->empty() && cons->size() == first);
        } else {
            // the old copies AND the new copies carry known hashes: all move
            CHECK(cons->empty() && null_bs->size() == 2 * first);
        }
    }
}

int main(int argc, char** argv) {
    const bool no_gpu = argc > 1 && std::strcmp(argv[1], "--no-gpu") == 0;
    names_and_options();
    empty_other_needs_no_device();
    if (!no_gpu) {
        try {
            hashes_and_moves();
            after_anchor_finder();
        } catch (const std::exception& e) {
            failures++;
            std::printf("FAIL exception: %s\n", e.what());
        }
    }
    std::printf("%d checks, %d failures%s\n", checks, failures,
                no_gpu ? " (no-gpu subset)" : "");
    return failures ? 1 : 0;
}
