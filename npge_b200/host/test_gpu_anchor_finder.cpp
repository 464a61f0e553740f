// test_gpu_anchor_finder.cpp -- the reference's unit tests of AnchorFinder
// (src/test/anchor_finder.cpp:18-110), run against npge::gpu::AnchorFinder.
// BOOST_CHECK -> CHECK (counted as failure), BOOST_WARN -> WARN (printed only).
//
//   test_gpu_anchor_finder            all cases (needs a CUDA device)
//   test_gpu_anchor_finder --no-gpu   only the cases that never reach the device
#include <malloc.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "gpu_anchor_finder.hpp"

using namespace npge;

static int failures = 0, warnings = 0, checks = 0;

#define CHECK(cond)                                                             \
    do {                                                                        \
        checks++;                                                               \
        if (!(cond)) {                                                          \
            failures++;                                                         \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);         \
        }                                                                       \
    } while (0)
#define WARN(cond)                                                              \
    do {                                                                        \
        if (!(cond)) {                                                          \
            warnings++;                                                         \
            std::printf("warn %s:%d: %s\n", __FILE__, __LINE__, #cond);         \
        }                                                                       \
    } while (0)

static BlockSetPtr blockset_of(std::initializer_list<const char*> texts) {
    BlockSetPtr bs = new_bs();
    for (const char* t : texts) bs->add_sequence(std::make_shared<InMemorySequence>(t));
    return bs;
}

// ---- cases that need no device ---------------------------------------------------
static void options_and_rules() {
    gpu::AnchorFinder af;
    CHECK(af.opt_value("anchor-size").as<int>() == 20);
    CHECK(af.opt_value("anchor-similar").as<bool>() == true);
    CHECK(af.opt_value("max-anchor-fragments").as<int>() == 100000);
    CHECK(af.opt_value("anchor-fp").as<double>() > 0.09);
    CHECK(af.has_opt("workers") && af.has_opt("timing"));
    CHECK(std::string(af.name()) == "Find anchors");
    af.set_block_set(blockset_of({"ATGC"}));
    for (int bad : {0, -5, 33}) {
        af.set_opt_value("anchor-size", bad);
        bool thrown = false;
        try {
            af.run();
        } catch (const Exception& e) {
            thrown = std::strstr(e.what(), "options") != nullptr;
        }
        CHECK(thrown);
    }
    bool thrown = false;
    try {
        af.set_opt_value("no-such-option", 1);
    } catch (const Exception&) {
        thrown = true;
    }
    CHECK(thrown);
    af.set_options("--anchor-size 7 --anchor-similar 0 --max-anchor-fragments 12");
    CHECK(af.opt_value("anchor-size").as<int>() == 7);
    CHECK(af.opt_value("anchor-similar").as<bool>() == false);
    CHECK(af.opt_value("max-anchor-fragments").as<int>() == 12);
}

static void workers_zero_skips_the_run() {  // Processor.cpp:667
    gpu::AnchorFinder af;
    BlockSetPtr bs = blockset_of({"GAAAGAAA"});
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 3);
    af.set_workers(0);
    af.run();  // must not touch the device
    CHECK(bs->size() == 0);
}

static void meta_registration() {
    Meta meta;
    gpu::register_processors(&meta);
    CHECK(meta.has("AnchorFinder"));
    std::shared_ptr<Processor> p = meta.get("AnchorFinder");
    CHECK(p->key() == "AnchorFinder");
    CHECK(p->has_opt("anchor-size"));
}

static void sequence_normalisation() {  // test-lua/sequence.lua:1-40
    InMemorySequence s("AT-G-C-A#T$N .C");
    CHECK(s.contents() == "ATGCATNC");
    InMemorySequence r("acgtRYn");
    CHECK(r.contents() == "ACGTNNN");
    InMemorySequence named("ATGC", "BRUAB&chr1&c");
    CHECK(named.genome() == "BRUAB");
    InMemorySequence plain("ATGC", "plasmid");
    CHECK(plain.genome() == "");
}

// ---- src/test/anchor_finder.cpp ----------------------------------------------------
static void AnchorFinder_main() {
    BlockSetPtr bs = blockset_of({"tgGTCCGagCGGACggcc"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 5);
    af.run();
    WARN(bs->size() == 1);
    if (bs->size() == 1) {
        Fragment* f = bs->front()->front();
        CHECK(f->str() == "GTCCG" || f->str() == "CGGAC");
    }
}

static void AnchorFinder_n_negative() {
    BlockSetPtr bs = blockset_of({"tgGTNCGagCGNACggcc"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 5);
    af.run();
    CHECK(bs->size() == 0);
}

static void AnchorFinder_n_positive() {
    BlockSetPtr bs = blockset_of({"GTNCGATAnnnGTNCGATA"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 5);
    af.run();
    CHECK(bs->size() > 0);
}

static void AnchorFinder_palindrome_elimination() {
    BlockSetPtr bs = blockset_of({"ATGCAT"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 6);
    af.run();
    CHECK(bs->size() == 0);
}

static void AnchorFinder_one_from_long_repeat() {
    BlockSetPtr bs = blockset_of({"GAAAGAAA"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 3);
    af.run();
    WARN(bs->size() == 1);
}

static void AnchorFinder_several_sequences(int workers) {
    BlockSetPtr bs = blockset_of({"GAAAGAAA", "GAAAGAAA"});
    gpu::AnchorFinder af;
    af.set_block_set(bs);
    af.set_opt_value("anchor-size", 3);
    if (workers) af.set_workers(workers);  // AnchorFinder_two_workers
    af.run();
    WARN(bs->size() >= 1 && bs->front()->size() == 4);
    for (BlockSet::const_iterator b = bs->begin(); b != bs->end(); ++b) {
        CHECK((*b)->size() >= 2);
        for (Block::const_iterator f = (*b)->begin(); f != (*b)->end(); ++f) {
            CHECK((*f)->length() == 3 && (*f)->str() == (*b)->front()->str());
        }
    }
}

// the processor's cross-run state: hashes already returned are not found again
// by the same instance (used_hashes_, AnchorFinder.cpp:385-389)
static void used_hashes_across_runs() {
    gpu::AnchorFinder af;
    af.set_opt_value("anchor-size", 4);
    af.set_opt_value("anchor-similar", false);
    BlockSetPtr bs1 = blockset_of({"ACGGTTCAACGGTTCA"});
    af.set_block_set(bs1);
    af.run();
    const size_t first = bs1->size();
    CHECK(first >= 2);
    BlockSetPtr bs2 = blockset_of({"ACGGTTCAACGGTTCA"});
    af.set_block_set(bs2);
    af.run();
    CHECK(bs2->size() <= 1 && bs2->size() < first);  // all but the first hash were used
    gpu::AnchorFinder fresh;
    fresh.set_opt_value("anchor-size", 4);
    fresh.set_opt_value("anchor-similar", false);
    BlockSetPtr bs3 = blockset_of({"ACGGTTCAACGGTTCA"});
    fresh.set_block_set(bs3);
    fresh.run();
    CHECK(bs3->size() == first);
}

static void truncation() {
    gpu::AnchorFinder af;
    af.set_opt_value("anchor-size", 4);
    af.set_opt_value("anchor-similar", false);
    af.set_opt_value("max-anchor-fragments", 4);
    BlockSetPtr bs = blockset_of({"ACGGTTCAACGGTTCAACGGTTCA"});
    af.set_block_set(bs);
    af.run();
    size_t frags = 0;
    for (BlockSet::const_iterator b = bs->begin(); b != bs->end(); ++b) frags += (*b)->size();
    CHECK(frags <= 4 && bs->size() >= 1);
}

// canonical text of a blockset: sorted blocks of sorted "seq:min:max:ori" tuples (a
// sequence is named by its size: the blockset orders its sequences by address)
static std::string canonical(const BlockSet& bs) {
    std::vector<std::string> blocks;
    for (BlockSet::const_iterator b = bs.begin(); b != bs.end(); ++b) {
        std::vector<std::string> fr;
        for (Block::const_iterator f = (*b)->begin(); f != (*b)->end(); ++f) {
            fr.push_back(std::to_string((*f)->seq()->size()) + ":" + std::to_string((*f)->min_pos()) + ":" +
                         std::to_string((*f)->max_pos()) + ":" + std::to_string((*f)->ori()));
        }
        std::sort(fr.begin(), fr.end());
        std::string t;
        for (size_t i = 0; i < fr.size(); i++) t += fr[i] + " ";
        blocks.push_back(t);
    }
    std::sort(blocks.begin(), blocks.end());
    std::string out;
    for (size_t i = 0; i < blocks.size(); i++) out += blocks[i] + "\n";
    return out;
}

static std::string pseudo_genome(unsigned seed, size_t n) {
    std::string s(n, 'A');
    unsigned x = seed * 2654435761u + 12345u;
    for (size_t i = 0; i < n; i++) {
        x = x * 1664525u + 1013904223u;
        s[i] = "ATGC"[(x >> 24) & 3];
    }
    return s;
}

// several GPUs behind the Processor (set_devices): the 2- and 3-rank jobs (ranks share
// the devices of this box) give the blockset of the one-device run
static void several_devices_equal_one_device() {
    const std::string a = pseudo_genome(1, 9000);
    std::string b = a.substr(2000, 5000) + pseudo_genome(2, 1500) + a.substr(100, 900);
    std::string c = pseudo_genome(3, 700) + a.substr(4000, 2500) + "NNNNNNNN" + b.substr(10, 600);
    for (int similar = 0; similar < 2; similar++) {
        std::string want;
        for (int ranks = 1; ranks <= 3; ranks++) {
            BlockSetPtr bs = blockset_of({a.c_str(), b.c_str(), c.c_str()});
            gpu::AnchorFinder af;
            af.set_devices(std::vector<int>(size_t(ranks), 0));
            af.set_block_set(bs);
            af.set_opt_value("anchor-size", 16);
            af.set_opt_value("anchor-similar", similar != 0);
            af.set_opt_value("max-anchor-fragments", similar ? 500 : -1);  // negative: no limit
            af.run();
            CHECK(bs->size() > 0);
            if (ranks == 1) want = canonical(*bs);
            CHECK(canonical(*bs) == want);
        }
    }
}

// the `workers` option sets the threads that build and check the Block / Fragment
// objects: the block set does not depend on it
static void workers_build_the_same_blocks() {
    const std::string a = pseudo_genome(21, 60000);
    const std::string b = a.substr(5000, 40000) + pseudo_genome(22, 3000);
    std::string want;
    for (int workers = 1; workers <= 5; workers += 2) {
        for (int ranks = 1; ranks <= 2; ranks++) {
            BlockSetPtr bs = blockset_of({a.c_str(), b.c_str()});
            gpu::AnchorFinder af;
            af.set_devices(std::vector<int>(size_t(ranks), 0));
            af.set_block_set(bs);
            af.set_workers(workers);
            af.set_opt_value("anchor-size", 14);
            af.set_opt_value("anchor-similar", false);
            af.set_opt_value("max-anchor-fragments", -1);
            af.run();
            CHECK(bs->size() > 20000);
            if (want.empty()) want = canonical(*bs);
            CHECK(canonical(*bs) == want);
        }
    }
}

// AnchorLoop re-runs one instance on the same sequences (MetaProcessor.cpp:32-38): the
// second run finds the packed genome still on the device and skips the upload
static void resident_genome_is_reused() {
    const std::string a = pseudo_genome(7, 6000);
    const std::string b = a.substr(1000, 3000) + pseudo_genome(8, 800);
    for (int ranks = 1; ranks <= 2; ranks++) {
        BlockSetPtr bs = blockset_of({a.c_str(), b.c_str()});
        gpu::AnchorFinder af;
        af.set_devices(std::vector<int>(size_t(ranks), 0));
        af.set_block_set(bs);
        af.set_opt_value("anchor-size", 12);
        af.set_opt_value("anchor-similar", false);
        af.run();
        CHECK(!af.last_run_reused_upload());
        const size_t first = bs->size();
        CHECK(first > 0);
        af.run();  // same blockset, same sequences: used hashes leave at most one block
        CHECK(af.last_run_reused_upload());
        CHECK(bs->size() <= first + 1);
        af.set_opt_value("anchor-size", 13);  // another k: the layout changes
        af.run();
        CHECK(!af.last_run_reused_upload());
        BlockSetPtr other = blockset_of({b.c_str(), a.c_str()});
        af.set_block_set(other);
        af.run();
        CHECK(!af.last_run_reused_upload());
    }
}

// ---- FragmentFinder ---------------------------------------------------------------
static void fragment_finder_options() {
    gpu::FragmentFinder ff;
    CHECK(ff.opt_value("max-matches").as<int>() == 1000);
    CHECK(std::string(ff.name()) == "Locate fragment by its sequence");
    ff.set_block_set(blockset_of({"ATGC"}));
    bool thrown = false;
    try {
        ff.run();  // empty pattern
    } catch (const Exception& e) {
        thrown = std::strstr(e.what(), "pattern") != nullptr;
    }
    CHECK(thrown);
    ff.set_opt_value("pattern", std::string("--"));  // nothing left after to_atgcn
    thrown = false;
    try {
        ff.run();
    } catch (const Exception&) {
        thrown = true;
    }
    CHECK(thrown);
    ff.set_opt_value("pattern", std::string("AT"));
    ff.set_opt_value("max-matches", 0);
    thrown = false;
    try {
        ff.run();
    } catch (const Exception&) {
        thrown = true;
    }
    CHECK(thrown);
    Meta meta;
    gpu::register_processors(&meta);
    CHECK(meta.has("FragmentFinder") && meta.get("FragmentFinder")->has_opt("pattern"));
}

static void fragment_finder_run() {
    BlockSetPtr bs = blockset_of({"tgGTCCGagCGGACggcc", "aaGTCCGnn"});
    gpu::FragmentFinder ff;
    ff.set_block_set(bs);
    ff.set_opt_value("pattern", std::string("gtccg"));
    ff.run();
    CHECK(bs->size() == 3);
    int forward = 0, reverse = 0;
    for (BlockSet::const_iterator b = bs->begin(); b != bs->end(); ++b) {
        CHECK((*b)->size() == 1);
        Fragment* f = (*b)->front();
        CHECK(f->str() == "GTCCG" && f->length() == 5);
        (f->ori() == 1 ? forward : reverse)++;
    }
    CHECK(forward == 2 && reverse == 1);
    BlockSetPtr cut = blockset_of({"ATATATATAT"});
    ff.set_block_set(cut);
    ff.set_opt_value("pattern", std::string("AT"));
    ff.set_opt_value("max-matches", 3);
    ff.run();
    CHECK(cut->size() == 3);
}

// ---- --bench: where the time of run() goes for a caller with real objects ---------------
// A C1-shaped input (3 strains x 3.3 Mbp, 0.5 % SNPs) through the Processor, parity
// mode: linearising the sequences, the library call, and building Fragment / Block
// objects one `new` at a time (the reference's ownership model, Block.cpp:45-50).
#include <chrono>
static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int bench_run(int n_devices, int workers) {
    const std::string anc = pseudo_genome(11, 3300000);
    std::vector<std::string> strains;
    for (unsigned st = 0; st < 3; st++) {
        std::string g = anc;
        unsigned x = st * 7919u + 17u;
        for (size_t i = 0; i < g.size(); i += 150 + (x >> 26)) {
            x = x * 1664525u + 1013904223u;
            g[i] = "ATGC"[(x >> 20) & 3];
        }
        // distinct sizes: the scan order (size descending) must not depend on the heap
        // addresses the blockset orders equal-sized sequences by (the result of the
        // anchor-similar rule depends on that order, in the reference too)
        g.resize(g.size() - 1000 * st);
        strains.push_back(g);
    }
    for (int mode = 0; mode < 2; mode++) {
        BlockSetPtr bs = blockset_of({strains[0].c_str(), strains[1].c_str(), strains[2].c_str()});
        gpu::AnchorFinder af;
        af.set_devices(std::vector<int>(size_t(n_devices), 0));
        af.set_block_set(bs);
        af.set_opt_value("anchor-similar", mode == 1);
        if (mode == 0) af.set_opt_value("max-anchor-fragments", -1);
        double best = 1e9, best2 = 1e9;
        size_t blocks = 0, frags = 0;
        for (int rep = 0; rep < 3; rep++) {
            bs->clear_blocks();
            gpu::AnchorFinder fresh;  // no used hashes from the previous repetition
            fresh.set_devices(std::vector<int>(size_t(n_devices), 0));
            fresh.set_block_set(bs);
            fresh.set_opt_value("anchor-similar", mode == 1);
            if (mode == 0) fresh.set_opt_value("max-anchor-fragments", -1);
            fresh.set_workers(workers);  // threads of the object build
            double t0 = now_s();
            fresh.run();
            if (now_s() - t0 < best) {
                const double* ms = fresh.last_run_ms();
                std::printf("  rep %d: linearise + upload %.1f ms, search + copy %.1f ms, "
                            "objects %.1f ms\n", rep, ms[0], ms[1], ms[2]);
            }
            best = std::min(best, now_s() - t0);
            blocks = bs->size();
            frags = 0;
            for (BlockSet::const_iterator b = bs->begin(); b != bs->end(); ++b) frags += (*b)->size();
            bs->clear_blocks();
            // the millions of freed Block / Fragment chunks sit in glibc's fast bins; the
            // next large malloc would consolidate them (~450 ms here) inside the timed run
            malloc_trim(0);
            t0 = now_s();
            fresh.run();  // resident genome, used hashes: mostly the fixed costs
            if (now_s() - t0 < best2) {
                const double* ms = fresh.last_run_ms();
                std::printf("  rep %d again (reused upload: %d): linearise + upload %.1f ms, "
                            "search + copy %.1f ms (search %.1f), objects %.1f ms\n", rep,
                            int(fresh.last_run_reused_upload()), ms[0], ms[1], ms[3], ms[2]);
            }
            best2 = std::min(best2, now_s() - t0);
        }
        std::printf("bench %s devices=%d workers=%d: run() %.1f ms (%zu blocks, %zu fragments); "
                    "second run on the resident genome %.1f ms\n",
                    mode == 0 ? "parity" : "default", n_devices, workers, best * 1e3, blocks, frags,
                    best2 * 1e3);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::strcmp(argv[1], "--bench") == 0) {
        // --bench [n_devices [workers]]
        return bench_run(argc > 2 ? std::atoi(argv[2]) : 1, argc > 3 ? std::atoi(argv[3]) : 1);
    }
    const bool no_gpu = argc > 1 && std::strcmp(argv[1], "--no-gpu") == 0;
    options_and_rules();
    workers_zero_skips_the_run();
    meta_registration();
    sequence_normalisation();
    fragment_finder_options();
    if (!no_gpu) {
        try {
            AnchorFinder_main();
            AnchorFinder_n_negative();
            AnchorFinder_n_positive();
            AnchorFinder_palindrome_elimination();
            AnchorFinder_one_from_long_repeat();
            AnchorFinder_several_sequences(0);
            AnchorFinder_several_sequences(2);
            used_hashes_across_runs();
            truncation();
            several_devices_equal_one_device();
            resident_genome_is_reused();
    workers_build_the_same_blocks();
            fragment_finder_run();
        } catch (const std::exception& e) {
            failures++;
            std::printf("FAIL exception: %s\n", e.what());
        }
    }
    std::printf("%d checks, %d failures, %d warnings%s\n", checks, failures, warnings,
                no_gpu ? " (no-gpu subset)" : "");
    return failures ? 1 : 0;
}
