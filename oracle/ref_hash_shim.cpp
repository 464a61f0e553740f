/*
 * oracle/ref_hash_shim.cpp -- C wrappers around the REFERENCE's own hash
 * arithmetic, compiled in place from /root/reference (never copied):
 *   src/util/make_hash.hpp, src/util/char_to_size.hpp,
 *   src/util/complement.hpp, src/util/complement.cpp
 * behind oracle/boost_shim.  The rest of AnchorFinder needs Boost, Lua and
 * luabind (src/CMakeLists.txt:13-60) and is unbuildable here, so this is the
 * one piece of the reference that runs: it pins the oracle's hash, roll and
 * reverse-complement functions (tests/test_oracle_vs_ref.py) and generates
 * tests/golden/hash_vectors.json.  TEST INFRASTRUCTURE ONLY; output goes to
 * oracle/_ref/ (git-ignored).
 */
#include "make_hash.hpp"
#include "complement.hpp"

extern "C" {

unsigned long long ref_make_hash(const char* start, int length, int ori) {
    return npge::make_hash(start, length, ori);
}

unsigned long long ref_reuse_hash(unsigned long long old_hash, int length,
                                  char remove_char, char add_char,
                                  int forward) {
    return npge::reuse_hash(old_hash, length, remove_char, add_char,
                            forward != 0);
}

unsigned long long ref_complement_hash(unsigned long long hash, int letters) {
    return npge::complement_hash(hash, letters);
}

char ref_complement_char(char c) {
    return npge::complement(c);
}

unsigned long ref_char_to_size(char c) {
    return npge::char_to_size(c);
}

}
