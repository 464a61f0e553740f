// block_hash_names.h -- host side of block_hash_core.h: the table that lets the
// device order fragment ids without building strings.
//
// An id is name + "_" + begin + "_" + last (Fragment::id,
// src/model/Fragment.cpp:173-183).  Two ids with different names are ordered by
// name + "_" alone provided neither name + "_" is a prefix of the other (then the
// first difference lies inside both prefixes).  That holds for every name
// without '_' and for NPGE's "GENOME&chr&c" names; the names of a set that breaks it
// (say "s1" next to "s1_2") are marked `slow`, and ids of two slow names are ordered
// by their regenerated strings (KeyOrder::string_less in block_hash_core.h).
#pragma once

#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

namespace npge_b200 {

struct NameTable {
    std::vector<uint32_t> seq_rank;  // [n_seqs] rank of the sequence's name
    std::vector<uint64_t> words;     // names in rank order, 8 bytes per word, little-endian,
                                     // every name starting a new word (zero padded)
    std::vector<uint32_t> pos;       // [ranks] first word of the name
    std::vector<uint32_t> len;       // [ranks] its length in bytes
    std::vector<unsigned char> slow; // [ranks] name + "_" is a prefix of another one, or
                                     // has one as a prefix; empty: no such name
};

// bcd4[v] = the decimal digits of v < 10000, one per nibble (block_hash_core.h: bcd10)
inline std::vector<uint16_t> make_bcd4() {
    std::vector<uint16_t> t(10000);
    for (uint32_t v = 0; v < 10000; v++) {
        t[v] = uint16_t(((v / 1000) << 12) | ((v / 100 % 10) << 8) | ((v / 10 % 10) << 4) | (v % 10));
    }
    return t;
}

// Returns "" or the reason the names cannot be used.
inline std::string build_name_table(const char* const* names, uint32_t n_seqs,
                                    NameTable& t) {
    std::vector<std::string> key(n_seqs);
    for (uint32_t i = 0; i < n_seqs; i++) {
        if (!names[i]) return "NULL sequence name";
        key[i] = std::string(names[i]) + "_";
        for (size_t j = 0; j + 1 < key[i].size(); j++) {
            // the joint string separates ids with ' ': comparing joint strings
            // equals comparing the lists only if every id byte is above ' '
            if (static_cast<unsigned char>(key[i][j]) <= 0x20) {
                return "sequence name with a space or control character: '" +
                       std::string(names[i]) + "'";
            }
        }
    }
    std::vector<std::string> sorted(key);
    std::sort(sorted.begin(), sorted.end());
    sorted.erase(std::unique(sorted.begin(), sorted.end()), sorted.end());
    // the keys that have a given key as a prefix follow it directly in sorted order
    std::vector<unsigned char> slow(sorted.size(), 0);
    bool any_slow = false;
    for (size_t i = 0; i < sorted.size(); i++) {
        const std::string& a = sorted[i];
        for (size_t j = i + 1; j < sorted.size(); j++) {
            const std::string& b = sorted[j];
            if (b.size() <= a.size() || b.compare(0, a.size(), a) != 0) break;
            slow[i] = slow[j] = 1;
            any_slow = true;
        }
    }
    t.slow.clear();
    if (any_slow) t.slow = slow;
    t.seq_rank.resize(n_seqs);
    for (uint32_t i = 0; i < n_seqs; i++) {
        t.seq_rank[i] = uint32_t(std::lower_bound(sorted.begin(), sorted.end(), key[i]) -
                                 sorted.begin());
    }
    t.words.clear();
    t.pos.clear();
    t.len.clear();
    for (const std::string& s : sorted) {
        const size_t n = s.size() - 1;  // without the '_'
        t.pos.push_back(uint32_t(t.words.size()));
        t.len.push_back(uint32_t(n));
        for (size_t i = 0; i < n; i += 8) {
            uint64_t w = 0;
            for (size_t j = i; j < n && j < i + 8; j++) {
                w |= uint64_t(static_cast<unsigned char>(s[j])) << (8 * (j - i));
            }
            t.words.push_back(w);
        }
    }
    if (t.words.empty()) t.words.push_back(0);  // never a zero-sized upload
    if (t.pos.empty()) {
        t.pos.push_back(0);
        t.len.push_back(0);
    }
    return "";
}

}  // namespace npge_b200
