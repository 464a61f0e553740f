// block_hash_names.h -- host side of block_hash_core.h: the table that lets the
// device order fragment ids without building strings.
//
// An id is name + "_" + begin + "_" + last (Fragment::id,
// src/model/Fragment.cpp:173-183).  Two ids with different names are ordered by
// name + "_" alone provided no name + "_" is a prefix of another one (then the
// first difference lies inside both prefixes).  That holds for every name
// without '_' and for NPGE's "GENOME&chr&c" names; a name set that breaks it
// (say "s1" next to "s1_2") is refused rather than hashed wrongly.
#pragma once

#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

namespace npge_b200 {

struct NameTable {
    std::vector<uint32_t> seq_rank;  // [n_seqs] rank of the sequence's name
    std::vector<uint8_t> bytes;      // names in rank order, concatenated
    std::vector<uint32_t> off;       // [ranks + 1]
};

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
    for (size_t i = 0; i + 1 < sorted.size(); i++) {
        const std::string &a = sorted[i], &b = sorted[i + 1];
        if (b.size() > a.size() && b.compare(0, a.size(), a) == 0) {
            return "sequence name '" + a.substr(0, a.size() - 1) + "' followed by '_' is a "
                   "prefix of '" + b.substr(0, b.size() - 1) + "': ids are not ordered by name";
        }
    }
    t.seq_rank.resize(n_seqs);
    for (uint32_t i = 0; i < n_seqs; i++) {
        t.seq_rank[i] = uint32_t(std::lower_bound(sorted.begin(), sorted.end(), key[i]) -
                                 sorted.begin());
    }
    t.bytes.clear();
    t.off.assign(1, 0);
    for (const std::string& s : sorted) {
        t.bytes.insert(t.bytes.end(), s.begin(), s.end() - 1);  // without the '_'
        t.off.push_back(uint32_t(t.bytes.size()));
    }
    if (t.bytes.empty()) t.bytes.push_back(0);  // never a zero-sized upload
    return "";
}

}  // namespace npge_b200
