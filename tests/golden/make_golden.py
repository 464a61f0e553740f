#!/usr/bin/env python
"""Generates tests/golden/*.json from the reference tree (dev container only).

Sources (all under /root/reference):
  * test-script/anchor_finder/<case>/{in,out}.fasta  -- the reference's golden
    files for the AnchorFinder script test (compared there by blockset_hash,
    src/test/meta_test.cxx:28-39,84-86); copied verbatim as data.
  * src/test/anchor_finder.cpp:18-110, src/test/hash.cpp:16-76,
    src/test/complement.cpp:55-90 -- the literal inputs of the unit tests,
    transcribed below with their line numbers.
  * hash known-answer vectors: OUTPUTS OF THE REFERENCE'S OWN CODE, produced by
    oracle/_ref/libnpge_ref_hash.so (src/util/make_hash.hpp, complement.cpp
    compiled in place, see oracle/Makefile).

Run:  python tests/golden/make_golden.py     (rewrites the json files)
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle  # noqa: E402


def read_fasta(path):
    recs = []
    name, chunks = None, []
    for line in open(path):
        line = line.rstrip("\n")
        if line.startswith(">"):
            if name is not None:
                recs.append((name, "".join(chunks)))
            name, chunks = line[1:], []
        else:
            chunks.append(line)
    if name is not None:
        recs.append((name, "".join(chunks)))
    return recs


def script_cases():
    base = os.path.join(REF, "test-script", "anchor_finder")
    cases = []
    for case in sorted(os.listdir(base)):
        d = os.path.join(base, case)
        if not os.path.isdir(d):
            continue
        seqs = [{"name": n.split()[0], "text": t}
                for n, t in read_fasta(os.path.join(d, "in.fasta"))]
        blocks = {}
        for header, text in read_fasta(os.path.join(d, "out.fasta")):
            parts = header.split()
            frag_id = parts[0]
            block = [p for p in parts if p.startswith("block=")][0][6:]
            blocks.setdefault(block, []).append({"id": frag_id, "text": text})
        cases.append({"case": case, "anchor_size": 20, "sequences": seqs,
                      "expected_blocks": list(blocks.values()),
                      "source": "test-script/anchor_finder/%s" % case})
    return cases


# src/test/anchor_finder.cpp -- literal inputs and what the test asserts.
UNIT_CASES = [
    {"name": "AnchorFinder_main", "line": "18-32", "k": 5, "workers": 1,
     "seqs": ["tgGTCCGagCGGACggcc"],
     "warn_blocks": 1, "first_fragment_in": ["GTCCG", "CGGAC"]},
    {"name": "AnchorFinder_n_negative", "line": "34-44", "k": 5, "workers": 1,
     "seqs": ["tgGTNCGagCGNACggcc"], "check_blocks": 0},
    {"name": "AnchorFinder_n_positive", "line": "46-57", "k": 5, "workers": 1,
     "seqs": ["GTNCGATAnnnGTNCGATA"], "check_blocks_gt": 0},
    {"name": "AnchorFinder_palindrome_elimination", "line": "59-69", "k": 6,
     "workers": 1, "seqs": ["ATGCAT"], "check_blocks": 0},
    {"name": "AnchorFinder_one_from_long_repeat", "line": "71-81", "k": 3,
     "workers": 1, "seqs": ["GAAAGAAA"], "warn_blocks": 1},
    {"name": "AnchorFinder_several_sequences", "line": "83-95", "k": 3,
     "workers": 1, "seqs": ["GAAAGAAA", "GAAAGAAA"], "warn_blocks": 1,
     "warn_front_size": 4},
    {"name": "AnchorFinder_two_workers", "line": "97-110", "k": 3,
     "workers": 2, "seqs": ["GAAAGAAA", "GAAAGAAA"], "warn_blocks_gte": 1,
     "warn_front_size": 4},
]

HASH_MAIN = "CGCAtacccTGCGgcaGGGTcaGGGC"          # src/test/hash.cpp:18
HASH_FULL = ("GATCCTCGATTAACAGTTTGGCCTGTTCCTATGTATGCCCTACTCCAAATGGT"
             "GCCAACTGGATCAATCCTCAGTGCCGCGGGAATCATGTCTTTATTTATGCTTT"
             "TCAGCTCTGCGAACTTAGGCTCAGCACAAGATTTAAGCGAGAAGCGAAAGCTG"
             "ACCGGCAGGGGGGGCACGGTTAATAACTAAGACTGTAGCGTGACAAACGGACC")  # :46-49
COMPLEMENT_STRINGS = [                             # src/test/complement.cpp:62-79
    "AAA", "ATGC", "ATGCAAAAG",
    "CTTTTGCAT", "CATATCGGTAAGGTGACATTA", "CGCCCGCGAGTGGAATACAACACCTCCACCTG",
    "CGCCCGCGAGTGGAATACAACACCTCCACCT", "CGCCCGCGAGTGGAATACAACACCTCCACC",
    "GCGCAAAAGAAAAAGGCCCCCAAACGTTGCCGT", "GCGCAAAAGAAAAAGGCCCCCAAACGTTGCGT",
    "ACGGCAACGTTTGGGGGCCTTTTTCTTTTGCGC", "CGCCCGCGAGTGGAATACAACACCTC",
]


def hash_vectors():
    import ctypes as C
    R = oracle.ref_lib()
    assert R is not None, "oracle/_ref not built"

    def mh(s, ori):
        b = s.encode()
        buf = C.create_string_buffer(b, len(b))
        if ori == 1:
            return R.ref_make_hash(buf, len(b), 1)
        return R.ref_make_hash(
            C.cast(C.addressof(buf) + len(b) - 1, C.c_char_p), len(b), -1)

    rng = random.Random(20261104)
    strings = list(COMPLEMENT_STRINGS) + [HASH_MAIN.upper(), HASH_FULL]
    # windows of the 212-bp test sequence, as hash_reuse_hash_full walks them
    for i in range(70, 100, 7):
        for length in (1, 2, 5, 16, 20, 31, 32, 33, 47, 64, 65, 99):
            strings.append(HASH_FULL[i:i + length])
    for length in list(range(1, 100)):
        strings.append("".join(rng.choice("ATGC") for _ in range(length)))
    for length in (5, 20, 32, 40):
        strings.append("".join(rng.choice("ATGCN") for _ in range(length)))
    vecs = []
    for s in strings:
        d = mh(s, 1)
        r = mh(s, -1)
        v = {"s": s, "dir": "%x" % d, "rev": "%x" % r,
             "complement_hash_of_dir": "%x" % R.ref_complement_hash(d, len(s))}
        vecs.append(v)
    rolls = []
    for _ in range(300):
        length = rng.randint(1, 99)
        s = "".join(rng.choice("ATGC") for _ in range(length + 1))
        h = mh(s[:length], 1)
        fwd = R.ref_reuse_hash(h, length, s[0].encode(), s[length].encode(), 1)
        h2 = mh(s[1:], 1)
        back = R.ref_reuse_hash(h2, length, s[length].encode(), s[0].encode(), 0)
        rolls.append({"s": s, "length": length, "h": "%x" % h,
                      "forward": "%x" % fwd, "h_shifted": "%x" % h2,
                      "backward": "%x" % back})
    return {"source": "outputs of /root/reference src/util/make_hash.hpp + "
                      "complement.cpp compiled in place (oracle/_ref)",
            "make_hash": vecs, "reuse_hash": rolls}


def main():
    json.dump({"script_cases": script_cases(), "unit_cases": UNIT_CASES,
               "source": "src/test/anchor_finder.cpp; test-script/anchor_finder"},
              open(os.path.join(HERE, "anchor_finder_cases.json"), "w"),
              indent=1)
    json.dump(hash_vectors(),
              open(os.path.join(HERE, "hash_vectors.json"), "w"), indent=0)
    print("golden files written")


if __name__ == "__main__":
    main()
