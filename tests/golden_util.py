"""Helpers shared by the golden-vector tests (oracle side and GPU side)."""
import json
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return json.load(open(os.path.join(GOLDEN, name)))


def parse_fragment_id(frag_id, names):
    """'abc_60_79' -> (seq_idx, min_pos, max_pos, ori); begin > last = reverse
    (Fragment::id, src/model/Fragment.cpp:173-183)."""
    name, a, b = frag_id.rsplit("_", 2)
    a, b = int(a), int(b)
    if b == -1:
        return names.index(name), a, a, -1
    if a <= b:
        return names.index(name), a, b, 1
    return names.index(name), b, a, -1


def expected_blocks(case):
    names = [s["name"] for s in case["sequences"]]
    return names, [[parse_fragment_id(f["id"], names) for f in block]
                   for block in case["expected_blocks"]]


def result_blocks4(res, k):
    """Result -> list of blocks of (seq, min, max, ori)."""
    out = []
    for b in range(res.n_blocks):
        lo, hi = int(res.block_offsets[b]), int(res.block_offsets[b + 1])
        out.append([(int(res.frag_seq[f]), int(res.frag_min_pos[f]),
                     int(res.frag_min_pos[f]) + k - 1, int(res.frag_ori[f]))
                    for f in range(lo, hi)])
    return out


def revcomp(s):
    return s[::-1].translate(str.maketrans("ATGC", "TACG"))
