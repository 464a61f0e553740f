"""Small, fast coverage of every kernel on the path for compute-sanitizer
(tools/sanitize.sh): each result is checked against the CPU oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402

import oracle  # noqa: E402
from golden_util import load  # noqa: E402
from npge_b200 import synth  # noqa: E402
from npge_b200.engine import AnchorEngine, MultiEngine  # noqa: E402

BIG = 2 ** 62


def same(got, want):
    assert got.blocks() == want.blocks()
    assert list(got.block_hash) == list(want.block_hash)
    assert got.digest == want.digest
    assert list(got.new_used) == list(want.new_used)


def main():
    eng = AnchorEngine(0)
    cases = load("anchor_finder_cases.json")
    for case in cases["unit_cases"]:
        seqs = [oracle.to_atgcn(s) for s in case["seqs"]]
        for similar in (False, True):
            same(eng.find(seqs, case["k"], similar=similar, max_frags=100000),
                 oracle.find_ideal(seqs, case["k"], similar=similar, max_frags=100000))
    names, seqs = synth.generate("tiny")
    seqs = seqs + [seqs[0][:11], seqs[1][100:700], seqs[2][:0], "A" * 9000]  # + an oversized unit
    for k in (12, 20, 27):
        base = oracle.find_ideal(seqs, k, similar=False)
        used = base.block_hash[::3].copy()
        for similar, limit, u in ((False, BIG, None), (True, 100000, None), (False, 777, None),
                                  (True, 9, None), (True, BIG, used)):
            same(eng.find(seqs, k, similar=similar, max_frags=limit, used=u),
                 oracle.find_ideal(seqs, k, similar=similar, max_frags=limit, used=u))
    names, seqs = synth.generate("tiny")
    eng.upload(seqs, 20)
    eng.find_device(20, similar=False, max_frags=BIG)
    eng.anchor_block_hashes(names)
    pat = bytes(seqs[0][40:60]).decode()
    want = eng.find_fragments_device(pat, max_matches=50)
    import torch
    out = [torch.zeros(64, dtype=t, device="cuda") for t in (torch.int32, torch.int32, torch.int8)]
    n, _ = eng.find_fragments_device_noalloc(pat, 50, out[0].data_ptr(), out[1].data_ptr(),
                                             out[2].data_ptr(), 64)
    assert n == len(want)
    assert eng.debug_check() == 0
    eng.close()
    for ranks in (2, 3):
        m = MultiEngine([0] * ranks)
        for k, similar, limit in ((16, False, BIG), (20, True, 100000), (27, False, 501)):
            same(m.find(seqs, k, similar=similar, max_frags=limit),
                 oracle.find_ideal(seqs, k, similar=similar, max_frags=limit))
        assert m.debug_check() == 0
        m.close()
    print("driver ok")


if __name__ == "__main__":
    main()
