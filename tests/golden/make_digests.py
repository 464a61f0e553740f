#!/usr/bin/env python
"""Writes tests/golden/digests.json: counts and the 128-bit digest of the anchors
of BASELINE.json's configurations C1-C5 (and of bench.py's multi-GPU jobs),
computed by the CPU oracle in hash-range slices (oracle_find_ideal_sliced, the
same exact semantics as find_ideal with O(N / slices) memory).

The GPU tests (tests/test_gpu_configs.py) and bench.py compare the digest of
every device result with these values; nothing here runs on the GPU box.

    python tests/golden/make_digests.py            # everything (~10 min on 8 cores)
    python tests/golden/make_digests.py brucella17 # one workload (merged into the file)
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle  # noqa: E402
from npge_b200 import synth  # noqa: E402

OUT = os.path.join(HERE, "digests.json")
BIG = 2 ** 62
MODES = {"parity": (False, BIG), "default": (True, 100000)}

# workload -> [(k, mode)]
PLAN = {
    "brucella3": [(20, "parity"), (20, "default")],
    "brucella17": [(20, "parity"), (20, "default")],
    "burkholderia": [(20, "parity"), (20, "default")],
    "yersinia": [(20, "parity"), (20, "default")],
    "pangenome1g": [(20, "parity"), (20, "default"), (24, "parity"), (28, "parity"),
                    (32, "parity")],
    # bench.py --gpus N: one 17-strain cluster per rank, sequences numbered in rank order
    "brucella17_x2": [(20, "parity"), (20, "default")],
    "brucella17_x4": [(20, "parity"), (20, "default")],
    "brucella17_x8": [(20, "parity"), (20, "default")],
}


def workload(name):
    """(names, seqs) of a digest workload; `<config>_x<N>` is the weak-scaling job
    of bench.py: the shards of ranks 0..N-1 one after the other."""
    if "_x" in name:
        base, n = name.rsplit("_x", 1)
        names, seqs = [], []
        for r in range(int(n)):
            nm, sq = synth.generate_shard(base, r, int(n))
            names += nm
            seqs += sq
        return names, seqs
    return synth.generate(name)


def key(name, k, mode):
    return "%s|k=%d|%s" % (name, k, mode)


def main():
    want = sys.argv[1:] or list(PLAN)
    table = {}
    if os.path.exists(OUT):
        with open(OUT) as f:
            table = json.load(f)
    for name in want:
        t0 = time.time()
        names, seqs = workload(name)
        bo = oracle.concat(seqs)
        n_bases = int(bo[1][-1])
        print("%s: %d sequences, %d bases (%.1f s)" % (name, len(seqs), n_bases,
                                                      time.time() - t0), flush=True)
        bits = 4 if n_bases < 2e8 else 6
        for k, mode in PLAN[name]:
            similar, limit = MODES[mode]
            r = oracle.find_ideal_sliced(bo, k, similar=similar, max_frags=limit,
                                         slice_bits=bits, threads=0)
            table[key(name, k, mode)] = {
                "workload": name, "anchor_size": k, "mode": mode,
                "anchor_similar": similar,
                "max_anchor_fragments": limit if limit < BIG else None,
                "n_sequences": len(seqs), "n_bases": n_bases,
                "n_fragments": r["n_fragments"], "n_blocks": r["n_blocks"],
                "digest": ["%016x" % r["digest"][0], "%016x" % r["digest"][1]],
                "oracle": "oracle_find_ideal_sliced, %d slices" % (1 << bits),
            }
            print("  k=%d %s: %d fragments, %d blocks, %.1f s"
                  % (k, mode, r["n_fragments"], r["n_blocks"], r["seconds"]), flush=True)
            with open(OUT, "w") as f:
                json.dump(table, f, indent=1, sort_keys=True)
                f.write("\n")


if __name__ == "__main__":
    main()
