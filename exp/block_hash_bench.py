"""Timing of the block_hash row: the reference's block_hash of
every anchor block of brucella17 (exact mode: 3.4 M blocks, 50.6 M fragments) on
the device-resident result, next to the oracle's string-based restatement on one
host core (bounded sample).  Checks the sample bit for bit."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "oracle")
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, concat
import oracle

config = sys.argv[1] if len(sys.argv) > 1 else "brucella17"
names, seqs = synth.generate(config)
eng = AnchorEngine(0)
eng.upload(concat(seqs), 20)
out = {"config": config}
for mode, similar in (("exact", False), ("anchor-similar", True)):
    st = eng.find_device(20, similar=similar, max_frags=2 ** 62)
    best = (1e9, 1e9)
    for it in range(4):
        t0 = time.perf_counter()
        hashes, total = eng.anchor_block_hashes(names)
        wall = time.perf_counter() - t0
        best = min(best, (eng.block_hashes_ms(), 1e3 * wall))
    r = eng.fetch()
    nb, nf = r.n_blocks, r.n_fragments
    sample = min(nb, 100000)
    offs = np.asarray(r.block_offsets[:sample + 1], np.uint64)
    n = int(offs[-1])
    mn = np.asarray(r.frag_min_pos[:n], np.uint32)
    t0 = time.perf_counter()
    want = oracle.block_hashes(names, offs, r.frag_seq[:n], mn, mn + np.uint32(19),
                               r.frag_ori[:n])
    cpu = time.perf_counter() - t0
    assert np.array_equal(hashes[:sample], want)
    out[mode] = {"blocks": nb, "fragments": nf, "device_ms": best[0], "call_ms": best[1],
                 "blocks_per_s": nb / (best[0] * 1e-3),
                 "fragments_per_s": nf / (best[0] * 1e-3),
                 "blockset_hash": "%016x" % total,
                 "cpu_oracle_blocks_per_s_1_thread": sample / cpu,
                 "cpu_sample_blocks": sample}
print(json.dumps(out))
