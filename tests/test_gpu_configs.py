"""GPU parity at BASELINE.json's full sizes (last row):
  C1 brucella3, C2 brucella17   tuple for tuple against oracle.find_ideal
  C3 burkholderia, C4 yersinia, C5 pangenome1g (k = 20 and 32)
                                counts + 128-bit digest against
                                tests/golden/digests.json, which the CPU oracle
                                wrote in hash-range slices (make_digests.py)
through the C ABI, in parity mode and with the default options."""
import numpy as np
import pytest

import oracle
from golden_util import load
from npge_b200 import synth
from npge_b200.engine import AnchorEngine

pytestmark = pytest.mark.gpu

BIG = 2 ** 62
DIGESTS = load("digests.json")
MODES = {"parity": (False, BIG), "default": (True, 100000)}


@pytest.fixture(scope="module")
def eng():
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


def golden(workload, k, mode):
    g = DIGESTS["%s|k=%d|%s" % (workload, k, mode)]
    return g["n_fragments"], g["n_blocks"], tuple(int(x, 16) for x in g["digest"])


def canonical(r):
    """(block index, seq, min_pos, ori) of every fragment, ordered inside a block
    by (seq, min_pos, ori): the oracle lists a block by (caller index, pos'), the
    device in scan order."""
    sizes = np.diff(r.block_offsets.astype(np.int64))
    blk = np.repeat(np.arange(len(sizes), dtype=np.int64), sizes)
    t = ((r.frag_seq.astype(np.uint64) << np.uint64(33)) |
         (r.frag_min_pos.astype(np.uint64) << np.uint64(1)) |
         (r.frag_ori == 1).astype(np.uint64))
    order = np.lexsort((t, blk))
    return blk[order], t[order]


def test_c2_brucella17_tuple_for_tuple(eng):
    """BASELINE.json configs[1], the benchmarked workload: every one of the
    50.6 M fragments equal to the oracle's, block by block."""
    names, seqs = synth.generate("brucella17")
    got = eng.find(seqs, 20, similar=False, max_frags=BIG)
    want = oracle.find_ideal(seqs, 20, similar=False)
    assert got.n_fragments == want.n_fragments and got.n_blocks == want.n_blocks
    assert np.array_equal(got.block_hash, want.block_hash)
    assert np.array_equal(got.block_offsets, want.block_offsets)
    gb, gt = canonical(got)
    wb, wt = canonical(want)
    assert np.array_equal(gb, wb) and np.array_equal(gt, wt)
    assert got.digest == want.digest
    assert np.array_equal(got.new_used, want.new_used)
    nf, nb, dg = golden("brucella17", 20, "parity")
    assert (got.n_fragments, got.n_blocks, got.digest) == (nf, nb, dg)
    # default options: cut at 100000 fragments inside a block
    got = eng.find(seqs, 20, similar=True, max_frags=100000)
    want = oracle.find_ideal(seqs, 20, similar=True, max_frags=100000)
    assert got.blocks() == want.blocks()
    assert np.array_equal(got.new_used, want.new_used)
    assert (got.n_fragments, got.n_blocks, got.digest) == golden("brucella17", 20, "default")


@pytest.mark.parametrize("workload", ["brucella3", "burkholderia", "yersinia"])
def test_digest_c1_c3_c4(eng, workload):
    names, seqs = synth.generate(workload)
    eng.upload(seqs, 20)
    for mode, (similar, limit) in MODES.items():
        st = eng.find_device(20, similar=similar, max_frags=limit)
        assert (st["n_fragments"], st["n_blocks"], st["digest"]) == golden(workload, 20, mode), mode
    # the host-buffer call gives the same answer as the resident one
    r = eng.find_counts(seqs, 20, similar=True, max_frags=100000)
    assert (r["n_fragments"], r["n_blocks"], r["digest"]) == golden(workload, 20, "default")


def test_digest_c5_pangenome1g(eng):
    """The north-star configuration: 1 Gbp, 600 sequences, on one GPU."""
    names, seqs = synth.generate("pangenome1g")
    bo = oracle.concat(seqs)
    del seqs
    for k, modes in ((20, ("parity", "default")), (32, ("parity",))):
        eng.upload(bo, k)
        for mode in modes:
            similar, limit = MODES[mode]
            st = eng.find_device(k, similar=similar, max_frags=limit)
            assert (st["n_fragments"], st["n_blocks"], st["digest"]) == golden(
                "pangenome1g", k, mode), (k, mode)
