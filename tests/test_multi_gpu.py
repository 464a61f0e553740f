"""AnchorFinder over several GPUs of ONE process (npge_anchor_multi_*, the host behind
npge::gpu::AnchorFinder::set_devices): G ranks, position ranges with halo, exchange
through peer pointers -- the joined result equals the oracle's, bit for bit.  With one
GPU in the box the ranks share it (device ids repeat)."""
import numpy as np
import pytest

import oracle
from npge_b200 import synth
from npge_b200.engine import MultiEngine
from test_gpu_parity import same

pytestmark = pytest.mark.gpu
BIG = 2 ** 62


def devices(n):
    import torch
    g = torch.cuda.device_count()
    return [r % g for r in range(n)]


@pytest.mark.parametrize("ranks", [1, 2, 3])
@pytest.mark.parametrize("k", [12, 20, 27])
def test_multi_equals_oracle(ranks, k):
    names, seqs = synth.generate("tiny")
    seqs = seqs + [seqs[0][:k - 1], seqs[1][100:700], seqs[2][:0]]
    m = MultiEngine(devices(ranks))
    try:
        base = oracle.find_ideal(seqs, k, similar=False)
        used = base.block_hash[::3].copy()
        for similar, limit, u in ((False, BIG, None), (True, BIG, None), (True, 1000, None),
                                  (False, 777, None), (True, BIG, used), (False, 9, None)):
            got = m.find(seqs, k, similar=similar, max_frags=limit, used=u)
            same(got, oracle.find_ideal(seqs, k, similar=similar, max_frags=limit, used=u))
        assert m.debug_check() == 0
    finally:
        m.close()


def test_one_chromosome_over_four_ranks():
    """Positions are sharded, not sequences: a single sequence over four ranks."""
    names, parts = synth.generate("small")
    seqs = [np.concatenate(parts)]
    m = MultiEngine(devices(4))
    try:
        for similar, limit in ((False, BIG), (True, 100000)):
            got = m.find(seqs, 20, similar=similar, max_frags=limit)
            same(got, oracle.find_ideal(seqs, 20, similar=similar, max_frags=limit))
        # resident genome: a second find on the same upload, with used hashes
        m.upload(seqs, 20)
        st = m.find_device(20, similar=False, max_frags=BIG)
        want = oracle.find_ideal(seqs, 20, similar=False)
        assert (st["n_fragments"], st["n_blocks"], st["digest"]) == (
            want.n_fragments, want.n_blocks, want.digest)
        assert sum(m.rank_stats(r)["n_fragments"] for r in range(4)) == want.n_fragments
        assert m.debug_check() == 0
    finally:
        m.close()


def test_multi_c1_brucella3_two_ranks():
    names, seqs = synth.generate("brucella3")
    m = MultiEngine(devices(2))
    try:
        m.upload(seqs, 20)
        st = m.find_device(20, similar=False, max_frags=BIG)
        want = oracle.find_ideal(seqs, 20, similar=False)
        assert (st["n_fragments"], st["n_blocks"], st["digest"]) == (
            want.n_fragments, want.n_blocks, want.digest)
        got = m.fetch()
        assert np.array_equal(got.block_hash, want.block_hash)
        assert np.array_equal(got.block_offsets, want.block_offsets)
        assert np.array_equal(got.new_used, want.new_used)
        assert m.debug_check() == 0
    finally:
        m.close()


@pytest.mark.parametrize("ranks", [1, 2, 3])
def test_multi_compact_layout_equals_oracle(ranks):
    """npge_anchor_multi_set_result_layout(COMPACT): every rank writes 4-byte
    (job-wide position, strand) payloads and multi_fetch_compact joins the slices;
    decoded with the layout table it is the oracle's result, also through the cut,
    with used hashes, and through the expanding npge_anchor_multi_fetch."""
    from npge_b200 import _abi
    names, seqs = synth.generate("tiny")
    seqs = seqs + [seqs[0][:19], seqs[1][100:700], seqs[2][:0]]
    k = 20
    m = MultiEngine(devices(ranks))
    try:
        base = oracle.find_ideal(seqs, k, similar=False)
        used = base.block_hash[::3].copy()
        m.set_result_layout(_abi.LAYOUT_COMPACT)
        m.upload(seqs, k)
        for similar, limit, u in ((False, BIG, None), (True, BIG, None), (True, 1000, None),
                                  (False, 777, None), (True, BIG, used), (False, 9, None)):
            want = oracle.find_ideal(seqs, k, similar=similar, max_frags=limit, used=u)
            st = m.find_device(k, similar=similar, max_frags=limit, used=u)
            assert st["n_fragments"] == want.n_fragments
            same(m.fetch_compact(), want)
            same(m.fetch(), want)  # the host expands the payloads
        m.set_result_layout(_abi.LAYOUT_FULL)
        m.find_device(k, similar=False, max_frags=BIG)
        with pytest.raises(Exception):
            m.fetch_compact()  # the resident result holds no payloads
        assert m.debug_check() == 0
    finally:
        m.close()
