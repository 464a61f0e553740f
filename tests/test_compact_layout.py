"""The compact result layout (npge_anchor_set_result_layout / fetch_compact /
find_compact): 4 bytes per fragment, decoded with the layout table, equals the full
result and the oracle -- also when max-anchor-fragments cuts inside a block."""
import numpy as np
import pytest

import oracle
from npge_b200 import _abi, synth
from npge_b200.engine import AnchorEngine, AnchorError
from test_gpu_parity import same

pytestmark = pytest.mark.gpu
BIG = 2 ** 62


@pytest.fixture(scope="module")
def eng():
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


@pytest.mark.parametrize("config,k", [("tiny", 12), ("tiny", 20), ("small", 20), ("small", 27)])
def test_find_compact_equals_oracle(eng, config, k):
    names, seqs = synth.generate(config)
    seqs = seqs + [seqs[0][:k - 1], seqs[1][100:700], seqs[2][:0]]
    base = oracle.find_ideal(seqs, k, similar=False)
    used = base.block_hash[::3].copy()
    for similar, limit, u in ((False, BIG, None), (True, 100000, None), (False, 777, None),
                              (True, 9, None), (False, 8, None), (True, BIG, used)):
        got = eng.find_compact(seqs, k, similar=similar, max_frags=limit, used=u)
        same(got, oracle.find_ideal(seqs, k, similar=similar, max_frags=limit, used=u))
    # the context's own layout is untouched by find_compact
    same(eng.find(seqs, k, similar=False, max_frags=BIG), base)


def test_layout_switch_on_a_resident_genome(eng):
    names, seqs = synth.generate("small")
    want = oracle.find_ideal(seqs, 20, similar=False)
    eng.upload(seqs, 20)
    eng.set_result_layout(_abi.LAYOUT_COMPACT)
    try:
        st = eng.find_device(20, similar=False, max_frags=BIG)
        assert (st["n_fragments"], st["n_blocks"], st["digest"]) == (
            want.n_fragments, want.n_blocks, want.digest)
        assert st["bytes_moved"] < 10 ** 12
        same(eng.fetch_compact(), want)
        same(eng.fetch(), want)            # the full arrays, expanded on the host
        with pytest.raises(AnchorError) as e:
            eng.anchor_block_hashes(names)  # needs the full arrays on the device
        assert e.value.code == _abi.ERR_STATE
    finally:
        eng.set_result_layout(_abi.LAYOUT_FULL)
    eng.find_device(20, similar=False, max_frags=BIG)
    with pytest.raises(AnchorError) as e:   # no payloads in the full layout
        eng.fetch_compact()
    assert e.value.code == _abi.ERR_STATE
    same(eng.fetch(), want)


def test_compact_c1_brucella3(eng):
    names, seqs = synth.generate("brucella3")
    want = oracle.find_ideal(seqs, 20, similar=False)
    got = eng.find_compact(seqs, 20, similar=False, max_frags=BIG)
    assert got.digest == want.digest and got.n_fragments == want.n_fragments
    assert np.array_equal(got.block_hash, want.block_hash)
    assert np.array_equal(got.block_offsets, want.block_offsets)
    assert got.blocks() == want.blocks()
