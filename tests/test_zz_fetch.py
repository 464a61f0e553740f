"""npge_anchor_fetch: the compact transfer (4-byte payloads over PCIe, expanded
to (sequence, position, strand) by host threads) returns exactly what the plain
copies of the device arrays return."""
import numpy as np
import pytest

import oracle
from npge_b200 import synth


def fetch_with(engine, monkeypatch, mode):
    monkeypatch.setenv("NPGE_FETCH", mode)
    return engine.fetch()


def same(a, b):
    return (np.array_equal(a.block_offsets, b.block_offsets) and
            np.array_equal(a.block_hash, b.block_hash) and
            np.array_equal(a.frag_seq, b.frag_seq) and
            np.array_equal(a.frag_min_pos, b.frag_min_pos) and
            np.array_equal(a.frag_ori, b.frag_ori) and
            np.array_equal(a.new_used, b.new_used) and a.digest == b.digest)


@pytest.fixture(scope="module")
def engine():
    from npge_b200.engine import AnchorEngine
    eng = AnchorEngine(0)
    yield eng
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("config,k", [("tiny", 20), ("tiny", 7), ("small", 20), ("brucella3", 20)])
def test_compact_fetch_equals_plain_fetch(engine, monkeypatch, config, k):
    names, seqs = synth.generate(config)
    engine.upload(seqs, k)
    for similar in (False, True):
        monkeypatch.setenv("NPGE_FETCH", "compact")   # the emit kernel leaves the payloads
        engine.find_device(k, similar=similar, max_frags=2 ** 62)
        plain = fetch_with(engine, monkeypatch, "plain")
        compact = fetch_with(engine, monkeypatch, "compact")
        assert plain.n_fragments > 0
        assert same(plain, compact)
        assert same(plain, fetch_with(engine, monkeypatch, "auto"))   # by size
        monkeypatch.delenv("NPGE_FETCH")
        engine.find_device(k, similar=similar, max_frags=2 ** 62)
        assert same(plain, engine.fetch())      # the default


@pytest.mark.gpu
def test_compact_fetch_against_the_oracle_with_short_sequences(engine, monkeypatch):
    # many sequences shorter than one lookup chunk (4096 positions): the host
    # expansion walks over several sequence starts inside one chunk
    rng = np.random.default_rng(5)
    base = "".join(rng.choice(list("ATGC"), 700))
    seqs = []
    for i in range(60):
        s = list(base[int(rng.integers(0, 80)):int(rng.integers(300, 700))])
        for j in rng.integers(0, len(s), 6):
            s[j] = "ATGCN"[int(rng.integers(0, 5))]
        seqs.append("".join(s))
    want = oracle.find_ideal(seqs, 12, similar=False, max_frags=2 ** 62)
    engine.upload(seqs, 12)
    monkeypatch.setenv("NPGE_FETCH", "compact")
    engine.find_device(12, similar=False, max_frags=2 ** 62)
    got = fetch_with(engine, monkeypatch, "compact")
    assert got.blocks() == want.blocks()
    assert got.digest == want.digest
    # a cut list always takes the plain path (the cut group is rebuilt on the host)
    engine.find_device(12, similar=False, max_frags=50)
    cut = fetch_with(engine, monkeypatch, "compact")
    want_cut = oracle.find_ideal(seqs, 12, similar=False, max_frags=50)
    assert cut.blocks() == want_cut.blocks()
