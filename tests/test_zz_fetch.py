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
    assert eng.debug_check() == 0   # no kernel stored past the end of a device buffer
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


# ---- the host expansion without a device ------------------------------------------------
def test_expand_payloads_on_the_host(tmp_path):
    """fetch_expand.h against a numpy statement of the same mapping, with
    sequences shorter than a lookup chunk and pieces that start off zero."""
    import ctypes as C
    import os
    import subprocess
    from conftest import ROOT
    so = str(tmp_path / "libfetch_expand.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-fPIC",
                           "-shared", "-o", so,
                           os.path.join(ROOT, "tests", "fetch_expand_host.cpp")])
    lib = C.CDLL(so)
    lib.fetch_expand_host.argtypes = ([C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int] +
                                      [C.c_void_p] * 5)
    rng = np.random.default_rng(9)
    shift = 12
    sizes = np.concatenate([rng.integers(1, 300, 40), rng.integers(3000, 20000, 12),
                            rng.integers(1, 50, 30)])
    sizes = np.sort(sizes)[::-1]                       # scan order: size descending
    start = np.zeros(len(sizes) + 1, np.uint32)
    start[1:] = np.cumsum(sizes + 1)                   # one pad after every sequence
    caller = rng.permutation(len(sizes)).astype(np.uint32)
    n_chunks = (int(start[-1]) >> shift) + 2
    lut = (np.searchsorted(start[:-1], np.arange(n_chunks, dtype=np.uint64) << shift,
                           side="right") - 1).astype(np.uint32)
    n = 100003
    s_true = rng.integers(0, len(sizes), n)
    p_true = (rng.random(n) * sizes[s_true]).astype(np.uint32)
    bit = rng.integers(0, 2, n).astype(np.uint32)
    pay = (((start[s_true] + p_true) << 1) | bit).astype(np.uint32)
    for lo, hi in ((0, n), (8, 8), (16, 23), (4096, n - 5)):
        seq = np.full(n, 0xFFFFFFFF, np.uint32)
        pos = np.full(n, 0xFFFFFFFF, np.uint32)
        ori = np.zeros(n, np.int8)
        lib.fetch_expand_host(pay.ctypes.data, lo, hi, lut.ctypes.data, shift, start.ctypes.data,
                              caller.ctypes.data, seq.ctypes.data, pos.ctypes.data,
                              ori.ctypes.data)
        assert np.array_equal(seq[lo:hi], caller[s_true[lo:hi]])
        assert np.array_equal(pos[lo:hi], p_true[lo:hi])
        assert np.array_equal(ori[lo:hi], np.where(bit[lo:hi] == 1, 1, -1))
        assert np.all(ori[:lo] == 0) and np.all(ori[hi:] == 0)      # nothing outside [lo, hi)
