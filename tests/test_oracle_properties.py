"""Invariants between the two CPU oracles, and the oracle's hash
arithmetic against the reference's own code (oracle/_ref).  CPU only."""
import ctypes as C
import random

import numpy as np
import pytest

import oracle
from npge_b200 import synth


def _random_seqs(rng, n_seqs, length, with_n=True, repeats=True):
    seqs = []
    pool = ["".join(rng.choice("ATGC") for _ in range(40)) for _ in range(6)]
    for _ in range(n_seqs):
        n = rng.randint(max(1, length // 2), length)
        s = [rng.choice("ATGC") for _ in range(n)]
        if repeats:
            for _ in range(rng.randint(0, 6)):
                piece = rng.choice(pool)
                if rng.random() < 0.5:
                    piece = piece[::-1].translate(str.maketrans("ATGC", "TACG"))
                p = rng.randint(0, max(0, n - len(piece)))
                s[p:p + len(piece)] = piece
        if with_n:
            for _ in range(rng.randint(0, 3)):
                p = rng.randint(0, n - 1)
                for j in range(p, min(n, p + rng.randint(1, 12))):
                    s[j] = "N"
        seqs.append("".join(s[:n]))
    return seqs


@pytest.mark.parametrize("k", [1, 3, 8, 16, 20, 31, 32])
def test_faithful_parity_mode_equals_ideal(k):
    rng = random.Random(100 + k)
    for trial in range(6):
        seqs = _random_seqs(rng, rng.randint(1, 5), 400)
        want = oracle.find_ideal(seqs, k, similar=False)
        for workers, seed in ((1, 1), (1, 99), (3, 5)):
            got = oracle.find_faithful(seqs, k, similar=False, max_frags=2**40,
                                       workers=workers, seed=seed)
            if workers > 1:
                # the reference's documented race (AnchorFinder.hpp:24-26):
                # two workers can both see a k-mer "first", losing the anchor
                assert set(got.blocks()) <= set(want.blocks())
                continue
            assert got.blocks() == want.blocks()
            assert got.digest == want.digest
            assert list(got.block_hash) == list(want.block_hash)
            # Bloom false positives add singleton candidates only
            assert set(want.new_used) <= set(got.new_used)


@pytest.mark.parametrize("k", [3, 12, 20])
def test_faithful_default_is_subset_of_parity(k):
    rng = random.Random(200 + k)
    for trial in range(6):
        seqs = _random_seqs(rng, rng.randint(1, 5), 500)
        all_blocks = set(oracle.find_ideal(seqs, k, similar=False).blocks())
        for seed in (1, 2, 3):
            got = oracle.find_faithful(seqs, k, similar=True, max_frags=2**40,
                                       seed=seed)
            assert set(got.blocks()) <= all_blocks


@pytest.mark.parametrize("k", [2, 5, 12, 20, 32])
def test_exact_filter_faithful_equals_ideal_similar(k):
    rng = random.Random(300 + k)
    for trial in range(8):
        seqs = _random_seqs(rng, rng.randint(1, 5), 500)
        used = []
        if trial % 2:
            full = oracle.find_ideal(seqs, k, similar=False)
            used = list(full.block_hash[::2])
        for max_frags in (2**40, 7, 2, 0):
            want = oracle.find_ideal(seqs, k, similar=True,
                                     max_frags=max_frags, used=used)
            got = oracle.find_faithful(seqs, k, similar=True, workers=1,
                                       max_frags=max_frags, exact_bloom=True,
                                       used=used)
            assert got.blocks() == want.blocks()
            assert list(got.new_used) == list(want.new_used)
            assert not set(used) & set(int(h) for h in want.block_hash)


def test_truncation_is_prefix_of_sorted_order():
    rng = random.Random(7)
    seqs = _random_seqs(rng, 3, 600, with_n=False)
    full = oracle.find_ideal(seqs, 4, similar=False)
    nf = full.n_fragments
    assert nf > 50
    for m in (0, 1, 2, 3, 10, nf - 1, nf, nf + 5):
        cut = oracle.find_ideal(seqs, 4, similar=False, max_frags=m)
        assert cut.n_fragments <= m
        assert set(cut.blocks()) <= set(
            tuple(b) for b in _prefix_blocks(full, m))


def _prefix_blocks(full, m):
    """blocks formed by the first m records of (h, seq, pos') order, given the
    full parity-mode result whose fragments are already in that order"""
    out, taken = [], 0
    for b in range(full.n_blocks):
        lo, hi = int(full.block_offsets[b]), int(full.block_offsets[b + 1])
        take = min(hi - lo, m - taken)
        taken += max(take, 0)
        if take >= 2:
            out.append(tuple(sorted(
                (int(full.frag_seq[f]), int(full.frag_min_pos[f]),
                 int(full.frag_ori[f])) for f in range(lo, lo + take))))
        if taken >= m:
            break
    return out


def test_used_hashes_across_runs():
    # F7: hashes recorded by run 1 (all but the first) are invisible to run 2
    rng = random.Random(11)
    seqs = _random_seqs(rng, 3, 500)
    r1 = oracle.find_ideal(seqs, 6, similar=False)
    assert r1.n_blocks > 3
    assert list(r1.new_used) == list(r1.block_hash[1:])
    r2 = oracle.find_ideal(seqs, 6, similar=False, used=r1.new_used)
    assert list(r2.block_hash) == [r1.block_hash[0]]
    assert len(r2.new_used) == 0


def test_soundness_on_synthetic():
    # src/test/test_repeats.sh.in: every anchor occurs >= 2x in text + revcomp
    names, seqs = synth.generate("tiny")
    res = oracle.find_ideal(seqs, 15, similar=True, max_frags=100000)
    assert res.n_blocks > 0
    text = "#".join(bytes(s).decode() for s in seqs)
    both = text + "#" + text[::-1].translate(str.maketrans("ATGC", "TACG"))
    for b in range(0, res.n_blocks, max(1, res.n_blocks // 50)):
        f = int(res.block_offsets[b])
        s, p = int(res.frag_seq[f]), int(res.frag_min_pos[f])
        word = bytes(seqs[s][p:p + 15]).decode()
        assert both.count(word) >= 2


def test_oracle_hash_matches_reference_code():
    R = oracle.ref_lib()
    if R is None:
        pytest.skip("oracle/_ref not available")
    L = oracle.lib()
    rng = random.Random(5)
    for _ in range(2000):
        n = rng.randint(1, 99)
        s = "".join(rng.choice("ATGCN" if rng.random() < 0.2 else "ATGC")
                    for _ in range(n + 1)).encode()
        buf = C.create_string_buffer(s, len(s))
        last = C.cast(C.addressof(buf) + n - 1, C.c_char_p)
        d = R.ref_make_hash(buf, n, 1)
        assert L.oracle_make_hash(buf, n, 1) == d
        assert L.oracle_make_hash(last, n, -1) == R.ref_make_hash(last, n, -1)
        assert L.oracle_complement_hash(d, n) == R.ref_complement_hash(d, n)
        for fwd in (0, 1):
            a, b = s[0:1], s[n:n + 1]
            assert (L.oracle_reuse_hash(d, n, a, b, fwd)
                    == R.ref_reuse_hash(d, n, a, b, fwd))


# ---- the sliced oracle behind tests/golden/digests.json ---------------------------------
@pytest.mark.parametrize("k", [1, 7, 20, 32])
def test_sliced_oracle_equals_find_ideal(k):
    """oracle_find_ideal_sliced (hash-range slices, threads) gives find_ideal's counts
    and digest in both modes, with cuts inside a group and with used hashes."""
    names, seqs = synth.generate("tiny")
    seqs = seqs + _random_seqs(random.Random(k), 5, 300)
    base = oracle.find_ideal(seqs, k, similar=False)
    used = base.block_hash[::4].copy()
    for similar, limit, u in ((False, 2**62, None), (True, 2**62, None), (True, 100000, None),
                              (False, 1001, None), (True, 3, None), (False, 0, None),
                              (False, 2**62, used), (True, 555, used)):
        want = oracle.find_ideal(seqs, k, similar=similar, max_frags=limit, used=u)
        for bits, threads in ((0, 1), (5, 3)):
            got = oracle.find_ideal_sliced(seqs, k, similar=similar, max_frags=limit, used=u,
                                           slice_bits=bits, threads=threads)
            assert got["n_fragments"] == want.n_fragments
            assert got["n_blocks"] == want.n_blocks
            assert got["digest"] == want.digest
            assert got["n_new_used"] == len(want.new_used)


def test_golden_digests_match_find_ideal_on_c1():
    """digests.json (written by tests/golden/make_digests.py) is what find_ideal says
    on the configuration small enough to run here: C1 brucella3."""
    from golden_util import load
    table = load("digests.json")
    names, seqs = synth.generate("brucella3")
    for mode, (similar, limit) in {"parity": (False, 2**62), "default": (True, 100000)}.items():
        g = table["brucella3|k=20|%s" % mode]
        want = oracle.find_ideal(seqs, 20, similar=similar, max_frags=limit)
        assert (g["n_fragments"], g["n_blocks"]) == (want.n_fragments, want.n_blocks)
        assert tuple(int(x, 16) for x in g["digest"]) == want.digest
