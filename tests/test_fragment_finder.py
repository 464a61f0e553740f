"""FragmentFinder: the oracle's restatement of
src/algo/FragmentFinder.cpp on CPU, and the CUDA path through the C ABI against
it (bit-exact, same order)."""
import random

import numpy as np
import pytest

import oracle
from npge_b200 import synth


def revcomp(s):
    return s[::-1].translate(str.maketrans("ATGC", "TACG"))


def brute(seqs, pattern, max_matches):
    """Independent statement of the semantics: text == P (+1), revcomp(text) == P (-1)."""
    L = len(pattern)
    order = sorted((i for i in range(len(seqs)) if len(seqs[i]) >= L),
                   key=lambda i: -len(seqs[i]))
    out = []
    for i in order:
        s = seqs[i]
        for p in range(len(s) - L + 1):
            w = s[p:p + L]
            if w == pattern:
                out.append((i, p, 1))
            if revcomp(w) == pattern:
                out.append((i, p, -1))
    return out[:max_matches]


def random_case(rng, alphabet="ATGC"):
    seqs = ["".join(rng.choice(alphabet) for _ in range(rng.randint(0, 300)))
            for _ in range(rng.randint(1, 6))]
    src = rng.choice([s for s in seqs if s] or ["ACGT"])
    L = rng.randint(1, min(70, len(src)))
    a = rng.randint(0, len(src) - L)
    pattern = src[a:a + L]
    if rng.random() < 0.3:
        pattern = revcomp(pattern)
    return seqs, pattern


# ---- oracle (no GPU) ------------------------------------------------------------------
def test_oracle_matches_brute_force():
    rng = random.Random(5)
    for trial in range(200):
        seqs, pattern = random_case(rng, "ATGC" if trial % 3 else "ATGCN")
        mm = rng.choice([1, 2, 5, 1000])
        assert oracle.fragment_find(seqs, pattern, mm) == brute(seqs, pattern, mm)


def test_oracle_documented_cases():
    s = oracle.to_atgcn("tgGTCCGagCGGACggcc")
    assert oracle.fragment_find([s], "GTCCG") == [(0, 2, 1), (0, 9, -1)]
    assert oracle.fragment_find(["ATATAT"], "AT") == [
        (0, 0, 1), (0, 0, -1), (0, 2, 1), (0, 2, -1), (0, 4, 1), (0, 4, -1)]
    assert oracle.fragment_find(["ACGT"], "ACGTA") == []       # longer than the sequence
    assert oracle.fragment_find(["AANTT", "ANT"], "ANT") == [  # N matches only N
        (0, 1, 1), (0, 1, -1), (1, 0, 1), (1, 0, -1)]
    with pytest.raises(ValueError):
        oracle.fragment_find(["ACGT"], "", 10)
    with pytest.raises(ValueError):
        oracle.fragment_find(["ACGT"], "A", 0)


# ---- CUDA path ------------------------------------------------------------------------
@pytest.fixture(scope="module")
def eng():
    from npge_b200.engine import AnchorEngine
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


@pytest.mark.gpu
def test_gpu_random_patterns(eng):
    rng = random.Random(11)
    for trial in range(150):
        seqs, pattern = random_case(rng, "ATGC" if trial % 3 else "ATGCN")
        mm = rng.choice([1, 3, 1000])
        got, total = eng.find_fragments(seqs, pattern, mm, with_total=True)
        assert got == oracle.fragment_find(seqs, pattern, mm), (seqs, pattern, mm)
        assert total == len(brute(seqs, pattern, 10**9))


@pytest.mark.gpu
def test_gpu_boundaries_and_errors(eng):
    from npge_b200.engine import AnchorError
    # a pattern with N must not match across the pad between two sequences
    assert eng.find_fragments(["CCA", "TGG"], "ANT") == []
    assert eng.find_fragments(["CCANTGG"], "ANT") == [(0, 2, 1), (0, 2, -1)]
    assert eng.find_fragments(["ACGT"], "ACGTA") == []
    assert eng.find_fragments([], "ACGT") == []
    with pytest.raises(AnchorError):
        eng.find_fragments(["ACGT"], "")
    with pytest.raises(AnchorError):
        eng.find_fragments(["ACGT"], "A", 0)
    with pytest.raises(AnchorError):
        eng.find_fragments(["ACGT"], "AXG")


@pytest.mark.gpu
@pytest.mark.parametrize("L", [12, 32, 33, 64, 200, 1500])
def test_gpu_genome_scale(eng, L):
    names, seqs = synth.generate("small")
    text = bytes(seqs[1][5000:5000 + L]).decode()
    texts = [bytes(s).decode() for s in seqs]
    want = oracle.fragment_find(texts, text, 100)
    assert len(want) >= 1
    eng.upload(seqs, 12)
    assert eng.find_fragments_device(text, 100) == want
    rc = revcomp(text)
    assert eng.find_fragments_device(rc, 100) == oracle.fragment_find(texts, rc, 100)
    # an IS-like repeat: many matches, the cut keeps the first ones in scan order
    assert eng.find_fragments_device(text[:12], 7) == oracle.fragment_find(texts, text[:12], 7)
