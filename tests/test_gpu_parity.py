"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle
(bit-exact: integer work).  Mirrors src/test/anchor_finder.cpp and
test-script/anchor_finder of the reference."""
import random

import numpy as np
import pytest

import oracle
from golden_util import expected_blocks, load, result_blocks4
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, AnchorError
from test_oracle_properties import _random_seqs

pytestmark = pytest.mark.gpu

CASES = load("anchor_finder_cases.json")
BIG = 2**62


@pytest.fixture(scope="module")
def eng():
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


def same(got, want, check_used=True):
    assert got.n_blocks == want.n_blocks
    assert got.n_fragments == want.n_fragments
    assert got.blocks() == want.blocks()
    assert list(got.block_hash) == list(want.block_hash)
    assert got.digest == want.digest
    if check_used:
        assert list(got.new_used) == list(want.new_used)


# ---- src/test/anchor_finder.cpp:18-110 -----------------------------------------
@pytest.mark.parametrize("similar", [True, False])
@pytest.mark.parametrize("case", CASES["unit_cases"], ids=lambda c: c["name"])
def test_unit_cases(eng, case, similar):
    seqs = [oracle.to_atgcn(s) for s in case["seqs"]]
    got = eng.find(seqs, case["k"], similar=similar, max_frags=100000)
    same(got, oracle.find_ideal(seqs, case["k"], similar=similar,
                                max_frags=100000))
    if "check_blocks" in case:
        assert got.n_blocks == case["check_blocks"]
    if "check_blocks_gt" in case:
        assert got.n_blocks > case["check_blocks_gt"]
    if similar:
        if "warn_blocks" in case:
            assert got.n_blocks == case["warn_blocks"]
        if "warn_front_size" in case:
            assert int(got.block_offsets[1]) == case["warn_front_size"]


def test_main_case_tuples(eng):
    got = eng.find(["TGGTCCGAGCGGACGGCC"], 5)
    assert got.blocks() == [((0, 2, 1), (0, 9, -1))]
    assert list(got.block_hash) == [0x2f6]


# ---- test-script/anchor_finder ----------------------------------------------------
@pytest.mark.parametrize("similar", [True, False])
@pytest.mark.parametrize("case", CASES["script_cases"], ids=lambda c: c["case"])
def test_script_cases(eng, case, similar):
    names, want = expected_blocks(case)
    seqs = [oracle.to_atgcn(s["text"]) for s in case["sequences"]]
    got = eng.find(seqs, 20, similar=similar)
    assert (oracle.blockset_hash(names, result_blocks4(got, 20))
            == oracle.blockset_hash(names, want))
    assert got.n_blocks == len(want)
    same(got, oracle.find_ideal(seqs, 20, similar=similar, max_frags=100000))


# ---- random inputs, every k class -------------------------------------------------
@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 8, 12, 13, 16, 17, 20, 24, 28, 31, 32])
def test_random_parity_and_similar(eng, k):
    rng = random.Random(1000 + k)
    for trial in range(4):
        seqs = _random_seqs(rng, rng.randint(1, 6), 700)
        for similar in (False, True):
            got = eng.find(seqs, k, similar=similar, max_frags=BIG)
            same(got, oracle.find_ideal(seqs, k, similar=similar))


@pytest.mark.parametrize("k", [3, 6, 20])
def test_truncation(eng, k):
    rng = random.Random(2000 + k)
    seqs = _random_seqs(rng, 4, 900, with_n=(k != 3))
    full = oracle.find_ideal(seqs, k, similar=False)
    nf = full.n_fragments
    assert nf > 10
    cuts = sorted(set([0, 1, 2, 3, 5, 8, nf // 3, nf // 2, nf - 2, nf - 1, nf,
                       nf + 1]))
    for similar in (False, True):
        for m in cuts:
            got = eng.find(seqs, k, similar=similar, max_frags=m)
            same(got, oracle.find_ideal(seqs, k, similar=similar, max_frags=m))


def test_truncation_inside_two_strand_group(eng):
    # one group with + and - members in the same sequence: the reference's
    # order is + strand first (pos' = pos + size for the reverse strand)
    a = "GATTACAGGCT"
    ra = a[::-1].translate(str.maketrans("ATGC", "TACG"))
    seqs = [a + "TT" + ra + "CC" + a + "AA" + ra, ra + "GG" + a]
    for m in range(0, 12):
        for similar in (False, True):
            got = eng.find(seqs, 11, similar=similar, max_frags=m)
            same(got, oracle.find_ideal(seqs, 11, similar=similar, max_frags=m))


@pytest.mark.parametrize("k", [4, 20])
def test_used_hashes(eng, k):
    rng = random.Random(3000 + k)
    seqs = _random_seqs(rng, 4, 900)
    for similar in (False, True):
        r1 = eng.find(seqs, k, similar=similar, max_frags=BIG)
        same(r1, oracle.find_ideal(seqs, k, similar=similar))
        used = np.sort(r1.new_used)
        r2 = eng.find(seqs, k, similar=similar, max_frags=BIG, used=used)
        same(r2, oracle.find_ideal(seqs, k, similar=similar, used=used))
        # a third run with everything used finds only the never-recorded first
        used3 = np.sort(np.concatenate([used, r2.new_used]))
        r3 = eng.find(seqs, k, similar=similar, max_frags=BIG, used=used3)
        same(r3, oracle.find_ideal(seqs, k, similar=similar, used=used3))


# ---- edge cases ---------------------------------------------------------------------
def test_edges(eng):
    for seqs, k in [([], 5), ([""], 5), (["ACG"], 5), (["NNNNNNNNNN"], 3),
                    (["ACGTA", "AC", "", "ACGTA"], 5), (["A" * 50], 1),
                    (["A" * 5000], 32), (["ACGT" * 2000], 7),
                    (["N" * 40 + "ACGTTGCA" + "N" * 40 + "ACGTTGCA"], 8)]:
        for similar in (False, True):
            got = eng.find(seqs, k, similar=similar, max_frags=BIG)
            same(got, oracle.find_ideal(seqs, k, similar=similar))


def test_sequence_boundaries_never_joined(eng):
    # a k-mer split across two sequences must not be found (no wrap)
    word = "ACGGTCATTGCAAGTC"
    seqs = [word[:8], word[8:], "TT" + word + "TT", word[8:] + word[:8]]
    got = eng.find(seqs, 16, similar=False, max_frags=BIG)
    same(got, oracle.find_ideal(seqs, 16, similar=False))
    assert got.n_blocks == 0


def test_option_rules(eng):
    for k in (0, -1, 33):
        with pytest.raises(AnchorError) as e:
            eng.find(["ACGTACGT"], k)
        assert e.value.code == 1
    # max-anchor-fragments has no rule: a negative value is compared as size_t in
    # the reference (AnchorFinder.cpp:361) and never cuts the list
    seqs = ["ACGTACGTTTGACCAACGTACGT", "GGACGTACGTTA"]
    same(eng.find(seqs, 4, similar=False, max_frags=-1),
         oracle.find_ideal(seqs, 4, similar=False, max_frags=-1))


def test_rejects_unnormalised_input(eng):
    with pytest.raises(AnchorError) as e:
        eng.find(["ACGTRYACGT-acgt"], 4)
    assert e.value.code == 5


# ---- synthetic genome sets -------------------------------------------------------------
@pytest.mark.parametrize("config,k", [("tiny", 15), ("tiny", 20), ("small", 20),
                                      ("small", 32), ("small", 11)])
def test_synthetic(eng, config, k):
    names, seqs = synth.generate(config)
    for similar, m in ((False, BIG), (True, 100000), (True, 1000), (False, 777)):
        got = eng.find(seqs, k, similar=similar, max_frags=m)
        same(got, oracle.find_ideal(seqs, k, similar=similar, max_frags=m))


def test_brucella3_full_size(eng):
    # BASELINE.json configs[0]: 3 Brucella-shaped genomes, 9.9 Mbp, k=20
    names, seqs = synth.generate("brucella3")
    want = oracle.find_ideal(seqs, 20, similar=False)
    got = eng.find(seqs, 20, similar=False, max_frags=BIG)
    assert got.n_fragments == want.n_fragments
    assert got.digest == want.digest
    assert np.array_equal(got.block_hash, want.block_hash)
    assert np.array_equal(got.block_offsets, want.block_offsets)
    # inside a block the oracle orders by (caller index, pos'), the device by
    # scan order: compare as sets
    assert got.blocks() == want.blocks()
    got = eng.find(seqs, 20, similar=True, max_frags=100000)
    same(got, oracle.find_ideal(seqs, 20, similar=True, max_frags=100000))


def test_device_resident_variant(eng):
    names, seqs = synth.generate("small")
    eng.upload(seqs, 20)
    st = eng.find_device(20, similar=False, max_frags=BIG)
    want = oracle.find_ideal(seqs, 20, similar=False)
    assert st["n_fragments"] == want.n_fragments
    assert st["n_blocks"] == want.n_blocks
    assert st["digest"] == want.digest
    assert st["gpu_launches"] >= 3 + st["radix_passes"]
    same(eng.fetch(), want)
    # a second run on the resident genome gives the same answer
    st2 = eng.find_device(20, similar=False, max_frags=BIG)
    assert st2["digest"] == want.digest


# ---- the digit plans of bigger jobs (512- and 1024-bin partition passes) -----------
@pytest.mark.parametrize("f,wa", [(17, None), (18, None), (19, None), (20, None), (16, 9), (18, 8)])
def test_wide_digit_plans(eng, f, wa):
    """npge_anchor_debug_plan forces the plan a 100 M - 1 G position job would
    get (wa, wb in 8..10), on an input small enough for the oracle."""
    eng.debug_plan(f, wa or 0)
    try:
        names, seqs = synth.generate("small")
        for k, similar, limit in ((20, False, BIG), (20, True, 100000), (27, False, BIG)):
            got = eng.find(seqs, k, similar=similar, max_frags=limit)
            same(got, oracle.find_ideal(seqs, k, similar=similar, max_frags=limit))
    finally:
        eng.debug_plan(0, 0)
