"""block_hash / blockset_hash: the reference's block
fingerprint (src/model/block_hash.cpp:29-54,125-130) that MoveUnchanged
(src/algo/MoveUnchanged.cpp:37-49) compares between the iterations of AnchorLoop.

CPU: the oracle against an independent Python statement of the definition, and
the code of the device kernel (npge_b200/csrc/block_hash_core.h, compiled for
the host by tests/block_hash_host.cpp) against the oracle.
GPU: the kernel through the C ABI against the oracle, bit for bit."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle
from conftest import ROOT
from npge_b200 import synth

NAMES = ["A&chr1&c", "A&chr2&c", "B&chr1&c", "s1", "s10", "s2", "x",
         "BRU17&chr10&c", "z_", "BRU17&chr1&c"]


def py_fragment_id(name, mn, mx, ori):
    a, b = (mn, mx) if ori == 1 else (mx, mn)
    if a == b and ori == -1:
        b = -1
    return "%s_%d_%d" % (name, a, b)


def py_block_hash(names, frags):
    """frags: [(seq, min, max, ori)]; block_hash.cpp:29-54 word for word."""
    ids_dir = sorted(py_fragment_id(names[s], a, z, o).encode() for s, a, z, o in frags)
    ids_inv = sorted(py_fragment_id(names[s], a, z, -o).encode() for s, a, z, o in frags)
    joint = b" ".join(ids_dir if ids_dir < ids_inv else ids_inv)
    joint += b" " * (-len(joint) % 16)
    a = 1
    for i in range(0, len(joint), 16):
        v0, v1 = struct.unpack("<QQ", joint[i:i + 16])
        a = (a * v0) & (2 ** 64 - 1)
        a ^= v1
    return a


def random_blocks(rng, names, sizes=(0, 1, 2, 3, 5, 17, 24, 25, 40, 200), max_blocks=30,
                  lengths=(1, 1, 2, 20, 32)):
    nb = int(rng.integers(1, max_blocks))
    offs = np.concatenate([[0], np.cumsum(rng.choice(sizes, nb))]).astype(np.uint64)
    n = int(offs[-1])
    seq = rng.integers(0, len(names), n).astype(np.uint32)
    hi = int(rng.choice([5, 30, 1000, 10 ** 6, 2 ** 31 - 40]))
    length = int(rng.choice(lengths))
    mn = rng.integers(0, hi, n).astype(np.uint32)
    ori = rng.choice([-1, 1], n).astype(np.int8)
    return offs, seq, mn, ori, length


def as_lists(offs, seq, mn, mx, ori):
    return [[(int(seq[f]), int(mn[f]), int(mx[f]), int(ori[f]))
             for f in range(int(offs[b]), int(offs[b + 1]))] for b in range(len(offs) - 1)]


# ---- oracle (no GPU) ------------------------------------------------------------------
def test_oracle_block_hash_matches_the_definition():
    rng = np.random.default_rng(11)
    for _ in range(60):
        offs, seq, mn, ori, length = random_blocks(rng, NAMES, sizes=(0, 1, 2, 3, 7, 30))
        mx = mn + rng.integers(0, 40, len(mn)).astype(np.uint32)
        got = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        want = [py_block_hash(NAMES, b) for b in as_lists(offs, seq, mn, mx, ori)]
        assert got.tolist() == want


def test_oracle_blockset_hash_is_the_xor_of_block_hashes():
    rng = np.random.default_rng(12)
    offs, seq, mn, ori, length = random_blocks(rng, NAMES)
    mx = mn + np.uint32(length - 1)
    hashes = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
    total = 0
    for b in range(len(offs) - 1):
        if offs[b + 1] - offs[b] > 1:      # HashGroup::create_task_impl, :101-113
            total ^= int(hashes[b])
    assert total == oracle.blockset_hash(NAMES, as_lists(offs, seq, mn, mx, ori))


# ---- the kernel's code on the host (no GPU) ---------------------------------------------
@pytest.fixture(scope="module")
def host_core(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("bh") / "libbh_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-fPIC",
                           "-shared", "-o", so,
                           os.path.join(ROOT, "tests", "block_hash_host.cpp")])
    lib = C.CDLL(so)
    lib.bh_host_error.restype = C.c_char_p
    lib.bh_host_set_fast_cap.argtypes = [C.c_uint32]
    lib.bh_host_block_hashes.argtypes = (
        [C.POINTER(C.c_char_p), C.c_uint32, C.c_uint64] + [C.c_void_p] * 4 +
        [C.c_uint32, C.c_void_p, C.c_void_p])

    def run(names, offs, seq, mn, mx, length, ori):
        out = np.zeros(max(1, len(offs) - 1), np.uint64)
        arr = (C.c_char_p * max(1, len(names)))(*[n.encode() for n in names])
        rc = lib.bh_host_block_hashes(
            arr, len(names), len(offs) - 1, offs.ctypes.data, seq.ctypes.data,
            mn.ctypes.data, mx.ctypes.data if mx is not None else None, length,
            ori.ctypes.data, out.ctypes.data)
        if rc:
            raise ValueError(lib.bh_host_error().decode())
        return out[:len(offs) - 1]
    run.set_fast_cap = lib.bh_host_set_fast_cap
    return run


def test_core_matches_oracle_fixed_length(host_core):
    rng = np.random.default_rng(1)
    for _ in range(150):
        offs, seq, mn, ori, length = random_blocks(rng, NAMES)
        mx = mn + np.uint32(length - 1)
        want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(host_core(NAMES, offs, seq, mn, None, length, ori), want)


def test_core_matches_oracle_any_length(host_core):
    rng = np.random.default_rng(2)
    for _ in range(150):
        offs, seq, mn, ori, _ = random_blocks(rng, NAMES)
        mx = mn + rng.integers(0, 50, len(mn)).astype(np.uint32)
        want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(host_core(NAMES, offs, seq, mn, mx, 0, ori), want)


def test_core_digit_boundaries_and_one_letter_fragments(host_core):
    # "12_" sorts after "123_" ('_' > digits); last = -1 for a one-letter reverse fragment
    pos = [0, 1, 9, 10, 11, 12, 99, 100, 101, 120, 123, 1234, 12345, 999999, 1000000,
           2 ** 31 - 2, 4294967295 - 40]
    for length in (1, 2, 20):
        frags = [(s, p, o) for s in (3, 4) for p in pos for o in (1, -1)]
        offs = np.array([0, len(frags)], np.uint64)
        seq = np.array([f[0] for f in frags], np.uint32)
        mn = np.array([f[1] for f in frags], np.uint32)
        ori = np.array([f[2] for f in frags], np.int8)
        mx = mn + np.uint32(length - 1)
        want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(host_core(NAMES, offs, seq, mn, None, length, ori), want)
        assert int(want[0]) == py_block_hash(NAMES, as_lists(offs, seq, mn, mx, ori)[0])


def test_core_ties_between_the_two_lists(host_core):
    # an id of one fragment equals the inverted id of another: the lists start
    # with the same id and the one-list shortcut must hand over to the full path
    frags = [(3, 5, 10, 1), (3, 5, 10, -1), (4, 7, 7, 1), (4, 7, 7, -1), (3, 0, 4, 1)]
    for take in (frags[:2], frags[2:4], frags, frags[::-1], frags[:3]):
        offs = np.array([0, len(take)], np.uint64)
        seq = np.array([f[0] for f in take], np.uint32)
        mn = np.array([f[1] for f in take], np.uint32)
        mx = np.array([f[2] for f in take], np.uint32)
        ori = np.array([f[3] for f in take], np.int8)
        want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        assert int(want[0]) == py_block_hash(NAMES, as_lists(offs, seq, mn, mx, ori)[0])
        assert np.array_equal(host_core(NAMES, offs, seq, mn, mx, 0, ori), want)


def test_core_fast_path_capacity(host_core):
    # the shared-memory capacity of the kernel's fast path does not change results
    rng = np.random.default_rng(4)
    offs, seq, mn, ori, length = random_blocks(rng, NAMES, max_blocks=60)
    mx = mn + np.uint32(length - 1)
    want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
    try:
        for cap in (1, 2, 5, 17, 64, 300):
            host_core.set_fast_cap(cap)
            assert np.array_equal(host_core(NAMES, offs, seq, mn, None, length, ori), want)
    finally:
        host_core.set_fast_cap(0)


def test_core_large_block_heap_sort(host_core):
    rng = np.random.default_rng(3)
    n = 5000
    offs = np.array([0, n, n + 2], np.uint64)
    seq = rng.integers(0, len(NAMES), n + 2).astype(np.uint32)
    mn = rng.integers(0, 10 ** 7, n + 2).astype(np.uint32)
    ori = rng.choice([-1, 1], n + 2).astype(np.int8)
    want = oracle.block_hashes(NAMES, offs, seq, mn, mn + np.uint32(19), ori)
    assert np.array_equal(host_core(NAMES, offs, seq, mn, None, 20, ori), want)


def test_core_refuses_names_with_spaces(host_core):
    empty = np.zeros(0, np.uint32)
    offs = np.array([0, 0], np.uint64)
    for bad in (["a b"], ["tab\tname"]):
        with pytest.raises(ValueError):
            host_core(bad, offs, empty, empty, None, 1, np.zeros(0, np.int8))
    # duplicates and names with '_' elsewhere are fine
    host_core(["s_1", "s_2", "s_1"], offs, empty, empty, None, 1, np.zeros(0, np.int8))


# name + "_" of one sequence a prefix of another's: the digits of one id meet the
# letters of the other ("a_5_10" vs "a_1_5_10"), rank order is not id order
PREFIX_NAMES = ["a", "a_1", "a_1_7", "b", "a_12", "s1", "s1_2", "zz", "a_"]


def test_core_names_where_one_is_a_prefix_of_another(host_core):
    rng = np.random.default_rng(5)
    for trial in range(120):
        offs, seq, mn, ori, length = random_blocks(rng, PREFIX_NAMES)
        if trial % 3 == 0:  # small positions: "a_1_5" against "a_15": every digit matters
            mn = (mn % np.uint32(130)).astype(np.uint32)
        mx = mn + rng.integers(0, 50, len(mn)).astype(np.uint32)
        want = oracle.block_hashes(PREFIX_NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(host_core(PREFIX_NAMES, offs, seq, mn, mx, 0, ori), want)
    # the textbook pair, checked against the independent Python statement too
    names = ["a", "a_1"]
    frags = [(0, 5, 10, 1), (1, 5, 10, 1), (0, 1, 5, -1), (1, 0, 9, -1), (0, 15, 15, 1)]
    offs = np.array([0, len(frags)], np.uint64)
    seq = np.array([f[0] for f in frags], np.uint32)
    mn = np.array([f[1] for f in frags], np.uint32)
    mx = np.array([f[2] for f in frags], np.uint32)
    ori = np.array([f[3] for f in frags], np.int8)
    want = oracle.block_hashes(names, offs, seq, mn, mx, ori)
    assert int(want[0]) == py_block_hash(names, as_lists(offs, seq, mn, mx, ori)[0])
    assert np.array_equal(host_core(names, offs, seq, mn, mx, 0, ori), want)


# ---- CUDA path through the C ABI ----------------------------------------------------------
@pytest.fixture(scope="module")
def engine():
    from npge_b200.engine import AnchorEngine
    eng = AnchorEngine(0)
    yield eng
    assert eng.debug_check() == 0   # no kernel stored past the end of a device buffer
    eng.close()


@pytest.mark.gpu
def test_gpu_block_hashes_random_lists(engine):
    rng = np.random.default_rng(21)
    for trial in range(40):
        offs, seq, mn, ori, length = random_blocks(rng, NAMES, max_blocks=400)
        if trial % 2:
            mx = mn + rng.integers(0, 50, len(mn)).astype(np.uint32)
            got, total = engine.block_hashes(NAMES, offs, seq, mn, ori, frag_max_pos=mx)
        else:
            mx = mn + np.uint32(length - 1)
            got, total = engine.block_hashes(NAMES, offs, seq, mn, ori, frag_len=length)
        want = oracle.block_hashes(NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(got, want)
        assert total == oracle.blockset_hash(NAMES, as_lists(offs, seq, mn, mx, ori))


@pytest.mark.gpu
def test_gpu_block_hashes_prefix_names(engine):
    """Names the first version refused ('s1' next to 's1_2'): the device orders such
    ids by their regenerated strings."""
    rng = np.random.default_rng(23)
    for trial in range(20):
        offs, seq, mn, ori, length = random_blocks(rng, PREFIX_NAMES, max_blocks=300)
        if trial % 2:
            mn = (mn % np.uint32(130)).astype(np.uint32)
        mx = mn + rng.integers(0, 50, len(mn)).astype(np.uint32)
        got, total = engine.block_hashes(PREFIX_NAMES, offs, seq, mn, ori, frag_max_pos=mx)
        want = oracle.block_hashes(PREFIX_NAMES, offs, seq, mn, mx, ori)
        assert np.array_equal(got, want)
        assert total == oracle.blockset_hash(PREFIX_NAMES, as_lists(offs, seq, mn, mx, ori))


@pytest.mark.gpu
def test_gpu_block_hashes_large_block_and_small_batches(engine, monkeypatch):
    rng = np.random.default_rng(22)
    sizes = [3, 5000, 2, 0, 700, 17, 17, 1, 64, 33]
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    n = int(offs[-1])
    seq = rng.integers(0, len(NAMES), n).astype(np.uint32)
    mn = rng.integers(0, 10 ** 7, n).astype(np.uint32)
    ori = rng.choice([-1, 1], n).astype(np.int8)
    want = oracle.block_hashes(NAMES, offs, seq, mn, mn + np.uint32(19), ori)
    got, _ = engine.block_hashes(NAMES, offs, seq, mn, ori, frag_len=20)
    assert np.array_equal(got, want)
    monkeypatch.setenv("NPGE_BLOCK_HASH_BATCH", "64")   # many launches, cuts inside blocks
    got, _ = engine.block_hashes(NAMES, offs, seq, mn, ori, frag_len=20)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_block_hashes_of_the_anchor_result(engine):
    names, seqs = synth.generate("tiny")
    for similar, max_frags in ((False, 2 ** 62), (True, 2 ** 62)):
        engine.upload(seqs, 20)
        engine.find_device(20, similar=similar, max_frags=max_frags)
        got, total = engine.anchor_block_hashes(names)
        r = engine.fetch()
        offs = np.asarray(r.block_offsets, np.uint64)
        mn = np.asarray(r.frag_min_pos, np.uint32)
        want = oracle.block_hashes(names, offs, r.frag_seq, mn, mn + np.uint32(19), r.frag_ori)
        assert len(got) == r.n_blocks and np.array_equal(got, want)
        assert total == oracle.blockset_hash(
            names, as_lists(offs, r.frag_seq, mn, mn + np.uint32(19), r.frag_ori))
        # the host-array entry point over the fetched arrays agrees
        got2, total2 = engine.block_hashes(names, offs, r.frag_seq, mn, r.frag_ori, frag_len=20)
        assert np.array_equal(got2, want) and total2 == total


@pytest.mark.gpu
def test_gpu_block_hashes_errors(engine):
    from npge_b200 import _abi
    from npge_b200.engine import AnchorError
    offs = np.array([0, 2], np.uint64)
    seq = np.array([0, 1], np.uint32)
    mn = np.array([5, 9], np.uint32)
    ori = np.array([1, -1], np.int8)
    with pytest.raises(AnchorError) as e:      # a space in a name
        engine.block_hashes(["s1", "s 2"], offs, seq, mn, ori, frag_len=4)
    assert e.value.code == _abi.ERR_PARAMS
    with pytest.raises(AnchorError) as e:      # sequence index out of range
        engine.block_hashes(["only"], offs, seq, mn, ori, frag_len=4)
    assert e.value.code == _abi.ERR_PARAMS
    with pytest.raises(AnchorError) as e:      # offsets not ascending
        engine.block_hashes(["a", "b"], np.array([0, 2, 1], np.uint64), seq, mn, ori, frag_len=4)
    assert e.value.code == _abi.ERR_PARAMS


@pytest.mark.gpu
@pytest.mark.parametrize("limit", [7, 8, 9, 300, 1001])
def test_gpu_block_hashes_after_a_cut(engine, limit):
    """max-anchor-fragments cuts the list inside a block: find_device patches the
    resident result (reference order inside the cut group, a lone tail dropped), so
    the device-side hashes equal those of the fetched arrays."""
    names, seqs = synth.generate("tiny")
    engine.upload(seqs, 20)
    for similar in (False, True):
        st = engine.find_device(20, similar=similar, max_frags=limit)
        want_r = oracle.find_ideal(seqs, 20, similar=similar, max_frags=limit)
        assert (st["n_fragments"], st["n_blocks"], st["digest"]) == (
            want_r.n_fragments, want_r.n_blocks, want_r.digest)
        got, total = engine.anchor_block_hashes(names)
        r = engine.fetch()
        assert r.blocks() == want_r.blocks() and r.digest == want_r.digest
        assert list(r.new_used) == list(want_r.new_used)
        offs = np.asarray(r.block_offsets, np.uint64)
        mn = np.asarray(r.frag_min_pos, np.uint32)
        want = oracle.block_hashes(names, offs, r.frag_seq, mn, mn + np.uint32(19), r.frag_ori)
        assert len(got) == r.n_blocks and np.array_equal(got, want)


# ---- the C++ processor npge::gpu::MoveUnchanged ---------------------------------------------
def run_move_unchanged(*args):
    from npge_b200 import build
    build.build_shim(force=False)
    p = subprocess.run([build.SHIM_TEST_MOVE] + list(args), capture_output=True, text=True,
                       timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert " 0 failures" in p.stdout
    return p.stdout


def test_move_unchanged_host_side_without_a_device():
    assert "no-gpu subset" in run_move_unchanged("--no-gpu")


@pytest.mark.gpu
def test_move_unchanged_processor_on_the_gpu():
    assert "no-gpu" not in run_move_unchanged()
