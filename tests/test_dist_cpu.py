"""Host logic of the multi-GPU path (npge_b200/dist.py) on CPU: the table
builders as pure functions, and a world_size-2 gloo job whose device stages are
replaced by a numpy stand-in (tests/dist_standin.py), checked against the
oracle on the whole job."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from npge_b200 import dist as D

HERE = os.path.dirname(os.path.abspath(__file__))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def launch(world, args, timeout=600, extra_env=None):
    port = free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, **(extra_env or {}))
        env.update(RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen(
            [sys.executable, os.path.join(HERE, "dist_worker.py")] + [str(a) for a in args],
            env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=timeout)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d failed:\n%s" % (r, out[-4000:])
    return outs


def test_scan_order_and_position_ranges():
    sizes = [10, 50, 3, 50, 20, 7]
    order = D.scan_order(sizes, 5)
    assert list(order) == [1, 3, 4, 0, 5]          # size desc, ties by index; 3 < k
    cuts = D.position_cuts(10000, 3)               # equal ranges on tile borders
    assert list(cuts) == [0, 4096, 8192, 10000]
    assert list(D.position_cuts(100, 4)) == [0, 100, 100, 100, 100]
    assert list(D.position_cuts(0, 2)) == [0, 0, 0]
    # sequences of 5 and 3 bases: job positions 0..4 | pad 5 | 6..8 | pad 9
    gs = D.global_starts([5, 3])
    assert D.range_pieces(gs, 0, 10) == [(0, 0, 5, 0), (1, 0, 3, 6)]
    assert D.range_pieces(gs, 3, 7) == [(0, 3, 5, 3), (1, 0, 1, 6)]   # a sequence is cut
    assert D.range_pieces(gs, 5, 6) == [] and D.range_pieces(gs, 4, 4) == []
    assert list(D.global_starts([5, 3])) == [0, 6, 10]


def test_bucket_ranges_balance():
    rng = np.random.default_rng(1)
    h = rng.integers(0, 1000, size=(4, 256))
    cuts = D.bucket_ranges(h, 4)
    assert cuts[0] == 0 and cuts[-1] == 256 and np.all(np.diff(cuts) > 0)
    per = [h[:, cuts[r]:cuts[r + 1]].sum() for r in range(4)]
    assert max(per) < 1.1 * h.sum() / 4
    # splits are each other's transposes
    send = [D.exchange_splits(h, cuts, r)[0] for r in range(4)]
    recv = [D.exchange_splits(h, cuts, r)[1] for r in range(4)]
    assert np.array_equal(np.asarray(send).T, np.asarray(recv))
    # everything in one bucket: one rank takes it all, the others nothing
    h1 = np.zeros((2, 256), dtype=np.int64)
    h1[:, 7] = 100
    c1 = D.bucket_ranges(h1, 2)
    t0, t1 = D.segment_tables(h1, c1, 0), D.segment_tables(h1, c1, 1)
    assert sorted([t0["n_recv"], t1["n_recv"]]) == [0, 200]


def test_segment_tables():
    h = np.zeros((3, 8), dtype=np.int64)
    h[0] = [5, 0, 4097, 1, 0, 0, 2, 0]
    h[1] = [0, 0, 3, 0, 9, 0, 0, 0]
    h[2] = [1, 0, 0, 0, 1, 0, 0, 8192]
    cuts = np.asarray([0, 3, 7, 8])
    t = D.segment_tables(h, cuts, 0)
    # receive buffer of rank 0: [src0: a0 x5, a2 x4097][src1: a2 x3][src2: a0 x1]
    assert t["n_recv"] == 5 + 4097 + 3 + 1
    assert list(t["seg_a"]) == [0, 0, 2, 2]
    assert list(t["seg_begin"]) == [0, 4105, 5, 4102]
    assert list(t["seg_end"]) == [5, 4106, 4102, 4105]
    assert list(t["seg_tile"]) == [0, 1, 2, 4, 5]
    assert list(t["a_tile0"]) == [0, 2, 2, 5, 5, 5, 5, 5, 5]
    assert list(t["a_start"]) == [0, 6, 6, 4106, 4106, 4106, 4106, 4106, 4106]
    t = D.segment_tables(h, cuts, 2)
    assert t["n_recv"] == 8192 and list(t["seg_tile"]) == [0, 2]
    empty = D.segment_tables(np.zeros((2, 8), dtype=np.int64), np.asarray([0, 4, 8]), 1)
    assert empty["n_recv"] == 0 and empty["n_seg"] == 1 and list(empty["seg_tile"]) == [0, 0]


def test_peer_destinations_match_the_receive_layout():
    rng = np.random.default_rng(5)
    h = rng.integers(0, 50, size=(3, 16))
    cuts = D.bucket_ranges(h, 3)
    # scatter element numbers through the destination tables and compare with
    # the layout an all-to-all of the A-sorted slices would produce
    recv = [dict() for _ in range(3)]
    for src in range(3):
        dest_rank, dest_off, n_recv = D.peer_destinations(h, cuts, src)
        a_start = np.concatenate([[0], np.cumsum(h[src])])
        for a in range(16):
            for i in range(int(a_start[a]), int(a_start[a + 1])):
                slot = int(dest_off[a]) + i
                assert slot not in recv[dest_rank[a]]
                recv[dest_rank[a]][slot] = (src, a, i - int(a_start[a]))
    for d in range(3):
        t = D.segment_tables(h, cuts, d)
        assert sorted(recv[d]) == list(range(t["n_recv"])) and n_recv[d] == t["n_recv"]
        for b, e, a in zip(t["seg_begin"], t["seg_end"], t["seg_a"]):
            got = [recv[d][i] for i in range(int(b), int(e))]
            assert len({g[0] for g in got}) <= 1 and all(g[1] == a for g in got)
            assert [g[2] for g in got] == list(range(len(got)))


def plan_tables_cxx(h, rank, chunks=1):
    """npge_anchor_dist_plan_tables (csrc/dist_plan.h) through the C ABI; host only.
    h: [world * chunks sources][bins]."""
    from npge_b200 import _abi
    L = _abi.load()
    h = np.ascontiguousarray(h, dtype=np.uint32)
    n_src, bins = h.shape
    world = n_src // chunks
    u32 = lambda n: np.zeros(n, dtype=np.uint32)
    cuts, n_recv = u32(world + 1), np.zeros(world, dtype=np.uint64)
    dest_rank, dest_off = u32(bins), np.zeros(bins, dtype=np.int64)
    n_seg = np.zeros(1, dtype=np.uint32)
    sb, se, sa, stl = u32(n_src * bins), u32(n_src * bins), u32(n_src * bins), u32(n_src * bins + 1)
    at0, ast = u32(bins + 1), u32(bins + 1)
    rc = L.npge_anchor_dist_plan_tables(
        h.ctypes.data, world, chunks, bins, rank, *[x.ctypes.data for x in (
            cuts, n_recv, dest_rank, dest_off, n_seg, sb, se, sa, stl, at0, ast)])
    assert rc == 0
    n = int(n_seg[0])
    return dict(cuts=cuts, n_recv_all=n_recv, dest_rank=dest_rank, dest_off=dest_off, n_seg=n,
                seg_begin=sb[:n], seg_end=se[:n], seg_a=sa[:n], seg_tile=stl[:n + 1],
                a_tile0=at0, a_start=ast)


@pytest.mark.parametrize("seed,world,bins", [(1, 2, 256), (2, 3, 16), (3, 8, 1024), (4, 4, 512)])
def test_cxx_plan_tables_equal_the_numpy_ones(seed, world, bins):
    """The library plans the exchange in C++ (no numpy on the step's path); the numpy
    functions of dist.py, which the stand-in job checks against the oracle, are the
    restatement it must agree with."""
    rng = np.random.default_rng(seed)
    h = rng.integers(0, 9000, size=(world, bins))
    h[rng.random((world, bins)) < 0.3] = 0
    if seed == 2:
        h[:, 5:] = 0                       # ranks that receive nothing
    cuts = D.bucket_ranges(h, world)
    for rank in range(world):
        t = plan_tables_cxx(h, rank)
        assert list(t["cuts"]) == list(cuts)
        want = D.segment_tables(h, cuts, rank)
        for key in ("n_seg", "seg_begin", "seg_end", "seg_a", "seg_tile", "a_tile0", "a_start"):
            assert np.array_equal(np.asarray(t[key]), np.asarray(want[key])), key
        dest_rank, dest_off, n_recv = D.peer_destinations(h, cuts, rank)
        assert np.array_equal(t["dest_rank"], dest_rank)
        assert np.array_equal(t["dest_off"], dest_off)
        assert [int(x) for x in t["n_recv_all"]] == n_recv


def test_cxx_plan_tables_with_chunked_sources():
    """Pass A in chunks: every (rank, chunk) is a source of the exchange; the receive
    layout is source-major, so the numpy tables over the virtual sources describe it."""
    rng = np.random.default_rng(9)
    world, chunks, bins = 3, 4, 64
    h = rng.integers(0, 6000, size=(world * chunks, bins))
    cuts = D.bucket_ranges(h, world)
    for rank in range(world):
        t = plan_tables_cxx(h, rank, chunks)
        assert list(t["cuts"]) == list(cuts)
        want = D.segment_tables(h, cuts, rank)
        for key in ("n_seg", "seg_begin", "seg_end", "seg_a", "seg_tile", "a_tile0", "a_start"):
            assert np.array_equal(np.asarray(t[key]), np.asarray(want[key])), key
        assert int(t["n_recv_all"][rank]) == want["n_recv"]


def test_local_limits():
    assert D.local_limits([10, 0, 7, 5], 12) == [10, 0, 2, 0]
    assert D.local_limits([10, 7], 100) == [10, 7]
    assert D.local_limits([10, 7], 0) == [0, 0]


@pytest.mark.parametrize("how", ["interleaved", "blocks"])
def test_two_ranks_gloo_standin(how):
    launch(2, ["standin", "tiny", 1.0, 12, how])


def test_three_ranks_gloo_standin_wide_key():
    launch(3, ["standin", "tiny", 1.0, 27, "interleaved"])


def test_one_sequence_over_three_ranks_standin():
    """Positions are sharded, not sequences: one chromosome is cut into
    three ranges with k - 1 bases of halo and the job equals the oracle's answer."""
    launch(3, ["standin", "oneseq", 1.0, 16, "interleaved"])
