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


def test_scan_order_and_shards():
    sizes = [10, 50, 3, 50, 20, 7]
    order = D.scan_order(sizes, 5)
    assert list(order) == [1, 3, 4, 0, 5]          # size desc, ties by index; 3 < k
    order, cuts = D.plan_shards(sizes, 5, 2)
    assert list(cuts) == [0, 1, 5] or list(cuts) == [0, 2, 5]
    order, cuts = D.plan_shards(sizes, 5, 8)       # more ranks than sequences
    assert cuts[0] == 0 and cuts[-1] == 5 and np.all(np.diff(cuts) >= 0)
    order, cuts = D.plan_shards([], 5, 3)
    assert list(cuts) == [0, 0, 0, 0]
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


def test_local_limits():
    assert D.local_limits([10, 0, 7, 5], 12) == [10, 0, 2, 0]
    assert D.local_limits([10, 7], 100) == [10, 7]
    assert D.local_limits([10, 7], 0) == [0, 0]


@pytest.mark.parametrize("how", ["interleaved", "blocks"])
def test_two_ranks_gloo_standin(how):
    launch(2, ["standin", "tiny", 1.0, 12, how])


def test_three_ranks_gloo_standin_wide_key():
    launch(3, ["standin", "tiny", 1.0, 27, "interleaved"])
