"""AnchorFinder over several GPUs: one process and one CUDA context per GPU.

SURVEY 8e: the reference parallelises AnchorFinder over sequences with a
thread pool that shares one Bloom filter and one hash -> positions map
(src/algo/AnchorFinder.cpp:252-292, 330-357).  Here positions are sharded
over the ranks and canonical k-mers are repartitioned by hash range:

  upload       every rank ends up with a contiguous range of the job-wide scan
               order (sequences by size, descending; ties by caller index),
               whatever the placement it was given (an all-to-all of the
               sequences moves what is misplaced)
  pass A       local: k-mers -> records partitioned by the top key bits
  all-gather   of the per-bucket counts; contiguous bucket ranges are cut so that
               every rank receives about the same number of records
  all-to-all   of the records (NCCL over NVLink; the only bulk exchange)
  pass B, bucket sort, count   local, over (bucket, source rank) segments
  all-reduce   of the found bitmap (anchor-similar only: found(p - 1) may
               belong to another rank's hash range)
  all-gather   of the totals -> every rank's share of max-anchor-fragments
  emit         local: rank r holds the r-th slice of the one-GPU result

The per-rank stages are the npge_anchor_dist_* entry points of the C ABI
(include/npge_anchor.h); this module owns the collectives and the small host
tables between them.  `stages` can be replaced by an object with the same
methods (the CPU tests drive the host logic with a numpy stand-in).
"""
import ctypes as C

import numpy as np

from . import _abi
from .engine import AnchorEngine, AnchorError, AnchorResult, _stats_dict, concat

TILE = 4096  # records per pass-B tile (kTile in csrc/common.cuh)


# ---- host tables (pure functions; tests/test_dist_cpu.py) --------------------------
def scan_order(sizes, k):
    """Job-wide scan order: indices of the sequences with size >= k, by size
    descending, ties by caller index (SeqI.hpp:45-57; api.cu upload)."""
    sizes = np.asarray(sizes, dtype=np.int64)
    idx = np.flatnonzero(sizes >= k)
    return idx[np.argsort(-sizes[idx], kind="stable")]


POS_ALIGN = 4096  # cuts of the position space fall on pass-A tile borders


def position_cuts(n_positions, world):
    """Equal ranges of the job's position space: rank r owns the windows
    that start in [cuts[r], cuts[r + 1]).  The same as dist_position_cuts in
    csrc/dist_plan.h."""
    per = -(-(-(-int(n_positions) // world)) // POS_ALIGN) * POS_ALIGN
    return np.minimum(np.arange(world + 1, dtype=np.int64) * per, int(n_positions))


def range_pieces(gstart, lo, hi):
    """The parts of the sequences (scan order, job-wide starts `gstart`, one pad after
    each) that lie in the position range [lo, hi): list of (scan index, first base,
    one past the last base, job-wide position of the first base)."""
    gstart = np.asarray(gstart, dtype=np.int64)
    out = []
    if hi <= lo:
        return out
    first = max(int(np.searchsorted(gstart, lo, side="right")) - 1, 0)
    for i in range(first, len(gstart) - 1):
        b, e = int(gstart[i]), int(gstart[i + 1]) - 1  # bases of sequence i
        if b >= hi:
            break
        x, y = max(b, lo), min(e, hi)
        if y > x:
            out.append((i, x - b, y - b, x))
    return out


def global_starts(sizes_in_scan_order):
    """First job-wide position of every sequence (+ the end): one pad position
    after each sequence, as in the one-GPU layout."""
    s = np.asarray(sizes_in_scan_order, dtype=np.uint64)
    out = np.zeros(len(s) + 1, dtype=np.uint64)
    out[1:] = np.cumsum(s + np.uint64(1))
    return out


def bucket_ranges(hist_all, world):
    """Contiguous A-bucket ranges, balanced by records.

    hist_all: [world, bins] records of every source rank per bucket.  Returns
    cuts[world + 1]: rank r receives the buckets cuts[r] .. cuts[r + 1] - 1."""
    tot = np.asarray(hist_all, dtype=np.int64).sum(axis=0)
    csum = np.concatenate([[0], np.cumsum(tot)])
    total = int(csum[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        j = int(np.searchsorted(csum, target))
        if j > 0 and target - csum[j - 1] < csum[min(j, len(tot))] - target:
            j -= 1
        cuts.append(min(max(j, cuts[-1]), len(tot)))
    cuts.append(len(tot))
    return np.asarray(cuts, dtype=np.int64)


def exchange_splits(hist_all, cuts, rank):
    """(send_splits, recv_splits) of the record all-to-all, in records."""
    h = np.asarray(hist_all, dtype=np.int64)
    world = h.shape[0]
    send = [int(h[rank, cuts[d]:cuts[d + 1]].sum()) for d in range(world)]
    recv = [int(h[s, cuts[rank]:cuts[rank + 1]].sum()) for s in range(world)]
    return send, recv


def segment_tables(hist_all, cuts, rank):
    """Pass-B tables of rank `rank`: one segment per (owned bucket, source rank)
    with records, bucket-major so that every bucket's tiles are consecutive
    and its sources come in rank (= position) order.

    The receive buffer holds source 0's slice, then source 1's, ...; inside a
    slice the buckets ascend.  Returns a dict of uint32 arrays (SegTables in
    csrc/kernels.h) and n_recv."""
    h = np.asarray(hist_all, dtype=np.int64)
    world, bins = h.shape
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    mine = h[:, lo:hi]                                   # [world, owned]
    src_base = np.concatenate([[0], np.cumsum(mine.sum(axis=1))])[:-1]
    inside = np.cumsum(mine, axis=1) - mine              # offset inside the slice
    begin = (src_base[:, None] + inside)                 # [world, owned]
    c = mine.T.reshape(-1)                               # bucket-major, then source
    keep = c > 0
    sb = begin.T.reshape(-1)[keep]
    se = sb + c[keep]
    seg_a = np.repeat(np.arange(lo, hi), world)[keep]
    a_count = np.zeros(bins, dtype=np.int64)
    a_count[lo:hi] = mine.sum(axis=0)
    if len(sb) == 0:  # nothing received: one empty segment keeps the tables valid
        sb, se = np.zeros(1, np.int64), np.zeros(1, np.int64)
        seg_a = np.asarray([lo if lo < bins else 0])
    tiles = (se - sb + TILE - 1) // TILE
    a_tiles = np.bincount(seg_a, weights=tiles, minlength=bins).astype(np.int64)
    u32 = lambda x: np.ascontiguousarray(np.asarray(x, dtype=np.uint32))
    return dict(
        n_seg=len(sb), seg_begin=u32(sb), seg_end=u32(se), seg_a=u32(seg_a),
        seg_tile=u32(np.concatenate([[0], np.cumsum(tiles)])),
        a_tile0=u32(np.concatenate([[0], np.cumsum(a_tiles)])),
        a_start=u32(np.concatenate([[0], np.cumsum(a_count)])),
        n_recv=int(mine.sum()))


def peer_destinations(hist_all, cuts, rank):
    """Per A bucket: the rank that owns it, and the record offset in that rank's
    receive buffer of element 0 of THIS rank's A-sorted order (so that element
    idx of the bucket's run lands at offset + idx; see segment_tables for the
    receive layout).  Also the records every rank receives."""
    h = np.asarray(hist_all, dtype=np.int64)
    world, bins = h.shape
    a_start = np.concatenate([[0], np.cumsum(h[rank])])
    dest_rank = np.zeros(bins, dtype=np.uint32)
    dest_off = np.zeros(bins, dtype=np.int64)
    n_recv = []
    for d in range(world):
        lo, hi = int(cuts[d]), int(cuts[d + 1])
        per_src = h[:, lo:hi].sum(axis=1)
        n_recv.append(int(per_src.sum()))
        dest_rank[lo:hi] = d
        dest_off[lo:hi] = int(per_src[:rank].sum()) - int(a_start[lo])
    return dest_rank, dest_off, n_recv


def local_limits(totals, limit):
    """Share of max-anchor-fragments of every rank: the cut runs over the list
    ordered by hash, and the ranks own ascending hash ranges."""
    out, before = [], 0
    for t in totals:
        out.append(int(min(max(limit - before, 0), t)))
        before += int(t)
    return out


# ---- the per-rank device stages (C ABI) ----------------------------------------------
class _DevMem:
    """__cuda_array_interface__ view of a device buffer the library owns."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {
            "shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
            "version": 3, "strides": None}


class CudaStages(AnchorEngine):
    """npge_anchor_dist_* on one GPU.  Tensors returned by buffers() alias the
    library's device memory and stay valid until the next upload / buffers()."""

    def __init__(self, device=0):
        AnchorEngine.__init__(self, device)
        self.device = device

    def dist_upload_range(self, k, bases, pieces, rng, global_start, global_caller):
        """bases: numpy array or CUDA uint8 tensor holding the pieces; pieces =
        (src_off, dst_pos, len) uint64 arrays; rng = (g0, n_own, n_have)."""
        base_ptr = bases.data_ptr() if hasattr(bases, "data_ptr") else bases.ctypes.data
        so, dp, ln = (np.ascontiguousarray(v, dtype=np.uint64) for v in pieces)
        gs = np.ascontiguousarray(global_start, dtype=np.uint64)
        gc = np.ascontiguousarray(global_caller, dtype=np.uint32)
        self._check(self._lib.npge_anchor_dist_upload_range(
            self._ctx, k, base_ptr, len(so), so.ctypes.data, dp.ctypes.data, ln.ctypes.data,
            int(rng[0]), int(rng[1]), int(rng[2]), gs.ctypes.data, gc.ctypes.data, len(gc)))

    def dist_exchange(self, params, hist_all, world, rank):
        """Plans the exchange from the all-gathered histograms and copies this rank's
        slices into the owners' receive buffers.  Returns the records every rank
        receives; raises AnchorError(ERR_CAPACITY) with .n_recv set when a receive
        buffer is too small."""
        h = np.ascontiguousarray(hist_all, dtype=np.uint32)
        n_recv = np.zeros(world, dtype=np.uint64)
        rc = self._lib.npge_anchor_dist_exchange(
            self._ctx, C.byref(params), h.ctypes.data, world, rank, n_recv.ctypes.data)
        if rc != _abi.OK:
            e = AnchorError(rc, self._lib.npge_anchor_last_error().decode())
            e.n_recv = [int(x) for x in n_recv]
            raise e
        return [int(x) for x in n_recv]

    def dist_sort_planned(self, params, stage, used):
        u, up = self._used(used)
        totals = (C.c_uint64 * 2)()
        self._check(self._lib.npge_anchor_dist_sort_planned(
            self._ctx, C.byref(params), stage, up, len(u), C.byref(totals)))
        return int(totals[0]), int(totals[1])

    def dist_pass_a(self, params):
        hist = np.zeros(1024, dtype=np.uint32)
        n_bins = C.c_uint32()
        self._check(self._lib.npge_anchor_dist_pass_a(
            self._ctx, C.byref(params), hist.ctypes.data, len(hist), C.byref(n_bins)))
        return hist[:n_bins.value].copy()

    _chunks = 1

    def dist_set_chunks(self, chunks):
        """Pass A of the following steps runs in `chunks` chunks of tiles; dist_count_a
        then returns chunks * bins counts ([chunk][bucket])."""
        self._check(self._lib.npge_anchor_dist_set_chunks(self._ctx, chunks))
        self._chunks = chunks

    def dist_count_a(self, params):
        hist = np.zeros(8 * 1024, dtype=np.uint32)
        n_bins = C.c_uint32()
        self._check(self._lib.npge_anchor_dist_count_a(
            self._ctx, C.byref(params), hist.ctypes.data, len(hist), C.byref(n_bins)))
        return hist[:n_bins.value * self._chunks].copy()

    def dist_peer_close(self):
        self._check(self._lib.npge_anchor_dist_peer_close(self._ctx))

    def dist_peer_alloc(self, capacity):
        h = np.zeros(_abi.IPC_BYTES, dtype=np.uint8)
        self._check(self._lib.npge_anchor_dist_peer_alloc(self._ctx, capacity,
                                                          h.ctypes.data))
        return h

    def dist_peer_open(self, world, rank, handles):
        h = np.ascontiguousarray(handles, dtype=np.uint8)
        assert h.size == world * _abi.IPC_BYTES
        self._check(self._lib.npge_anchor_dist_peer_open(self._ctx, world, rank,
                                                         h.ctypes.data))

    def dist_scatter_peer(self, params, dest_rank, dest_off):
        dr = np.ascontiguousarray(dest_rank, dtype=np.uint32)
        do = np.ascontiguousarray(dest_off, dtype=np.int64)
        self._check(self._lib.npge_anchor_dist_scatter_peer(
            self._ctx, C.byref(params), dr.ctypes.data, do.ctypes.data, len(dr)))

    def dist_scatter_local(self, params):
        self._check(self._lib.npge_anchor_dist_scatter_local(self._ctx, C.byref(params)))

    def dist_buffers(self, n_send, n_recv):
        import torch
        p = _abi.DistPtrs()
        self._check(self._lib.npge_anchor_dist_buffers(self._ctx, n_recv, C.byref(p)))
        dev = torch.device("cuda", self.device)

        def view(ptr, n, typestr):
            if not ptr:
                return None
            if n == 0:
                return torch.zeros(0, device=dev,
                                   dtype=torch.int64 if typestr == "<i8" else torch.int32)
            return torch.as_tensor(_DevMem(ptr, n, typestr), device=dev)

        return dict(send_rec=view(p.send_rec, n_send, "<i8"),
                    send_hi=view(p.send_hi, n_send, "<i4") if p.wide else None,
                    recv_rec=view(p.recv_rec, n_recv, "<i8"),
                    recv_hi=view(p.recv_hi, n_recv, "<i4") if p.wide else None,
                    found=view(p.found_bits, p.found_words, "<i4"))

    def dist_sort(self, params, stage, tables, used):
        u, up = self._used(used)
        totals = (C.c_uint64 * 2)()
        t = tables
        self._check(self._lib.npge_anchor_dist_sort(
            self._ctx, C.byref(params), stage, t["n_recv"], t["n_seg"],
            t["seg_begin"].ctypes.data, t["seg_end"].ctypes.data,
            t["seg_a"].ctypes.data, t["seg_tile"].ctypes.data,
            t["a_tile0"].ctypes.data, t["a_start"].ctypes.data, up, len(u),
            C.byref(totals)))
        return int(totals[0]), int(totals[1])

    def dist_emit(self, params, local_limit):
        s = _abi.Stats()
        self._check(self._lib.npge_anchor_dist_emit(
            self._ctx, C.byref(params), local_limit, C.byref(s)))
        return _stats_dict(s)


# ---- collectives -----------------------------------------------------------------------
class _ShmGather:
    """All-gather of small int64 arrays between the ranks of ONE host through a
    file in /dev/shm: no GPU kernel (NCCL's cannot start while the library's
    kernels hold every register of every SM, so a collective that should
    overlap pass A would wait for it), no sockets, ~10 us.

    Layout [2 buffers][world][1 + CAP] int64; word 0 of a row is the sequence
    number, written after the payload (x86 keeps the store order).  Buffers
    alternate; a rank cannot run two gathers ahead of the slowest one, because
    it needs everybody's row of the gather in between."""
    CAP = 8 * 1024 + 64

    def __init__(self, comm):
        import os
        import socket
        self.rank, self.world = comm.rank, comm.world
        path = None
        if self.rank == 0:
            path = "/dev/shm/npge_anchor_%d_%x" % (os.getpid(), id(self) & 0xffffff)
            try:
                with open(path, "wb") as f:
                    f.truncate(2 * self.world * (self.CAP + 1) * 8)
            except OSError:
                path = None
        info = comm.gather_objects((socket.gethostname(), path))
        path = info[0][1]
        self.ok = path is not None and len({h for h, _ in info}) == 1
        self.mm = None
        if self.ok:
            try:
                self.mm = np.memmap(path, dtype=np.int64, mode="r+",
                                    shape=(2, self.world, self.CAP + 1))
            except OSError:
                self.ok = False
        self.ok = all(comm.gather_objects(self.ok))
        if self.rank == 0 and path is not None:
            try:
                os.unlink(path)  # the mappings keep it alive
            except OSError:
                pass
        self.step = 0

    def gather(self, arr, timeout=120.0):
        import time
        a = np.ascontiguousarray(arr, dtype=np.int64).reshape(-1)
        n = len(a)
        assert n <= self.CAP
        self.step += 1
        step, buf = self.step, self.mm[self.step & 1]
        buf[self.rank, 1:1 + n] = a
        buf[self.rank, 0] = step
        out = np.empty((self.world, n), dtype=np.int64)
        t0 = None
        for r in range(self.world):
            while buf[r, 0] != step:
                if t0 is None:
                    t0 = time.monotonic()
                elif time.monotonic() - t0 > timeout:
                    raise AnchorError(_abi.ERR_STATE, "rank %d did not reach the gather" % r)
            out[r] = buf[r, 1:1 + n]
        return out


class _Comm:
    """torch.distributed with the tensors where the backend wants them: NCCL
    works on the device buffers in place; gloo (CPU tests, or several ranks on
    one GPU) goes through host copies."""

    def __init__(self, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.nccl = dist.get_backend(group) == "nccl"

    def enable_shm(self):
        """Collective.  Small gathers and barriers go through shared memory when
        all ranks are on one host."""
        shm = _ShmGather(self)
        self.shm = shm if shm.ok else None

    def gather_objects(self, obj):
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def gather_array(self, arr, device=None):
        """all-gather of equal-length int64 numpy arrays -> [world, n]"""
        if getattr(self, "shm", None) is not None and np.size(arr) <= _ShmGather.CAP:
            a = np.asarray(arr)
            return self.shm.gather(a).reshape((self.world,) + a.shape)
        t = self.torch.from_numpy(np.ascontiguousarray(arr, dtype=np.int64))
        if self.nccl:
            t = t.to(device)
        out = self.torch.empty(self.world * t.numel(), dtype=t.dtype, device=t.device)
        self.dist.all_gather_into_tensor(out, t.reshape(-1), group=self.group)
        return out.cpu().numpy().reshape((self.world,) + tuple(t.shape))

    def gather_varlen(self, arr, device=None):
        """all-gather of int64 arrays of different lengths -> list of `world` arrays
        (two fixed-length gathers: no pickling, and through /dev/shm when it is there)"""
        a = np.asarray(arr, dtype=np.int64).reshape(-1)
        counts = self.gather_array(np.asarray([len(a)], dtype=np.int64), device)[:, 0]
        m = int(counts.max())
        if m == 0:
            return [np.zeros(0, np.int64) for _ in range(self.world)]
        pad = np.zeros(m, dtype=np.int64)
        pad[:len(a)] = a
        rows = self.gather_array(pad, device)
        return [np.asarray(rows[r, :int(counts[r])], dtype=np.int64) for r in range(self.world)]

    def _done(self, t):
        # the library's kernels run on its own stream: the collective must have
        # finished (not just been enqueued) before the next stage is launched
        if t.is_cuda:
            self.torch.cuda.current_stream(t.device).synchronize()

    def all_to_all(self, out, inp, out_splits, in_splits):
        if self.nccl or not out.is_cuda:
            self.dist.all_to_all_single(out, inp, list(out_splits), list(in_splits),
                                        group=self.group)
            self._done(out)
            return
        o, i = out.cpu(), inp.cpu()
        self._a2a_host(o, i, out_splits, in_splits)
        out.copy_(o)

    def _a2a_host(self, o, i, out_splits, in_splits):
        try:
            self.dist.all_to_all_single(o, i, list(out_splits), list(in_splits),
                                        group=self.group)
        except RuntimeError:  # gloo builds without alltoall: pairwise exchange
            ins = list(i.split(list(in_splits)))
            outs = list(o.split(list(out_splits)))
            outs[self.rank].copy_(ins[self.rank])
            for step in range(1, self.world):
                dst = (self.rank + step) % self.world
                src = (self.rank - step) % self.world
                req = self.dist.isend(ins[dst].contiguous(), dst, group=self.group)
                self.dist.recv(outs[src], src, group=self.group)
                req.wait()

    def sum_(self, t):
        if self.nccl or not t.is_cuda:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
            self._done(t)
            return
        h = t.cpu()
        self.dist.all_reduce(h, op=self.dist.ReduceOp.SUM, group=self.group)
        t.copy_(h)

    def barrier(self, device=None):
        """Every rank has arrived (and, for a CUDA device, the host has seen it:
        the next kernels go to the library's own stream)."""
        if getattr(self, "shm", None) is not None:
            self.shm.gather(np.zeros(1, dtype=np.int64))
            return
        if self.nccl and device is not None:
            if getattr(self, "_one", None) is None:
                self._one = self.torch.zeros(1, dtype=self.torch.int32, device=device)
            self.dist.all_reduce(self._one, group=self.group)
            self.torch.cuda.current_stream(device).synchronize()
            return
        self.dist.barrier(group=self.group)


class DistAnchorFinder:
    """AnchorFinder of one job over all ranks of a process group.

    Every rank calls the same methods with its own sequences (any placement);
    fetch() returns the rank's slice of the job's result (ascending hash
    ranges), with frag_seq = job-wide caller index: the ranks' sequences
    numbered one after the other in rank order."""

    def __init__(self, device=0, group=None, stages=None, exchange=None):
        """exchange: how the records reach the rank that owns their hash range.
        "bulk" (the default): local pass A, then the library copies every
        destination's contiguous slice into its receive buffer, mapped through
        CUDA IPC: one asynchronous device-to-device copy per destination, all in
        flight at once, planned in C++ (csrc/dist_plan.h); "peer": pass A itself
        scatters every bucket into the owners' HBM over NVLink (no extra pass, but
        32-byte runs); "collective": local pass A + all_to_all_single, for process
        groups whose ranks cannot map each other's memory (and the CPU tests)."""
        import os
        self.comm = _Comm(group)
        self.rank, self.world = self.comm.rank, self.comm.world
        if os.environ.get("NPGE_ANCHOR_SHM", "1") != "0":
            self.comm.enable_shm()
        self.stages = stages if stages is not None else CudaStages(device)
        self.device = device
        self._k = None
        self.exchange_ms = {}
        want = exchange or os.environ.get("NPGE_ANCHOR_EXCHANGE") or "auto"
        if want not in ("auto", "peer", "bulk", "collective"):
            raise ValueError("exchange must be peer, bulk, collective or auto")
        if not hasattr(self.stages, "dist_peer_alloc"):
            if want in ("peer", "bulk"):
                raise ValueError("these stages have no peer exchange")
            want = "collective"
        self._want = want
        self.exchange = "collective"
        self._capacity = 0
        # chunks of pass A in the bulk exchange: the copies of a finished chunk overlap
        # the ranking of the next
        self.chunks = max(1, min(8, int(os.environ.get("NPGE_ANCHOR_CHUNKS", "4"))))

    def _peer_setup(self, capacity):
        """(Re)allocates the receive buffers and maps them on every rank.
        Collective.  Falls back to the all-to-all when a rank cannot map a peer
        (only when exchange was left to "auto")."""
        st = self.stages
        st.dist_peer_close()
        self.comm.barrier()  # nobody still has the old buffers mapped
        handle = st.dist_peer_alloc(capacity)
        handles = np.concatenate(self.comm.gather_objects(handle))
        err = None
        try:
            st.dist_peer_open(self.world, self.rank, handles)
        except AnchorError as e:
            err = str(e)
        errs = [e for e in self.comm.gather_objects(err) if e]
        if errs:
            st.dist_peer_close()
            if self._want in ("peer", "bulk"):
                raise AnchorError(_abi.ERR_CUDA, errs[0])
            self.exchange = "collective"
            return
        self.exchange = self._want if self._want != "auto" else "bulk"
        self._capacity = capacity

    # -- upload ------------------------------------------------------------------------
    def _rebalance(self, bases, offsets, k):
        """Gives every rank the bases of ITS range of the job-wide position space
        (sequences in scan order, one pad after each, cut into `world` equal ranges;
        plus the k - 1 bases after the range: the halo), wherever the sequences were
        loaded: one all-to-all of bytes (on the device with NCCL).  Returns
        (buffer, pieces, range, gstart, job_caller): `pieces` = arrays (src_off,
        dst_pos, len) placing parts of `buffer` at local positions; range = (g0, n_own,
        n_have)."""
        torch = self.comm.torch
        off = np.asarray(offsets, dtype=np.int64)
        dev = torch.device("cuda", self.device) if self.comm.nccl else None
        sizes_by_rank = self.comm.gather_varlen(np.diff(off), dev)
        counts = [len(x) for x in sizes_by_rank]
        base = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        sizes = np.asarray([x for r in sizes_by_rank for x in r], dtype=np.int64)
        origin = np.repeat(np.arange(self.world), counts)
        order = scan_order(sizes, k)              # job-wide caller indices in scan order
        size_at = sizes[order] if len(order) else np.zeros(0, np.int64)
        gstart = global_starts(size_at).astype(np.int64)
        total = int(gstart[-1])
        cuts = position_cuts(total, self.world)
        have_end = np.minimum(cuts[1:] + (k - 1), total)  # own range + halo
        me = self.rank
        src_of = origin[order] if len(order) else np.zeros(0, np.int64)
        # what rank d needs: pieces of [cuts[d], have_end[d]), in scan order
        need = [range_pieces(gstart, int(cuts[d]), int(have_end[d])) for d in range(self.world)]
        send_spans, in_splits = [], []
        for d in range(self.world):
            n = 0
            for (i, x, y, g) in need[d]:
                if src_of[i] == me:
                    local = int(order[i] - base[me])   # my caller index of the sequence
                    send_spans.append((int(off[local]) + x, y - x))
                    n += y - x
            in_splits.append(n)
        recv_from = [[(i, x, y, g) for (i, x, y, g) in need[me] if src_of[i] == s_]
                     for s_ in range(self.world)]
        out_splits = [sum(y - x for (_, x, y, _) in lst) for lst in recv_from]
        on_device = self.comm.nccl and isinstance(self.stages, CudaStages)
        src = torch.from_numpy(np.ascontiguousarray(bases))
        if on_device:
            dev = torch.device("cuda", self.device)
            src = src.to(dev, non_blocking=True)
        spans = [(a_, n) for a_, n in send_spans if n]
        if not spans:
            send = src[:0]
        elif all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(len(spans) - 1)):
            send = src[spans[0][0]:spans[-1][0] + spans[-1][1]]   # already contiguous
        else:
            send = torch.cat([src[a_:a_ + n] for a_, n in spans])
        recv = torch.empty(sum(out_splits), dtype=torch.uint8, device=src.device)
        if on_device:
            self.comm.all_to_all(recv, send.contiguous(), out_splits, in_splits)
            torch.cuda.current_stream(dev).synchronize()  # the library reads it on its own stream
        else:
            self.comm._a2a_host(recv, send.contiguous(), out_splits, in_splits)
            recv = recv.numpy()
        g0 = int(cuts[me])
        src_off, dst_pos, lens, at = [], [], [], 0
        for lst in recv_from:           # the receive buffer: source by source, scan order
            for (i, x, y, g) in lst:
                src_off.append(at)
                dst_pos.append(g - g0)
                lens.append(y - x)
                at += y - x
        pieces = tuple(np.asarray(v, dtype=np.uint64) for v in (src_off, dst_pos, lens))
        rng = (g0, int(cuts[me + 1]) - g0, int(have_end[me]) - g0)
        return recv, pieces, rng, gstart.astype(np.uint64), order

    def upload(self, seqs, k):
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        buf, pieces, rng, gstart, job_caller = self._rebalance(bases, offsets, k)
        self.n_job_bases = int(gstart[-1]) - len(job_caller)
        self.stages.dist_upload_range(k, buf, pieces, rng, gstart, job_caller)
        self._k = k
        if self._want != "collective":
            # room for an uneven split of the job's windows; grown on demand
            need = int(gstart[-1]) // self.world * 5 // 4 + 65536
            if need > self._capacity or self.exchange == "collective":
                self._peer_setup(need)

    # -- find --------------------------------------------------------------------------
    def _timed(self, name, fn):
        """Device time of a collective (CUDA events when the tensors are on a GPU)."""
        torch = self.comm.torch
        if not (self.comm.nccl and torch.cuda.is_available()):
            fn()
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        self.exchange_ms[name] = self.exchange_ms.get(name, 0.0) + e0.elapsed_time(e1)

    def find_device(self, k=None, similar=True, max_frags=100000, used=None,
                    job_stats=True):
        """All stages up to the anchors in device memory.  Returns the job's
        fragment count + this rank's stage times; with job_stats also the job's
        block count and digest (one more small all-gather; job_stats() does it
        later)."""
        import time
        k = self._k if k is None else k
        if k != self._k:
            raise AnchorError(_abi.ERR_STATE, "anchor-size differs from the uploaded one")
        st, comm = self.stages, self.comm
        params = _abi.Params(k, int(bool(similar)), int(max_frags))
        self.exchange_ms = {}
        trace, t_prev = {}, [time.perf_counter()]

        def lap(name):  # host wall clock of every phase of the step, ms
            now = time.perf_counter()
            trace[name] = trace.get(name, 0.0) + 1e3 * (now - t_prev[0])
            t_prev[0] = now

        dev = comm.torch.device("cuda", self.device) if comm.nccl else None
        fused = self.exchange == "peer"
        bulk = self.exchange == "bulk"
        if fused or bulk:
            chunks = self.chunks if bulk else 1
            if getattr(st, "_chunks", 1) != chunks:
                st.dist_set_chunks(chunks)
            hist = st.dist_count_a(params)
            if bulk:  # pass A runs while the histograms are gathered
                st.dist_scatter_local(params)
            lap("count_a")
        else:
            hist = st.dist_pass_a(params)
            lap("pass_a")
        # (also the step's barrier: no rank writes into a receive buffer whose
        # owner has not finished its previous step)
        hist_all = comm.gather_array(hist, dev)
        lap("gather_hist")
        if bulk:
            # tables and copies in the library (csrc/dist_plan.h): no numpy on the path
            try:
                n_recv = st.dist_exchange(params, hist_all, self.world, self.rank)
            except AnchorError as e:
                if e.code != _abi.ERR_CAPACITY:
                    raise
                # the same decision on every rank (same histograms, same capacity)
                self._peer_setup(max(e.n_recv) * 9 // 8 + 65536)
                if self.exchange == "collective":
                    raise AnchorError(_abi.ERR_CUDA, "peer buffers could not be regrown")
                n_recv = st.dist_exchange(params, hist_all, self.world, self.rank)
            lap("exchange")
            self._timed("barrier", lambda: comm.barrier(dev))  # every copy has landed
            lap("barrier")
            totals = st.dist_sort_planned(params, 0, used)
            lap("sort")
            if similar:
                found = st.dist_buffers(0, n_recv[self.rank])["found"]
                self._timed("found_bits", lambda: comm.sum_(found))
                lap("reduce_found")
                totals = st.dist_sort_planned(params, 1, used)
                lap("count")
        else:
            cuts = bucket_ranges(hist_all, self.world)
            tables = segment_tables(hist_all, cuts, self.rank)
            if fused:
                dest_rank, dest_off, n_recv = peer_destinations(hist_all, cuts, self.rank)
                lap("tables")
                if max(n_recv) > self._capacity:  # the same decision on every rank
                    self._peer_setup(max(n_recv) * 9 // 8 + 65536)
                    if self.exchange == "collective":
                        raise AnchorError(_abi.ERR_CUDA, "peer buffers could not be regrown")
                st.dist_scatter_peer(params, dest_rank, dest_off)
                lap("scatter_peer")
                self._timed("barrier", lambda: comm.barrier(dev))  # every scatter has landed
                lap("barrier")
                buf = st.dist_buffers(0, tables["n_recv"])
            else:
                send_splits, recv_splits = exchange_splits(hist_all, cuts, self.rank)
                lap("tables")
                buf = st.dist_buffers(sum(send_splits), tables["n_recv"])
                self._timed("records", lambda: comm.all_to_all(
                    buf["recv_rec"], buf["send_rec"], recv_splits, send_splits))
                if buf["send_hi"] is not None:
                    self._timed("records", lambda: comm.all_to_all(
                        buf["recv_hi"], buf["send_hi"], recv_splits, send_splits))
                lap("all_to_all")
            totals = st.dist_sort(params, 0, tables, used)
            lap("sort")
            if similar:
                self._timed("found_bits", lambda: comm.sum_(buf["found"]))
                lap("reduce_found")
                totals = st.dist_sort(params, 1, tables, used)
                lap("count")
        n_windows = int(np.asarray(hist_all, dtype=np.int64).sum())
        if int(max_frags) >= n_windows:
            # max-anchor-fragments cannot bite (every rank sees the same histograms): no
            # rank needs the others' totals, one rendezvous less in the step
            s = st.dist_emit(params, int(max_frags))
            lap("emit")
            n_frag_job = n_blocks_job = total_selected = None  # job_stats() fills them
        else:
            tot_all = comm.gather_array(np.asarray(totals, dtype=np.int64), dev)
            lap("gather_totals")
            limits = local_limits(tot_all[:, 0], int(max_frags))
            s = st.dist_emit(params, limits[self.rank])
            lap("emit")
            total_selected = int(tot_all[:, 0].sum())
            n_frag_job = int(sum(limits))
            n_blocks_job = None if total_selected > int(max_frags) else int(tot_all[:, 1].sum())
        self.local_stats = s
        out = dict(s)
        out.update(
            n_fragments=n_frag_job, n_blocks=n_blocks_job,
            n_windows=n_windows, digest=None,
            n_positions=self.n_job_bases, local_fragments=s["n_fragments"],
            local_records=s["n_windows"], total_selected=total_selected,
            exchange_ms=dict(self.exchange_ms), exchange=self.exchange,
            chunks=(self.chunks if bulk else 1), host_ms=trace)
        self._hist_all, self._hist = hist_all, hist
        self._last = out
        if job_stats:
            self.job_stats()
        return out

    def job_stats(self):
        """Completes the last find_device's dict with the job's block count,
        launches and digest (collective)."""
        s, out = self.local_stats, self._last
        dev = self.comm.torch.device("cuda", self.device) if self.comm.nccl else None
        d = np.asarray([s["n_blocks"], s["gpu_launches"], s.get("bytes_algorithmic", 0),
                        s.get("bytes_moved", 0), s["n_fragments"]], dtype=np.int64)
        dg = np.asarray([s["digest"][0], s["digest"][1]], dtype=np.uint64).view(np.int64)
        job = self.comm.gather_array(np.concatenate([d, dg]), dev)
        mask = (1 << 64) - 1
        # records this rank wrote into other ranks' HBM, and the device time of the
        # stage that moved them (bulk: the copies; peer: pass A itself)
        bins = len(self._hist) // max(1, getattr(self.stages, "_chunks", 1) if self.exchange == "bulk" else 1)
        cuts = bucket_ranges(np.asarray(self._hist_all).reshape(-1, bins), self.world)
        hist = np.asarray(self._hist, dtype=np.int64).reshape(-1, bins).sum(axis=0)
        kept = int(hist[int(cuts[self.rank]):int(cuts[self.rank + 1])].sum())
        out.update(bucket_cuts=[int(c) for c in cuts],
                   nvlink_bytes_sent=(int(hist.sum()) - kept) * (12 if self._k > 21 else 8),
                   # bulk: ms_pass[5] is what the copies add after the last chunk of pass A
                   nvlink_ms=(s["ms_pass"][5] if self.exchange == "bulk"
                              else s["ms_pass"][0] if self.exchange == "peer"
                              else out.get("exchange_ms", {}).get("records", 0.0)))
        out.update(n_blocks=int(job[:, 0].sum()), job_launches=int(job[:, 1].sum()),
                   bytes_algorithmic_job=int(job[:, 2].sum()),
                   bytes_moved_job=int(job[:, 3].sum()),
                   n_fragments=int(job[:, 4].sum()),
                   digest=(sum(int(x) & mask for x in job[:, 5]) & mask,
                           sum(int(x) & mask for x in job[:, 6]) & mask))
        return out

    def fetch(self):
        """This rank's slice of the job's anchors.  new_used is the job's list
        (every distinct hash but the job's first), the same on every rank."""
        r = self.stages.fetch()
        parts = self.comm.gather_objects((r.new_used, r.digest))
        alln = np.concatenate([np.asarray(p[0], dtype=np.uint64) for p in parts])
        r.new_used = alln[1:]
        mask = (1 << 64) - 1
        r.job_digest = (sum(p[1][0] for p in parts) & mask,
                        sum(p[1][1] for p in parts) & mask)
        return r

    def find_counts(self, seqs, k, similar=True, max_frags=100000, used=None):
        """find(), keeping only the sizes of this rank's host arrays (benchmarks)."""
        self.upload(seqs, k)
        stats = self.find_device(k, similar=similar, max_frags=max_frags, used=used)
        stats["result_bytes"] = self.stages.fetch_counts()["result_bytes"]
        return stats

    def find_compact_counts(self, seqs, k, similar=True, max_frags=100000, used=None):
        """find_counts() with the compact result layout: every rank's slice crosses
        PCIe as 4 bytes per fragment (job-wide position << 1 | strand)."""
        import time
        self.stages.set_result_layout(_abi.LAYOUT_COMPACT)
        try:
            t0 = time.perf_counter()
            self.upload(seqs, k)
            t1 = time.perf_counter()
            stats = self.find_device(k, similar=similar, max_frags=max_frags, used=used)
            t2 = time.perf_counter()
            stats["result_bytes"] = self.stages.fetch_compact_counts()["result_bytes"]
            t3 = time.perf_counter()
        finally:
            self.stages.set_result_layout(_abi.LAYOUT_FULL)
        # this rank's wall time by phase (upload = pieces to their owners + H2D + pack;
        # find = the step + the job's counts; fetch = this rank's slice over PCIe)
        stats["host_phases_ms"] = {"upload": 1e3 * (t1 - t0), "find_device": 1e3 * (t2 - t1),
                                   "fetch": 1e3 * (t3 - t2)}
        return stats

    def fetch_compact(self):
        """This rank's slice in the compact layout (after a find_device under
        stages.set_result_layout(LAYOUT_COMPACT)); decoded like AnchorResult."""
        return self.stages.fetch_compact()

    def find(self, seqs, k, similar=True, max_frags=100000, used=None):
        """upload + find_device + fetch (host buffers in, host arrays out)."""
        self.upload(seqs, k)
        stats = self.find_device(k, similar=similar, max_frags=max_frags, used=used)
        r = self.fetch()
        r.job = stats
        return r

    def gather(self, r):
        """The whole job's result on every rank (tests, small jobs)."""
        parts = self.comm.gather_objects(
            (r.block_offsets, r.block_hash, r.frag_seq, r.frag_min_pos, r.frag_ori))
        out = AnchorResult.__new__(AnchorResult)
        offs, base = [np.zeros(1, np.uint64)], 0
        for p in parts:
            offs.append(p[0][1:] + np.uint64(base))
            base += int(p[0][-1])
        out.block_offsets = np.concatenate(offs)
        out.block_hash = np.concatenate([p[1] for p in parts])
        out.frag_seq = np.concatenate([p[2] for p in parts])
        out.frag_min_pos = np.concatenate([p[3] for p in parts])
        out.frag_ori = np.concatenate([p[4] for p in parts])
        out.new_used = r.new_used
        out.digest = getattr(r, "job_digest", None)
        return out
