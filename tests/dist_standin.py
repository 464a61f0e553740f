"""numpy stand-in for the per-rank device stages of npge_b200.dist (TEST
INFRASTRUCTURE): lets the world_size-2 gloo tests drive the host logic of the
multi-GPU path (sequence rebalance, bucket ranges, exchange splits, segment
tables, found-bitmap reduction, limits, new_used) without a GPU.  It restates
oracle/anchor_oracle.cpp find_ideal on records, stage by stage; it is never
imported by the product package."""
import numpy as np
import torch

BINS = 256
_CODE = np.full(256, 255, dtype=np.uint8)
for _c, _v in zip(b"ATGC", range(4)):
    _CODE[_c] = _v


def _windows(seq, k):
    """(hash, pos, direct) of every N-free window of one sequence."""
    code = _CODE[np.asarray(seq, dtype=np.uint8)]
    n = len(code) - k + 1
    if n <= 0:
        return (np.zeros(0, np.uint64),) * 2 + (np.zeros(0, bool),)
    bad = (code > 3).astype(np.int64)
    csum = np.concatenate([[0], np.cumsum(bad)])
    ok = (csum[k:] - csum[:-k]) == 0
    c = (code & 3).astype(np.uint64)
    d = np.zeros(n, dtype=np.uint64)
    r = np.zeros(n, dtype=np.uint64)
    for j in range(k):
        d |= c[j:j + n] << np.uint64(2 * j)
        r |= (c[j:j + n] ^ np.uint64(1)) << np.uint64(2 * (k - 1 - j))
    h = np.minimum(d, r)
    pos = np.flatnonzero(ok)
    return h[pos], pos.astype(np.uint64), (h == d)[pos]


class Result:
    def __init__(self):
        self.block_offsets = np.zeros(1, np.uint64)
        self.block_hash = np.zeros(0, np.uint64)
        self.frag_seq = np.zeros(0, np.uint32)
        self.frag_min_pos = np.zeros(0, np.uint32)
        self.frag_ori = np.zeros(0, np.int8)
        self.new_used = np.zeros(0, np.uint64)
        self.digest = (0, 0)


class NumpyStages:
    def dist_upload_range(self, k, bases, pieces, rng, global_start, global_caller):
        """This rank's range of the job-wide position space: local position p is
        job-wide position g0 + p; positions no piece covers are pads."""
        self.k = k
        g0, n_own, n_have = (int(x) for x in rng)
        buf = np.asarray(bases, dtype=np.uint8)
        local = np.full(n_have, ord("N"), dtype=np.uint8)
        for so, dp, ln in zip(*pieces):
            so, dp, ln = int(so), int(dp), int(ln)
            assert dp + ln <= n_have
            local[dp:dp + ln] = buf[so:so + ln]
        self.local, self.g0, self.n_own = local, g0, n_own
        self.gstart = np.asarray(global_start, dtype=np.uint64)
        self.gcaller = np.asarray(global_caller, dtype=np.uint32)
        # the range must be what the job-wide layout says: pads exactly at the pads
        gs = self.gstart.astype(np.int64)
        pads = gs[1:] - 1
        pads = pads[(pads >= g0) & (pads < g0 + n_have)] - g0
        assert np.all(local[pads] == ord("N"))

    def dist_pass_a(self, params):
        k = self.k
        # the pads keep windows from crossing a sequence end; windows starting in the
        # halo belong to the next rank
        h, pos, d = _windows(self.local, k)
        own = pos < np.uint64(self.n_own)
        h, g, d = h[own], pos[own] + np.uint64(self.g0), d[own]
        shift = max(2 * k - 8, 0)
        a = (h >> np.uint64(shift)).astype(np.int64)
        order = np.argsort(a, kind="stable")
        self.send_h = h[order]
        self.send_p = (g[order] << np.uint64(1)) | d[order].astype(np.uint64)
        return np.bincount(a, minlength=BINS).astype(np.uint32)

    def dist_buffers(self, n_send, n_recv):
        assert n_send == len(self.send_h)
        words = int(self.gstart[-1]) // 32 + 8
        self.buf = dict(
            send_rec=torch.from_numpy(self.send_p.view(np.int64).copy()),
            send_hi=torch.from_numpy(self.send_h.view(np.int64).copy()),
            recv_rec=torch.zeros(n_recv, dtype=torch.int64),
            recv_hi=torch.zeros(n_recv, dtype=torch.int64),
            found=torch.zeros(words, dtype=torch.int32))
        return self.buf

    def _groups(self):
        h = self.h
        new = np.ones(len(h), dtype=bool)
        new[1:] = h[1:] != h[:-1]
        start = np.flatnonzero(new)
        return start, np.append(start[1:], len(h))

    def dist_sort(self, params, stage, t, used):
        used = set(int(x) for x in (used if used is not None else []))
        similar = bool(params.anchor_similar)
        if stage == 0:
            p = self.buf["recv_rec"].numpy().view(np.uint64)
            h = self.buf["recv_hi"].numpy().view(np.uint64)
            idx = np.concatenate(
                [np.arange(int(b), int(e)) for b, e in zip(t["seg_begin"], t["seg_end"])]
                + [np.zeros(0, np.int64)]).astype(np.int64)
            assert len(idx) == t["n_recv"] == len(set(idx.tolist()))
            # the tile tables must describe the same segments
            tiles = (t["seg_end"].astype(np.int64) - t["seg_begin"] + 4095) // 4096
            assert list(np.cumsum(tiles)) == list(t["seg_tile"][1:])
            for a in range(BINS):
                seg = np.flatnonzero(t["seg_a"] == a)
                got = int((t["seg_end"][seg].astype(np.int64) - t["seg_begin"][seg]).sum())
                if t["n_recv"]:
                    assert got == int(t["a_start"][a + 1]) - int(t["a_start"][a])
            p, h = p[idx], h[idx]
            order = np.argsort(h, kind="stable")  # source order = position order
            self.h, self.p = h[order], p[order]
            s, e = self._groups()
            for a, b in zip(s, e):  # inside a group the positions ascend
                assert np.all(np.diff(self.p[a:b].astype(np.int64)) > 0)
            rep = (e - s >= 2) & np.asarray(
                [int(x) not in used for x in self.h[s]], dtype=bool)
            self.rep, self.ok = rep, rep.copy()
            if similar:
                f = self.buf["found"].numpy().view(np.uint32)
                f[:] = 0
                for a, b in zip(s[rep], e[rep]):
                    g = (self.p[a + 1:b] >> np.uint64(1)).astype(np.int64)
                    np.bitwise_or.at(f, g >> 5, np.uint32(1) << (g & 31).astype(np.uint32))
                return 0, 0
        else:
            f = self.buf["found"].numpy().view(np.uint32)
            s, e = self._groups()
            g = (self.p >> np.uint64(1)).astype(np.int64)
            bit = lambda x: ((f[x >> 5] >> (x & 31).astype(np.uint32)) & 1).astype(bool)
            is_start = np.isin(g, self.gstart.astype(np.int64))
            cand = bit(g) & (is_start | ~bit(np.maximum(g - 1, 0)))
            self.ok = np.asarray([cand[a:b].any() for a, b in zip(s, e)], dtype=bool)
        s, e = self._groups()
        return int((e - s)[self.ok].sum()), int(self.ok.sum())

    def dist_emit(self, params, limit):
        s, e = self._groups()
        r = Result()
        offs, hashes, rows = [0], [], []
        for a, b in zip(s[self.ok], e[self.ok]):
            if offs[-1] + (b - a) > limit:
                assert offs[-1] == limit, "the stand-in cuts at block boundaries only"
                break
            rows.append(np.arange(a, b))
            offs.append(offs[-1] + int(b - a))
            hashes.append(self.h[a])
        rows = np.concatenate(rows) if rows else np.zeros(0, np.int64)
        g = self.p[rows] >> np.uint64(1)
        si = np.searchsorted(self.gstart, g, side="right") - 1
        r.block_offsets = np.asarray(offs, dtype=np.uint64)
        r.block_hash = np.asarray(hashes, dtype=np.uint64)
        r.frag_seq = self.gcaller[si].astype(np.uint32)
        r.frag_min_pos = (g - self.gstart[si]).astype(np.uint32)
        r.frag_ori = np.where(self.p[rows] & np.uint64(1), 1, -1).astype(np.int8)
        r.new_used = r.block_hash.copy()
        self.result = r
        return dict(n_fragments=len(rows), n_blocks=len(hashes), n_windows=len(self.h),
                    gpu_launches=0, digest=(0, 0))

    def fetch(self):
        return self.result
