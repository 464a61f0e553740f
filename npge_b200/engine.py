"""Thin Python driver over the C ABI (include/npge_anchor.h): used by the tests,
bench.py and the Processor mirror.  All compute happens in the CUDA library."""
import ctypes as C

import numpy as np

from . import _abi


class AnchorError(RuntimeError):
    """npge::Exception equivalent: carries the C ABI status code."""

    def __init__(self, code, message):
        RuntimeError.__init__(self, message)
        self.code = code


def concat(seqs):
    """list of str / bytes / uint8 arrays -> (uint8 array, uint64 offsets[S+1])"""
    arrs = []
    for s in seqs:
        if isinstance(s, str):
            s = s.encode()
        if isinstance(s, (bytes, bytearray)):
            s = np.frombuffer(bytes(s), dtype=np.uint8)
        arrs.append(np.asarray(s, dtype=np.uint8))
    offsets = np.zeros(len(arrs) + 1, dtype=np.uint64)
    if arrs:
        offsets[1:] = np.cumsum([len(a) for a in arrs])
        bases = np.concatenate(arrs) if len(arrs) > 1 else arrs[0].copy()
    else:
        bases = np.zeros(0, dtype=np.uint8)
    return np.ascontiguousarray(bases), offsets


class AnchorResult:
    """Flat anchor blocks: block b owns fragments block_offsets[b]:block_offsets[b+1]."""

    def __init__(self, r):
        nb, nf, nu = int(r.n_blocks), int(r.n_fragments), int(r.n_new_used)

        def arr(p, n, dtype):
            if n == 0:
                return np.zeros(0, dtype)
            return np.ctypeslib.as_array(p, (n,)).astype(dtype, copy=True)

        self.block_offsets = (arr(r.block_offsets, nb + 1, np.uint64)
                              if r.block_offsets else np.zeros(1, np.uint64))
        self.block_hash = arr(r.block_hash, nb, np.uint64)
        self.frag_seq = arr(r.frag_seq, nf, np.uint32)
        self.frag_min_pos = arr(r.frag_min_pos, nf, np.uint32)
        self.frag_ori = arr(r.frag_ori, nf, np.int8)
        self.new_used = arr(r.new_used, nu, np.uint64)
        self.digest = (int(r.digest[0]), int(r.digest[1]))

    @property
    def n_blocks(self):
        return len(self.block_hash)

    @property
    def n_fragments(self):
        return len(self.frag_seq)

    def blocks(self):
        """Canonical form: sorted list of blocks, each a sorted tuple of
        (seq, min_pos, ori)."""
        out = []
        for b in range(self.n_blocks):
            lo, hi = int(self.block_offsets[b]), int(self.block_offsets[b + 1])
            out.append(tuple(sorted(
                (int(self.frag_seq[f]), int(self.frag_min_pos[f]),
                 int(self.frag_ori[f])) for f in range(lo, hi))))
        return sorted(out)


class CompactAnchorResult(AnchorResult):
    """A compact result (npge_anchor_compact_result) decoded into the arrays of
    AnchorResult with the layout table, as npge_decode_compact does."""

    def __init__(self, r):
        nb, nf, nu, nl = int(r.n_blocks), int(r.n_fragments), int(r.n_new_used), int(r.n_layout)

        def arr(p, n, dtype):
            if n == 0:
                return np.zeros(0, dtype)
            return np.ctypeslib.as_array(p, (n,)).astype(dtype, copy=True)

        self.block_offsets = (arr(r.block_offsets, nb + 1, np.uint64)
                              if r.block_offsets else np.zeros(1, np.uint64))
        self.block_hash = arr(r.block_hash, nb, np.uint64)
        self.frag_pos_ori = arr(r.frag_pos_ori, nf, np.uint32)
        self.layout_start = arr(r.layout_start, nl + 1, np.uint64)
        self.layout_seq = arr(r.layout_seq, nl, np.uint32)
        self.new_used = arr(r.new_used, nu, np.uint64)
        self.digest = (int(r.digest[0]), int(r.digest[1]))
        g = (self.frag_pos_ori >> np.uint32(1)).astype(np.uint64)
        if nf:
            i = np.searchsorted(self.layout_start, g, side="right") - 1
            self.frag_seq = self.layout_seq[i].astype(np.uint32)
            self.frag_min_pos = (g - self.layout_start[i]).astype(np.uint32)
        else:
            self.frag_seq = np.zeros(0, np.uint32)
            self.frag_min_pos = np.zeros(0, np.uint32)
        self.frag_ori = np.where(self.frag_pos_ori & np.uint32(1), 1, -1).astype(np.int8)


def _stats_dict(s):
    d = {name: getattr(s, name) for name, _ in s._fields_
         if name not in ("digest", "ms_pass")}
    d["ms_pass"] = [float(x) for x in s.ms_pass][:8]
    d["digest"] = (int(s.digest[0]), int(s.digest[1]))
    return d


class AnchorEngine:
    """One CUDA context of the anchor finder (not thread-safe, like the
    reference's Processor instances)."""

    def __init__(self, device=0):
        self._lib = _abi.load()
        self._ctx = C.c_void_p()
        self._check(self._lib.npge_anchor_create(C.byref(self._ctx), device))

    def close(self):
        if self._ctx:
            self._lib.npge_anchor_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != _abi.OK:
            raise AnchorError(rc, self._lib.npge_anchor_last_error().decode())

    @staticmethod
    def _used(used):
        u = np.ascontiguousarray(np.sort(np.asarray(
            used if used is not None else [], dtype=np.uint64)))
        return u, (u.ctypes.data if len(u) else None)

    def debug_check(self):
        """npge_anchor_debug_check: number of device buffers whose guard bytes were
        overwritten (0 = no kernel stored past the end of a buffer)."""
        n = C.c_uint32(0)
        self._check(self._lib.npge_anchor_debug_check(self._ctx, C.byref(n)))
        if n.value:
            print(self._lib.npge_anchor_last_error().decode())
        return int(n.value)

    def debug_plan(self, fine_bits=0, first_split=0):
        """Test hook (npge_anchor_debug_plan): force the digit plan of a larger job."""
        self._check(self._lib.npge_anchor_debug_plan(self._ctx, fine_bits, first_split))

    def find(self, seqs, k, similar=True, max_frags=100000, used=None):
        """npge_anchor_find with host buffers (the drop-in call)."""
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        u, up = self._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        r = _abi.Result()
        self._check(self._lib.npge_anchor_find(
            self._ctx, C.byref(p), bases.ctypes.data, offsets.ctypes.data,
            len(offsets) - 1, up, len(u), C.byref(r)))
        try:
            return AnchorResult(r)
        finally:
            self._lib.npge_anchor_result_free(C.byref(r))

    def set_result_layout(self, layout):
        self._check(self._lib.npge_anchor_set_result_layout(self._ctx, layout))

    def fetch_compact(self):
        r = _abi.CompactResult()
        self._check(self._lib.npge_anchor_fetch_compact(self._ctx, C.byref(r)))
        try:
            return CompactAnchorResult(r)
        finally:
            self._lib.npge_anchor_compact_result_free(C.byref(r))

    def find_compact(self, seqs, k, similar=True, max_frags=100000, used=None, counts=False):
        """npge_anchor_find_compact with host buffers: the drop-in call with the
        4-bytes-per-fragment result.  counts=True keeps only the sizes (benchmarks)."""
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        u, up = self._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        r = _abi.CompactResult()
        self._check(self._lib.npge_anchor_find_compact(
            self._ctx, C.byref(p), bases.ctypes.data, offsets.ctypes.data,
            len(offsets) - 1, up, len(u), C.byref(r)))
        try:
            if counts:
                nb, nf = int(r.n_blocks), int(r.n_fragments)
                return {"n_blocks": nb, "n_fragments": nf,
                        "result_bytes": nf * 4 + (nb + 1) * 8 + nb * 8,
                        "digest": (int(r.digest[0]), int(r.digest[1]))}
            return CompactAnchorResult(r)
        finally:
            self._lib.npge_anchor_compact_result_free(C.byref(r))

    def find_counts(self, seqs, k, similar=True, max_frags=100000, used=None):
        """npge_anchor_find, but only the sizes are kept: the host arrays the
        library filled are released at once (benchmarks of the drop-in call,
        without the numpy copies AnchorResult makes)."""
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        u, up = self._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        r = _abi.Result()
        self._check(self._lib.npge_anchor_find(
            self._ctx, C.byref(p), bases.ctypes.data, offsets.ctypes.data,
            len(offsets) - 1, up, len(u), C.byref(r)))
        nb, nf = int(r.n_blocks), int(r.n_fragments)
        out = {"n_blocks": nb, "n_fragments": nf,
               "result_bytes": nf * 9 + (nb + 1) * 8 + nb * 8,
               "digest": (int(r.digest[0]), int(r.digest[1]))}
        self._lib.npge_anchor_result_free(C.byref(r))
        return out

    def upload(self, seqs, k):
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        self._check(self._lib.npge_anchor_upload(
            self._ctx, k, bases.ctypes.data, offsets.ctypes.data,
            len(offsets) - 1))

    def upload_raw(self, raw_seqs, k):
        """upload() of raw sequence text (FASTA body lines, any case, gaps, IUPAC
        codes): Sequence::to_atgcn runs on the device.  Returns the sizes of the
        cleaned sequences."""
        raw, offsets = raw_seqs if isinstance(raw_seqs, tuple) else concat(raw_seqs)
        sizes = np.zeros(len(offsets) - 1, dtype=np.uint64)
        self._check(self._lib.npge_anchor_upload_raw(
            self._ctx, k, raw.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
            sizes.ctypes.data if len(sizes) else None))
        return sizes

    def find_device(self, k, similar=True, max_frags=100000, used=None):
        u, up = self._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        s = _abi.Stats()
        self._check(self._lib.npge_anchor_find_device(
            self._ctx, C.byref(p), up, len(u), C.byref(s)))
        return _stats_dict(s)

    def fetch(self):
        r = _abi.Result()
        self._check(self._lib.npge_anchor_fetch(self._ctx, C.byref(r)))
        try:
            return AnchorResult(r)
        finally:
            self._lib.npge_anchor_result_free(C.byref(r))

    def fetch_counts(self):
        """npge_anchor_fetch, keeping only the sizes (see find_counts)."""
        r = _abi.Result()
        self._check(self._lib.npge_anchor_fetch(self._ctx, C.byref(r)))
        nb, nf = int(r.n_blocks), int(r.n_fragments)
        self._lib.npge_anchor_result_free(C.byref(r))
        return {"n_blocks": nb, "n_fragments": nf,
                "result_bytes": nf * 9 + (nb + 1) * 8 + nb * 8}

    def fetch_compact_counts(self):
        """npge_anchor_fetch_compact, keeping only the sizes and the digest."""
        r = _abi.CompactResult()
        self._check(self._lib.npge_anchor_fetch_compact(self._ctx, C.byref(r)))
        nb, nf = int(r.n_blocks), int(r.n_fragments)
        dg = (int(r.digest[0]), int(r.digest[1]))
        self._lib.npge_anchor_compact_result_free(C.byref(r))
        return {"n_blocks": nb, "n_fragments": nf, "digest": dg,
                "result_bytes": nf * 4 + (nb + 1) * 8 + nb * 8}

    # ---- FragmentFinder (src/algo/FragmentFinder.cpp) --------------------------------
    def _matches(self, r):
        n = int(r.n_matches)
        try:
            if n == 0:
                out = []
            else:
                seq = np.ctypeslib.as_array(r.frag_seq, (n,))
                pos = np.ctypeslib.as_array(r.frag_min_pos, (n,))
                ori = np.ctypeslib.as_array(r.frag_ori, (n,))
                out = list(zip(seq.tolist(), pos.tolist(), ori.tolist()))
            return out, int(r.total_matches)
        finally:
            self._lib.npge_fragment_result_free(C.byref(r))

    def find_fragments(self, seqs, pattern, max_matches=100, with_total=False):
        """Every occurrence of `pattern` (ATGCN text) on either strand:
        [(seq, min_pos, ori)] in scan order, cut to max_matches."""
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        p = pattern.encode() if isinstance(pattern, str) else bytes(pattern)
        r = _abi.FragmentResult()
        self._check(self._lib.npge_fragment_find(
            self._ctx, p, len(p), int(max_matches), bases.ctypes.data,
            offsets.ctypes.data, len(offsets) - 1, C.byref(r)))
        out, total = self._matches(r)
        return (out, total) if with_total else out

    def find_fragments_device(self, pattern, max_matches=100, with_total=False):
        """The same on the sequences of the last upload()."""
        p = pattern.encode() if isinstance(pattern, str) else bytes(pattern)
        r = _abi.FragmentResult()
        self._check(self._lib.npge_fragment_find_device(
            self._ctx, p, len(p), int(max_matches), C.byref(r)))
        out, total = self._matches(r)
        return (out, total) if with_total else out

    def find_fragments_device_noalloc(self, pattern, max_matches, d_seq, d_pos, d_ori,
                                      capacity):
        """npge_fragment_find_device_noalloc: the library allocates nothing; d_seq /
        d_pos / d_ori are DEVICE pointers (ints) to arrays of `capacity` elements the
        caller owns.  Returns (matches written, matches found)."""
        p = pattern.encode() if isinstance(pattern, str) else bytes(pattern)
        n, total = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.npge_fragment_find_device_noalloc(
            self._ctx, p, len(p), int(max_matches), d_seq, d_pos, d_ori, int(capacity),
            C.byref(n), C.byref(total)))
        return int(n.value), int(total.value)

    # ---- block_hash / blockset_hash (src/model/block_hash.cpp) -----------------------
    @staticmethod
    def _names(names):
        return (C.c_char_p * max(1, len(names)))(
            *[n.encode() if isinstance(n, str) else bytes(n) for n in names])

    def block_hashes(self, names, block_offsets, frag_seq, frag_min_pos, frag_ori,
                     frag_max_pos=None, frag_len=0):
        """The reference's block_hash of every block of a flat block list and
        their XOR over the blocks of more than one fragment (blockset_hash):
        (hashes u64 [B], blockset_hash).  frag_max_pos None: every fragment is
        frag_len long."""
        offs = np.ascontiguousarray(block_offsets, dtype=np.uint64)
        seq = np.ascontiguousarray(frag_seq, dtype=np.uint32)
        mn = np.ascontiguousarray(frag_min_pos, dtype=np.uint32)
        ori = np.ascontiguousarray(frag_ori, dtype=np.int8)
        mx = (np.ascontiguousarray(frag_max_pos, dtype=np.uint32)
              if frag_max_pos is not None else None)
        nb = len(offs) - 1
        out = np.zeros(max(nb, 1), dtype=np.uint64)
        total = C.c_uint64(0)
        self._check(self._lib.npge_block_hashes(
            self._ctx, self._names(names), len(names), nb, offs.ctypes.data,
            seq.ctypes.data, mn.ctypes.data, mx.ctypes.data if mx is not None else None,
            int(frag_len), ori.ctypes.data, out.ctypes.data, C.byref(total)))
        return out[:nb], int(total.value)

    def anchor_block_hashes(self, names, want_hashes=True):
        """The same over the device-resident result of the last find_device():
        only the hashes cross PCIe."""
        nb = int(self.stats()["n_blocks"])
        out = np.zeros(max(nb, 1), dtype=np.uint64) if want_hashes else None
        total = C.c_uint64(0)
        self._check(self._lib.npge_anchor_block_hashes(
            self._ctx, self._names(names), len(names),
            out.ctypes.data if want_hashes else None, C.byref(total)))
        return (out[:nb] if want_hashes else None), int(total.value)

    def block_hashes_ms(self):
        return float(self._lib.npge_block_hashes_last_ms(self._ctx))

    def stats(self):
        s = _abi.Stats()
        self._check(self._lib.npge_anchor_get_stats(self._ctx, C.byref(s)))
        return _stats_dict(s)


class MultiEngine:
    """AnchorFinder over several GPUs of this process (npge_anchor_multi_*): what
    npge::gpu::AnchorFinder::set_devices drives.  `devices` may repeat a device
    (several ranks on one GPU)."""

    def __init__(self, devices):
        self._lib = _abi.load()
        self._m = C.c_void_p()
        ids = (C.c_int * len(devices))(*devices)
        self._check(self._lib.npge_anchor_multi_create(C.byref(self._m), ids, len(devices)))

    def close(self):
        if self._m:
            self._lib.npge_anchor_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != _abi.OK:
            raise AnchorError(rc, self._lib.npge_anchor_last_error().decode())

    def upload(self, seqs, k):
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        self._check(self._lib.npge_anchor_multi_upload(
            self._m, k, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1))

    def find_device(self, k, similar=True, max_frags=100000, used=None):
        u, up = AnchorEngine._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        s = _abi.Stats()
        self._check(self._lib.npge_anchor_multi_find_device(
            self._m, C.byref(p), up, len(u), C.byref(s)))
        return _stats_dict(s)

    def fetch(self):
        r = _abi.Result()
        self._check(self._lib.npge_anchor_multi_fetch(self._m, C.byref(r)))
        try:
            return AnchorResult(r)
        finally:
            self._lib.npge_anchor_multi_result_free(C.byref(r))

    def set_result_layout(self, layout):
        self._check(self._lib.npge_anchor_multi_set_result_layout(self._m, int(layout)))

    def fetch_compact(self):
        r = _abi.CompactResult()
        self._check(self._lib.npge_anchor_multi_fetch_compact(self._m, C.byref(r)))
        try:
            return CompactAnchorResult(r)
        finally:
            self._lib.npge_anchor_multi_compact_result_free(C.byref(r))

    def find(self, seqs, k, similar=True, max_frags=100000, used=None):
        bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
        u, up = AnchorEngine._used(used)
        p = _abi.Params(k, int(bool(similar)), int(max_frags))
        r = _abi.Result()
        self._check(self._lib.npge_anchor_multi_find(
            self._m, C.byref(p), bases.ctypes.data, offsets.ctypes.data,
            len(offsets) - 1, up, len(u), C.byref(r)))
        try:
            return AnchorResult(r)
        finally:
            self._lib.npge_anchor_multi_result_free(C.byref(r))

    def debug_check(self):
        n = C.c_uint32(0)
        self._check(self._lib.npge_anchor_multi_debug_check(self._m, C.byref(n)))
        return int(n.value)

    def rank_stats(self, rank):
        s = _abi.Stats()
        self._check(self._lib.npge_anchor_multi_get_stats(self._m, rank, C.byref(s)))
        return _stats_dict(s)
