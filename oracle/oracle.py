"""ctypes loader for the CPU oracle (oracle/liboracle.so, oracle/_ref).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(npge_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class OracleParams(C.Structure):
    _fields_ = [
        ("anchor_size", C.c_int32),
        ("anchor_similar", C.c_int32),
        ("max_anchor_fragments", C.c_int64),
        ("anchor_fp", C.c_double),
        ("workers", C.c_int32),
        ("bloom_seed", C.c_uint32),
        ("exact_bloom", C.c_int32),
        ("asserts", C.c_int32),
    ]


class OracleResult(C.Structure):
    _fields_ = [
        ("n_blocks", C.c_uint64),
        ("n_fragments", C.c_uint64),
        ("block_offsets", C.POINTER(C.c_uint64)),
        ("block_hash", C.POINTER(C.c_uint64)),
        ("frag_seq", C.POINTER(C.c_uint32)),
        ("frag_min_pos", C.POINTER(C.c_uint32)),
        ("frag_ori", C.POINTER(C.c_int8)),
        ("n_new_used", C.c_uint64),
        ("new_used", C.POINTER(C.c_uint64)),
        ("seconds", C.c_double),
    ]


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    so = os.path.join(_HERE, "liboracle.so")
    if force or not os.path.exists(so) or (
            os.path.getmtime(so) < os.path.getmtime(
                os.path.join(_HERE, "anchor_oracle.cpp"))):
        subprocess.check_call(["make", "-C", _HERE, "all"],
                              stdout=subprocess.DEVNULL)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.oracle_make_hash.restype = C.c_uint64
        L.oracle_make_hash.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.oracle_reuse_hash.restype = C.c_uint64
        L.oracle_reuse_hash.argtypes = [C.c_uint64, C.c_int, C.c_char,
                                        C.c_char, C.c_int]
        L.oracle_complement_hash.restype = C.c_uint64
        L.oracle_complement_hash.argtypes = [C.c_uint64, C.c_int]
        L.oracle_to_atgcn.restype = C.c_size_t
        L.oracle_to_atgcn.argtypes = [C.c_char_p, C.c_size_t]
        L.oracle_bloom_optimal_bits.restype = C.c_size_t
        L.oracle_bloom_optimal_bits.argtypes = [C.c_size_t, C.c_double]
        L.oracle_bloom_optimal_hashes.restype = C.c_size_t
        L.oracle_bloom_optimal_hashes.argtypes = [C.c_size_t, C.c_size_t]
        L.oracle_find_faithful.restype = C.c_int
        L.oracle_find_faithful.argtypes = [
            C.POINTER(OracleParams), C.c_void_p, C.c_void_p, C.c_uint32,
            C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(OracleResult)]
        L.oracle_find_ideal.restype = C.c_int
        L.oracle_find_ideal.argtypes = [
            C.POINTER(OracleParams), C.c_void_p, C.c_void_p, C.c_uint32,
            C.c_void_p, C.c_uint64, C.POINTER(OracleResult)]
        L.oracle_result_free.argtypes = [C.POINTER(OracleResult)]
        L.oracle_blockset_hash.restype = C.c_uint64
        L.oracle_blockset_hash.argtypes = [
            C.POINTER(C.c_char_p), C.c_uint64, C.c_void_p, C.c_void_p,
            C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_digest.argtypes = [C.POINTER(OracleResult),
                                    C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


_ref = None


def ref_lib():
    """The reference's own hash headers compiled in place (oracle/_ref)."""
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libnpge_ref_hash.so")
        if not os.path.exists(path):
            build(force=True)
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        R.ref_make_hash.restype = C.c_uint64
        R.ref_make_hash.argtypes = [C.c_char_p, C.c_int, C.c_int]
        R.ref_reuse_hash.restype = C.c_uint64
        R.ref_reuse_hash.argtypes = [C.c_uint64, C.c_int, C.c_char, C.c_char,
                                     C.c_int]
        R.ref_complement_hash.restype = C.c_uint64
        R.ref_complement_hash.argtypes = [C.c_uint64, C.c_int]
        _ref = R
    return _ref


def make_hash(s, ori=1):
    """make_hash over the whole of `s` (str). ori=-1 reads from the last letter."""
    b = s.encode()
    buf = C.create_string_buffer(b, len(b))
    if ori == 1:
        return lib().oracle_make_hash(buf, len(b), 1)
    addr = C.addressof(buf) + len(b) - 1
    return lib().oracle_make_hash(C.cast(addr, C.c_char_p), len(b), -1)


def to_atgcn(s):
    b = s.encode()
    buf = C.create_string_buffer(b, len(b) + 1)
    n = lib().oracle_to_atgcn(buf, len(b))
    return buf.raw[:n].decode()


def concat(seqs):
    """list of str/bytes/uint8 arrays -> (uint8 array, uint64 offsets[S+1])"""
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


class Result:
    """Flat anchor result; the same shape as npge_b200.AnchorResult."""

    def __init__(self, r, free=True):
        nb, nf = int(r.n_blocks), int(r.n_fragments)
        self.block_offsets = np.ctypeslib.as_array(
            r.block_offsets, (nb + 1,)).copy()
        self.block_hash = (np.ctypeslib.as_array(r.block_hash, (nb,)).copy()
                           if nb else np.zeros(0, np.uint64))
        if nf:
            self.frag_seq = np.ctypeslib.as_array(r.frag_seq, (nf,)).copy()
            self.frag_min_pos = np.ctypeslib.as_array(
                r.frag_min_pos, (nf,)).copy()
            self.frag_ori = np.ctypeslib.as_array(r.frag_ori, (nf,)).copy()
        else:
            self.frag_seq = np.zeros(0, np.uint32)
            self.frag_min_pos = np.zeros(0, np.uint32)
            self.frag_ori = np.zeros(0, np.int8)
        nu = int(r.n_new_used)
        self.new_used = (np.ctypeslib.as_array(r.new_used, (nu,)).copy()
                         if nu else np.zeros(0, np.uint64))
        self.seconds = float(r.seconds)
        d = (C.c_uint64 * 2)()
        lib().oracle_digest(C.byref(r), d)
        self.digest = (int(d[0]), int(d[1]))
        if free:
            lib().oracle_result_free(C.byref(r))

    @property
    def n_blocks(self):
        return len(self.block_hash)

    @property
    def n_fragments(self):
        return len(self.frag_seq)

    def blocks(self):
        """Canonical form: sorted list of blocks, each a sorted
        tuple of (seq, min_pos, ori)."""
        out = []
        for b in range(self.n_blocks):
            lo, hi = int(self.block_offsets[b]), int(self.block_offsets[b + 1])
            out.append(tuple(sorted(
                (int(self.frag_seq[f]), int(self.frag_min_pos[f]),
                 int(self.frag_ori[f])) for f in range(lo, hi))))
        return sorted(out)


def _params(k, similar, max_frags, fp=0.1, workers=1, seed=1, exact_bloom=0,
            asserts=1):
    return OracleParams(k, int(similar), int(max_frags), fp, workers, seed,
                        int(exact_bloom), int(asserts))


def find_ideal(seqs, k, similar=False, max_frags=2**62, used=None):
    bases, offsets = concat(seqs)
    used = np.ascontiguousarray(
        np.sort(np.asarray(used if used is not None else [], dtype=np.uint64)))
    p = _params(k, similar, max_frags)
    r = OracleResult()
    rc = lib().oracle_find_ideal(
        C.byref(p), bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
        used.ctypes.data, len(used), C.byref(r))
    if rc != 0:
        raise ValueError("oracle_find_ideal failed: %d" % rc)
    return Result(r)


class OracleSummary(C.Structure):
    _fields_ = [("n_windows", C.c_uint64), ("n_fragments", C.c_uint64),
                ("n_blocks", C.c_uint64), ("total_selected", C.c_uint64),
                ("n_new_used", C.c_uint64), ("digest", C.c_uint64 * 2),
                ("seconds", C.c_double)]


def find_ideal_sliced(seqs, k, similar=False, max_frags=2**62, used=None,
                      slice_bits=6, threads=0):
    """find_ideal in 2**slice_bits hash-range slices: counts + digest only, O(N /
    slices) memory per thread.  seqs: list of sequences or (bases, offsets)."""
    bases, offsets = seqs if isinstance(seqs, tuple) else concat(seqs)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    used = np.ascontiguousarray(
        np.sort(np.asarray(used if used is not None else [], dtype=np.uint64)))
    p = _params(k, similar, max_frags)
    r = OracleSummary()
    L = lib()
    L.oracle_find_ideal_sliced.restype = C.c_int
    L.oracle_find_ideal_sliced.argtypes = [
        C.POINTER(OracleParams), C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
        C.c_uint64, C.c_int32, C.c_int32, C.POINTER(OracleSummary)]
    rc = L.oracle_find_ideal_sliced(
        C.byref(p), bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
        used.ctypes.data, len(used), slice_bits, threads, C.byref(r))
    if rc != 0:
        raise ValueError("oracle_find_ideal_sliced failed: %d" % rc)
    return dict(n_windows=int(r.n_windows), n_fragments=int(r.n_fragments),
                n_blocks=int(r.n_blocks), total_selected=int(r.total_selected),
                n_new_used=int(r.n_new_used),
                digest=(int(r.digest[0]), int(r.digest[1])), seconds=float(r.seconds))


def find_faithful(seqs, k, similar=True, max_frags=100000, fp=0.1, workers=1,
                  seed=1, exact_bloom=False, asserts=True, genome_ids=None,
                  used=None):
    bases, offsets = concat(seqs)
    used = np.ascontiguousarray(
        np.sort(np.asarray(used if used is not None else [], dtype=np.uint64)))
    gid = None
    if genome_ids is not None:
        gid = np.ascontiguousarray(np.asarray(genome_ids, dtype=np.int32))
    p = _params(k, similar, max_frags, fp, workers, seed, exact_bloom, asserts)
    r = OracleResult()
    rc = lib().oracle_find_faithful(
        C.byref(p), bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
        gid.ctypes.data if gid is not None else None,
        used.ctypes.data, len(used), C.byref(r))
    if rc != 0:
        raise ValueError("oracle_find_faithful failed: %d" % rc)
    return Result(r)


def blockset_hash(names, blocks):
    """blocks: list of lists of (seq_idx, min_pos, max_pos, ori)."""
    offs = np.zeros(len(blocks) + 1, dtype=np.uint64)
    seq, mn, mx, ori = [], [], [], []
    for i, b in enumerate(blocks):
        for (s, a, z, o) in b:
            seq.append(s)
            mn.append(a)
            mx.append(z)
            ori.append(o)
        offs[i + 1] = len(seq)
    seq = np.asarray(seq, dtype=np.uint32)
    mn = np.asarray(mn, dtype=np.uint32)
    mx = np.asarray(mx, dtype=np.uint32)
    ori = np.asarray(ori, dtype=np.int8)
    arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
    return lib().oracle_blockset_hash(
        arr, len(blocks), offs.ctypes.data, seq.ctypes.data, mn.ctypes.data,
        mx.ctypes.data, ori.ctypes.data)


def block_hashes(names, offsets, seq, min_pos, max_pos, ori):
    """Reference block_hash of every block of a flat block list (numpy arrays:
    offsets u64 [B+1], seq / min_pos / max_pos u32 [F], ori i8 [F])."""
    offs = np.ascontiguousarray(offsets, dtype=np.uint64)
    seq = np.ascontiguousarray(seq, dtype=np.uint32)
    mn = np.ascontiguousarray(min_pos, dtype=np.uint32)
    mx = np.ascontiguousarray(max_pos, dtype=np.uint32)
    ori = np.ascontiguousarray(ori, dtype=np.int8)
    out = np.zeros(len(offs) - 1, dtype=np.uint64)
    arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
    L = lib()
    L.oracle_block_hashes.restype = None
    L.oracle_block_hashes.argtypes = [C.POINTER(C.c_char_p), C.c_uint64] + [C.c_void_p] * 6
    L.oracle_block_hashes(arr, len(offs) - 1, offs.ctypes.data, seq.ctypes.data,
                          mn.ctypes.data, mx.ctypes.data, ori.ctypes.data,
                          out.ctypes.data)
    return out


def fragment_find(seqs, pattern, max_matches=100):
    """FragmentFinder restatement: list of (seq, min_pos, ori), in the order of
    a one-worker run of the reference, cut to max_matches."""
    bases, offsets = concat(seqs)
    p = to_atgcn(pattern).encode()
    n = C.c_uint64()
    ps, pp, po = C.POINTER(C.c_uint32)(), C.POINTER(C.c_uint32)(), C.POINTER(C.c_int8)()
    L = lib()
    L.oracle_fragment_find.argtypes = [
        C.c_char_p, C.c_uint32, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint32,
        C.POINTER(C.c_uint64), C.POINTER(C.POINTER(C.c_uint32)),
        C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.POINTER(C.c_int8))]
    L.oracle_free.argtypes = [C.c_void_p]
    rc = L.oracle_fragment_find(p, len(p), int(max_matches), bases.ctypes.data,
                                offsets.ctypes.data, len(offsets) - 1, C.byref(n),
                                C.byref(ps), C.byref(pp), C.byref(po))
    if rc != 0:
        raise ValueError("oracle_fragment_find failed: %d" % rc)
    out = [(int(ps[i]), int(pp[i]), int(po[i])) for i in range(n.value)]
    for q in (ps, pp, po):
        L.oracle_free(q)
    return out
