"""ctypes binding of include/npge_anchor.h.  The library is required: if it is
missing or has no CUDA device the calls fail loudly (no CPU fallback)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NPGE_ANCHOR_LIB: another build of the same library (kernel experiments, build.py
# build_variant); the product default is the in-tree libnpge_anchor.so
LIB_PATH = os.environ.get("NPGE_ANCHOR_LIB") or os.path.join(_HERE, "libnpge_anchor.so")

EXPORTS = [
    "npge_anchor_create", "npge_anchor_destroy", "npge_anchor_last_error",
    "npge_anchor_find", "npge_anchor_result_free", "npge_anchor_upload",
    "npge_anchor_find_device", "npge_anchor_fetch", "npge_anchor_get_stats",
    "npge_anchor_version",
    "npge_anchor_dist_upload_range", "npge_anchor_dist_pass_a",
    "npge_anchor_dist_exchange", "npge_anchor_dist_sort_planned",
    "npge_anchor_dist_plan_tables", "npge_anchor_dist_set_chunks",
    "npge_anchor_dist_buffers", "npge_anchor_dist_sort", "npge_anchor_dist_emit",
    "npge_anchor_dist_peer_alloc", "npge_anchor_dist_peer_open",
    "npge_anchor_dist_peer_close", "npge_anchor_dist_count_a",
    "npge_anchor_dist_scatter_peer",
    "npge_fragment_find", "npge_fragment_find_device", "npge_fragment_result_free",
    "npge_anchor_upload_raw", "npge_anchor_dist_scatter_local",
    "npge_block_hashes", "npge_anchor_block_hashes", "npge_block_hashes_last_ms",
    "npge_anchor_debug_plan", "npge_anchor_debug_plan_of", "npge_anchor_debug_check", "npge_fragment_find_device_noalloc",
    "npge_anchor_set_result_layout", "npge_anchor_fetch_compact", "npge_anchor_find_compact",
    "npge_anchor_compact_result_free",
    "npge_anchor_dist_peer_attach", "npge_anchor_dist_found_or", "npge_anchor_dist_found_swap",
    "npge_anchor_multi_create", "npge_anchor_multi_destroy", "npge_anchor_multi_devices",
    "npge_anchor_multi_find", "npge_anchor_multi_upload", "npge_anchor_multi_find_device",
    "npge_anchor_multi_fetch", "npge_anchor_multi_result_free", "npge_anchor_multi_get_stats",
    "npge_anchor_multi_set_result_layout", "npge_anchor_multi_fetch_compact",
    "npge_anchor_multi_compact_result_free",
    "npge_anchor_multi_debug_check",
]
IPC_BYTES = 128

OK, ERR_PARAMS, ERR_CUDA, ERR_MEMORY, ERR_CAPACITY, ERR_INPUT, ERR_STATE = range(7)
LAYOUT_FULL, LAYOUT_COMPACT = 0, 1


class Params(C.Structure):
    _fields_ = [("anchor_size", C.c_int32), ("anchor_similar", C.c_int32),
                ("max_anchor_fragments", C.c_int64)]


class Result(C.Structure):
    _fields_ = [
        ("n_blocks", C.c_uint64), ("n_fragments", C.c_uint64),
        ("block_offsets", C.POINTER(C.c_uint64)),
        ("block_hash", C.POINTER(C.c_uint64)),
        ("frag_seq", C.POINTER(C.c_uint32)),
        ("frag_min_pos", C.POINTER(C.c_uint32)),
        ("frag_ori", C.POINTER(C.c_int8)),
        ("n_new_used", C.c_uint64), ("new_used", C.POINTER(C.c_uint64)),
        ("digest", C.c_uint64 * 2), ("owner_", C.c_void_p),
    ]


class CompactResult(C.Structure):
    _fields_ = [
        ("n_blocks", C.c_uint64), ("n_fragments", C.c_uint64),
        ("block_offsets", C.POINTER(C.c_uint64)),
        ("block_hash", C.POINTER(C.c_uint64)),
        ("frag_pos_ori", C.POINTER(C.c_uint32)),
        ("n_layout", C.c_uint32), ("anchor_size", C.c_int32),
        ("layout_start", C.POINTER(C.c_uint64)),
        ("layout_seq", C.POINTER(C.c_uint32)),
        ("n_new_used", C.c_uint64), ("new_used", C.POINTER(C.c_uint64)),
        ("digest", C.c_uint64 * 2), ("owner_", C.c_void_p),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("n_positions", C.c_uint64), ("n_windows", C.c_uint64),
        ("n_fragments", C.c_uint64), ("n_blocks", C.c_uint64),
        ("digest", C.c_uint64 * 2), ("radix_passes", C.c_uint32),
        ("gpu_launches", C.c_uint32), ("ms_hist", C.c_float),
        ("ms_sort", C.c_float), ("ms_group", C.c_float),
        ("ms_total", C.c_float), ("ms_pass", C.c_float * 16),
        ("bytes_algorithmic", C.c_uint64), ("bytes_moved", C.c_uint64),
    ]


class FragmentResult(C.Structure):
    _fields_ = [
        ("n_matches", C.c_uint64), ("total_matches", C.c_uint64),
        ("frag_seq", C.POINTER(C.c_uint32)), ("frag_min_pos", C.POINTER(C.c_uint32)),
        ("frag_ori", C.POINTER(C.c_int8)), ("owner_", C.c_void_p),
    ]


class DistPtrs(C.Structure):
    _fields_ = [
        ("send_rec", C.c_void_p), ("send_hi", C.c_void_p),
        ("recv_rec", C.c_void_p), ("recv_hi", C.c_void_p),
        ("found_bits", C.c_void_p), ("found_words", C.c_uint64),
        ("wide", C.c_int32), ("n_bins", C.c_int32),
    ]


_lib = None


def load():
    """Loads libnpge_anchor.so; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "npge_b200/libnpge_anchor.so is missing: run "
            "`python -m npge_b200.build` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    ctx = C.c_void_p
    L.npge_anchor_create.restype = C.c_int
    L.npge_anchor_create.argtypes = [C.POINTER(ctx), C.c_int]
    L.npge_anchor_destroy.restype = None
    L.npge_anchor_destroy.argtypes = [ctx]
    L.npge_anchor_last_error.restype = C.c_char_p
    L.npge_anchor_version.restype = C.c_char_p
    L.npge_anchor_find.restype = C.c_int
    L.npge_anchor_find.argtypes = [ctx, C.POINTER(Params), C.c_void_p,
                                   C.c_void_p, C.c_uint32, C.c_void_p,
                                   C.c_uint64, C.POINTER(Result)]
    L.npge_anchor_result_free.restype = None
    L.npge_anchor_result_free.argtypes = [C.POINTER(Result)]
    L.npge_anchor_upload.restype = C.c_int
    L.npge_anchor_upload.argtypes = [ctx, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_uint32]
    L.npge_anchor_find_device.restype = C.c_int
    L.npge_anchor_find_device.argtypes = [ctx, C.POINTER(Params), C.c_void_p,
                                          C.c_uint64, C.POINTER(Stats)]
    L.npge_anchor_fetch.restype = C.c_int
    L.npge_anchor_fetch.argtypes = [ctx, C.POINTER(Result)]
    L.npge_anchor_set_result_layout.restype = C.c_int
    L.npge_anchor_set_result_layout.argtypes = [ctx, C.c_int32]
    L.npge_anchor_fetch_compact.restype = C.c_int
    L.npge_anchor_fetch_compact.argtypes = [ctx, C.POINTER(CompactResult)]
    L.npge_anchor_find_compact.restype = C.c_int
    L.npge_anchor_find_compact.argtypes = [ctx, C.POINTER(Params), C.c_void_p, C.c_void_p,
                                           C.c_uint32, C.c_void_p, C.c_uint64,
                                           C.POINTER(CompactResult)]
    L.npge_anchor_compact_result_free.restype = None
    L.npge_anchor_compact_result_free.argtypes = [C.POINTER(CompactResult)]
    L.npge_anchor_debug_check.restype = C.c_int
    L.npge_anchor_debug_check.argtypes = [ctx, C.POINTER(C.c_uint32)]
    L.npge_anchor_debug_plan_of.restype = C.c_int
    L.npge_anchor_debug_plan_of.argtypes = [C.c_int32, C.c_uint64] + [C.POINTER(C.c_int32)] * 3
    L.npge_anchor_get_stats.restype = C.c_int
    L.npge_anchor_get_stats.argtypes = [ctx, C.POINTER(Stats)]
    L.npge_anchor_dist_upload_range.restype = C.c_int
    L.npge_anchor_dist_upload_range.argtypes = [
        ctx, C.c_int32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
        C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32]
    L.npge_anchor_dist_exchange.restype = C.c_int
    L.npge_anchor_dist_exchange.argtypes = [ctx, C.POINTER(Params), C.c_void_p, C.c_uint32,
                                            C.c_uint32, C.c_void_p]
    L.npge_anchor_dist_sort_planned.restype = C.c_int
    L.npge_anchor_dist_sort_planned.argtypes = [ctx, C.POINTER(Params), C.c_int32, C.c_void_p,
                                                C.c_uint64, C.POINTER(C.c_uint64 * 2)]
    L.npge_anchor_dist_plan_tables.restype = C.c_int
    L.npge_anchor_dist_plan_tables.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32,
                                               C.c_uint32, C.c_uint32] + [C.c_void_p] * 11
    L.npge_anchor_dist_set_chunks.restype = C.c_int
    L.npge_anchor_dist_set_chunks.argtypes = [ctx, C.c_int32]
    L.npge_anchor_dist_pass_a.restype = C.c_int
    L.npge_anchor_dist_pass_a.argtypes = [ctx, C.POINTER(Params), C.c_void_p,
                                          C.c_uint32, C.POINTER(C.c_uint32)]
    L.npge_anchor_dist_count_a.restype = C.c_int
    L.npge_anchor_dist_count_a.argtypes = L.npge_anchor_dist_pass_a.argtypes
    L.npge_anchor_dist_peer_alloc.restype = C.c_int
    L.npge_anchor_dist_peer_alloc.argtypes = [ctx, C.c_uint64, C.c_void_p]
    L.npge_anchor_dist_peer_open.restype = C.c_int
    L.npge_anchor_dist_peer_open.argtypes = [ctx, C.c_uint32, C.c_uint32, C.c_void_p]
    L.npge_anchor_dist_peer_close.restype = C.c_int
    L.npge_anchor_dist_peer_close.argtypes = [ctx]
    L.npge_anchor_dist_scatter_peer.restype = C.c_int
    L.npge_anchor_dist_scatter_peer.argtypes = [ctx, C.POINTER(Params), C.c_void_p,
                                                C.c_void_p, C.c_uint32]
    L.npge_anchor_dist_scatter_local.restype = C.c_int
    L.npge_anchor_dist_scatter_local.argtypes = [ctx, C.POINTER(Params)]
    L.npge_anchor_dist_buffers.restype = C.c_int
    L.npge_anchor_dist_buffers.argtypes = [ctx, C.c_uint64, C.POINTER(DistPtrs)]
    L.npge_anchor_dist_sort.restype = C.c_int
    L.npge_anchor_dist_sort.argtypes = [ctx, C.POINTER(Params), C.c_int32,
                                        C.c_uint64, C.c_uint32] + [C.c_void_p] * 7 + [
                                            C.c_uint64, C.POINTER(C.c_uint64 * 2)]
    L.npge_anchor_dist_emit.restype = C.c_int
    L.npge_anchor_dist_emit.argtypes = [ctx, C.POINTER(Params), C.c_uint64,
                                        C.POINTER(Stats)]
    L.npge_anchor_upload_raw.restype = C.c_int
    L.npge_anchor_upload_raw.argtypes = [ctx, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_uint32, C.c_void_p]
    L.npge_fragment_find.restype = C.c_int
    L.npge_fragment_find.argtypes = [ctx, C.c_char_p, C.c_uint32, C.c_int64, C.c_void_p,
                                     C.c_void_p, C.c_uint32, C.POINTER(FragmentResult)]
    L.npge_fragment_find_device.restype = C.c_int
    L.npge_fragment_find_device.argtypes = [ctx, C.c_char_p, C.c_uint32, C.c_int64,
                                            C.POINTER(FragmentResult)]
    L.npge_fragment_find_device_noalloc.restype = C.c_int
    L.npge_fragment_find_device_noalloc.argtypes = [
        ctx, C.c_char_p, C.c_uint32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
        C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.npge_fragment_result_free.restype = None
    L.npge_fragment_result_free.argtypes = [C.POINTER(FragmentResult)]
    L.npge_block_hashes.restype = C.c_int
    L.npge_block_hashes.argtypes = [ctx, C.POINTER(C.c_char_p), C.c_uint32, C.c_uint64,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_uint32, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_uint64)]
    L.npge_anchor_block_hashes.restype = C.c_int
    L.npge_anchor_block_hashes.argtypes = [ctx, C.POINTER(C.c_char_p), C.c_uint32,
                                           C.c_void_p, C.POINTER(C.c_uint64)]
    L.npge_block_hashes_last_ms.restype = C.c_float
    L.npge_block_hashes_last_ms.argtypes = [ctx]
    L.npge_anchor_multi_create.restype = C.c_int
    L.npge_anchor_multi_create.argtypes = [C.POINTER(ctx), C.c_void_p, C.c_int]
    L.npge_anchor_multi_destroy.restype = None
    L.npge_anchor_multi_destroy.argtypes = [ctx]
    L.npge_anchor_multi_devices.restype = C.c_int
    L.npge_anchor_multi_devices.argtypes = [ctx]
    L.npge_anchor_multi_find.restype = C.c_int
    L.npge_anchor_multi_find.argtypes = [ctx, C.POINTER(Params), C.c_void_p, C.c_void_p,
                                         C.c_uint32, C.c_void_p, C.c_uint64, C.POINTER(Result)]
    L.npge_anchor_multi_upload.restype = C.c_int
    L.npge_anchor_multi_upload.argtypes = [ctx, C.c_int32, C.c_void_p, C.c_void_p, C.c_uint32]
    L.npge_anchor_multi_find_device.restype = C.c_int
    L.npge_anchor_multi_find_device.argtypes = [ctx, C.POINTER(Params), C.c_void_p, C.c_uint64,
                                                C.POINTER(Stats)]
    L.npge_anchor_multi_fetch.restype = C.c_int
    L.npge_anchor_multi_fetch.argtypes = [ctx, C.POINTER(Result)]
    L.npge_anchor_multi_result_free.restype = None
    L.npge_anchor_multi_result_free.argtypes = [C.POINTER(Result)]
    L.npge_anchor_multi_set_result_layout.restype = C.c_int
    L.npge_anchor_multi_set_result_layout.argtypes = [ctx, C.c_int32]
    L.npge_anchor_multi_fetch_compact.restype = C.c_int
    L.npge_anchor_multi_fetch_compact.argtypes = [ctx, C.POINTER(CompactResult)]
    L.npge_anchor_multi_compact_result_free.restype = None
    L.npge_anchor_multi_compact_result_free.argtypes = [C.POINTER(CompactResult)]
    L.npge_anchor_multi_get_stats.restype = C.c_int
    L.npge_anchor_multi_get_stats.argtypes = [ctx, C.c_int, C.POINTER(Stats)]
    L.npge_anchor_multi_debug_check.restype = C.c_int
    L.npge_anchor_multi_debug_check.argtypes = [ctx, C.POINTER(C.c_uint32)]
    _lib = L
    return L
