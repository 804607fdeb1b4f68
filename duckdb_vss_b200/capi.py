"""ctypes declarations of the C ABI in include/gvss.h. Loading fails loudly if libgvss.so is missing."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GVSS_LIB") or os.path.join(HERE, "libgvss.so")  # GVSS_LIB: A/B builds of the same library


class GvssError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("dimensions", C.c_uint32), ("metric", C.c_char), ("connectivity", C.c_uint32),
                ("connectivity_base", C.c_uint32), ("expansion_add", C.c_uint32), ("expansion_search", C.c_uint32),
                ("device", C.c_int32)]


class LevelStats(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("edges", C.c_uint64), ("max_edges", C.c_uint64),
                ("allocated_bytes", C.c_uint64)]


class Info(C.Structure):
    _fields_ = [("size", C.c_uint64), ("capacity", C.c_uint64), ("node_count", C.c_uint64),
                ("max_level", C.c_uint64), ("memory_usage", C.c_uint64), ("dimensions", C.c_uint64),
                ("connectivity", C.c_uint32), ("connectivity_base", C.c_uint32), ("expansion_add", C.c_uint32),
                ("expansion_search", C.c_uint32), ("metric", C.c_char), ("concurrent_searches_peak", C.c_uint32)]


class BuildStats(C.Structure):
    _fields_ = [("inserts", C.c_uint64), ("query_distances", C.c_uint64), ("pair_distances", C.c_uint64),
                ("base_expansions", C.c_uint64), ("upper_expansions", C.c_uint64), ("rows_written", C.c_uint64)]


class SearchStats(C.Structure):
    _fields_ = [("computed_distances", C.c_uint32), ("visited_members", C.c_uint32),
                ("base_expansions", C.c_uint32), ("flags", C.c_uint32)]


WRITE_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t)
READ_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t)

# every symbol include/gvss.h declares (tests/test_abi.py checks the header against this table)
SIGNATURES = {
    "gvss_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "gvss_create_sharded": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_int32), C.c_uint32, C.POINTER(C.c_void_p)]),
    "gvss_destroy": (None, [C.c_void_p]),
    "gvss_reset": (C.c_int, [C.c_void_p]),
    "gvss_reserve": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32]),
    "gvss_add": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32]),
    "gvss_add_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "gvss_add_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "gvss_flush": (C.c_int, [C.c_void_p]),
    "gvss_remove": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_uint64)]),
    "gvss_compact": (C.c_int, [C.c_void_p]),
    "gvss_compact_rebuild": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "gvss_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p]),
    "gvss_search_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gvss_search_device_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gvss_save": (C.c_int, [C.c_void_p, WRITE_CB, C.c_void_p]),
    "gvss_load": (C.c_int, [C.c_void_p, READ_CB, C.c_void_p]),
    "gvss_serialized_length": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "gvss_get_info": (C.c_int, [C.c_void_p, C.POINTER(Info)]),
    "gvss_get_level_stats": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(LevelStats)]),
    "gvss_get_build_stats": (C.c_int, [C.c_void_p, C.POINTER(BuildStats)]),
    "gvss_exact_topk": (C.c_int, [C.c_char, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32,
                                  C.c_uint32, C.c_void_p, C.c_void_p]),
    "gvss_exact_topk_device": (C.c_int, [C.c_char, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64,
                                         C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gvss_exact_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gvss_merge_topk_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p,
                                         C.c_void_p, C.c_void_p]),
    "gvss_merge_topk_peers_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                               C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gvss_peer_exchange_status": (C.c_int, [C.POINTER(C.c_uint32)]),
    "gvss_set_build_concurrency": (C.c_int, [C.c_void_p, C.c_uint32]),
    "gvss_set_rows_only": (C.c_int, [C.c_void_p, C.c_int]),
    "gvss_warmup": (C.c_int, [C.c_int32]),
    "gvss_set_background_build": (C.c_int, [C.c_void_p, C.c_uint64]),
    "gvss_last_error": (C.c_char_p, []),
    "gvss_version": (C.c_char_p, []),
    "gvss_kernel_launches": (C.c_uint64, []),
}

ANY_THREAD = 0xFFFFFFFF  # GVSS_ANY_THREAD

_lib = None


def lib():
    """The loaded C-ABI library. Raises (no fallback) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GvssError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc:
        raise GvssError(lib().gvss_last_error().decode())
