"""ctypes binding of ``librad_b200.so`` (the C ABI declared in ``include/rad_b200.h``).

This is the only place the native library is loaded.  It fails loudly when the library is missing: there is no
Python or CPU fallback for the hot path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RAD_B200_LIB: development aid for A/B timing of two builds of the SAME library (never a different implementation)
LIB_PATH = os.environ.get("RAD_B200_LIB") or os.path.join(_HERE, "librad_b200.so")


class InitOptions(C.Structure):
    _fields_ = [
        ("dimensions", C.c_uint64),
        ("connectivity", C.c_uint32),
        ("expansion_add", C.c_uint32),
        ("expansion_search", C.c_uint32),
        ("device", C.c_int32),
        ("flags", C.c_uint32),
        ("batch_growth_den", C.c_uint32),
        ("batch_max", C.c_uint32),
        ("bloom_bits_log2", C.c_uint32),
        ("debug_max_batches", C.c_uint32),
        ("no_batch_tile", C.c_uint32),
        ("connectivity_base", C.c_uint32),
    ]


class LevelStats(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("edges", C.c_uint64), ("max_edges", C.c_uint64),
                ("allocated_bytes", C.c_uint64)]


class Metadata(C.Structure):
    _fields_ = [
        ("version_major", C.c_uint16), ("version_minor", C.c_uint16), ("version_patch", C.c_uint16),
        ("kind_metric", C.c_uint8), ("kind_scalar", C.c_uint8), ("kind_key", C.c_uint8),
        ("kind_compressed_slot", C.c_uint8),
        ("count_present", C.c_uint64), ("count_deleted", C.c_uint64), ("dimensions", C.c_uint64),
        ("multi", C.c_uint8), ("has_vectors", C.c_uint8),
        ("matrix_rows", C.c_uint64), ("matrix_cols", C.c_uint64),
    ]


class Counters(C.Structure):
    _fields_ = [("computed_distances", C.c_uint64), ("visited_members", C.c_uint64), ("batches", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("device_ms", C.c_double), ("insert_ms", C.c_double),
                ("sort_ms", C.c_double), ("reverse_ms", C.c_double), ("insert_launches", C.c_uint64),
                ("tile_ms", C.c_double), ("select_ms", C.c_double)]


PROGRESS = C.CFUNCTYPE(C.c_int, C.c_uint64, C.c_uint64, C.c_void_p)
RB_FLAG_LAZY_HOST_SYNC = 1
RB_FLAG_TIME_KERNELS = 2

_ERR = C.POINTER(C.c_char_p)
_H = C.c_void_p

# name -> (restype, argtypes); every symbol include/rad_b200.h declares
SIGNATURES = {
    "rb_version": (C.c_char_p, []),
    "rb_device_count": (C.c_int, []),
    "rb_init": (_H, [C.POINTER(InitOptions), _ERR]),
    "rb_free": (None, [_H]),
    "rb_release_device": (None, [_H, C.c_int, _ERR]),
    "rb_reserve": (None, [_H, C.c_size_t, _ERR]),
    "rb_add": (None, [_H, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, PROGRESS, C.c_void_p, _ERR]),
    "rb_add_device": (None, [_H, C.c_void_p, C.c_void_p, C.c_size_t, _ERR]),
    "rb_sync_host": (None, [_H, _ERR]),
    "rb_clear": (None, [_H, _ERR]),
    "rb_stream": (C.c_size_t, [_H]),
    "rb_search": (None, [_H, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                         C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), _ERR]),
    "rb_search_device": (None, [_H, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                _ERR]),
    "rb_exact_topk_device": (None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint32,
                                    C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, _ERR]),
    "rb_exact_topk_workspace_bytes": (C.c_size_t, [C.c_uint64, C.c_uint64, C.c_uint32]),
    "rb_exact_topk_noalloc_max_bytes": (C.c_size_t, []),
    "rb_exact_topk_device_noalloc": (None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint32,
                                            C.c_void_p, C.c_size_t, _ERR]),
    "rb_merge_topk_device": (None, [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_size_t, _ERR]),
    "rb_unpack_topk_device": (None, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_size_t, _ERR]),
    "rb_extract_level": (None, [C.c_void_p, C.c_void_p, _ERR]),
    "rb_dedup_workspace_bytes": (C.c_size_t, [C.c_uint64]),
    "rb_dedup_device": (None, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, _ERR]),
    "rb_popcounts_device": (None, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_size_t, _ERR]),
    "rb_sharded_search_workspace_bytes": (C.c_size_t, [C.c_size_t, C.c_size_t, C.c_uint32]),
    "rb_sharded_search_device": (None, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_size_t,
                                        C.c_void_p, C.c_void_p, C.c_size_t, _ERR]),
    "rb_cluster_device": (None, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, _ERR]),
    "rb_stable_match_workspace_bytes": (C.c_size_t, [C.c_uint64, C.c_uint64]),
    "rb_stable_match_device": (None, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.POINTER(C.c_uint64),
                                      C.POINTER(C.c_uint64), _ERR]),
    "rb_synth_fingerprints_device": (None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32,
                                            C.c_int, C.c_size_t, _ERR]),
    "rb_synth_fingerprints_device_ex": (None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32,
                                               C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_size_t,
                                               _ERR]),
    "rb_add_device_levels": (None, [_H, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, _ERR]),
    "rb_seek_level_stream": (None, [_H, C.c_uint64, _ERR]),
    "rb_gather_vectors_device": (None, [_H, C.c_void_p, C.c_uint64, C.c_void_p, _ERR]),
    "rb_copy_vectors_device": (None, [_H, C.c_uint64, C.c_uint64, C.c_void_p, _ERR]),
    "rb_stitch_device": (None, [_H, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                C.c_void_p, _ERR]),
    "rb_import_graph": (None, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                               C.c_uint64, _ERR]),
    "rb_save": (None, [_H, C.c_char_p, C.c_int, _ERR]),
    "rb_load": (None, [_H, C.c_char_p, C.c_int, _ERR]),
    "rb_view": (None, [_H, C.c_char_p, C.c_int, _ERR]),
    "rb_metadata": (None, [C.c_char_p, C.POINTER(Metadata), _ERR]),
    "rb_neighbors": (C.c_size_t, [_H, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, _ERR]),
    "rb_neighbors_many": (C.c_size_t, [_H, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p,
                                       _ERR]),
    "rb_top_level_nodes": (C.c_size_t, [_H, C.c_void_p, C.c_size_t, _ERR]),
    "rb_keys_from_slots": (None, [_H, C.c_void_p, C.c_size_t, C.c_void_p, _ERR]),
    "rb_slots_from_keys": (None, [_H, C.c_void_p, C.c_size_t, C.c_void_p, _ERR]),
    "rb_contains": (C.c_int, [_H, C.c_uint64]),
    "rb_get_vectors": (None, [_H, C.c_void_p, C.c_size_t, C.c_void_p, _ERR]),
    "rb_size": (C.c_size_t, [_H]),
    "rb_capacity": (C.c_size_t, [_H]),
    "rb_max_level": (C.c_int32, [_H]),
    "rb_entry_slot": (C.c_uint64, [_H]),
    "rb_connectivity": (C.c_size_t, [_H]),
    "rb_dimensions": (C.c_size_t, [_H]),
    "rb_memory_usage": (C.c_size_t, [_H]),
    "rb_serialized_length": (C.c_size_t, [_H, C.c_int]),
    "rb_expansion_add": (C.c_size_t, [_H]),
    "rb_expansion_search": (C.c_size_t, [_H]),
    "rb_set_expansion_add": (None, [_H, C.c_size_t]),
    "rb_set_expansion_search": (None, [_H, C.c_size_t]),
    "rb_level_stats": (None, [_H, C.c_size_t, C.POINTER(LevelStats), _ERR]),
    "rb_levels": (None, [_H, C.c_void_p, C.c_size_t]),
    "rb_last_counters": (None, [_H, C.POINTER(Counters)]),
    "rb_debug_adjacency": (C.c_uint64, [_H, C.c_void_p, C.c_void_p]),
}

_LIB = None


def lib() -> C.CDLL:
    """The loaded library.  Raises ImportError with build instructions when it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C rad_b200/csrc` (nvcc, sm_100a).  rad_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _LIB = L
    return _LIB


class NativeError(RuntimeError):
    pass


def check(err: C.c_char_p, value_error_prefixes=()) -> None:
    """Raise for a non-NULL error: ValueError for argument problems, RuntimeError otherwise
    (usearch/python/lib.cpp:232-256 -> std::invalid_argument -> ValueError; :216-220 -> RuntimeError)."""
    if err.value is None:
        return
    msg = err.value.decode("utf-8", "replace")
    for prefix in value_error_prefixes:
        if msg.startswith(prefix):
            raise ValueError(msg)
    raise RuntimeError(msg)
