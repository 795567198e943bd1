"""ctypes front-end to ``librad_oracle.so`` (the plain-C restatement in ``rad_oracle.c``) and a loader for the
compiled, unmodified reference binding in ``oracle/_ref`` -- TEST INFRASTRUCTURE ONLY (see ``__init__``).

Reference anchors are listed function by function in ``rad_oracle.c``'s header.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB: Optional[C.CDLL] = None

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(dtype=np.int16, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    """Compile the C restatement (a second or two of gcc)."""
    so = os.path.join(_HERE, "librad_oracle.so")
    src = os.path.join(_HERE, "rad_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "librad_oracle.so"])
    return so


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.ro_tanimoto.restype = C.c_float
        L.ro_tanimoto.argtypes = [_u8p, _u8p, C.c_size_t]
        L.ro_levels.restype = None
        L.ro_levels.argtypes = [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint32), _i16p]
        L.ro_exact_search.restype = None
        L.ro_exact_search.argtypes = [_u8p, C.c_uint64, C.c_size_t, _u8p, C.c_uint64, C.c_size_t, _u64p, _f32p, _i64p]
        L.ro_index_new.restype = C.c_void_p
        L.ro_index_new.argtypes = [C.c_size_t, C.c_uint32, C.c_size_t, C.c_size_t]
        L.ro_index_free.restype = None
        L.ro_index_free.argtypes = [C.c_void_p]
        L.ro_index_reserve.restype = None
        L.ro_index_reserve.argtypes = [C.c_void_p, C.c_uint64]
        L.ro_index_add_many.restype = None
        L.ro_index_add_many.argtypes = [C.c_void_p, _u64p, _u8p, C.c_uint64]
        L.ro_index_search.restype = None
        L.ro_index_search.argtypes = [C.c_void_p, _u8p, C.c_uint64, C.c_size_t, _u64p, _u64p, _f32p, _i64p]
        L.ro_index_size.restype = C.c_uint64
        L.ro_index_size.argtypes = [C.c_void_p]
        L.ro_index_max_level.restype = C.c_int32
        L.ro_index_max_level.argtypes = [C.c_void_p]
        L.ro_index_entry.restype = C.c_uint64
        L.ro_index_entry.argtypes = [C.c_void_p]
        L.ro_index_set_ef.restype = None
        L.ro_index_set_ef.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t]
        L.ro_index_levels.restype = None
        L.ro_index_levels.argtypes = [C.c_void_p, _i16p]
        L.ro_index_counters.restype = None
        L.ro_index_counters.argtypes = [C.c_void_p, _u64p]
        L.ro_index_neighbors.restype = C.c_int64
        L.ro_index_neighbors.argtypes = [C.c_void_p, C.c_uint64, C.c_int, _u32p, C.c_size_t]
        L.ro_index_save.restype = C.c_int
        L.ro_index_save.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
        _LIB = L
    return _LIB


# --------------------------------------------------------------------------- numpy restatement (vectorised)

_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.uint16)


def tanimoto_matrix(queries: np.ndarray, db: np.ndarray) -> np.ndarray:
    """float32 [Q, N] distances, the f64 formula of simsimd binary.h:99-105 narrowed like index_plugins.hpp:1892."""
    q = np.ascontiguousarray(queries, dtype=np.uint8)
    d = np.ascontiguousarray(db, dtype=np.uint8)
    inter = _POP8[q[:, None, :] & d[None, :, :]].sum(axis=2, dtype=np.int64)
    union = _POP8[q[:, None, :] | d[None, :, :]].sum(axis=2, dtype=np.int64)
    out = np.ones(inter.shape, dtype=np.float64)
    nz = union != 0
    out[nz] = 1.0 - inter[nz].astype(np.float64) / union[nz].astype(np.float64)
    return out.astype(np.float32)


def exact_topk_numpy(queries: np.ndarray, db: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Reference ordering of Index.search(exact=True): ascending distance, ties by DESCENDING slot
    (index.hpp:923-934 -- lower_bound puts a later slot before equal distances)."""
    dist = tanimoto_matrix(queries, db)
    n = db.shape[0]
    slots = np.arange(n, dtype=np.int64)
    keys = np.empty((len(queries), min(k, n)), dtype=np.uint64)
    dd = np.empty((len(queries), min(k, n)), dtype=np.float32)
    for i in range(len(queries)):
        order = np.lexsort((-slots, dist[i]))[: min(k, n)]
        keys[i] = order
        dd[i] = dist[i, order]
    return keys, dd


def tanimoto(a: np.ndarray, b: np.ndarray) -> np.float32:
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    return np.float32(lib().ro_tanimoto(a, b, a.size))


def levels(n: int, connectivity: int, state: int = 1) -> Tuple[np.ndarray, int]:
    """The level sequence of a single-threaded reference build (index.hpp:3867-3871); returns (levels, new state)."""
    out = np.empty(n, dtype=np.int16)
    st = C.c_uint32(state)
    lib().ro_levels(n, connectivity, C.byref(st), out)
    return out, int(st.value)


def exact_search(db: np.ndarray, queries: np.ndarray, k: int):
    """(slots u64 [Q,k], distances f32 [Q,k], counts i64 [Q]) -- the reference's Index.search(exact=True)."""
    db = np.ascontiguousarray(db, dtype=np.uint8)
    queries = np.ascontiguousarray(queries, dtype=np.uint8)
    if queries.ndim == 1:
        queries = queries[None, :]
    nq = queries.shape[0]
    slots = np.zeros((nq, k), dtype=np.uint64)
    dists = np.zeros((nq, k), dtype=np.float32)
    counts = np.zeros(nq, dtype=np.int64)
    lib().ro_exact_search(db, db.shape[0], db.shape[1], queries, nq, k, slots, dists, counts)
    return slots, dists, counts


class OracleIndex:
    """Single-threaded restatement of usearch ``Index(dtype='b1', metric='tanimoto')`` (add / search / save)."""

    def __init__(self, ndim: int = 1024, connectivity: int = 16, expansion_add: int = 128, expansion_search: int = 64):
        assert ndim % 8 == 0
        self.ndim = ndim
        self.bytes = ndim // 8
        self.connectivity = connectivity
        self.expansion_add = expansion_add
        self.expansion_search = expansion_search
        self._h = C.c_void_p(lib().ro_index_new(self.bytes, connectivity, expansion_add, expansion_search))

    def __del__(self):
        try:
            if self._h:
                lib().ro_index_free(self._h)
                self._h = None
        except Exception:
            pass

    def __len__(self) -> int:
        return int(lib().ro_index_size(self._h))

    @property
    def max_level(self) -> int:
        return int(lib().ro_index_max_level(self._h))

    @property
    def entry_slot(self) -> int:
        return int(lib().ro_index_entry(self._h))

    def add(self, keys: np.ndarray, vectors: np.ndarray) -> None:
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        vectors = np.ascontiguousarray(vectors, dtype=np.uint8)
        assert vectors.shape == (len(keys), self.bytes)
        lib().ro_index_add_many(self._h, keys, vectors, len(keys))

    def search(self, queries: np.ndarray, k: int = 10, expansion: Optional[int] = None):
        queries = np.ascontiguousarray(queries, dtype=np.uint8)
        if queries.ndim == 1:
            queries = queries[None, :]
        if expansion is not None:
            lib().ro_index_set_ef(self._h, self.expansion_add, expansion)
        nq = queries.shape[0]
        keys = np.zeros((nq, k), dtype=np.uint64)
        slots = np.zeros((nq, k), dtype=np.uint64)
        dists = np.zeros((nq, k), dtype=np.float32)
        counts = np.zeros(nq, dtype=np.int64)
        lib().ro_index_search(self._h, queries, nq, k, keys, slots, dists, counts)
        return keys, slots, dists, counts

    def levels(self) -> np.ndarray:
        out = np.empty(len(self), dtype=np.int16)
        lib().ro_index_levels(self._h, out)
        return out

    def counters(self) -> dict:
        out = np.zeros(4, dtype=np.uint64)
        lib().ro_index_counters(self._h, out)
        return dict(computed_distances=int(out[0]), iteration_cycles=int(out[1]),
                    distances_in_refines=int(out[2]), distances_in_reverse_refines=int(out[3]))

    def get_neighbors(self, slot: int, level: int) -> np.ndarray:
        out = np.zeros(2 * self.connectivity, dtype=np.uint32)
        cnt = lib().ro_index_neighbors(self._h, slot, level, out, out.size)
        if cnt < 0:
            raise ValueError("out of range")
        return out[:cnt].copy()

    def save(self, path: str) -> None:
        if lib().ro_index_save(self._h, os.fsencode(path), self.ndim) != 0:
            raise RuntimeError("save failed")


# --------------------------------------------------------------------------- the compiled, unmodified reference

def ref_available() -> bool:
    import glob
    return bool(glob.glob(os.path.join(_HERE, "_ref", "usearch", "compiled*.so")))


def ref_index_module():
    """Import ``usearch.index`` of the UNMODIFIED reference from ``oracle/_ref`` (built by ``build_ref.sh``).

    The reference package is imported under its own name ``usearch``; callers that also use the drop-in shim must
    not have the shim directory on ``sys.path`` at the same time."""
    if not ref_available():
        raise ImportError("oracle/_ref is not built (run oracle/build_ref.sh where /root/reference exists)")
    root = os.path.join(_HERE, "_ref")
    if root not in sys.path:
        sys.path.insert(0, root)
    import importlib
    mod = importlib.import_module("usearch.index")
    if not os.path.abspath(mod.__file__).startswith(root):
        raise ImportError(f"another 'usearch' is already imported from {mod.__file__}")
    return mod
