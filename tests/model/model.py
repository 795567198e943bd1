"""ctypes front-end to batched_model.c (CPU model of the batched GPU build) -- TEST INFRASTRUCTURE."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
EMPTY = 0xFFFFFFFF

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(dtype=np.int16, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libbatched_model.so")
        src = os.path.join(_HERE, "batched_model.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O3", "-march=x86-64-v2", "-mpopcnt", "-fopenmp", "-fPIC", "-shared",
                                   "-fvisibility=hidden", "-o", so, src, "-lm"])
        L = C.CDLL(so)
        L.bm_new.restype = C.c_void_p
        L.bm_new.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _u8p, _i16p,
                             C.c_uint64]
        L.bm_free.argtypes = [C.c_void_p]
        L.bm_build.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
        L.bm_search.argtypes = [C.c_void_p, _u8p, C.c_uint64, C.c_uint32, C.c_uint32, _u32p, _f32p, _i64p]
        for name, rt in [("bm_size", C.c_uint64), ("bm_max_level", C.c_int32), ("bm_entry", C.c_uint32),
                         ("bm_batches", C.c_uint64), ("bm_upper_rows", C.c_uint64),
                         ("bm_adj0", C.POINTER(C.c_uint32)), ("bm_adju", C.POINTER(C.c_uint32))]:
            getattr(L, name).restype = rt
            getattr(L, name).argtypes = [C.c_void_p]
        L.bm_counters.argtypes = [C.c_void_p, _u64p]
        L.bm_stitch_rows.argtypes = [_u8p, C.c_uint64, _u32p, _u32p, C.c_uint32, C.c_uint32, _u32p]
        _LIB = L
    return _LIB


class BatchedModel:
    def __init__(self, vecs: np.ndarray, levels: np.ndarray, *, connectivity=8, expansion_add=400,
                 bloom_bits_log2=15, growth_den=8, batch_max=65536, batch_k=16):
        self.vecs = np.ascontiguousarray(vecs, dtype=np.uint8)
        assert self.vecs.shape[1] == 128
        self.levels = np.ascontiguousarray(levels, dtype=np.int16)
        self.m = connectivity
        self.m0 = 2 * connectivity
        self._h = C.c_void_p(lib().bm_new(connectivity, expansion_add, bloom_bits_log2, growth_den, batch_max,
                                          batch_k, self.vecs, self.levels, len(self.levels)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().bm_free(self._h)
            self._h = None

    def build(self, upto=None, max_batches=0):
        lib().bm_build(self._h, len(self.levels) if upto is None else upto, max_batches)

    def __len__(self):
        return int(lib().bm_size(self._h))

    @property
    def max_level(self):
        return int(lib().bm_max_level(self._h))

    @property
    def entry(self):
        return int(lib().bm_entry(self._h))

    @property
    def batches(self):
        return int(lib().bm_batches(self._h))

    def adj0(self) -> np.ndarray:
        n = len(self.levels)
        return np.ctypeslib.as_array(lib().bm_adj0(self._h), shape=(n, self.m0)).copy()

    def adju(self) -> np.ndarray:
        rows = int(lib().bm_upper_rows(self._h))
        if rows == 0:
            return np.zeros((0, self.m), dtype=np.uint32)
        return np.ctypeslib.as_array(lib().bm_adju(self._h), shape=(rows, self.m)).copy()

    def counters(self):
        out = np.zeros(2, dtype=np.uint64)
        lib().bm_counters(self._h, out)
        return dict(computed_distances=int(out[0]), expansions=int(out[1]))

    def search(self, queries, k=10, ef=64):
        queries = np.ascontiguousarray(queries, dtype=np.uint8)
        nq = len(queries)
        slots = np.zeros((nq, k), dtype=np.uint32)
        dists = np.zeros((nq, k), dtype=np.float32)
        counts = np.zeros(nq, dtype=np.int64)
        lib().bm_search(self._h, queries, nq, k, ef, slots, dists, counts)
        return slots, dists, counts


def write_usearch(path, vecs, keys, levels, adj0, adju, m, max_level, entry, ndim=1024):
    """numpy writer of the .usearch v2.17 layout (index_dense.hpp:1025-1093, index.hpp:3236-3277) from the
    struct-of-arrays graph -- an independent check of the C++ writer in rad_b200/csrc."""
    n = len(levels)
    m0 = 2 * m
    levels = np.asarray(levels, dtype=np.int16)
    node_bytes = 10 + 4 * (1 + m0) + levels.astype(np.int64) * 4 * (1 + m)
    offs = np.concatenate([[0], np.cumsum(node_bytes)])
    tape = np.zeros(int(offs[-1]), dtype=np.uint8)
    keys = np.asarray(keys, dtype=np.uint64)
    upper_off = np.concatenate([[0], np.cumsum(levels.astype(np.int64))])[:-1]
    for s in range(n):
        o = int(offs[s])
        tape[o:o + 8] = np.frombuffer(keys[s].tobytes(), dtype=np.uint8)
        tape[o + 8:o + 10] = np.frombuffer(levels[s].tobytes(), dtype=np.uint8)
        row = adj0[s]
        cnt = int((row != EMPTY).sum())
        blk = np.zeros(1 + m0, dtype=np.uint32)
        blk[0] = cnt
        blk[1:1 + cnt] = row[:cnt]
        tape[o + 10:o + 10 + 4 * (1 + m0)] = blk.view(np.uint8)
        o2 = o + 10 + 4 * (1 + m0)
        for l in range(1, int(levels[s]) + 1):
            row = adju[int(upper_off[s]) + l - 1]
            cnt = int((row != EMPTY).sum())
            blk = np.zeros(1 + m, dtype=np.uint32)
            blk[0] = cnt
            blk[1:1 + cnt] = row[:cnt]
            tape[o2:o2 + 4 * (1 + m)] = blk.view(np.uint8)
            o2 += 4 * (1 + m)
    head = bytearray(64)
    head[0:7] = b"usearch"
    head[7:13] = np.array([2, 17, 2], dtype=np.uint16).tobytes()
    head[13] = ord("t")
    head[14] = 1
    head[15] = 14
    head[16] = 15
    head[17:25] = np.uint64(n).tobytes()
    head[25:33] = np.uint64(0).tobytes()
    head[33:41] = np.uint64(ndim).tobytes()
    with open(path, "wb") as f:
        f.write(np.array([n, vecs.shape[1]], dtype=np.uint32).tobytes())
        f.write(np.ascontiguousarray(vecs).tobytes())
        f.write(bytes(head))
        f.write(np.array([n, m, m0, max_level, entry], dtype=np.uint64).tobytes())
        f.write(levels.tobytes())
        f.write(tape.tobytes())


def stitch_rows(vecs, qslots, cands, needed):
    """heuristic over (own links + remote candidates) per node -> uint32 [n, needed] (batched_model.c: bm_stitch_rows)"""
    vecs = np.ascontiguousarray(vecs, dtype=np.uint8)
    qslots = np.ascontiguousarray(qslots, dtype=np.uint32)
    cands = np.ascontiguousarray(cands, dtype=np.uint32)
    out = np.empty((len(qslots), needed), dtype=np.uint32)
    lib().bm_stitch_rows(vecs, len(qslots), qslots, cands, cands.shape[1], needed, out)
    return out
