"""Synthetic clustered 1024-bit fingerprints: the workload generator shared by tests and bench.

C = max(1000, N/100) cluster centres, each bit of a centre set with probability p_c ~ U[0.03, 0.08]; a row is its
centre with 30 % of the set bits dropped and fresh bits added with probability 0.015 (about 5 % density, 40-80 bits
set, many exact distance ties and near-duplicates).  Row 0 is all-zero and row 1 duplicates row 2 (the u == 0 and
d == 0 edge cases every test set must contain).

Everything below the centre table is COUNTER-BASED: bit b of row r depends only on (seed, r, b) through a
splitmix64 finaliser, so any slice can be regenerated independently on the CPU (this file, numpy) and on the GPU
(``rb_synth_fingerprints`` in csrc/synth.cu, same arithmetic) -- the CPU oracle can therefore check any shard of a
100M-1B row device-generated set without materialising it on the host.
Bit order inside a byte is MSB-first (``numpy.packbits``), which is what the .usearch matrix blob stores.
"""
from __future__ import annotations

import numpy as np

NDIM = 1024
ROW_BYTES = NDIM // 8
KEEP_T = 19661   # round(0.30 * 65536): a centre bit survives when keep16 >= KEEP_T
NOISE_T = 983    # round(0.015 * 65536): a fresh bit appears when noise16 < NOISE_T
_GOLD = np.uint64(0x9E3779B97F4A7C15)
_ROWS_PER_CHUNK = 2048


def _mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array (wraps modulo 2^64)."""
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    return x


def n_centres(n_total: int) -> int:
    return max(1000, n_total // 100)


def centre_table(seed: int, n_total: int, ndim: int = NDIM, *, count: int | None = None) -> np.ndarray:
    """uint8 [C, ndim/8] packed cluster centres (host-side numpy; uploaded as-is for the device generator).
    ``count`` overrides C = max(1000, n_total / 100): the first C rows of the table do not depend on C."""
    c = n_centres(n_total) if count is None else int(count)
    rng = np.random.default_rng([seed, 0xC3])
    p = rng.uniform(0.03, 0.08, size=(c, 1)).astype(np.float32)
    out = np.empty((c, ndim // 8), dtype=np.uint8)
    step = 8192
    for lo in range(0, c, step):
        hi = min(c, lo + step)
        out[lo:hi] = np.packbits(rng.random((hi - lo, ndim), dtype=np.float32) < p[lo:hi], axis=1)
    return out


def row_assignment(rows: np.ndarray, seed: int, c: int) -> np.ndarray:
    """centre index of each row: splitmix64(seed * GOLD + row + 0xA55A) mod C."""
    with np.errstate(over="ignore"):
        h = _mix64(np.uint64(seed) * _GOLD + rows.astype(np.uint64) + np.uint64(0xA55A))
    return (h % np.uint64(c)).astype(np.int64)


def source_rows(rows: np.ndarray, dup_fold: int = 0, dup_start: int = 0, dup_rows: int = 0) -> np.ndarray:
    """The row whose bits row r carries: r itself, except inside the duplicate region [dup_start, dup_start + dup_rows)
    where every run of ``dup_fold`` consecutive rows repeats the first row of the run (exact duplicates, d = 0)."""
    rows = rows.astype(np.uint64)
    if dup_fold > 1 and dup_rows:
        inside = (rows >= np.uint64(dup_start)) & (rows < np.uint64(dup_start + dup_rows))
        rel = rows - np.uint64(dup_start)
        rows = np.where(inside, np.uint64(dup_start) + rel - rel % np.uint64(dup_fold), rows)
    return rows


def clustered_fingerprints(n: int, seed: int = 0, *, start: int = 0, n_total: int | None = None,
                           ndim: int = NDIM, edge_cases: bool = True, centres: np.ndarray | None = None,
                           rows_per_centre: int = 0, dup_fold: int = 0, dup_start: int = 0,
                           dup_rows: int = 0) -> np.ndarray:
    """Rows [start, start + n) of the synthetic set of ``n_total`` rows (default ``start + n``) -> uint8 [n, ndim/8].

    ``rows_per_centre`` > 0 selects the ANALOGUE-SERIES order: row r belongs to centre (r // rows_per_centre) mod C, so
    the rows of one cluster are contiguous in insertion order (a build batch then holds whole clusters whose members
    have never seen each other); 0 = the hashed assignment.  ``dup_fold`` / ``dup_start`` / ``dup_rows`` plant runs
    of exact duplicates (``source_rows``)."""
    if n_total is None:
        n_total = start + n
    if centres is None:
        centres = centre_table(seed, n_total, ndim)
    c = centres.shape[0]
    out = np.empty((n, ndim // 8), dtype=np.uint8)
    bit_ids = np.arange(ndim, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        base = np.uint64(seed) * _GOLD
        for lo in range(0, n, _ROWS_PER_CHUNK):
            hi = min(n, lo + _ROWS_PER_CHUNK)
            rows = source_rows(np.arange(start + lo, start + hi, dtype=np.uint64), dup_fold, dup_start, dup_rows)
            h = _mix64(base + ((rows[:, None] << np.uint64(10)) | bit_ids))
            keep = (h & np.uint64(0xFFFF)) >= np.uint64(KEEP_T)
            noise = ((h >> np.uint64(16)) & np.uint64(0xFFFF)) < np.uint64(NOISE_T)
            if rows_per_centre > 0:
                assign = ((rows // np.uint64(rows_per_centre)) % np.uint64(c)).astype(np.int64)
            else:
                assign = row_assignment(rows, seed, c)
            cbits = np.unpackbits(centres[assign], axis=1).astype(bool)
            out[lo:hi] = np.packbits((cbits & keep) | noise, axis=1)
    if edge_cases and n_total >= 3:
        if start <= 0 < start + n:
            out[0 - start] = 0
        if start <= 1 < start + n:
            out[1 - start] = clustered_fingerprints(1, seed, start=2, n_total=n_total, ndim=ndim,
                                                    edge_cases=False, centres=centres,
                                                    rows_per_centre=rows_per_centre, dup_fold=dup_fold,
                                                    dup_start=dup_start, dup_rows=dup_rows)[0]
    return out


def dense_set(n: int, seed: int = 0, **kw) -> np.ndarray:
    """The ROUTABLE set (SURVEY.md §6 probe): a fixed table of 1000 centres whatever n, hashed assignment -- n/1000 rows
    per cluster; the reference graph reaches recall@10 0.99 at ef=64 on it at 100 k rows."""
    return clustered_fingerprints(n, seed, centres=centre_table(seed, n, count=1000), **kw)


ANALOGUE_ROWS_PER_CENTRE = 200
ANALOGUE_DUP_FOLD = 400


def analogue_params(n_total: int) -> dict:
    """Generator arguments of the ADVERSARIAL analogue-series set: n/200 centres, the 200 rows of a centre contiguous
    in insertion order, and 1/50 of the rows (from n/2 on) in runs of 400 exact duplicates."""
    dup_rows = (n_total // 50) // ANALOGUE_DUP_FOLD * ANALOGUE_DUP_FOLD
    return dict(rows_per_centre=ANALOGUE_ROWS_PER_CENTRE, dup_fold=ANALOGUE_DUP_FOLD,
                dup_start=(n_total // 2) // ANALOGUE_ROWS_PER_CENTRE * ANALOGUE_ROWS_PER_CENTRE, dup_rows=dup_rows)


def analogue_set(n: int, seed: int = 0, *, start: int = 0, n_total: int | None = None, **kw) -> np.ndarray:
    n_total = start + n if n_total is None else n_total
    centres = centre_table(seed, n_total, count=max(1, n_total // ANALOGUE_ROWS_PER_CENTRE))
    return clustered_fingerprints(n, seed, start=start, n_total=n_total, centres=centres,
                                  **analogue_params(n_total), **kw)


def queries_from_db(db: np.ndarray, q: int, seed: int = 1) -> np.ndarray:
    """Q rows sampled without replacement from the database (configs 1-2)."""
    rng = np.random.default_rng([seed, 0x51])
    idx = rng.choice(db.shape[0], size=min(q, db.shape[0]), replace=False)
    return np.ascontiguousarray(db[np.sort(idx)])


def heldout_queries(q: int, seed: int = 0, *, n_total: int, ndim: int = NDIM) -> np.ndarray:
    """Q held-out rows drawn from the same centres as the database of ``n_total`` rows (config 3): they are rows
    [2^40, 2^40 + q) of the same counter-based stream, which no database shard ever reaches."""
    return clustered_fingerprints(q, seed, start=1 << 40, n_total=n_total, ndim=ndim, edge_cases=False)


def set_spec(kind: str, n_total: int, seed: int = 0) -> tuple[np.ndarray, dict]:
    """(centre table, generator arguments) of a named workload: 'sparse' = SURVEY.md §8d (n/100 centres, hashed),
    'dense' = the routable set (1000 centres), 'analogue' = the adversarial analogue-series set."""
    if kind == "sparse":
        return centre_table(seed, n_total), {}
    if kind == "dense":
        return centre_table(seed, n_total, count=1000), {}
    if kind == "analogue":
        return (centre_table(seed, n_total, count=max(1, n_total // ANALOGUE_ROWS_PER_CENTRE)),
                analogue_params(n_total))
    raise ValueError(f"unknown workload {kind!r}")


def named_set(kind: str, n: int, seed: int = 0, *, start: int = 0, n_total: int | None = None) -> np.ndarray:
    """numpy rows [start, start + n) of a named workload."""
    n_total = start + n if n_total is None else n_total
    centres, modes = set_spec(kind, n_total, seed)
    return clustered_fingerprints(n, seed, start=start, n_total=n_total, centres=centres, **modes)


def named_set_device(kind: str, n: int, seed: int = 0, *, start: int = 0, n_total: int | None = None, device=None):
    """The same rows generated ON THE DEVICE (rb_synth_fingerprints_device_ex) -> torch.uint8 [n, 128] cuda tensor."""
    import ctypes as C

    import torch

    from . import _native
    n_total = start + n if n_total is None else n_total
    centres, modes = set_spec(kind, n_total, seed)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    d_c = torch.from_numpy(centres).to(dev)
    out = torch.empty((n, ROW_BYTES), dtype=torch.uint8, device=dev)
    err = C.c_char_p()
    with torch.cuda.device(dev):
        _native.lib().rb_synth_fingerprints_device_ex(
            out.data_ptr(), start, n, seed, d_c.data_ptr(), len(centres), 1, modes.get("rows_per_centre", 0),
            modes.get("dup_fold", 0), modes.get("dup_start", 0), modes.get("dup_rows", 0),
            torch.cuda.current_stream(dev).cuda_stream, C.byref(err))
        _native.check(err)
        torch.cuda.synchronize(dev)
    return out
