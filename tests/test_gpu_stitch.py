"""One HNSW graph from G row shards: every shard is built independently (here: G handles on ONE
GPU, one thread per shard, the collectives of rad_b200.sharded.LocalComm standing in for NCCL -- the multi-process
form is the same code over TorchComm, bench.py --gpus N), then stitched.

  * rb_stitch_device is bit-exact with its CPU statement (tests/model/batched_model.c: bm_stitch_rows);
  * the stitched graph has global slots, the levels of the reference's single level stream, valid structure, loads in
    the UNMODIFIED reference binding, and its recall@10 is within 0.5 points of the one-piece graph at equal M / ef."""
import threading

import numpy as np
import pytest

from oracle import oracle as O
from rad_b200 import sharded, synth
from rad_b200.index import Index
from tests.model import model as M

pytestmark = pytest.mark.gpu
EMPTY = 0xFFFFFFFF


def _recall(dists, exact_dists, k=10):
    return float((dists[:, :k] <= exact_dists[:, k - 1:k]).mean())


def _build_shards(db, bounds, m, ef_add):
    shards = []
    for g in range(len(bounds) - 1):
        lo, hi = bounds[g], bounds[g + 1]
        ix = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
        ix.seek_level_stream(lo)
        ix.add(np.arange(lo, hi, dtype=np.uint64), db[lo:hi])
        shards.append(ix)
    return shards


def test_stitch_kernel_is_bit_exact_with_its_cpu_statement():
    import torch
    n, m, ef_add, kr, efm = 12000, 8, 64, 8, 64
    db = synth.named_set("dense", n, seed=5)
    bounds = [0, 5000, 12000]
    a, b = _build_shards(db, bounds, m, ef_add)
    sa, sb = sharded.DeviceShard(a), sharded.DeviceShard(b)
    for me, other, lo, hi, obase in ((sa, sb, 0, 5000, 5000), (sb, sa, 5000, 12000, 0)):
        cnt = hi - lo
        q = torch.from_numpy(db[lo:hi]).cuda()
        slots, fp = other.remote_candidates(q, kr, efm)                     # local slots of the OTHER shard
        gslots = torch.where(slots != -1, slots + obase, slots).contiguous()
        rows = torch.full((cnt, 2 * m), -1, dtype=torch.int32, device="cuda")
        me.stitch(0, cnt, lo, gslots, fp.contiguous(), rows)
        got = rows.cpu().numpy().view(np.uint32)
        own, _ = me.index._debug_adjacency()
        own = np.where(own == EMPTY, EMPTY, own.astype(np.int64) + lo).astype(np.uint32)
        cand = np.concatenate([own, gslots.cpu().numpy().view(np.uint32)], axis=1)
        want = M.stitch_rows(db, np.arange(lo, hi, dtype=np.uint32), cand, 2 * m)
        assert np.array_equal(got, want)
        # the remote answers carry the right fingerprints
        s = slots.cpu().numpy()
        f = fp.cpu().numpy()
        for r, c in np.argwhere(s != -1)[::97]:
            assert np.array_equal(f[r, c], db[obase + s[r, c]])


def _stitch_on_one_gpu(parts, bounds, ef_add, batch=16384, gather_below=262144):
    shards = len(parts)
    hub = sharded.LocalComm(shards)
    results, errors = [None] * shards, []

    def run(g):
        try:
            import torch
            torch.cuda.set_device(0)
            results[g] = sharded.build_global_graph(sharded.DeviceShard(parts[g]), bounds[g], expansion_add=ef_add,
                                                    batch=batch, comm=hub.rank_view(g), gather_below=gather_below)
        except Exception as e:  # noqa: BLE001 -- surface the failure in the main thread
            errors.append(e)
            hub._bar.abort()

    threads = [threading.Thread(target=run, args=(g,)) for g in range(shards)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert all(r is None for r in results[1:])
    return results[0]


@pytest.mark.parametrize("kind,shards,gather_below", [("dense", 3, 262144), ("analogue", 2, 262144),
                                                      ("dense", 2, 2000), ("dense", 3, 0)])
def test_stitched_graph_is_one_valid_graph_with_single_graph_recall(kind, shards, gather_below, tmp_path):
    """gather_below = the number of upper-level nodes from which the remaining levels are built on rank 0: the default
    (everything above level 0 on rank 0 at this size), 2000 (level 1 stitched across the shards like level 0, levels
    2.. on rank 0) and 0 (EVERY level stitched)."""
    n, m, ef_add = 90_000, 8, 400
    db = synth.named_set(kind, n, seed=17)
    bounds = [n * g // shards for g in range(shards + 1)]
    parts = _build_shards(db, bounds, m, ef_add)
    res = _stitch_on_one_gpu(parts, bounds, ef_add, gather_below=gather_below)
    lv, _ = O.levels(n, m)
    assert np.array_equal(res["levels"], lv)  # every shard drew ITS rows of the reference's single level stream
    adj0, adju = res["adj0"], res["adj_upper"]
    # structure: global slots in range, no self loops, no duplicates, upper links only to nodes of that level
    valid = adj0 != EMPTY
    assert (adj0[valid] < n).all() and valid.sum(axis=1).min() >= 1
    assert not (adj0 == np.arange(n, dtype=np.uint32)[:, None]).any()
    srt = np.sort(adj0, axis=1)
    assert not ((srt[:, 1:] == srt[:, :-1]) & (srt[:, 1:] != EMPTY)).any()
    upper_off = np.concatenate([[0], np.cumsum(lv.astype(np.int64))])
    owner = np.repeat(np.arange(n), lv.astype(np.int64))
    row_level = np.arange(len(owner)) - upper_off[owner] + 1
    uv = adju != EMPTY
    assert (adju[uv] < n).all() and not (adju == owner[:, None].astype(np.uint32)).any()
    assert (lv[adju[uv]] >= np.broadcast_to(row_level[:, None], adju.shape)[uv]).all()
    assert int(lv[res["entry"]]) == res["max_level"] == int(lv.max())
    # cross-shard links exist: that is the point
    shard_of = np.searchsorted(np.asarray(bounds[1:]), np.arange(n), side="right")
    cross = (shard_of[np.where(valid, adj0, 0)] != shard_of[:, None]) & valid
    assert cross.mean() > 0.05
    # one index, one file
    whole = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
    whole.import_graph(np.arange(n, dtype=np.uint64), res["levels"], adj0, adju, db, res["entry"])
    assert len(whole) == n and whole.max_level == res["max_level"] and whole.entry_slot == res["entry"]
    path = str(tmp_path / "stitched.usearch")
    whole.save(path)
    q = synth.queries_from_db(db, 1000, seed=9)
    exact = whole.search(q, 10, exact=True)
    one = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
    one.add(np.arange(n, dtype=np.uint64), db)
    report = {}
    for ef in (64, 128):
        whole.expansion_search = ef
        one.expansion_search = ef
        report[ef] = (_recall(whole.search(q, 10).distances, exact.distances),
                      _recall(one.search(q, 10).distances, exact.distances))
    print(f"\n[stitch] {kind} {n} rows in {shards} shards, gather_below={gather_below}: " +
          ", ".join(f"ef={ef}: stitched {a:.4f} vs one-piece {b:.4f}" for ef, (a, b) in report.items()))
    for ef, (a, b) in report.items():
        assert a >= b - 0.005, (kind, ef, a, b)
    if O.ref_available():
        ref = O.ref_index_module().Index(path=path)
        assert len(ref) == n and ref.max_level == whole.max_level
        for slot in list(range(0, n, 997)) + [res["entry"]]:
            for level in range(int(lv[slot]) + 1):
                assert np.array_equal(np.asarray(ref.get_neighbors(slot, level), dtype=np.uint64),
                                      np.asarray(whole.get_neighbors(slot, level), dtype=np.uint64))
        ref.expansion_search = 64
        theirs = _recall(ref.search(q, 10).distances, exact.distances)
        assert theirs >= report[64][0] - 0.01  # the reference's own search walks the stitched graph just as well


def test_extracted_upper_levels_are_the_links_of_the_build():
    """rb_extract_level: the sub-index of the nodes of level >= 1, one level down, equals a numpy remap of the source's
    own upper-level rows; it answers searches and extracts again (the recursion of the upper-level stitch)."""
    n, m = 40000, 8
    db = synth.named_set("dense", n, seed=6)
    src = Index(ndim=1024, connectivity=m, expansion_add=64, expansion_search=64)
    src.add(np.arange(n, dtype=np.uint64), db)
    lv = np.asarray(src.levels, dtype=np.int64)
    _, au = src._debug_adjacency()
    upper_off = np.concatenate([[0], np.cumsum(lv)])
    for depth in range(2):
        sub = Index(ndim=1024, connectivity=m, expansion_add=64, expansion_search=64, _connectivity_base=m)
        sub.extract_upper_levels(src)
        up = np.flatnonzero(lv >= 1)
        assert len(sub) == len(up) and sub.max_level == src.max_level - 1
        assert np.array_equal(np.asarray(sub.levels, dtype=np.int64), lv[up] - 1)
        table = np.full(len(lv) + 1, EMPTY, dtype=np.uint32)
        table[up] = np.arange(len(up), dtype=np.uint32)

        def remap(rows):
            return table[np.where(rows == EMPTY, len(lv), rows).astype(np.int64)]

        sub.sync_host()
        s0, su = sub._debug_adjacency()
        assert s0.shape == (len(up), m)
        assert np.array_equal(s0, remap(au[upper_off[up]]))
        sub_off = np.concatenate([[0], np.cumsum(lv[up] - 1)])
        deeper = np.flatnonzero(lv[up] >= 2)
        assert np.array_equal(su[sub_off[deeper]], remap(au[upper_off[up[deeper]] + 1]))
        assert int(sub.entry_slot) == int(table[src.entry_slot])
        vec = sub.get_vectors(np.arange(0, len(up), 97, dtype=np.uint64)) if hasattr(sub, "get_vectors") else None
        q = np.ascontiguousarray(db[up[::53]][:64]) if depth == 0 else None
        if q is not None:
            got = sub.search(q, 5, exact=True)  # keys of the sub-index are its own slots
            slots, dists, _ = O.exact_search(np.ascontiguousarray(db[up]), q, 5)
            assert np.array_equal(got.keys, slots) and np.array_equal(got.distances.view(np.uint32), dists.view(np.uint32))
            approx = sub.search(q, 5)
            assert _recall(approx.distances, dists, 5) > 0.9
        del vec
        # once more, from the extracted index
        db, src, lv, au = db[up], sub, lv[up] - 1, su
        upper_off = np.concatenate([[0], np.cumsum(lv)])
