"""CPU probe of the cross-shard STITCH: G row shards built independently by the batched
algorithm (tests/model), every node then searches the OTHER shards' graphs, the remote candidates are merged into its
level-0 row, the upper levels are rebuilt over the gathered upper-level nodes, and the stitched graph -- one `.usearch`
file with global slots -- is searched by the UNMODIFIED reference.  Test infrastructure.
  python scripts/stitch_probe.py dense 200000 4"""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from rad_b200 import synth  # noqa: E402
from tests.model import model as M  # noqa: E402

EMPTY = 0xFFFFFFFF


def main():
    kind, n, G = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    k_remote = int(sys.argv[4]) if len(sys.argv) > 4 else 8
    ef_merge = int(sys.argv[5]) if len(sys.argv) > 5 else 64
    keep_own = int(sys.argv[6]) if len(sys.argv) > 6 else 8
    m, ef_add = 8, 400
    db = synth.named_set(kind, n, seed=3)
    q = synth.queries_from_db(db, 1000, seed=9)
    _, exact_d, _ = O.exact_search(db, q, 10)
    kth = exact_d[:, 9:10]
    lv, _ = O.levels(n, m)
    ui = O.ref_index_module()

    def recall_of_file(path):
        ix = ui.Index(path=path)
        out = []
        for ef in (64, 128):
            ix.expansion_search = ef
            out.append(float((ix.search(q, 10).distances <= kth).mean()))
        return out

    tmp = tempfile.mkdtemp()
    # single graph (the yardstick)
    t0 = time.time()
    one = M.BatchedModel(db, lv, connectivity=m, expansion_add=ef_add)
    one.build()
    p1 = os.path.join(tmp, "one.usearch")
    M.write_usearch(p1, db, np.arange(n, dtype=np.uint64), lv, one.adj0(), one.adju(), m, one.max_level, one.entry)
    print("single graph", recall_of_file(p1), f"{time.time() - t0:.0f}s")

    # shards
    bounds = [n * g // G for g in range(G + 1)]
    shards = []
    for g in range(G):
        lo, hi = bounds[g], bounds[g + 1]
        sm = M.BatchedModel(db[lo:hi], lv[lo:hi], connectivity=m, expansion_add=ef_add)
        sm.build()
        shards.append(sm)
    adj0 = np.full((n, 2 * m), EMPTY, dtype=np.uint32)
    for g in range(G):
        lo, hi = bounds[g], bounds[g + 1]
        a = shards[g].adj0().astype(np.int64)
        a[a != EMPTY] += lo
        own = a
        own_d = np.full(own.shape, np.inf, dtype=np.float32)
        # distances of own links
        for r0 in range(0, hi - lo, 4096):
            r1 = min(hi - lo, r0 + 4096)
            blk = own[r0:r1]
            valid = blk != EMPTY
            fp = db[np.where(valid, blk, 0)]
            qq = db[lo + r0:lo + r1, None, :]
            inter = np.unpackbits(fp & qq, axis=2).sum(axis=2).astype(np.float64)
            union = np.unpackbits(fp | qq, axis=2).sum(axis=2).astype(np.float64)
            d = np.where(union > 0, 1.0 - inter / np.maximum(union, 1), 1.0).astype(np.float32)
            own_d[r0:r1] = np.where(valid, d, np.inf)
        cand_s = [own]
        cand_d = [own_d]
        for h in range(G):
            if h == g:
                continue
            s, d, c = shards[h].search(db[lo:hi], k_remote, ef_merge)
            s = s.astype(np.int64) + bounds[h]
            d = d.copy()
            for j in range(k_remote):
                bad = c <= j
                d[bad, j] = np.inf
                s[bad, j] = EMPTY
            cand_s.append(s)
            cand_d.append(d)
        # own links: the closest `keep_own` stay whatever happens; the rest compete with the remote candidates
        order_own = np.argsort(own_d, axis=1, kind="stable")
        own_sorted = np.take_along_axis(own, order_own, axis=1)
        own_d_sorted = np.take_along_axis(own_d, order_own, axis=1)
        keep_s, keep_d = own_sorted[:, :keep_own], own_d_sorted[:, :keep_own]
        rest_s = np.concatenate([own_sorted[:, keep_own:]] + cand_s[1:], axis=1)
        rest_d = np.concatenate([own_d_sorted[:, keep_own:]] + cand_d[1:], axis=1)
        o2 = np.argsort(rest_d, axis=1, kind="stable")[:, :2 * m - keep_own]
        fin_s = np.concatenate([keep_s, np.take_along_axis(rest_s, o2, axis=1)], axis=1)
        fin_d = np.concatenate([keep_d, np.take_along_axis(rest_d, o2, axis=1)], axis=1)
        o3 = np.argsort(fin_d, axis=1, kind="stable")
        fin_s = np.take_along_axis(fin_s, o3, axis=1)
        fin_d = np.take_along_axis(fin_d, o3, axis=1)
        fin_s[~np.isfinite(fin_d)] = EMPTY
        adj0[lo:hi] = fin_s.astype(np.uint32)
    # upper levels: one build over the gathered upper-level nodes, one level down
    up = np.flatnonzero(lv >= 1)
    um = M.BatchedModel(db[up], (lv[up] - 1).astype(np.int16), connectivity=m, expansion_add=ef_add)
    um.build()
    ua0, uau = um.adj0(), um.adju()  # level 0 of the sub-build = level 1 of the graph (rows of 2m: keep the closest m)
    upper_off = np.concatenate([[0], np.cumsum(lv.astype(np.int64))])
    adju = np.full((int(upper_off[-1]), m), EMPTY, dtype=np.uint32)
    sub_off = np.concatenate([[0], np.cumsum((lv[up] - 1).astype(np.int64))])
    for j, s in enumerate(up):
        row = ua0[j]
        cnt = int((row != EMPTY).sum())
        if cnt > m:  # a level-1 row holds m links: the closest m of the sub-build's 2m
            fp = db[up[row[:cnt]]]
            inter = np.unpackbits(fp & db[s], axis=1).sum(axis=1)
            union = np.unpackbits(fp | db[s], axis=1).sum(axis=1)
            dd = 1.0 - inter / np.maximum(union, 1)
            row = row[:cnt][np.argsort(dd, kind="stable")[:m]]
            cnt = m
        adju[upper_off[s], :cnt] = up[row[:cnt]]
        for l in range(2, int(lv[s]) + 1):
            r = uau[sub_off[j] + l - 2]
            c = int((r != EMPTY).sum())
            adju[upper_off[s] + l - 1, :c] = up[r[:c]]
    entry = int(up[um.entry])
    p2 = os.path.join(tmp, "stitched.usearch")
    M.write_usearch(p2, db, np.arange(n, dtype=np.uint64), lv, adj0, adju, m, int(lv.max()), entry)
    print(f"stitched G={G} k_remote={k_remote} ef_merge={ef_merge} keep_own={keep_own}", recall_of_file(p2))


if __name__ == "__main__":
    main()
