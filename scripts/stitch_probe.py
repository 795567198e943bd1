"""CPU probe of the cross-shard STITCH: G row shards built independently by the batched
algorithm (tests/model), every node then searches the OTHER shards' graphs, own links + remote candidates are pruned by
the heuristic into its level-0 row (tests/model: stitch_rows), the upper levels are rebuilt over the gathered
upper-level nodes, and the stitched graph -- one `.usearch` file with global slots -- is searched by the UNMODIFIED
reference.  Test infrastructure.
  python scripts/stitch_probe.py dense 200000 4 [k_remote] [ef_merge]"""
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
    t0 = time.time()
    one = M.BatchedModel(db, lv, connectivity=m, expansion_add=ef_add)
    one.build()
    p1 = os.path.join(tmp, "one.usearch")
    M.write_usearch(p1, db, np.arange(n, dtype=np.uint64), lv, one.adj0(), one.adju(), m, one.max_level, one.entry)
    print("single graph", recall_of_file(p1), f"{time.time() - t0:.0f}s")

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
        cands = [a]
        for h in range(G):
            if h == g:
                continue
            sl, _, c = shards[h].search(db[lo:hi], k_remote, ef_merge)
            sl = sl.astype(np.int64) + bounds[h]
            for j in range(k_remote):
                sl[c <= j, j] = EMPTY
            cands.append(sl)
        cand = np.concatenate(cands, axis=1).astype(np.uint32)
        adj0[lo:hi] = M.stitch_rows(db, np.arange(lo, hi, dtype=np.uint32), cand, 2 * m)
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
        if cnt > m:
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
    print(f"stitched G={G} k_remote={k_remote} ef_merge={ef_merge}", recall_of_file(p2))


if __name__ == "__main__":
    main()
