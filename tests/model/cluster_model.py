"""CPU statement of ``cluster(query, level)`` (usearch/include/usearch/index.hpp:3046-3079, 3935-3975): the greedy
descent from the entry point down to and including ``level``.  Test infrastructure: reads the graph through the index's
host calls (``get_neighbors``), never the GPU."""
import numpy as np

_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.int64)


def _dist(vectors, slot, q):
    inter = int(_POP8[vectors[slot] & q].sum())
    union = int(_POP8[vectors[slot] | q].sum())
    return np.float32(1.0 - inter / union) if union else np.float32(0.0)


def descend(index, vectors, queries, level):
    """``vectors``: the indexed fingerprints by slot -> (slots uint32 [Q], distances float32 [Q], visited, computed)"""
    rows = {}

    def row(slot, lv):
        if (slot, lv) not in rows:
            rows[(slot, lv)] = [int(x) for x in list(index.get_neighbors(slot, lv))[0::2]]
        return rows[(slot, lv)]

    slots = np.empty(len(queries), dtype=np.uint32)
    dists = np.empty(len(queries), dtype=np.float32)
    visited = computed = 0
    for qi, q in enumerate(queries):
        cur = index.entry_slot
        cur_d = _dist(vectors, cur, q)
        computed += 1
        for lv in range(index.max_level, (level if level > 0 else 1) - 1, -1):
            changed = True
            while changed:
                changed = False
                for cand in row(cur, lv):
                    d = _dist(vectors, cand, q)
                    computed += 1
                    if d < cur_d:
                        cur, cur_d, changed = cand, d, True
                visited += 1
        slots[qi], dists[qi] = cur, cur_d
        computed += 1
    return slots, dists, visited, computed
