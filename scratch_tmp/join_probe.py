import sys, math, time
import numpy as np
sys.path.insert(0, "/root/repo")
from oracle import oracle as O
from rad_b200 import synth
U = O.ref_index_module()

def build(vecs, keys):
    ix = U.Index(ndim=1024, metric="tanimoto", dtype="b1", connectivity=8, expansion_add=400, expansion_search=64)
    ix.add(keys, vecs)
    return ix

def tanimoto(a, b):
    inter = np.unpackbits(a & b).sum(); uni = np.unpackbits(a | b).sum()
    return np.float32(1.0 - inter / uni) if uni else np.float32(0.0)

def seq_model(men_v, women_v, lists_slots, lists_d, P):
    """sequential statement of index.hpp:4330-4500 (one thread)"""
    from collections import deque
    n, w = len(men_v), len(women_v)
    free = deque(range(n)); count = np.zeros(n, int); wife = -np.ones(n, int); husb = -np.ones(w, int); hd = np.zeros(w, np.float32)
    while free:
        m = free.popleft()
        if count[m] >= P: continue
        j = count[m]; count[m] += 1
        wo, d = lists_slots[m, j], lists_d[m, j]
        if husb[wo] < 0:
            husb[wo], hd[wo], wife[m] = m, d, wo
        elif hd[wo] > d:
            old = husb[wo]; wife[old] = -1; free.append(old)
            husb[wo], hd[wo], wife[m] = m, d, wo
        else:
            free.append(m)
    return wife

def par_model(n, w, lists_slots, lists_d, P):
    count = np.zeros(n, int); wife = -np.ones(n, int); husb = -np.ones(w, int); hd = np.zeros(w, np.float32)
    rounds = 0
    while True:
        act = np.flatnonzero((wife < 0) & (count < P))
        if not len(act): break
        rounds += 1
        j = count[act]; count[act] += 1
        wo, d = lists_slots[act, j], lists_d[act, j]
        key = (d.view(np.uint32).astype(np.uint64) << 32) | act.astype(np.uint64)
        order = np.lexsort((key, wo))
        wo_s, key_s = wo[order], key[order]
        first = np.append(True, wo_s[1:] != wo_s[:-1])
        bw, bk = wo_s[first], key_s[first]
        bm = (bk & 0xFFFFFFFF).astype(int); bd = (bk >> 32).astype(np.uint32).view(np.float32)
        take = (husb[bw] < 0) | (bd < hd[bw])
        bw, bm, bd = bw[take], bm[take], bd[take]
        old = husb[bw]; wife[old[old >= 0]] = -1
        husb[bw], hd[bw], wife[bm] = bm, bd, bw
    return wife, rounds

n_men, n_women = int(sys.argv[1]), int(sys.argv[2])
kind = sys.argv[3] if len(sys.argv) > 3 else "dense"
db = synth.named_set(kind, n_men + n_women, seed=5)
rng = np.random.default_rng(1); perm = rng.permutation(len(db)); db = db[perm]
men_v, women_v = db[:n_men], db[n_men:]
a = build(men_v, np.arange(n_men, dtype=np.uint64)); b = build(women_v, np.arange(n_women, dtype=np.uint64) + 1000000)
t = time.time(); ref = a.join(b, exact=True); print("ref join", len(ref), time.time() - t)
P = min(n_men, int(math.log(n_men) + a.threads if hasattr(a, "threads") else 1))
print("threads?", getattr(a, "threads", None))
for P in (4, 12):
    ref = a.join(b, max_proposals=P, exact=True)
    res = b.search(men_v, P, exact=True)
    ls = (np.asarray(res.keys).astype(np.int64) - 1000000); ld = np.asarray(res.distances, dtype=np.float32)
    ws = seq_model(men_v, women_v, ls, ld, P)
    wp, rounds = par_model(n_men, n_women, ls, ld, P)
    refw = -np.ones(n_men, int)
    for k, v in ref.items(): refw[int(k)] = int(v) - 1000000
    print(f"P={P}: ref pairs {len(ref)}, seq pairs {(ws>=0).sum()}, par pairs {(wp>=0).sum()} rounds {rounds}; ref==seq {(refw==ws).mean():.4f} ref==par {(refw==wp).mean():.4f} seq==par {(ws==wp).mean():.4f}")
