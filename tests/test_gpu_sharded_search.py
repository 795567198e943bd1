"""The device-resident merge behind ShardedIndex.search (rb_search_device -> keys on the GPU -> rb_merge_topk_device ->
user keys), exercised with two shards held by one process on one GPU; the NCCL all-gather between them is replaced by
torch.stack (bench.py --gpus N runs the same path over NCCL and checks it against the exact answers)."""
import numpy as np
import pytest

from rad_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("exact", [True, False])
def test_two_shards_merge_on_the_device_like_one_index(exact):
    import torch
    from rad_b200.index import Index
    from rad_b200.sharded import ShardedIndex, merge_topk_host, pack_topk_keys, unpack_topk_keys
    n, cut, k = 30000, 17000, 10
    db = synth.dense_set(n, seed=8)
    db[cut + 5] = db[3]  # a tie across the shard boundary: the higher GLOBAL slot comes first
    keys = np.arange(n, dtype=np.uint64) * 7 + 11
    q = synth.queries_from_db(db, 300, seed=1)
    q[0] = 0  # the all-zero query
    shards, bases = [], [0, cut]
    for lo, hi in ((0, cut), (cut, n)):
        ix = Index(ndim=1024, connectivity=8, expansion_add=128, expansion_search=64)
        ix.add(keys[lo:hi], db[lo:hi])
        shards.append(ShardedIndex(ix, lo))
    d_q = torch.from_numpy(q).cuda()
    parts = [s.local_answers_device(d_q, k, exact) for s in shards]
    mk = torch.stack([p[0] for p in parts])
    uk = torch.stack([p[1] for p in parts])
    got_keys, got_d, got_c = ShardedIndex.merge_answers_device(mk, uk, k)
    got_keys = got_keys.cpu().numpy().view(np.uint64)
    got_d = got_d.cpu().numpy()
    # the numpy statement of the same merge over the shards' host answers
    host_parts, host_user = [], []
    for s, base in zip(shards, bases):
        r = s.local.search(q, k, exact=exact)
        slots = np.asarray(s.local.get_node_ids_from_keys(r.keys.reshape(-1)), dtype=np.uint64).reshape(len(q), k)
        host_parts.append(pack_topk_keys(r.distances, slots + np.uint64(base), r.counts))
        host_user.append(r.keys)
    want = merge_topk_host(np.stack(host_parts), k)
    want_slots, want_d, want_c = unpack_topk_keys(want)
    assert np.array_equal(got_d.view(np.uint32), want_d.view(np.uint32))
    assert np.array_equal(got_c.cpu().numpy(), want_c)
    assert np.array_equal(got_keys, np.where(want != np.uint64(0xFFFFFFFFFFFFFFFF), want_slots * 7 + 11, 0))
    if exact:  # ... and, for the exact search, the answers of ONE index over all rows
        whole = Index(ndim=1024, connectivity=8, expansion_add=128)
        whole.add(keys, db)
        ref = whole.search(q, k, exact=True)
        assert np.array_equal(got_keys, ref.keys) and np.array_equal(got_d.view(np.uint32), ref.distances.view(np.uint32))
