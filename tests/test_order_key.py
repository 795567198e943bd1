"""The f32 order key used inside GPU graph traversal (device_common.cuh: order_key_bits = 1.0f - RN(i/u)) must order
every pair of (intersection, union) counts exactly like the reference's distance and be equal for equal rationals.
Exhaustive over u <= 1024 (a union of two 1024-bit fingerprints cannot exceed 1024)."""
import numpy as np


def test_order_key_is_exact_order():
    ii, uu = np.meshgrid(np.arange(0, 1025), np.arange(1, 1025), indexing="ij")
    m = ii <= uu
    i = ii[m].astype(np.int64)
    u = uu[m].astype(np.int64)
    exact = (i << 40) // u
    sur = (np.float32(1.0) - i.astype(np.float32) / u.astype(np.float32)).view(np.uint32).astype(np.int64)
    ref = (1.0 - i.astype(np.float64) / u.astype(np.float64)).astype(np.float32).view(np.uint32).astype(np.int64)
    order = np.argsort(exact, kind="stable")
    e, s, r = exact[order], sur[order], ref[order]
    same = np.diff(e) == 0
    assert not (np.diff(s)[same] != 0).any() and not (np.diff(r)[same] != 0).any()
    assert not (np.diff(s)[~same] >= 0).any() and not (np.diff(r)[~same] >= 0).any()
