"""Host logic of the ingest path (no GPU): the duplicate chains of rad_b200.ingest.chain_duplicates."""
import numpy as np

from rad_b200.ingest import EMPTY, chain_duplicates


def test_duplicates_hang_behind_their_representative_as_a_chain():
    # rows 0..9; 3 and 7 copy row 1, 8 copies row 2, 9 copies row 1; row 2's row is full, row 1's has room
    rep = np.array([0, 1, 2, 1, 4, 5, 6, 1, 2, 1], dtype=np.uint32)
    m0 = 4
    adj0 = np.full((10, m0), EMPTY, dtype=np.uint32)
    adj0[0, :2] = [1, 2]
    adj0[1, :2] = [0, 4]
    adj0[2] = [0, 4, 5, 6]
    adj0[4, :1] = [1]
    adj0[5, :1] = [2]
    adj0[6, :1] = [2]
    before = adj0.copy()
    chain_duplicates(rep, adj0)
    # the chain of row 1: 1 -> 3 -> 7 -> 9, every copy links back to the representative first
    assert adj0[1].tolist() == [0, 4, 3, int(EMPTY)]
    assert adj0[3].tolist()[:2] == [1, 7] and adj0[7].tolist()[:2] == [1, 9] and adj0[9].tolist()[:2] == [1, int(EMPTY)]
    # row 2 was full: its LAST link makes room for the head of its chain
    assert adj0[2].tolist() == [0, 4, 5, 8] and adj0[8].tolist()[:2] == [2, int(EMPTY)]
    # nobody else changed, and a copy spends at most two links
    for r in (0, 4, 5, 6):
        assert np.array_equal(adj0[r], before[r])
    assert ((adj0[[3, 7, 8, 9]] != EMPTY).sum(axis=1) <= 2).all()


def test_no_duplicates_is_a_no_op():
    rep = np.arange(5, dtype=np.uint32)
    adj0 = np.arange(20, dtype=np.uint32).reshape(5, 4)
    want = adj0.copy()
    chain_duplicates(rep, adj0)
    assert np.array_equal(adj0, want)
