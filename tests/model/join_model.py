"""CPU statement of the reference's join loop on ONE thread (test infrastructure: only tests/ use it).

usearch/include/usearch/index.hpp:4330-4515: a ring of free men; the man at the front proposes to his next candidate
(``women.search(man, k)`` -> the k-th answer); a free woman accepts, an engaged one leaves her husband only when the
proposer is STRICTLY nearer (:4467); the loser goes to the back of the ring; a man who has made ``proposals`` proposals
leaves the ring when he is popped (:4440-4442).  ``rb_stable_match_device`` must produce this matching bit for bit."""
from collections import deque

import numpy as np


def sequential_join(cand_slots, cand_dist, cand_counts, women: int, proposals: int):
    """cand_slots int [men, proposals] (nearest first), cand_dist float32, cand_counts int [men] ->
    (man_to_woman int64 [men], woman_to_man int64 [women], engagements); -1 = unmatched."""
    men = len(cand_slots)
    free = deque(range(men))
    made = np.zeros(men, dtype=np.int64)
    wife = -np.ones(men, dtype=np.int64)
    husband = -np.ones(women, dtype=np.int64)
    husband_dist = np.zeros(women, dtype=np.float32)
    engagements = 0
    while free:
        m = free.popleft()
        if made[m] >= proposals:
            continue
        j = made[m]
        made[m] += 1
        if j >= cand_counts[m]:
            free.append(m)  # the last woman again (candidates.back()): her husband is at least as near by now
            continue
        w, d = int(cand_slots[m, j]), np.float32(cand_dist[m, j])
        if husband[w] < 0:
            husband[w], husband_dist[w], wife[m] = m, d, w
            engagements += 1
        elif husband_dist[w] > d:
            old = husband[w]
            wife[old] = -1
            free.append(old)
            husband[w], husband_dist[w], wife[m] = m, d, w
            engagements += 1
        else:
            free.append(m)
    return wife, husband, engagements
