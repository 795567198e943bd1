// ingest_kernels.cuh -- fingerprint ingest pre-passes on the device: exact duplicates.
//
// The reference has no such pass: identical fingerprints (stereo-isomers under an achiral Morgan fingerprint, salts of
// one parent) are inserted like any other row and form d = 0 cliques that fill the M links of their members and
// inflate every ef-bounded search that touches them (usearch/include/usearch/index.hpp:2734-2834 treats every row
// alike).  rb_dedup_device names, for every row, the LOWEST row with the same 128 bytes:
//   row_hash_kernel        one 64-bit hash per row (8 lanes per row, one 128-byte line)
//   cub radix sort         (hash, row) pairs, stable -> rows ascending inside a run of equal hashes (library code)
//   run_start_kernel + cub inclusive max-scan   first sorted position of every run
//   representative_kernel  row -> first row of its run with IDENTICAL BYTES (a 64-bit collision between different
//                          fingerprints only costs extra compares; the answer never depends on the hash)
#pragma once
#include "device_common.cuh"

namespace rb {

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33, x *= 0xff51afd7ed558ccdull, x ^= x >> 33, x *= 0xc4ceb9fe1a85ec53ull, x ^= x >> 33;
    return x;
}

__global__ void __launch_bounds__(256) row_hash_kernel(const uint4* __restrict__ fp, uint64_t rows,
                                                       uint64_t* __restrict__ hashes, uint32_t* __restrict__ ids) {
    const uint32_t sub = threadIdx.x & 7;
    for (uint64_t row = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;; row += ((uint64_t)gridDim.x * blockDim.x) >> 3) {
        if (!__any_sync(RB_FULL_MASK, row < rows))  // a warp leaves as a whole: every lane takes part in the shuffles
            break;
        uint64_t h = 0;
        if (row < rows) {
            const uint4 v = __ldg(fp + row * 8 + sub);
            h = mix64(((uint64_t)v.y << 32 | v.x) + 0x9E3779B97F4A7C15ull * (2 * sub + 1));
            h ^= mix64(((uint64_t)v.w << 32 | v.z) + 0x9E3779B97F4A7C15ull * (2 * sub + 2));
        }
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 1);
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 2);
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 4);
        if (row < rows && sub == 0) {
            hashes[row] = mix64(h);
            ids[row] = (uint32_t)row;
        }
    }
}

// start[p] = p where a run of equal hashes begins, 0 elsewhere (an inclusive max-scan then spreads it over the run)
__global__ void __launch_bounds__(256) run_start_kernel(const uint64_t* __restrict__ sorted_hashes, uint64_t rows,
                                                        uint32_t* __restrict__ start) {
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < rows; p += (uint64_t)gridDim.x * blockDim.x)
        start[p] = (p && sorted_hashes[p] != sorted_hashes[p - 1]) ? (uint32_t)p : 0u;
}

// 8 lanes per sorted position: the first row of the run whose bytes equal this row's
__global__ void __launch_bounds__(256) representative_kernel(const uint4* __restrict__ fp, uint64_t rows,
                                                             const uint32_t* __restrict__ sorted_ids,
                                                             const uint32_t* __restrict__ run_start,
                                                             uint32_t* __restrict__ representative) {
    const uint32_t sub = threadIdx.x & 7;
    for (uint64_t p = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;; p += ((uint64_t)gridDim.x * blockDim.x) >> 3) {
        if (!__any_sync(RB_FULL_MASK, p < rows))
            break;
        const bool on = p < rows;
        const uint32_t me = on ? sorted_ids[p] : 0u;
        const uint4 mine = on ? __ldg(fp + (size_t)me * 8 + sub) : make_uint4(0, 0, 0, 0);
        uint64_t c = on ? run_start[p] : 0;
        uint32_t rep = me;
        bool done = !on;
        // the groups of a warp walk runs of different lengths: loop while any group still looks
        while (__any_sync(RB_FULL_MASK, !done)) {
            const bool look = !done && c < p;
            const uint32_t other = look ? sorted_ids[c] : me;
            const uint4 v = look ? __ldg(fp + (size_t)other * 8 + sub) : mine;
            bool same = v.x == mine.x && v.y == mine.y && v.z == mine.z && v.w == mine.w;
            uint32_t agree = __ballot_sync(RB_FULL_MASK, same);
            same = ((agree >> (threadIdx.x & 24)) & 0xFFu) == 0xFFu;  // all 8 lanes of this group
            if (!done) {
                if (!look) {
                    done = true;  // reached itself: the first of its kind
                } else if (same) {
                    rep = other;
                    done = true;
                } else {
                    ++c;
                }
            }
        }
        if (on && sub == 0)
            representative[me] = rep;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Index.join: one-to-one matching of two fingerprint sets (usearch/include/usearch/index.hpp:4330-4515).  The reference
// keeps a ring of free "men"; a popped man proposes to his next-nearest "woman" (the k-th answer of a search), a free
// woman accepts, an engaged one leaves her husband only for a STRICTLY nearer man, and whoever loses goes to the back
// of the ring.  With one thread the ring is processed in GENERATIONS: everybody appended while generation g is being
// processed forms generation g + 1, in the order of the proposals that sent them there.  Women are independent inside
// a generation, so a generation is three data-parallel steps here, and the matching is the one the reference's loop
// produces on one thread (tests/model/join_model.py states that loop):
//   join_propose_kernel   position i of the generation -> sort key (woman << 32 | i), value = distance bits
//   cub radix sort        proposals grouped by woman, in ring order inside a group (library code)
//   join_settle_kernel    one thread per woman walks her proposals in ring order; the loser of proposal i (the
//                         proposer or the husband he replaces) is written to next[i]
//   cub select            next[] without its holes = the next generation, in ring order (library code)
constexpr uint32_t kJoinNone = 0xFFFFFFFFu;

struct JoinParams {
    const uint32_t* cand_slots;  // [men x proposals] women, nearest first (rb_search_device)
    const uint32_t* cand_dist;   // [men x proposals] f32 bits
    const int32_t* cand_counts;  // [men] valid candidates
    uint32_t proposals;
    uint32_t* made;              // [men] proposals made so far
    uint32_t* man_to_woman;      // [men]
    uint32_t* woman_to_man;      // [women]
    uint32_t* husband_dist;      // [women] f32 bits
    const uint32_t* ring;        // this generation, [length]
    uint32_t length;
    uint64_t* keys;              // [length]
    uint32_t* values;            // [length]
    uint32_t* next;              // [length]
    unsigned long long* engagements;
};

__global__ void __launch_bounds__(256) join_propose_kernel(JoinParams p) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.length)
        return;
    const uint32_t man = p.ring[i];
    const uint32_t c = p.made[man];
    uint64_t key = ~0ull;
    uint32_t value = 0, next = kJoinNone;
    if (c < p.proposals) {  // a man who has used up his proposals leaves the ring (index.hpp:4440-4442)
        p.made[man] = c + 1;
        if ((int32_t)c < p.cand_counts[man]) {
            key = (uint64_t)p.cand_slots[(size_t)man * p.proposals + c] << 32 | i;
            value = p.cand_dist[(size_t)man * p.proposals + c];
        } else {
            // fewer answers than proposals: the reference asks the last woman again (candidates.back()), whose husband
            // is by now at least as near as this man -- he is sent to the back of the ring
            next = man;
        }
    }
    p.keys[i] = key;
    p.values[i] = value;
    p.next[i] = next;
}

// keys / values: sorted.  The thread at the first proposal of a woman settles all of them.
__global__ void __launch_bounds__(256) join_settle_kernel(JoinParams p) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.length)
        return;
    const uint64_t first = p.keys[t];
    if (first == ~0ull)
        return;
    const uint32_t woman = (uint32_t)(first >> 32);
    if (t && (uint32_t)(p.keys[t - 1] >> 32) == woman)
        return;
    uint32_t holder = p.woman_to_man[woman], holder_dist = p.husband_dist[woman];
    unsigned long long engaged = 0;
    for (uint32_t u = t; u < p.length; ++u) {
        const uint64_t key = p.keys[u];
        if ((uint32_t)(key >> 32) != woman)
            break;
        const uint32_t i = (uint32_t)key, man = p.ring[i], dist = p.values[u];
        // non-negative floats order like their bits; strict: a tie keeps the husband (index.hpp:4467)
        if (holder == kJoinNone || holder_dist > dist) {
            if (holder != kJoinNone) {
                p.man_to_woman[holder] = kJoinNone;
                p.next[i] = holder;
            }
            holder = man, holder_dist = dist;
            ++engaged;
        } else {
            p.next[i] = man;
        }
    }
    p.woman_to_man[woman] = holder;
    p.husband_dist[woman] = holder_dist;
    p.man_to_woman[holder] = woman;
    if (engaged)
        atomicAdd(p.engagements, engaged);
}

__global__ void __launch_bounds__(256) join_init_kernel(uint32_t* ring, uint32_t men) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < men)
        ring[i] = i;
}

struct JoinIsMan {
    __host__ __device__ bool operator()(uint32_t v) const { return v != kJoinNone; }
};

}  // namespace rb
