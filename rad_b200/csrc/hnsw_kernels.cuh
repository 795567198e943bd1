// hnsw_kernels.cuh -- warp-cooperative HNSW traversal, neighbour selection and batched link updates (sm_100a).
//
// What these kernels replace in the reference (usearch/include/usearch/index.hpp):
//   warp_greedy_descent   search_for_one_            3935-3975
//   warp_beam_search      search_to_insert_ / search_to_find_in_base_   3982-4051 / 4147-4217
//                         (max_heap_gt `next` 659-830 + sorted_buffer_gt `top` 840-951 + growing_hash_set_gt visits
//                          1080-1206 collapse into ONE sorted shared-memory list with an expanded bit per entry and
//                          a shared-memory Bloom filter)
//   warp_select_neighbors refine_                    4246-4289  (via form_links_to_closest_ 3793-3814)
//   insert_kernel         index_gt::add              2734-2834  for a whole batch of nodes at once
//   reverse_link_kernel   form_reverse_links_        3816-3865  merged per (level, target) instead of per-node locks
//   search_kernel         index_gt::search           2970-3029  (approximate branch)
//
// One warp owns one query.  A fingerprint is fetched by a group of 8 lanes (one 128-byte line, LDG.128 per lane),
// so every warp-wide load instruction brings 4 complete fingerprints; up to 16 are in flight per lane-group round.
// The executable specification of these kernels is tests/model/batched_model.c (bit-exact graph parity).
#pragma once
#include "device_common.cuh"

namespace rb {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxWidth = 32;      // 2 * max connectivity
constexpr int kCandMax = 64;       // candidates considered when re-pruning a saturated list
constexpr int kIncomingCap = 256;  // incoming reverse links examined per (level, target) and batch

// key = order_key_bits << 32 | slot << 1 | expanded
__device__ __forceinline__ uint64_t make_key(uint32_t bits, uint32_t slot) {
    return ((uint64_t)bits << 32) | ((uint64_t)slot << 1);
}
__device__ __forceinline__ uint32_t key_slot(uint64_t key) { return (uint32_t)(key >> 1) & 0x7FFFFFFFu; }
__device__ __forceinline__ uint32_t key_bits(uint64_t key) { return (uint32_t)(key >> 32); }

struct WarpScratch {
    uint64_t* list;    // ef entries
    uint32_t* bloom;   // 2^bloom_log2 bits
    uint4* kept;       // kept_rows fingerprints for the selection heuristic   } alias the Bloom filter: the filter is
    uint64_t* sel;     // kMaxWidth selected keys                              } dead once a beam search has ended
    uint64_t* pend;    // 32 qualified candidates not yet merged into the list
    uint2* ids;        // 32 staging (id, link popcount) pairs
    uint32_t bloom_log2;
};

// kept_rows = fingerprints the selection heuristic may keep (= connectivity in the insert / search kernels)
__host__ __device__ inline size_t warp_scratch_bytes(uint32_t ef, uint32_t bloom_log2, uint32_t kept_rows) {
    size_t list = ((size_t)ef * 8 + 15) & ~(size_t)15;
    size_t bloom = ((size_t)1 << bloom_log2) / 8, select = (size_t)kept_rows * 128 + kMaxWidth * 8;
    return list + (bloom > select ? bloom : select) + 32 * 8 + 32 * 8;
}

__device__ __forceinline__ WarpScratch carve_scratch(unsigned char* base, uint32_t ef, uint32_t bloom_log2,
                                                     uint32_t kept_rows) {
    WarpScratch w;
    size_t list = ((size_t)ef * 8 + 15) & ~(size_t)15;
    size_t bloom = ((size_t)1 << bloom_log2) / 8, select = (size_t)kept_rows * 128 + kMaxWidth * 8;
    w.list = reinterpret_cast<uint64_t*>(base);
    base += list;
    w.pend = reinterpret_cast<uint64_t*>(base);
    base += 32 * 8;
    w.bloom = reinterpret_cast<uint32_t*>(base);
    w.kept = reinterpret_cast<uint4*>(base);
    w.sel = reinterpret_cast<uint64_t*>(base + (size_t)kept_rows * 128);
    base += bloom > select ? bloom : select;
    w.ids = reinterpret_cast<uint2*>(base);
    w.bloom_log2 = bloom_log2;
    return w;
}

// ------------------------------------------------------------------------------------------------ distances
// Lane j < cnt holds ids_j; returns to lane j the order key bits of d(query, ids_j).  cnt is warp-uniform, <= 4 * ROUNDS.
// All ROUNDS loads of a lane are issued before the first is consumed: 4 * ROUNDS 128-byte lines in flight per warp.
template <int ROUNDS>
__device__ __forceinline__ uint32_t warp_eval_n(const uint4* __restrict__ fp, uint4 q4, uint32_t ids, uint32_t cnt) {
    const int lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
    uint4 a[ROUNDS];
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        uint32_t j = r * 4 + grp;
        uint32_t id = __shfl_sync(RB_FULL_MASK, ids, j & 31);
        a[r] = (j < cnt) ? load_slice(fp, id, sub) : make_uint4(0, 0, 0, 0);
    }
    uint32_t mine = 0;  // i | u << 16 of this lane's fingerprint
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        uint32_t v = group8_inter_union(a[r], q4);
        uint32_t got = __shfl_sync(RB_FULL_MASK, v, (lane & 3) * 8);
        if ((lane >> 2) == r)
            mine = got;
    }
    // ONE divide per lane for up to 32 fingerprints (instead of one per group of four)
    return order_key_bits(mine & 0xFFFFu, mine >> 16);
}

__device__ __forceinline__ uint32_t warp_eval(const uint4* __restrict__ fp, uint4 q4, uint32_t ids, uint32_t cnt) {
    return cnt <= 16 ? warp_eval_n<4>(fp, q4, ids, cnt) : warp_eval_n<8>(fp, q4, ids, cnt);
}

// The beam-search form: a fingerprint is fetched by a group of FOUR lanes (two LDG.128 per lane, slices sub and sub + 4),
// so one round covers 8 candidates and the per-round overhead (id broadcast, address, reduction) is paid half as often.
// The per-round partial sums are reduced with a transposing butterfly: after it lane (grp, sub) holds the complete
// counts of candidate `cand` = sub * 8 + grp (sub < ROUNDS), 3 shuffles for 4 rounds instead of 4 x 3.
// qa / qb = query slices (lane & 3) and (lane & 3) + 4.  Returns the order key bits of candidate `cand` (0xFFFFFFFF: none).
// Only the intersection is counted here: lane j holds |ids_j| in `pcs` (it came with the link, GraphView::pc0 / pcu),
// pq = |query|, union = |ids_j| + |query| - intersection.
template <int ROUNDS>
__device__ __forceinline__ uint32_t warp_eval4_n(const uint4* __restrict__ fp, uint4 qa, uint4 qb, uint32_t ids,
                                                 uint32_t pcs, uint32_t pq, uint32_t cnt, uint32_t& cand) {
    static_assert(ROUNDS == 1 || ROUNDS == 2 || ROUNDS == 4, "rounds");
    const int lane = threadIdx.x & 31, sub = lane & 3, grp = lane >> 2;
    uint4 a[ROUNDS], b[ROUNDS];
    // A group without a candidate (j >= cnt) re-reads the last candidate's line -- the same line another group of this
    // very instruction requests, so it costs no traffic -- instead of zero-filling eight registers under a predicate;
    // its result is discarded below (cand >= cnt).  Rounds without any candidate are skipped (warp-uniform).
    const uint32_t jmax = cnt - 1;  // cnt >= 1
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        if (r == 0 || (uint32_t)(r * 8) < cnt) {
            uint32_t j = r * 8 + grp;
            uint32_t id = __shfl_sync(RB_FULL_MASK, ids, (j < jmax ? j : jmax) & 31);
            a[r] = load_slice(fp, id, sub);
            b[r] = load_slice(fp, id, sub + 4);
        }
    }
    uint32_t v[ROUNDS];
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        v[r] = 0;
        if (r == 0 || (uint32_t)(r * 8) < cnt)  // warp-uniform: a round without candidates costs nothing
            v[r] = popc4_and(a[r], qa) + popc4_and(b[r], qb);
    }
    uint32_t k;
    if (ROUNDS == 4) {
        const bool b0 = sub & 1, b1 = sub & 2;
        uint32_t k0 = (b0 ? v[1] : v[0]) + __shfl_xor_sync(RB_FULL_MASK, b0 ? v[0] : v[1], 1);
        uint32_t k1 = (b0 ? v[3] : v[2]) + __shfl_xor_sync(RB_FULL_MASK, b0 ? v[2] : v[3], 1);
        k = (b1 ? k1 : k0) + __shfl_xor_sync(RB_FULL_MASK, b1 ? k0 : k1, 2);
        cand = (uint32_t)(sub * 8 + grp);
    } else if (ROUNDS == 2) {
        const bool b0 = sub & 1;
        k = (b0 ? v[1] : v[0]) + __shfl_xor_sync(RB_FULL_MASK, b0 ? v[0] : v[1], 1);
        k += __shfl_xor_sync(RB_FULL_MASK, k, 2);
        cand = sub < 2 ? (uint32_t)(sub * 8 + grp) : 0xFFFFFFFFu;
    } else {
        k = v[0] + __shfl_xor_sync(RB_FULL_MASK, v[0], 1);
        k += __shfl_xor_sync(RB_FULL_MASK, k, 2);
        cand = sub == 0 ? (uint32_t)grp : 0xFFFFFFFFu;
    }
    if (cand >= cnt)
        cand = 0xFFFFFFFFu;
    const uint32_t pc = __shfl_sync(RB_FULL_MASK, pcs, cand & 31);
    return order_key_bits(k, pc + pq - k);
}

__device__ __forceinline__ uint32_t warp_eval4(const uint4* __restrict__ fp, uint4 qa, uint4 qb, uint32_t ids,
                                               uint32_t pcs, uint32_t pq, uint32_t cnt, uint32_t& cand) {
    if (cnt <= 8)
        return warp_eval4_n<1>(fp, qa, qb, ids, pcs, pq, cnt, cand);
    if (cnt <= 16)
        return warp_eval4_n<2>(fp, qa, qb, ids, pcs, pq, cnt, cand);
    return warp_eval4_n<4>(fp, qa, qb, ids, pcs, pq, cnt, cand);
}

// distance of the query to ONE slot, result in every lane
__device__ __forceinline__ uint32_t warp_eval_one(const uint4* __restrict__ fp, uint4 q4, uint32_t slot) {
    uint4 a = load_slice(fp, slot, threadIdx.x & 7);
    uint32_t v = group8_inter_union(a, q4);
    return order_key_bits(v & 0xFFFFu, v >> 16);
}

// ------------------------------------------------------------------------------------------------ Bloom filter
__device__ __forceinline__ void bloom_positions(uint32_t x, uint32_t log2bits, uint32_t h[3]) {
    uint32_t a = x * 0x9E3779B1u, b = (x * 0x85EBCA77u) | 1u;
#pragma unroll
    for (int j = 0; j < 3; ++j)
        h[j] = (a + (uint32_t)j * b) >> (32 - log2bits);
}
__device__ __forceinline__ bool bloom_test(const uint32_t* bl, uint32_t x, uint32_t log2bits) {
    uint32_t h[3];
    bloom_positions(x, log2bits, h);
    return ((bl[h[0] >> 5] >> (h[0] & 31)) & (bl[h[1] >> 5] >> (h[1] & 31)) & (bl[h[2] >> 5] >> (h[2] & 31)) & 1u) != 0;
}
__device__ __forceinline__ void bloom_set(uint32_t* bl, uint32_t x, uint32_t log2bits) {
    uint32_t h[3];
    bloom_positions(x, log2bits, h);
#pragma unroll
    for (int j = 0; j < 3; ++j)
        atomicOr(&bl[h[j] >> 5], 1u << (h[j] & 31));
}

// ------------------------------------------------------------------------------------------------ sorted list
// Insert `key` (expanded bit clear) into the ascending list of at most `ef` keys held in shared memory.
// Warp-uniform arguments; returns the insert position or 0xFFFFFFFF when the key does not qualify.
__device__ __forceinline__ uint32_t list_insert(uint64_t* list, uint32_t& len, uint32_t ef, uint64_t key) {
    const int lane = threadIdx.x & 31;
    if (len == ef && key >= (list[ef - 1] & ~1ull))
        return 0xFFFFFFFFu;
    // position = number of entries smaller than key: coarse pass over 32 pivots, fine pass inside one segment
    uint32_t step = (len + 31) >> 5;
    uint32_t pos = 0;
    if (len) {
        uint32_t pidx = (uint32_t)lane * step;
        bool below = pidx < len && (list[pidx] & ~1ull) < key;
        uint32_t seg = __popc(__ballot_sync(RB_FULL_MASK, below));  // pivots 0..seg-1 are smaller
        if (seg) {
            uint32_t base = (seg - 1) * step;  // entries [0, base] are smaller; look inside (base, base + step)
            uint32_t idx = base + 1 + lane;
            bool below2 = (uint32_t)lane + 1 < step && idx < len && (list[idx] & ~1ull) < key;
            pos = base + 1 + __popc(__ballot_sync(RB_FULL_MASK, below2));
        }
    }
    uint32_t end = (len == ef) ? ef - 1 : len;  // entries [pos, end) move up by one
    for (uint32_t hi = end; hi > pos;) {
        uint32_t lo = (hi - pos > 32) ? hi - 32 : pos;
        uint32_t idx = lo + lane;
        uint64_t v = 0;
        bool act = idx < hi;
        if (act)
            v = list[idx];
        __syncwarp();
        if (act)
            list[idx + 1] = v;
        __syncwarp();
        hi = lo;
    }
    if (lane == 0)
        list[pos] = key;
    __syncwarp();
    if (len < ef)
        ++len;
    return pos;
}

// ascending bitonic sort of one u64 per lane; `top` = false skips the final 32-wide merge (valid when lanes 16..31
// all hold ~0 and only the lower half matters)
__device__ __forceinline__ uint64_t warp_sort32(uint64_t v, bool top) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
        if (k == 32 && !top)
            break;
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            uint64_t o = __shfl_xor_sync(RB_FULL_MASK, v, j);
            bool up = (lane & k) == 0 || k == 32;
            bool lower = (lane & j) == 0;
            uint64_t lo = v < o ? v : o, hi = v < o ? o : v;
            v = (up == lower) ? lo : hi;
        }
    }
    return v;
}

// Merge the candidates of one expansion (lane j holds `key` when `valid`; expanded bit clear) into the ascending
// list in ONE pass: the result is exactly what inserting them one after the other with list_insert would leave
// (the `ef` smallest of list + candidates; keys are unique because the visited filter has no false negatives).
// Returns the smallest insert position, or 0xFFFFFFFF when nothing qualified.
__device__ __forceinline__ uint32_t list_merge(uint64_t* list, uint32_t& len, uint32_t ef, uint64_t key, bool valid) {
    const int lane = threadIdx.x & 31;
    if (len == ef)
        valid = valid && key < (list[ef - 1] & ~1ull);
    const uint32_t mask = __ballot_sync(RB_FULL_MASK, valid);
    if (!mask)
        return 0xFFFFFFFFu;
    const uint32_t cnt = __popc(mask);
    if (cnt == 1) {
        uint64_t k = __shfl_sync(RB_FULL_MASK, key, __ffs((int)mask) - 1);
        return list_insert(list, len, ef, k);
    }
    // candidates ascending in lanes 0 .. cnt-1
    uint64_t k = warp_sort32(valid ? key : ~0ull, (mask >> 16) != 0);
    // lower bound of every candidate in the list: branch-free binary search, one per lane
    uint32_t pos = 0;
    if ((uint32_t)lane < cnt) {
#pragma unroll
        for (uint32_t step = 512; step; step >>= 1) {
            uint32_t probe = pos + step;
            if (probe <= len && (list[probe - 1] & ~1ull) < k)
                pos = probe;
        }
    } else {
        pos = 0xFFFFFFFFu;
    }
    const uint32_t pmin = __shfl_sync(RB_FULL_MASK, pos, 0);
    // Candidate j lands at fin_j = pos_j + j (strictly ascending over the lanes); every other place o of the merged
    // list takes list[o - c(o)], c(o) = number of candidates that land below o.  The places [pmin, newlen) are written
    // 32 at a time from the top down -- a chunk reads at most 32 entries below itself, which the chunks after it have
    // not written yet -- and per chunk the candidates' places are ONE bit mask (REDUX.OR), so c(o) is two popcounts.
    const uint32_t newlen = (len + cnt < ef) ? len + cnt : ef;
    const uint32_t fin = (uint32_t)lane < cnt ? pos + (uint32_t)lane : 0xFFFFFFFFu;
    const uint32_t below_lane = (1u << lane) - 1u;
    for (uint32_t hi = newlen; hi > pmin;) {
        const uint32_t lo = (hi - pmin > 32) ? hi - 32 : pmin;
        const uint32_t o = lo + lane;
        const uint32_t cb = __popc(__ballot_sync(RB_FULL_MASK, fin < lo));
        const uint32_t rel = fin - lo;  // < 32: lands in this chunk
        const uint32_t cmask = __reduce_or_sync(RB_FULL_MASK, rel < 32u ? 1u << rel : 0u);
        const uint32_t c = cb + __popc(cmask & below_lane);
        const bool is_cand = (cmask >> lane) & 1u;
        const uint64_t ck = __shfl_sync(RB_FULL_MASK, k, (int)(c & 31u));
        const bool act = o < hi;
        uint64_t v = ck;
        if (act && !is_cand)
            v = list[o - c];
        __syncwarp();
        if (act)
            list[o] = v;
        __syncwarp();
        hi = lo;
    }
    len = newlen;
    return pmin;
}

struct TraversalCounters {
    uint32_t distances;
    uint32_t expansions;
};

// RB_PHASE_TIMING (development build, scripts/build_variant.sh phase -DRB_PHASE_TIMING): where a warp's cycles go inside
// warp_beam_search, summed over all warps into g_phase (printed by rb_add when RB_PHASE_PRINT is set).
#ifdef RB_PHASE_TIMING
__device__ unsigned long long g_phase[8];
#define RB_PHASE(k)                                                         \
    do {                                                                    \
        const long long now_ = clock64();                                   \
        if (lane == 0)                                                      \
            phase_smem[(threadIdx.x >> 5) * 8 + (k)] += (uint32_t)(now_ - phase_t); \
        phase_t = now_;                                                     \
    } while (0)
#else
#define RB_PHASE(k) \
    do {            \
    } while (0)
#endif

// ------------------------------------------------------------------------------------------------ greedy descent
// index.hpp:3935-3975: move to the best neighbour until no neighbour improves, on levels (from, to].
__device__ __forceinline__ void warp_greedy_descent(const GraphView& g, uint4 q4, uint32_t& cur, uint32_t& cur_bits,
                                                    int from_level, int to_level, TraversalCounters& tc) {
    const int lane = threadIdx.x & 31;
    for (int level = from_level; level > to_level; --level) {
        bool changed;
        do {
            changed = false;
            uint32_t width;
            const uint32_t* row = graph_row(g, cur, level, width);
            uint32_t nid = (uint32_t)lane < width ? __ldg(row + lane) : RB_EMPTY;
            uint32_t cnt = __popc(__ballot_sync(RB_FULL_MASK, nid != RB_EMPTY));  // rows are packed at the front
            uint32_t bits = warp_eval(g.fp, q4, nid, cnt);
            tc.distances += cnt;
            tc.expansions += 1;
            // first strict improvement in list order wins ties exactly like the sequential scan: min over (bits, lane)
            uint64_t mine = (uint32_t)lane < cnt ? (((uint64_t)bits << 32) | (uint32_t)lane) : ~0ull;
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                uint64_t other = __shfl_xor_sync(RB_FULL_MASK, mine, o);
                mine = other < mine ? other : mine;
            }
            uint32_t best_bits = (uint32_t)(mine >> 32);
            if (mine != ~0ull && best_bits < cur_bits) {
                cur = __shfl_sync(RB_FULL_MASK, nid, (int)(mine & 31));
                cur_bits = best_bits;
                changed = true;
            }
        } while (changed);
    }
}

// ------------------------------------------------------------------------------------------------ beam search
// ef-bounded best-first search on one level.  Leaves the ascending candidate list in w.list and returns its length.
__device__ __forceinline__ uint32_t warp_beam_search(const GraphView& g, const WarpScratch& w, uint4 q4, uint32_t start,
                                                     uint32_t start_bits, int level, uint32_t ef,
                                                     TraversalCounters& tc) {
    const int lane = threadIdx.x & 31;
#ifdef RB_PHASE_TIMING
    __shared__ uint32_t phase_smem[kWarpsPerBlock * 8];
    if (lane < 8)
        phase_smem[(threadIdx.x >> 5) * 8 + lane] = 0;
    __syncwarp();
    long long phase_t = clock64();
#endif
    // clear the filter
    {
        uint4* b4 = reinterpret_cast<uint4*>(w.bloom);
        uint32_t n4 = (1u << w.bloom_log2) / 128;
        for (uint32_t i = lane; i < n4; i += 32)
            b4[i] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    uint32_t len = 0;
    if (lane == 0) {
        w.list[0] = make_key(start_bits, start);
        bloom_set(w.bloom, start, w.bloom_log2);
    }
    len = 1;
    __syncwarp();
    uint32_t hint = 0;  // every entry before `hint` is expanded
    // Candidates that passed the filter wait in w.pend (lane j's copy in `pk`) and are merged 32 at a time.  The
    // traversal is exactly the one of a list kept merged after every expansion: the entries expanded next are the
    // smallest unexpanded keys of list + pending, and pending is merged first whenever membership of the `ef` best
    // could depend on it (see the flush rule below).
    uint32_t pn = 0;
    uint64_t pk = ~0ull;
    // query slices for the 4-lane evaluation (q4 = slice lane & 7 is held by lanes 0-7)
    uint4 qa, qb;
    qa.x = __shfl_sync(RB_FULL_MASK, q4.x, lane & 3), qa.y = __shfl_sync(RB_FULL_MASK, q4.y, lane & 3);
    qa.z = __shfl_sync(RB_FULL_MASK, q4.z, lane & 3), qa.w = __shfl_sync(RB_FULL_MASK, q4.w, lane & 3);
    qb.x = __shfl_sync(RB_FULL_MASK, q4.x, (lane & 3) + 4), qb.y = __shfl_sync(RB_FULL_MASK, q4.y, (lane & 3) + 4);
    qb.z = __shfl_sync(RB_FULL_MASK, q4.z, (lane & 3) + 4), qb.w = __shfl_sync(RB_FULL_MASK, q4.w, (lane & 3) + 4);
    uint32_t pq = popc4(qa) + popc4(qb);  // |query| in every lane
    pq += __shfl_xor_sync(RB_FULL_MASK, pq, 1);
    pq += __shfl_xor_sync(RB_FULL_MASK, pq, 2);
    // rows of at most 16 links are expanded TWO list entries at a time: lanes 0-15 read one row, lanes 16-31 the other
    const uint32_t row_width = level == 0 ? g.m0 : g.m;
    const bool two = row_width <= 16;
    for (;;) {
        RB_PHASE(4);  // keep + park (and the set-up before the first iteration)
        // first (two) unexpanded entries of the list
        uint32_t pos1 = 0xFFFFFFFFu, pos2 = 0xFFFFFFFFu;
        for (uint32_t base = hint; base < len; base += 32) {
            uint32_t idx = base + lane;
            bool open = idx < len && !(w.list[idx] & 1ull);
            uint32_t mask = __ballot_sync(RB_FULL_MASK, open);
            if (mask) {
                if (pos1 == 0xFFFFFFFFu) {
                    pos1 = base + (uint32_t)__ffs((int)mask) - 1;
                    mask &= mask - 1;
                }
                if (two && mask)
                    pos2 = base + (uint32_t)__ffs((int)mask) - 1;
                if (!two || pos2 != 0xFFFFFFFFu)
                    break;
            }
        }
        if (pn) {
            // An entry at list position p has at most p + pn smaller keys, a pending key below it fewer: while the
            // last scanned position + pn < ef, everything at or below that entry is among the ef best.  Otherwise
            // (or when the list ran out of open entries) merge now and look again.
            const uint32_t last = two ? pos2 : pos1;
            if (last == 0xFFFFFFFFu || last + pn >= ef) {
                RB_PHASE(0);
                uint32_t at = list_merge(w.list, len, ef, pk, (uint32_t)lane < pn);
                if (at < hint)
                    hint = at;
                pn = 0;
                pk = ~0ull;
                RB_PHASE(5);  // merges
                continue;
            }
        }
        if (pos1 == 0xFFFFFFFFu)
            break;
        bool both = pos2 != 0xFFFFFFFFu;
        uint64_t key1 = w.list[pos1];
        uint64_t key2 = both ? w.list[pos2] : ~0ull;
        uint32_t src1 = 32, src2 = 33;  // 32/33 = list entry pos1/pos2, < 32 = pending lane
        if (pn) {
            // pending keys below the last scanned entry compete for the two places (pn > 0 implies a full scan):
            // the two smallest of them by warp-wide minimum (REDUX on the high word, then on the low word among
            // its holders; keys are unique), then a 2 + 2 merge with the list entries
            const bool comp = !(pk & 1ull) && pk < (two ? key2 : key1);
            if (__ballot_sync(RB_FULL_MASK, comp)) {
                const uint32_t hi = (uint32_t)(pk >> 32), lo = (uint32_t)pk;
                uint32_t mh = __reduce_min_sync(RB_FULL_MASK, comp ? hi : 0xFFFFFFFFu);
                uint32_t ml = __reduce_min_sync(RB_FULL_MASK, (comp && hi == mh) ? lo : 0xFFFFFFFFu);
                const uint32_t l1 = (uint32_t)__ffs((int)__ballot_sync(RB_FULL_MASK, comp && hi == mh && lo == ml)) - 1;
                const uint64_t x1 = ((uint64_t)mh << 32) | ml;
                uint64_t x2 = ~0ull;
                uint32_t l2 = 0;
                const bool comp2 = comp && (uint32_t)lane != l1;
                if (two && __ballot_sync(RB_FULL_MASK, comp2)) {
                    mh = __reduce_min_sync(RB_FULL_MASK, comp2 ? hi : 0xFFFFFFFFu);
                    ml = __reduce_min_sync(RB_FULL_MASK, (comp2 && hi == mh) ? lo : 0xFFFFFFFFu);
                    l2 = (uint32_t)__ffs((int)__ballot_sync(RB_FULL_MASK, comp2 && hi == mh && lo == ml)) - 1;
                    x2 = ((uint64_t)mh << 32) | ml;
                }
                if (x1 < key1) {
                    if (x2 < key1) {
                        key2 = x2;
                        src2 = l2;
                    } else {
                        key2 = key1;
                        src2 = src1;
                    }
                    key1 = x1;
                    src1 = l1;
                } else {  // key1 < x1 < key2
                    key2 = x1;
                    src2 = l1;
                }
            }
        }
        __syncwarp();
        // mark the chosen entries expanded, wherever they live
        if ((uint32_t)lane == src1 || ((uint32_t)lane == src2 && both)) {
            pk |= 1ull;
            w.pend[lane] = pk;
        }
        if (lane == 0) {
            if (src1 == 32)
                w.list[pos1] = key1 | 1ull;
            if (both && src2 == 33)
                w.list[pos2] = key2 | 1ull;
            else if (both && src2 == 32)
                w.list[pos1] = key2 | 1ull;
        }
        {
            const bool took1 = src1 == 32 || (both && src2 == 32), took2 = both && src2 == 33;
            hint = took2 ? pos2 + 1 : took1 ? pos1 + 1 : pos1;
        }
        tc.expansions += both ? 2 : 1;
        RB_PHASE(0);  // scan + flush check + pick + mark
        uint32_t nid = RB_EMPTY, npc = 0;
        {
            const bool upper_half = two && lane >= 16;
            const uint32_t col = two ? (uint32_t)(lane & 15) : (uint32_t)lane;
            if (col < row_width && (!upper_half || both)) {
                uint32_t width;
                const uint32_t* row = graph_row(g, key_slot(upper_half ? key2 : key1), level, width);
                nid = __ldg(row + col);
                npc = __ldg(graph_row_pc(g, row, level) + col);
            }
        }
        bool fresh = nid != RB_EMPTY && !bloom_test(w.bloom, nid, w.bloom_log2);
        if (both) {
            // a neighbour listed in both rows is taken once (the lowest lane holding it)
            uint32_t same = __match_any_sync(RB_FULL_MASK, nid);
            fresh = fresh && (uint32_t)(__ffs((int)same) - 1) == (uint32_t)lane;
        }
        __syncwarp();  // every lane has tested against the state before this iteration
        RB_PHASE(1);  // link-row loads until the filter has been consulted
        if (fresh)
            bloom_set(w.bloom, nid, w.bloom_log2);
        uint32_t mask = __ballot_sync(RB_FULL_MASK, fresh);
        uint32_t cnt = __popc(mask);
        if (!cnt)
            continue;
        if (fresh)
            w.ids[__popc(mask & ((1u << lane) - 1u))] = make_uint2(nid, npc);
        __syncwarp();
        const uint2 staged = (uint32_t)lane < cnt ? w.ids[lane] : make_uint2(0u, 0u);
        const uint32_t id = staged.x;
        __syncwarp();
        RB_PHASE(2);  // filter update + staging
        uint32_t cand;  // the candidate whose distance this lane received
        uint32_t bits = warp_eval4(g.fp, qa, qb, id, staged.y, pq, cnt, cand);
        tc.distances += cnt;
        RB_PHASE(3);  // evaluation (fingerprint loads + counts + divide)
        // qualify against the (possibly stale, hence only looser) worst list entry and park in pending
        const uint64_t key = make_key(bits, __shfl_sync(RB_FULL_MASK, id, cand & 31));
        bool keep = cand != 0xFFFFFFFFu && (len < ef || key < (w.list[ef - 1] & ~1ull));
        const uint32_t kmask = __ballot_sync(RB_FULL_MASK, keep);
        const uint32_t kc = __popc(kmask);
        if (!kc)
            continue;
        if (pn + kc > 32) {
            RB_PHASE(4);
            uint32_t at = list_merge(w.list, len, ef, pk, (uint32_t)lane < pn);
            if (at < hint)
                hint = at;
            pn = 0;
            RB_PHASE(5);
        }
        if (keep)
            w.pend[pn + __popc(kmask & ((1u << lane) - 1u))] = key;
        pn += kc;
        __syncwarp();
        pk = (uint32_t)lane < pn ? w.pend[lane] : ~0ull;
    }
#ifdef RB_PHASE_TIMING
    RB_PHASE(0);
    __syncwarp();
    if (lane < 8)
        atomicAdd(&g_phase[lane], (unsigned long long)phase_smem[(threadIdx.x >> 5) * 8 + lane]);
    if (lane == 0)
        atomicAdd(&g_phase[6], 1ull);  // searches (index 6 also took the unused slot's zero above)
#endif
    return len;
}

// ------------------------------------------------------------------------------------------------ selection heuristic
// index.hpp:4246-4289 over `count` ascending keys in shared memory `cand`; writes the selected keys to w.sel and
// returns how many (<= needed <= kMaxWidth).  Keys keep their low (expanded) bit; callers mask it.
// `q4` is the slice (lane & 7) of the fingerprint the candidates' keys were measured from: the test
// d(c, s) < d(c, q) is done on the integer counts, i_cs * u_cq > i_cq * u_cs, which is the same decision as comparing
// the order keys (they are strictly monotone in i / u and equal for equal ratios) without a divide per pair.
// `fetch(key, sub)` returns slice `sub` of the fingerprint a candidate key stands for (the graph's own rows by default;
// the cross-shard stitch also feeds candidates that live in a staging buffer).
template <typename Fetch>
__device__ __forceinline__ uint32_t warp_select_neighbors_from(const WarpScratch& w, const uint64_t* cand, uint32_t count,
                                                               uint32_t needed, TraversalCounters& tc, uint4 q4,
                                                               Fetch fetch) {
    const int lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
    if (count < needed) {
        for (uint32_t i = lane; i < count; i += 32)
            w.sel[i] = cand[i];
        __syncwarp();
        return count;
    }
    // Unions come from popcounts (|a| + |b| - intersection): |c| is counted once per candidate, |s| once per kept
    // fingerprint (lane j keeps |kept_j| in `kept_pc`), so a (candidate, kept) pair costs only the AND-POPC half.
    const uint32_t pq = group8_sum(popc4(q4));
    uint32_t kept_pc = 0;
    // the closest candidate is always kept
    {
        uint64_t k0 = cand[0];
        const uint4 a0 = fetch(k0, sub);
        if (lane < 8)
            w.kept[sub] = a0;
        if (lane == 0)
            w.sel[0] = k0;
        const uint32_t p0 = group8_sum(popc4(a0));
        if (lane == 0)
            kept_pc = p0;
    }
    __syncwarp();
    uint32_t kept = 1, c = 1;
    while (kept < needed && c < count) {
        uint32_t ci = c + grp;  // each group of 8 lanes examines one candidate
        bool have = ci < count;
        uint64_t ck = have ? cand[ci] : 0ull;
        uint4 a = have ? fetch(ck, sub) : make_uint4(0, 0, 0, 0);
        bool good = have;
        const uint32_t vq = group8_sum(popc4_and(a, q4) | (popc4(a) << 16));
        const uint32_t icq = vq & 0xFFFFu, pa = vq >> 16;
        const uint32_t ucq = (pa + pq - icq) ? (pa + pq - icq) : 1u;  // u == 0: ratio 0
        for (uint32_t j = 0; j < kept; ++j) {
            uint4 s = w.kept[j * 8 + sub];
            const uint32_t ics = group8_sum(popc4_and(a, s));
            const uint32_t u = pa + __shfl_sync(RB_FULL_MASK, kept_pc, (int)j) - ics;
            const uint32_t ucs = u ? u : 1u;
            good = good && !(ics * ucq > icq * ucs);  // counts <= 1024, unions <= 2048: no overflow
        }
        tc.distances += kept * (count - c < 4 ? count - c : 4);
        uint32_t gm = __ballot_sync(RB_FULL_MASK, good && sub == 0);  // bit 8*grp set when that candidate passes
        if (!gm) {
            c += 4;
            continue;
        }
        int first = (__ffs((int)gm) - 1) >> 3;
        if (grp == first) {
            w.kept[kept * 8 + sub] = a;
            if (sub == 0)
                w.sel[kept] = ck;
        }
        {
            const uint32_t pf = __shfl_sync(RB_FULL_MASK, pa, first * 8);
            if ((uint32_t)lane == kept)
                kept_pc = pf;
        }
        __syncwarp();
        ++kept;
        c += (uint32_t)first + 1;
    }
    return kept;
}

__device__ __forceinline__ uint32_t warp_select_neighbors(const GraphView& g, const WarpScratch& w, const uint64_t* cand,
                                                          uint32_t count, uint32_t needed, TraversalCounters& tc,
                                                          uint4 q4) {
    const uint4* fp = g.fp;
    return warp_select_neighbors_from(w, cand, count, needed, tc, q4,
                                      [fp](uint64_t key, int sub) { return load_slice(fp, key_slot(key), sub); });
}

// ------------------------------------------------------------------------------------------------ insertion
// A batch is inserted in three steps (the executable specification is tests/model/batched_model.c):
//   insert_search_kernel   every node's greedy descent and ef-bounded searches against the FROZEN graph of the nodes
//                          before the batch; the sorted candidate list of every (node, level) task is kept in global
//                          memory, and the entry at rank min(ef, kSeedRank) becomes the SEED bound of the batch tile
//   batch tile             exact_topk_i8_kernel_t<0> in its triangular form (exact_i8_kernels.cuh): the nodes of a batch
//                          never meet in the frozen graph, so every task is offered its kBatchMates nearest EARLIER
//                          batch-mates (work order: more levels first, then ascending slot), exact, on the tensor cores
//   insert_select_kernel   batch-mates merged into the list, the reference's heuristic (index.hpp:4246-4289), forward
//                          links, link popcounts and the reverse-link records
constexpr int kBatchMates = 16;
// A batch-mate is offered to a task only if it is strictly closer than the task's candidate at this rank (or at rank ef
// when ef is smaller): the heuristic keeps M <= 16 links from the front of the list, and what lies behind rank 128
// does not change the graph's recall (tests/model, scripts/model_recall_probe.py) but triples the tile's survivors.
constexpr int kSeedRank = 128;
constexpr int kMaxLevels = 32;  // levels 0 .. 31 (P(level >= 32) = M^-32 per node)

struct InsertParams {
    GraphView g;
    const int8_t* levels;      // per slot
    const uint32_t* task_rel;  // per slot: first task of the node, relative to its batch
    uint64_t* pair_keys;       // [tasks x m]  level << 32 | target      (~0 = unused)
    uint64_t* pair_vals;       // [tasks x m]  order bits << 32 | source
    uint32_t batch_start, batch_size;
    uint32_t entry;
    int32_t max_level;
    uint32_t ef;
    uint32_t bloom_log2;
    unsigned int* work_counter;
    unsigned long long* counters;  // [0] distances, [1] expansions
    const uint32_t* order;         // the t-th unit of work of the batch is node batch_start + order[t]
    // per-(level, position) arrays, level-major: entry level_off[level] + t belongs to the node at position t
    uint64_t* lists;               // [tasks x ef]   sorted candidate keys of the task
    uint32_t* lens;                // [tasks]
    unsigned long long* seeds;     // [tasks]        bound word of the batch tile (exact_i8_kernels.cuh), ~0 = none
    const uint64_t* mate_keys;     // [tasks x kBatchMates]  tile output: distance bits << 32 | ~position, ~0 padding
    uint4* batch_fp;               // [batch_size x 8]  the batch's fingerprints in work order (the tile's operand)
    uint32_t level_off[kMaxLevels];
    uint32_t use_mates;            // 0: no batch tile ran (select uses the lists alone)
};

// MINB = resident CTAs per SM the register budget is sized for: 4 (<= 128 registers, 8 loads in flight per lane) when
// shared memory allows no more, 6 or 8 when the per-warp scratch is small enough to fit them
template <int MINB>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, MINB) insert_search_kernel(InsertParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7;
    WarpScratch w = carve_scratch(smem + (size_t)warp * warp_scratch_bytes(p.ef, p.bloom_log2, p.g.m), p.ef,
                                  p.bloom_log2, p.g.m);
    TraversalCounters tc{0, 0};
    for (;;) {
        uint32_t t = 0;
        if (lane == 0)
            t = atomicAdd(p.work_counter, 1u);
        t = __shfl_sync(RB_FULL_MASK, t, 0);
        if (t >= p.batch_size)
            break;
        const uint32_t slot = p.batch_start + __ldg(p.order + t);
        const uint4 q4 = load_slice(p.g.fp, slot, sub);
        if (lane < 8)
            p.batch_fp[(size_t)t * 8 + sub] = q4;
        const int lq = p.levels[slot];
        uint32_t cur = p.entry;
        uint32_t cur_bits = warp_eval_one(p.g.fp, q4, cur);
        tc.distances += 1;
        warp_greedy_descent(p.g, q4, cur, cur_bits, p.max_level, lq, tc);
        for (int level = lq < p.max_level ? lq : p.max_level; level >= 0; --level) {
            const uint32_t len = warp_beam_search(p.g, w, q4, cur, cur_bits, level, p.ef, tc);
            const size_t idx = (size_t)p.level_off[level] + t;
            uint64_t* out = p.lists + idx * p.ef;
            // the sorted list leaves shared memory as ONE bulk-asynchronous store (cp.async.bulk, 16-byte granules: an
            // even number of entries, which an even ef always leaves room for) instead of a copy loop of the warp
            const bool bulk_out = (p.ef & 1u) == 0;
            if (bulk_out) {
                fence_async_proxy();  // every lane's list writes -> visible to the copy engine
                __syncwarp();
                if (lane == 0) {
                    bulk_store(out, w.list, ((len + 1u) & ~1u) * 8u);
                    bulk_store_commit();
                }
            } else {
                for (uint32_t i = lane; i < len; i += 32)
                    out[i] = w.list[i];
            }
            // seed of the batch tile: the exact counts of the list entry at rank min(ef, kSeedRank)
            unsigned long long seed = ~0ull;
            const uint32_t seed_rank = p.ef < (uint32_t)kSeedRank ? p.ef : (uint32_t)kSeedRank;
            if (len >= seed_rank) {
                const uint32_t v = group8_inter_union(load_slice(p.g.fp, key_slot(w.list[seed_rank - 1]), sub), q4);
                const uint32_t i_t = v & 0xFFFFu, u_t = v >> 16;
                seed = ((unsigned long long)exact_distance_bits(i_t, u_t) << 32) | (0x1FFull << 23) |
                       ((unsigned long long)i_t << 12) | u_t;
            }
            if (lane == 0) {
                p.lens[idx] = len;
                p.seeds[idx] = seed;
            }
            const uint64_t k0 = w.list[0];
            cur = key_slot(k0);
            cur_bits = key_bits(k0);
            if (bulk_out && lane == 0)
                bulk_store_wait_read();  // the list may be overwritten by the next search
            __syncwarp();
        }
    }
    if (lane == 0) {
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
        atomicAdd(p.counters + 1, (unsigned long long)tc.expansions);
    }
}

// per-warp scratch of insert_select_kernel: the list, the kept fingerprints of the heuristic and its selection
__host__ __device__ inline size_t select_scratch_bytes(uint32_t ef, uint32_t kept_rows) {
    return (((size_t)ef * 8 + 15) & ~(size_t)15) + (size_t)kept_rows * 128 + kMaxWidth * 8;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) insert_select_kernel(InsertParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7;
    WarpScratch w;
    {
        unsigned char* base = smem + (size_t)warp * select_scratch_bytes(p.ef, p.g.m);
        w.list = reinterpret_cast<uint64_t*>(base);
        base += ((size_t)p.ef * 8 + 15) & ~(size_t)15;
        w.kept = reinterpret_cast<uint4*>(base);
        w.sel = reinterpret_cast<uint64_t*>(base + (size_t)p.g.m * 128);
        w.bloom = nullptr;
        w.pend = nullptr;
        w.ids = nullptr;
        w.bloom_log2 = 0;
    }
    __shared__ unsigned long long list_bar[kWarpsPerBlock];  // one mbarrier per warp for its bulk list loads
    uint32_t bar_parity = 0;
    if (lane == 0) {
        mbar_init(&list_bar[warp], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    TraversalCounters tc{0, 0};
    for (;;) {
        uint32_t t = 0;
        if (lane == 0)
            t = atomicAdd(p.work_counter, 1u);
        t = __shfl_sync(RB_FULL_MASK, t, 0);
        if (t >= p.batch_size)
            break;
        const uint32_t slot = p.batch_start + __ldg(p.order + t);
        const uint4 q4 = load_slice(p.g.fp, slot, sub);
        const int lq = p.levels[slot];
        uint32_t task = p.task_rel[slot];
        for (int level = lq < p.max_level ? lq : p.max_level; level >= 0; --level, ++task) {
            const size_t idx = (size_t)p.level_off[level] + t;
            uint32_t len = p.lens[idx];
            {
                const uint64_t* in = p.lists + idx * p.ef;
                if ((p.ef & 1u) == 0 && len) {  // one bulk-asynchronous copy, completion counted on the warp's mbarrier
                    const uint32_t bytes = ((len + 1u) & ~1u) * 8u;
                    fence_async_proxy();
                    __syncwarp();  // the previous task's reads and writes of the list are done
                    if (lane == 0) {
                        mbar_arrive_expect_tx(&list_bar[warp], bytes);
                        bulk_load(w.list, in, bytes, &list_bar[warp]);
                    }
                    mbar_wait(&list_bar[warp], bar_parity);
                    bar_parity ^= 1u;
                } else {
                    for (uint32_t i = lane; i < len; i += 32)
                        w.list[i] = in[i];
                }
            }
            __syncwarp();
            if (p.use_mates) {
                // the tile's answers are positions in the batch's work order, ascending by (distance, -position)
                const uint64_t mk = lane < kBatchMates ? p.mate_keys[idx * kBatchMates + lane] : ~0ull;
                const uint32_t cnt = __popc(__ballot_sync(RB_FULL_MASK, mk != ~0ull));  // a prefix of the lanes
                if (cnt) {
                    const uint32_t mslot =
                        (uint32_t)lane < cnt ? p.batch_start + __ldg(p.order + (~(uint32_t)mk)) : 0u;
                    const uint32_t bits = warp_eval(p.g.fp, q4, mslot, cnt);
                    tc.distances += cnt;
                    list_merge(w.list, len, p.ef, make_key(bits, mslot), (uint32_t)lane < cnt);
                }
            }
            uint32_t ns = warp_select_neighbors(p.g, w, w.list, len, p.g.m, tc, q4);
            uint32_t width;
            uint32_t* mine = graph_row(p.g, slot, level, width);
            uint64_t sk = (uint32_t)lane < ns ? w.sel[lane] : 0ull;
            const uint32_t spc = warp_popcounts(p.g.fp, key_slot(sk), ns);
            if ((uint32_t)lane < width) {
                mine[lane] = (uint32_t)lane < ns ? key_slot(sk) : RB_EMPTY;
                graph_row_pc(p.g, mine, level)[lane] = (uint16_t)spc;
            }
            if ((uint32_t)lane < p.g.m) {
                size_t at = (size_t)task * p.g.m + lane;
                bool on = (uint32_t)lane < ns;
                p.pair_keys[at] = on ? (((uint64_t)(uint32_t)level << 32) | key_slot(sk)) : ~0ull;
                p.pair_vals[at] = on ? (((uint64_t)key_bits(sk) << 32) | slot) : 0ull;
            }
            __syncwarp();
        }
    }
    if (lane == 0)
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
}

// ------------------------------------------------------------------------------------------------ reverse links
struct ReverseParams {
    GraphView g;
    const uint64_t* pair_keys;  // sorted ascending; ~0 padding at the end
    const uint64_t* pair_vals;
    uint32_t pair_count;
    unsigned long long* counters;
    unsigned int* work_counter;  // zeroed before the launch: warps take 32 pairs at a time
};

// bitonic sort of 64 keys held two per lane (a: index lane, b: index lane + 32), ascending
__device__ __forceinline__ void warp_sort64(uint64_t& a, uint64_t& b) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j == 32) {
                // partner of index i is i ^ 32: a <-> b of the same lane; k == 64 -> ascending everywhere
                uint64_t lo = a < b ? a : b, hi = a < b ? b : a;
                a = lo;
                b = hi;
            } else {
                uint64_t pa = __shfl_xor_sync(RB_FULL_MASK, a, j);
                uint64_t pb = __shfl_xor_sync(RB_FULL_MASK, b, j);
                bool upper = (lane & j) != 0;
                bool asc_a = ((lane & k) == 0) || k == 64;         // index = lane
                bool asc_b = (((lane + 32) & k) == 0) || k == 64;  // index = lane + 32
                bool take_min_a = (asc_a != upper);
                bool take_min_b = (asc_b != upper);
                a = take_min_a ? (a < pa ? a : pa) : (a < pa ? pa : a);
                b = take_min_b ? (b < pb ? b : pb) : (b < pb ? pb : b);
            }
        }
    }
}

// one (level, target) segment of the sorted pair list, starting at pair i (its head), by one warp
__device__ __forceinline__ void reverse_link_segment(const ReverseParams& p, const WarpScratch& w, uint64_t* cand,
                                                     uint64_t* inc, uint32_t i) {
    const int lane = threadIdx.x & 31, sub = lane & 7;
    const uint64_t key = p.pair_keys[i];
    const int level = (int)(key >> 32);
    const uint32_t target = (uint32_t)key;
    TraversalCounters tc{0, 0};

    uint32_t width;
    uint32_t* row = graph_row(p.g, target, level, width);
    uint32_t have = (uint32_t)lane < width ? row[lane] : RB_EMPTY;
    uint32_t deg = __popc(__ballot_sync(RB_FULL_MASK, have != RB_EMPTY));

    // gather the incoming sources of this segment (source order), skipping anything already linked
    uint32_t ni = 0;
    for (uint32_t off = 0;; off += 32) {
        uint32_t idx = i + off + lane;
        bool in_seg = idx < p.pair_count && p.pair_keys[idx] == key;
        uint32_t seg_mask = __ballot_sync(RB_FULL_MASK, in_seg);
        uint64_t val = in_seg ? p.pair_vals[idx] : 0ull;
        bool dup = false;
        for (uint32_t e = 0; e < deg; ++e)
            dup |= (__shfl_sync(RB_FULL_MASK, have, (int)e) == (uint32_t)val);
        bool take = in_seg && !dup;
        uint32_t tmask = __ballot_sync(RB_FULL_MASK, take);
        uint32_t at = ni + __popc(tmask & ((1u << lane) - 1u));
        if (take && at < kIncomingCap)
            inc[at] = val;
        ni += __popc(tmask);
        if (seg_mask != RB_FULL_MASK)
            break;  // the segment ended inside this chunk
    }
    if (ni > kIncomingCap)
        ni = kIncomingCap;
    __syncwarp();
    if (!ni)
        return;

    if (deg + ni <= width) {
        const uint32_t src = (uint32_t)lane < ni ? (uint32_t)inc[lane] : 0u;
        const uint32_t spc = warp_popcounts(p.g.fp, src, ni);
        if ((uint32_t)lane < ni) {
            row[deg + lane] = src;
            graph_row_pc(p.g, row, level)[deg + lane] = (uint16_t)spc;
        }
        return;
    }

    // saturated: re-prune existing + the closest incoming, at most kCandMax candidates in total
    uint32_t room = kCandMax - deg;
    if (ni > room) {
        // keep the `room` smallest incoming keys (order bits << 32 | source): rank by counting, ni <= kIncomingCap
        for (uint32_t base_i = 0; base_i < ni; base_i += 32) {
            uint32_t idx = base_i + lane;
            uint64_t mine = idx < ni ? inc[idx] : ~0ull;
            uint32_t rank = 0;
            for (uint32_t j = 0; j < ni; ++j)
                rank += inc[j] < mine;
            __syncwarp();
            if (idx < ni && rank < room)
                cand[deg + rank] = ((mine >> 32) << 32) | ((uint64_t)(uint32_t)mine << 1);
        }
        ni = room;
    } else {
        if ((uint32_t)lane < ni)
            cand[deg + lane] = ((inc[lane] >> 32) << 32) | ((uint64_t)(uint32_t)inc[lane] << 1);
        if ((uint32_t)lane + 32 < ni)
            cand[deg + lane + 32] = ((inc[lane + 32] >> 32) << 32) | ((uint64_t)(uint32_t)inc[lane + 32] << 1);
    }
    // distances from the target to its current neighbours
    const uint4 t4 = load_slice(p.g.fp, target, sub);
    {
        uint32_t bits = warp_eval(p.g.fp, t4, have, deg);
        tc.distances += deg;
        if ((uint32_t)lane < deg)
            cand[lane] = make_key(bits, have);
    }
    __syncwarp();
    uint32_t nc = deg + ni;
    uint64_t a = (uint32_t)lane < nc ? cand[lane] : ~0ull;
    uint64_t b = (uint32_t)lane + 32 < nc ? cand[lane + 32] : ~0ull;
    __syncwarp();
    warp_sort64(a, b);
    cand[lane] = a;
    cand[lane + 32] = b;
    __syncwarp();
    uint32_t ns = warp_select_neighbors(p.g, w, cand, nc, width, tc, t4);
    {
        const uint32_t nb = (uint32_t)lane < ns ? key_slot(w.sel[lane]) : RB_EMPTY;
        const uint32_t spc = warp_popcounts(p.g.fp, nb, ns);
        if ((uint32_t)lane < width) {
            row[lane] = nb;
            graph_row_pc(p.g, row, level)[lane] = (uint16_t)spc;
        }
    }
    if (lane == 0)
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
}

// Persistent warps: a warp takes 32 consecutive pairs at a time, finds the segment heads among them with one
// coalesced read, and works through those segments.  (One warp per PAIR left most warps of a CTA idle from their
// first instruction -- only heads have work -- while the CTA kept its slots until its busiest warp finished.)
__global__ void __launch_bounds__(kWarpsPerBlock * 32) reverse_link_kernel(ReverseParams p) {
    __shared__ __align__(16) unsigned char smem[kWarpsPerBlock * (kMaxWidth * 128 + kMaxWidth * 8 + kCandMax * 8 +
                                                                   kIncomingCap * 8)];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* base = smem + (size_t)warp * (kMaxWidth * 128 + kMaxWidth * 8 + kCandMax * 8 + kIncomingCap * 8);
    WarpScratch w;
    w.kept = reinterpret_cast<uint4*>(base);
    w.sel = reinterpret_cast<uint64_t*>(base + kMaxWidth * 128);
    uint64_t* cand = reinterpret_cast<uint64_t*>(base + kMaxWidth * 128 + kMaxWidth * 8);
    uint64_t* inc = cand + kCandMax;
    w.list = nullptr;
    w.bloom = nullptr;
    w.ids = nullptr;
    w.bloom_log2 = 0;
    for (;;) {
        uint32_t first = 0;
        if (lane == 0)
            first = atomicAdd(p.work_counter, 32u);
        first = __shfl_sync(RB_FULL_MASK, first, 0);
        if (first >= p.pair_count)
            break;
        const uint32_t i = first + lane;
        bool head = false;
        if (i < p.pair_count) {
            const uint64_t key = p.pair_keys[i];
            head = key != ~0ull && (i == 0 || p.pair_keys[i - 1] != key);
        }
        for (uint32_t heads = __ballot_sync(RB_FULL_MASK, head); heads; heads &= heads - 1) {
            reverse_link_segment(p, w, cand, inc, first + (uint32_t)__ffs((int)heads) - 1);
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------ link popcounts
// For a graph that arrives from a file (rb_load): popcount of every fingerprint (streaming, one 8-lane group per row),
// then pc[link] = node_pc[adj[link]] for every link.  Graphs built here get their popcounts from the kernels above.
__global__ void __launch_bounds__(256) node_popcount_kernel(const uint4* __restrict__ fp, uint32_t rows,
                                                            uint16_t* __restrict__ out) {
    const uint32_t sub = threadIdx.x & 7;
    for (size_t row = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;; row += ((size_t)gridDim.x * blockDim.x) >> 3) {
        // a warp leaves the loop as a whole, so that every lane takes part in the shuffles
        const bool any = __any_sync(RB_FULL_MASK, row < rows);
        if (!any)
            break;
        uint32_t v = group8_sum(row < rows ? popc4(__ldg(fp + row * 8 + sub)) : 0u);
        if (row < rows && sub == 0)
            out[row] = (uint16_t)v;
    }
}

__global__ void __launch_bounds__(256) link_popcount_kernel(const uint32_t* __restrict__ adj, size_t links,
                                                            const uint16_t* __restrict__ node_pc,
                                                            uint16_t* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < links; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t id = __ldg(adj + i);
        out[i] = id == RB_EMPTY ? (uint16_t)0 : __ldg(node_pc + id);
    }
}

// ------------------------------------------------------------------------------------------------ cross-shard stitch
// G row shards are built independently (one per GPU); afterwards every node has searched the OTHER shards' graphs and
// received their nearest rows (global slots + fingerprints).  One warp per node: the node's own
// level-0 links and the remote candidates are ordered by (distance, global slot) and pruned by the reference's heuristic
// (index.hpp:4246-4289), needed = the row width -- what form_reverse_links_ does when a full row receives new links
// (index.hpp:3816-3865) -- and the row is emitted in GLOBAL slots.  The shard's own graph is NOT modified: the other
// shards are still searching it.  CPU statement: tests/model/batched_model.c: bm_stitch_rows.
constexpr int kStitchMax = 128;  // own links + remote candidates per node

struct StitchParams {
    GraphView g;                 // this shard: fingerprints and level-0 rows in LOCAL slots
    uint32_t first, count;       // local nodes [first, first + count)
    uint32_t slot_base;          // global slot of local slot 0
    const uint32_t* cand_slots;  // [count x cands]  GLOBAL slots of the remote candidates, RB_EMPTY = none
    const uint4* cand_fp;        // [count x cands x 8]  their fingerprints
    uint32_t cands;              // <= kStitchMax - m0
    uint32_t* rows_out;          // [count x m0]  GLOBAL slots, RB_EMPTY padded
    unsigned long long* counters;
};

__global__ void __launch_bounds__(kWarpsPerBlock * 32) stitch_kernel(StitchParams p) {
    constexpr int kPerWarp = kStitchMax * 8 * 2 + kMaxWidth * 128 + kMaxWidth * 8 + kStitchMax * 2;
    __shared__ __align__(16) unsigned char smem[kWarpsPerBlock * kPerWarp];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7, grp = lane >> 3;
    unsigned char* base = smem + (size_t)warp * kPerWarp;
    uint64_t* ckey = reinterpret_cast<uint64_t*>(base);  // order bits << 32 | global slot, unsorted
    uint64_t* sorted = ckey + kStitchMax;                // order bits << 32 | candidate index << 1, ascending
    WarpScratch w;
    w.kept = reinterpret_cast<uint4*>(base + kStitchMax * 16);
    w.sel = reinterpret_cast<uint64_t*>(base + kStitchMax * 16 + kMaxWidth * 128);
    // column of cand_slots a compacted remote entry came from
    uint16_t* origin = reinterpret_cast<uint16_t*>(base + kStitchMax * 16 + kMaxWidth * 128 + kMaxWidth * 8);
    w.list = nullptr, w.bloom = nullptr, w.pend = nullptr, w.ids = nullptr, w.bloom_log2 = 0;
    TraversalCounters tc{0, 0};
    for (uint32_t t = blockIdx.x * kWarpsPerBlock + warp; t < p.count; t += gridDim.x * kWarpsPerBlock) {
        const uint32_t node = p.first + t;
        const uint4 q4 = load_slice(p.g.fp, node, sub);
        const uint32_t m0 = p.g.m0;
        const uint32_t* row = p.g.adj0 + (size_t)node * m0;
        const uint32_t own = (uint32_t)lane < m0 ? row[lane] : RB_EMPTY;
        const uint32_t deg = __popc(__ballot_sync(RB_FULL_MASK, own != RB_EMPTY));  // rows are packed at the front
        {
            const uint32_t bits = warp_eval(p.g.fp, q4, own, deg);
            if ((uint32_t)lane < deg)
                ckey[lane] = ((uint64_t)bits << 32) | (uint64_t)(own + p.slot_base);
        }
        // remote candidates: their fingerprints travelled with them; four per round, 8 lanes each
        uint32_t n = deg;
        const uint32_t* cs = p.cand_slots + (size_t)t * p.cands;
        const uint4* cf = p.cand_fp + (size_t)t * p.cands * 8;
        for (uint32_t c0 = 0; c0 < p.cands; c0 += 4) {
            const uint32_t c = c0 + grp;
            const uint32_t gs = c < p.cands ? cs[c] : RB_EMPTY;
            const uint4 a = gs != RB_EMPTY ? __ldg(cf + (size_t)c * 8 + sub) : make_uint4(0, 0, 0, 0);
            const uint32_t v = group8_inter_union(a, q4);
            const uint32_t bits = order_key_bits(v & 0xFFFFu, v >> 16);
            const uint32_t mask = __ballot_sync(RB_FULL_MASK, gs != RB_EMPTY && sub == 0);  // bit 8 * grp
            if (gs != RB_EMPTY && sub == 0) {
                const uint32_t at = n + __popc(mask & ((1u << lane) - 1u));
                ckey[at] = ((uint64_t)bits << 32) | gs;
                origin[at] = (uint16_t)c;
            }
            n += __popc(mask);
        }
        tc.distances += n;
        __syncwarp();
        // rank by counting (keys are unique: a global slot appears once), carrying the candidate's index along
        for (uint32_t i = lane; i < n; i += 32) {
            const uint64_t mine = ckey[i];
            uint32_t rank = 0;
            for (uint32_t j = 0; j < n; ++j)
                rank += ckey[j] < mine;
            sorted[rank] = (mine & 0xFFFFFFFF00000000ull) | ((uint64_t)i << 1);
        }
        __syncwarp();
        const uint4* own_fp = p.g.fp;
        const uint32_t base_slot = p.slot_base;
        auto fetch = [own_fp, base_slot, ckey, origin, cf, deg](uint64_t key, int sl) -> uint4 {
            const uint32_t idx = (uint32_t)(key >> 1) & 0x7FFFFFFFu;
            if (idx < deg)
                return load_slice(own_fp, (uint32_t)ckey[idx] - base_slot, sl);
            return __ldg(cf + (size_t)origin[idx] * 8 + sl);
        };
        const uint32_t ns = warp_select_neighbors_from(w, sorted, n, m0, tc, q4, fetch);
        if ((uint32_t)lane < m0) {
            uint32_t out = RB_EMPTY;
            if ((uint32_t)lane < ns)
                out = (uint32_t)ckey[(uint32_t)(w.sel[lane] >> 1) & 0x7FFFFFFFu];
            p.rows_out[(size_t)t * m0 + lane] = out;
        }
        __syncwarp();
    }
    if (lane == 0 && p.counters)
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
}

// out[i] = fingerprint of local slot slots[i] (RB_EMPTY -> zeros): 8 lanes per row
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint4* __restrict__ fp, const uint32_t* __restrict__ slots,
                                                          uint64_t count, uint4* __restrict__ out) {
    const uint32_t sub = threadIdx.x & 7;
    for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3; i < count;
         i += ((uint64_t)gridDim.x * blockDim.x) >> 3) {
        const uint32_t s = __ldg(slots + i);
        out[i * 8 + sub] = s == RB_EMPTY ? make_uint4(0, 0, 0, 0) : __ldg(fp + (size_t)s * 8 + sub);
    }
}

// ------------------------------------------------------------------------------------------------ level extraction
// dst = the sub-index of src over src's nodes of level >= 1, ONE LEVEL DOWN: graph level l + 1 of src becomes level l of
// dst (its level 0 is M wide), slots renumbered by `table` (src slot -> dst slot).  The upper levels of the cross-shard
// stitch start from these links -- the ones the build already made -- instead of a second build.  One warp per node.
struct ExtractParams {
    GraphView src;
    GraphView dst;              // dst.m0 == dst.m == src.m
    uint4* dst_fp;
    const uint32_t* src_slots;  // [count] ascending
    const uint32_t* table;      // [src rows]
    const int8_t* dst_levels;   // [count] = src level - 1
    uint32_t count;
};

__global__ void __launch_bounds__(256) extract_level_kernel(ExtractParams p) {
    const int lane = threadIdx.x & 31;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < p.count; i += (gridDim.x * blockDim.x) >> 5) {
        const uint32_t s = __ldg(p.src_slots + i);
        if (lane < 8)
            p.dst_fp[(size_t)i * 8 + lane] = load_slice(p.src.fp, s, lane);
        const int lv = p.dst_levels[i];
        const size_t src_row0 = (size_t)__ldg(p.src.upper_off + s);
        const size_t dst_row0 = lv ? (size_t)__ldg(p.dst.upper_off + i) : 0;
        for (int l = 0; l <= lv; ++l) {
            if ((uint32_t)lane < p.src.m) {
                const size_t from = (src_row0 + (uint32_t)l) * p.src.m + lane;
                const uint32_t v = p.src.adju[from];
                const uint32_t w = v == RB_EMPTY ? RB_EMPTY : __ldg(p.table + v);
                const uint16_t pc = p.src.pcu[from];
                if (l == 0) {
                    p.dst.adj0[(size_t)i * p.dst.m0 + lane] = w;
                    p.dst.pc0[(size_t)i * p.dst.m0 + lane] = pc;
                } else {
                    const size_t to = (dst_row0 + (uint32_t)(l - 1)) * p.dst.m + lane;
                    p.dst.adju[to] = w;
                    p.dst.pcu[to] = pc;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ level statistics
// Edges per level of the whole graph (index.hpp:3147-3179 counts them by walking every node on the host): counted here,
// where the adjacency was produced, and handed to the host with it at rb_sync_host.  One thread per node.
__global__ void __launch_bounds__(256) edge_count_kernel(GraphView g, const int8_t* __restrict__ levels, uint32_t rows,
                                                         unsigned long long* __restrict__ out /* [kMaxLevels] */) {
    __shared__ unsigned int local[kMaxLevels];
    if (threadIdx.x < kMaxLevels)
        local[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < rows; slot += gridDim.x * blockDim.x) {
        const int lv = levels[slot];
        for (int level = 0; level <= lv && level < kMaxLevels; ++level) {
            uint32_t width;
            const uint32_t* row = graph_row(g, slot, level, width);
            uint32_t cnt = 0;
            for (uint32_t j = 0; j < width; ++j)
                cnt += row[j] != RB_EMPTY;
            if (cnt)
                atomicAdd(&local[level], cnt);
        }
    }
    __syncthreads();
    if (threadIdx.x < kMaxLevels && local[threadIdx.x])
        atomicAdd(out + threadIdx.x, (unsigned long long)local[threadIdx.x]);
}

// ------------------------------------------------------------------------------------------------ approximate search
struct SearchParams {
    GraphView g;
    const uint4* queries;  // Q x 8 uint4
    uint32_t query_count;
    uint32_t entry;
    int32_t max_level;
    uint32_t ef, k;
    uint32_t bloom_log2;
    uint32_t* slots_out;   // Q x k
    float* dists_out;      // Q x k  (reference-exact bits)
    int32_t* counts_out;   // Q
    unsigned int* work_counter;
    unsigned long long* counters;
};

template <int MINB>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, MINB) search_kernel(SearchParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7;
    WarpScratch w = carve_scratch(smem + (size_t)warp * warp_scratch_bytes(p.ef, p.bloom_log2, p.g.m), p.ef,
                                  p.bloom_log2, p.g.m);
    TraversalCounters tc{0, 0};
    for (;;) {
        uint32_t qi = 0;
        if (lane == 0)
            qi = atomicAdd(p.work_counter, 1u);
        qi = __shfl_sync(RB_FULL_MASK, qi, 0);
        if (qi >= p.query_count)
            break;
        const uint4 q4 = __ldg(p.queries + (size_t)qi * 8 + sub);
        uint32_t cur = p.entry;
        uint32_t cur_bits = warp_eval_one(p.g.fp, q4, cur);
        tc.distances += 1;
        warp_greedy_descent(p.g, q4, cur, cur_bits, p.max_level, 0, tc);
        uint32_t len = warp_beam_search(p.g, w, q4, cur, cur_bits, 0, p.ef, tc);
        uint32_t cnt = len < p.k ? len : p.k;
        // reference-exact distances for what is returned (the traversal ran on order keys)
        for (uint32_t base = 0; base < cnt; base += 4) {
            uint32_t j = base + (lane >> 3);
            uint32_t id = j < cnt ? key_slot(w.list[j]) : 0u;
            uint4 a = j < cnt ? load_slice(p.g.fp, id, sub) : make_uint4(0, 0, 0, 0);
            uint32_t v = group8_inter_union(a, q4);
            if (j < cnt && sub == 0) {
                p.slots_out[(size_t)qi * p.k + j] = id;
                p.dists_out[(size_t)qi * p.k + j] = __uint_as_float(exact_distance_bits(v & 0xFFFFu, v >> 16));
            }
        }
        if (lane == 0)
            p.counts_out[qi] = (int32_t)cnt;
        __syncwarp();
    }
    if (lane == 0) {
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
        atomicAdd(p.counters + 1, (unsigned long long)tc.expansions);
    }
}

// ------------------------------------------------------------------------------------------------ cluster_kernel
// index.hpp:3046-3079 (`cluster(query, level)`): the greedy descent alone, from the entry point down to and including
// `level` (never below level 1: neighbors_non_base_), and the reference-exact distance of where it ends.  One warp per
// query, no shared memory.  Deterministic given the graph: on a graph the reference built it is the reference's answer.
struct ClusterParams {
    GraphView g;
    const uint4* queries;
    uint32_t query_count, entry;
    int max_level, level;
    uint32_t* slots_out;  // Q
    float* dists_out;     // Q
    unsigned long long* counters;
};

__global__ void __launch_bounds__(256) cluster_kernel(ClusterParams p) {
    const int lane = threadIdx.x & 31, sub = lane & 7;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    TraversalCounters tc{0, 0};
    for (uint32_t qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; qi < p.query_count; qi += warps) {
        const uint4 q4 = __ldg(p.queries + (size_t)qi * 8 + sub);
        uint32_t cur = p.entry;
        uint32_t cur_bits = warp_eval_one(p.g.fp, q4, cur);
        tc.distances += 1;
        warp_greedy_descent(p.g, q4, cur, cur_bits, p.max_level, (p.level <= 0 ? 1 : p.level) - 1, tc);
        const uint32_t v = group8_inter_union(load_slice(p.g.fp, cur, sub), q4);
        tc.distances += 1;  // the reference measures the member once more (index.hpp:3071)
        if (lane == 0) {
            p.slots_out[qi] = cur;
            p.dists_out[qi] = __uint_as_float(exact_distance_bits(v & 0xFFFFu, v >> 16));
        }
    }
    if (lane == 0) {
        atomicAdd(p.counters + 0, (unsigned long long)tc.distances);
        atomicAdd(p.counters + 1, (unsigned long long)tc.expansions);
    }
}

}  // namespace rb
