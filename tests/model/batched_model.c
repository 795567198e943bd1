/*
 * batched_model.c -- sequential CPU model of rad_b200's BATCHED GPU HNSW construction.  TEST INFRASTRUCTURE.
 *
 * This is NOT the reference's algorithm (that is oracle/rad_oracle.c); it is an executable specification of what
 * the CUDA kernels in rad_b200/csrc/ compute, statement by statement, so that the GPU-built graph can be compared
 * BIT-EXACTLY with a CPU result (tests/test_gpu_build.py) and so that algorithmic choices (batch growth, caps) can
 * be evaluated for recall against the reference without a GPU (tests/test_model_recall.py).
 *
 * Algorithm (see DESIGN.md §4 for the rationale):
 *   levels    : the reference's single-stream level sequence (minstd_rand0; index.hpp:3867-3871), known up front.
 *   batches   : nodes [s, s+B) are inserted together against the frozen graph of nodes [0, s);
 *               B = clamp(s / growth_den, 1, batch_max), cut so that a batch holds at most one node that raises
 *               the maximum level.
 *   search    : per node, greedy descent on the levels above its own (index.hpp:3935-3975), then on each level
 *               l = min(level, max_level) .. 0 an ef-bounded best-first search with ONE sorted candidate list
 *               (key = f32 distance bits << 32 | slot), an "expanded" flag per entry and a Bloom filter
 *               (bloom_bits, 3 probes) as the visited set.  Each iteration expands the first TWO unexpanded entries
 *               together when rows hold at most 16 links (one otherwise); all their neighbours are tested against
 *               the filter state BEFORE any of them is added, duplicates between the two rows taken once.
 *   select    : the reference's heuristic (index.hpp:4246-4289) over the sorted list, needed = M.
 *   reverse   : per (level, target) the incoming sources of the whole batch are merged at once: appended in source
 *               order while the list has room, otherwise existing + incoming (closest 64 - deg by (distance, slot))
 *               are re-pruned with the same heuristic, needed = M0 (level 0) or M.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define BM_EXPORT __attribute__((visibility("default")))
#define EMPTY 0xFFFFFFFFu
#define CAND_MAX 64
#define INCOMING_CAP 256
#define SEED_RANK 128u

typedef struct {
    uint32_t m, m0, ef;
    uint32_t bloom_bits_log2;
    uint32_t growth_den, batch_max;
    uint32_t batch_k;     /* batch-mates offered to every (node, level) task; 0 = none */
    uint64_t n, cap;
    const uint8_t* vecs;  /* borrowed: cap x 128 bytes */
    int16_t* levels;
    uint32_t* adj0;       /* cap x m0 */
    uint32_t* adju;       /* upper rows x m */
    uint64_t* upper_off;  /* cap */
    uint64_t upper_rows;
    int32_t max_level;
    uint32_t entry;
    uint64_t computed_distances, expansions;
    uint64_t batches;
} bm_index;

static inline uint32_t pop64(uint64_t x) { return (uint32_t)__builtin_popcountll(x); }

static inline uint32_t dist_bits(const uint8_t* a, const uint8_t* b) {
    uint32_t i = 0, u = 0;
    for (int w = 0; w < 128; w += 8) {
        uint64_t x, y;
        memcpy(&x, a + w, 8);
        memcpy(&y, b + w, 8);
        i += pop64(x & y);
        u += pop64(x | y);
    }
    float f = (float)((u != 0) ? 1 - (double)i / (double)u : 1);
    uint32_t bits;
    memcpy(&bits, &f, 4);
    return bits;
}

static inline uint32_t* row_of(const bm_index* ix, uint32_t slot, int level, uint32_t* width) {
    if (level == 0) {
        *width = ix->m0;
        return ix->adj0 + (size_t)slot * ix->m0;
    }
    *width = ix->m;
    return ix->adju + (ix->upper_off[slot] + (uint64_t)(level - 1)) * ix->m;
}

static inline uint32_t row_count(const uint32_t* row, uint32_t width) {
    uint32_t c = 0;
    while (c < width && row[c] != EMPTY)
        ++c;
    return c;
}

BM_EXPORT bm_index* bm_new(uint32_t m, uint32_t ef, uint32_t bloom_bits_log2, uint32_t growth_den, uint32_t batch_max,
                           uint32_t batch_k, const uint8_t* vecs, const int16_t* levels, uint64_t n) {
    bm_index* ix = (bm_index*)calloc(1, sizeof(bm_index));
    ix->m = m;
    ix->m0 = 2 * m;
    ix->ef = ef;
    ix->bloom_bits_log2 = bloom_bits_log2;
    ix->growth_den = growth_den;
    ix->batch_max = batch_max;
    ix->batch_k = batch_k > 64 ? 64 : batch_k;
    ix->cap = n;
    ix->vecs = vecs;
    ix->levels = (int16_t*)malloc(n * sizeof(int16_t));
    memcpy(ix->levels, levels, n * sizeof(int16_t));
    ix->upper_off = (uint64_t*)malloc(n * sizeof(uint64_t));
    uint64_t rows = 0;
    for (uint64_t i = 0; i < n; ++i) {
        ix->upper_off[i] = rows;
        rows += (uint64_t)levels[i];
    }
    ix->upper_rows = rows;
    ix->adj0 = (uint32_t*)malloc(n * ix->m0 * sizeof(uint32_t));
    memset(ix->adj0, 0xFF, n * ix->m0 * sizeof(uint32_t));
    ix->adju = (uint32_t*)malloc((rows ? rows : 1) * ix->m * sizeof(uint32_t));
    memset(ix->adju, 0xFF, (rows ? rows : 1) * ix->m * sizeof(uint32_t));
    ix->max_level = -1;
    return ix;
}

BM_EXPORT void bm_free(bm_index* ix) {
    if (!ix)
        return;
    free(ix->levels);
    free(ix->upper_off);
    free(ix->adj0);
    free(ix->adju);
    free(ix);
}

/* ---- per-thread scratch */
typedef struct {
    uint64_t* list;     /* ef + 1 keys: dist_bits << 32 | slot */
    uint8_t* expanded;  /* ef + 1 */
    uint32_t len;
    uint32_t* bloom;
    uint64_t distances, expansions;
} scratch_t;

static inline void bloom_bits_of(uint32_t x, uint32_t log2bits, uint32_t out[3]) {
    uint32_t a = x * 0x9E3779B1u, b = x * 0x85EBCA77u | 1u;
    for (int j = 0; j < 3; ++j)
        out[j] = (a + (uint32_t)j * b) >> (32 - log2bits);
}
static inline int bloom_test(const uint32_t* bl, uint32_t x, uint32_t log2bits) {
    uint32_t h[3];
    bloom_bits_of(x, log2bits, h);
    return ((bl[h[0] >> 5] >> (h[0] & 31)) & (bl[h[1] >> 5] >> (h[1] & 31)) & (bl[h[2] >> 5] >> (h[2] & 31)) & 1u);
}
static inline void bloom_set(uint32_t* bl, uint32_t x, uint32_t log2bits) {
    uint32_t h[3];
    bloom_bits_of(x, log2bits, h);
    for (int j = 0; j < 3; ++j)
        bl[h[j] >> 5] |= 1u << (h[j] & 31);
}

/* insert into the bounded sorted list; returns 1 if inserted */
static int list_insert(scratch_t* sc, uint32_t ef, uint64_t key) {
    if (sc->len == ef && key >= sc->list[ef - 1])
        return 0;
    uint32_t lo = 0, hi = sc->len;
    while (lo < hi) {
        uint32_t mid = (lo + hi) / 2;
        if (sc->list[mid] < key)
            lo = mid + 1;
        else
            hi = mid;
    }
    uint32_t end = sc->len == ef ? ef - 1 : sc->len;
    memmove(sc->list + lo + 1, sc->list + lo, (size_t)(end - lo) * 8);
    memmove(sc->expanded + lo + 1, sc->expanded + lo, (size_t)(end - lo));
    sc->list[lo] = key;
    sc->expanded[lo] = 0;
    if (sc->len < ef)
        sc->len++;
    return 1;
}

static void beam_search(const bm_index* ix, scratch_t* sc, const uint8_t* q, uint32_t start, uint32_t start_bits,
                        int level, uint32_t ef) {
    uint32_t words = (1u << ix->bloom_bits_log2) / 32;
    memset(sc->bloom, 0, words * 4);
    sc->len = 0;
    list_insert(sc, ef, ((uint64_t)start_bits << 32) | start);
    bloom_set(sc->bloom, start, ix->bloom_bits_log2);
    /* rows of at most 16 links are expanded TWO list entries at a time (one warp reads both rows with one load) */
    const uint32_t row_width = level == 0 ? ix->m0 : ix->m;
    const uint32_t per_iter = row_width <= 16 ? 2 : 1;
    for (;;) {
        uint32_t open[2], no = 0;
        for (uint32_t p = 0; p < sc->len && no < per_iter; ++p)
            if (!sc->expanded[p])
                open[no++] = p;
        if (!no)
            break;
        uint32_t fresh[64], nf = 0;
        for (uint32_t e = 0; e < no; ++e) {
            sc->expanded[open[e]] = 1;
            sc->expansions++;
            uint32_t v = (uint32_t)sc->list[open[e]];
            uint32_t width;
            const uint32_t* row = row_of(ix, v, level, &width);
            /* every neighbour is tested against the filter state BEFORE this iteration; a neighbour that appears in
             * both rows is taken once */
            for (uint32_t j = 0; j < width && row[j] != EMPTY; ++j) {
                if (bloom_test(sc->bloom, row[j], ix->bloom_bits_log2))
                    continue;
                int dup = 0;
                for (uint32_t t = 0; t < nf; ++t)
                    dup |= (fresh[t] == row[j]);
                if (!dup)
                    fresh[nf++] = row[j];
            }
        }
        for (uint32_t j = 0; j < nf; ++j)
            bloom_set(sc->bloom, fresh[j], ix->bloom_bits_log2);
        for (uint32_t j = 0; j < nf; ++j) {
            uint32_t bits = dist_bits(q, ix->vecs + (size_t)fresh[j] * 128);
            sc->distances++;
            list_insert(sc, ef, ((uint64_t)bits << 32) | fresh[j]);
        }
    }
}

/* heuristic over keys[0..count) sorted ascending; writes selected keys to out, returns how many */
static uint32_t select_neighbors(const bm_index* ix, const uint64_t* keys, uint32_t count, uint32_t needed,
                                 uint64_t* out, uint64_t* distances) {
    if (count < needed) {
        memcpy(out, keys, (size_t)count * 8);
        return count;
    }
    uint32_t kept = 0;
    out[kept++] = keys[0];
    for (uint32_t c = 1; c < count && kept < needed; ++c) {
        uint32_t cs = (uint32_t)keys[c], cd = (uint32_t)(keys[c] >> 32);
        int good = 1;
        for (uint32_t j = 0; j < kept; ++j) {
            uint32_t between = dist_bits(ix->vecs + (size_t)cs * 128, ix->vecs + (size_t)(uint32_t)out[j] * 128);
            (*distances)++;
            if (between < cd) { /* non-negative floats order like their bit patterns */
                good = 0;
                break;
            }
        }
        if (good)
            out[kept++] = keys[c];
    }
    return kept;
}

typedef struct {
    uint64_t key; /* level << 32 | target */
    uint64_t val; /* dist_bits << 32 | source */
} pair_t;

static int pair_cmp(const void* a, const void* b) {
    const pair_t* x = (const pair_t*)a;
    const pair_t* y = (const pair_t*)b;
    if (x->key != y->key)
        return x->key < y->key ? -1 : 1;
    /* stable order = emission order = ascending source slot for one (level, target) */
    uint32_t sx = (uint32_t)x->val, sy = (uint32_t)y->val;
    return sx < sy ? -1 : (sx > sy);
}

static int u64_cmp(const void* a, const void* b) {
    uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
    return x < y ? -1 : (x > y);
}

static void reverse_links(bm_index* ix, pair_t* pairs, size_t np) {
    qsort(pairs, np, sizeof(pair_t), pair_cmp);
    size_t i = 0;
    while (i < np) {
        size_t j = i;
        while (j < np && pairs[j].key == pairs[i].key)
            ++j;
        int level = (int)(pairs[i].key >> 32);
        uint32_t target = (uint32_t)pairs[i].key;
        uint32_t width;
        uint32_t* row = row_of(ix, target, level, &width);
        uint32_t deg = row_count(row, width);
        /* incoming, without anything already linked */
        uint64_t inc[INCOMING_CAP];
        uint32_t ni = 0;
        for (size_t t = i; t < j && ni < INCOMING_CAP; ++t) {
            uint32_t src = (uint32_t)pairs[t].val;
            int dup = 0;
            for (uint32_t e = 0; e < deg; ++e)
                dup |= (row[e] == src);
            if (!dup)
                inc[ni++] = pairs[t].val;
        }
        if (deg + ni <= width) {
            for (uint32_t t = 0; t < ni; ++t)
                row[deg++] = (uint32_t)inc[t];
        } else {
            uint64_t cand[CAND_MAX];
            uint32_t nc = 0;
            if (ni > CAND_MAX - deg) {
                qsort(inc, ni, 8, u64_cmp);
                ni = CAND_MAX - deg;
            }
            for (uint32_t e = 0; e < deg; ++e) {
                uint32_t bits = dist_bits(ix->vecs + (size_t)target * 128, ix->vecs + (size_t)row[e] * 128);
                ix->computed_distances++;
                cand[nc++] = ((uint64_t)bits << 32) | row[e];
            }
            for (uint32_t t = 0; t < ni; ++t)
                cand[nc++] = inc[t];
            qsort(cand, nc, 8, u64_cmp);
            uint64_t sel[CAND_MAX];
            uint32_t ns = select_neighbors(ix, cand, nc, width, sel, &ix->computed_distances);
            for (uint32_t t = 0; t < width; ++t)
                row[t] = t < ns ? (uint32_t)sel[t] : EMPTY;
        }
        i = j;
    }
}

/* insert nodes [ix->n, upto) ; stops after max_batches batches when max_batches > 0 */
BM_EXPORT void bm_build(bm_index* ix, uint64_t upto, uint64_t max_batches) {
    uint64_t done_batches = 0;
    if (upto > ix->cap)
        upto = ix->cap;
    while (ix->n < upto) {
        if (max_batches && done_batches >= max_batches)
            break;
        uint64_t s = ix->n;
        if (s == 0) {
            ix->entry = 0;
            ix->max_level = ix->levels[0];
            ix->n = 1;
            ix->batches++;
            done_batches++;
            continue;
        }
        uint64_t b = s / ix->growth_den;
        if (b < 1)
            b = 1;
        if (b > ix->batch_max)
            b = ix->batch_max;
        if (s + b > upto)
            b = upto - s;
        /* at most one level-raising node per batch */
        int raised = 0;
        int32_t run_max = ix->max_level;
        for (uint64_t t = 0; t < b; ++t) {
            if (ix->levels[s + t] > run_max) {
                if (raised) {
                    b = t;
                    break;
                }
                raised = 1;
                run_max = ix->levels[s + t];
            }
        }
        /* tasks */
        size_t ntasks = 0;
        for (uint64_t t = 0; t < b; ++t) {
            int l = ix->levels[s + t];
            ntasks += (size_t)((l < ix->max_level ? l : ix->max_level) + 1);
        }
        pair_t* pairs = (pair_t*)malloc(ntasks * ix->m * sizeof(pair_t));
        size_t* task_base = (size_t*)malloc(b * sizeof(size_t));
        {
            size_t acc = 0;
            for (uint64_t t = 0; t < b; ++t) {
                task_base[t] = acc;
                int l = ix->levels[s + t];
                acc += (size_t)((l < ix->max_level ? l : ix->max_level) + 1);
            }
        }
        /* phase 1 (search_kernel of the build): every task's ef-bounded search against the frozen graph; the sorted
         * candidate lists are kept per task.  The search of the next level down starts from the closest entry found
         * in the OLD graph (batch-mates join only in phase 3). */
        const uint32_t ef = ix->ef;
        uint64_t* lists = (uint64_t*)malloc((ntasks ? ntasks : 1) * (size_t)ef * 8);
        uint32_t* lens = (uint32_t*)malloc((ntasks ? ntasks : 1) * 4);
        uint64_t dsum = 0, esum = 0;
#pragma omp parallel reduction(+ : dsum, esum)
        {
            scratch_t sc;
            sc.list = (uint64_t*)malloc((ix->ef + 2) * 8);
            sc.expanded = (uint8_t*)malloc(ix->ef + 2);
            sc.bloom = (uint32_t*)malloc((1u << ix->bloom_bits_log2) / 8);
            sc.distances = sc.expansions = 0;
#pragma omp for schedule(dynamic, 16)
            for (uint64_t t = 0; t < b; ++t) {
                uint32_t slot = (uint32_t)(s + t);
                const uint8_t* q = ix->vecs + (size_t)slot * 128;
                int lq = ix->levels[slot];
                uint32_t cur = ix->entry;
                uint32_t cur_bits = dist_bits(q, ix->vecs + (size_t)cur * 128);
                sc.distances++;
                for (int l = ix->max_level; l > lq; --l) {
                    int changed;
                    do {
                        changed = 0;
                        uint32_t width;
                        const uint32_t* row = row_of(ix, cur, l, &width);
                        uint32_t best = cur, best_bits = cur_bits;
                        for (uint32_t j = 0; j < width && row[j] != EMPTY; ++j) {
                            uint32_t bits = dist_bits(q, ix->vecs + (size_t)row[j] * 128);
                            sc.distances++;
                            if (bits < best_bits) {
                                best_bits = bits;
                                best = row[j];
                                changed = 1;
                            }
                        }
                        cur = best;
                        cur_bits = best_bits;
                        sc.expansions++;
                    } while (changed);
                }
                size_t task = task_base[t];
                for (int l = lq < ix->max_level ? lq : ix->max_level; l >= 0; --l, ++task) {
                    beam_search(ix, &sc, q, cur, cur_bits, l, ef);
                    memcpy(lists + task * ef, sc.list, (size_t)sc.len * 8);
                    lens[task] = sc.len;
                    cur = (uint32_t)sc.list[0];
                    cur_bits = (uint32_t)(sc.list[0] >> 32);
                }
            }
            dsum += sc.distances;
            esum += sc.expansions;
            free(sc.list);
            free(sc.expanded);
            free(sc.bloom);
        }
        /* phase 2 (the batch tile): the nodes of a batch never meet in the frozen graph, so every task is offered its
         * batch_k nearest EARLIER BATCH-MATES.  "Earlier" is the order in which the kernels hand out the work of a
         * batch: nodes with more levels to search first, equal counts by ascending slot (a stable counting sort) --
         * so every batch-mate in front of a node exists on all of that node's levels.  Per task: the exact top-batch_k
         * of the rows in front of the node by (distance ascending, position DESCENDING) -- the exact-search order --
         * among the rows strictly closer than the task's list entry at rank min(ef, SEED_RANK) (when the list is that
         * long). */
        const uint32_t bk = ix->batch_k;
        uint64_t* mates = NULL;  /* per task: bk keys (distance bits << 32 | slot), ascending; ~0 padding */
        if (bk) {
            mates = (uint64_t*)malloc((ntasks ? ntasks : 1) * (size_t)bk * 8);
            uint32_t* order = (uint32_t*)malloc(b * 4); /* position -> node of the batch */
            {
                size_t count[64] = {0}, first[64];
                for (uint64_t t = 0; t < b; ++t) {
                    int l = ix->levels[s + t];
                    ++count[(l < ix->max_level ? l : ix->max_level) & 63];
                }
                size_t acc = 0;
                for (int l = 63; l >= 0; --l) {
                    first[l] = acc;
                    acc += count[l];
                }
                for (uint64_t t = 0; t < b; ++t) {
                    int l = ix->levels[s + t];
                    order[first[(l < ix->max_level ? l : ix->max_level) & 63]++] = (uint32_t)t;
                }
            }
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : dsum)
            for (uint64_t pos = 0; pos < b; ++pos) {
                const uint64_t t = order[pos];
                uint32_t slot = (uint32_t)(s + t);
                const uint8_t* q = ix->vecs + (size_t)slot * 128;
                int lq = ix->levels[slot];
                size_t task = task_base[t];
                uint64_t top[64]; /* keys: distance bits << 32 | ~position, ascending */
                for (int l = lq < ix->max_level ? lq : ix->max_level; l >= 0; --l, ++task) {
                    const uint32_t seed_rank = ef < SEED_RANK ? ef : SEED_RANK;
                    uint32_t tau = lens[task] >= seed_rank ? (uint32_t)(lists[task * ef + seed_rank - 1] >> 32) : 0xFFFFFFFFu;
                    uint32_t nt = 0;
                    for (uint64_t p2 = 0; p2 < pos; ++p2) {
                        uint32_t bits = dist_bits(q, ix->vecs + (size_t)(s + order[p2]) * 128);
                        dsum++;
                        if (bits >= tau)
                            continue;
                        uint64_t key = ((uint64_t)bits << 32) | (uint32_t)~(uint32_t)p2;
                        if (nt == bk && key >= top[nt - 1])
                            continue;
                        uint32_t at = nt < bk ? nt : nt - 1;
                        while (at > 0 && top[at - 1] > key) {
                            top[at] = top[at - 1];
                            --at;
                        }
                        top[at] = key;
                        if (nt < bk)
                            ++nt;
                    }
                    uint64_t* out = mates + task * bk;
                    for (uint32_t j = 0; j < nt; ++j)
                        out[j] = (top[j] & 0xFFFFFFFF00000000ull) | (uint32_t)(s + order[~(uint32_t)top[j]]);
                    qsort(out, nt, 8, u64_cmp);
                    for (uint32_t j = nt; j < bk; ++j)
                        out[j] = ~0ull;
                }
            }
            free(order);
        }
        /* phase 3 (select_kernel): batch-mates merged into the list (the ef smallest keys survive), the heuristic,
         * forward links and the reverse-link records */
#pragma omp parallel reduction(+ : dsum)
        {
            uint64_t* merged = (uint64_t*)malloc((ef + 2) * 8);
            uint64_t* sel = (uint64_t*)malloc((ef + 2) * 8);
            uint64_t nd = 0;
#pragma omp for schedule(dynamic, 16)
            for (uint64_t t = 0; t < b; ++t) {
                uint32_t slot = (uint32_t)(s + t);
                int lq = ix->levels[slot];
                size_t task = task_base[t];
                for (int l = lq < ix->max_level ? lq : ix->max_level; l >= 0; --l, ++task) {
                    uint32_t len = lens[task];
                    memcpy(merged, lists + task * ef, (size_t)len * 8);
                    for (uint32_t j = 0; bk && j < bk && mates[task * bk + j] != ~0ull; ++j) {
                        uint64_t key = mates[task * bk + j];
                        if (len == ef && key >= merged[ef - 1])
                            continue;
                        uint32_t pos = len < ef ? len : ef - 1;
                        while (pos > 0 && merged[pos - 1] > key) {
                            merged[pos] = merged[pos - 1];
                            --pos;
                        }
                        merged[pos] = key;
                        if (len < ef)
                            ++len;
                    }
                    uint32_t ns = select_neighbors(ix, merged, len, ix->m, sel, &nd);
                    uint32_t width;
                    uint32_t* mine = row_of(ix, slot, l, &width);
                    for (uint32_t j = 0; j < width; ++j)
                        mine[j] = j < ns ? (uint32_t)sel[j] : EMPTY;
                    for (uint32_t j = 0; j < ix->m; ++j) {
                        pair_t* p = &pairs[task * ix->m + j];
                        if (j < ns) {
                            p->key = ((uint64_t)l << 32) | (uint32_t)sel[j];
                            p->val = (sel[j] & 0xFFFFFFFF00000000ull) | slot;
                        } else {
                            p->key = ~0ull;
                            p->val = 0;
                        }
                    }
                }
            }
            dsum += nd;
            free(merged);
            free(sel);
        }
        free(lists);
        free(lens);
        free(mates);
        ix->computed_distances += dsum;
        ix->expansions += esum;
        /* drop the padding pairs, then merge the reverse links */
        size_t np = 0, total = ntasks * ix->m;
        for (size_t t = 0; t < total; ++t)
            if (pairs[t].key != ~0ull)
                pairs[np++] = pairs[t];
        reverse_links(ix, pairs, np);
        free(pairs);
        free(task_base);
        for (uint64_t t = 0; t < b; ++t)
            if (ix->levels[s + t] > ix->max_level) {
                ix->max_level = ix->levels[s + t];
                ix->entry = (uint32_t)(s + t);
            }
        ix->n = s + b;
        ix->batches++;
        done_batches++;
    }
}

/* approximate search with the same kernel semantics (ef-bounded, Bloom visited) */
BM_EXPORT void bm_search(const bm_index* ix, const uint8_t* queries, uint64_t nq, uint32_t k, uint32_t ef,
                         uint32_t* out_slots, float* out_dists, int64_t* out_counts) {
    if (ef < k)
        ef = k;
#pragma omp parallel
    {
        scratch_t sc;
        sc.list = (uint64_t*)malloc((ef + 2) * 8);
        sc.expanded = (uint8_t*)malloc(ef + 2);
        sc.bloom = (uint32_t*)malloc((1u << ix->bloom_bits_log2) / 8);
        sc.distances = sc.expansions = 0;
#pragma omp for schedule(dynamic, 16)
        for (uint64_t qi = 0; qi < nq; ++qi) {
            const uint8_t* q = queries + qi * 128;
            if (!ix->n) {
                out_counts[qi] = 0;
                continue;
            }
            uint32_t cur = ix->entry;
            uint32_t cur_bits = dist_bits(q, ix->vecs + (size_t)cur * 128);
            for (int l = ix->max_level; l > 0; --l) {
                int changed;
                do {
                    changed = 0;
                    uint32_t width;
                    const uint32_t* row = row_of(ix, cur, l, &width);
                    uint32_t best = cur, best_bits = cur_bits;
                    for (uint32_t j = 0; j < width && row[j] != EMPTY; ++j) {
                        uint32_t bits = dist_bits(q, ix->vecs + (size_t)row[j] * 128);
                        if (bits < best_bits) {
                            best_bits = bits;
                            best = row[j];
                            changed = 1;
                        }
                    }
                    cur = best;
                    cur_bits = best_bits;
                } while (changed);
            }
            beam_search(ix, &sc, q, cur, cur_bits, 0, ef);
            uint32_t cnt = sc.len < k ? sc.len : k;
            for (uint32_t j = 0; j < cnt; ++j) {
                out_slots[qi * k + j] = (uint32_t)sc.list[j];
                uint32_t bits = (uint32_t)(sc.list[j] >> 32);
                memcpy(&out_dists[qi * k + j], &bits, 4);
            }
            out_counts[qi] = cnt;
        }
        free(sc.list);
        free(sc.expanded);
        free(sc.bloom);
    }
}

BM_EXPORT uint64_t bm_size(const bm_index* ix) { return ix->n; }
BM_EXPORT int32_t bm_max_level(const bm_index* ix) { return ix->max_level; }
BM_EXPORT uint32_t bm_entry(const bm_index* ix) { return ix->entry; }
BM_EXPORT uint64_t bm_batches(const bm_index* ix) { return ix->batches; }
BM_EXPORT uint64_t bm_upper_rows(const bm_index* ix) { return ix->upper_rows; }
BM_EXPORT const uint32_t* bm_adj0(const bm_index* ix) { return ix->adj0; }
BM_EXPORT const uint32_t* bm_adju(const bm_index* ix) { return ix->adju; }
BM_EXPORT void bm_counters(const bm_index* ix, uint64_t* out2) {
    out2[0] = ix->computed_distances;
    out2[1] = ix->expansions;
}

/* The cross-shard stitch of one level-0 row (rad_b200/csrc: stitch_kernel): candidates = the node's own links and the
 * remote shards' answers (global slots, EMPTY padded, duplicates allowed), ordered by (distance, slot) and pruned by
 * the same heuristic as everything else, needed = row width.  vecs holds ALL rows (global slots). */
BM_EXPORT void bm_stitch_rows(const uint8_t* vecs, uint64_t nq, const uint32_t* qslots, const uint32_t* cands,
                              uint32_t ncand, uint32_t needed, uint32_t* out) {
    bm_index fake;
    memset(&fake, 0, sizeof(fake));
    fake.vecs = vecs;
#pragma omp parallel
    {
        uint64_t* keys = (uint64_t*)malloc((size_t)ncand * 8);
        uint64_t* sel = (uint64_t*)malloc((size_t)(ncand + 1) * 8);
        uint64_t nd = 0;
#pragma omp for schedule(dynamic, 64)
        for (uint64_t r = 0; r < nq; ++r) {
            const uint8_t* q = vecs + (size_t)qslots[r] * 128;
            uint32_t nk = 0;
            for (uint32_t j = 0; j < ncand; ++j) {
                uint32_t c = cands[r * ncand + j];
                if (c == EMPTY || c == qslots[r])
                    continue;
                keys[nk++] = ((uint64_t)dist_bits(q, vecs + (size_t)c * 128) << 32) | c;
            }
            qsort(keys, nk, 8, u64_cmp);
            uint32_t uniq = 0;
            for (uint32_t j = 0; j < nk; ++j)
                if (!uniq || keys[j] != keys[uniq - 1])
                    keys[uniq++] = keys[j];
            uint32_t ns = select_neighbors(&fake, keys, uniq, needed, sel, &nd);
            for (uint32_t j = 0; j < needed; ++j)
                out[r * needed + j] = j < ns ? (uint32_t)sel[j] : EMPTY;
        }
        free(keys);
        free(sel);
    }
}
