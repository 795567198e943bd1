/*
 * rad_oracle.c -- CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (rad_b200/) never links, imports or calls it.
 *
 * What it restates (reference = keiserlab/rad @ a21d60b, vendored usearch fork 2.17.2):
 *   ro_tanimoto        simsimd/include/simsimd/binary.h:99-105 (jaccard_b8: popcount of AND and OR,
 *                      1 - (f64)i/(f64)u, u==0 -> 1) followed by the f64->f32 narrowing of
 *                      include/usearch/index_plugins.hpp:1886-1893 (invoke_simsimd).
 *   ro_levels          include/usearch/index.hpp:3867-3871 (choose_random_level_) over
 *                      std::default_random_engine (= minstd_rand0, seed 1, index.hpp:2187) and
 *                      libstdc++'s uniform_real_distribution<double> (two engine draws per level).
 *   ro_exact_search    include/usearch/index.hpp:4222-4239 (search_exact_) with the bounded sorted
 *                      insert of index.hpp:923-934 (lower_bound => a later slot wins a distance tie).
 *   ro_index_*         include/usearch/index.hpp:2734-2834 (add), 3935-3975 (search_for_one_),
 *                      3982-4051 (search_to_insert_), 4147-4217 (search_to_find_in_base_),
 *                      4246-4289 (refine_), 3793-3814 (form_links_to_closest_), 3816-3865
 *                      (form_reverse_links_), 2970-3029 (search); heaps per index.hpp:659-951.
 *                      Single-threaded (the reference with threads=1).  This restatement keeps the
 *                      reference's buffer aliasing in form_reverse_links_: the view of the freshly
 *                      selected neighbours lives in the same `top` buffer that the re-prune of a
 *                      saturated neighbour overwrites, so later reverse links go to whatever that
 *                      buffer holds.  Reproducing it is what makes the graph bit-identical to the
 *                      reference build (pinned by tests/test_oracle_vs_ref.py and tests/golden/).
 *   ro_index_save      include/usearch/index_dense.hpp:1025-1093 + index.hpp:3236-3277 (.usearch v2.17).
 *
 * Parity status: PINNED -- checked against the compiled reference (oracle/_ref, built by
 * oracle/build_ref.sh from /root/reference) and against golden vectors generated from it
 * (tests/golden/, scripts/make_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define RO_EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------ distance */

static inline uint32_t pop64(uint64_t x) { return (uint32_t)__builtin_popcountll(x); }

static inline void ro_counts(const uint8_t* a, const uint8_t* b, size_t bytes, uint32_t* inter, uint32_t* uni) {
    uint32_t i = 0, u = 0;
    size_t w = 0;
    for (; w + 8 <= bytes; w += 8) {
        uint64_t x, y;
        memcpy(&x, a + w, 8);
        memcpy(&y, b + w, 8);
        i += pop64(x & y);
        u += pop64(x | y);
    }
    for (; w < bytes; ++w) {
        i += pop64((uint64_t)(a[w] & b[w]));
        u += pop64((uint64_t)(a[w] | b[w]));
    }
    *inter = i;
    *uni = u;
}

RO_EXPORT float ro_tanimoto(const uint8_t* a, const uint8_t* b, size_t bytes) {
    uint32_t i, u;
    ro_counts(a, b, bytes, &i, &u);
    double d = (u != 0) ? 1 - (double)i / (double)u : 1;
    return (float)d;
}

RO_EXPORT void ro_tanimoto_counts(const uint8_t* a, const uint8_t* b, size_t bytes, uint32_t* inter, uint32_t* uni) {
    ro_counts(a, b, bytes, inter, uni);
}

/* ------------------------------------------------------------------ level draw */

/* minstd_rand0: x <- 16807 x mod (2^31 - 1); min 1, max 2^31 - 2. */
static inline uint32_t minstd_next(uint32_t* state) {
    uint64_t x = (uint64_t)(*state) * 16807ull % 2147483647ull;
    *state = (uint32_t)x;
    return (uint32_t)x;
}

/* libstdc++ generate_canonical<double,53>(minstd_rand0): two draws, low digit first. */
static inline double canonical_double(uint32_t* state) {
    const double range = 2147483646.0; /* max - min + 1 */
    double sum = 0, scale = 1;
    for (int k = 0; k < 2; ++k) {
        sum += (double)(minstd_next(state) - 1u) * scale;
        scale *= range;
    }
    double r = sum / scale;
    if (r >= 1.0)
        r = nextafter(1.0, 0.0);
    return r;
}

static inline int16_t draw_level(uint32_t* state, double inverse_log_connectivity) {
    double r = -log(canonical_double(state)) * inverse_log_connectivity;
    return (int16_t)r;
}

/* state: in/out engine state (start with 1). */
RO_EXPORT void ro_levels(uint64_t n, uint32_t connectivity, uint32_t* state, int16_t* out) {
    double inv = 1.0 / log((double)connectivity);
    for (uint64_t i = 0; i < n; ++i)
        out[i] = draw_level(state, inv);
}

/* ------------------------------------------------------------------ candidate containers */

typedef struct {
    float d;
    uint32_t slot;
} cand_t;

typedef struct {
    cand_t* e;
    size_t n, cap;
} buf_t;

static void buf_reserve(buf_t* b, size_t cap) {
    if (cap <= b->cap)
        return;
    b->e = (cand_t*)realloc(b->e, cap * sizeof(cand_t));
    b->cap = cap;
}

/* first index whose distance is not smaller than d (ties: new element goes BEFORE equals) */
static size_t sorted_lower_bound(const buf_t* b, float d) {
    size_t lo = 0, hi = b->n;
    while (lo < hi) {
        size_t mid = lo + (hi - lo) / 2;
        if (b->e[mid].d < d)
            lo = mid + 1;
        else
            hi = mid;
    }
    return lo;
}

static void sorted_insert_reserved(buf_t* b, cand_t c) {
    size_t at = b->n ? sorted_lower_bound(b, c.d) : 0;
    memmove(b->e + at + 1, b->e + at, (b->n - at) * sizeof(cand_t));
    b->e[at] = c;
    b->n++;
}

static int sorted_insert_limited(buf_t* b, cand_t c, size_t limit) {
    size_t at = b->n ? sorted_lower_bound(b, c.d) : 0;
    if (at == limit)
        return 0;
    size_t full = (b->n == limit);
    size_t to_move = b->n - at - full;
    memmove(b->e + at + 1, b->e + at, to_move * sizeof(cand_t));
    b->e[at] = c;
    b->n += !full;
    return 1;
}

/* binary max-heap on .d (used with negated distances => pops the closest) */
static void heap_push(buf_t* h, cand_t c) {
    if (h->n + 1 > h->cap)
        buf_reserve(h, h->cap ? h->cap * 2 : 64);
    size_t i = h->n++;
    h->e[i] = c;
    while (i) {
        size_t p = (i - 1) / 2;
        if (!(h->e[p].d < h->e[i].d))
            break;
        cand_t t = h->e[p];
        h->e[p] = h->e[i];
        h->e[i] = t;
        i = p;
    }
}

static cand_t heap_pop(buf_t* h) {
    cand_t top = h->e[0];
    h->e[0] = h->e[h->n - 1];
    h->e[h->n - 1] = top;
    h->n--;
    size_t i = 0;
    for (;;) {
        size_t best = i, l = 2 * i + 1, r = 2 * i + 2;
        if (l < h->n && h->e[best].d < h->e[l].d)
            best = l;
        if (r < h->n && h->e[best].d < h->e[r].d)
            best = r;
        if (best == i)
            break;
        cand_t t = h->e[i];
        h->e[i] = h->e[best];
        h->e[best] = t;
        i = best;
    }
    return top;
}

/* ------------------------------------------------------------------ exact search */

RO_EXPORT void ro_exact_search(const uint8_t* db, uint64_t n, size_t bytes, const uint8_t* queries, uint64_t nq,
                               size_t k, uint64_t* out_slots, float* out_dists, int64_t* out_counts) {
    buf_t top = {0};
    buf_reserve(&top, k + 1);
    for (uint64_t q = 0; q < nq; ++q) {
        top.n = 0;
        const uint8_t* qv = queries + q * bytes;
        for (uint64_t s = 0; s < n; ++s) {
            cand_t c = {ro_tanimoto(qv, db + s * bytes, bytes), (uint32_t)s};
            sorted_insert_limited(&top, c, k);
        }
        for (size_t j = 0; j < top.n; ++j) {
            out_slots[q * k + j] = top.e[j].slot;
            out_dists[q * k + j] = top.e[j].d;
        }
        out_counts[q] = (int64_t)top.n;
    }
    free(top.e);
}

/* ------------------------------------------------------------------ HNSW (single-threaded) */

typedef struct {
    size_t bytes;         /* bytes per fingerprint */
    uint32_t m, m0;       /* connectivity, connectivity_base (= 2 m) */
    size_t ef_add, ef_search;
    uint64_t n, cap;
    uint8_t* vecs;        /* n x bytes */
    uint64_t* keys;
    int16_t* levels;
    uint32_t** links;     /* per node: [1 + m0] then level x [1 + m]  (count first) */
    int32_t max_level;
    uint64_t entry;
    uint32_t rng;
    double inv_log_m;
    /* scratch */
    buf_t top, next;
    uint8_t* visited;
    uint32_t* touched;
    size_t n_touched;
    /* the reference's own work counters (index.hpp:2183-2191) */
    uint64_t computed_distances, iteration_cycles, distances_in_refines, distances_in_reverse_refines;
} ro_index;

static uint32_t* links_at(const ro_index* ix, uint32_t slot, int level) {
    uint32_t* t = ix->links[slot];
    return level ? t + (1 + ix->m0) + (size_t)(level - 1) * (1 + ix->m) : t;
}

static inline float measure_q(ro_index* ix, const uint8_t* q, uint32_t slot) {
    ix->computed_distances++;
    return ro_tanimoto(q, ix->vecs + (size_t)slot * ix->bytes, ix->bytes);
}
static inline float measure_ss(ro_index* ix, uint32_t a, uint32_t b) {
    ix->computed_distances++;
    return ro_tanimoto(ix->vecs + (size_t)a * ix->bytes, ix->vecs + (size_t)b * ix->bytes, ix->bytes);
}

static int visit_set(ro_index* ix, uint32_t slot) { /* returns previous state */
    if (ix->visited[slot])
        return 1;
    ix->visited[slot] = 1;
    ix->touched[ix->n_touched++] = slot;
    return 0;
}
static void visit_clear(ro_index* ix) {
    for (size_t i = 0; i < ix->n_touched; ++i)
        ix->visited[ix->touched[i]] = 0;
    ix->n_touched = 0;
}

RO_EXPORT ro_index* ro_index_new(size_t bytes, uint32_t connectivity, size_t ef_add, size_t ef_search) {
    ro_index* ix = (ro_index*)calloc(1, sizeof(ro_index));
    ix->bytes = bytes;
    ix->m = connectivity;
    ix->m0 = connectivity * 2;
    ix->ef_add = ef_add;
    ix->ef_search = ef_search;
    ix->rng = 1;
    ix->inv_log_m = 1.0 / log((double)connectivity);
    return ix;
}

RO_EXPORT void ro_index_free(ro_index* ix) {
    if (!ix)
        return;
    for (uint64_t i = 0; i < ix->n; ++i)
        free(ix->links[i]);
    free(ix->links);
    free(ix->vecs);
    free(ix->keys);
    free(ix->levels);
    free(ix->visited);
    free(ix->touched);
    free(ix->top.e);
    free(ix->next.e);
    free(ix);
}

RO_EXPORT void ro_index_reserve(ro_index* ix, uint64_t cap) {
    if (cap <= ix->cap)
        return;
    ix->vecs = (uint8_t*)realloc(ix->vecs, cap * ix->bytes);
    ix->keys = (uint64_t*)realloc(ix->keys, cap * sizeof(uint64_t));
    ix->levels = (int16_t*)realloc(ix->levels, cap * sizeof(int16_t));
    ix->links = (uint32_t**)realloc(ix->links, cap * sizeof(uint32_t*));
    ix->visited = (uint8_t*)realloc(ix->visited, cap);
    memset(ix->visited + ix->cap, 0, cap - ix->cap);
    ix->touched = (uint32_t*)realloc(ix->touched, cap * sizeof(uint32_t));
    ix->cap = cap;
}

/* greedy descent on the levels above the target (index.hpp:3935-3975) */
static uint32_t search_for_one(ro_index* ix, const uint8_t* q, uint32_t closest, int begin_level, int end_level) {
    float closest_d = measure_q(ix, q, closest);
    for (int level = begin_level; level > end_level; --level) {
        int changed;
        do {
            changed = 0;
            const uint32_t* nb = links_at(ix, closest, level);
            uint32_t cnt = nb[0];
            for (uint32_t j = 0; j < cnt; ++j) {
                uint32_t c = nb[1 + j];
                float d = measure_q(ix, q, c);
                if (d < closest_d) {
                    closest_d = d;
                    closest = c;
                    changed = 1;
                }
            }
            ix->iteration_cycles++;
        } while (changed);
    }
    return closest;
}

/* ef-bounded best-first search on one level (index.hpp:3982-4051 and its read-only twin 4147-4217) */
static void search_level(ro_index* ix, const uint8_t* q, uint32_t start, int level, size_t top_limit) {
    visit_clear(ix);
    ix->next.n = 0;
    ix->top.n = 0;
    float radius = measure_q(ix, q, start);
    heap_push(&ix->next, (cand_t){-radius, start});
    sorted_insert_reserved(&ix->top, (cand_t){radius, start});
    visit_set(ix, start);
    while (ix->next.n) {
        cand_t best = ix->next.e[0];
        if ((-best.d) > radius && ix->top.n == top_limit)
            break;
        heap_pop(&ix->next);
        ix->iteration_cycles++;
        const uint32_t* nb = links_at(ix, best.slot, level);
        uint32_t cnt = nb[0];
        for (uint32_t j = 0; j < cnt; ++j) {
            uint32_t s = nb[1 + j];
            if (visit_set(ix, s))
                continue;
            float d = measure_q(ix, q, s);
            if (ix->top.n < top_limit || d < radius) {
                heap_push(&ix->next, (cand_t){-d, s});
                sorted_insert_limited(&ix->top, (cand_t){d, s}, top_limit);
                radius = ix->top.e[ix->top.n - 1].d;
            }
        }
    }
}

/* neighbour-selection heuristic, in place on ix->top (index.hpp:4246-4289) */
static size_t refine(ro_index* ix, size_t needed, uint64_t* counter) {
    cand_t* t = ix->top.e;
    size_t count = ix->top.n;
    if (count < needed)
        return count;
    size_t kept = 1, seen = 1;
    while (kept < needed && seen < count) {
        cand_t c = t[seen];
        int good = 1;
        size_t j = 0;
        for (; j < kept; ++j) {
            float between = measure_ss(ix, c.slot, t[j].slot);
            if (between < c.d) {
                good = 0;
                break;
            }
        }
        *counter += j;
        if (good)
            t[kept++] = t[seen];
        seen++;
    }
    ix->top.n = kept;
    return kept;
}

RO_EXPORT int ro_index_add(ro_index* ix, uint64_t key, const uint8_t* vec) {
    if (ix->n >= ix->cap)
        ro_index_reserve(ix, ix->cap ? ix->cap * 2 : 1024);
    size_t top_limit = ix->ef_add > (size_t)ix->m0 + 1 ? ix->ef_add : (size_t)ix->m0 + 1;
    buf_reserve(&ix->top, top_limit + 2);
    buf_reserve(&ix->next, top_limit + 2);

    int max_level = ix->max_level;
    uint32_t entry = (uint32_t)ix->entry;
    int level = draw_level(&ix->rng, ix->inv_log_m);

    uint32_t slot = (uint32_t)ix->n++;
    size_t words = (1 + ix->m0) + (size_t)level * (1 + ix->m);
    ix->links[slot] = (uint32_t*)calloc(words, sizeof(uint32_t));
    ix->keys[slot] = key;
    ix->levels[slot] = (int16_t)level;
    memcpy(ix->vecs + (size_t)slot * ix->bytes, vec, ix->bytes);
    if (slot == 0) {
        ix->entry = 0;
        ix->max_level = level;
        return 0;
    }

    const uint8_t* q = ix->vecs + (size_t)slot * ix->bytes;
    uint32_t closest = search_for_one(ix, q, entry, max_level, level);
    for (int l = level < max_level ? level : max_level; l >= 0; --l) {
        search_level(ix, q, closest, l, ix->ef_add);
        /* forward links */
        size_t kept = refine(ix, ix->m, &ix->distances_in_refines);
        uint32_t* mine = links_at(ix, slot, l);
        mine[0] = 0;
        for (size_t j = 0; j < kept; ++j)
            mine[1 + mine[0]++] = ix->top.e[j].slot;
        closest = ix->top.e[0].slot;
        /* reverse links; `view` aliases ix->top.e on purpose (see file header) */
        const cand_t* view = ix->top.e;
        size_t cap = l ? ix->m : ix->m0;
        for (size_t j = 0; j < kept; ++j) {
            uint32_t close_slot = view[j].slot;
            if (close_slot == slot)
                continue;
            uint32_t* theirs = links_at(ix, close_slot, l);
            if (theirs[0] < cap) {
                theirs[1 + theirs[0]++] = slot;
                continue;
            }
            ix->top.n = 0;
            sorted_insert_reserved(&ix->top, (cand_t){measure_q(ix, q, close_slot), slot});
            for (uint32_t t = 0; t < theirs[0]; ++t)
                sorted_insert_reserved(&ix->top, (cand_t){measure_ss(ix, close_slot, theirs[1 + t]), theirs[1 + t]});
            uint32_t old = theirs[0];
            memset(theirs, 0, (1 + old) * sizeof(uint32_t));
            size_t keep2 = refine(ix, cap, &ix->distances_in_reverse_refines);
            for (size_t t = 0; t < keep2; ++t)
                theirs[1 + theirs[0]++] = ix->top.e[t].slot;
        }
    }
    if (level > max_level) {
        ix->entry = slot;
        ix->max_level = level;
    }
    return 0;
}

RO_EXPORT void ro_index_add_many(ro_index* ix, const uint64_t* keys, const uint8_t* vecs, uint64_t n) {
    ro_index_reserve(ix, ix->n + n);
    for (uint64_t i = 0; i < n; ++i)
        ro_index_add(ix, keys[i], vecs + i * ix->bytes);
}

/* approximate k-NN (index.hpp:2970-3029): rows beyond counts[q] are left untouched */
RO_EXPORT void ro_index_search(ro_index* ix, const uint8_t* queries, uint64_t nq, size_t k, uint64_t* out_keys,
                               uint64_t* out_slots, float* out_dists, int64_t* out_counts) {
    size_t ef = ix->ef_search > k ? ix->ef_search : k;
    buf_reserve(&ix->top, ef + 2);
    buf_reserve(&ix->next, ef + 2);
    for (uint64_t qi = 0; qi < nq; ++qi) {
        const uint8_t* q = queries + qi * ix->bytes;
        if (!ix->n) {
            out_counts[qi] = 0;
            continue;
        }
        uint32_t closest = search_for_one(ix, q, (uint32_t)ix->entry, ix->max_level, 0);
        search_level(ix, q, closest, 0, ef);
        size_t cnt = ix->top.n < k ? ix->top.n : k;
        for (size_t j = 0; j < cnt; ++j) {
            out_slots[qi * k + j] = ix->top.e[j].slot;
            out_keys[qi * k + j] = ix->keys[ix->top.e[j].slot];
            out_dists[qi * k + j] = ix->top.e[j].d;
        }
        out_counts[qi] = (int64_t)cnt;
    }
}

RO_EXPORT uint64_t ro_index_size(const ro_index* ix) { return ix->n; }
RO_EXPORT int32_t ro_index_max_level(const ro_index* ix) { return ix->max_level; }
RO_EXPORT uint64_t ro_index_entry(const ro_index* ix) { return ix->entry; }
RO_EXPORT void ro_index_set_ef(ro_index* ix, size_t ef_add, size_t ef_search) {
    ix->ef_add = ef_add;
    ix->ef_search = ef_search;
}
RO_EXPORT void ro_index_levels(const ro_index* ix, int16_t* out) { memcpy(out, ix->levels, ix->n * sizeof(int16_t)); }
RO_EXPORT void ro_index_counters(const ro_index* ix, uint64_t* out4) {
    out4[0] = ix->computed_distances;
    out4[1] = ix->iteration_cycles;
    out4[2] = ix->distances_in_refines;
    out4[3] = ix->distances_in_reverse_refines;
}

/* neighbours of (slot, level) as slots; returns count or -1 when out of range (python/lib.cpp:270-288) */
RO_EXPORT int64_t ro_index_neighbors(const ro_index* ix, uint64_t slot, int level, uint32_t* out, size_t cap) {
    if (slot >= ix->n || level < 0 || level > ix->levels[slot])
        return -1;
    const uint32_t* nb = links_at(ix, (uint32_t)slot, level);
    uint32_t cnt = nb[0];
    for (uint32_t j = 0; j < cnt && j < cap; ++j)
        out[j] = nb[1 + j];
    return cnt;
}

/* .usearch v2.17 writer (index_dense.hpp:1025-1093, index.hpp:3236-3277) */
RO_EXPORT int ro_index_save(const ro_index* ix, const char* path, uint64_t dimensions_bits) {
    FILE* f = fopen(path, "wb");
    if (!f)
        return -1;
    uint32_t dims[2] = {(uint32_t)ix->n, (uint32_t)ix->bytes};
    fwrite(dims, sizeof(dims), 1, f);
    fwrite(ix->vecs, ix->bytes, ix->n, f);
    uint8_t head[64];
    memset(head, 0, sizeof(head));
    memcpy(head, "usearch", 7);
    uint16_t ver[3] = {2, 17, 2};
    memcpy(head + 7, ver, 6);
    head[13] = 't'; /* metric_kind_t::tanimoto_k */
    head[14] = 1;   /* scalar_kind_t::b1x8_k */
    head[15] = 14;  /* u64 keys */
    head[16] = 15;  /* u32 slots */
    uint64_t present = ix->n, deleted = 0;
    memcpy(head + 17, &present, 8);
    memcpy(head + 25, &deleted, 8);
    memcpy(head + 33, &dimensions_bits, 8);
    head[41] = 0; /* multi */
    fwrite(head, 64, 1, f);
    uint64_t core[5] = {ix->n, ix->m, ix->m0, (uint64_t)ix->max_level, ix->entry};
    fwrite(core, sizeof(core), 1, f);
    fwrite(ix->levels, sizeof(int16_t), ix->n, f);
    for (uint64_t s = 0; s < ix->n; ++s) {
        fwrite(&ix->keys[s], 8, 1, f);
        fwrite(&ix->levels[s], 2, 1, f);
        size_t words = (1 + ix->m0) + (size_t)ix->levels[s] * (1 + ix->m);
        fwrite(ix->links[s], 4, words, f);
    }
    fclose(f);
    return 0;
}
