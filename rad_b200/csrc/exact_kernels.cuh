// exact_kernels.cuh -- tiled all-pairs Tanimoto with a streaming per-query top-k (sm_100a, integer POPC pipe).
//
// Replaces the reference's exact search: index_gt::search_exact_ (usearch/include/usearch/index.hpp:4222-4239) with
// the bounded sorted insert of sorted_buffer_gt (index.hpp:923-934), i.e. ascending distance, ties by DESCENDING
// slot, the highest slots kept at the k-th boundary; and the shard merge rule of index.hpp:2625-2645.
//
// Work decomposition: grid = (database chunks, query tiles).  A CTA keeps a 64-query tile in shared memory and
// streams its chunk of the database through a double-buffered 64-row stage, from the HIGHEST rows to the lowest, so
// that a row which merely ties the current k-th distance can never displace it (lower slot loses) and the per-pair
// filter can be strict.  Each thread owns a 4 x 4 block of (query, row) pairs and accumulates the intersection
// counts with LOP3.AND + POPC over the 32 words; the union comes from popcounts computed once per row on load:
// u = |q| + |d| - i.  The per-pair epilogue is an INTEGER test against the query's current k-th best
//     i / u > i_t / u_t   <=>   i * (u_t + i_t) > i_t * |q| + i_t * |d|
// (2 IMAD + 1 ISETP).  Survivors (about k ln(n/k) per query) go through a shared-memory queue to a slow path that
// forms the reference-exact f32 distance bits with an f64 divide and inserts into the CTA's sorted partial list.
// Shared-memory rows are padded to 144 bytes so that the 128-bit reads of 8 consecutive lanes hit distinct banks.
#pragma once
#include "device_common.cuh"

namespace rb {

constexpr int kXQ = 64;         // queries per CTA
constexpr int kXD = 64;         // database rows per stage
constexpr int kXThreads = 256;  // 16 x 16 threads, 4 x 4 pairs each
constexpr int kXStride = 36;    // words per padded shared-memory row
constexpr uint32_t kXInvalid = 0xFFFFu;

struct ExactParams {
    const uint4* db;        // rows x 8 uint4
    const uint4* queries;   // Q x 8 uint4
    uint64_t rows;
    uint64_t slot_base;     // global slot of row 0 (row-sharded databases)
    uint32_t query_count;
    uint32_t k;
    uint32_t chunks;
    uint64_t chunk_rows;    // multiple of kXD
    uint64_t* partial_keys; // [chunks x Q x k]  distance bits << 32 | ~slot ; ~0 = empty
    uint32_t* partial_iu;   // [chunks x Q x k]  i | u << 16 of each entry
    unsigned long long* bounds;        // [Q], ~0 at launch (tensor-core kernels): the smallest k-th-best key any CTA has
                                       // reached for the query so far, distance bits << 32 | i << 16 | u
    unsigned long long* debug_cycles;  // optional [32] cycle counters of the tensor kernel's roles (profiling aid)
    // TRIANGULAR form (the batch tile of the build, exact_topk_i8_kernel_t<0> only): queries == db == one array in
    // work order; grid.y = level; level l covers the first tri_cnt[l] rows, a query only sees the rows IN FRONT of it,
    // its list and bound word live at entry tri_off[l] + query of partial_keys / bounds (one "chunk"), and the bound
    // words arrive SEEDED (insert_search_kernel) instead of ~0.
    uint32_t tri;
    uint32_t tri_cnt[32];
    uint32_t tri_off[32];
};

struct ExactSmem {
    uint32_t q[kXQ * kXStride];
    uint32_t d[2][kXD * kXStride];
    uint32_t pq[kXQ];
    uint32_t pd[2][kXD];
    uint4 thr[kXQ];  // x = A = u_t + i_t, y = i_t, z = B = i_t * |q|, w = slack (1 while the list is not full)
    uint32_t queue[kXQ * kXD];
    uint32_t queue_count;
    unsigned long long full[2];  // mbarriers: the bulk copies of database stage 0 / 1 have landed
};

// insert (key, iu) into an ascending list of exactly k entries (~0 = empty tail) in GLOBAL memory; warp-uniform args
__device__ __forceinline__ void glist_insert(uint64_t* keys, uint32_t* ius, uint32_t k, uint64_t key, uint32_t iu) {
    const int lane = threadIdx.x & 31;
    if (key >= keys[k - 1])
        return;
    uint32_t pos = 0;
    for (uint32_t base = 0; base < k; base += 32) {
        uint32_t idx = base + lane;
        bool below = idx < k && keys[idx] < key;
        pos += __popc(__ballot_sync(RB_FULL_MASK, below));
    }
    for (uint32_t hi = k - 1; hi > pos;) {
        uint32_t lo = (hi - pos > 32) ? hi - 32 : pos;
        uint32_t idx = lo + lane;
        bool act = idx < hi;
        uint64_t kv = 0;
        uint32_t iv = 0;
        if (act) {
            kv = keys[idx];
            iv = ius[idx];
        }
        __syncwarp();
        if (act) {
            keys[idx + 1] = kv;
            ius[idx + 1] = iv;
        }
        __syncwarp();
        hi = lo;
    }
    if (lane == 0) {
        keys[pos] = key;
        ius[pos] = iu;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kXThreads, 2) exact_topk_kernel(ExactParams p) {
    extern __shared__ __align__(16) unsigned char raw[];
    ExactSmem& s = *reinterpret_cast<ExactSmem*>(raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & 15, ty = tid >> 4;
    const uint32_t chunk = blockIdx.x;
    const uint32_t q0 = blockIdx.y * kXQ;
    const uint64_t row_begin = (uint64_t)chunk * p.chunk_rows;
    const uint64_t row_end = row_begin + p.chunk_rows < p.rows ? row_begin + p.chunk_rows : p.rows;
    // loader mapping: 8 threads per row (one uint4 each), 32 rows per pass, 2 passes per 64-row tile
    const int lrow = tid >> 3, lsub = tid & 7;

    // ---- query tile + its popcounts
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        int r = lrow + 32 * pass;
        uint32_t q = q0 + r;
        uint4 v = q < p.query_count ? __ldg(p.queries + (size_t)q * 8 + lsub) : make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(&s.q[r * kXStride + 4 * lsub]) = v;
        uint32_t pc = group8_sum(popc4(v));
        if (lsub == 0) {
            s.pq[r] = q < p.query_count ? pc : kXInvalid;
            s.thr[r] = make_uint4(0, 0, 0, 1);
        }
    }
    if (tid == 0)
        s.queue_count = 0;
    // ---- this CTA's partial lists start empty
    {
        uint32_t nq = p.query_count - q0 < (uint32_t)kXQ ? p.query_count - q0 : (uint32_t)kXQ;
        size_t base = ((size_t)chunk * p.query_count + q0) * p.k;
        for (size_t i = tid; i < (size_t)nq * p.k; i += kXThreads)
            p.partial_keys[base + i] = ~0ull;
    }

    if (row_begin >= row_end)
        return;  // (uniform) an empty chunk still publishes empty partial lists
    const uint64_t n_tiles = (row_end - row_begin + kXD - 1) / kXD;
    // The database stage is brought in by the bulk-asynchronous copy engine (cp.async.bulk, the linear form of TMA):
    // one 128-byte copy per row straight into its padded shared-memory row, completion counted in bytes on the
    // stage's mbarrier -- no registers and no LDG/STS issue slots on the POPC-bound threads.  Rows past the end of the
    // chunk are zero-filled by the threads themselves.
    if (tid == 0) {
        mbar_init(&s.full[0], 1);
        mbar_init(&s.full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto fetch = [&](int buf, uint64_t tile) {  // (the stage was last read before the preceding __syncthreads)
        if (tid < kXD) {
            const uint64_t first = row_begin + tile * kXD, row = first + tid;
            if (tid == 0) {
                const uint64_t valid = row_end - first < (uint64_t)kXD ? row_end - first : (uint64_t)kXD;
                mbar_arrive_expect_tx(&s.full[buf], (uint32_t)valid * 128u);
            }
            uint32_t* dst = &s.d[buf][tid * kXStride];
            if (row < row_end) {
                bulk_load(dst, p.db + row * 8, 128u, &s.full[buf]);
            } else {
#pragma unroll
                for (int w = 0; w < 8; ++w)
                    reinterpret_cast<uint4*>(dst)[w] = make_uint4(0, 0, 0, 0);
            }
        }
    };
    // row popcounts of a landed stage (8 threads per row)
    auto commit = [&](int buf, uint64_t tile, uint32_t parity) {
        mbar_wait(&s.full[buf], parity);
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {
            int r = lrow + 32 * pass;
            uint64_t row = row_begin + tile * kXD + r;
            const uint4 v = *reinterpret_cast<const uint4*>(&s.d[buf][r * kXStride + 4 * lsub]);
            uint32_t pc = group8_sum(popc4(v));
            if (lsub == 0)
                s.pd[buf][r] = row < row_end ? pc : kXInvalid;
        }
    };

    fetch(0, n_tiles - 1);
    commit(0, n_tiles - 1, 0);
    __syncthreads();

    int buf = 0;
    for (uint64_t it = 0; it < n_tiles; ++it) {
        const uint64_t tile = n_tiles - 1 - it;  // highest rows first
        if (it + 1 < n_tiles)
            fetch(buf ^ 1, tile - 1);

        // ---- 4 x 4 intersections over 32 words
        uint32_t acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b)
                acc[a][b] = 0;
#pragma unroll 2
        for (int w4 = 0; w4 < 8; ++w4) {
            uint4 qv[4], dv[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                qv[r] = *reinterpret_cast<const uint4*>(&s.q[(ty + 16 * r) * kXStride + 4 * w4]);
                dv[r] = *reinterpret_cast<const uint4*>(&s.d[buf][(tx + 16 * r) * kXStride + 4 * w4]);
            }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b)
                    acc[a][b] += popc4_and(qv[a], dv[b]);
        }

        // ---- integer filter against each query's current k-th best
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int ql = ty + 16 * a;
            if (s.pq[ql] == kXInvalid)
                continue;
            const uint4 t = s.thr[ql];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int dl = tx + 16 * b;
                const uint32_t pdv = s.pd[buf][dl];
                const uint32_t i = acc[a][b];
                if (pdv != kXInvalid && i * t.x + t.w > t.z + t.y * pdv) {
                    uint32_t at = atomicAdd(&s.queue_count, 1u);
                    s.queue[at] = ((uint32_t)ql << 24) | ((uint32_t)dl << 16) | i;
                }
            }
        }
        __syncthreads();

        // ---- slow path: warp w owns the queries with (ql & 7) == w
        const uint32_t qn = s.queue_count;
        if (qn) {
            for (uint32_t base = 0; base < qn; base += 32) {
                uint32_t e = base + lane < qn ? s.queue[base + lane] : 0xFFFFFFFFu;
                uint32_t mine = __ballot_sync(RB_FULL_MASK, base + lane < qn && (int)((e >> 24) & 7) == warp);
                while (mine) {
                    int j = __ffs((int)mine) - 1;
                    mine &= mine - 1;
                    uint32_t ev = __shfl_sync(RB_FULL_MASK, e, j);
                    uint32_t ql = ev >> 24, dl = (ev >> 16) & 0xFFu, i = ev & 0xFFFFu;
                    uint32_t u = s.pq[ql] + s.pd[buf][dl] - i;  // |q or d|
                    uint32_t slot = (uint32_t)(p.slot_base + row_begin + tile * kXD + dl);
                    uint64_t key = ((uint64_t)exact_distance_bits(i, u) << 32) | (uint32_t)~slot;
                    size_t off = ((size_t)chunk * p.query_count + q0 + ql) * p.k;
                    glist_insert(p.partial_keys + off, p.partial_iu + off, p.k, key, i | ((u ? u : 1u) << 16));
                    if (p.partial_keys[off + p.k - 1] != ~0ull) {  // full: publish the new k-th best
                        uint32_t iu = p.partial_iu[off + p.k - 1];
                        uint32_t it_ = iu & 0xFFFFu, ut = iu >> 16;
                        if (lane == 0)
                            s.thr[ql] = make_uint4(ut + it_, it_, it_ * s.pq[ql], 0);
                    }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
        if (tid == 0)
            s.queue_count = 0;
        if (it + 1 < n_tiles)
            commit(buf ^ 1, tile - 1, (uint32_t)(((it + 1) >> 1) & 1));  // the (it + 1) / 2-th use of that stage
        __syncthreads();
        buf ^= 1;
    }
}

// ---- k-way merge of `parts` sorted partial lists per query (index.hpp:2625-2645): one warp per query
struct MergeParams {
    const uint64_t* partial_keys;  // [parts x Q x k]
    uint32_t parts;
    uint32_t query_count;
    uint32_t k;
    uint64_t* keys_out;            // [Q x k]
};

__global__ void __launch_bounds__(128) merge_topk_kernel(MergeParams p) {
    extern __shared__ __align__(16) unsigned char raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t* list = reinterpret_cast<uint64_t*>(raw) + (size_t)warp * p.k;
    const uint32_t q = blockIdx.x * 4 + warp;
    if (q >= p.query_count)
        return;
    for (uint32_t i = lane; i < p.k; i += 32)
        list[i] = ~0ull;
    __syncwarp();
    uint32_t len = p.k;  // a full list of +inf entries
    const uint32_t per = p.k;
    const uint32_t total = p.parts * per;
    for (uint32_t base = 0; base < total; base += 32) {
        uint32_t idx = base + lane;
        uint64_t key = ~0ull;
        if (idx < total) {
            uint32_t part = idx / per, j = idx % per;
            key = p.partial_keys[((size_t)part * p.query_count + q) * p.k + j];
        }
        uint64_t worst = list[p.k - 1];
        uint32_t acc = __ballot_sync(RB_FULL_MASK, key < worst);
        while (acc) {
            int j = __ffs((int)acc) - 1;
            acc &= acc - 1;
            uint64_t kv = __shfl_sync(RB_FULL_MASK, key, j);
            // same sorted insert as the traversal list, on plain u64 keys
            if (kv < list[p.k - 1]) {
                uint32_t pos = 0;
                for (uint32_t b2 = 0; b2 < p.k; b2 += 32) {
                    uint32_t i2 = b2 + lane;
                    pos += __popc(__ballot_sync(RB_FULL_MASK, i2 < p.k && list[i2] < kv));
                }
                for (uint32_t hi = p.k - 1; hi > pos;) {
                    uint32_t lo = (hi - pos > 32) ? hi - 32 : pos;
                    uint32_t i2 = lo + lane;
                    bool act = i2 < hi;
                    uint64_t v = act ? list[i2] : 0ull;
                    __syncwarp();
                    if (act)
                        list[i2 + 1] = v;
                    __syncwarp();
                    hi = lo;
                }
                if (lane == 0)
                    list[pos] = kv;
                __syncwarp();
            }
        }
    }
    (void)len;
    for (uint32_t i = lane; i < p.k; i += 32)
        p.keys_out[(size_t)q * p.k + i] = list[i];
}

// ---- keys -> (slot, distance, count)
__global__ void unpack_topk_kernel(const uint64_t* keys, uint64_t query_count, uint32_t k, uint64_t* slots,
                                   float* dists, int64_t* counts) {
    uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= query_count)
        return;
    uint32_t cnt = 0;
    for (uint32_t j = 0; j < k; ++j) {
        uint64_t key = keys[q * k + j];
        if (key == ~0ull)
            break;
        slots[q * k + j] = (uint32_t)~(uint32_t)key;
        dists[q * k + j] = __uint_as_float((uint32_t)(key >> 32));
        ++cnt;
    }
    counts[q] = cnt;
}

}  // namespace rb
