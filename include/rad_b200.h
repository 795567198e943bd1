/*
 * rad_b200.h -- C ABI of the B200-native drop-in for RAD's data-parallel hot path:
 * HNSW build over packed (<= 1024-bit) fingerprints under Tanimoto and the batched Tanimoto k-NN search.
 *
 * This is the boundary the reference's FFI for this path binds (keiserlab/rad @ a21d60b, vendored usearch fork
 * 2.17.2).  The reference's own native boundary is the pybind11 module `usearch.compiled`
 * (usearch/python/lib.cpp); its C flavour is usearch/c/usearch.h, whose conventions are kept here:
 * an opaque handle, plain pointers and sizes, and a `rb_error_t* error` out-parameter that stays NULL on success
 * and otherwise points at a NUL-terminated message owned by the library (usearch/c/usearch.h:20-28,123).
 * No torch / numpy / C++ types cross this boundary.  Every entry point cites the reference interface it replaces.
 *
 * Threading: one handle may be used from one thread at a time.  All read-side calls (rb_neighbors,
 * rb_top_level_nodes, rb_keys_from_slots, rb_slots_from_keys, rb_level_stats, rb_size ...) touch HOST memory only
 * and make no CUDA call, so they keep working in a fork()ed child (rad/hnsw_service.py:129-134 forks after the
 * index exists).  rb_add / rb_search / rb_*_device need a CUDA device and fail loudly without one: there is no
 * CPU fallback.
 */
#ifndef RAD_B200_H
#define RAD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RB_EXPORT __attribute__((visibility("default")))

typedef struct rb_index rb_index_t;
typedef char const* rb_error_t;

/* progress(processed, total, user) -> nonzero to continue, 0 to cancel (python/lib.cpp:204-208) */
typedef int (*rb_progress_t)(uint64_t processed, uint64_t total, void* user);

enum {
    RB_FLAG_LAZY_HOST_SYNC = 1u, /* do not copy the graph back to the host after every rb_add; call rb_sync_host */
    RB_FLAG_TIME_KERNELS = 2u,   /* CUDA events around every build kernel; fills rb_counters_t::*_ms (bench / roofline) */
};

/* usearch/python/lib.cpp:116-147 (make_index) / usearch/c/usearch.h:94-123 (usearch_init_options_t) */
typedef struct rb_init_options {
    uint64_t dimensions;       /* bits per fingerprint (ndim); multiple of 8, 8..1024                         */
    uint32_t connectivity;     /* M; level-0 lists hold 2M (index.hpp:1375-1376); 0 -> 16, max 16             */
    uint32_t expansion_add;    /* ef of insertions; 0 -> 128; max 1024                                        */
    uint32_t expansion_search; /* ef of queries;    0 -> 64;  max 1024                                        */
    int32_t device;            /* CUDA ordinal; -1 = the calling thread's current device                      */
    uint32_t flags;            /* RB_FLAG_*                                                                   */
    uint32_t batch_growth_den; /* a batch holds at most size/den nodes; 0 -> 8                                */
    uint32_t batch_max;        /* upper bound of a batch; 0 -> 65536                                          */
    uint32_t bloom_bits_log2;  /* visited filter size per traversal (bits, log2); 0 -> 15                     */
    uint32_t debug_max_batches;/* stop an rb_add after this many batches (0 = unlimited); tests only          */
    uint32_t no_batch_tile;    /* 1: skip the batch tile (the nodes of a batch stay blind to each other); A/B only */
    uint32_t connectivity_base;/* level-0 row width; 0 -> 2 x connectivity (what the file format and the reference use).
                                * Another value is only meant for the upper-level sub-build of the cross-shard stitch
                                * (level l of the graph = level l - 1 of that build) and cannot be saved                */
} rb_init_options_t;

/* per-level statistics: usearch python/lib.cpp:1007-1013 (IndexStats), index.hpp:3087-3179 */
typedef struct rb_level_stats {
    uint64_t nodes, edges, max_edges, allocated_bytes;
} rb_level_stats_t;

/* Index.metadata(path): index_dense.hpp:253-319 (index_dense_metadata_from_path) */
typedef struct rb_metadata {
    uint16_t version_major, version_minor, version_patch;
    uint8_t kind_metric, kind_scalar, kind_key, kind_compressed_slot;
    uint64_t count_present, count_deleted, dimensions;
    uint8_t multi;
    uint8_t has_vectors;   /* the matrix blob precedes the head */
    uint64_t matrix_rows, matrix_cols;
} rb_metadata_t;

/* kernel-side work counters of the last rb_add / rb_search (index.hpp:2183-2191, 2523-2531, 2580-2584) */
typedef struct rb_counters {
    uint64_t computed_distances;   /* fingerprints fetched and compared in graph traversal                    */
    uint64_t visited_members;      /* node expansions (the reference's iteration_cycles)                      */
    uint64_t batches;              /* insertion batches launched                                              */
    uint64_t kernel_launches;      /* CUDA kernels of this library launched                                   */
    double device_ms;              /* CUDA-event time of the device work, on the library's stream             */
    /* with RB_FLAG_TIME_KERNELS: summed CUDA-event durations of each kernel class of the last rb_add */
    double insert_ms;              /* insert_search_kernel (greedy descent + ef-bounded searches of a batch)  */
    double sort_ms;                /* radix sort of the (level, target) reverse pairs                         */
    double reverse_ms;             /* reverse_link_kernel                                                     */
    uint64_t insert_launches;      /* insert_search_kernel launches (= batches that did graph work)           */
    double tile_ms;                /* the batch tile (exact nearest earlier batch-mates, tensor cores)        */
    double select_ms;              /* insert_select_kernel (merge batch-mates, heuristic, forward links)      */
} rb_counters_t;

/* ---------------------------------------------------------------- lifetime: python/lib.cpp:116-147, index.py:1154-1162 */
RB_EXPORT rb_index_t* rb_init(rb_init_options_t const* options, rb_error_t* error);
RB_EXPORT void rb_free(rb_index_t* index);
/* drop every device allocation (fingerprints stay on the host only if keep_host_vectors); read side keeps working */
RB_EXPORT void rb_release_device(rb_index_t* index, int keep_host_vectors, rb_error_t* error);

/* ---------------------------------------------------------------- build: python/lib.cpp:223-268 (add_many_to_index),
 * index_dense.hpp:2101-2150 (add_), index.hpp:2734-2834 (add), 897-935 (try_reserve) */
RB_EXPORT void rb_reserve(rb_index_t* index, size_t capacity, rb_error_t* error);
RB_EXPORT void rb_add(rb_index_t* index, uint64_t const* keys, uint8_t const* vectors, size_t count,
                      size_t row_stride_bytes, int copy, rb_progress_t progress, void* progress_user,
                      rb_error_t* error);
/* same insertion with the fingerprints ALREADY in HBM: `d_vectors` is a device pointer to count x 128 bytes,
 * 16-byte aligned, zero-padded past dimensions/8; `keys` is a host pointer or NULL (size .. size+count-1). */
RB_EXPORT void rb_add_device(rb_index_t* index, uint64_t const* keys, void const* d_vectors, size_t count,
                             rb_error_t* error);
/* rb_add_device with PRESCRIBED levels (host pointer, one per row) instead of the level stream: the upper-level sub-build
 * of the cross-shard stitch */
RB_EXPORT void rb_add_device_levels(rb_index_t* index, uint64_t const* keys, void const* d_vectors, size_t count,
                                    int16_t const* levels, rb_error_t* error);
/* a row shard of a larger build: this EMPTY index starts the reference's single level stream (index.hpp:3867-3871) at
 * row `first_row`, so its nodes get the levels a one-piece build would give those rows */
RB_EXPORT void rb_seek_level_stream(rb_index_t* index, uint64_t first_row, rb_error_t* error);
RB_EXPORT void rb_sync_host(rb_index_t* index, rb_error_t* error);
/* Index.clear(): forget every node but keep the reserved host and device memory (index.py:1150-1152 ->
 * index_dense.hpp:989-998 clear(): "erases all vectors from the index, preserving the allocated space") */
RB_EXPORT void rb_clear(rb_index_t* index, rb_error_t* error);
/* the cudaStream_t (as an integer) every kernel of this handle is launched on, so that callers can bracket calls
 * with their own CUDA events; 0 until the first device call */
RB_EXPORT uintptr_t rb_stream(rb_index_t const* index);

/* ---------------------------------------------------------------- search: python/lib.cpp:511-557 (search_many_in_index),
 * index.hpp:2970-3029 (search), 4222-4239 (search_exact_: ties by descending slot, 923-934).
 * keys_out/distances_out are [query_count x k]; rows beyond counts_out[q] are left untouched. */
RB_EXPORT void rb_search(rb_index_t* index, uint8_t const* queries, size_t query_count, size_t row_stride_bytes,
                         size_t k, int exact, uint64_t* keys_out, float* distances_out, int64_t* counts_out,
                         uint64_t* visited_members, uint64_t* computed_distances, rb_error_t* error);
/* device-resident twin: queries and outputs are device pointers (queries: query_count x 128 bytes padded;
 * slots_out: u32 [Q x k]; distances_out: f32 [Q x k]; counts_out: i32 [Q]); runs on the library's stream and
 * returns after it has been synchronised. */
RB_EXPORT void rb_search_device(rb_index_t* index, void const* d_queries, size_t query_count, size_t k, int exact,
                                void* d_slots_out, void* d_distances_out, void* d_counts_out, rb_error_t* error);

/* `cluster(query, level)` of usearch/include/usearch/index.hpp:3046-3079 for a batch: the greedy descent from the entry
 * point down to and including `level` (level 0 stops at level 1, like the reference) -> the node where it ends
 * (slots_out u32 [Q]) and its distance (distances_out f32 [Q]); device pointers, queries padded to 128 bytes.  The
 * building block of Index.cluster (usearch/include/usearch/index_dense.hpp:1919-2080; rad_b200/cluster.py). */
RB_EXPORT void rb_cluster_device(rb_index_t* index, void const* d_queries, size_t query_count, int level,
                                 void* d_slots_out, void* d_distances_out, rb_error_t* error);

/* ---------------------------------------------------------------- stateless device kernels (multi-GPU plumbing, bench)
 * Exact Tanimoto top-k of Q queries against a shard of `rows` fingerprints resident in HBM.
 * d_keys_out: u64 [Q x k], key = f32 distance bits << 32 | ~(slot_base + row) so that ascending key order IS the
 * reference order (ascending distance, ties by descending slot); unused tail entries are ~0.
 * `stream` is a cudaStream_t passed as an integer (0 = legacy default stream). */
RB_EXPORT void rb_exact_topk_device(void const* d_database, uint64_t rows, uint64_t slot_base, void const* d_queries,
                                    uint64_t query_count, uint32_t k, void* d_keys_out, void* d_workspace,
                                    size_t workspace_bytes, uintptr_t stream, rb_error_t* error);
RB_EXPORT size_t rb_exact_topk_workspace_bytes(uint64_t rows, uint64_t query_count, uint32_t k);
/* k-way merge of `parts` partial results [parts x Q x k] (e.g. the all-gathered per-shard keys; the merge rule of
 * index.hpp:2625-2645) into d_keys_out [Q x k]. */
/* rb_exact_topk_device for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs, same
 * d_keys_out, bit-identical results; the workspace is a file-scope static device array of RB_NOALLOC_SCRATCH_BYTES
 * (16 MiB, held by every GPU that loads the library until the process exits), so the call makes no cudaMalloc /
 * cudaFree and uses no allocating library routine.  A search whose workspace (rb_exact_topk_workspace_bytes) exceeds
 * the array is REJECTED with an error naming both sizes.  All calls on one GPU share the array: calls that overlap on
 * the GPU (different streams) race -- serialise them.  The array keeps the last call's contents; every call
 * re-initialises what it reads.  The device of `stream` must be current. */
#define RB_NOALLOC_SCRATCH_BYTES (16u << 20)
RB_EXPORT size_t rb_exact_topk_noalloc_max_bytes(void);
RB_EXPORT void rb_exact_topk_device_noalloc(void const* d_database, uint64_t rows, uint64_t slot_base,
                                            void const* d_queries, uint64_t query_count, uint32_t k, void* d_keys_out,
                                            uintptr_t stream, rb_error_t* error);
RB_EXPORT void rb_merge_topk_device(void const* d_partial_keys, uint32_t parts, uint64_t query_count, uint32_t k,
                                    void* d_keys_out, uintptr_t stream, rb_error_t* error);
/* Multi-GPU search behind the C ABI (the reference's `Indexes`, usearch/python/lib.cpp:417-498, searches its shards side
 * by side and merges by index.hpp:2625-2645).  Every rank calls this with its shard `index` (rows in global slots
 * slot_base ..), the SAME queries (device, 128 bytes each) and its NCCL communicator (an `ncclComm_t` passed as an
 * integer): local search (exact or ef-bounded) -> keys in global slots -> ncclAllGather on the handle's stream ->
 * k-way merge.  d_keys_out: u64 [Q x k] on every rank, the format of rb_exact_topk_device (rb_unpack_topk_device
 * splits it).  NCCL is loaded with dlopen at the first call (the copy already in the process, else libnccl.so.2):
 * the library does not link against it.  An empty shard takes part with no candidates. */
RB_EXPORT size_t rb_sharded_search_workspace_bytes(size_t query_count, size_t k, uint32_t world);
RB_EXPORT void rb_sharded_search_device(rb_index_t* index, uint64_t slot_base, void const* d_queries,
                                        size_t query_count, size_t k, int exact, uintptr_t nccl_comm, void* d_keys_out,
                                        void* d_workspace, size_t workspace_bytes, rb_error_t* error);
/* split merged keys into slots (u64), distances (f32) and counts (i64) -- all device pointers */
RB_EXPORT void rb_unpack_topk_device(void const* d_keys, uint64_t query_count, uint32_t k, void* d_slots_out,
                                     void* d_distances_out, void* d_counts_out, uintptr_t stream, rb_error_t* error);
/* counter-based synthetic fingerprints (rad_b200/synth.py), rows [first_row, first_row + rows) -> d_out */
RB_EXPORT void rb_synth_fingerprints_device(void* d_out, uint64_t first_row, uint64_t rows, uint64_t seed,
                                            void const* d_centres, uint32_t centre_count, int edge_cases,
                                            uintptr_t stream, rb_error_t* error);
/* the same generator with its optional modes: rows_per_centre > 0 = analogue-series order (the rows of a cluster are
 * contiguous: row r belongs to centre (r / rows_per_centre) mod centre_count); dup_fold > 1 = runs of dup_fold exact
 * duplicates inside [dup_start, dup_start + dup_rows) */
RB_EXPORT void rb_synth_fingerprints_device_ex(void* d_out, uint64_t first_row, uint64_t rows, uint64_t seed,
                                               void const* d_centres, uint32_t centre_count, int edge_cases,
                                               uint64_t rows_per_centre, uint64_t dup_fold, uint64_t dup_start,
                                               uint64_t dup_rows, uintptr_t stream, rb_error_t* error);

/* ---------------------------------------------------------------- fingerprint ingest.
 * The reference has no pre-pass: identical fingerprints are inserted like any other row (index.hpp:2734-2834) and form
 * d = 0 cliques that fill their members' M links and inflate every ef-bounded search that meets them; its clustering /
 * join helpers (python/usearch/index.py:1182-1273) start from the same rows.
 * rb_dedup_device: d_representative[i] (u32) = the LOWEST row whose 128 bytes equal row i's (== i for a fingerprint seen
 * for the first time).  d_vectors: rows x 128 bytes on the device; the workspace is rb_dedup_workspace_bytes(rows)
 * (about 28 bytes per row).  Runs on `stream`, no synchronisation; the answer never depends on the internal hash.
 * rb_popcounts_device: d_popcounts[i] (u16) = number of set bits of row i (the |a| of union = |a| + |q| - intersection). */
RB_EXPORT size_t rb_dedup_workspace_bytes(uint64_t rows);
RB_EXPORT void rb_dedup_device(void const* d_vectors, uint64_t rows, void* d_representative, void* d_workspace,
                               size_t workspace_bytes, uintptr_t stream, rb_error_t* error);
RB_EXPORT void rb_popcounts_device(void const* d_vectors, uint64_t rows, void* d_popcounts, uintptr_t stream,
                                   rb_error_t* error);
/* Index.join: one-to-one ("stable marriage") matching of two fingerprint sets -- usearch/include/usearch/index.hpp:
 * 4330-4515 (the algorithm), usearch/python/lib.cpp:876-894 (the binding), usearch/python/usearch/index.py:1182-1212.
 * The candidates of every "man" (row of the SMALLER set) among the "women" come from rb_search_device on the women's
 * index (k = proposals; exact or ef-bounded): d_candidate_slots u32 [men x proposals] nearest first,
 * d_candidate_distances f32 [men x proposals], d_candidate_counts i32 [men].  The matching is the one the reference's
 * loop produces on ONE thread (ring of free men processed front to back; a woman leaves her husband only for a
 * strictly nearer man), computed generation by generation on the device.  d_man_to_woman u32 [men] and
 * d_woman_to_man u32 [women] receive slots, 0xFFFFFFFF = unmatched.  Synchronises `stream` once per generation. */
RB_EXPORT size_t rb_stable_match_workspace_bytes(uint64_t men, uint64_t women);
RB_EXPORT void rb_stable_match_device(void const* d_candidate_slots, void const* d_candidate_distances,
                                      void const* d_candidate_counts, uint64_t men, uint64_t women, uint32_t proposals,
                                      void* d_man_to_woman, void* d_woman_to_man, void* d_workspace,
                                      size_t workspace_bytes, uintptr_t stream, uint64_t* generations_out,
                                      uint64_t* engagements_out, rb_error_t* error);

/* ---------------------------------------------------------------- one graph from G row shards (SURVEY.md section 8e row 3;
 * usearch has no counterpart: its multi-index `Indexes`, python/lib.cpp:417-498, only searches shards side by side).
 * Every shard is built on its own GPU; then every node searches the OTHER shards (rb_search_device on their handles),
 * the answers travel with their fingerprints (rb_gather_vectors_device; NCCL all-gather / all-to-all between the
 * processes -- rad_b200/sharded.py), and rb_stitch_device prunes own links + remote candidates with the reference's
 * heuristic (index.hpp:4246-4289) into level-0 rows in GLOBAL slots.  rb_import_graph assembles the global index.
 * All pointers named d_* are device pointers; the calls run on the handle's stream and return when done. */
/* d_out[i] = fingerprint (128 bytes) of local slot d_slots[i] (u32; 0xFFFFFFFF -> zeros) */
RB_EXPORT void rb_gather_vectors_device(rb_index_t* index, void const* d_slots, uint64_t count, void* d_out,
                                        rb_error_t* error);
/* d_out = fingerprints of local slots [first, first + count), 128 bytes each */
RB_EXPORT void rb_copy_vectors_device(rb_index_t* index, uint64_t first, uint64_t count, void* d_out, rb_error_t* error);
/* level-0 rows of local nodes [first, first + count) after the stitch -> d_rows_out (u32 [count x 2M], GLOBAL slots,
 * 0xFFFFFFFF padded).  d_cand_slots: u32 [count x cands] GLOBAL slots of the remote candidates (0xFFFFFFFF = none),
 * d_cand_vectors: their fingerprints [count x cands x 128 bytes]; cands <= 128 - 2M.  slot_base = global slot of this
 * shard's local slot 0.  The shard's own graph is not modified. */
RB_EXPORT void rb_stitch_device(rb_index_t* index, uint64_t first, uint64_t count, uint64_t slot_base,
                                void const* d_cand_slots, void const* d_cand_vectors, uint32_t cands, void* d_rows_out,
                                rb_error_t* error);
/* destination (EMPTY, created with connectivity = connectivity_base = the source's connectivity) becomes the sub-index
 * of `source` over its nodes of level >= 1, one level down: graph level l + 1 of the source is level l of the
 * destination, slots renumbered in ascending source order, fingerprints copied -- all on the device.  The upper levels
 * of the stitch start from the links the build already made (index.hpp:2734-2834 builds every level in the one add). */
RB_EXPORT void rb_extract_level(rb_index_t* source, rb_index_t* destination, rb_error_t* error);
/* replace the graph of `index` by one assembled elsewhere: host arrays keys u64[n], levels i16[n], adj0 u32[n x 2M],
 * adj_upper u32[sum(levels) x M] (a node's rows consecutive, level 1 first), vectors u8[n x dimensions/8] or NULL,
 * entry = a node of the top level.  Afterwards the index saves / serves / searches like one built in one piece. */
RB_EXPORT void rb_import_graph(rb_index_t* index, uint64_t const* keys, int16_t const* levels, uint32_t const* adj0,
                               uint32_t const* adj_upper, uint8_t const* vectors, size_t n, uint64_t entry,
                               rb_error_t* error);

/* ---------------------------------------------------------------- serialisation: python/lib.cpp:940-954,
 * index_dense.hpp:1025-1093 (save), 1114-1259 (load), 1267-1415 (view), index.hpp:3236-3277, 3463-3532 */
RB_EXPORT void rb_save(rb_index_t* index, char const* path, int exclude_vectors, rb_error_t* error);
RB_EXPORT void rb_load(rb_index_t* index, char const* path, int exclude_vectors, rb_error_t* error);
RB_EXPORT void rb_view(rb_index_t* index, char const* path, int exclude_vectors, rb_error_t* error);
RB_EXPORT void rb_metadata(char const* path, rb_metadata_t* out, rb_error_t* error);

/* ---------------------------------------------------------------- graph reads: python/lib.cpp:270-305,
 * index_dense.hpp:942-981, index.hpp:2690-2707.  Output is [slot, key, slot, key, ...]; the return value is the
 * number of u64 written (2 x neighbours).  Out-of-range slot/level sets *error (-> ValueError). */
RB_EXPORT size_t rb_neighbors(rb_index_t const* index, size_t slot, int level, uint64_t* out_pairs, size_t capacity,
                              rb_error_t* error);
/* Batched form (additive): neighbours of `count` (slot, level) pairs in one call, for a service
 * that prefetches the lists of many work items per round trip (rad/coordination_service.py:308-328 asks one node at
 * a time through a queue).  Lists are written back to back as [slot, key, ...]; offsets[i] .. offsets[i + 1] is the
 * u64 range of request i (offsets has count + 1 entries).  levels == NULL means level 0 for every slot.  An
 * out-of-range request yields an EMPTY list (offsets[i + 1] == offsets[i]) and is counted in the return value, the
 * number of failed requests; *error is set only when `capacity` is too small (2 x 2 x connectivity x count suffices). */
RB_EXPORT size_t rb_neighbors_many(rb_index_t const* index, uint64_t const* slots, int32_t const* levels, size_t count,
                                   uint64_t* out_pairs, size_t capacity, uint64_t* offsets, rb_error_t* error);
/* with out_pairs == NULL returns the required capacity (2 x nodes on the top level) */
RB_EXPORT size_t rb_top_level_nodes(rb_index_t const* index, uint64_t* out_pairs, size_t capacity, rb_error_t* error);
RB_EXPORT void rb_keys_from_slots(rb_index_t const* index, uint64_t const* slots, size_t count, uint64_t* keys_out,
                                  rb_error_t* error);
/* slots_out[i] = -1 when the key is absent (python/lib.cpp:307-330) */
RB_EXPORT void rb_slots_from_keys(rb_index_t* index, uint64_t const* keys, size_t count, int64_t* slots_out,
                                  rb_error_t* error);
RB_EXPORT int rb_contains(rb_index_t* index, uint64_t key);
/* fingerprints of `count` slots -> out [count x dimensions/8] (Index.get; test_sparse.py:12-33) */
RB_EXPORT void rb_get_vectors(rb_index_t* index, uint64_t const* slots, size_t count, uint8_t* out, rb_error_t* error);

/* ---------------------------------------------------------------- metadata: python/lib.cpp:1331-1342, 1465-1475,
 * index.hpp:2274 (max_level), 3087-3201 (stats, memory_usage) */
RB_EXPORT size_t rb_size(rb_index_t const* index);
RB_EXPORT size_t rb_capacity(rb_index_t const* index);
RB_EXPORT int32_t rb_max_level(rb_index_t const* index);
RB_EXPORT uint64_t rb_entry_slot(rb_index_t const* index);
RB_EXPORT size_t rb_connectivity(rb_index_t const* index);
RB_EXPORT size_t rb_dimensions(rb_index_t const* index);
RB_EXPORT size_t rb_memory_usage(rb_index_t const* index);
RB_EXPORT size_t rb_serialized_length(rb_index_t const* index, int exclude_vectors);
RB_EXPORT size_t rb_expansion_add(rb_index_t const* index);
RB_EXPORT size_t rb_expansion_search(rb_index_t const* index);
RB_EXPORT void rb_set_expansion_add(rb_index_t* index, size_t expansion);
RB_EXPORT void rb_set_expansion_search(rb_index_t* index, size_t expansion);
RB_EXPORT void rb_level_stats(rb_index_t const* index, size_t level, rb_level_stats_t* out, rb_error_t* error);
RB_EXPORT void rb_levels(rb_index_t const* index, int16_t* out, size_t capacity);
RB_EXPORT void rb_last_counters(rb_index_t const* index, rb_counters_t* out);
/* raw struct-of-arrays adjacency for tests: level-0 rows [size x 2M] and upper rows [upper_rows x M], 0xFFFFFFFF
 * = empty.  Returns the number of upper rows; either pointer may be NULL. */
RB_EXPORT uint64_t rb_debug_adjacency(rb_index_t const* index, uint32_t* adj0_out, uint32_t* adj_upper_out);
RB_EXPORT char const* rb_version(void);
RB_EXPORT int rb_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* RAD_B200_H */
