// host_index.hpp -- host-side container of the index: keys, levels, adjacency (struct-of-arrays or an mmap'ed
// .usearch node tape), key -> slot lookups, level statistics and the .usearch v2.17 reader / writer.
//
// Everything here is plain host memory: the read-side entry points RAD calls at run time (get_neighbors,
// get_top_level_nodes, len, max_level, levels_stats, ... -- rad/hnsw_service.py:219-232, 400-412) work in a
// fork()ed child without a CUDA context.
//
// Reference behaviour mirrored (usearch fork 2.17.2 under /root/reference/usearch):
//   node tape        include/usearch/index.hpp:2111-2176, 3705-3730  (u64 key | i16 level | [u32 n | u32 x M0] |
//                    level x [u32 n | u32 x M], packed, zero-filled tail)
//   file layout      include/usearch/index_dense.hpp:42-79, 1025-1093 and index.hpp:1855-1861, 3236-3277
//   load / view      index_dense.hpp:1114-1259, 1267-1415; index.hpp:3463-3532 (matrix blob auto-detected, optional
//                    exclude_vectors, key -> slot map NOT rebuilt eagerly after load)
//   level draw       index.hpp:3867-3871 over minstd_rand0 (index.hpp:2187), one stream (= a threads=1 build)
//   stats            index.hpp:3087-3179, memory_usage index_dense.hpp:737-743
#pragma once
#include <cstddef>
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

namespace rbh {

constexpr uint32_t kEmpty = 0xFFFFFFFFu;

struct LevelStats {
    uint64_t nodes = 0, edges = 0, max_edges = 0, allocated_bytes = 0;
};

struct FileMetadata {
    uint16_t version_major = 0, version_minor = 0, version_patch = 0;
    uint8_t kind_metric = 0, kind_scalar = 0, kind_key = 0, kind_compressed_slot = 0;
    uint64_t count_present = 0, count_deleted = 0, dimensions = 0;
    uint8_t multi = 0, has_vectors = 0;
    uint64_t matrix_rows = 0, matrix_cols = 0;
    uint64_t head_offset = 0;  // file offset of the 64-byte head
};

class KeyMap {  // open-addressing u64 -> u32, never erases
  public:
    void clear();
    void reserve(size_t n);
    bool insert(uint64_t key, uint32_t slot);  // false when the key is already present
    int64_t find(uint64_t key) const;          // -1 when absent
    size_t size() const { return count_; }
    size_t memory() const { return keys_.capacity() * 8 + vals_.capacity() * 4; }

  private:
    void grow(size_t cap);
    std::vector<uint64_t> keys_;
    std::vector<uint32_t> vals_;  // kEmpty = free cell
    size_t count_ = 0, mask_ = 0;
};

class HostIndex {
  public:
    // configuration
    uint64_t ndim = 1024;
    uint32_t vec_bytes = 128;
    uint32_t m = 16, m0 = 32;
    uint32_t ef_add = 128, ef_search = 64;

    // graph, struct-of-arrays (authoritative unless `viewing`)
    std::vector<uint64_t> keys;
    std::vector<int16_t> levels;
    std::vector<uint32_t> adj0;       // size x m0
    std::vector<uint32_t> adju;       // upper_rows x m
    std::vector<uint64_t> upper_off;  // size + 1 prefix sums of levels
    int32_t max_level = 0;
    uint64_t entry = 0;
    uint32_t level_rng = 1;           // minstd_rand0 state of the single level stream
    size_t capacity = 0;

    // fingerprints that currently live on the host (after load, or kept on request); size x vec_bytes or empty
    std::vector<uint8_t> host_vectors;

    // mmap view (immutable)
    bool viewing = false;
    bool vectors_excluded = false;

    ~HostIndex();
    size_t size() const { return viewing ? view_size_ : keys.size(); }
    bool immutable() const { return viewing; }

    // draws the levels of the next `n` nodes from the single reference stream
    void draw_levels(size_t n, std::vector<int16_t>& out);
    // an EMPTY index whose first node will be row `first_row` of the single stream: a row shard of a larger build
    // draws exactly the levels a one-piece build gives those rows
    void seek_level_stream(uint64_t first_row);
    // replace the graph by one assembled elsewhere (the cross-shard stitch): arrays in this class's layout
    std::string import_graph(uint64_t const* new_keys, int16_t const* new_levels, uint32_t const* new_adj0,
                             uint32_t const* new_adju, uint8_t const* vectors, size_t n, uint64_t new_entry);

    uint64_t key_of(size_t slot) const;
    int level_of(size_t slot) const;
    // neighbours of (slot, level) -> out (slots); returns count, or -1 / -2 for a bad slot / level
    int neighbors(size_t slot, int level, uint32_t* out, size_t cap) const;
    size_t top_level_nodes(std::vector<uint32_t>& out) const;
    bool level_stats(size_t level, LevelStats& out) const;
    size_t memory_usage() const;
    size_t serialized_length(bool exclude_vectors) const;

    // key lookups (lazy map; see note on load/view in the header comment)
    int64_t slot_of_key(uint64_t key);
    // returns the offending key's position or -1 when every new key is unique and absent
    int64_t check_new_keys(uint64_t const* new_keys, size_t n);
    void note_new_keys(uint64_t const* new_keys, size_t n);
    // undo an append that was committed before the device confirmed it: keep the first `rows` nodes
    void truncate_rows(size_t rows, int32_t old_max_level, uint64_t old_entry);
    // per-level population (kept from the levels alone, see below): call after `levels` grew by `count` rows
    void note_new_levels(size_t first_slot, int16_t const* new_levels, size_t count);
    // edges per level counted where the adjacency was produced (the device, at rb_sync_host): no host scan needed
    void set_level_edges(std::vector<uint64_t> const& edges);
    void invalidate_level_edges();

    // serialisation; `vectors` may be null only when exclude_vectors
    std::string save(char const* path, uint8_t const* vectors, bool exclude_vectors) const;
    std::string load(char const* path, bool exclude_vectors);
    std::string view(char const* path, bool exclude_vectors);
    static std::string metadata(char const* path, FileMetadata& out);
    uint8_t const* viewed_vectors() const { return view_vectors_; }

    void reset_graph();

  private:
    void close_view();
    void rebuild_keymap();
    // view state
    void* map_ = nullptr;
    size_t map_len_ = 0;
    size_t view_size_ = 0;
    uint8_t const* view_tape_ = nullptr;
    int16_t const* view_levels_ = nullptr;
    uint8_t const* view_vectors_ = nullptr;
    std::vector<uint64_t> view_off_;  // size + 1 node offsets into the tape

    // lookups
    KeyMap map_keys_;
    bool map_valid_ = false;
    bool keys_ascending_ = true;  // keys were added in strictly ascending order so far
    uint64_t max_key_ = 0;

    // Per-level metadata.  The reference rescans every node for get_top_level_nodes (python/lib.cpp:290-305) and every
    // node AND list for stats (index.hpp:3147-3179) on each call.  Here the node counts per level and the list of
    // top-level nodes are maintained from the LEVELS alone, as they arrive (add / load / view) -- they never touch the
    // adjacency, so they are valid even while the host copy of a freshly built graph is stale.  Edge counts come with
    // the adjacency: from the device at rb_sync_host, or from one parallel pass over a loaded / viewed tape on first
    // use.  A mutex guards the lazily built parts (RAD's threaded hnsw_server calls these concurrently).
    std::vector<uint64_t> level_nodes_;  // [l] = nodes with level >= l
    std::vector<uint32_t> top_nodes_;    // slots of the nodes at max level, ascending
    mutable std::vector<uint64_t> level_edges_;
    mutable bool edges_valid_ = false;
    mutable std::mutex mu_;
    void rebuild_level_index();
    void ensure_edges() const;
};

}  // namespace rbh
