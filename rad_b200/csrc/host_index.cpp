// host_index.cpp -- see host_index.hpp.  Host-only: no CUDA in this translation unit.
#include "host_index.hpp"

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace rbh {

// ------------------------------------------------------------------------------------------------ key map
static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xFF51AFD7ED558CCDull;
    x ^= x >> 33;
    x *= 0xC4CEB9FE1A85EC53ull;
    x ^= x >> 33;
    return x;
}

void KeyMap::clear() {
    keys_.clear();
    vals_.clear();
    count_ = 0;
    mask_ = 0;
}

void KeyMap::grow(size_t cap) {
    std::vector<uint64_t> old_keys;
    std::vector<uint32_t> old_vals;
    old_keys.swap(keys_);
    old_vals.swap(vals_);
    keys_.assign(cap, 0);
    vals_.assign(cap, kEmpty);
    mask_ = cap - 1;
    count_ = 0;
    for (size_t i = 0; i < old_vals.size(); ++i)
        if (old_vals[i] != kEmpty)
            insert(old_keys[i], old_vals[i]);
}

void KeyMap::reserve(size_t n) {
    size_t cap = 16;
    while (cap < n * 2)
        cap <<= 1;
    if (cap > vals_.size())
        grow(cap);
}

bool KeyMap::insert(uint64_t key, uint32_t slot) {
    if ((count_ + 1) * 2 > vals_.size())
        grow(vals_.empty() ? 16 : vals_.size() * 2);
    size_t i = mix64(key) & mask_;
    while (vals_[i] != kEmpty) {
        if (keys_[i] == key)
            return false;
        i = (i + 1) & mask_;
    }
    keys_[i] = key;
    vals_[i] = slot;
    ++count_;
    return true;
}

int64_t KeyMap::find(uint64_t key) const {
    if (vals_.empty())
        return -1;
    size_t i = mix64(key) & mask_;
    while (vals_[i] != kEmpty) {
        if (keys_[i] == key)
            return vals_[i];
        i = (i + 1) & mask_;
    }
    return -1;
}

// ------------------------------------------------------------------------------------------------ level stream
// minstd_rand0 + libstdc++ generate_canonical<double, 53> (two draws, low digit first), index.hpp:3867-3871
static inline uint32_t minstd_next(uint32_t& state) {
    // x * 16807 mod (2^31 - 1) without a division: fold the bits above 31 back in (2^31 == 1 mod p), twice at most
    uint64_t v = (uint64_t)state * 16807ull;
    v = (v & 0x7FFFFFFFull) + (v >> 31);
    v = (v & 0x7FFFFFFFull) + (v >> 31);
    state = (uint32_t)(v == 0x7FFFFFFFull ? 0 : v);
    return state;
}

// state after `steps` more draws: x * 16807^steps mod (2^31 - 1)
static inline uint32_t minstd_skip(uint32_t state, uint64_t steps) {
    uint64_t base = 16807ull, acc = state;
    const uint64_t p = 2147483647ull;
    for (; steps; steps >>= 1) {
        if (steps & 1)
            acc = acc * base % p;
        base = base * base % p;
    }
    return (uint32_t)acc;
}

static void draw_level_range(uint32_t state, uint32_t m, int16_t* out, size_t n) {
    const double inverse_log_connectivity = 1.0 / std::log((double)m);
    const double range = 2147483646.0;
    // 1 - 1/m of all draws give level 0: r safely above 1/m needs no logarithm (the margin keeps every r whose
    // -log(r) / log(m) could round to 1.0 on the exact path)
    const double level0_above = (1.0 / (double)m) * (1.0 + 1e-9);
    for (size_t i = 0; i < n; ++i) {
        double low = (double)(minstd_next(state) - 1u);
        double high = (double)(minstd_next(state) - 1u);
        double r = (low + high * range) / (range * range);
        if (r > level0_above && r < 1.0) {
            out[i] = 0;
            continue;
        }
        if (r >= 1.0)
            r = std::nextafter(1.0, 0.0);
        double l = -std::log(r) * inverse_log_connectivity;
        if (!(l < 30.0))
            l = 30.0;  // r == 0 would be +inf; never happens with minstd, keeps the cast defined
        out[i] = (int16_t)l;
    }
}

void HostIndex::draw_levels(size_t n, std::vector<int16_t>& out) {
    out.resize(n);
    // every node consumes exactly two draws, so the stream can be cut anywhere: large requests are split over
    // threads, each starting from the jumped-ahead state (10 M levels: 110 ms on one core)
    size_t threads = n >= (1u << 20) ? std::min<size_t>(16, std::max(1u, std::thread::hardware_concurrency())) : 1;
    if (threads <= 1) {
        draw_level_range(level_rng, m, out.data(), n);
    } else {
        std::vector<std::thread> pool;
        const size_t per = (n + threads - 1) / threads;
        for (size_t t = 0; t < threads; ++t) {
            const size_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
            if (lo == hi)
                continue;
            const uint32_t start = minstd_skip(level_rng, 2 * (uint64_t)lo);
            pool.emplace_back(draw_level_range, start, m, out.data() + lo, hi - lo);
        }
        for (std::thread& th : pool)
            th.join();
    }
    level_rng = minstd_skip(level_rng, 2 * (uint64_t)n);
}

void HostIndex::seek_level_stream(uint64_t first_row) { level_rng = minstd_skip(1u, 2 * first_row); }

std::string HostIndex::import_graph(uint64_t const* new_keys, int16_t const* new_levels, uint32_t const* new_adj0,
                                    uint32_t const* new_adju, uint8_t const* vectors, size_t n, uint64_t new_entry) {
    if (n && new_entry >= n)
        return "import: the entry slot is out of range";
    int top = 0;
    uint64_t upper = 0;
    for (size_t i = 0; i < n; ++i) {
        if (new_levels[i] < 0)
            return "import: negative level";
        upper += (uint64_t)new_levels[i];
        top = std::max<int>(top, new_levels[i]);
    }
    if (n && new_levels[new_entry] != top)
        return "import: the entry slot is not a node of the top level";
    for (size_t i = 0; i < n * (size_t)m0; ++i)
        if (new_adj0[i] != kEmpty && new_adj0[i] >= n)
            return "import: a level-0 link points outside the graph";
    for (size_t i = 0; i < upper * (size_t)m; ++i)
        if (new_adju[i] != kEmpty && new_adju[i] >= n)
            return "import: an upper-level link points outside the graph";
    // everything checked: only now is the old graph dropped
    const uint32_t keep_m = m, keep_m0 = m0, keep_ef_add = ef_add, keep_ef_search = ef_search, keep_vb = vec_bytes;
    const uint64_t keep_ndim = ndim;
    reset_graph();
    m = keep_m, m0 = keep_m0, ef_add = keep_ef_add, ef_search = keep_ef_search, vec_bytes = keep_vb, ndim = keep_ndim;
    keys.assign(new_keys, new_keys + n);
    levels.assign(new_levels, new_levels + n);
    upper_off.assign(n + 1, 0);
    for (size_t i = 0; i < n; ++i)
        upper_off[i + 1] = upper_off[i] + (uint64_t)levels[i];
    adj0.assign(new_adj0, new_adj0 + n * (size_t)m0);
    adju.assign(new_adju, new_adju + upper * (size_t)m);
    if (vectors && n)
        host_vectors.assign(vectors, vectors + n * (size_t)vec_bytes);
    vectors_excluded = vectors == nullptr;
    max_level = top;
    entry = new_entry;
    capacity = n;
    std::vector<int16_t> scratch;
    level_rng = 1;
    draw_levels(n, scratch);  // further insertions continue the single stream
    keys_ascending_ = true;
    max_key_ = 0;
    for (size_t i = 0; i < n; ++i) {
        if (i && keys[i] <= max_key_)
            keys_ascending_ = false;
        max_key_ = std::max(max_key_, keys[i]);
    }
    rebuild_level_index();
    return {};
}

// ------------------------------------------------------------------------------------------------ accessors
static inline size_t node_bytes(uint32_t m, uint32_t m0, int level) {
    return 10 + 4 * (1 + (size_t)m0) + (size_t)level * 4 * (1 + (size_t)m);
}

HostIndex::~HostIndex() { close_view(); }

void HostIndex::close_view() {
    if (map_) {
        munmap(map_, map_len_);
        map_ = nullptr;
        map_len_ = 0;
    }
    view_tape_ = nullptr;
    view_levels_ = nullptr;
    view_vectors_ = nullptr;
    view_off_.clear();
    view_off_.shrink_to_fit();
    view_size_ = 0;
    viewing = false;
}

void HostIndex::reset_graph() {
    close_view();
    keys.clear();
    levels.clear();
    adj0.clear();
    adju.clear();
    upper_off.assign(1, 0);
    host_vectors.clear();
    max_level = 0;
    entry = 0;
    level_rng = 1;
    map_keys_.clear();
    map_valid_ = false;
    keys_ascending_ = true;
    max_key_ = 0;
    level_nodes_.clear();
    top_nodes_.clear();
    edges_valid_ = false;
    vectors_excluded = false;
}

uint64_t HostIndex::key_of(size_t slot) const {
    if (viewing) {
        uint64_t k;
        std::memcpy(&k, view_tape_ + view_off_[slot], 8);
        return k;
    }
    return keys[slot];
}

int HostIndex::level_of(size_t slot) const {
    if (viewing) {
        int16_t l;
        std::memcpy(&l, reinterpret_cast<uint8_t const*>(view_levels_) + slot * 2, 2);
        return l;
    }
    return levels[slot];
}

int HostIndex::neighbors(size_t slot, int level, uint32_t* out, size_t cap) const {
    if (slot >= size())
        return -1;
    if (level < 0 || level > level_of(slot))
        return -2;
    if (viewing) {
        uint8_t const* p = view_tape_ + view_off_[slot] + 10;
        if (level)
            p += 4 * (1 + (size_t)m0) + (size_t)(level - 1) * 4 * (1 + (size_t)m);
        uint32_t cnt;
        std::memcpy(&cnt, p, 4);
        uint32_t width = level ? m : m0;
        if (cnt > width)
            cnt = width;
        size_t n = std::min<size_t>(cnt, cap);
        std::memcpy(out, p + 4, n * 4);
        return (int)cnt;
    }
    uint32_t const* row;
    uint32_t width;
    if (level == 0) {
        row = adj0.data() + slot * m0;
        width = m0;
    } else {
        row = adju.data() + (upper_off[slot] + (uint64_t)(level - 1)) * m;
        width = m;
    }
    uint32_t cnt = 0;
    while (cnt < width && row[cnt] != kEmpty) {
        if (cnt < cap)
            out[cnt] = row[cnt];
        ++cnt;
    }
    return (int)cnt;
}

void HostIndex::note_new_levels(size_t first_slot, int16_t const* new_levels, size_t count) {
    for (size_t i = 0; i < count; ++i) {
        const int l = new_levels[i];
        if ((size_t)l + 1 > level_nodes_.size())
            level_nodes_.resize((size_t)l + 1, 0);
        for (int lv = 0; lv <= l; ++lv)
            ++level_nodes_[lv];
        // the top level of the graph is the highest level seen so far (index.hpp:2761-2781: a node above it becomes
        // the entry point); nodes AT that level are what get_top_level_nodes returns
        const int top = (int)level_nodes_.size() - 1;
        if (l == top) {
            if (level_nodes_[top] == 1)
                top_nodes_.clear();
            top_nodes_.push_back((uint32_t)(first_slot + i));
        }
    }
}

void HostIndex::rebuild_level_index() {
    level_nodes_.clear();
    top_nodes_.clear();
    const size_t n = size();
    // one sequential pass over the level array (2 bytes per node), in blocks so that the histogram stays in registers
    uint64_t hist[64] = {0};
    int top = 0;
    for (size_t i = 0; i < n; ++i) {
        const int l = level_of(i);
        ++hist[std::min(l, 63)];
        top = std::max(top, l);
    }
    if (n) {
        level_nodes_.assign((size_t)top + 1, 0);
        uint64_t acc = 0;
        for (int l = std::min(top, 63); l >= 0; --l) {
            acc += hist[l];
            level_nodes_[l] = acc;
        }
        for (size_t i = 0; i < n; ++i)
            if (level_of(i) == top)
                top_nodes_.push_back((uint32_t)i);
    }
    std::lock_guard<std::mutex> lock(mu_);
    edges_valid_ = false;
}

void HostIndex::set_level_edges(std::vector<uint64_t> const& edges) {
    std::lock_guard<std::mutex> lock(mu_);
    level_edges_ = edges;
    level_edges_.resize(std::max(level_edges_.size(), level_nodes_.size()), 0);
    edges_valid_ = true;
}

void HostIndex::invalidate_level_edges() {
    std::lock_guard<std::mutex> lock(mu_);
    edges_valid_ = false;
}

// edges per level of a loaded / viewed graph: one pass over the lists, split over host threads, once
void HostIndex::ensure_edges() const {
    std::lock_guard<std::mutex> lock(mu_);
    if (edges_valid_)
        return;
    const size_t n = size(), nl = level_nodes_.size();
    const size_t threads = std::min<size_t>({16, std::max(1u, std::thread::hardware_concurrency()), n / 65536 + 1});
    std::vector<std::vector<uint64_t>> part(threads, std::vector<uint64_t>(nl, 0));
    auto work = [&](size_t t) {
        uint32_t tmp[64];
        const size_t lo = n * t / threads, hi = n * (t + 1) / threads;
        for (size_t i = lo; i < hi; ++i) {
            const int l = level_of(i);
            for (int lv = 0; lv <= l && (size_t)lv < nl; ++lv)
                part[t][lv] += (uint64_t)std::max(neighbors(i, lv, tmp, 64), 0);
        }
    };
    if (threads <= 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (size_t t = 0; t < threads; ++t)
            pool.emplace_back(work, t);
        for (std::thread& th : pool)
            th.join();
    }
    level_edges_.assign(nl, 0);
    for (size_t t = 0; t < threads; ++t)
        for (size_t lv = 0; lv < nl; ++lv)
            level_edges_[lv] += part[t][lv];
    edges_valid_ = true;
}

bool HostIndex::level_stats(size_t level, LevelStats& out) const {
    out = LevelStats{};
    if (level >= level_nodes_.size())
        return level <= (size_t)max_level || level == 0;
    ensure_edges();
    out.nodes = level_nodes_[level];
    out.edges = level_edges_[level];
    out.max_edges = out.nodes * (level ? m : m0);
    out.allocated_bytes = out.nodes * (level ? 4 * (1 + (size_t)m) : 10 + 4 * (1 + (size_t)m0));
    return true;
}

size_t HostIndex::top_level_nodes(std::vector<uint32_t>& out) const {
    out = top_nodes_;
    return out.size();
}

size_t HostIndex::memory_usage() const {
    size_t total = 0;
    if (viewing)
        total += view_off_.capacity() * 8;
    else
        total += keys.capacity() * 8 + levels.capacity() * 2 + adj0.capacity() * 4 + adju.capacity() * 4 +
                 upper_off.capacity() * 8 + host_vectors.capacity();
    total += map_keys_.memory();
    return total;
}

size_t HostIndex::serialized_length(bool exclude_vectors) const {
    size_t n = size();
    size_t total = 64 + 40 + n * 2;
    if (!exclude_vectors)
        total += 8 + n * (size_t)vec_bytes;
    for (size_t i = 0; i < n; ++i)
        total += node_bytes(m, m0, level_of(i));
    return total;
}

// ------------------------------------------------------------------------------------------------ key lookups
void HostIndex::rebuild_keymap() {
    map_keys_.clear();
    size_t n = size();
    map_keys_.reserve(n);
    for (size_t i = 0; i < n; ++i)
        map_keys_.insert(key_of(i), (uint32_t)i);
    map_valid_ = true;
}

int64_t HostIndex::slot_of_key(uint64_t key) {
    std::lock_guard<std::mutex> lock(mu_);  // the map is built on first use; concurrent readers must not race on it
    if (!map_valid_)
        rebuild_keymap();
    return map_keys_.find(key);
}

int64_t HostIndex::check_new_keys(uint64_t const* new_keys, size_t n) {
    // fast path: strictly ascending keys above everything seen so far cannot collide
    bool ascending = keys_ascending_;
    if (ascending && n) {
        if (size() && new_keys[0] <= max_key_)
            ascending = false;
        for (size_t i = 1; ascending && i < n; ++i)
            ascending = new_keys[i] > new_keys[i - 1];
    }
    if (ascending)
        return -1;
    std::lock_guard<std::mutex> lock(mu_);  // may run beside readers (and on the add's helper thread)
    if (!map_valid_)
        rebuild_keymap();
    KeyMap fresh;
    fresh.reserve(n);
    for (size_t i = 0; i < n; ++i) {
        if (map_keys_.find(new_keys[i]) >= 0)
            return (int64_t)i;
        if (!fresh.insert(new_keys[i], (uint32_t)i))
            return (int64_t)i;
    }
    return -1;
}

void HostIndex::note_new_keys(uint64_t const* new_keys, size_t n) {
    // called after `keys` has been extended by the new keys
    size_t first = keys.size() - n;
    for (size_t i = 0; i < n; ++i) {
        if (keys_ascending_ && (first + i) && new_keys[i] <= max_key_)
            keys_ascending_ = false;
        max_key_ = std::max(max_key_, new_keys[i]);
        if (map_valid_)
            map_keys_.insert(new_keys[i], (uint32_t)(first + i));
    }
}

void HostIndex::truncate_rows(size_t rows, int32_t old_max_level, uint64_t old_entry) {
    keys.resize(rows);
    levels.resize(rows);
    upper_off.resize(rows + 1);
    max_level = old_max_level;
    entry = old_entry;
    map_keys_.clear();
    map_valid_ = false;
    keys_ascending_ = true;
    max_key_ = 0;
    for (size_t i = 0; i < rows; ++i) {
        if (i && keys[i] <= max_key_)
            keys_ascending_ = false;
        max_key_ = std::max(max_key_, keys[i]);
    }
    rebuild_level_index();
}

// ------------------------------------------------------------------------------------------------ file format
namespace {

struct Head {
    uint8_t raw[64];
};

void fill_head(Head& h, uint64_t present, uint64_t dimensions) {
    std::memset(h.raw, 0, 64);
    std::memcpy(h.raw, "usearch", 7);
    uint16_t ver[3] = {2, 17, 2};
    std::memcpy(h.raw + 7, ver, 6);
    h.raw[13] = 't';  // metric_kind_t::tanimoto_k
    h.raw[14] = 1;    // scalar_kind_t::b1x8_k
    h.raw[15] = 14;   // scalar_kind_t::u64_k  (keys)
    h.raw[16] = 15;   // scalar_kind_t::u32_k  (slots)
    uint64_t deleted = 0;
    std::memcpy(h.raw + 17, &present, 8);
    std::memcpy(h.raw + 25, &deleted, 8);
    std::memcpy(h.raw + 33, &dimensions, 8);
    h.raw[41] = 0;  // multi
}

std::string parse_head(uint8_t const* raw, FileMetadata& md) {
    if (std::memcmp(raw, "usearch", 7) != 0)
        return "Magic header mismatch - the file isn't an index";
    std::memcpy(&md.version_major, raw + 7, 2);
    std::memcpy(&md.version_minor, raw + 9, 2);
    std::memcpy(&md.version_patch, raw + 11, 2);
    md.kind_metric = raw[13];
    md.kind_scalar = raw[14];
    md.kind_key = raw[15];
    md.kind_compressed_slot = raw[16];
    std::memcpy(&md.count_present, raw + 17, 8);
    std::memcpy(&md.count_deleted, raw + 25, 8);
    std::memcpy(&md.dimensions, raw + 33, 8);
    md.multi = raw[41];
    return {};
}

// locate the head: the file either starts with it, or with a u32[2] / u64[2] matrix header (index_dense.hpp:253-319)
std::string locate_head(uint8_t const* data, size_t size, FileMetadata& md) {
    if (size < 64)
        return "End of file reached!";
    if (std::memcmp(data, "usearch", 7) == 0) {
        md.has_vectors = 0;
        md.head_offset = 0;
        return parse_head(data, md);
    }
    uint32_t d32[2];
    std::memcpy(d32, data, 8);
    uint64_t off32 = (uint64_t)d32[0] * d32[1] + 8;
    if (off32 + 64 <= size && std::memcmp(data + off32, "usearch", 7) == 0) {
        md.has_vectors = 1;
        md.matrix_rows = d32[0];
        md.matrix_cols = d32[1];
        md.head_offset = off32;
        return parse_head(data + off32, md);
    }
    uint64_t d64[2];
    std::memcpy(d64, data, 16);
    uint64_t off64 = d64[0] * d64[1] + 16;
    if (d64[1] && d64[0] <= size / d64[1] && off64 + 64 <= size && std::memcmp(data + off64, "usearch", 7) == 0) {
        md.has_vectors = 1;
        md.matrix_rows = d64[0];
        md.matrix_cols = d64[1];
        md.head_offset = off64;
        return parse_head(data + off64, md);
    }
    return "Not a dense USearch index!";
}

struct Mapping {
    void* ptr = nullptr;
    size_t len = 0;
    std::string open(char const* path) {
        int fd = ::open(path, O_RDONLY);
        if (fd < 0)
            return std::string(std::strerror(errno));
        struct stat st;
        if (fstat(fd, &st) != 0) {
            ::close(fd);
            return std::string(std::strerror(errno));
        }
        len = (size_t)st.st_size;
        if (!len) {
            ::close(fd);
            return "End of file reached!";
        }
        ptr = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
        ::close(fd);
        if (ptr == MAP_FAILED) {
            ptr = nullptr;
            return std::string(std::strerror(errno));
        }
        return {};
    }
    void release() {
        if (ptr)
            munmap(ptr, len);
        ptr = nullptr;
    }
};

std::string validate(FileMetadata const& md) {
    if (md.version_major != 2)
        return "File format may be different, please rebuild";
    if (md.kind_key != 14)
        return "Key type doesn't match, consider rebuilding";
    if (md.kind_compressed_slot != 15)
        return "Slot type doesn't match, consider rebuilding";
    if (md.kind_metric != 't' || md.kind_scalar != 1)
        return "rad_b200 only serves b1 fingerprints under the Tanimoto metric";
    if (md.dimensions == 0 || md.dimensions % 8 != 0 || md.dimensions > 1024)
        return "rad_b200 supports 8..1024-bit fingerprints (multiples of 8)";
    return {};
}

}  // namespace

std::string HostIndex::metadata(char const* path, FileMetadata& out) {
    Mapping mp;
    std::string err = mp.open(path);
    if (!err.empty())
        return err;
    err = locate_head(static_cast<uint8_t const*>(mp.ptr), mp.len, out);
    mp.release();
    return err;
}

std::string HostIndex::save(char const* path, uint8_t const* vectors, bool exclude_vectors) const {
    std::FILE* f = std::fopen(path, "wb");
    if (!f)
        return std::string(std::strerror(errno));
    auto put = [&](void const* p, size_t bytes) { return bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes; };
    bool ok = true;
    const size_t n = size();
    if (!exclude_vectors) {
        if (!vectors && !viewed_vectors() && n) {
            std::fclose(f);
            return "The fingerprints are not available for saving";
        }
        uint8_t const* src = vectors ? vectors : viewed_vectors();
        uint32_t dims[2] = {(uint32_t)n, vec_bytes};
        ok = ok && put(dims, 8) && put(src, n * (size_t)vec_bytes);
    }
    Head head;
    fill_head(head, n, ndim);
    ok = ok && put(head.raw, 64);
    uint64_t core[5] = {n, m, m0, (uint64_t)max_level, entry};
    ok = ok && put(core, 40);
    if (viewing) {
        ok = ok && put(view_levels_, n * 2) && put(view_tape_ + (n ? view_off_[0] : 0), n ? view_off_[n] - view_off_[0] : 0);
    } else {
        ok = ok && put(levels.data(), n * 2);
        // pack the node tapes chunk by chunk, in parallel, and stream them out in large writes (§8f item 2)
        const size_t chunk_nodes = 1u << 20;
        std::vector<uint64_t> off;
        std::vector<uint8_t> tape;
        const unsigned workers = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
        for (size_t lo = 0; ok && lo < n; lo += chunk_nodes) {
            size_t hi = std::min(n, lo + chunk_nodes);
            off.assign(hi - lo + 1, 0);
            for (size_t i = lo; i < hi; ++i)
                off[i - lo + 1] = off[i - lo] + node_bytes(m, m0, levels[i]);
            tape.assign(off[hi - lo], 0);
            auto pack = [&](size_t a, size_t b) {
                for (size_t i = a; i < b; ++i) {
                    uint8_t* p = tape.data() + off[i - lo];
                    std::memcpy(p, &keys[i], 8);
                    std::memcpy(p + 8, &levels[i], 2);
                    p += 10;
                    for (int lv = 0; lv <= levels[i]; ++lv) {
                        uint32_t width = lv ? m : m0;
                        uint32_t const* row = lv ? adju.data() + (upper_off[i] + (uint64_t)(lv - 1)) * m
                                                 : adj0.data() + i * (size_t)m0;
                        uint32_t cnt = 0;
                        while (cnt < width && row[cnt] != kEmpty)
                            ++cnt;
                        std::memcpy(p, &cnt, 4);
                        std::memcpy(p + 4, row, (size_t)cnt * 4);
                        p += 4 * (1 + (size_t)width);
                    }
                }
            };
            if (hi - lo < 65536 || workers == 1) {
                pack(lo, hi);
            } else {
                std::vector<std::thread> pool;
                size_t per = (hi - lo + workers - 1) / workers;
                for (unsigned w = 0; w < workers; ++w) {
                    size_t a = lo + w * per, b = std::min(hi, a + per);
                    if (a < b)
                        pool.emplace_back(pack, a, b);
                }
                for (auto& t : pool)
                    t.join();
            }
            ok = ok && put(tape.data(), tape.size());
        }
    }
    if (std::fclose(f) != 0)
        ok = false;
    return ok ? std::string() : std::string("Failed to serialize into stream");
}

std::string HostIndex::view(char const* path, bool exclude_vectors) {
    Mapping mp;
    std::string err = mp.open(path);
    if (!err.empty())
        return err;
    FileMetadata md;
    uint8_t const* data = static_cast<uint8_t const*>(mp.ptr);
    err = locate_head(data, mp.len, md);
    if (err.empty())
        err = validate(md);
    if (!err.empty()) {
        mp.release();
        return err;
    }
    size_t at = md.head_offset + 64;
    if (mp.len < at + 40) {
        mp.release();
        return "File is corrupted and lacks a header";
    }
    uint64_t core[5];
    std::memcpy(core, data + at, 40);
    at += 40;
    const uint64_t n = core[0];
    if (md.has_vectors && md.matrix_rows != n) {
        mp.release();
        return "Index size and the number of vectors doesn't match";
    }
    if (core[1] < 2 || core[1] > 16 || core[2] != 2 * core[1]) {
        mp.release();
        return "rad_b200 supports connectivity 2..16 with connectivity_base = 2 x connectivity";
    }
    if (mp.len < at + n * 2) {
        mp.release();
        return "File is corrupted and can't fit all the nodes";
    }
    reset_graph();
    ndim = md.dimensions;
    vec_bytes = (uint32_t)(md.dimensions / 8);
    m = (uint32_t)core[1];
    m0 = (uint32_t)core[2];
    view_levels_ = reinterpret_cast<int16_t const*>(data + at);
    view_off_.resize(n + 1);
    uint64_t off = at + n * 2;
    for (uint64_t i = 0; i < n; ++i) {
        int16_t l;
        std::memcpy(&l, data + at + i * 2, 2);
        view_off_[i] = off;
        off += node_bytes(m, m0, l);
    }
    view_off_[n] = off;
    if (mp.len < off) {
        view_off_.clear();
        mp.release();
        return "File is corrupted and can't fit all the nodes";
    }
    map_ = mp.ptr;
    map_len_ = mp.len;
    view_tape_ = data;  // offsets are absolute file offsets
    view_size_ = n;
    viewing = true;
    max_level = (int32_t)core[3];
    entry = core[4];
    capacity = n;
    vectors_excluded = exclude_vectors || !md.has_vectors;
    view_vectors_ = vectors_excluded ? nullptr : data + (md.head_offset - md.matrix_rows * md.matrix_cols);
    map_valid_ = false;
    rebuild_level_index();
    return {};
}

std::string HostIndex::load(char const* path, bool exclude_vectors) {
    // parse through a temporary view, then unpack the tape into the struct-of-arrays form used by the GPU build
    HostIndex tmp;
    std::string err = tmp.view(path, exclude_vectors);
    if (!err.empty())
        return err;
    const size_t n = tmp.size();
    reset_graph();
    ndim = tmp.ndim;
    vec_bytes = tmp.vec_bytes;
    m = tmp.m;
    m0 = tmp.m0;
    max_level = tmp.max_level;
    entry = tmp.entry;
    vectors_excluded = tmp.vectors_excluded;
    keys.resize(n);
    levels.resize(n);
    upper_off.assign(n + 1, 0);
    for (size_t i = 0; i < n; ++i) {
        keys[i] = tmp.key_of(i);
        levels[i] = (int16_t)tmp.level_of(i);
        upper_off[i + 1] = upper_off[i] + (uint64_t)levels[i];
    }
    adj0.assign(n * (size_t)m0, kEmpty);
    adju.assign(upper_off[n] * (size_t)m, kEmpty);
    for (size_t i = 0; i < n; ++i) {
        tmp.neighbors(i, 0, adj0.data() + i * (size_t)m0, m0);
        for (int lv = 1; lv <= levels[i]; ++lv)
            tmp.neighbors(i, lv, adju.data() + (upper_off[i] + (uint64_t)(lv - 1)) * m, m);
    }
    if (!vectors_excluded && n) {
        host_vectors.assign(tmp.viewed_vectors(), tmp.viewed_vectors() + n * (size_t)vec_bytes);
    }
    capacity = n;
    // replay the level stream so that further insertions continue the reference's single-thread sequence
    std::vector<int16_t> scratch;
    level_rng = 1;
    draw_levels(n, scratch);
    keys_ascending_ = true;
    max_key_ = 0;
    for (size_t i = 0; i < n; ++i) {
        if (i && keys[i] <= max_key_)
            keys_ascending_ = false;
        max_key_ = std::max(max_key_, keys[i]);
    }
    rebuild_level_index();
    return {};
}

}  // namespace rbh
