// api.cu -- the C ABI of include/rad_b200.h: device state, batched build orchestration, search, stateless kernels.
//
// Host orchestration replaces usearch/python/lib.cpp:177-268 (add_typed_to_index / add_many_to_index: the
// executor-driven per-vector loop) and lib.cpp:357-415, 511-557 (search_typed / search_many_in_index).  Instead of a
// thread pool with per-node locks, an rb_add is ONE stream of kernel launches whose shape (levels, batch boundaries,
// entry point and max level per batch) is known on the host before the first launch, so there is no device->host
// synchronisation inside a build.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <dlfcn.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cuda_runtime.h>

#include "../../include/rad_b200.h"
#include "exact_kernels.cuh"
#include "exact_i8_kernels.cuh"
#include "exact_f4_kernels.cuh"
#include "hnsw_kernels.cuh"
#include "host_index.hpp"
#include "ingest_kernels.cuh"
#include "synth_kernels.cuh"

namespace {

thread_local std::string g_error;

char const* set_error(rb_error_t* error, std::string msg) {
    g_error = std::move(msg);
    if (error)
        *error = g_error.c_str();
    return g_error.c_str();
}

std::string format(char const* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return buf;
}

#define RB_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return format("CUDA error %s at %s:%d (%s)", cudaGetErrorName(e_), __FILE__, __LINE__,      \
                          cudaGetErrorString(e_));                                                      \
    } while (0)

// RB_HOST_TIMING=1 prints the host-side phases of an add to stderr (tuning aid)
struct PhaseTimer {
    bool on = std::getenv("RB_HOST_TIMING") != nullptr;
    std::chrono::steady_clock::time_point last = std::chrono::steady_clock::now();
    void mark(char const* what) {
        if (!on)
            return;
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[rb host] %-28s %8.2f ms\n", what,
                     std::chrono::duration<double, std::milli>(now - last).count());
        last = now;
    }
};

size_t ceil2(size_t v) {
    size_t r = 1;
    while (r < v)
        r <<= 1;
    return r;
}

struct DeviceState {
    int device = -1;
    cudaStream_t stream = nullptr;
    int sm_count = 0;
    // graph + fingerprints
    uint4* fp = nullptr;
    uint32_t* adj0 = nullptr;
    uint32_t* adju = nullptr;
    uint32_t* upper_off = nullptr;
    uint16_t* pc0 = nullptr;  // link popcounts mirroring adj0 / adju (device_common.cuh)
    uint16_t* pcu = nullptr;
    int8_t* levels = nullptr;
    uint32_t* task_rel = nullptr;
    uint32_t* task_order = nullptr;  // per slot: the t-th unit of work of a batch is node batch_start + task_order[batch_start + t]
    size_t cap_rows = 0, cap_upper = 0;
    size_t rows = 0, upper_rows = 0;  // what the device currently holds
    // per-batch scratch
    uint64_t* pair_keys[2] = {nullptr, nullptr};
    uint64_t* pair_vals[2] = {nullptr, nullptr};
    size_t pair_cap = 0;
    // per-(level, position) task arrays of one batch (hnsw_kernels.cuh: InsertParams)
    uint64_t* task_lists = nullptr;
    uint32_t* task_lens = nullptr;
    unsigned long long* task_seeds = nullptr;
    uint64_t* mate_keys = nullptr;
    uint32_t* mate_iu = nullptr;
    size_t task_cap = 0, task_cap_ef = 0;
    uint4* batch_fp = nullptr;
    size_t batch_fp_cap = 0;
    void* sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    unsigned int* work_counters = nullptr;
    size_t work_counters_cap = 0;
    unsigned long long* counters = nullptr;  // [0] distances [1] expansions
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    cudaStream_t copy_stream = nullptr;  // metadata of the second segment of a large add (build_new_rows)
    cudaEvent_t ev_copy = nullptr;
    int insert_blocks_per_sm = 0, search_blocks_per_sm = 0, select_blocks_per_sm = 0;
    int insert_variant = 4, search_variant = 4;
    uint32_t insert_cfg_ef = 0, search_cfg_ef = 0;
    void* scratch = nullptr;  // grow-only arena for query / result staging
    size_t scratch_bytes = 0;

    ~DeviceState() { release(); }
    void release() {
        if (device < 0)
            return;
        cudaSetDevice(device);
        cudaFree(fp);
        cudaFree(adj0);
        cudaFree(adju);
        cudaFree(upper_off);
        cudaFree(pc0);
        cudaFree(pcu);
        cudaFree(levels);
        cudaFree(task_rel);
        cudaFree(task_order);
        for (int i = 0; i < 2; ++i) {
            cudaFree(pair_keys[i]);
            cudaFree(pair_vals[i]);
        }
        cudaFree(task_lists);
        cudaFree(task_lens);
        cudaFree(task_seeds);
        cudaFree(mate_keys);
        cudaFree(mate_iu);
        cudaFree(batch_fp);
        cudaFree(sort_tmp);
        cudaFree(work_counters);
        cudaFree(counters);
        cudaFree(scratch);
        if (ev_begin)
            cudaEventDestroy(ev_begin);
        if (ev_end)
            cudaEventDestroy(ev_end);
        if (ev_copy)
            cudaEventDestroy(ev_copy);
        if (copy_stream)
            cudaStreamDestroy(copy_stream);
        if (stream)
            cudaStreamDestroy(stream);
        *this = DeviceState{};
    }
    DeviceState() = default;
    DeviceState(DeviceState const&) = default;
    DeviceState& operator=(DeviceState const&) = default;
};

}  // namespace

struct rb_index {
    rbh::HostIndex host;
    DeviceState dev;
    rb_init_options_t opts{};
    bool device_graph_valid = false;  // the device holds the current graph and fingerprints
    bool host_graph_valid = true;     // the host arrays hold the current graph
    rb_counters_t last{};
};

namespace {

int exact_kernel_override();  // RB_EXACT_KERNEL (defined with plan_exact)

std::string ensure_device(rb_index* ix) {
    DeviceState& d = ix->dev;
    if (d.device >= 0) {
        RB_CUDA(cudaSetDevice(d.device));
        return {};
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return "rad_b200 needs a CUDA device for add/search and found none (there is no CPU fallback)";
    int dev = ix->opts.device;
    if (dev < 0)
        RB_CUDA(cudaGetDevice(&dev));
    if (dev >= count)
        return format("CUDA device %d requested but only %d present", dev, count);
    RB_CUDA(cudaSetDevice(dev));
    cudaDeviceProp prop;
    RB_CUDA(cudaGetDeviceProperties(&prop, dev));
    d.sm_count = prop.multiProcessorCount;
    RB_CUDA(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    RB_CUDA(cudaEventCreate(&d.ev_begin));
    RB_CUDA(cudaEventCreate(&d.ev_end));
    RB_CUDA(cudaMalloc(&d.counters, 4 * sizeof(unsigned long long)));
    d.device = dev;
    return {};
}

template <typename T>
std::string grow_array(T*& ptr, size_t old_count, size_t new_count, int fill_byte, cudaStream_t stream) {
    T* fresh = nullptr;
    RB_CUDA(cudaMalloc(&fresh, std::max<size_t>(new_count, 1) * sizeof(T)));
    if (old_count)
        RB_CUDA(cudaMemcpyAsync(fresh, ptr, old_count * sizeof(T), cudaMemcpyDeviceToDevice, stream));
    if (new_count > old_count)
        RB_CUDA(cudaMemsetAsync(fresh + old_count, fill_byte, (new_count - old_count) * sizeof(T), stream));
    RB_CUDA(cudaStreamSynchronize(stream));
    cudaFree(ptr);
    ptr = fresh;
    return {};
}

std::string ensure_capacity(rb_index* ix, size_t rows, size_t upper) {
    DeviceState& d = ix->dev;
    const uint32_t m = ix->host.m, m0 = ix->host.m0;
    if (rows > d.cap_rows) {
        size_t cap = std::max(ceil2(rows), (size_t)1024);
        std::string err;
        if (!(err = grow_array(d.fp, d.cap_rows * 8, cap * 8, 0, d.stream)).empty()) return err;
        if (!(err = grow_array(d.adj0, d.cap_rows * m0, cap * m0, 0xFF, d.stream)).empty()) return err;
        if (!(err = grow_array(d.upper_off, d.cap_rows, cap, 0, d.stream)).empty()) return err;
        if (!(err = grow_array(d.pc0, d.cap_rows * m0, cap * m0, 0, d.stream)).empty()) return err;
        if (!(err = grow_array(d.levels, d.cap_rows, cap, 0, d.stream)).empty()) return err;
        if (!(err = grow_array(d.task_rel, d.cap_rows, cap, 0, d.stream)).empty()) return err;
        if (!(err = grow_array(d.task_order, d.cap_rows, cap, 0, d.stream)).empty()) return err;
        d.cap_rows = cap;
    }
    if (upper > d.cap_upper) {
        size_t cap = std::max(ceil2(upper), (size_t)1024);
        std::string err = grow_array(d.adju, d.cap_upper * m, cap * m, 0xFF, d.stream);
        if (!err.empty()) return err;
        if (!(err = grow_array(d.pcu, d.cap_upper * m, cap * m, 0, d.stream)).empty()) return err;
        d.cap_upper = cap;
    }
    ix->host.capacity = std::max(ix->host.capacity, d.cap_rows);
    return {};
}

// copy host rows [first, first + count) of a (count x vec_bytes, stride) matrix into 128-byte device rows
std::string upload_rows(rb_index* ix, uint8_t const* src, size_t count, size_t stride, size_t first) {
    DeviceState& d = ix->dev;
    if (!count)
        return {};
    uint8_t* dst = reinterpret_cast<uint8_t*>(d.fp) + first * 128;
    const size_t vb = ix->host.vec_bytes;
    if (vb == 128 && stride == 128) {
        RB_CUDA(cudaMemcpyAsync(dst, src, count * 128, cudaMemcpyHostToDevice, d.stream));
    } else {
        RB_CUDA(cudaMemsetAsync(dst, 0, count * 128, d.stream));
        RB_CUDA(cudaMemcpy2DAsync(dst, 128, src, stride, vb, count, cudaMemcpyHostToDevice, d.stream));
    }
    return {};
}

// after rb_load: push the host graph and fingerprints to the device
// the graph of `n` nodes given as host arrays -> device (fingerprints, link rows, offsets, levels, link popcounts)
std::string upload_arrays(rb_index* ix, size_t n, uint8_t const* vectors, uint32_t const* adj0, uint32_t const* adju,
                          uint64_t const* upper_off, int16_t const* levels) {
    rbh::HostIndex& h = ix->host;
    DeviceState& d = ix->dev;
    std::string err = ensure_capacity(ix, n, upper_off[n]);
    if (!err.empty())
        return err;
    if (n) {
        if (!(err = upload_rows(ix, vectors, n, h.vec_bytes, 0)).empty())
            return err;
        RB_CUDA(cudaMemcpyAsync(d.adj0, adj0, n * h.m0 * 4, cudaMemcpyHostToDevice, d.stream));
        if (upper_off[n])
            RB_CUDA(cudaMemcpyAsync(d.adju, adju, upper_off[n] * h.m * 4, cudaMemcpyHostToDevice, d.stream));
        std::vector<uint32_t> off32(n);
        std::vector<int8_t> lv8(n);
        for (size_t i = 0; i < n; ++i) {
            off32[i] = (uint32_t)upper_off[i];
            lv8[i] = (int8_t)levels[i];
        }
        RB_CUDA(cudaMemcpyAsync(d.upper_off, off32.data(), n * 4, cudaMemcpyHostToDevice, d.stream));
        RB_CUDA(cudaMemcpyAsync(d.levels, lv8.data(), n, cudaMemcpyHostToDevice, d.stream));
        // link popcounts of the loaded graph
        uint16_t* node_pc = nullptr;
        RB_CUDA(cudaMalloc(&node_pc, n * sizeof(uint16_t)));
        const unsigned grid = (unsigned)std::min<size_t>((n * 8 + 255) / 256, (size_t)d.sm_count * 16);
        rb::node_popcount_kernel<<<grid, 256, 0, d.stream>>>(d.fp, (uint32_t)n, node_pc);
        rb::link_popcount_kernel<<<(unsigned)d.sm_count * 8, 256, 0, d.stream>>>(d.adj0, n * (size_t)h.m0, node_pc, d.pc0);
        if (upper_off[n])
            rb::link_popcount_kernel<<<(unsigned)d.sm_count * 8, 256, 0, d.stream>>>(
                d.adju, upper_off[n] * (size_t)h.m, node_pc, d.pcu);
        cudaError_t le = cudaGetLastError();
        cudaError_t se = cudaStreamSynchronize(d.stream);
        cudaFree(node_pc);
        RB_CUDA(le);
        RB_CUDA(se);
    }
    d.rows = n;
    d.upper_rows = upper_off[n];
    ix->device_graph_valid = true;
    return {};
}

std::string upload_graph(rb_index* ix) {
    rbh::HostIndex& h = ix->host;
    const size_t n = h.keys.size();
    if (ix->device_graph_valid)
        return {};
    if (n && h.host_vectors.empty())
        return "The fingerprints of this index are not available (loaded with exclude_vectors)";
    std::string err = upload_arrays(ix, n, h.host_vectors.data(), h.adj0.data(), h.adju.data(), h.upper_off.data(),
                                    h.levels.data());
    if (!err.empty())
        return err;
    h.host_vectors.clear();
    h.host_vectors.shrink_to_fit();
    return {};
}

// A VIEWED index (mmap of a .usearch file, index_dense.hpp:1267-1415) is searched like the reference searches it: the
// graph and the fingerprints are read from the mapping -- here once, into HBM; the host side keeps serving from the file.
std::string upload_view(rb_index* ix) {
    rbh::HostIndex& h = ix->host;
    if (ix->device_graph_valid)
        return {};
    const size_t n = h.size();
    if (n && !h.viewed_vectors())
        return "The fingerprints of this index are not available (viewed with exclude_vectors)";
    std::vector<uint64_t> upper_off(n + 1, 0);
    std::vector<int16_t> levels(n);
    for (size_t i = 0; i < n; ++i) {
        levels[i] = (int16_t)h.level_of(i);
        upper_off[i + 1] = upper_off[i] + (uint64_t)levels[i];
    }
    std::vector<uint32_t> adj0(n * (size_t)h.m0, RB_EMPTY), adju(upper_off[n] * (size_t)h.m, RB_EMPTY);
    const unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < nt; ++t)
        pool.emplace_back([&, t] {
            for (size_t i = n * t / nt; i < n * (t + 1) / nt; ++i) {
                h.neighbors(i, 0, adj0.data() + i * (size_t)h.m0, h.m0);
                for (int lv = 1; lv <= levels[i]; ++lv)
                    h.neighbors(i, lv, adju.data() + (upper_off[i] + (uint64_t)(lv - 1)) * h.m, h.m);
            }
        });
    for (auto& th : pool)
        th.join();
    return upload_arrays(ix, n, h.viewed_vectors(), adj0.data(), adju.data(), upper_off.data(), levels.data());
}

rb::GraphView graph_view(rb_index* ix) {
    rb::GraphView g;
    g.fp = ix->dev.fp;
    g.adj0 = ix->dev.adj0;
    g.adju = ix->dev.adju;
    g.upper_off = ix->dev.upper_off;
    g.pc0 = ix->dev.pc0;
    g.pcu = ix->dev.pcu;
    g.m = ix->host.m;
    g.m0 = ix->host.m0;
    return g;
}

std::string sync_host(rb_index* ix) {
    if (ix->host_graph_valid)
        return {};
    rbh::HostIndex& h = ix->host;
    DeviceState& d = ix->dev;
    const size_t n = h.keys.size();
    RB_CUDA(cudaSetDevice(d.device));
    h.adj0.resize(n * (size_t)h.m0);
    h.adju.resize(h.upper_off[n] * (size_t)h.m);
    if (n)
        RB_CUDA(cudaMemcpyAsync(h.adj0.data(), d.adj0, h.adj0.size() * 4, cudaMemcpyDeviceToHost, d.stream));
    if (!h.adju.empty())
        RB_CUDA(cudaMemcpyAsync(h.adju.data(), d.adju, h.adju.size() * 4, cudaMemcpyDeviceToHost, d.stream));
    // the edge counts per level travel with the adjacency (HostIndex: per-level metadata)
    unsigned long long* d_edges = nullptr;
    unsigned long long edges[rb::kMaxLevels] = {0};
    if (n) {
        RB_CUDA(cudaMalloc(&d_edges, sizeof(edges)));
        RB_CUDA(cudaMemsetAsync(d_edges, 0, sizeof(edges), d.stream));
        const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, (size_t)d.sm_count * 8);
        rb::edge_count_kernel<<<grid, 256, 0, d.stream>>>(graph_view(ix), d.levels, (uint32_t)n, d_edges);
        RB_CUDA(cudaMemcpyAsync(edges, d_edges, sizeof(edges), cudaMemcpyDeviceToHost, d.stream));
    }
    cudaError_t se = cudaStreamSynchronize(d.stream);
    cudaFree(d_edges);
    RB_CUDA(se);
    h.set_level_edges(std::vector<uint64_t>(edges, edges + rb::kMaxLevels));
    ix->host_graph_valid = true;
    return {};
}

}  // namespace

// ================================================================================================ build
namespace {

struct Batch {
    uint32_t start, size;
    uint32_t entry;
    int32_t max_level;
    uint32_t tasks;
    uint32_t tasks_bound;  // size + sum of levels: known after pass 1, >= tasks
    uint32_t levels;       // levels with work in this batch: 1 + max over its nodes of min(level, max_level)
    uint32_t level_cnt[rb::kMaxLevels];  // nodes of the batch that have a task on level l = a prefix of the work order
};

// run f(i) for i in [0, count) on up to 16 host threads (static interleaved split; f must be thread-safe per i)
template <typename F>
void parallel_for(size_t count, size_t min_per_thread, F&& f) {
    size_t threads = std::min<size_t>({16, std::max(1u, std::thread::hardware_concurrency()),
                                       std::max<size_t>(1, count / std::max<size_t>(1, min_per_thread))});
    if (threads <= 1) {
        for (size_t i = 0; i < count; ++i)
            f(i);
        return;
    }
    std::vector<std::thread> pool;
    for (size_t t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            for (size_t i = t; i < count; i += threads)
                f(i);
        });
    for (std::thread& th : pool)
        th.join();
}

// The schedule is a pure function of (current size, levels): mirrored by tests/model/batched_model.c.
// Pass 1 (sequential, cheap): batch boundaries, and the entry point / top level each batch starts from.
void plan_batch_bounds(rb_index* ix, size_t old_size, std::vector<int16_t> const& all_levels /* old + new */,
                       std::vector<Batch>& batches, int32_t& max_level, uint64_t& entry) {
    const size_t total = all_levels.size();
    const uint32_t den = ix->opts.batch_growth_den, bmax = ix->opts.batch_max;
    size_t s = old_size;
    while (s < total) {
        if (ix->opts.debug_max_batches && batches.size() >= ix->opts.debug_max_batches)
            break;
        if (s == 0) {
            batches.push_back(Batch{0, 1, 0, -1, 0, 0, 0, {0}});  // the first node only becomes the entry point
            entry = 0;
            max_level = all_levels[0];
            s = 1;
            continue;
        }
        size_t b = std::max<size_t>(1, s / den);
        b = std::min<size_t>(b, bmax);
        b = std::min(b, total - s);
        bool raised = false;
        int32_t run_max = max_level;
        uint32_t bound = 0;
        for (size_t t = 0; t < b; ++t) {
            if (all_levels[s + t] > run_max) {
                if (raised) {
                    b = t;
                    break;
                }
                raised = true;
                run_max = all_levels[s + t];
            }
            bound += 1u + (uint32_t)all_levels[s + t];
        }
        batches.push_back(Batch{(uint32_t)s, (uint32_t)b, (uint32_t)entry, max_level, 0, bound, 0, {0}});
        if (raised)
            for (size_t t = 0; t < b; ++t)
                if (all_levels[s + t] > max_level) {
                    max_level = all_levels[s + t];
                    entry = s + t;
                }
        s += b;
    }
}

// Pass 2 for batches [b0, b1) (one batch per task, on host threads): first task of every node, and the order of work
// inside the batch -- nodes with more levels to search first (longest job first), so the last wave of a launch holds
// only the short level-0 jobs; the nodes are independent, so any order gives the same graph.  Stable counting sort
// by the number of searched levels, descending.  task_rel / task_order are indexed by new node.
void plan_batch_tasks(size_t old_size, std::vector<int16_t> const& all_levels, std::vector<Batch>& batches, size_t b0,
                      size_t b1, std::vector<uint32_t>& task_rel, std::vector<uint32_t>& task_order) {
    parallel_for(b1 - b0, 4, [&](size_t rel) {
        const size_t bi = b0 + rel;
        Batch& bt = batches[bi];
        if (bt.max_level < 0)
            return;
        const size_t s0 = bt.start, b = bt.size;
        const int32_t top = bt.max_level;
        uint32_t tasks = 0;
        size_t count[64] = {0};
        for (size_t t = 0; t < b; ++t) {
            const size_t jobs = (size_t)std::min<int32_t>(all_levels[s0 + t], top);
            task_rel[s0 + t - old_size] = tasks;
            tasks += (uint32_t)jobs + 1;
            ++count[std::min<size_t>(jobs, 63)];
        }
        bt.tasks = tasks;
        size_t start[64];
        size_t acc = 0;
        for (int l = 63; l >= 0; --l) {
            start[l] = acc;
            acc += count[l];
            if (l < rb::kMaxLevels) {
                bt.level_cnt[l] = (uint32_t)acc;  // nodes with at least l searched levels above level 0
                if (acc && !bt.levels)
                    bt.levels = (uint32_t)l + 1;
            }
        }
        for (size_t t = 0; t < b; ++t) {
            const size_t jobs = (size_t)std::min<int32_t>(all_levels[s0 + t], top);
            task_order[s0 - old_size + start[std::min<size_t>(jobs, 63)]++] = (uint32_t)t;
        }
    });
}

// the exact-search kernels use more than 48 KB of dynamic shared memory: opt in once per DEVICE (the attribute belongs
// to the device's context, not to the process)
std::string configure_exact_kernels() {
    static std::mutex mu;
    static bool configured[64] = {false};
    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && configured[dev])
        return {};
    RB_CUDA(cudaFuncSetAttribute(rb::exact_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)sizeof(rb::ExactSmem)));
    RB_CUDA(cudaFuncSetAttribute(rb::exact_topk_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)rb::kTensorSmemBytes));
    RB_CUDA(cudaFuncSetAttribute(rb::exact_topk_f4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)rb::kF4SmemBytes));
    if (dev >= 0 && dev < 64)
        configured[dev] = true;
    return {};
}

// resident CTAs per SM that `smem` bytes of dynamic shared memory per CTA allow, quantised to the compiled variants
int traversal_variant(size_t smem) {
    const size_t per_sm = 227 * 1024;
    size_t fit = per_sm / (smem + 1024);  // 1 KB of system-reserved shared memory per CTA
    return fit >= 8 ? 8 : fit >= 4 ? (int)fit : 4;
}

// expands to KERNEL<4..8> according to the run-time variant
#define RB_FOR_VARIANT(variant, STMT)                       \
    switch (variant) {                                      \
        case 8: { constexpr int V = 8; STMT; } break;       \
        case 7: { constexpr int V = 7; STMT; } break;       \
        case 6: { constexpr int V = 6; STMT; } break;       \
        case 5: { constexpr int V = 5; STMT; } break;       \
        default: { constexpr int V = 4; STMT; } break;      \
    }

std::string configure_variant(const void* fn, size_t smem, int& blocks_per_sm) {
    RB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    RB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, fn, rb::kWarpsPerBlock * 32, smem));
    return {};
}

std::string configure_traversal_kernels(rb_index* ix) {
    DeviceState& d = ix->dev;
    const uint32_t bloom = ix->opts.bloom_bits_log2, m = ix->host.m;
    if (d.insert_cfg_ef != ix->host.ef_add) {
        size_t smem = rb::warp_scratch_bytes(ix->host.ef_add, bloom, m) * rb::kWarpsPerBlock;
        d.insert_variant = traversal_variant(smem);
        const void* fn = nullptr;
        RB_FOR_VARIANT(d.insert_variant, fn = (const void*)rb::insert_search_kernel<V>);
        std::string err = configure_variant(fn, smem, d.insert_blocks_per_sm);
        if (!err.empty())
            return err;
        if (d.insert_blocks_per_sm < 1)
            return "expansion_add does not fit the shared memory of one SM";
        err = configure_variant((const void*)rb::insert_select_kernel,
                                rb::select_scratch_bytes(ix->host.ef_add, m) * rb::kWarpsPerBlock,
                                d.select_blocks_per_sm);
        if (!err.empty())
            return err;
        if (d.select_blocks_per_sm < 1)
            return "expansion_add does not fit the shared memory of one SM";
        d.insert_cfg_ef = ix->host.ef_add;
    }
    return {};
}

std::string configure_search_kernel(rb_index* ix, uint32_t ef) {
    DeviceState& d = ix->dev;
    if (d.search_cfg_ef != ef) {
        size_t smem = rb::warp_scratch_bytes(ef, ix->opts.bloom_bits_log2, ix->host.m) * rb::kWarpsPerBlock;
        d.search_variant = traversal_variant(smem);
        const void* fn = nullptr;
        RB_FOR_VARIANT(d.search_variant, fn = (const void*)rb::search_kernel<V>);
        std::string err = configure_variant(fn, smem, d.search_blocks_per_sm);
        if (!err.empty())
            return err;
        if (d.search_blocks_per_sm < 1)
            return "expansion_search does not fit the shared memory of one SM";
        d.search_cfg_ef = ef;
    }
    return {};
}

// shared tail of rb_add / rb_add_device: the fingerprints of the new rows are already enqueued on the stream
std::string build_new_rows(rb_index* ix, uint64_t const* keys, size_t n, std::vector<int16_t>& new_levels,
                           rb_progress_t progress, void* user, size_t& accepted) {
    rbh::HostIndex& h = ix->host;
    DeviceState& d = ix->dev;
    const size_t old_size = h.keys.size();
    accepted = 0;

    std::vector<int16_t> all_levels(h.levels);
    all_levels.insert(all_levels.end(), new_levels.begin(), new_levels.end());
    std::vector<Batch> batches;
    std::vector<uint32_t> task_rel(n, 0), task_order(n, 0);
    int32_t max_level = h.max_level;
    uint64_t entry = h.entry;
    PhaseTimer pt;
    plan_batch_bounds(ix, old_size, all_levels, batches, max_level, entry);
    pt.mark("  plan_batch_bounds");

    // Per-node metadata of the new rows (upper-row offset, level, task tables) is prepared and uploaded in TWO
    // segments: the batches over the first 1/16 of a large add are launched as soon as their own metadata is on the
    // device, and the rest is prepared while the device works through them (uploaded on a second stream, so the copy
    // does not queue behind those kernels).  The schedule and the graph do not depend on the split.
    std::vector<uint32_t> off32(n);
    std::vector<int8_t> lv8(n);
    uint64_t upper_run = h.upper_off[old_size];
    size_t split_batch = batches.size(), split_row = n;  // batches [0, split_batch) = rows [0, split_row) go first
    size_t segment_min_rows = (size_t)1 << 21;
    if (char const* e = std::getenv("RB_SEGMENT_MIN_ROWS"))  // tests exercise the split on small adds
        segment_min_rows = std::max<size_t>(16, std::strtoull(e, nullptr, 10));
    if (n >= segment_min_rows)
        for (size_t bi = 0; bi < batches.size(); ++bi)
            if (batches[bi].start - old_size >= n / 16) {
                split_batch = bi;
                split_row = batches[bi].start - old_size;
                break;
            }
    auto prepare_segment = [&](size_t b0, size_t b1, size_t r0, size_t r1, cudaStream_t stream) -> std::string {
        if (r1 <= r0)
            return {};
        plan_batch_tasks(old_size, all_levels, batches, b0, b1, task_rel, task_order);
        for (size_t i = r0; i < r1; ++i) {
            off32[i] = (uint32_t)upper_run;
            upper_run += (uint64_t)new_levels[i];
        }
        parallel_for((r1 - r0 + 65535) / 65536, 8, [&](size_t c) {
            for (size_t i = r0 + c * 65536, e = std::min(r1, i + 65536); i < e; ++i)
                lv8[i] = (int8_t)new_levels[i];
        });
        const size_t at = old_size + r0, cnt = r1 - r0;
        RB_CUDA(cudaMemcpyAsync(d.upper_off + at, off32.data() + r0, cnt * 4, cudaMemcpyHostToDevice, stream));
        RB_CUDA(cudaMemcpyAsync(d.levels + at, lv8.data() + r0, cnt, cudaMemcpyHostToDevice, stream));
        RB_CUDA(cudaMemcpyAsync(d.task_rel + at, task_rel.data() + r0, cnt * 4, cudaMemcpyHostToDevice, stream));
        RB_CUDA(cudaMemcpyAsync(d.task_order + at, task_order.data() + r0, cnt * 4, cudaMemcpyHostToDevice, stream));
        return {};
    };
    {
        std::string perr = prepare_segment(0, split_batch, 0, split_row, d.stream);
        if (!perr.empty())
            return perr;
    }

    // scratch sized for the largest batch (tasks_bound >= tasks is known for every batch after pass 1)
    size_t max_pairs = 0;
    for (Batch const& b : batches)
        max_pairs = std::max(max_pairs, (size_t)b.tasks_bound * h.m);
    if (max_pairs > d.pair_cap) {
        size_t cap = ceil2(max_pairs);
        for (int i = 0; i < 2; ++i) {
            cudaFree(d.pair_keys[i]);
            cudaFree(d.pair_vals[i]);
            RB_CUDA(cudaMalloc(&d.pair_keys[i], cap * 8));
            RB_CUDA(cudaMalloc(&d.pair_vals[i], cap * 8));
        }
        size_t tmp = 0;
        RB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, d.pair_keys[0], d.pair_keys[1], d.pair_vals[0],
                                                d.pair_vals[1], (int64_t)cap, 0, 40, d.stream));
        if (tmp > d.sort_tmp_bytes) {
            cudaFree(d.sort_tmp);
            RB_CUDA(cudaMalloc(&d.sort_tmp, tmp));
            d.sort_tmp_bytes = tmp;
        }
        d.pair_cap = cap;
    }
    // the per-task arrays of a batch: candidate lists, their lengths, the batch tile's seeds and answers
    {
        size_t max_tasks = 1, max_batch = 1;
        for (Batch const& b : batches) {
            max_tasks = std::max(max_tasks, (size_t)b.tasks_bound);
            max_batch = std::max(max_batch, (size_t)b.size);
        }
        if (max_tasks > d.task_cap || h.ef_add > d.task_cap_ef) {
            const size_t cap = std::max(ceil2(max_tasks), d.task_cap), ef = std::max<size_t>(h.ef_add, d.task_cap_ef);
            cudaFree(d.task_lists);
            cudaFree(d.task_lens);
            cudaFree(d.task_seeds);
            cudaFree(d.mate_keys);
            cudaFree(d.mate_iu);
            d.task_lists = nullptr, d.task_lens = nullptr, d.task_seeds = nullptr, d.mate_keys = nullptr;
            d.mate_iu = nullptr;
            d.task_cap = d.task_cap_ef = 0;
            RB_CUDA(cudaMalloc(&d.task_lists, cap * ef * 8));
            RB_CUDA(cudaMalloc(&d.task_lens, cap * 4));
            RB_CUDA(cudaMalloc(&d.task_seeds, cap * 8));
            RB_CUDA(cudaMalloc(&d.mate_keys, cap * rb::kBatchMates * 8));
            RB_CUDA(cudaMalloc(&d.mate_iu, cap * rb::kBatchMates * 4));
            d.task_cap = cap;
            d.task_cap_ef = ef;
        }
        if (max_batch > d.batch_fp_cap) {
            cudaFree(d.batch_fp);
            d.batch_fp = nullptr;
            d.batch_fp_cap = 0;
            const size_t cap = ceil2(max_batch);
            RB_CUDA(cudaMalloc(&d.batch_fp, cap * 128));
            d.batch_fp_cap = cap;
        }
    }
    // three work counters per batch: insert_search_kernel's (nodes) at [bi], reverse_link_kernel's (pairs) at
    // [batches + bi], insert_select_kernel's (nodes) at [2 batches + bi]
    if (3 * batches.size() > d.work_counters_cap) {
        cudaFree(d.work_counters);
        d.work_counters_cap = ceil2(3 * batches.size());
        RB_CUDA(cudaMalloc(&d.work_counters, d.work_counters_cap * sizeof(unsigned int)));
    }
    RB_CUDA(cudaMemsetAsync(d.work_counters, 0, std::max<size_t>(3 * batches.size(), 1) * sizeof(unsigned int),
                            d.stream));
    RB_CUDA(cudaMemsetAsync(d.counters, 0, 4 * sizeof(unsigned long long), d.stream));

    std::string err = configure_traversal_kernels(ix);
    if (!err.empty())
        return err;
    const size_t smem = rb::warp_scratch_bytes(h.ef_add, ix->opts.bloom_bits_log2, h.m) * rb::kWarpsPerBlock;
    const size_t select_smem = rb::select_scratch_bytes(h.ef_add, h.m) * rb::kWarpsPerBlock;
    const bool use_mates = !ix->opts.no_batch_tile;
    if (use_mates && !(err = configure_exact_kernels()).empty())
        return err;
    const rb::GraphView g = graph_view(ix);
    uint64_t launches = 0;
    size_t inserted = 0;
    bool cancelled = false;
    std::vector<cudaEvent_t> marks;  // progress marks
    size_t next_mark = 0;
    const bool time_kernels = (ix->opts.flags & RB_FLAG_TIME_KERNELS) != 0;
    // 6 per batch with graph work: before search, after search, after the batch tile, after select, after sort,
    // after reverse
    std::vector<cudaEvent_t> ticks;
    auto tick = [&]() -> cudaError_t {
        cudaEvent_t ev;
        cudaError_t e = cudaEventCreate(&ev);
        if (e != cudaSuccess)
            return e;
        ticks.push_back(ev);
        return cudaEventRecord(ev, d.stream);
    };
    pt.mark("  first segment uploads + scratch");
    RB_CUDA(cudaEventRecord(d.ev_begin, d.stream));
    auto launch_batches = [&](size_t first_batch, size_t last_batch) -> std::string {
    for (size_t bi = first_batch; bi < last_batch && !cancelled; ++bi) {
        Batch const& b = batches[bi];
        if (b.max_level >= 0) {
            rb::InsertParams ip;
            ip.g = g;
            ip.levels = d.levels;
            ip.task_rel = d.task_rel;
            ip.order = d.task_order + b.start;
            ip.pair_keys = d.pair_keys[0];
            ip.pair_vals = d.pair_vals[0];
            ip.batch_start = b.start;
            ip.batch_size = b.size;
            ip.entry = b.entry;
            ip.max_level = b.max_level;
            ip.ef = h.ef_add;
            ip.bloom_log2 = ix->opts.bloom_bits_log2;
            ip.work_counter = d.work_counters + bi;
            ip.counters = d.counters;
            ip.lists = d.task_lists;
            ip.lens = d.task_lens;
            ip.seeds = d.task_seeds;
            ip.mate_keys = d.mate_keys;
            ip.batch_fp = d.batch_fp;
            ip.use_mates = use_mates ? 1u : 0u;
            {
                uint32_t off = 0;
                for (int l = 0; l < rb::kMaxLevels; ++l) {
                    ip.level_off[l] = off;
                    off += b.level_cnt[l];
                }
            }
            unsigned blocks = (unsigned)std::min<size_t>((b.size + rb::kWarpsPerBlock - 1) / rb::kWarpsPerBlock,
                                                         (size_t)d.sm_count * d.insert_blocks_per_sm);
            if (time_kernels)
                RB_CUDA(tick());
            RB_FOR_VARIANT(d.insert_variant,
                           (rb::insert_search_kernel<V><<<blocks, rb::kWarpsPerBlock * 32, smem, d.stream>>>(ip)));
            ++launches;
            if (time_kernels)
                RB_CUDA(tick());
            if (use_mates) {
                // the batch tile: exact nearest EARLIER batch-mates of every task, all levels in one launch
                rb::ExactParams ep{};
                ep.db = ep.queries = d.batch_fp;
                ep.rows = b.size;
                ep.slot_base = 0;
                ep.query_count = b.size;
                ep.k = rb::kBatchMates;
                ep.chunks = 1;
                ep.chunk_rows = ((uint64_t)b.size + rb::kTD - 1) / rb::kTD * rb::kTD;
                ep.partial_keys = d.mate_keys;
                ep.partial_iu = d.mate_iu;
                ep.bounds = d.task_seeds;
                ep.debug_cycles = nullptr;
                static unsigned long long* tile_dbg = nullptr;
                const bool tile_debug = std::getenv("RB_TILE_DEBUG") && bi + 2 == batches.size();
                if (tile_debug) {
                    if (!tile_dbg)
                        RB_CUDA(cudaMalloc(&tile_dbg, 32 * sizeof(unsigned long long)));
                    RB_CUDA(cudaMemsetAsync(tile_dbg, 0, 32 * sizeof(unsigned long long), d.stream));
                    ep.debug_cycles = tile_dbg;
                }
                ep.tri = 1;
                for (int l = 0; l < rb::kMaxLevels; ++l) {
                    ep.tri_cnt[l] = b.level_cnt[l];
                    ep.tri_off[l] = ip.level_off[l];
                }
                dim3 grid((unsigned)((b.size + rb::kTQ - 1) / rb::kTQ), b.levels);
                const int forced_kernel = exact_kernel_override();  // (the FP4 kernel has no cycle counters)
                if (!tile_debug && (forced_kernel == 0 || forced_kernel == 5))
                    rb::exact_topk_f4_kernel<<<grid, rb::kFThreads, rb::kF4SmemBytes, d.stream>>>(ep);
                else
                    rb::exact_topk_i8_kernel<<<grid, rb::kTThreads, rb::kTensorSmemBytes, d.stream>>>(ep);
                ++launches;
                if (tile_debug) {
                    unsigned long long hh[32];
                    RB_CUDA(cudaStreamSynchronize(d.stream));
                    RB_CUDA(cudaMemcpy(hh, tile_dbg, sizeof(hh), cudaMemcpyDeviceToHost));
                    double tt = (double)std::max<unsigned long long>(hh[31], 1);
                    fprintf(stderr, "[rb tile] batch %zu size %u: longest CTA tiles=%llu, cycles per tile: producer %.0f (gload %.0f, "
                                    "wait empty %.0f, wait pd_empty %.0f) mma %.0f (wait tmem_empty %.0f, wait full %.0f) "
                                    "epilogue %.0f (wait pd_full %.0f, wait tmem_full %.0f, tmem ld %.0f, filter %.0f, slow %.0f, "
                                    "lane-0 inserts/tile %.1f; whole CTA: %llu filter hits, %llu insertions)\n", bi, b.size, hh[31], hh[0] / tt, hh[1] / tt, hh[2] / tt, hh[3] / tt,
                            hh[10] / tt, hh[12] / tt, hh[13] / tt, hh[20] / tt, hh[21] / tt, hh[22] / tt, hh[23] / tt, hh[24] / tt,
                            hh[26] / tt, hh[28] / tt, hh[29], hh[30]);
                }
            }
            if (time_kernels)
                RB_CUDA(tick());
            {
                rb::InsertParams sp = ip;
                sp.work_counter = d.work_counters + 2 * batches.size() + bi;
                unsigned sblocks = (unsigned)std::min<size_t>((b.size + rb::kWarpsPerBlock - 1) / rb::kWarpsPerBlock,
                                                              (size_t)d.sm_count * d.select_blocks_per_sm);
                rb::insert_select_kernel<<<sblocks, rb::kWarpsPerBlock * 32, select_smem, d.stream>>>(sp);
                ++launches;
            }
            if (time_kernels)
                RB_CUDA(tick());
            const size_t pairs = (size_t)b.tasks * h.m;
            size_t tmp = d.sort_tmp_bytes;
            RB_CUDA(cub::DeviceRadixSort::SortPairs(d.sort_tmp, tmp, d.pair_keys[0], d.pair_keys[1], d.pair_vals[0],
                                                    d.pair_vals[1], (int64_t)pairs, 0, 40, d.stream));
            rb::ReverseParams rp;
            rp.g = g;
            rp.pair_keys = d.pair_keys[1];
            rp.pair_vals = d.pair_vals[1];
            rp.pair_count = (uint32_t)pairs;
            rp.counters = d.counters;
            rp.work_counter = d.work_counters + batches.size() + bi;
            // persistent warps taking 32 pairs at a time: no more CTAs than fit the SMs at once
            unsigned rblocks = (unsigned)std::min<size_t>((pairs + 32 * rb::kWarpsPerBlock - 1) / (32 * rb::kWarpsPerBlock),
                                                          (size_t)d.sm_count * 8);
            if (time_kernels)
                RB_CUDA(tick());
            rb::reverse_link_kernel<<<rblocks, rb::kWarpsPerBlock * 32, 0, d.stream>>>(rp);
            ++launches;  // (the radix-sort passes between the two are library code and not counted)
            if (time_kernels)
                RB_CUDA(tick());
            RB_CUDA(cudaGetLastError());
        }
        inserted += b.size;
        if (progress && (bi % 8 == 7 || bi + 1 == batches.size())) {
            cudaEvent_t ev;
            RB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            RB_CUDA(cudaEventRecord(ev, d.stream));
            marks.push_back(ev);
            // report whatever has completed so far without blocking the launch loop
            while (next_mark < marks.size() && cudaEventQuery(marks[next_mark]) == cudaSuccess)
                ++next_mark;
            size_t done_batches = next_mark ? std::min(batches.size(), next_mark * 8) : 0;
            size_t done_rows = 0;
            for (size_t t = 0; t < done_batches; ++t)
                done_rows += batches[t].size;
            if (!progress(done_rows, n, user)) {
                cancelled = true;
                break;
            }
        }
    }
    return {};
    };
    {
        std::string lerr = launch_batches(0, split_batch);
        if (lerr.empty() && split_batch < batches.size() && !cancelled) {
            // second segment: prepared while the device runs the first, uploaded beside it, then ordered after it
            pt.mark("  first segment launched");
            if (!d.copy_stream)
                RB_CUDA(cudaStreamCreateWithFlags(&d.copy_stream, cudaStreamNonBlocking));
            if (!d.ev_copy)
                RB_CUDA(cudaEventCreateWithFlags(&d.ev_copy, cudaEventDisableTiming));
            lerr = prepare_segment(split_batch, batches.size(), split_row, n, d.copy_stream);
            if (lerr.empty()) {
                RB_CUDA(cudaEventRecord(d.ev_copy, d.copy_stream));
                RB_CUDA(cudaStreamWaitEvent(d.stream, d.ev_copy, 0));
                pt.mark("  second segment prepared");
                lerr = launch_batches(split_batch, batches.size());
            }
        }
        if (!lerr.empty()) {
            cudaStreamSynchronize(d.stream);
            if (d.copy_stream)
                cudaStreamSynchronize(d.copy_stream);
            for (cudaEvent_t ev : marks)
                cudaEventDestroy(ev);
            for (cudaEvent_t ev : ticks)
                cudaEventDestroy(ev);
            return lerr;
        }
    }
    RB_CUDA(cudaEventRecord(d.ev_end, d.stream));
    // Host bookkeeping of the new rows.  Without a progress callback nothing can cut the schedule short, so it runs
    // HERE, while the device is still working through the batches, and is undone if the device reports an error.
    const int32_t old_max_level = h.max_level;
    const uint64_t old_entry = h.entry;
    auto commit = [&](size_t count) {
        int32_t top = h.max_level;
        uint64_t top_slot = h.entry;
        for (size_t i = 0; i < count; ++i) {
            size_t slot = old_size + i;
            if (slot == 0 || new_levels[i] > top) {
                top = new_levels[i];
                top_slot = slot;
            }
        }
        h.keys.insert(h.keys.end(), keys, keys + count);
        h.levels.insert(h.levels.end(), new_levels.begin(), new_levels.begin() + count);
        h.upper_off.resize(old_size + count + 1);
        for (size_t i = 0; i < count; ++i)
            h.upper_off[old_size + i + 1] = h.upper_off[old_size + i] + (uint64_t)new_levels[i];
        h.max_level = top;
        h.entry = top_slot;
        h.note_new_keys(keys, count);
        h.note_new_levels(old_size, new_levels.data(), count);
        h.invalidate_level_edges();
    };
    const bool early = !progress && !cancelled && inserted == n;
    pt.mark("  launch loop");
    if (early)
        commit(n);
    // the host copy of the adjacency is read back right after the build unless the caller asked for a lazy sync:
    // size (and page in) its destination now, while the device is busy, instead of inside the timed copy
    const bool presize = early && !(ix->opts.flags & RB_FLAG_LAZY_HOST_SYNC);
    if (presize) {
        h.adj0.resize(h.keys.size() * (size_t)h.m0);
        h.adju.resize(h.upper_off[h.keys.size()] * (size_t)h.m);
    }
    pt.mark("  commit (device busy)");
    {
        cudaError_t e = cudaStreamSynchronize(d.stream);
        if (e == cudaSuccess)
            e = cudaGetLastError();
        if (e != cudaSuccess) {
            if (early)
                h.truncate_rows(old_size, old_max_level, old_entry);
            if (presize) {
                h.adj0.resize(old_size * (size_t)h.m0);
                h.adju.resize(h.upper_off[old_size] * (size_t)h.m);
            }
            for (cudaEvent_t ev : marks)
                cudaEventDestroy(ev);
            for (cudaEvent_t ev : ticks)
                cudaEventDestroy(ev);
            return std::string("CUDA error: ") + cudaGetErrorString(e);
        }
    }
    for (cudaEvent_t ev : marks)
        cudaEventDestroy(ev);
    pt.mark("  wait for the device");
    RB_CUDA(cudaGetLastError());
    float ms = 0;
    RB_CUDA(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
    unsigned long long counters[4];
    RB_CUDA(cudaMemcpy(counters, d.counters, sizeof(counters), cudaMemcpyDeviceToHost));
    ix->last.computed_distances = counters[0];
    ix->last.visited_members = counters[1];
#ifdef RB_PHASE_TIMING
    {
        unsigned long long ph[8], zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        RB_CUDA(cudaMemcpyFromSymbol(ph, rb::g_phase, sizeof(ph)));
        RB_CUDA(cudaMemcpyToSymbol(rb::g_phase, zero, sizeof(zero)));
        const double s = ph[6] ? (double)ph[6] : 1.0, e = counters[1] ? (double)counters[1] : 1.0;
        std::fprintf(stderr,
                     "[rb phase] searches %llu expansions %llu | cycles per search: pick %.0f rows %.0f stage %.0f eval %.0f "
                     "park %.0f merge %.0f | per expansion: pick %.0f rows %.0f stage %.0f eval %.0f park %.0f merge %.0f\n",
                     ph[6], counters[1], ph[0] / s, ph[1] / s, ph[2] / s, ph[3] / s, ph[4] / s, ph[5] / s, ph[0] / e,
                     ph[1] / e, ph[2] / e, ph[3] / e, ph[4] / e, ph[5] / e);
    }
#endif
    ix->last.batches = batches.size();
    ix->last.kernel_launches = launches;
    ix->last.device_ms = ms;
    ix->last.insert_ms = ix->last.sort_ms = ix->last.reverse_ms = ix->last.tile_ms = ix->last.select_ms = 0;
    ix->last.insert_launches = 0;
    for (size_t t = 0; t + 5 < ticks.size(); t += 6) {
        float dt[5] = {0, 0, 0, 0, 0};
        for (int j = 0; j < 5; ++j)
            cudaEventElapsedTime(&dt[j], ticks[t + j], ticks[t + j + 1]);
        ix->last.insert_ms += dt[0];
        ix->last.tile_ms += dt[1];
        ix->last.select_ms += dt[2];
        ix->last.sort_ms += dt[3];
        ix->last.reverse_ms += dt[4];
        ix->last.insert_launches += 1;
        if (pt.on && (t + 11 >= ticks.size() || std::getenv("RB_BATCH_TIMING")))  // the last batch (or every batch)
            std::fprintf(stderr, "[rb batch] %zu: search %.3f  tile %.3f  select %.3f  sort %.3f  reverse %.3f ms\n",
                         t / 6, dt[0], dt[1], dt[2], dt[3], dt[4]);
    }
    for (cudaEvent_t ev : ticks)
        cudaEventDestroy(ev);

    // commit what was inserted (everything unless a debug cap or a cancellation cut the schedule short)
    accepted = inserted;
    if (!early)
        commit(accepted);
    d.rows = old_size + accepted;
    d.upper_rows = h.upper_off[old_size + accepted];
    ix->device_graph_valid = true;
    ix->host_graph_valid = false;
    if (progress)
        progress(accepted, n, user);
    if (cancelled)
        return "Operation has been terminated";
    return {};
}

std::string add_common(rb_index* ix, uint64_t const* keys, uint8_t const* host_vectors, size_t stride,
                       void const* device_vectors, size_t n, rb_progress_t progress, void* user,
                       int16_t const* forced_levels = nullptr) {
    rbh::HostIndex& h = ix->host;
    if (h.immutable())
        return "Can't add to an immutable index";
    if (!n)
        return {};
    const size_t old_size = h.keys.size();
    if (old_size + n >= (1ull << 31))
        return "rad_b200 addresses at most 2^31 - 1 fingerprints per index (shard larger sets by rows)";
    std::vector<uint64_t> generated;
    if (!keys) {
        generated.resize(n);
        for (size_t i = 0; i < n; ++i)
            generated[i] = old_size + i;
        keys = generated.data();
    }
    PhaseTimer pt;
    if (h.check_new_keys(keys, n) >= 0)
        return "Duplicate keys not allowed in high-level wrappers";
    pt.mark("keys generated + checked");
    std::string err = ensure_device(ix);
    if (!err.empty())
        return err;
    if (!(err = upload_graph(ix)).empty())
        return err;
    std::vector<int16_t> new_levels;
    const uint32_t rng_before = h.level_rng;
    h.draw_levels(n, new_levels);  // (the stream advances either way: an index stays in step with its size)
    if (forced_levels) {
        for (size_t i = 0; i < n; ++i)
            if (forced_levels[i] < 0) {
                h.level_rng = rng_before;
                return "levels must not be negative";
            }
        new_levels.assign(forced_levels, forced_levels + n);
    }
    pt.mark("device ready + level stream");
    uint64_t upper = h.upper_off[old_size];
    int16_t top_new = 0;
    for (size_t i = 0; i < n; ++i) {
        upper += (uint64_t)new_levels[i];
        top_new = std::max(top_new, new_levels[i]);
    }
    if (upper >= (1ull << 32) || top_new >= rb::kMaxLevels) {
        h.level_rng = rng_before;
        return upper >= (1ull << 32) ? "too many upper-level rows" : "levels above 31 are not supported";
    }
    if (!(err = ensure_capacity(ix, ceil2(old_size + n), upper)).empty()) {
        h.level_rng = rng_before;
        return err;
    }
    DeviceState& d = ix->dev;
    if (device_vectors) {
        RB_CUDA(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(d.fp) + old_size * 128, device_vectors, n * 128,
                                cudaMemcpyDeviceToDevice, d.stream));
    } else if (!(err = upload_rows(ix, host_vectors, n, stride, old_size)).empty()) {
        return err;
    }
    // rows of the new nodes must start empty (a re-used capacity region may hold an aborted build)
    RB_CUDA(cudaMemsetAsync(d.adj0 + old_size * h.m0, 0xFF, n * (size_t)h.m0 * 4, d.stream));
    if (upper > h.upper_off[old_size])
        RB_CUDA(cudaMemsetAsync(d.adju + h.upper_off[old_size] * h.m, 0xFF,
                                (upper - h.upper_off[old_size]) * (size_t)h.m * 4, d.stream));
    size_t accepted = 0;
    pt.mark("capacity + copies enqueued");
    err = build_new_rows(ix, keys, n, new_levels, progress, user, accepted);
    pt.mark("build_new_rows");
    if (accepted < n) {
        // the level stream only advances over nodes that exist
        h.level_rng = rng_before;
        std::vector<int16_t> replay;
        h.draw_levels(accepted, replay);
    }
    if (!err.empty())
        return err;
    if (!(ix->opts.flags & RB_FLAG_LAZY_HOST_SYNC))
        return sync_host(ix);
    return {};
}

}  // namespace

// ================================================================================================ search
namespace {

struct ExactPlan {
    bool tensor;  // tcgen05 int8 kernel (128-query tiles, 256-row stages) instead of the POPC kernel (64 x 64)
    bool f4;      // the FP4 tensor-core kernel (exact_topk_f4_kernel: kind::mxf4, 224-row tiles) -- the default tensor path
    uint32_t chunks;
    uint64_t chunk_rows;
    size_t keys_bytes, iu_bytes;  // iu_bytes includes the per-query bound words at bounds_off
    size_t bounds_off;            // u64[Q] shared k-th-distance bounds (tensor-core kernels), from the workspace start
};

// RB_EXACT_KERNEL=popc|f4|i8 forces one of the bit-identical kernels (tests, profiling)
int exact_kernel_override() {
    char const* e = std::getenv("RB_EXACT_KERNEL");  // read per call: tests flip it between calls
    if (!e)
        return 0;
    return std::strcmp(e, "popc") == 0 ? 1 : std::strcmp(e, "i8") == 0 ? 2 : std::strcmp(e, "f4") == 0 ? 5 : 0;
}

// SMs of the current device (148 on a B200; the same figure when no device is visible, so that workspace sizes can
// still be asked for)
uint64_t current_sm_count() {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        sms <= 0) {
        cudaGetLastError();
        return 148;
    }
    return (uint64_t)sms;
}

ExactPlan plan_exact(uint64_t rows, uint64_t nq, uint32_t k) {
    ExactPlan p;
    const int forced = exact_kernel_override();
    // the tensor-core tile is 128 queries wide: worth it once at least half a tile is used and there are enough
    // pairs per CTA to amortise unpacking the queries and warming up the per-CTA top-k lists (measured crossover
    // with the POPC kernel: about 1e9 pairs)
    p.tensor = forced >= 2 || (forced == 0 && nq >= 64 && rows * nq >= 2000000000ull);
    p.f4 = forced == 5 || (forced == 0 && p.tensor);  // the FP4 kernel unless kind::i8 is asked for
    const uint64_t tq = p.tensor ? rb::kTQ : rb::kXQ, td = p.f4 ? rb::kFD : p.tensor ? rb::kTD : rb::kXD;
    const uint64_t sms = current_sm_count();
    const uint64_t resident = p.tensor ? sms : 2 * sms;  // resident CTAs: one (tensor kernel) or two (POPC kernel) per SM
    uint64_t qtiles = std::max<uint64_t>(1, (nq + tq - 1) / tq);
    uint64_t row_tiles = std::max<uint64_t>(1, (rows + td - 1) / td);
    // Database chunks: 3 to 8 waves of CTAs, and among those the count that leaves the LAST wave fullest -- all CTAs of a
    // launch run equally long, so 632 CTAs on 148 SMs (4.27 waves) cost five full waves (measured: 15 % of the launch)
    uint64_t best_chunks = 1, best_tpc = row_tiles;
    double best_util = -1.0;
    const uint64_t c_lo = std::max<uint64_t>(1, (3 * resident + qtiles - 1) / qtiles);
    const uint64_t c_hi = std::max<uint64_t>(c_lo, (8 * resident) / qtiles);
    for (uint64_t c = c_lo; c <= c_hi; ++c) {
        const uint64_t cc = std::min(c, row_tiles);
        const uint64_t tpc = (row_tiles + cc - 1) / cc;
        const uint64_t real = (row_tiles + tpc - 1) / tpc;  // chunks after rounding to whole tiles
        const uint64_t ctas = real * qtiles, waves = (ctas + resident - 1) / resident;
        const double util = (double)ctas / (double)(waves * resident);
        if (util > best_util + 1e-9) {
            best_util = util;
            best_chunks = real;
            best_tpc = tpc;
        }
    }
    if (char const* e = std::getenv("RB_EXACT_CHUNKS")) {  // profiling aid: force the chunk count
        const uint64_t cc = std::min<uint64_t>(std::max<uint64_t>(1, std::strtoull(e, nullptr, 10)), row_tiles);
        best_tpc = (row_tiles + cc - 1) / cc;
        best_chunks = (row_tiles + best_tpc - 1) / best_tpc;
    }
    const uint64_t tiles_per_chunk = best_tpc;
    p.chunk_rows = tiles_per_chunk * td;
    p.chunks = (uint32_t)best_chunks;
    p.keys_bytes = ((size_t)p.chunks * nq * k * 8 + 255) & ~(size_t)255;
    p.iu_bytes = ((size_t)p.chunks * nq * k * 4 + 255) & ~(size_t)255;
    p.bounds_off = p.keys_bytes + p.iu_bytes;
    p.iu_bytes += ((size_t)nq * 8 + 255) & ~(size_t)255;
    return p;
}

std::string launch_exact(uint4 const* db, uint64_t rows, uint64_t slot_base, uint4 const* queries, uint64_t nq,
                         uint32_t k, uint64_t* keys_out, void* workspace, size_t workspace_bytes,
                         cudaStream_t stream, bool allow_debug_counters = true) {
    if (!nq)
        return {};
    if (k == 0 || k > 1024)
        return "k must be in 1..1024";
    if (slot_base + rows > (1ull << 32))
        return "exact search addresses at most 2^32 slots";
    ExactPlan pl = plan_exact(rows, nq, k);
    if (workspace_bytes < pl.keys_bytes + pl.iu_bytes)
        return "workspace too small (see rb_exact_topk_workspace_bytes)";
    {
        std::string cerr = configure_exact_kernels();
        if (!cerr.empty())
            return cerr;
    }
    rb::ExactParams ep{};
    ep.db = db;
    ep.queries = queries;
    ep.rows = rows;
    ep.slot_base = slot_base;
    ep.query_count = (uint32_t)nq;
    ep.k = k;
    ep.chunks = pl.chunks;
    ep.chunk_rows = pl.chunk_rows;
    ep.partial_keys = static_cast<uint64_t*>(workspace);
    ep.partial_iu = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(workspace) + pl.keys_bytes);
    ep.bounds = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(workspace) + pl.bounds_off);
    if (pl.tensor && pl.chunks <= 512)  // the bound word has 9 bits for the chunk
        RB_CUDA(cudaMemsetAsync(ep.bounds, 0xFF, (size_t)nq * 8, stream));
    else
        ep.bounds = nullptr;
    ep.debug_cycles = nullptr;
    static unsigned long long* dbg_dev = nullptr;
    if (allow_debug_counters && pl.tensor && std::getenv("RB_EXACT_DEBUG")) {  // (this aid allocates its counters)
        if (!dbg_dev)
            RB_CUDA(cudaMalloc(&dbg_dev, 32 * sizeof(unsigned long long)));
        RB_CUDA(cudaMemsetAsync(dbg_dev, 0, 32 * sizeof(unsigned long long), stream));
        ep.debug_cycles = dbg_dev;
    }
    if (pl.tensor) {
        dim3 grid((unsigned)((nq + rb::kTQ - 1) / rb::kTQ), pl.chunks);  // query tiles fastest: CTAs resident
        if (pl.f4)
            rb::exact_topk_f4_kernel<<<grid, rb::kFThreads, rb::kF4SmemBytes, stream>>>(ep);
        else
            rb::exact_topk_i8_kernel<<<grid, rb::kTThreads, rb::kTensorSmemBytes, stream>>>(ep);  // together share rows
    } else {
        dim3 grid(pl.chunks, (unsigned)((nq + rb::kXQ - 1) / rb::kXQ));
        rb::exact_topk_kernel<<<grid, rb::kXThreads, sizeof(rb::ExactSmem), stream>>>(ep);
    }
    rb::MergeParams mp;
    mp.partial_keys = ep.partial_keys;
    mp.parts = pl.chunks;
    mp.query_count = (uint32_t)nq;
    mp.k = k;
    mp.keys_out = keys_out;
    rb::merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 4 * (size_t)k * 8, stream>>>(mp);
    RB_CUDA(cudaGetLastError());
    if (ep.debug_cycles) {
        unsigned long long h[32];
        RB_CUDA(cudaStreamSynchronize(stream));
        RB_CUDA(cudaMemcpy(h, dbg_dev, sizeof(h), cudaMemcpyDeviceToHost));
        double t = (double)std::max<unsigned long long>(h[31], 1);
        fprintf(stderr, "[rb exact i8] CTA(0,0) tiles=%llu, cycles per tile\n  producer total %.0f: gload %.0f, wait empty %.0f, "
                        "wait pd_empty %.0f\n  mma total %.0f: wait tmem_empty %.0f, wait full %.0f\n  epilogue total %.0f: "
                        "wait pd_full %.0f, wait tmem_full %.0f, tmem ld %.0f, filter %.0f, bar1 %.0f, queue %.0f, bar2+3 %.0f, "
                        "survivors/tile %.1f\n", h[31], h[0] / t, h[1] / t, h[2] / t, h[3] / t, h[10] / t, h[12] / t, h[13] / t,
                h[20] / t, h[21] / t, h[22] / t, h[23] / t, h[24] / t, h[25] / t, h[26] / t, h[27] / t, h[28] / t);
    }
    return {};
}

__global__ void unpack_topk32_kernel(const uint64_t* keys, uint64_t query_count, uint32_t k, uint64_t slot_base,
                                     uint32_t* slots, float* dists, int32_t* counts) {
    uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= query_count)
        return;
    uint32_t cnt = 0;
    for (uint32_t j = 0; j < k; ++j) {
        uint64_t key = keys[q * k + j];
        if (key == ~0ull)
            break;
        slots[q * k + j] = (uint32_t)(~(uint32_t)key - (uint32_t)slot_base);
        dists[q * k + j] = __uint_as_float((uint32_t)(key >> 32));
        ++cnt;
    }
    counts[q] = (int32_t)cnt;
}

std::string ensure_scratch(DeviceState& d, size_t bytes) {
    if (bytes <= d.scratch_bytes)
        return {};
    cudaFree(d.scratch);
    d.scratch = nullptr;
    d.scratch_bytes = 0;
    RB_CUDA(cudaMalloc(&d.scratch, bytes));
    d.scratch_bytes = bytes;
    return {};
}

// queries already on the device (padded rows); results into device buffers: slots u32, dists f32, counts i32
std::string search_on_device(rb_index* ix, uint4 const* d_queries, size_t nq, size_t k, bool exact, uint32_t* d_slots,
                             float* d_dists, int32_t* d_counts, void* extra_scratch, size_t extra_bytes) {
    rbh::HostIndex& h = ix->host;
    DeviceState& d = ix->dev;
    const size_t n = h.size();
    RB_CUDA(cudaMemsetAsync(d.counters, 0, 4 * sizeof(unsigned long long), d.stream));
    RB_CUDA(cudaEventRecord(d.ev_begin, d.stream));
    uint64_t launches = 0;
    if (exact) {
        ExactPlan pl = plan_exact(n, nq, (uint32_t)k);
        size_t need = pl.keys_bytes + pl.iu_bytes + nq * k * 8;
        if (extra_bytes < need)
            return "internal: scratch too small";
        uint64_t* merged = reinterpret_cast<uint64_t*>(static_cast<uint8_t*>(extra_scratch) + pl.keys_bytes + pl.iu_bytes);
        std::string err = launch_exact(d.fp, n, 0, d_queries, nq, (uint32_t)k, merged, extra_scratch,
                                       pl.keys_bytes + pl.iu_bytes, d.stream);
        if (!err.empty())
            return err;
        unpack_topk32_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, d.stream>>>(merged, nq, (uint32_t)k, 0, d_slots,
                                                                                 d_dists, d_counts);
        launches = 3;
    } else {
        uint32_t ef = (uint32_t)std::max<size_t>(h.ef_search, k);
        if (ef > 1024)
            return "expansion_search (or k) above 1024 is not supported";
        std::string err = configure_search_kernel(ix, ef);
        if (!err.empty())
            return err;
        unsigned int* wc = reinterpret_cast<unsigned int*>(d.counters + 3);
        rb::SearchParams sp;
        sp.g = graph_view(ix);
        sp.queries = d_queries;
        sp.query_count = (uint32_t)nq;
        sp.entry = (uint32_t)h.entry;
        sp.max_level = h.max_level;
        sp.ef = ef;
        sp.k = (uint32_t)k;
        sp.bloom_log2 = ix->opts.bloom_bits_log2;
        sp.slots_out = d_slots;
        sp.dists_out = d_dists;
        sp.counts_out = d_counts;
        sp.work_counter = wc;
        sp.counters = d.counters;
        size_t smem = rb::warp_scratch_bytes(ef, ix->opts.bloom_bits_log2, h.m) * rb::kWarpsPerBlock;
        unsigned blocks = (unsigned)std::min<size_t>((nq + rb::kWarpsPerBlock - 1) / rb::kWarpsPerBlock,
                                                     (size_t)d.sm_count * d.search_blocks_per_sm);
        RB_FOR_VARIANT(d.search_variant,
                       (rb::search_kernel<V><<<blocks, rb::kWarpsPerBlock * 32, smem, d.stream>>>(sp)));
        launches = 1;
    }
    RB_CUDA(cudaGetLastError());
    RB_CUDA(cudaEventRecord(d.ev_end, d.stream));
    RB_CUDA(cudaStreamSynchronize(d.stream));
    float ms = 0;
    RB_CUDA(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
    unsigned long long counters[4];
    RB_CUDA(cudaMemcpy(counters, d.counters, sizeof(counters), cudaMemcpyDeviceToHost));
    ix->last.computed_distances = exact ? (uint64_t)nq * n : counters[0];
    ix->last.visited_members = exact ? 0 : counters[1];
    ix->last.batches = 0;
    ix->last.kernel_launches = launches;
    ix->last.device_ms = ms;
    return {};
}

size_t exact_scratch_bytes(size_t n, size_t nq, size_t k) {
    ExactPlan pl = plan_exact(n, nq, (uint32_t)k);
    return pl.keys_bytes + pl.iu_bytes + nq * k * 8 + 256;
}

std::string prepare_search(rb_index* ix, size_t k) {
    rbh::HostIndex& h = ix->host;
    if (!k)
        return "k must be positive";
    std::string err = ensure_device(ix);
    if (!err.empty())
        return err;
    return h.viewing ? upload_view(ix) : upload_graph(ix);
}

}  // namespace

// ================================================================================================ C ABI
#define RB_FAIL(msg)              \
    do {                          \
        set_error(error, (msg));  \
        return;                   \
    } while (0)

extern "C" {

char const* rb_version(void) { return "rad_b200 0.1 (usearch-file 2.17.2, sm_100a)"; }

int rb_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess)
        return 0;
    return count;
}

rb_index_t* rb_init(rb_init_options_t const* options, rb_error_t* error) {
    if (error)
        *error = nullptr;
    rb_init_options_t o{};
    if (options)
        o = *options;
    if (!o.dimensions)
        o.dimensions = 1024;
    if (o.dimensions % 8 != 0 || o.dimensions > 1024) {
        set_error(error, "rad_b200 supports 8..1024-bit fingerprints (multiples of 8)");
        return nullptr;
    }
    if (!o.connectivity)
        o.connectivity = 16;
    if (o.connectivity < 2 || o.connectivity > 16) {
        set_error(error, "connectivity must be in 2..16");
        return nullptr;
    }
    if (o.connectivity_base && (o.connectivity_base < o.connectivity || o.connectivity_base > (uint32_t)rb::kMaxWidth)) {
        set_error(error, "connectivity_base must be in connectivity..32");
        return nullptr;
    }
    if (!o.expansion_add)
        o.expansion_add = 128;
    if (!o.expansion_search)
        o.expansion_search = 64;
    if (o.expansion_add > 1024 || o.expansion_search > 1024) {
        set_error(error, "expansion factors above 1024 are not supported");
        return nullptr;
    }
    if (!o.batch_growth_den)
        o.batch_growth_den = 8;
    if (!o.batch_max)
        o.batch_max = 65536;
    if (!o.bloom_bits_log2)
        o.bloom_bits_log2 = 15;
    if (o.bloom_bits_log2 < 10 || o.bloom_bits_log2 > 18) {
        set_error(error, "bloom_bits_log2 must be in 10..18");
        return nullptr;
    }
    rb_index* ix = new rb_index();
    ix->opts = o;
    ix->host.ndim = o.dimensions;
    ix->host.vec_bytes = (uint32_t)(o.dimensions / 8);
    ix->host.m = o.connectivity;
    ix->host.m0 = o.connectivity_base ? o.connectivity_base : 2 * o.connectivity;
    ix->host.ef_add = o.expansion_add;
    ix->host.ef_search = o.expansion_search;
    ix->host.upper_off.assign(1, 0);
    return ix;
}

void rb_free(rb_index_t* index) { delete index; }

void rb_release_device(rb_index_t* ix, int keep_host_vectors, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (ix->dev.device < 0)
        return;
    std::string err = sync_host(ix);
    if (!err.empty())
        RB_FAIL(err);
    if (keep_host_vectors && ix->device_graph_valid && ix->host.keys.size()) {
        rbh::HostIndex& h = ix->host;
        h.host_vectors.resize(h.keys.size() * (size_t)h.vec_bytes);
        cudaError_t e = cudaMemcpy2D(h.host_vectors.data(), h.vec_bytes, ix->dev.fp, 128, h.vec_bytes, h.keys.size(),
                                     cudaMemcpyDeviceToHost);
        if (e != cudaSuccess)
            RB_FAIL(std::string("CUDA error while fetching fingerprints: ") + cudaGetErrorString(e));
    }
    ix->dev.release();
    ix->device_graph_valid = false;
}

void rb_reserve(rb_index_t* ix, size_t capacity, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (ix->host.immutable())
        RB_FAIL("Can't reserve on an immutable index");
    std::string err = ensure_device(ix);
    if (err.empty())
        err = upload_graph(ix);
    if (err.empty())
        err = ensure_capacity(ix, capacity, std::max<size_t>(ix->host.upper_off.back(), capacity / 4));
    if (!err.empty())
        RB_FAIL(err);
    ix->host.capacity = std::max(ix->host.capacity, capacity);
}

void rb_add(rb_index_t* ix, uint64_t const* keys, uint8_t const* vectors, size_t count, size_t row_stride_bytes,
            int copy, rb_progress_t progress, void* progress_user, rb_error_t* error) {
    (void)copy;  // the index always owns a device copy (index_dense.hpp:2134-2139 allows either)
    if (error)
        *error = nullptr;
    if (count && !vectors)
        RB_FAIL("Expects a matrix of vectors to add!");
    if (row_stride_bytes < ix->host.vec_bytes)
        RB_FAIL("The number of vector dimensions doesn't match!");
    std::string err = add_common(ix, keys, vectors, row_stride_bytes, nullptr, count, progress, progress_user);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_add_device(rb_index_t* ix, uint64_t const* keys, void const* d_vectors, size_t count, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (count && !d_vectors)
        RB_FAIL("Expects a matrix of vectors to add!");
    std::string err = add_common(ix, keys, nullptr, 128, d_vectors, count, nullptr, nullptr);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_add_device_levels(rb_index_t* ix, uint64_t const* keys, void const* d_vectors, size_t count,
                          int16_t const* levels, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (count && (!d_vectors || !levels))
        RB_FAIL("Expects a matrix of vectors and their levels!");
    std::string err = add_common(ix, keys, nullptr, 128, d_vectors, count, nullptr, nullptr, levels);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_seek_level_stream(rb_index_t* ix, uint64_t first_row, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (ix->host.size())
        RB_FAIL("the level stream can only be positioned on an empty index");
    ix->host.seek_level_stream(first_row);
}

void rb_gather_vectors_device(rb_index_t* ix, void const* d_slots, uint64_t count, void* d_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!count)
        return;
    std::string err = prepare_search(ix, 1);
    if (!err.empty())
        RB_FAIL(err);
    DeviceState& d = ix->dev;
    const unsigned grid = (unsigned)std::min<uint64_t>((count * 8 + 255) / 256, (uint64_t)d.sm_count * 16);
    rb::gather_rows_kernel<<<grid, 256, 0, d.stream>>>(d.fp, static_cast<uint32_t const*>(d_slots), count,
                                                       static_cast<uint4*>(d_out));
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(d.stream);
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in gather: ") + cudaGetErrorString(e));
}

void rb_copy_vectors_device(rb_index_t* ix, uint64_t first, uint64_t count, void* d_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!count)
        return;
    std::string err = prepare_search(ix, 1);
    if (!err.empty())
        RB_FAIL(err);
    if (first + count > ix->host.size())
        RB_FAIL("rows out of range");
    DeviceState& d = ix->dev;
    cudaError_t e = cudaMemcpyAsync(d_out, reinterpret_cast<uint8_t const*>(d.fp) + first * 128, count * 128,
                                    cudaMemcpyDeviceToDevice, d.stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(d.stream);
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in copy: ") + cudaGetErrorString(e));
}

void rb_stitch_device(rb_index_t* ix, uint64_t first, uint64_t count, uint64_t slot_base, void const* d_cand_slots,
                      void const* d_cand_vectors, uint32_t cands, void* d_rows_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!count)
        return;
    std::string err = prepare_search(ix, 1);
    if (!err.empty())
        RB_FAIL(err);
    if (first + count > ix->host.size())
        RB_FAIL("rows out of range");
    if (cands + ix->host.m0 > (uint32_t)rb::kStitchMax)
        RB_FAIL("too many remote candidates per node (own links + candidates must fit 128)");
    if (slot_base + ix->host.size() > (1ull << 32))
        RB_FAIL("global slots must fit 32 bits");
    DeviceState& d = ix->dev;
    rb::StitchParams sp;
    sp.g = graph_view(ix);
    sp.first = (uint32_t)first;
    sp.count = (uint32_t)count;
    sp.slot_base = (uint32_t)slot_base;
    sp.cand_slots = static_cast<uint32_t const*>(d_cand_slots);
    sp.cand_fp = static_cast<uint4 const*>(d_cand_vectors);
    sp.cands = cands;
    sp.rows_out = static_cast<uint32_t*>(d_rows_out);
    sp.counters = nullptr;
    const unsigned grid = (unsigned)std::min<uint64_t>((count + rb::kWarpsPerBlock - 1) / rb::kWarpsPerBlock,
                                                       (uint64_t)d.sm_count * 8);
    rb::stitch_kernel<<<grid, rb::kWarpsPerBlock * 32, 0, d.stream>>>(sp);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(d.stream);
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in stitch: ") + cudaGetErrorString(e));
}

void rb_import_graph(rb_index_t* ix, uint64_t const* keys, int16_t const* levels, uint32_t const* adj0,
                     uint32_t const* adj_upper, uint8_t const* vectors, size_t n, uint64_t entry, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (ix->host.viewing)
        RB_FAIL("Can't import into an immutable index");
    std::string err = ix->host.import_graph(keys, levels, adj0, adj_upper, vectors, n, entry);
    if (!err.empty())
        RB_FAIL(err);
    // the device copy (if any) belongs to the old graph: start from fresh arrays at the next device call
    if (ix->dev.device >= 0) {
        int dev = ix->dev.device;
        ix->dev.release();
        ix->opts.device = dev;
    }
    ix->device_graph_valid = false;
    ix->host_graph_valid = true;
    ix->last = rb_counters_t{};
}

void rb_extract_level(rb_index_t* src, rb_index_t* dst, rb_error_t* error) {
    if (error)
        *error = nullptr;
    rbh::HostIndex& hs = src->host;
    rbh::HostIndex& hd = dst->host;
    if (hd.viewing || hd.size())
        RB_FAIL("rb_extract_level fills an empty index");
    if (hs.viewing)
        RB_FAIL("rb_extract_level reads a built or loaded index, not a view");
    if (hd.m != hs.m || hd.m0 != hs.m || hd.ndim != hs.ndim)
        RB_FAIL("the destination needs the source's connectivity as connectivity AND connectivity_base, and its dimensions");
    std::string err = prepare_search(src, 1);
    if (!err.empty())
        RB_FAIL(err);
    dst->opts.device = src->dev.device;
    if (!(err = ensure_device(dst)).empty())
        RB_FAIL(err);
    const size_t n = hs.keys.size();
    std::vector<uint32_t> slots, table(n, RB_EMPTY);
    for (size_t i = 0; i < n; ++i)
        if (hs.levels[i] >= 1) {
            table[i] = (uint32_t)slots.size();
            slots.push_back((uint32_t)i);
        }
    const size_t cnt = slots.size();
    if (!cnt)
        return;
    std::vector<int16_t> lv(cnt);
    std::vector<uint64_t> keys(cnt);
    std::vector<uint32_t> off32(cnt);
    std::vector<int8_t> lv8(cnt);
    uint64_t upper = 0;
    for (size_t i = 0; i < cnt; ++i) {
        lv[i] = (int16_t)(hs.levels[slots[i]] - 1);
        lv8[i] = (int8_t)lv[i];
        keys[i] = i;
        off32[i] = (uint32_t)upper;
        upper += (uint64_t)lv[i];
    }
    if (!(err = ensure_capacity(dst, cnt, upper)).empty())
        RB_FAIL(err);
    DeviceState& d = dst->dev;
    uint32_t *d_slots = nullptr, *d_table = nullptr;
    cudaError_t e = cudaMalloc(&d_slots, cnt * 4);
    if (e == cudaSuccess)
        e = cudaMalloc(&d_table, n * 4);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_slots, slots.data(), cnt * 4, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_table, table.data(), n * 4, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d.upper_off, off32.data(), cnt * 4, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d.levels, lv8.data(), cnt, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess) {
        rb::ExtractParams ep;
        ep.src = graph_view(src);
        ep.dst = graph_view(dst);
        ep.dst_fp = d.fp;
        ep.src_slots = d_slots;
        ep.table = d_table;
        ep.dst_levels = d.levels;
        ep.count = (uint32_t)cnt;
        const unsigned grid = (unsigned)std::min<size_t>((cnt * 32 + 255) / 256, (size_t)d.sm_count * 16);
        rb::extract_level_kernel<<<grid, 256, 0, d.stream>>>(ep);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(d.stream);
    cudaFree(d_slots);
    cudaFree(d_table);
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in extract_level: ") + cudaGetErrorString(e));
    // host bookkeeping of the new index (the adjacency stays on the device until rb_sync_host)
    hd.keys = keys;
    hd.levels = lv;
    hd.upper_off.assign(cnt + 1, 0);
    for (size_t i = 0; i < cnt; ++i)
        hd.upper_off[i + 1] = hd.upper_off[i] + (uint64_t)lv[i];
    hd.max_level = hs.max_level - 1;
    hd.entry = table[hs.entry];
    hd.note_new_keys(keys.data(), cnt);
    hd.note_new_levels(0, lv.data(), cnt);
    hd.invalidate_level_edges();
    d.rows = cnt;
    d.upper_rows = upper;
    dst->device_graph_valid = true;
    dst->host_graph_valid = false;
    dst->last = rb_counters_t{};
}

void rb_clear(rb_index_t* ix, rb_error_t* error) {
    if (error)
        *error = nullptr;
    const size_t capacity = ix->host.capacity;
    ix->host.reset_graph();
    ix->host.viewing = false;
    ix->host.capacity = capacity;
    ix->dev.rows = ix->dev.upper_rows = 0;
    ix->device_graph_valid = ix->dev.device >= 0;  // an empty graph; the arrays keep their capacity
    ix->host_graph_valid = true;
    ix->last = rb_counters_t{};
}

uintptr_t rb_stream(rb_index_t const* ix) { return reinterpret_cast<uintptr_t>(ix->dev.stream); }

void rb_sync_host(rb_index_t* ix, rb_error_t* error) {
    if (error)
        *error = nullptr;
    std::string err = sync_host(ix);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_search(rb_index_t* ix, uint8_t const* queries, size_t nq, size_t stride, size_t k, int exact,
               uint64_t* keys_out, float* distances_out, int64_t* counts_out, uint64_t* visited_members,
               uint64_t* computed_distances, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (visited_members)
        *visited_members = 0;
    if (computed_distances)
        *computed_distances = 0;
    if (!nq)
        return;
    rbh::HostIndex& h = ix->host;
    if (stride < h.vec_bytes)
        RB_FAIL("The number of vector dimensions doesn't match!");
    if (!h.size()) {
        for (size_t q = 0; q < nq; ++q)
            counts_out[q] = 0;
        return;
    }
    std::string err = prepare_search(ix, k);
    if (!err.empty())
        RB_FAIL(err);
    DeviceState& d = ix->dev;
    // arena: [queries | slots | dists | counts | exact workspace]
    size_t q_bytes = (nq * 128 + 255) & ~(size_t)255;
    size_t s_bytes = (nq * k * 4 + 255) & ~(size_t)255;
    size_t c_bytes = (nq * 4 + 255) & ~(size_t)255;
    size_t x_bytes = exact ? exact_scratch_bytes(h.size(), nq, k) : 0;
    if (!(err = ensure_scratch(d, q_bytes + 2 * s_bytes + c_bytes + x_bytes)).empty())
        RB_FAIL(err);
    uint8_t* base = static_cast<uint8_t*>(d.scratch);
    uint4* dq = reinterpret_cast<uint4*>(base);
    uint32_t* ds = reinterpret_cast<uint32_t*>(base + q_bytes);
    float* dd = reinterpret_cast<float*>(base + q_bytes + s_bytes);
    int32_t* dc = reinterpret_cast<int32_t*>(base + q_bytes + 2 * s_bytes);
    void* dx = base + q_bytes + 2 * s_bytes + c_bytes;
    cudaError_t e;
    if (h.vec_bytes == 128 && stride == 128) {
        e = cudaMemcpyAsync(dq, queries, nq * 128, cudaMemcpyHostToDevice, d.stream);
    } else {
        e = cudaMemsetAsync(dq, 0, nq * 128, d.stream);
        if (e == cudaSuccess)
            e = cudaMemcpy2DAsync(dq, 128, queries, stride, h.vec_bytes, nq, cudaMemcpyHostToDevice, d.stream);
    }
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error while uploading queries: ") + cudaGetErrorString(e));
    if (!(err = search_on_device(ix, dq, nq, k, exact != 0, ds, dd, dc, dx, x_bytes)).empty())
        RB_FAIL(err);
    std::vector<uint32_t> slots(nq * k);
    std::vector<float> dists(nq * k);
    std::vector<int32_t> counts(nq);
    e = cudaMemcpy(slots.data(), ds, nq * k * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess)
        e = cudaMemcpy(dists.data(), dd, nq * k * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess)
        e = cudaMemcpy(counts.data(), dc, nq * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error while fetching results: ") + cudaGetErrorString(e));
    for (size_t q = 0; q < nq; ++q) {
        int32_t c = counts[q];
        counts_out[q] = c;
        for (int32_t j = 0; j < c; ++j) {
            keys_out[q * k + j] = h.key_of(slots[q * k + j]);
            distances_out[q * k + j] = dists[q * k + j];
        }
    }
    if (visited_members)
        *visited_members = ix->last.visited_members;
    if (computed_distances)
        *computed_distances = ix->last.computed_distances;
}

void rb_search_device(rb_index_t* ix, void const* d_queries, size_t nq, size_t k, int exact, void* d_slots_out,
                      void* d_distances_out, void* d_counts_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!nq)
        return;
    std::string err = prepare_search(ix, k);
    if (!err.empty())
        RB_FAIL(err);
    if (!ix->host.size())
        RB_FAIL("The index is empty");
    size_t x_bytes = exact ? exact_scratch_bytes(ix->host.size(), nq, k) : 0;
    if (!(err = ensure_scratch(ix->dev, x_bytes)).empty())
        RB_FAIL(err);
    err = search_on_device(ix, static_cast<uint4 const*>(d_queries), nq, k, exact != 0,
                           static_cast<uint32_t*>(d_slots_out), static_cast<float*>(d_distances_out),
                           static_cast<int32_t*>(d_counts_out), ix->dev.scratch, x_bytes);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_cluster_device(rb_index_t* ix, void const* d_queries, size_t nq, int level, void* d_slots_out,
                       void* d_distances_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!nq)
        return;
    if (!ix->host.size())
        RB_FAIL("No clusters to identify");  // index.hpp:3056-3057
    std::string err = prepare_search(ix, 1);
    if (!err.empty())
        RB_FAIL(err);
    if (level < 0 || level > ix->host.max_level)
        RB_FAIL("level outside 0..max_level");
    err = [&]() -> std::string {
        DeviceState& d = ix->dev;
        RB_CUDA(cudaMemsetAsync(d.counters, 0, 4 * sizeof(unsigned long long), d.stream));
        RB_CUDA(cudaEventRecord(d.ev_begin, d.stream));
        rb::ClusterParams cp;
        cp.g = graph_view(ix);
        cp.queries = static_cast<uint4 const*>(d_queries);
        cp.query_count = (uint32_t)nq;
        cp.entry = (uint32_t)ix->host.entry;
        cp.max_level = ix->host.max_level;
        cp.level = level;
        cp.slots_out = static_cast<uint32_t*>(d_slots_out);
        cp.dists_out = static_cast<float*>(d_distances_out);
        cp.counters = d.counters;
        unsigned blocks = (unsigned)std::min<size_t>((nq + 7) / 8, (size_t)d.sm_count * 8);
        rb::cluster_kernel<<<blocks, 256, 0, d.stream>>>(cp);
        RB_CUDA(cudaGetLastError());
        RB_CUDA(cudaEventRecord(d.ev_end, d.stream));
        RB_CUDA(cudaStreamSynchronize(d.stream));
        float ms = 0;
        RB_CUDA(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
        unsigned long long counters[4];
        RB_CUDA(cudaMemcpy(counters, d.counters, sizeof(counters), cudaMemcpyDeviceToHost));
        ix->last.computed_distances = counters[0];
        ix->last.visited_members = counters[1];
        ix->last.batches = 0;
        ix->last.kernel_launches = 1;
        ix->last.device_ms = ms;
        return {};
    }();
    if (!err.empty())
        RB_FAIL(err);
}

size_t rb_exact_topk_workspace_bytes(uint64_t rows, uint64_t query_count, uint32_t k) {
    ExactPlan pl = plan_exact(rows, query_count, k);
    return pl.keys_bytes + pl.iu_bytes;
}

void rb_exact_topk_device(void const* d_database, uint64_t rows, uint64_t slot_base, void const* d_queries,
                          uint64_t query_count, uint32_t k, void* d_keys_out, void* d_workspace,
                          size_t workspace_bytes, uintptr_t stream, rb_error_t* error) {
    if (error)
        *error = nullptr;
    std::string err = launch_exact(static_cast<uint4 const*>(d_database), rows, slot_base,
                                   static_cast<uint4 const*>(d_queries), query_count, k,
                                   static_cast<uint64_t*>(d_keys_out), d_workspace, workspace_bytes,
                                   reinterpret_cast<cudaStream_t>(stream));
    if (!err.empty())
        RB_FAIL(err);
}

}  // extern "C"

// ---- rb_exact_topk_device_noalloc: the workspace of the exact search in a STATIC device array (see include/rad_b200.h).
// Every GPU that loads the library holds these RB_NOALLOC_SCRATCH_BYTES until the process exits.
static __device__ __align__(256) unsigned char g_noalloc_scratch[RB_NOALLOC_SCRATCH_BYTES];

namespace {

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current, and cached.
std::string noalloc_scratch_address(void*& out) {
    static std::mutex mu;
    static void* address[64] = {nullptr};
    int dev = 0;
    RB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64)
        return "rb_exact_topk_device_noalloc supports device ordinals below 64";
    std::lock_guard<std::mutex> lock(mu);
    if (!address[dev]) {
        void* ptr = nullptr;
        RB_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        address[dev] = ptr;
    }
    out = address[dev];
    return {};
}

}  // namespace

extern "C" {

size_t rb_exact_topk_noalloc_max_bytes(void) { return RB_NOALLOC_SCRATCH_BYTES; }

void rb_exact_topk_device_noalloc(void const* d_database, uint64_t rows, uint64_t slot_base, void const* d_queries,
                                  uint64_t query_count, uint32_t k, void* d_keys_out, uintptr_t stream,
                                  rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!query_count)
        return;
    if (k == 0 || k > 1024)
        RB_FAIL("k must be in 1..1024");
    const ExactPlan pl = plan_exact(rows, query_count, k);
    const size_t need = pl.keys_bytes + pl.iu_bytes;
    if (need > (size_t)RB_NOALLOC_SCRATCH_BYTES)
        RB_FAIL(format("rb_exact_topk_device_noalloc: this search needs %zu bytes of scratch, the static array holds %zu "
                       "(fewer queries per call or a smaller k; or rb_exact_topk_device with a caller workspace)",
                       need, (size_t)RB_NOALLOC_SCRATCH_BYTES));
    void* scratch = nullptr;
    std::string err = noalloc_scratch_address(scratch);
    // the kernels initialise every list of the workspace they use and launch_exact resets the bound words, so the
    // contents the previous call left behind are never read
    if (err.empty())
        err = launch_exact(static_cast<uint4 const*>(d_database), rows, slot_base, static_cast<uint4 const*>(d_queries),
                           query_count, k, static_cast<uint64_t*>(d_keys_out), scratch, need,
                           reinterpret_cast<cudaStream_t>(stream), /*allow_debug_counters=*/false);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_merge_topk_device(void const* d_partial_keys, uint32_t parts, uint64_t query_count, uint32_t k,
                          void* d_keys_out, uintptr_t stream, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!query_count)
        return;
    rb::MergeParams mp;
    mp.partial_keys = static_cast<uint64_t const*>(d_partial_keys);
    mp.parts = parts;
    mp.query_count = (uint32_t)query_count;
    mp.k = k;
    mp.keys_out = static_cast<uint64_t*>(d_keys_out);
    rb::merge_topk_kernel<<<(unsigned)((query_count + 3) / 4), 128, 4 * (size_t)k * 8,
                            reinterpret_cast<cudaStream_t>(stream)>>>(mp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in merge: ") + cudaGetErrorString(e));
}

// ---------------------------------------------------------------------------------------------- sharded search over NCCL
// The exchange of a multi-GPU search behind the C ABI (no host framework needed): NCCL is reached through dlopen at the
// first call -- the copy already loaded in the process if there is one (e.g. PyTorch's), else libnccl.so.2 from the
// loader's path -- so the library has no link-time dependency on it and single-GPU users never load it.
namespace {
struct NcclApi {
    // ncclResult_t ncclAllGather(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) etc. (nccl.h)
    int (*all_gather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*comm_count)(void*, int*) = nullptr;
    const char* (*error_string)(int) = nullptr;
    std::string problem;
};
constexpr int kNcclUint64 = 5;  // ncclDataType_t: ncclUint64 (nccl.h)

NcclApi const& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h)
            h = dlopen("libnccl.so.2", RTLD_NOW);
        if (!h) {
            a.problem = std::string("libnccl.so.2 cannot be loaded: ") + dlerror();
            return a;
        }
        a.all_gather = reinterpret_cast<decltype(a.all_gather)>(dlsym(h, "ncclAllGather"));
        a.comm_count = reinterpret_cast<decltype(a.comm_count)>(dlsym(h, "ncclCommCount"));
        a.error_string = reinterpret_cast<decltype(a.error_string)>(dlsym(h, "ncclGetErrorString"));
        if (!a.all_gather || !a.comm_count || !a.error_string)
            a.problem = "libnccl.so.2 lacks ncclAllGather / ncclCommCount / ncclGetErrorString";
        return a;
    }();
    return api;
}

struct ShardedSearchLayout {
    size_t slots, dists, counts, local_keys, gathered, total;
};
ShardedSearchLayout sharded_search_layout(size_t nq, size_t k, uint32_t world) {
    ShardedSearchLayout L{};
    auto take = [&](size_t bytes) {
        size_t at = L.total;
        L.total += (bytes + 255) & ~(size_t)255;
        return at;
    };
    L.slots = take(nq * k * 4), L.dists = take(nq * k * 4), L.counts = take(nq * 4);
    L.local_keys = take(nq * k * 8), L.gathered = take((size_t)world * nq * k * 8);
    return L;
}

// (slot, distance, count) answers of one shard -> merge keys in GLOBAL slots (distance bits << 32 | ~slot; ~0 = none)
__global__ void pack_topk_kernel(const uint32_t* slots, const float* dists, const int32_t* counts, uint64_t query_count,
                                 uint32_t k, uint64_t slot_base, uint64_t* keys) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= query_count * k)
        return;
    const uint64_t q = i / k;
    const uint32_t j = (uint32_t)(i % k);
    keys[i] = (int32_t)j < counts[q]
                  ? ((uint64_t)__float_as_uint(dists[i]) << 32) | (uint32_t)~(uint32_t)(slot_base + slots[i])
                  : ~0ull;
}

std::string sharded_search(rb_index* ix, uint64_t slot_base, void const* d_queries, size_t nq, size_t k, bool exact,
                           void* comm, uint64_t* d_keys_out, void* d_workspace, size_t workspace_bytes) {
    NcclApi const& nccl = nccl_api();
    if (!nccl.problem.empty())
        return nccl.problem;
    int world = 0;
    int rc = nccl.comm_count(comm, &world);
    if (rc != 0)
        return std::string("ncclCommCount: ") + nccl.error_string(rc);
    if (slot_base + ix->host.size() > (1ull << 32))
        return "sharded search addresses at most 2^32 global slots";
    const ShardedSearchLayout L = sharded_search_layout(nq, k, (uint32_t)world);
    if (workspace_bytes < L.total)
        return "workspace smaller than rb_sharded_search_workspace_bytes(query_count, k, world)";
    std::string err = ensure_device(ix);
    if (!err.empty())
        return err;
    DeviceState& d = ix->dev;
    uint8_t* ws = static_cast<uint8_t*>(d_workspace);
    uint32_t* slots = reinterpret_cast<uint32_t*>(ws + L.slots);
    float* dists = reinterpret_cast<float*>(ws + L.dists);
    int32_t* counts = reinterpret_cast<int32_t*>(ws + L.counts);
    uint64_t* local_keys = reinterpret_cast<uint64_t*>(ws + L.local_keys);
    uint64_t* gathered = reinterpret_cast<uint64_t*>(ws + L.gathered);
    if (ix->host.size()) {  // an empty shard answers with no candidates
        if (!(err = prepare_search(ix, k)).empty())
            return err;
        size_t x_bytes = exact ? exact_scratch_bytes(ix->host.size(), nq, k) : 0;
        if (!(err = ensure_scratch(d, x_bytes)).empty())
            return err;
        err = search_on_device(ix, static_cast<uint4 const*>(d_queries), nq, k, exact, slots, dists, counts, d.scratch,
                               x_bytes);
        if (!err.empty())
            return err;
    } else {
        RB_CUDA(cudaMemsetAsync(counts, 0, nq * 4, d.stream));
    }
    pack_topk_kernel<<<(unsigned)((nq * k + 255) / 256), 256, 0, d.stream>>>(slots, dists, counts, nq, (uint32_t)k,
                                                                             slot_base, local_keys);
    RB_CUDA(cudaGetLastError());
    rc = nccl.all_gather(local_keys, gathered, nq * k, kNcclUint64, comm, d.stream);
    if (rc != 0)
        return std::string("ncclAllGather: ") + nccl.error_string(rc);
    rb::MergeParams mp;
    mp.partial_keys = gathered;
    mp.parts = (uint32_t)world;
    mp.query_count = (uint32_t)nq;
    mp.k = (uint32_t)k;
    mp.keys_out = d_keys_out;
    rb::merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 4 * k * 8, d.stream>>>(mp);
    RB_CUDA(cudaGetLastError());
    RB_CUDA(cudaStreamSynchronize(d.stream));
    ix->last.kernel_launches += 2;
    return {};
}
}  // namespace

size_t rb_sharded_search_workspace_bytes(size_t query_count, size_t k, uint32_t world) {
    return sharded_search_layout(query_count, k, world ? world : 1).total;
}

void rb_sharded_search_device(rb_index_t* ix, uint64_t slot_base, void const* d_queries, size_t query_count, size_t k,
                              int exact, uintptr_t nccl_comm, void* d_keys_out, void* d_workspace,
                              size_t workspace_bytes, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!query_count)
        return;
    if (!k || k > 1024)
        RB_FAIL("k must be in 1..1024");
    std::string err = sharded_search(ix, slot_base, d_queries, query_count, k, exact != 0,
                                     reinterpret_cast<void*>(nccl_comm), static_cast<uint64_t*>(d_keys_out),
                                     d_workspace, workspace_bytes);
    if (!err.empty())
        RB_FAIL(err);
}

// ---------------------------------------------------------------------------------------------- ingest: duplicates
namespace {
struct DedupLayout {
    size_t hashes[2], ids[2], start, tmp, tmp_bytes, total;
};
DedupLayout dedup_layout(uint64_t rows) {
    DedupLayout L{};
    auto take = [&](size_t bytes) {
        size_t at = L.total;
        L.total += (bytes + 255) & ~(size_t)255;
        return at;
    };
    L.hashes[0] = take(rows * 8), L.hashes[1] = take(rows * 8);
    L.ids[0] = take(rows * 4), L.ids[1] = take(rows * 4);
    L.start = take(rows * 4);
    size_t sort_tmp = 0, scan_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (int64_t)rows);
    cub::DeviceScan::InclusiveScan(nullptr, scan_tmp, (uint32_t*)nullptr, (uint32_t*)nullptr, cub::Max(), (int64_t)rows);
    L.tmp_bytes = std::max(sort_tmp, scan_tmp);
    L.tmp = take(L.tmp_bytes);
    return L;
}
}  // namespace

size_t rb_dedup_workspace_bytes(uint64_t rows) { return rows ? dedup_layout(rows).total : 0; }

void rb_dedup_device(void const* d_vectors, uint64_t rows, void* d_representative, void* d_workspace,
                     size_t workspace_bytes, uintptr_t stream_, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!rows)
        return;
    if (rows >= (1ull << 32))
        RB_FAIL("rb_dedup_device addresses at most 2^32 - 1 rows");
    const DedupLayout L = dedup_layout(rows);
    if (workspace_bytes < L.total)
        RB_FAIL("workspace smaller than rb_dedup_workspace_bytes(rows)");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    uint8_t* ws = static_cast<uint8_t*>(d_workspace);
    uint64_t* h0 = reinterpret_cast<uint64_t*>(ws + L.hashes[0]);
    uint64_t* h1 = reinterpret_cast<uint64_t*>(ws + L.hashes[1]);
    uint32_t* i0 = reinterpret_cast<uint32_t*>(ws + L.ids[0]);
    uint32_t* i1 = reinterpret_cast<uint32_t*>(ws + L.ids[1]);
    uint32_t* start = reinterpret_cast<uint32_t*>(ws + L.start);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid8 = (unsigned)std::min<uint64_t>((rows * 8 + 255) / 256, (uint64_t)sms * 16);
    const unsigned grid1 = (unsigned)std::min<uint64_t>((rows + 255) / 256, (uint64_t)sms * 16);
    const uint4* fp = static_cast<const uint4*>(d_vectors);
    rb::row_hash_kernel<<<grid8, 256, 0, stream>>>(fp, rows, h0, i0);
    size_t tmp = L.tmp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(ws + L.tmp, tmp, h0, h1, i0, i1, (int64_t)rows, 0, 64, stream);
    if (e == cudaSuccess) {
        rb::run_start_kernel<<<grid1, 256, 0, stream>>>(h1, rows, start);
        tmp = L.tmp_bytes;
        e = cub::DeviceScan::InclusiveScan(ws + L.tmp, tmp, start, start, cub::Max(), (int64_t)rows, stream);
    }
    if (e == cudaSuccess) {
        rb::representative_kernel<<<grid8, 256, 0, stream>>>(fp, rows, i1, start, static_cast<uint32_t*>(d_representative));
        e = cudaGetLastError();
    }
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in dedup: ") + cudaGetErrorString(e));
}

// ---------------------------------------------------------------------------------------------- join: stable matching
namespace {
struct JoinLayout {
    size_t made, husband_dist, ring[2], next, keys[2], values[2], counters, tmp, tmp_bytes, total;
};
JoinLayout join_layout(uint64_t men, uint64_t women) {
    JoinLayout L{};
    auto take = [&](size_t bytes) {
        size_t at = L.total;
        L.total += (bytes + 255) & ~(size_t)255;
        return at;
    };
    L.made = take(men * 4), L.husband_dist = take(women * 4);
    L.ring[0] = take(men * 4), L.ring[1] = take(men * 4), L.next = take(men * 4);
    L.keys[0] = take(men * 8), L.keys[1] = take(men * 8);
    L.values[0] = take(men * 4), L.values[1] = take(men * 4);
    L.counters = take(16);
    size_t sort_tmp = 0, select_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (int64_t)men);
    cub::DeviceSelect::If(nullptr, select_tmp, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)men,
                          rb::JoinIsMan());
    L.tmp_bytes = std::max(sort_tmp, select_tmp);
    L.tmp = take(L.tmp_bytes);
    return L;
}
}  // namespace

size_t rb_stable_match_workspace_bytes(uint64_t men, uint64_t women) { return men ? join_layout(men, women).total : 0; }

namespace {
std::string stable_match(void const* d_candidate_slots, void const* d_candidate_distances,
                         void const* d_candidate_counts, uint64_t men, uint64_t women, uint32_t proposals,
                         void* d_man_to_woman, void* d_woman_to_man, void* d_workspace, size_t workspace_bytes,
                         cudaStream_t stream, uint64_t* generations_out, uint64_t* engagements_out) {
    if (men >= (1ull << 32) - 1 || women >= (1ull << 32) - 1)
        return "rb_stable_match_device addresses at most 2^32 - 2 rows on either side";
    if (men)
        RB_CUDA(cudaMemsetAsync(d_man_to_woman, 0xFF, men * 4, stream));
    if (women)
        RB_CUDA(cudaMemsetAsync(d_woman_to_man, 0xFF, women * 4, stream));
    if (!men || !women || !proposals)
        return {};
    const JoinLayout L = join_layout(men, women);
    if (workspace_bytes < L.total)
        return "workspace smaller than rb_stable_match_workspace_bytes(men, women)";
    uint8_t* ws = static_cast<uint8_t*>(d_workspace);
    uint32_t* ring[2] = {reinterpret_cast<uint32_t*>(ws + L.ring[0]), reinterpret_cast<uint32_t*>(ws + L.ring[1])};
    uint64_t* keys[2] = {reinterpret_cast<uint64_t*>(ws + L.keys[0]), reinterpret_cast<uint64_t*>(ws + L.keys[1])};
    uint32_t* values[2] = {reinterpret_cast<uint32_t*>(ws + L.values[0]), reinterpret_cast<uint32_t*>(ws + L.values[1])};
    uint32_t* selected = reinterpret_cast<uint32_t*>(ws + L.counters);
    rb::JoinParams p{};
    p.cand_slots = static_cast<const uint32_t*>(d_candidate_slots);
    p.cand_dist = static_cast<const uint32_t*>(d_candidate_distances);
    p.cand_counts = static_cast<const int32_t*>(d_candidate_counts);
    p.proposals = proposals;
    p.made = reinterpret_cast<uint32_t*>(ws + L.made);
    p.man_to_woman = static_cast<uint32_t*>(d_man_to_woman);
    p.woman_to_man = static_cast<uint32_t*>(d_woman_to_man);
    p.husband_dist = reinterpret_cast<uint32_t*>(ws + L.husband_dist);
    p.next = reinterpret_cast<uint32_t*>(ws + L.next);
    p.engagements = reinterpret_cast<unsigned long long*>(ws + L.counters + 8);
    RB_CUDA(cudaMemsetAsync(p.made, 0, men * 4, stream));
    RB_CUDA(cudaMemsetAsync(ws + L.counters, 0, 16, stream));
    rb::join_init_kernel<<<(unsigned)((men + 255) / 256), 256, 0, stream>>>(ring[0], (uint32_t)men);
    uint32_t length = (uint32_t)men;
    uint64_t generations = 0;
    for (int cur = 0; length; cur ^= 1, ++generations) {
        p.ring = ring[cur];
        p.length = length;
        p.keys = keys[0], p.values = values[0];
        const unsigned grid = (length + 255) / 256;
        rb::join_propose_kernel<<<grid, 256, 0, stream>>>(p);
        size_t tmp = L.tmp_bytes;
        cudaError_t e = cub::DeviceRadixSort::SortPairs(ws + L.tmp, tmp, keys[0], keys[1], values[0], values[1],
                                                        (int64_t)length, 0, 64, stream);
        if (e == cudaSuccess) {
            p.keys = keys[1], p.values = values[1];
            rb::join_settle_kernel<<<grid, 256, 0, stream>>>(p);
            tmp = L.tmp_bytes;
            e = cub::DeviceSelect::If(ws + L.tmp, tmp, p.next, ring[cur ^ 1], selected, (int64_t)length, rb::JoinIsMan(),
                                      stream);
        }
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(&length, selected, 4, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess)
            return std::string("CUDA error in the matching: ") + cudaGetErrorString(e);
    }
    unsigned long long engaged = 0;
    RB_CUDA(cudaMemcpyAsync(&engaged, p.engagements, 8, cudaMemcpyDeviceToHost, stream));
    RB_CUDA(cudaStreamSynchronize(stream));
    if (generations_out)
        *generations_out = generations;
    if (engagements_out)
        *engagements_out = engaged;
    return {};
}
}  // namespace

void rb_stable_match_device(void const* d_candidate_slots, void const* d_candidate_distances,
                            void const* d_candidate_counts, uint64_t men, uint64_t women, uint32_t proposals,
                            void* d_man_to_woman, void* d_woman_to_man, void* d_workspace, size_t workspace_bytes,
                            uintptr_t stream, uint64_t* generations_out, uint64_t* engagements_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (generations_out)
        *generations_out = 0;
    if (engagements_out)
        *engagements_out = 0;
    std::string msg = stable_match(d_candidate_slots, d_candidate_distances, d_candidate_counts, men, women, proposals,
                                   d_man_to_woman, d_woman_to_man, d_workspace, workspace_bytes,
                                   reinterpret_cast<cudaStream_t>(stream), generations_out, engagements_out);
    if (!msg.empty())
        RB_FAIL(msg);
}

void rb_popcounts_device(void const* d_vectors, uint64_t rows, void* d_popcounts, uintptr_t stream_, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!rows)
        return;
    if (rows >= (1ull << 32))
        RB_FAIL("rb_popcounts_device addresses at most 2^32 - 1 rows");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)std::min<uint64_t>((rows * 8 + 255) / 256, (uint64_t)sms * 16);
    rb::node_popcount_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(
        static_cast<const uint4*>(d_vectors), (uint32_t)rows, static_cast<uint16_t*>(d_popcounts));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in popcounts: ") + cudaGetErrorString(e));
}

void rb_unpack_topk_device(void const* d_keys, uint64_t query_count, uint32_t k, void* d_slots_out,
                           void* d_distances_out, void* d_counts_out, uintptr_t stream, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!query_count)
        return;
    rb::unpack_topk_kernel<<<(unsigned)((query_count + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        static_cast<uint64_t const*>(d_keys), query_count, k, static_cast<uint64_t*>(d_slots_out),
        static_cast<float*>(d_distances_out), static_cast<int64_t*>(d_counts_out));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in unpack: ") + cudaGetErrorString(e));
}

void rb_synth_fingerprints_device(void* d_out, uint64_t first_row, uint64_t rows, uint64_t seed, void const* d_centres,
                                  uint32_t centre_count, int edge_cases, uintptr_t stream, rb_error_t* error) {
    rb_synth_fingerprints_device_ex(d_out, first_row, rows, seed, d_centres, centre_count, edge_cases, 0, 0, 0, 0, stream,
                                    error);
}

void rb_synth_fingerprints_device_ex(void* d_out, uint64_t first_row, uint64_t rows, uint64_t seed, void const* d_centres,
                                     uint32_t centre_count, int edge_cases, uint64_t rows_per_centre, uint64_t dup_fold,
                                     uint64_t dup_start, uint64_t dup_rows, uintptr_t stream, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!rows)
        return;
    if (!centre_count)
        RB_FAIL("the centre table is empty");
    uint64_t threads = rows * 32;
    rb::synth_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        static_cast<uint32_t*>(d_out), first_row, rows, seed, static_cast<uint32_t const*>(d_centres), centre_count,
        edge_cases, rb::SynthModes{rows_per_centre, dup_fold, dup_start, dup_rows});
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        RB_FAIL(std::string("CUDA error in synth: ") + cudaGetErrorString(e));
}

}  // extern "C"

// ================================================================================================ C ABI, host side
namespace {

// fingerprints of the whole index as a host matrix (size x vec_bytes); empty when they live nowhere
std::string fetch_vectors(rb_index* ix, std::vector<uint8_t>& out) {
    rbh::HostIndex& h = ix->host;
    const size_t n = h.size();
    out.clear();
    if (!n)
        return {};
    if (!h.host_vectors.empty()) {
        out = h.host_vectors;
        return {};
    }
    if (ix->device_graph_valid && ix->dev.device >= 0) {
        RB_CUDA(cudaSetDevice(ix->dev.device));
        out.resize(n * (size_t)h.vec_bytes);
        RB_CUDA(cudaMemcpy2D(out.data(), h.vec_bytes, ix->dev.fp, 128, h.vec_bytes, n, cudaMemcpyDeviceToHost));
        return {};
    }
    return "The fingerprints are not available";
}

}  // namespace

extern "C" {

void rb_save(rb_index_t* ix, char const* path, int exclude_vectors, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (ix->host.m0 != 2 * ix->host.m)
        RB_FAIL("an index with a non-standard connectivity_base cannot be saved (the reader expects 2 x connectivity)");
    std::string err = sync_host(ix);
    if (!err.empty())
        RB_FAIL(err);
    std::vector<uint8_t> vectors;
    uint8_t const* vptr = nullptr;
    if (!exclude_vectors && !ix->host.viewing) {
        if (!(err = fetch_vectors(ix, vectors)).empty())
            RB_FAIL(err);
        vptr = vectors.data();
    }
    err = ix->host.save(path, vptr, exclude_vectors != 0);
    if (!err.empty())
        RB_FAIL(err);
}

void rb_load(rb_index_t* ix, char const* path, int exclude_vectors, rb_error_t* error) {
    if (error)
        *error = nullptr;
    const uint32_t ef_add = ix->host.ef_add, ef_search = ix->host.ef_search;
    std::string err = ix->host.load(path, exclude_vectors != 0);
    ix->host.ef_add = ef_add;
    ix->host.ef_search = ef_search;
    if (!err.empty())
        RB_FAIL(err);  // a file that does not parse leaves the index -- host and device -- as it was
    ix->device_graph_valid = false;
    ix->host_graph_valid = true;
    ix->dev.rows = 0;
    ix->dev.insert_cfg_ef = ix->dev.search_cfg_ef = 0;  // connectivity may have changed
    if (ix->dev.device >= 0 && (ix->dev.cap_rows || ix->dev.cap_upper)) {
        // row widths depend on the connectivity stored in the file: start from fresh device arrays
        int dev = ix->dev.device;
        ix->dev.release();
        ix->opts.device = dev;
    }
}

void rb_view(rb_index_t* ix, char const* path, int exclude_vectors, rb_error_t* error) {
    if (error)
        *error = nullptr;
    const uint32_t ef_add = ix->host.ef_add, ef_search = ix->host.ef_search;
    std::string err = ix->host.view(path, exclude_vectors != 0);
    ix->host.ef_add = ef_add;
    ix->host.ef_search = ef_search;
    if (!err.empty())
        RB_FAIL(err);  // a file that does not parse leaves the index as it was
    ix->device_graph_valid = false;
    ix->host_graph_valid = true;
}

void rb_metadata(char const* path, rb_metadata_t* out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    rbh::FileMetadata md;
    std::string err = rbh::HostIndex::metadata(path, md);
    if (!err.empty())
        RB_FAIL(err);
    out->version_major = md.version_major;
    out->version_minor = md.version_minor;
    out->version_patch = md.version_patch;
    out->kind_metric = md.kind_metric;
    out->kind_scalar = md.kind_scalar;
    out->kind_key = md.kind_key;
    out->kind_compressed_slot = md.kind_compressed_slot;
    out->count_present = md.count_present;
    out->count_deleted = md.count_deleted;
    out->dimensions = md.dimensions;
    out->multi = md.multi;
    out->has_vectors = md.has_vectors;
    out->matrix_rows = md.matrix_rows;
    out->matrix_cols = md.matrix_cols;
}

size_t rb_neighbors(rb_index_t const* ix, size_t slot, int level, uint64_t* out_pairs, size_t capacity,
                    rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!ix->host_graph_valid) {
        set_error(error, "the host copy of the graph is stale: call rb_sync_host first");
        return 0;
    }
    uint32_t tmp[64];
    int cnt = ix->host.neighbors(slot, level, tmp, 64);
    if (cnt == -1) {
        set_error(error, format("Node %zu is out of range (size: %zu)", slot, ix->host.size()));
        return 0;
    }
    if (cnt == -2) {
        int node_level = slot < ix->host.size() ? ix->host.level_of(slot) : 0;
        set_error(error, format("Level %d is out of range for node (max level: %d)", level, node_level));
        return 0;
    }
    size_t written = 0;
    for (int j = 0; j < cnt && written + 2 <= capacity; ++j) {
        out_pairs[written++] = tmp[j];
        out_pairs[written++] = ix->host.key_of(tmp[j]);
    }
    return written;
}

size_t rb_neighbors_many(rb_index_t const* ix, uint64_t const* slots, int32_t const* levels, size_t count,
                         uint64_t* out_pairs, size_t capacity, uint64_t* offsets, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!ix->host_graph_valid) {
        set_error(error, "the host copy of the graph is stale: call rb_sync_host first");
        return 0;
    }
    size_t written = 0, failed = 0;
    offsets[0] = 0;
    uint32_t tmp[64];
    for (size_t i = 0; i < count; ++i) {
        int cnt = ix->host.neighbors((size_t)slots[i], levels ? (int)levels[i] : 0, tmp, 64);
        if (cnt < 0) {
            ++failed;
            cnt = 0;
        }
        if (written + 2 * (size_t)cnt > capacity) {
            set_error(error, "output capacity too small for rb_neighbors_many");
            return failed;
        }
        for (int j = 0; j < cnt; ++j) {
            out_pairs[written++] = tmp[j];
            out_pairs[written++] = ix->host.key_of(tmp[j]);
        }
        offsets[i + 1] = written;
    }
    return failed;
}

size_t rb_top_level_nodes(rb_index_t const* ix, uint64_t* out_pairs, size_t capacity, rb_error_t* error) {
    if (error)
        *error = nullptr;
    std::vector<uint32_t> top;
    ix->host.top_level_nodes(top);
    if (!out_pairs)
        return top.size() * 2;
    size_t written = 0;
    for (size_t j = 0; j < top.size() && written + 2 <= capacity; ++j) {
        out_pairs[written++] = top[j];
        out_pairs[written++] = ix->host.key_of(top[j]);
    }
    return written;
}

void rb_keys_from_slots(rb_index_t const* ix, uint64_t const* slots, size_t count, uint64_t* keys_out,
                        rb_error_t* error) {
    if (error)
        *error = nullptr;
    const size_t n = ix->host.size();
    for (size_t i = 0; i < count; ++i) {
        if (slots[i] >= n)
            RB_FAIL(format("Node %llu is out of range (size: %zu)", (unsigned long long)slots[i], n));
        keys_out[i] = ix->host.key_of(slots[i]);
    }
}

void rb_slots_from_keys(rb_index_t* ix, uint64_t const* keys, size_t count, int64_t* slots_out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    for (size_t i = 0; i < count; ++i)
        slots_out[i] = ix->host.slot_of_key(keys[i]);
}

int rb_contains(rb_index_t* ix, uint64_t key) { return ix->host.slot_of_key(key) >= 0; }

void rb_get_vectors(rb_index_t* ix, uint64_t const* slots, size_t count, uint8_t* out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    rbh::HostIndex& h = ix->host;
    const size_t n = h.size(), vb = h.vec_bytes;
    for (size_t i = 0; i < count; ++i)
        if (slots[i] >= n)
            RB_FAIL(format("Node %llu is out of range (size: %zu)", (unsigned long long)slots[i], n));
    if (h.viewing) {
        if (!h.viewed_vectors())
            RB_FAIL("The fingerprints of this index are not available (viewed with exclude_vectors)");
        for (size_t i = 0; i < count; ++i)
            std::memcpy(out + i * vb, h.viewed_vectors() + slots[i] * vb, vb);
        return;
    }
    if (!h.host_vectors.empty()) {
        for (size_t i = 0; i < count; ++i)
            std::memcpy(out + i * vb, h.host_vectors.data() + slots[i] * vb, vb);
        return;
    }
    if (!ix->device_graph_valid || ix->dev.device < 0)
        RB_FAIL("The fingerprints are not available");
    cudaSetDevice(ix->dev.device);
    for (size_t i = 0; i < count; ++i) {
        cudaError_t e = cudaMemcpy(out + i * vb, reinterpret_cast<uint8_t const*>(ix->dev.fp) + slots[i] * 128, vb,
                                   cudaMemcpyDeviceToHost);
        if (e != cudaSuccess)
            RB_FAIL(std::string("CUDA error while fetching fingerprints: ") + cudaGetErrorString(e));
    }
}

size_t rb_size(rb_index_t const* ix) { return ix->host.size(); }
size_t rb_capacity(rb_index_t const* ix) { return std::max(ix->host.capacity, ix->host.size()); }
int32_t rb_max_level(rb_index_t const* ix) { return ix->host.max_level; }
uint64_t rb_entry_slot(rb_index_t const* ix) { return ix->host.entry; }
size_t rb_connectivity(rb_index_t const* ix) { return ix->host.m; }
size_t rb_dimensions(rb_index_t const* ix) { return ix->host.ndim; }
size_t rb_memory_usage(rb_index_t const* ix) {
    size_t device = 0;
    DeviceState const& d = ix->dev;
    if (d.device >= 0)
        device = d.cap_rows * (128 + (size_t)ix->host.m0 * 4 + 4 + 1 + 4) + d.cap_upper * (size_t)ix->host.m * 4 +
                 d.pair_cap * 32 + d.sort_tmp_bytes + d.scratch_bytes;
    return ix->host.memory_usage() + device;
}
size_t rb_serialized_length(rb_index_t const* ix, int exclude_vectors) {
    return ix->host.serialized_length(exclude_vectors != 0);
}
size_t rb_expansion_add(rb_index_t const* ix) { return ix->host.ef_add; }
size_t rb_expansion_search(rb_index_t const* ix) { return ix->host.ef_search; }
void rb_set_expansion_add(rb_index_t* ix, size_t expansion) {
    ix->host.ef_add = (uint32_t)std::min<size_t>(std::max<size_t>(expansion, 1), 1024);
}
void rb_set_expansion_search(rb_index_t* ix, size_t expansion) {
    ix->host.ef_search = (uint32_t)std::min<size_t>(std::max<size_t>(expansion, 1), 1024);
}

void rb_level_stats(rb_index_t const* ix, size_t level, rb_level_stats_t* out, rb_error_t* error) {
    if (error)
        *error = nullptr;
    if (!ix->host_graph_valid)
        RB_FAIL("the host copy of the graph is stale: call rb_sync_host first");
    rbh::LevelStats st;
    if (!ix->host.level_stats(level, st))
        RB_FAIL("The level is out of range");
    out->nodes = st.nodes;
    out->edges = st.edges;
    out->max_edges = st.max_edges;
    out->allocated_bytes = st.allocated_bytes;
}

void rb_levels(rb_index_t const* ix, int16_t* out, size_t capacity) {
    size_t n = std::min(capacity, ix->host.size());
    for (size_t i = 0; i < n; ++i)
        out[i] = (int16_t)ix->host.level_of(i);
}

void rb_last_counters(rb_index_t const* ix, rb_counters_t* out) { *out = ix->last; }

uint64_t rb_debug_adjacency(rb_index_t const* ix, uint32_t* adj0_out, uint32_t* adj_upper_out) {
    rbh::HostIndex const& h = ix->host;
    if (h.viewing || !ix->host_graph_valid)
        return 0;
    if (adj0_out)
        std::memcpy(adj0_out, h.adj0.data(), h.adj0.size() * 4);
    if (adj_upper_out)
        std::memcpy(adj_upper_out, h.adju.data(), h.adju.size() * 4);
    return h.upper_off[h.keys.size()];
}

}  // extern "C"
