// device_common.cuh -- shared device primitives: Tanimoto counts on 128-byte fingerprints, order keys, warp helpers.
//
// Replaces the reference's innermost kernel simsimd_jaccard_b8_{serial,haswell,ice}
// (usearch/simsimd/include/simsimd/binary.h:99-105, 349-444, 465-474) reached through metric_punned_t
// (usearch/include/usearch/index_plugins.hpp:1691-1693, 1886-1893).
//
// Layout: one fingerprint = 128 bytes = 8 x uint4, 128-byte aligned (one L2 line).  A group of 8 lanes owns one
// fingerprint (lane & 7 selects the 16-byte slice), so a warp reads 4 complete lines per LDG.128 instruction.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define RB_EMPTY 0xFFFFFFFFu
#define RB_FULL_MASK 0xFFFFFFFFu

namespace rb {

// Reference-exact distance bits: (float)(1.0 - (double)i / (double)u), u == 0 -> 1  (binary.h:104, index_plugins.hpp:1892).
__device__ __forceinline__ uint32_t exact_distance_bits(uint32_t inter, uint32_t uni) {
    double d = uni ? __dsub_rn(1.0, __ddiv_rn((double)inter, (double)uni)) : 1.0;
    return __float_as_uint(__double2float_rn(d));
}

// Order key used inside graph traversal: 1.0f - RN(i / u) in f32.  Strictly monotone in the exact rational i/u and
// equal for equal rationals (checked exhaustively for u <= 1024 in tests/test_order_key.py), so every comparison
// agrees with the reference-exact bits, at a fraction of the cost of the f64 divide.  Never returned to callers.
__device__ __forceinline__ uint32_t order_key_bits(uint32_t inter, uint32_t uni) {
    float s = uni ? __fdiv_rn((float)inter, (float)uni) : 0.0f;
    return __float_as_uint(__fsub_rn(1.0f, s));
}

__device__ __forceinline__ uint32_t popc4(uint4 v) { return __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w); }
__device__ __forceinline__ uint32_t popc4_and(uint4 a, uint4 b) {
    return __popc(a.x & b.x) + __popc(a.y & b.y) + __popc(a.z & b.z) + __popc(a.w & b.w);
}
__device__ __forceinline__ uint32_t popc4_or(uint4 a, uint4 b) {
    return __popc(a.x | b.x) + __popc(a.y | b.y) + __popc(a.z | b.z) + __popc(a.w | b.w);
}

// sum over the 8 lanes of a group (all 32 lanes must call)
__device__ __forceinline__ uint32_t group8_sum(uint32_t v) {
    v += __shfl_xor_sync(RB_FULL_MASK, v, 1);
    v += __shfl_xor_sync(RB_FULL_MASK, v, 2);
    v += __shfl_xor_sync(RB_FULL_MASK, v, 4);
    return v;
}

// intersection | union << 16 of two fingerprints held slice-wise by a group of 8 lanes
__device__ __forceinline__ uint32_t group8_inter_union(uint4 a, uint4 b) {
    return group8_sum(popc4_and(a, b) | (popc4_or(a, b) << 16));
}

__device__ __forceinline__ uint4 load_slice(const uint4* __restrict__ fp, uint32_t slot, int sub) {
    return __ldg(fp + (size_t)slot * 8 + sub);
}

struct GraphView {
    const uint4* fp;            // rows x 8 uint4
    uint32_t* adj0;             // rows x m0, RB_EMPTY padded
    uint32_t* adju;             // upper rows x m
    const uint32_t* upper_off;  // rows: first upper row of a node (levels 1..level are consecutive)
    uint16_t* pc0;              // rows x m0: popcount of the fingerprint each adj0 entry points at (device-only side
    uint16_t* pcu;              // upper rows x m: the same for adju                                  array, see below)
    uint32_t m, m0;
};

// Link popcounts.  union = |a| + |q| - intersection, so a traversal that knows |a| for every link it follows needs
// only the AND-POPC half of the Tanimoto counts.  |a| travels WITH the link: pc0 / pcu mirror adj0 / adju entry for
// entry (2 bytes per link, read by the same lanes in the same instruction slot as the link row: one extra 32-byte
// sector per expansion instead of one gather per candidate).  Every writer of a link row writes its popcounts.

__device__ __forceinline__ uint32_t* graph_row(const GraphView& g, uint32_t slot, int level, uint32_t& width) {
    if (level == 0) {
        width = g.m0;
        return g.adj0 + (size_t)slot * g.m0;
    }
    width = g.m;
    return g.adju + ((size_t)__ldg(g.upper_off + slot) + (uint32_t)(level - 1)) * g.m;
}

// the popcount row that mirrors `row` (a pointer returned by graph_row for the same level)
__device__ __forceinline__ uint16_t* graph_row_pc(const GraphView& g, const uint32_t* row, int level) {
    return level == 0 ? g.pc0 + (row - g.adj0) : g.pcu + (row - g.adju);
}

// Lane j < cnt holds ids_j (cnt <= 32, warp-uniform); returns to lane j the popcount of fingerprint ids_j.
__device__ __noinline__ uint32_t warp_popcounts(const uint4* __restrict__ fp, uint32_t ids, uint32_t cnt) {
    const int lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
    uint32_t mine = 0;
#pragma unroll 1
    for (uint32_t base = 0; base < cnt; base += 4) {
        uint32_t j = base + grp;
        uint32_t id = __shfl_sync(RB_FULL_MASK, ids, j & 31);
        uint32_t v = group8_sum(j < cnt ? popc4(load_slice(fp, id, sub)) : 0u);
        uint32_t got = __shfl_sync(RB_FULL_MASK, v, (lane & 3) * 8);
        if ((uint32_t)(lane >> 2) == (base >> 2))
            mine = got;
    }
    return mine;
}

// ------------------------------------------------------------------------------------------------ mbarriers, bulk copies
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(void* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"  // %3: suspend-time hint, the thread
            "selp.b32 %0, 1, 0, p;\n\t}"                                        // sleeps instead of re-polling
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u)
            : "memory");
    } while (!done);
}
// one arrival that also announces `bytes` of asynchronous copies to come (complete_tx) on this barrier
__device__ __forceinline__ void mbar_arrive_expect_tx(void* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bulk-asynchronous copies (the TMA engine's linear form, cp.async.bulk): `bytes` is a multiple of 16 and both
// addresses are 16-byte aligned.  global -> shared completes on an mbarrier (complete_tx::bytes); shared -> global is
// tracked by the issuing thread's bulk group.
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the issuing thread's bulk stores have finished READING shared memory (the source may be overwritten)
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// shared-memory writes made through the generic proxy (st.shared) become visible to the asynchronous proxy
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace rb
