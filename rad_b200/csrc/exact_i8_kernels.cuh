// exact_i8_kernels.cuh -- exact Tanimoto top-k on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Same contract as exact_topk_kernel (exact_kernels.cuh): replaces index_gt::search_exact_
// (usearch/include/usearch/index.hpp:4222-4239) with the tie rule of sorted_buffer_gt::insert (index.hpp:923-934).
//
// Blackwell has no 1-bit MMA (ptxas expands mma.sync...b1.and.popc into 8 x IMMA.16832.U8 plus ~100 unpack
// instructions), so the intersection counts are formed as an unsigned-8-bit GEMM over 0/1 bytes:
//     inter[q][d] = sum_k bit_k(q) * bit_k(d)        (tcgen05.mma kind::i8, M = 128 queries, N = 256 rows, K = 32)
// Each fingerprint is unpacked bit -> byte ONCE per CTA tile by CUDA cores straight into the shared-memory operand
// layout, and then re-used against 128 (database rows) or every database tile of the chunk (queries):
//   * the CTA's 128 queries are unpacked once into a resident 128 KB A operand (K-major, no swizzle);
//   * 8 producer warps stream the database: one thread per row keeps its 128-byte line in registers and unpacks it
//     16 bytes (= one 128-bit K chunk) at a time into a 2-stage 32 KB B ring;
//   * one elected thread issues 4 MMAs per K chunk; accumulators (128 x 256 x s32) are double-buffered in TMEM;
//   * 4 epilogue warps read the counts back with tcgen05.ld (one query per thread) and run the same INTEGER filter
//     and the same slow path as the POPC kernel, so the results are bit-identical to it.
// Synchronisation is by mbarriers: full/empty per B stage (producers <-> MMA via tcgen05.commit), tmem_full /
// tmem_empty per accumulator stage (MMA <-> epilogue), pd_full for the per-row popcounts.
#pragma once
#include "exact_kernels.cuh"

namespace rb {

constexpr int kTQ = 128;             // queries per CTA (UMMA M)
constexpr int kTD = 256;             // database rows per tile (UMMA N)
constexpr int kTChunkBytes = 128;    // unpacked K bytes per stage (= 128 fingerprint bits = one uint4 per row)
constexpr int kTStages = 2;
constexpr int kTEpiWarps = 4, kTMmaWarp = 4, kTProdWarp0 = 5, kTProdWarps = 8;
constexpr int kTThreads = (kTEpiWarps + 1 + kTProdWarps) * 32;  // 416
constexpr int kTQueueCap = kTQ * 32;  // survivors of one 32-column block, worst case

// shared-memory operand layout (K-major, SWIZZLE_NONE): 16-byte K pieces of 8 consecutive rows form one contiguous
// 128-byte core matrix; core matrices step by LBO along K and by SBO along the rows
//   address(row, k16) = (row / 8) * SBO + k16 * LBO + (row % 8) * 16
constexpr uint32_t kLBO = 128;
constexpr uint32_t kSBO_A = 64 * kLBO;  // A holds all 64 K pieces of a row group
constexpr uint32_t kSBO_B = 8 * kLBO;   // a B stage holds 8
constexpr uint32_t kASmemBytes = (kTQ / 8) * kSBO_A;          // 131072
constexpr uint32_t kBStageBytes = (kTD / 8) * kSBO_B;         // 32768

struct TensorSmemTail {
    uint32_t pq[kTQ];
    uint32_t pd[2][kTD];
    uint4 thr[kTQ];
    uint32_t queue[kTQueueCap];
    uint32_t queue_count;
    uint32_t tmem_base;
    unsigned long long full[kTStages], empty[kTStages], tmem_full[2], tmem_empty[2], pd_full[2], pd_empty[2];
};
constexpr size_t kTensorSmemBytes = kASmemBytes + kTStages * kBStageBytes + sizeof(TensorSmemTail) + 1024;

// ---------------------------------------------------------------------------------------------- PTX helpers
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
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.b32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(void* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], unsigned 8-bit operands, s32 accumulate
__device__ __forceinline__ void tc_mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp: SmemDescriptor): start address, LBO and SBO in
// 16-byte units at bits [0,14), [16,30), [32,46); version 1 at bits [46,48); layout type 0 = no swizzle
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// instruction descriptor (InstrDescriptor): c_format = S32 (2) at [4,6), a/b format = UINT8 (0), both K-major,
// N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t kIdescI8 = (2u << 4) | ((uint32_t)(kTD >> 3) << 17) | ((uint32_t)(kTQ >> 4) << 24);

// 16 fingerprint bits -> 16 bytes of 0/1 (any fixed bit -> K position mapping works: both operands use this one)
__device__ __forceinline__ uint4 unpack16(uint32_t h) {
    uint4 o;
    o.x = ((h & 0xFu) * 0x00204081u) & 0x01010101u;
    o.y = (((h >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
    o.z = (((h >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
    o.w = (((h >> 12) & 0xFu) * 0x00204081u) & 0x01010101u;
    return o;
}

#define RB_TMEM_LD32(acc, taddr)                                                                                   \
    asm volatile(                                                                                                   \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                   \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                   \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                   \
        : "=r"(acc[0]), "=r"(acc[1]), "=r"(acc[2]), "=r"(acc[3]), "=r"(acc[4]), "=r"(acc[5]), "=r"(acc[6]),         \
          "=r"(acc[7]), "=r"(acc[8]), "=r"(acc[9]), "=r"(acc[10]), "=r"(acc[11]), "=r"(acc[12]), "=r"(acc[13]),     \
          "=r"(acc[14]), "=r"(acc[15]), "=r"(acc[16]), "=r"(acc[17]), "=r"(acc[18]), "=r"(acc[19]), "=r"(acc[20]),  \
          "=r"(acc[21]), "=r"(acc[22]), "=r"(acc[23]), "=r"(acc[24]), "=r"(acc[25]), "=r"(acc[26]), "=r"(acc[27]),  \
          "=r"(acc[28]), "=r"(acc[29]), "=r"(acc[30]), "=r"(acc[31])                                                \
        : "r"(taddr)                                                                                                \
        : "memory")

// grid = (query tiles of 128, database chunks); ExactParams as for exact_topk_kernel with chunk_rows a multiple of kTD
__global__ void __launch_bounds__(kTThreads, 1) exact_topk_i8_kernel(ExactParams p) {
    extern __shared__ unsigned char raw_unaligned[];
    unsigned char* raw = reinterpret_cast<unsigned char*>(((uintptr_t)raw_unaligned + 1023) & ~(uintptr_t)1023);
    unsigned char* smem_a = raw;
    unsigned char* smem_b = raw + kASmemBytes;
    TensorSmemTail& s = *reinterpret_cast<TensorSmemTail*>(raw + kASmemBytes + kTStages * kBStageBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q0 = blockIdx.x * kTQ;
    const uint32_t chunk = blockIdx.y;
    const uint64_t row_begin = (uint64_t)chunk * p.chunk_rows;
    const uint64_t row_end = row_begin + p.chunk_rows < p.rows ? row_begin + p.chunk_rows : p.rows;
    const uint64_t n_tiles = row_begin < row_end ? (row_end - row_begin + kTD - 1) / kTD : 0;

    // ---- the CTA's partial lists start empty
    {
        uint32_t nq = p.query_count - q0 < (uint32_t)kTQ ? p.query_count - q0 : (uint32_t)kTQ;
        size_t base = ((size_t)chunk * p.query_count + q0) * p.k;
        for (size_t i = tid; i < (size_t)nq * p.k; i += kTThreads)
            p.partial_keys[base + i] = ~0ull;
    }
    if (n_tiles == 0)
        return;  // uniform

    // ---- query tile: popcounts, thresholds, and the resident unpacked A operand
    for (int r = tid; r < kTQ; r += kTThreads) {
        uint32_t q = q0 + r, pc = 0;
        if (q < p.query_count)
            for (int w = 0; w < 8; ++w)
                pc += popc4(__ldg(p.queries + (size_t)q * 8 + w));
        s.pq[r] = q < p.query_count ? pc : kXInvalid;
        s.thr[r] = make_uint4(0, 0, 0, 1);
    }
    {
        const uint16_t* q16 = reinterpret_cast<const uint16_t*>(p.queries);
        for (int idx = tid; idx < kTQ * 64; idx += kTThreads) {
            int r = idx & (kTQ - 1), k16 = idx >> 7;
            uint32_t q = q0 + r;
            uint32_t h = q < p.query_count ? (uint32_t)__ldg(q16 + (size_t)q * 64 + k16) : 0u;
            *reinterpret_cast<uint4*>(smem_a + (r >> 3) * kSBO_A + k16 * kLBO + (r & 7) * 16) = unpack16(h);
        }
    }
    if (tid == 0) {
        for (int i = 0; i < kTStages; ++i) {
            mbar_init(&s.full[i], kTProdWarps * 32);
            mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s.tmem_full[i], 1);
            mbar_init(&s.tmem_empty[i], kTEpiWarps * 32);
            mbar_init(&s.pd_full[i], kTProdWarps * 32);
            mbar_init(&s.pd_empty[i], kTEpiWarps * 32);
        }
        s.queue_count = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s.tmem_base))
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();  // the A operand was written through the generic proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = s.tmem_base;

    if (warp >= kTProdWarp0) {
        // ================================================================ producers: one database row per thread
        const int pr = tid - kTProdWarp0 * 32;  // 0 .. 255
        unsigned char* my_b = smem_b + (pr >> 3) * kSBO_B + (pr & 7) * 16;
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;  // highest rows first (see exact_kernels.cuh)
            const uint64_t row = row_begin + tile * kTD + pr;
            uint4 v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c)
                v[c] = row < row_end ? __ldg(p.db + row * 8 + c) : make_uint4(0, 0, 0, 0);
            uint32_t pc = 0;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                pc += popc4(v[c]);
                const uint64_t g = t * 8 + c;
                const int st = (int)(g & (kTStages - 1));
                mbar_wait(&s.empty[st], (uint32_t)(((g / kTStages) & 1) ^ 1));
                unsigned char* dst = my_b + st * kBStageBytes;
                const uint32_t w[4] = {v[c].x, v[c].y, v[c].z, v[c].w};
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint32_t h = (j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xFFFFu);
                    *reinterpret_cast<uint4*>(dst + j * kLBO) = unpack16(h);
                }
                fence_proxy_async();
                mbar_arrive(&s.full[st]);
            }
            const int a = (int)(t & 1);
            mbar_wait(&s.pd_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1));  // the epilogue of tile t-2 is done with pd[a]
            s.pd[a][pr] = row < row_end ? pc : kXInvalid;
            mbar_arrive(&s.pd_full[a]);
        }
    } else if (warp == kTMmaWarp) {
        // ================================================================ MMA issuer
        const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), kLBO, kSBO_A);
        const uint64_t b_desc0 = make_smem_desc(smem_u32(smem_b), kLBO, kSBO_B);
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const int a = (int)(t & 1);
            mbar_wait(&s.tmem_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1));
            tc_fence_after();
            for (int c = 0; c < 8; ++c) {
                const uint64_t g = t * 8 + c;
                const int st = (int)(g & (kTStages - 1));
                mbar_wait(&s.full[st], (uint32_t)((g / kTStages) & 1));
                tc_fence_after();
                if (lane == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        // K = 32 bytes per instruction = two 16-byte K pieces
                        uint64_t ad = a_desc0 + (uint64_t)(((c * 8 + j * 2) * kLBO) >> 4);
                        uint64_t bd = b_desc0 + (uint64_t)((st * kBStageBytes + j * 2 * kLBO) >> 4);
                        tc_mma_i8(tmem_base + (uint32_t)a * kTD, ad, bd, kIdescI8, (uint32_t)((c | j) != 0));
                    }
                    tc_commit(&s.empty[st]);
                    if (c == 7)
                        tc_commit(&s.tmem_full[a]);
                }
                __syncwarp();
            }
        }
    } else {
        // ================================================================ epilogue: thread = one query of the tile
        const int ql = warp * 32 + lane;
        const uint32_t pqv = s.pq[ql];
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;
            const int a = (int)(t & 1);
            mbar_wait(&s.pd_full[a], (uint32_t)((t >> 1) & 1));
            mbar_wait(&s.tmem_full[a], (uint32_t)((t >> 1) & 1));
            tc_fence_after();
            // 32-column blocks from the HIGHEST rows down: the thresholds move between blocks, and a row that only ties
            // the k-th distance must never displace a higher slot (same invariant as the tile order)
            for (int b = kTD / 32 - 1; b >= 0; --b) {
                uint32_t acc[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(a * kTD + b * 32);
                RB_TMEM_LD32(acc, taddr);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (b == 0) {  // the accumulator stage is free as soon as the last block is in registers
                    tc_fence_before();
                    mbar_arrive(&s.tmem_empty[a]);
                }
                const uint4 th = s.thr[ql];
                if (pqv != kXInvalid) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const uint32_t pdv = s.pd[a][b * 32 + j];
                        const uint32_t i = acc[j];
                        if (pdv != kXInvalid && i * th.x + th.w > th.z + th.y * pdv) {
                            uint32_t at = atomicAdd(&s.queue_count, 1u);
                            s.queue[at] = ((uint32_t)ql << 24) | ((uint32_t)(b * 32 + j) << 16) | i;
                        }
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                const uint32_t qn = s.queue_count;
                for (uint32_t base = 0; base < qn; base += 32) {
                    uint32_t e = base + lane < qn ? s.queue[base + lane] : 0xFFFFFFFFu;
                    uint32_t mine = __ballot_sync(RB_FULL_MASK, base + lane < qn && (int)((e >> 24) & 3) == warp);
                    while (mine) {
                        int jj = __ffs((int)mine) - 1;
                        mine &= mine - 1;
                        uint32_t ev = __shfl_sync(RB_FULL_MASK, e, jj);
                        uint32_t eq = ev >> 24, dl = (ev >> 16) & 0xFFu, i = ev & 0xFFFFu;
                        uint32_t u = s.pq[eq] + s.pd[a][dl] - i;
                        uint32_t slot = (uint32_t)(p.slot_base + row_begin + tile * kTD + dl);
                        uint64_t key = ((uint64_t)exact_distance_bits(i, u) << 32) | (uint32_t)~slot;
                        size_t off = ((size_t)chunk * p.query_count + q0 + eq) * p.k;
                        glist_insert(p.partial_keys + off, p.partial_iu + off, p.k, key, i | ((u ? u : 1u) << 16));
                        if (p.partial_keys[off + p.k - 1] != ~0ull) {
                            uint32_t iu = p.partial_iu[off + p.k - 1];
                            uint32_t it_ = iu & 0xFFFFu, ut = iu >> 16;
                            if (lane == 0)
                                s.thr[eq] = make_uint4(ut + it_, it_, it_ * s.pq[eq], 0);
                        }
                        __syncwarp();
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (tid == 0)
                    s.queue_count = 0;
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            mbar_arrive(&s.pd_empty[a]);
        }
    }
    __syncthreads();
    if (warp == kTMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

}  // namespace rb
