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
#ifndef RB_I8_PIECES
#define RB_I8_PIECES 8
#endif
constexpr int kTPieces = RB_I8_PIECES;         // 16-byte K pieces (= 16 fingerprint bits each) per B stage: 8 or 4
constexpr int kTStages = 16 / kTPieces;        // the B ring is 64 KB: 2 stages of 128 bits or 4 stages of 64 bits
                                               // (measured: 4 x 64 bits is 20 % slower -- two MMAs per commit)
constexpr int kTSteps = 64 / kTPieces;         // B stages per database tile
constexpr int kTEpiWarps = 4, kTMmaWarp = 4, kTProdWarp0 = 5, kTProdWarps = 8;
constexpr int kTThreads = (kTEpiWarps + 1 + kTProdWarps) * 32;  // 416
constexpr int kTQueueCap = kTQ * 8;   // survivors of one 8-column round of a block, worst case (32 queries x 8 per warp)
constexpr int kTListK = 16;           // top-k lists of up to this k live in shared memory for the whole kernel

// shared-memory operand layout (K-major, SWIZZLE_NONE): 16-byte K pieces of 8 consecutive rows form one contiguous
// 128-byte core matrix; core matrices step by LBO along K and by SBO along the rows
//   address(row, k16) = (row / 8) * SBO + k16 * LBO + (row % 8) * 16
constexpr uint32_t kLBO = 128;
constexpr uint32_t kSBO_A = 64 * kLBO;  // A holds all 64 K pieces of a row group
constexpr uint32_t kSBO_B = kTPieces * kLBO;   // a B stage holds kTPieces
constexpr uint32_t kASmemBytes = (kTQ / 8) * kSBO_A;          // 131072
constexpr uint32_t kBStageBytes = (kTD / 8) * kSBO_B;         // 32768

struct TensorSmemTail {
    uint32_t pq[kTQ];
    uint32_t pd[2][kTD];
    uint4 thr[kTQ];
    uint32_t queue[kTEpiWarps][kTQueueCap / kTEpiWarps];  // survivors of one warp's 32 queries x 8 columns
    uint32_t queue_count[kTEpiWarps];
    // per-query sorted partial lists (k <= kTListK), written back at the end; thread = query in the epilogue, so rows
    // are padded by one entry: the 32 lanes of a warp touching the same column hit distinct banks
    uint32_t lnslot[kTQ][kTListK + 1];  // ~slot of each entry
    uint32_t liu[kTQ][kTListK + 1];     // i | u << 16 of each entry: lists are ordered on the exact rationals
    uint32_t tmem_base;
    unsigned long long full[8], empty[8], tmem_full[2], tmem_empty[2], pd_full[2], pd_empty[2];
};
constexpr size_t kTensorSmemBytes = kASmemBytes + kTStages * kBStageBytes + sizeof(TensorSmemTail) + 128;
static_assert(kTStages <= 8, "barrier arrays");
static_assert(kTensorSmemBytes <= 232448, "227 KB of shared memory per CTA");

// ---------------------------------------------------------------------------------------------- PTX helpers
// (smem_u32, mbar_init, mbar_arrive, mbar_wait: device_common.cuh)
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

// 32 fingerprint bits -> 32 bytes as two 16-byte K pieces.  Any fixed bit -> K position mapping works as long as both
// operands use it, so the cheapest one is taken: output word j holds bits j, j+8, j+16, j+24 of the input word.
// The DATABASE side, unpacked for every tile, leaves each bit where it is inside its byte: w & (0x01010101 << j), ONE
// instruction per four bits, byte value 2^j.  The QUERY side, unpacked once per CTA, compensates with byte value
// 2^(7-j): every matching bit contributes 2^j * 2^(7-j) = 128, so the accumulators hold 128 x intersection
// (<= 2^17, exact in s32) and the epilogue's thresholds are scaled by 128 instead of the counts being shifted.
constexpr int kTScaleLog2 = 7;
__device__ __forceinline__ uint4 unpack_half_db(uint32_t w, int half) {
    const uint32_t m = 0x01010101u;
    const int j = half * 4;
    return make_uint4(w & (m << j), w & (m << (j + 1)), w & (m << (j + 2)), w & (m << (j + 3)));
}
__device__ __forceinline__ uint4 unpack_half_query(uint32_t w, int half) {
    const uint32_t m = 0x01010101u;
    const int j = half * 4;
    return make_uint4(((w >> j) & m) << (7 - j), ((w >> (j + 1)) & m) << (6 - j), ((w >> (j + 2)) & m) << (5 - j),
                      ((w >> (j + 3)) & m) << (4 - j));
}
// a[c] for a run-time c without moving the array out of registers: a binary tree of 31 selects
__device__ __forceinline__ uint32_t select32(const uint32_t (&a)[32], uint32_t c) {
    uint32_t t16[16], t8[8], t4[4];
#pragma unroll
    for (int k = 0; k < 16; ++k)
        t16[k] = (c & 16u) ? a[k + 16] : a[k];
#pragma unroll
    for (int k = 0; k < 8; ++k)
        t8[k] = (c & 8u) ? t16[k + 8] : t16[k];
#pragma unroll
    for (int k = 0; k < 4; ++k)
        t4[k] = (c & 4u) ? t8[k + 4] : t8[k];
    const uint32_t t2a = (c & 2u) ? t4[2] : t4[0], t2b = (c & 2u) ? t4[3] : t4[1];
    return (c & 1u) ? t2b : t2a;
}
// padding rows carry this "popcount": >= 2048 keeps them out while a list is not full (see the filter), and small
// enough that 128 * y * |d| stays inside 32 bits
constexpr uint32_t kTPadRow = 4096;

#define RB_TICK(slot, expr)                                      \
    do {                                                         \
        long long t0_ = dbg ? clock64() : 0;                     \
        expr;                                                    \
        if (dbg)                                                 \
            dbg_acc[slot] += (unsigned long long)(clock64() - t0_); \
    } while (0)

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
//
// (Round 1 also carried a CTA-pair form -- cta_group::2, half of every database tile unpacked per CTA -- and a form with
// the query operand in tensor memory; both were bit-identical and no faster (DESIGN.md section 4 has the measurements)
// and went when exact_topk_f4_kernel became the default tensor path.)
__global__ void __launch_bounds__(kTThreads, 1) exact_topk_i8_kernel(ExactParams p) {
    constexpr int STAGES = kTStages;                        // B ring stages
    constexpr int PIECES = kTPieces;                        // K pieces per stage
    constexpr int STEPS = 64 / PIECES;                      // stages per database tile
    constexpr uint32_t SBO_B = PIECES * kLBO;
    constexpr uint32_t STAGE_BYTES = (kTD / 8) * SBO_B;
    // no swizzle: the operand descriptors only need 16-byte alignment, which dynamic shared memory guarantees; the
    // array is used directly so that every access stays in the shared address space (LDS / STS, not generic LD / ST)
    extern __shared__ __align__(1024) unsigned char raw[];
    unsigned char* smem_a = raw;
    unsigned char* smem_b = raw + kASmemBytes;
    TensorSmemTail& s = *reinterpret_cast<TensorSmemTail*>(raw + kASmemBytes + kTStages * kBStageBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Triangular form (the batch tile of the build, ExactParams::tri): blockIdx.y is a LEVEL, not a chunk; the level's
    // queries are the first tri_cnt[level] rows of the one array, query q sees rows [0, q), and the CTAs with the most
    // rows in front of them come first in launch order.
    const bool tri = p.tri != 0;
    const uint32_t qcount = tri ? p.tri_cnt[blockIdx.y] : p.query_count;
    const uint32_t q0 = (tri ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * kTQ;  // may lie past the last query (the odd
                                                                                // CTA of the last pair): all rows invalid
    if (tri && q0 >= qcount)
        return;  // uniform: the launch is sized for the widest level
    const uint32_t chunk = tri ? 0u : blockIdx.y;
    const size_t out_base = tri ? (size_t)p.tri_off[blockIdx.y] : (size_t)chunk * p.query_count;  // first list of this CTA's
                                                                                                 // level / chunk
    const size_t bound_base = tri ? (size_t)p.tri_off[blockIdx.y] : 0;
    const uint64_t row_begin = (uint64_t)chunk * p.chunk_rows;
    const uint64_t row_end = tri ? (qcount < q0 + kTQ - 1 ? qcount : q0 + kTQ - 1)
                                 : (row_begin + p.chunk_rows < p.rows ? row_begin + p.chunk_rows : p.rows);
    const uint64_t n_tiles = row_begin < row_end ? (row_end - row_begin + kTD - 1) / kTD : 0;

    // ---- the CTA's partial lists start empty
    {
        uint32_t nq = q0 >= qcount ? 0u : qcount - q0 < (uint32_t)kTQ ? qcount - q0 : (uint32_t)kTQ;
        size_t base = (out_base + q0) * p.k;
        for (size_t i = tid; i < (size_t)nq * p.k; i += kTThreads)
            p.partial_keys[base + i] = ~0ull;
    }
    if (n_tiles == 0)
        return;  // uniform

    // ---- query tile: popcounts, thresholds, and the resident unpacked A operand
    for (int r = tid; r < kTQ; r += kTThreads) {
        uint32_t q = q0 + r, pc = 0;
        if (q < qcount)
            for (int w = 0; w < 8; ++w)
                pc += popc4(__ldg(p.queries + (size_t)q * 8 + w));
        s.pq[r] = q < qcount ? pc : kXInvalid;
        s.thr[r] = make_uint4(0, 1, 0, 2048);  // list not full: every valid row (|d| < 2048) passes, padding rows do not
    }
    {
        const uint32_t* q32 = reinterpret_cast<const uint32_t*>(p.queries);
        for (int idx = tid; idx < kTQ * 64; idx += kTThreads) {
            int r = idx & (kTQ - 1), k16 = idx >> 7;
            uint32_t q = q0 + r;
            uint32_t w = q < qcount ? __ldg(q32 + (size_t)q * 32 + (k16 >> 1)) : 0u;
            *reinterpret_cast<uint4*>(smem_a + (r >> 3) * kSBO_A + k16 * kLBO + (r & 7) * 16) =
                unpack_half_query(w, k16 & 1);
        }
    }
    if (tid == 0) {
        constexpr uint32_t ctas = 1;
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&s.full[i], ctas * kTProdWarps);  // one arrival per producer WARP (256 arrivals on one word serialise)
            mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s.tmem_full[i], 1);
            mbar_init(&s.tmem_empty[i], ctas * kTEpiWarps);
            mbar_init(&s.pd_full[i], ctas * kTProdWarps);
            mbar_init(&s.pd_empty[i], ctas * kTEpiWarps);
        }
        for (int i = 0; i < kTEpiWarps; ++i)
            s.queue_count[i] = 0;
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
    const bool dbg = p.debug_cycles != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
    unsigned long long dbg_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long dbg_start = clock64();

    if (warp >= kTProdWarp0) {
        // ================================================================ producers: one database row per thread
        const int pr = tid - kTProdWarp0 * 32;  // 0 .. 255
        unsigned char* my_b = smem_b + (pr >> 3) * kSBO_B + (pr & 7) * 16;
        // the row of tile t + 1 is requested while tile t is unpacked (its L2 latency is off the hand-off path), and the
        // pieces of a stage are formed in registers BEFORE the wait for the stage, so that only stores follow the wait
        uint4 v[8];
        {
            const uint64_t row0 = row_begin + (n_tiles - 1) * kTD + pr;
#pragma unroll
            for (int c = 0; c < 8; ++c)
                v[c] = row0 < row_end ? __ldg(p.db + row0 * 8 + c) : make_uint4(0, 0, 0, 0);
        }
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;  // highest rows first (see exact_kernels.cuh)
            const uint64_t row = row_begin + tile * kTD + pr;
            uint4 vn[8];
            {
                const uint64_t rown = row - kTD;  // tile - 1 (only used when t + 1 < n_tiles; always < row_end then)
                const bool more = t + 1 < n_tiles;
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    vn[c] = more ? __ldg(p.db + rown * 8 + c) : make_uint4(0, 0, 0, 0);
            }
            uint32_t pc = 0;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                pc += popc4(v[c]);
                const uint32_t w[4] = {v[c].x, v[c].y, v[c].z, v[c].w};
#pragma unroll
                for (int h = 0; h < 8 / kTPieces; ++h) {  // K piece c * 8 + h * kTPieces + j  <->  A piece of the same index
                    const uint64_t g = t * kTSteps + c * (8 / kTPieces) + h;
                    const int st = (int)(g % STAGES);
                    uint4 u[kTPieces];
#pragma unroll
                    for (int j = 0; j < kTPieces; ++j)
                        u[j] = unpack_half_db(w[(h * kTPieces + j) >> 1], j & 1);
                    RB_TICK(1, mbar_wait(&s.empty[st], (uint32_t)(((g / STAGES) & 1) ^ 1)));
                    unsigned char* dst = my_b + st * STAGE_BYTES;
#pragma unroll
                    for (int j = 0; j < kTPieces; ++j)
                        *reinterpret_cast<uint4*>(dst + j * kLBO) = u[j];
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0)
                        mbar_arrive(&s.full[st]);
                }
            }
#pragma unroll
            for (int c = 0; c < 8; ++c)
                v[c] = vn[c];
            const int a = (int)(t & 1);
            RB_TICK(2, mbar_wait(&s.pd_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1)));  // epilogue of tile t-2 done with pd[a]
            s.pd[a][pr] = row < row_end ? pc : kTPadRow;
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&s.pd_full[a]);
        }
    } else if (warp == kTMmaWarp) {
        // ================================================================ MMA issuer
        const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), kLBO, kSBO_A);
        const uint64_t b_desc0 = make_smem_desc(smem_u32(smem_b), kLBO, SBO_B);
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const int a = (int)(t & 1);
            const uint32_t tph = (uint32_t)((t >> 1) & 1);
            RB_TICK(1, mbar_wait(&s.tmem_empty[a], tph ^ 1));
            tc_fence_after();
            for (int c = 0; c < STEPS; ++c) {
                const uint64_t g = t * STEPS + c;
                const int st = (int)(g % STAGES);
                RB_TICK(2, mbar_wait(&s.full[st], (uint32_t)((g / STAGES) & 1)));
                tc_fence_after();
                if (lane == 0) {
#pragma unroll
                    for (int j = 0; j < PIECES / 2; ++j) {
                        // K = 32 bytes per instruction = two 16-byte K pieces
                        uint64_t ad = a_desc0 + (uint64_t)(((c * PIECES + j * 2) * kLBO) >> 4);
                        uint64_t bd = b_desc0 + (uint64_t)((st * STAGE_BYTES + j * 2 * kLBO) >> 4);
                        tc_mma_i8(tmem_base + (uint32_t)a * kTD, ad, bd, kIdescI8, (uint32_t)((c | j) != 0));
                    }
                    tc_commit(&s.empty[st]);
                    if (c == STEPS - 1)
                        tc_commit(&s.tmem_full[a]);
                }
                __syncwarp();
            }
        }
    } else {
        // ================================================================ epilogue: thread = one query of the tile
        const int ql = warp * 32 + lane;
        const uint32_t pqv = s.pq[ql];
        const bool smem_lists = p.k <= (uint32_t)kTListK;
        const bool share_bounds = smem_lists && p.bounds != nullptr;  // bound word: d bits << 32 | chunk << 23 | i << 12 | u
        // this thread's own list (k <= kTListK): length, and its worst entry once it is full -- counts, ~slot, and the
        // reference-exact distance bits that bound words are compared on
        const uint32_t kk = p.k;
        uint32_t llen = 0, w_i = 0, w_u = 1, w_ns = 0, w_pos = 0, own_bits = 0xFFFFFFFFu;
        unsigned long long n_hits = 0, n_ins = 0;  // RB_EXACT_DEBUG / RB_TILE_DEBUG: filter survivors and list insertions
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;
            const int a = (int)(t & 1);            // popcount side array of this tile
            const int acc_i = a;                   // accumulator stage
            const uint32_t acc_ph = (uint32_t)((t >> 1) & 1);
            RB_TICK(0, mbar_wait(&s.pd_full[a], acc_ph));
            RB_TICK(1, mbar_wait(&s.tmem_full[acc_i], acc_ph));
            tc_fence_after();
            // Shared bounds.  Every CTA starts with empty lists, and until a list is full and tight nearly every pair
            // survives the filter (about k ln(rows/k) survivors per query and CTA, ~1.5 k cycles each on the slow path:
            // 2.9 ms per CTA at k = 10).  The k-th best distance ANY CTA has reached for a query bounds the final k-th
            // distance from above, so it is published (atomicMin below) and adopted here whenever it is tighter than
            // this CTA's own.  The bound word carries the CHUNK it was found in: chunks are ascending slot ranges, so a
            // row of a lower chunk that merely ties the bound loses to the k rows behind it (strict test, w = 0), while
            // a row of a higher chunk wins the tie on its slot (non-strict, w = 1).  Which survivors take the slow path
            // depends on timing; the result does not.
            if (share_bounds && pqv != kXInvalid) {
                const unsigned long long g = __ldcg(p.bounds + bound_base + q0 + ql);
                if (g != ~0ull && (llen < kk || (!tri && (uint32_t)(g >> 32) < own_bits))) {  // (a triangular list,
                    // once full, is tighter than its seed: every entry passed it)
                    const uint32_t holder = (uint32_t)(g >> 23) & 0x1FFu;
                    const uint32_t it_ = (uint32_t)(g >> 12) & 0x7FFu, ut = (uint32_t)g & 0xFFFu;
                    s.thr[ql] = make_uint4(ut + it_, it_, it_ * pqv, chunk < holder ? 0u : 1u);
                }
            }
            // 32-column blocks from the HIGHEST rows down: the thresholds move between blocks, and a row that only ties
            // the k-th distance must never displace a higher slot (same invariant as the tile order)
            // (Keeping the TMEM read of block b-1 in flight while block b is filtered -- two register buffers -- was
            // measured 5 % SLOWER: 128 registers with spills, and the read competes with the MMAs' accumulator traffic.)
            for (int b = kTD / 32 - 1; b >= 0; --b) {
                uint32_t acc[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc_i * kTD + b * 32);
                long long te0 = dbg ? clock64() : 0;
                RB_TMEM_LD32(acc, taddr);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                long long te1 = dbg ? clock64() : 0;
                if (b == 0) {  // the accumulator stage is free as soon as the last block is in registers
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0)
                        mbar_arrive(&s.tmem_empty[acc_i]);
                }
                // exact integer filter, 3 instructions per pair: the pair passes iff
                //     i * A + w > B + y * |d|      (A = u_t + i_t, y = i_t, B = i_t |q|, w = 0; see exact_kernels.cuh)
                // i.e. iff  dlt = (y |d| + B - w) - i A  is negative; its sign bit is funnel-shifted into one of four
                // independent 8-bit masks (pair 8 r + j of the block ends up at bit 7 - j of mask r).  Padding rows
                // (|d| = kTPadRow, i = 0) never pass.  The accumulators hold 128 i: y, B and w are scaled to match.
                uint32_t hm[4] = {0, 0, 0, 0};
                {
                    const uint4 th = s.thr[ql];
                    const uint32_t neg_a = 0u - th.x, bw = (th.z - th.w) << kTScaleLog2, y = th.y << kTScaleLog2;
                    const uint4* pd4 = reinterpret_cast<const uint4*>(&s.pd[a][b * 32]);
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        const uint4 pv = pd4[g];
                        const uint32_t pdv[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            uint32_t r1, dlt;  // two multiply-adds, kept from being re-associated into three
                            asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r1) : "r"(y), "r"(pdv[e]), "r"(bw));
                            asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(dlt) : "r"(acc[g * 4 + e]), "r"(neg_a), "r"(r1));
                            hm[g >> 1] = __funnelshift_l(dlt, hm[g >> 1], 1);
                        }
                    }
                }
                const uint32_t any_hit = pqv == kXInvalid ? 0u : (hm[0] | hm[1] | hm[2] | hm[3]);
                long long te2 = dbg ? clock64() : 0, te3 = te2;
                uint32_t qn_total = 0;
                if (smem_lists) {
                    // Survivors, lists in shared memory (k <= kTListK): thread = query, so every thread works through
                    // ITS OWN survivors while the other lanes do the same for theirs.  Entries are compared on the exact
                    // rationals (i1 u2 > i2 u1: closer; equal ratios: the higher slot first -- the order of the keys
                    // distance bits << 32 | ~slot without forming them).  The list is UNSORTED with its worst entry
                    // tracked in registers: a survivor costs two multiplies against the worst entry and, when it
                    // enters, overwrites it and rescans the k entries for the new worst (independent loads; a sorted
                    // insert is a chain of dependent shared-memory round trips, measured ~1.7 k cycles per survivor).
                    // The list is sorted, and the f64 divide of the reference-exact bits paid, once, when it is
                    // published.  The final list is the top-k of all survivors under a total order.
                    // (The survivor code is kept SMALL and rolled: it runs a few times per tile from a cold instruction
                    // cache, and the four-fold unrolled version of it -- 2800 instructions -- cost ~1.7 k cycles per
                    // survivor whatever the list algorithm was.)
                    if (any_hit) {
                        // bit 8 r + 7 - j of `hits` <-> column 8 r + j of the block
                        uint32_t hits = (hm[3] << 24) | (hm[2] << 16) | (hm[1] << 8) | hm[0];
                        n_hits += __popc(hits);
                        uint32_t* ls = &s.lnslot[ql][0];
                        uint32_t* li = &s.liu[ql][0];
                        bool changed = false;
#pragma unroll 1
                        while (hits) {
                            const uint32_t bit = 31u - (uint32_t)__clz((int)hits);
                            hits &= ~(1u << bit);
                            const uint32_t col = (bit & 24u) | (7u - (bit & 7u));
                            const uint32_t iv = select32(acc, col) >> kTScaleLog2;
                            const uint32_t dl = (uint32_t)(b * 32) + col;
                            const uint32_t pdv = s.pd[a][dl];
                            if (pdv >= kTPadRow)  // a padding row can pass a non-strict bound with i_t = 0
                                continue;
                            const uint64_t row = row_begin + tile * kTD + dl;
                            if (tri && row >= (uint64_t)q0 + (uint32_t)ql)
                                continue;         // triangular form: only the rows in front of the query
                            const uint32_t u = pqv + pdv - iv, uu = u ? u : 1u;  // u == 0: distance 1 = ratio 0
                            const uint32_t ns = ~(uint32_t)(p.slot_base + row);
                            if (llen == kk) {
                                const uint32_t lhs = iv * w_u, rhs = w_i * uu;  // counts <= 1024, unions <= 2048
                                if (lhs < rhs || (lhs == rhs && ns > w_ns))
                                    continue;
                                // replace the worst entry, then find the new worst: independent loads, no chain
                                li[w_pos] = iv | (uu << 16);
                                ls[w_pos] = ns;
                                uint32_t ci = iv, cu = uu, cns = ns, cpos = w_pos;
#pragma unroll 1
                                for (uint32_t e0 = 0; e0 < kk; e0 += 4) {  // four entries per round: their loads overlap
                                    uint32_t eiu[4], ens[4];
#pragma unroll
                                    for (uint32_t t4 = 0; t4 < 4; ++t4) {  // (rows hold kTListK + 1 >= e0 + 4 entries)
                                        eiu[t4] = li[e0 + t4];
                                        ens[t4] = ls[e0 + t4];
                                    }
#pragma unroll
                                    for (uint32_t t4 = 0; t4 < 4; ++t4) {
                                        const uint32_t l2 = (eiu[t4] & 0xFFFFu) * cu, r2 = ci * (eiu[t4] >> 16);
                                        // entry e0 + t4 is farther than the candidate
                                        if (e0 + t4 < kk && (l2 < r2 || (l2 == r2 && ens[t4] > cns)))
                                            ci = eiu[t4] & 0xFFFFu, cu = eiu[t4] >> 16, cns = ens[t4], cpos = e0 + t4;
                                    }
                                }
                                w_i = ci, w_u = cu, w_ns = cns, w_pos = cpos;
                            } else {
                                li[llen] = iv | (uu << 16);
                                ls[llen] = ns;
                                // the worst of a list that is not full yet only matters when it becomes full
                                if (llen == 0 || iv * w_u < w_i * uu || (iv * w_u == w_i * uu && ns > w_ns))
                                    w_i = iv, w_u = uu, w_ns = ns, w_pos = llen;
                                ++llen;
                            }
                            changed = true;
                            ++qn_total;
                            ++n_ins;
                        }
                        if (changed && llen == kk) {
                            s.thr[ql] = make_uint4(w_u + w_i, w_i, w_i * pqv, 0);
                            if (share_bounds && !tri) {  // (triangular form: one CTA per query, nobody to share with)
                                own_bits = exact_distance_bits(w_i, w_u);
                                atomicMin(p.bounds + bound_base + q0 + ql,
                                          ((unsigned long long)own_bits << 32) | ((unsigned long long)chunk << 23) |
                                              ((unsigned long long)w_i << 12) | w_u);
                            }
                        }
                    }
                } else
                if (__any_sync(RB_FULL_MASK, any_hit != 0)) {
                    // rare: rounds of 8 columns from the HIGHEST rows down (round r = columns 8 r .. 8 r + 7); a survivor
                    // always belongs to one of this warp's own 32 queries, so the slow path is warp-local
                    const uint32_t hits = pqv == kXInvalid ? 0u : ((hm[3] << 24) | (hm[2] << 16) | (hm[1] << 8) | hm[0]);
#pragma unroll 1
                    for (int r = 3; r >= 0; --r) {
                        uint32_t hit = (hits >> (8 * r)) & 0xFFu;
                        if (!__any_sync(RB_FULL_MASK, hit != 0))
                            continue;
                        while (hit) {
                            const int j = 7 - (31 - __clz((int)hit));  // bit 7 - j <-> column 8 r + j
                            hit &= ~(0x80u >> j);
                            const uint32_t iv = select32(acc, (uint32_t)(r * 8 + j)) >> kTScaleLog2;
                            uint32_t at = atomicAdd(&s.queue_count[warp], 1u);
                            s.queue[warp][at] = ((uint32_t)ql << 24) | ((uint32_t)(b * 32 + r * 8 + j) << 16) | iv;
                        }
                        __syncwarp();
                        const uint32_t qn = s.queue_count[warp];
                        qn_total += qn;
                        for (uint32_t base = 0; base < qn; base += 32) {
                            uint32_t e = base + lane < qn ? s.queue[warp][base + lane] : 0xFFFFFFFFu;
                            for (uint32_t jj = 0; jj < 32 && base + jj < qn; ++jj) {
                                uint32_t ev = __shfl_sync(RB_FULL_MASK, e, (int)jj);
                                uint32_t eq = ev >> 24, dl = (ev >> 16) & 0xFFu, i = ev & 0xFFFFu;
                                if (s.pd[a][dl] >= kTPadRow)  // a padding row can pass a non-strict bound with i_t = 0
                                    continue;                 // (warp-uniform: every lane looks at the same entry)
                                if (tri && row_begin + tile * kTD + dl >= (uint64_t)q0 + eq)
                                    continue;                 // triangular form: only the rows in front of the query
                                uint32_t u = s.pq[eq] + s.pd[a][dl] - i;
                                uint32_t slot = (uint32_t)(p.slot_base + row_begin + tile * kTD + dl);
                                uint64_t key = ((uint64_t)exact_distance_bits(i, u) << 32) | (uint32_t)~slot;
                                const size_t off = (out_base + q0 + eq) * p.k;  // k > kTListK: lists in global memory
                                uint64_t* lk = p.partial_keys + off;
                                uint32_t* li = p.partial_iu + off;
                                glist_insert(lk, li, p.k, key, i | ((u ? u : 1u) << 16));
                                if (lk[p.k - 1] != ~0ull) {
                                    uint32_t iu = li[p.k - 1];
                                    uint32_t it_ = iu & 0xFFFFu, ut = iu >> 16;
                                    if (lane == 0)
                                        s.thr[eq] = make_uint4(ut + it_, it_, it_ * s.pq[eq], 0);
                                }
                                __syncwarp();
                            }
                        }
                        if (lane == 0)
                            s.queue_count[warp] = 0;
                        __syncwarp();
                    }
                }
                if (dbg) {
                    long long te5 = clock64();
                    dbg_acc[2] += (unsigned long long)(te1 - te0);
                    dbg_acc[3] += (unsigned long long)(te2 - te1);
                    dbg_acc[4] += (unsigned long long)(te3 - te2);
                    dbg_acc[5] += (unsigned long long)(te5 - te3);
                    dbg_acc[7] += qn_total;
                }
            }
            if (lane == 0)
                mbar_arrive(&s.pd_empty[a]);
        }
        if (p.debug_cycles != nullptr && blockIdx.x == 0 && blockIdx.y == 0) {
            atomicAdd(p.debug_cycles + 29, n_hits);
            atomicAdd(p.debug_cycles + 30, n_ins);
        }
        if (smem_lists && pqv != kXInvalid) {  // publish this query's partial list: sorted, reference-exact bits at last
            size_t off = (out_base + q0 + ql) * p.k;
            uint32_t* ls = &s.lnslot[ql][0];
            uint32_t* li = &s.liu[ql][0];
            for (uint32_t j = 1; j < llen; ++j) {  // insertion sort, closest first
                const uint32_t eiu = li[j], ens = ls[j];
                uint32_t pos = j;
                while (pos > 0) {
                    const uint32_t piu = li[pos - 1], pns = ls[pos - 1];
                    const uint32_t lhs = (eiu & 0xFFFFu) * (piu >> 16), rhs = (piu & 0xFFFFu) * (eiu >> 16);
                    if (!(lhs > rhs || (lhs == rhs && ens < pns)))
                        break;
                    li[pos] = piu;
                    ls[pos] = pns;
                    --pos;
                }
                li[pos] = eiu;
                ls[pos] = ens;
            }
            for (uint32_t j = 0; j < kk; ++j) {
                const uint32_t eiu = li[j];
                p.partial_keys[off + j] =
                    j < llen ? (((unsigned long long)exact_distance_bits(eiu & 0xFFFFu, eiu >> 16) << 32) | ls[j]) : ~0ull;
            }
        }
    }
    if (dbg && (warp == kTProdWarp0 || warp == kTMmaWarp || warp == 0)) {
        int role = warp == kTProdWarp0 ? 0 : warp == kTMmaWarp ? 1 : 2;
        p.debug_cycles[role * 10 + 0] = (unsigned long long)(clock64() - dbg_start);
        for (int i = 0; i < 8; ++i)
            p.debug_cycles[role * 10 + 1 + i] = dbg_acc[i];
        p.debug_cycles[31] = n_tiles;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kTMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

}  // namespace rb
