// exact_f4_kernels.cuh -- exact Tanimoto top-k on the FP4 path of the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Same contract and the same structure as exact_topk_i8_kernel_t<0> (exact_i8_kernels.cuh; replaces
// index_gt::search_exact_, usearch/include/usearch/index.hpp:4222-4239, tie rule index.hpp:923-934), with the
// intersection counts formed by `tcgen05.mma kind::mxf4.block_scale` instead of `kind::i8`:
//     inter[q][d] = sum_k bit_k(q) * bit_k(d),   bits as e2m1 values, every block scale = 1.0 (ue8m0 127), f32 accumulate
// An instruction covers K = 64 fingerprint bits at the cycles kind::i8 needs for 32 (scripts/mma_f4_microbench.cu:
// 166.8 vs 166.0 cycles at M128 N256; sums of <= 1024 products of small powers of two are exact in f32), a fingerprint
// unpacks to 512 bytes instead of 1024 (half the producers' stores, half the operand bytes through the shared-memory
// port per pair), and the query operand shrinks to 64 KB, which pays for a 4-stage B ring: one whole database tile in
// flight.  Tensor memory holds two accumulators AND the scale factors, so the tile is 224 rows wide (2 x 224 + 64 of
// the 512 columns).
//   unpack (any fixed bit -> K position works as long as both operands use it): a 32-bit word w becomes the four
//   words of one 16-byte K piece, eight 4-bit elements each --
//     database:  w & 0x11111111 (e2m1 0b0001 = 0.5), w & 0x22222222 (1.0), w & 0x44444444 (2.0), (w >> 3) & 0x11111111 (0.5)
//     query:     the reciprocal values 2.0, 1.0, 0.5, 2.0 -- every matching bit contributes exactly 1.0
// The epilogue's filter is the integer test of the other kernels carried out in f32 (all its terms stay below 2^24).
// Results are bit-identical to exact_topk_kernel and exact_topk_i8_kernel_t (tests/test_gpu_exact.py runs every case
// on all of them).
#pragma once
#include <cuda_fp16.h>

#include "exact_i8_kernels.cuh"

namespace rb {

constexpr int kFQ = 128;               // queries per CTA (UMMA M)
constexpr int kFD = 224;               // database rows per tile (UMMA N)
#ifndef RB_F4_PIECES
#define RB_F4_PIECES 16
#endif
constexpr int kFPieces = RB_F4_PIECES; // 16-byte K pieces (32 fingerprint bits each) per B stage: 16 (8: 2 % slower --
                                       // one fence.proxy.async + hand-off per 4 MMAs instead of per 8)
constexpr int kFSteps = 32 / kFPieces; // stages per database tile
constexpr int kFStages = kFSteps;      // B ring = the 1024 bits of one tile: 2 stages of 512 bits
// TWO epilogue groups of four warps (a warp reaches the TMEM lanes 32 (warp % 4) ..: warps 0-3 and 4-7 both cover all
// 128 queries).  Both groups work on EVERY tile, group 0 on its columns 112..223, group 1 on 0..111, so an accumulator
// stage is held for half an epilogue and the epilogues of consecutive tiles run back to back.  (Four warps took 6.6 k
// cycles per tile and bounded the kernel at 80 ms for 10 k x 10 M; the two groups on ALTERNATE tiles: 66 ms, bounded by
// (MMA + epilogue) / 2 because an accumulator stayed occupied for a whole epilogue.)
constexpr int kFEpiGroups = 2, kFEpiWarps = 4 * kFEpiGroups, kFMmaWarp = kFEpiWarps, kFProdWarp0 = kFMmaWarp + 1;
constexpr int kFProdWarps = kFD / 32;                            // 7 producer warps
constexpr int kFThreads = (kFEpiWarps + 1 + kFProdWarps) * 32;  // 512
constexpr uint32_t kFSBO_A = 32 * kLBO;                  // A holds all 32 K pieces of a row group
constexpr uint32_t kFSBO_B = kFPieces * kLBO;
constexpr uint32_t kFASmemBytes = (kFQ / 8) * kFSBO_A;   // 65536
constexpr uint32_t kFStageBytes = (kFD / 8) * kFSBO_B;   // 28672
constexpr uint32_t kFSfColumn = 2 * kFD;                 // scale factors: TMEM columns [448, 512)
constexpr int kFPdRing = 4;                              // row-popcount side arrays in flight (tiles)
constexpr int kFGroupCols = kFD / kFEpiGroups;           // 112 columns of a tile per epilogue group

struct F4SmemTail {
    uint32_t pq[kFQ];
    __half pdf[kFPdRing][kFD];  // row popcounts as f16 (kTPadRow = 4096 for padding rows; integers <= 2048 and 4096 are
                                // exact): the filter's operand -- eight per 16-byte broadcast read, the shared-memory
                                // data pipe is what bounds this kernel
    uint4 thr[kFEpiGroups][kFQ];       // the filter thresholds of thread (group, query): written by that thread only
    uint32_t worst[kFEpiGroups][kFQ];  // i | u << 16 of the worst entry of that thread's FULL list, 0 before (read by
                                       // the other group: the k-th best of the union is at least as good as either)
    uint32_t llen[kFEpiGroups][kFQ];
    uint32_t queue_count[4];
    // per-(group, query) partial lists (k <= kTListK), merged when they are published; rows padded by one entry: the 32
    // lanes of a warp touching the same column hit distinct banks.  For k > kTListK the lists live in global memory,
    // group 0 works alone, and its survivor queues (4 x 256 words) lie over lnslot[1].
    uint32_t lnslot[kFEpiGroups][kFQ][kTListK + 1];  // ~slot of each entry
    uint32_t liu[kFEpiGroups][kFQ][kTListK + 1];     // i | u << 16 of each entry
    uint32_t tmem_base;
    unsigned long long full[kFStages], empty[kFStages], tmem_full[2], tmem_empty[2], pd_full[kFPdRing], pd_empty[kFPdRing];
};
static_assert(sizeof(uint32_t) * kFQ * (kTListK + 1) >= sizeof(uint32_t) * kTQueueCap, "the queues fit over lnslot[1]");
constexpr size_t kF4SmemBytes = kFASmemBytes + kFStages * kFStageBytes + sizeof(F4SmemTail) + 128;
static_assert(kF4SmemBytes <= 232448, "227 KB of shared memory per CTA");

// instruction descriptor of the block-scaled kinds (cute/arch/mma_sm100_desc.hpp: InstrDescriptorBlockScaled): a / b
// format E2M1 = 1 at [7,10) / [10,13), both K-major, N >> 3 at [17,23), scale format UE8M0 = 1 at bit 23, M >> 4 at
// [24,29), scale-factor ids 0, K = 64 (bit 31 clear)
constexpr uint32_t kIdescF4 = (1u << 7) | (1u << 10) | ((uint32_t)(kFD >> 3) << 17) | (1u << 23) | ((uint32_t)(kFQ >> 4) << 24);

__device__ __forceinline__ void tc_mma_f4(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t sfa,
                                          uint32_t sfb, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
        : "memory");
}
__device__ __forceinline__ uint4 unpack_f4_db(uint32_t w) {
    return make_uint4(w & 0x11111111u, w & 0x22222222u, w & 0x44444444u, (w >> 3) & 0x11111111u);
}
__device__ __forceinline__ uint4 unpack_f4_query(uint32_t w) {
    return make_uint4((w & 0x11111111u) << 2, w & 0x22222222u, (w & 0x44444444u) >> 2, ((w >> 3) & 0x11111111u) << 2);
}

// grid = (query tiles of 128, database chunks) or, in the triangular form, (query tiles, levels); ExactParams as for
// the other kernels with chunk_rows a multiple of kFD
__global__ void __launch_bounds__(kFThreads, 1) exact_topk_f4_kernel(ExactParams p) {
    extern __shared__ __align__(1024) unsigned char raw[];
    unsigned char* smem_a = raw;
    unsigned char* smem_b = raw + kFASmemBytes;
    F4SmemTail& s = *reinterpret_cast<F4SmemTail*>(raw + kFASmemBytes + kFStages * kFStageBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool tri = p.tri != 0;
    const uint32_t qcount = tri ? p.tri_cnt[blockIdx.y] : p.query_count;
    const uint32_t q0 = (tri ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * kFQ;
    if (tri && q0 >= qcount)
        return;  // uniform: the launch is sized for the widest level
    const uint32_t chunk = tri ? 0u : blockIdx.y;
    const size_t out_base = tri ? (size_t)p.tri_off[blockIdx.y] : (size_t)chunk * p.query_count;
    const size_t bound_base = tri ? (size_t)p.tri_off[blockIdx.y] : 0;
    const uint64_t row_begin = (uint64_t)chunk * p.chunk_rows;
    const uint64_t row_end = tri ? (qcount < q0 + kFQ - 1 ? qcount : q0 + kFQ - 1)
                                 : (row_begin + p.chunk_rows < p.rows ? row_begin + p.chunk_rows : p.rows);
    const uint64_t n_tiles = row_begin < row_end ? (row_end - row_begin + kFD - 1) / kFD : 0;

    // ---- the CTA's partial lists start empty
    {
        uint32_t nq = q0 >= qcount ? 0u : qcount - q0 < (uint32_t)kFQ ? qcount - q0 : (uint32_t)kFQ;
        size_t base = (out_base + q0) * p.k;
        for (size_t i = tid; i < (size_t)nq * p.k; i += kFThreads)
            p.partial_keys[base + i] = ~0ull;
    }
    if (n_tiles == 0)
        return;  // uniform

    // ---- query tile: popcounts, thresholds, and the resident unpacked A operand (one piece per 32-bit word)
    for (int r = tid; r < kFQ; r += kFThreads) {
        uint32_t q = q0 + r, pc = 0;
        if (q < qcount)
            for (int w = 0; w < 8; ++w)
                pc += popc4(__ldg(p.queries + (size_t)q * 8 + w));
        s.pq[r] = q < qcount ? pc : kXInvalid;
        for (int g = 0; g < kFEpiGroups; ++g) {
            s.thr[g][r] = make_uint4(0, 1, 0, 2048);  // list not full: every valid row (|d| < 2048) passes, padding rows do not
            s.worst[g][r] = 0;
            s.llen[g][r] = 0;
        }
    }
    {
        const uint32_t* q32 = reinterpret_cast<const uint32_t*>(p.queries);
        for (int idx = tid; idx < kFQ * 32; idx += kFThreads) {
            int r = idx & (kFQ - 1), piece = idx >> 7;
            uint32_t q = q0 + r;
            uint32_t w = q < qcount ? __ldg(q32 + (size_t)q * 32 + piece) : 0u;
            *reinterpret_cast<uint4*>(smem_a + (r >> 3) * kFSBO_A + piece * kLBO + (r & 7) * 16) = unpack_f4_query(w);
        }
    }
    if (tid == 0) {
        for (int i = 0; i < kFStages; ++i) {
            mbar_init(&s.full[i], kFProdWarps);  // one arrival per producer warp
            mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s.tmem_full[i], 1);
            mbar_init(&s.tmem_empty[i], kFEpiWarps);
        }
        for (int i = 0; i < kFPdRing; ++i) {
            mbar_init(&s.pd_full[i], kFProdWarps);
            mbar_init(&s.pd_empty[i], kFEpiWarps);
        }
        for (int i = 0; i < 4; ++i)
            s.queue_count[i] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kFMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s.tmem_base))
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();  // the A operand was written through the generic proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = s.tmem_base;
    if (warp < 4) {
        // every scale factor is 1.0: thread = TMEM lane, columns [448, 512) = 0x7F7F7F7F (ue8m0 127 in every byte, so the
        // instruction reads 1.0 whatever the layout of the factors inside those columns)
        const uint32_t one = 0x7F7F7F7Fu;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + kFSfColumn + (uint32_t)(h * 32);
            asm volatile(
                "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
                "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(one)
                : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp >= kFProdWarp0) {
        // ================================================================ producers: one database row per thread
        const int pr = tid - kFProdWarp0 * 32;  // 0 .. 223
        unsigned char* my_b = smem_b + (pr >> 3) * kFSBO_B + (pr & 7) * 16;
        uint4 v[8];
        {
            const uint64_t row0 = row_begin + (n_tiles - 1) * kFD + pr;
#pragma unroll
            for (int c = 0; c < 8; ++c)
                v[c] = row0 < row_end ? __ldg(p.db + row0 * 8 + c) : make_uint4(0, 0, 0, 0);
        }
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;  // highest rows first (see exact_kernels.cuh)
            const uint64_t row = row_begin + tile * kFD + pr;
            uint4 vn[8];
            {
                const uint64_t rown = row - kFD;  // tile - 1 (only used when t + 1 < n_tiles; always < row_end then)
                const bool more = t + 1 < n_tiles;
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    vn[c] = more ? __ldg(p.db + rown * 8 + c) : make_uint4(0, 0, 0, 0);
            }
            uint32_t pc = 0;
#pragma unroll
            for (int h = 0; h < kFSteps; ++h) {  // stage h of the tile = fingerprint words kFPieces h .. kFPieces (h + 1) - 1
                const uint64_t g = t * kFSteps + h;
                const int st = (int)(g % kFStages);
                unsigned char* dst = my_b + st * kFStageBytes;
#pragma unroll
                for (int sub = 0; sub < kFPieces / 8; ++sub) {  // eight words (two uint4) at a time
                    const int c4 = h * (kFPieces / 4) + 2 * sub;
                    pc += popc4(v[c4]) + popc4(v[c4 + 1]);
                    const uint32_t w[8] = {v[c4].x, v[c4].y, v[c4].z, v[c4].w, v[c4 + 1].x, v[c4 + 1].y, v[c4 + 1].z, v[c4 + 1].w};
                    uint4 u[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j)  // K piece kFPieces h + 8 sub + j  <->  A piece of the same index
                        u[j] = unpack_f4_db(w[j]);
                    if (sub == 0)
                        mbar_wait(&s.empty[st], (uint32_t)(((g / kFStages) & 1) ^ 1));
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(dst + (sub * 8 + j) * kLBO) = u[j];
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(&s.full[st]);
            }
#pragma unroll
            for (int c = 0; c < 8; ++c)
                v[c] = vn[c];
            const int a4 = (int)(t % kFPdRing);
            mbar_wait(&s.pd_empty[a4], (uint32_t)(((t / kFPdRing) & 1) ^ 1));  // epilogue of tile t - 4 done with it
            s.pdf[a4][pr] = __float2half_rn((float)(row < row_end ? pc : kTPadRow));
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&s.pd_full[a4]);
        }
    } else if (warp == kFMmaWarp) {
        // ================================================================ MMA issuer
        const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), kLBO, kFSBO_A);
        const uint64_t b_desc0 = make_smem_desc(smem_u32(smem_b), kLBO, kFSBO_B);
        const uint32_t sfa = tmem_base + kFSfColumn, sfb = tmem_base + kFSfColumn + 32;
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const int a = (int)(t & 1);
            mbar_wait(&s.tmem_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1));
            tc_fence_after();
            for (int c = 0; c < kFSteps; ++c) {
                const uint64_t g = t * kFSteps + c;
                const int st = (int)(g % kFStages);
                mbar_wait(&s.full[st], (uint32_t)((g / kFStages) & 1));
                tc_fence_after();
                if (lane == 0) {
#pragma unroll
                    for (int j = 0; j < kFPieces / 2; ++j) {
                        // K = 64 elements per instruction = two 16-byte K pieces
                        const uint64_t ad = a_desc0 + (uint64_t)(((c * kFPieces + j * 2) * kLBO) >> 4);
                        const uint64_t bd = b_desc0 + (uint64_t)((st * kFStageBytes + j * 2 * kLBO) >> 4);
                        tc_mma_f4(tmem_base + (uint32_t)a * kFD, ad, bd, kIdescF4, sfa, sfb, (uint32_t)((c | j) != 0));
                    }
                    tc_commit(&s.empty[st]);
                    if (c == kFSteps - 1)
                        tc_commit(&s.tmem_full[a]);
                }
                __syncwarp();
            }
        }
    } else {
        // ================================================================ epilogue: thread = (group, one query of the tile)
        const int grp = warp >> 2, wq = warp & 3;  // group = column half of every tile; wq = TMEM lane quadrant
        const int ql = wq * 32 + lane;
        const uint32_t pqv = s.pq[ql];
        const bool smem_lists = p.k <= (uint32_t)kTListK;
        const bool share_bounds = smem_lists && p.bounds != nullptr;
        // k > kTListK: one list per query in GLOBAL memory -- group 0 takes both column halves, group 1 only keeps the
        // barrier counts right
        const bool both_halves = !smem_lists;
        const int groups = smem_lists ? kFEpiGroups : 1;
        uint32_t (*queue)[kTQueueCap / 4] = reinterpret_cast<uint32_t (*)[kTQueueCap / 4]>(&s.lnslot[1][0][0]);
        const uint32_t kk = p.k;
        // this thread's own list: length, and its worst entry once it is full (counts, ~slot, position)
        uint32_t llen = 0, w_i = 0, w_u = 1, w_ns = 0, w_pos = 0;
        for (uint64_t t = 0; t < n_tiles; ++t) {
            const uint64_t tile = n_tiles - 1 - t;
            const int a = (int)(t & 1), a4 = (int)(t % kFPdRing);
            if (both_halves && grp == 1) {  // nothing to do but to release what group 0 does not wait for
                mbar_wait(&s.tmem_full[a], (uint32_t)((t >> 1) & 1));
                if (lane == 0) {
                    mbar_arrive(&s.tmem_empty[a]);
                    mbar_arrive(&s.pd_empty[a4]);
                }
                continue;
            }
            mbar_wait(&s.pd_full[a4], (uint32_t)((t / kFPdRing) & 1));
            mbar_wait(&s.tmem_full[a], (uint32_t)((t >> 1) & 1));
            tc_fence_after();
            // The filter bound of this tile: the tightest of (1) the worst entry of this thread's own full list -- strict:
            // its rows arrive from the highest slot down, so a row that merely ties loses; (2) the worst entry of the OTHER
            // group's full list -- the k-th best of the union is at least as good as it; non-strict, that list's slots are
            // interleaved with ours; (3) the bound word other CTAs published (see exact_topk_i8_kernel_t), strict for a
            // lower chunk, non-strict for a higher one.  Ratios are compared by cross-multiplication (counts <= 1024,
            // unions <= 2048).  Which survivors take the slow path depends on timing; the result does not.
            if (pqv != kXInvalid) {
                bool have = llen == kk;
                uint32_t bi = w_i, bu = w_u, bw = 0;
                if (groups > 1) {
                    const uint32_t ow = *reinterpret_cast<volatile uint32_t*>(&s.worst[grp ^ 1][ql]);
                    const uint32_t oi = ow & 0xFFFFu, ou = ow >> 16;
                    if (ow && (!have || oi * bu > bi * ou))
                        have = true, bi = oi, bu = ou, bw = 1;
                }
                if (share_bounds) {
                    const unsigned long long g = __ldcg(p.bounds + bound_base + q0 + ql);
                    if (g != ~0ull) {
                        const uint32_t holder = (uint32_t)(g >> 23) & 0x1FFu;
                        const uint32_t it_ = (uint32_t)(g >> 12) & 0x7FFu, ut = (uint32_t)g & 0xFFFu;
                        // a seed of the triangular form only matters until a list is full (every entry passed it)
                        if (tri ? llen < kk && !have : (!have || it_ * bu > bi * ut))
                            have = true, bi = it_, bu = ut, bw = chunk < holder ? 0u : 1u;
                    }
                }
                if (have)
                    s.thr[grp][ql] = make_uint4(bu + bi, bi, bi * pqv, bw);
            }
            // This group's 112 columns in 32-column blocks from the HIGHEST rows down: columns base + 80, 48, 16, and the
            // block at base + 0 of which only the lower 16 columns are new (the read is 32 wide; the rest is masked).
            // (k > kTListK: group 0 walks both halves, the upper one first.)
            const int n_blocks = both_halves ? 8 : 4;
            float neg_a = 0.f, bw = 0.f, y = 0.f;  // the filter's constants of this tile (see below)
            bool reload_thr = true;
#pragma unroll 1
            for (int bi = 0; bi < n_blocks; ++bi) {
                const int base = both_halves ? (bi < 4 ? kFGroupCols : 0) : (grp == 0 ? kFGroupCols : 0);
                const int bq = bi & 3;
                const int c0 = base + (bq < 3 ? 80 - 32 * bq : 0);       // first column of the block in the tile
                const uint32_t new_cols = bq < 3 ? 0xFFFFFFFFu : 0x0000FFFFu;  // hits bit 8 r + 7 - j <-> column 8 r + j
                uint32_t acc[32];  // f32 bit patterns of the counts
                const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a * kFD + c0);
                RB_TMEM_LD32(acc, taddr);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (bi == n_blocks - 1) {  // the accumulator stage is free as soon as the last block is in registers
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0)
                        mbar_arrive(&s.tmem_empty[a]);
                }
                // the integer filter in f32: the pair passes iff  dlt = (y |d| + B - w) - i A  is negative; every term is
                // an integer below 2^24, so the f32 arithmetic is exact and +0 (not -0) stands for "equal"
                uint32_t hm[4] = {0, 0, 0, 0};
                {
                    // (the thresholds change only at the start of a tile and when this thread's own list does -- or, for
                    // k > kTListK, when another lane of the warp updates them: read them again only then)
                    if (reload_thr || !smem_lists) {
                        const uint4 th = s.thr[smem_lists ? grp : 0][ql];
                        neg_a = -(float)th.x, bw = (float)(int)(th.z - th.w), y = (float)th.y;
                        reload_thr = false;
                    }
                    const uint4* pd8 = reinterpret_cast<const uint4*>(&s.pdf[a4][c0]);  // eight f16 per read
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint4 pv = pd8[g];
                        const uint32_t pw[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
                        for (int h2 = 0; h2 < 4; ++h2) {
                            const float2 pf = __half22float2(*reinterpret_cast<const __half2*>(&pw[h2]));
                            const float pdv[2] = {pf.x, pf.y};
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int col = g * 8 + h2 * 2 + e;
                                const float r1 = __fmaf_rn(y, pdv[e], bw);
                                const float dlt = __fmaf_rn(__uint_as_float(acc[col]), neg_a, r1);
                                hm[col >> 3] = __funnelshift_l(__float_as_uint(dlt), hm[col >> 3], 1);
                            }
                        }
                    }
                }
                const uint32_t all_hits = pqv == kXInvalid ? 0u : (((hm[3] << 24) | (hm[2] << 16) | (hm[1] << 8) | hm[0]) & new_cols);
                const uint32_t any_hit = all_hits;
                if (smem_lists) {
                    if (any_hit) {
                        // bit 8 r + 7 - j of `hits` <-> column 8 r + j of the block (exact_topk_i8_kernel_t has the notes)
                        uint32_t hits = all_hits;
                        uint32_t* ls = &s.lnslot[grp][ql][0];
                        uint32_t* li = &s.liu[grp][ql][0];
                        bool changed = false;
#pragma unroll 1
                        while (hits) {
                            const uint32_t bit = 31u - (uint32_t)__clz((int)hits);
                            hits &= ~(1u << bit);
                            const uint32_t col = (bit & 24u) | (7u - (bit & 7u));
                            const uint32_t iv = (uint32_t)__float2uint_rn(__uint_as_float(select32(acc, col)));
                            const uint32_t dl = (uint32_t)c0 + col;
                            const uint32_t pdv = (uint32_t)__half2float(s.pdf[a4][dl]);
                            if (pdv >= kTPadRow)  // a padding row can pass a non-strict bound with i_t = 0
                                continue;
                            const uint64_t row = row_begin + tile * kFD + dl;
                            if (tri && row >= (uint64_t)q0 + (uint32_t)ql)
                                continue;         // triangular form: only the rows in front of the query
                            const uint32_t u = pqv + pdv - iv, uu = u ? u : 1u;  // u == 0: distance 1 = ratio 0
                            const uint32_t ns = ~(uint32_t)(p.slot_base + row);
                            if (llen == kk) {
                                const uint32_t lhs = iv * w_u, rhs = w_i * uu;  // counts <= 1024, unions <= 2048
                                if (lhs < rhs || (lhs == rhs && ns > w_ns))
                                    continue;
                                li[w_pos] = iv | (uu << 16);
                                ls[w_pos] = ns;
                                uint32_t ci = iv, cu = uu, cns = ns, cpos = w_pos;
#pragma unroll 1
                                for (uint32_t e0 = 0; e0 < kk; e0 += 4) {
                                    uint32_t eiu[4], ens[4];
#pragma unroll
                                    for (uint32_t t4 = 0; t4 < 4; ++t4) {
                                        eiu[t4] = li[e0 + t4];
                                        ens[t4] = ls[e0 + t4];
                                    }
#pragma unroll
                                    for (uint32_t t4 = 0; t4 < 4; ++t4) {
                                        const uint32_t l2 = (eiu[t4] & 0xFFFFu) * cu, r2 = ci * (eiu[t4] >> 16);
                                        if (e0 + t4 < kk && (l2 < r2 || (l2 == r2 && ens[t4] > cns)))
                                            ci = eiu[t4] & 0xFFFFu, cu = eiu[t4] >> 16, cns = ens[t4], cpos = e0 + t4;
                                    }
                                }
                                w_i = ci, w_u = cu, w_ns = cns, w_pos = cpos;
                            } else {
                                li[llen] = iv | (uu << 16);
                                ls[llen] = ns;
                                if (llen == 0 || iv * w_u < w_i * uu || (iv * w_u == w_i * uu && ns > w_ns))
                                    w_i = iv, w_u = uu, w_ns = ns, w_pos = llen;
                                ++llen;
                            }
                            changed = true;
                        }
                        if (changed && llen == kk) {
                            // own list tighter: its worst entry becomes this thread's bound if it beats the one in use
                            const uint4 cur = s.thr[grp][ql];  // (x - y, y) = (u, i) of the bound in use; y = 1, x = 0 before any
                            if (cur.x == 0 || w_i * (cur.x - cur.y) >= cur.y * w_u) {
                                s.thr[grp][ql] = make_uint4(w_u + w_i, w_i, w_i * pqv, 0);
                                reload_thr = true;
                            }
                            *reinterpret_cast<volatile uint32_t*>(&s.worst[grp][ql]) = w_i | (w_u << 16);
                            if (share_bounds && !tri)
                                atomicMin(p.bounds + bound_base + q0 + ql,
                                          ((unsigned long long)exact_distance_bits(w_i, w_u) << 32) |
                                              ((unsigned long long)chunk << 23) | ((unsigned long long)w_i << 12) | w_u);
                        }
                    }
                } else if (__any_sync(RB_FULL_MASK, any_hit != 0)) {
                    // k > kTListK: lists in global memory, warp-local slow path (rounds of 8 columns, highest rows first)
                    const uint32_t hits = all_hits;
#pragma unroll 1
                    for (int r = 3; r >= 0; --r) {
                        uint32_t hit = (hits >> (8 * r)) & 0xFFu;
                        if (!__any_sync(RB_FULL_MASK, hit != 0))
                            continue;
                        while (hit) {
                            const int j = 7 - (31 - __clz((int)hit));  // bit 7 - j <-> column 8 r + j
                            hit &= ~(0x80u >> j);
                            const uint32_t iv = (uint32_t)__float2uint_rn(__uint_as_float(select32(acc, (uint32_t)(r * 8 + j))));
                            uint32_t at = atomicAdd(&s.queue_count[wq], 1u);
                            queue[wq][at] = ((uint32_t)ql << 24) | ((uint32_t)(c0 + r * 8 + j) << 16) | iv;
                        }
                        __syncwarp();
                        const uint32_t qn = s.queue_count[wq];
                        for (uint32_t base = 0; base < qn; base += 32) {
                            uint32_t e = base + lane < qn ? queue[wq][base + lane] : 0xFFFFFFFFu;
                            for (uint32_t jj = 0; jj < 32 && base + jj < qn; ++jj) {
                                uint32_t ev = __shfl_sync(RB_FULL_MASK, e, (int)jj);
                                uint32_t eq = ev >> 24, dl = (ev >> 16) & 0xFFu, i = ev & 0xFFFFu;
                                const uint32_t pdq = (uint32_t)__half2float(s.pdf[a4][dl]);
                                if (pdq >= kTPadRow)
                                    continue;
                                if (tri && row_begin + tile * kFD + dl >= (uint64_t)q0 + eq)
                                    continue;
                                uint32_t u = s.pq[eq] + pdq - i;
                                uint32_t slot = (uint32_t)(p.slot_base + row_begin + tile * kFD + dl);
                                uint64_t key = ((uint64_t)exact_distance_bits(i, u) << 32) | (uint32_t)~slot;
                                const size_t off = (out_base + q0 + eq) * p.k;
                                uint64_t* lk = p.partial_keys + off;
                                uint32_t* li = p.partial_iu + off;
                                glist_insert(lk, li, p.k, key, i | ((u ? u : 1u) << 16));
                                if (lk[p.k - 1] != ~0ull) {
                                    uint32_t iu = li[p.k - 1];
                                    uint32_t it_ = iu & 0xFFFFu, ut = iu >> 16;
                                    if (lane == 0)
                                        s.thr[0][eq] = make_uint4(ut + it_, it_, it_ * s.pq[eq], 0);
                                }
                                __syncwarp();
                            }
                        }
                        if (lane == 0)
                            s.queue_count[wq] = 0;
                        __syncwarp();
                    }
                }
            }
            if (lane == 0)
                mbar_arrive(&s.pd_empty[a4]);
        }
        if (smem_lists) {
            // publish: every thread sorts its own list (closest first), the two groups meet, group 0 merges the two
            // sorted lists of its query and pays the f64 divide of the reference-exact bits for the k entries kept
            uint32_t* ls = &s.lnslot[grp][ql][0];
            uint32_t* li = &s.liu[grp][ql][0];
            for (uint32_t j = 1; j < llen; ++j) {  // insertion sort
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
            s.llen[grp][ql] = llen;
            asm volatile("bar.sync 1, %0;" ::"r"(kFEpiWarps * 32) : "memory");  // the eight epilogue warps only
            if (grp == 0 && pqv != kXInvalid) {
                const size_t off = (out_base + q0 + ql) * p.k;
                const uint32_t n0 = llen, n1 = s.llen[1][ql];
                const uint32_t* ls1 = &s.lnslot[1][ql][0];
                const uint32_t* li1 = &s.liu[1][ql][0];
                uint32_t i0 = 0, i1 = 0;
                for (uint32_t j = 0; j < kk; ++j) {
                    uint32_t eiu = 0, ens = 0;
                    bool any = false;
                    if (i0 < n0 && i1 < n1) {
                        const uint32_t a_iu = li[i0], b_iu = li1[i1];
                        const uint32_t lhs = (a_iu & 0xFFFFu) * (b_iu >> 16), rhs = (b_iu & 0xFFFFu) * (a_iu >> 16);
                        const bool first = lhs > rhs || (lhs == rhs && ls[i0] < ls1[i1]);  // closer, or the higher slot
                        eiu = first ? a_iu : b_iu, ens = first ? ls[i0] : ls1[i1];
                        i0 += first ? 1u : 0u, i1 += first ? 0u : 1u;
                        any = true;
                    } else if (i0 < n0) {
                        eiu = li[i0], ens = ls[i0], ++i0, any = true;
                    } else if (i1 < n1) {
                        eiu = li1[i1], ens = ls1[i1], ++i1, any = true;
                    }
                    p.partial_keys[off + j] =
                        any ? (((unsigned long long)exact_distance_bits(eiu & 0xFFFFu, eiu >> 16) << 32) | ens) : ~0ull;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kFMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

}  // namespace rb
