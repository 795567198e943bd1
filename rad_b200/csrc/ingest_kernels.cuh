// ingest_kernels.cuh -- fingerprint ingest pre-passes on the device: exact duplicates.
//
// The reference has no such pass: identical fingerprints (stereo-isomers under an achiral Morgan fingerprint, salts of
// one parent) are inserted like any other row and form d = 0 cliques that fill the M links of their members and
// inflate every ef-bounded search that touches them (usearch/include/usearch/index.hpp:2734-2834 treats every row
// alike).  rb_dedup_device names, for every row, the LOWEST row with the same 128 bytes:
//   row_hash_kernel        one 64-bit hash per row (8 lanes per row, one 128-byte line)
//   cub radix sort         (hash, row) pairs, stable -> rows ascending inside a run of equal hashes (library code)
//   run_start_kernel + cub inclusive max-scan   first sorted position of every run
//   representative_kernel  row -> first row of its run with IDENTICAL BYTES (a 64-bit collision between different
//                          fingerprints only costs extra compares; the answer never depends on the hash)
#pragma once
#include "device_common.cuh"

namespace rb {

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33, x *= 0xff51afd7ed558ccdull, x ^= x >> 33, x *= 0xc4ceb9fe1a85ec53ull, x ^= x >> 33;
    return x;
}

__global__ void __launch_bounds__(256) row_hash_kernel(const uint4* __restrict__ fp, uint64_t rows,
                                                       uint64_t* __restrict__ hashes, uint32_t* __restrict__ ids) {
    const uint32_t sub = threadIdx.x & 7;
    for (uint64_t row = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;; row += ((uint64_t)gridDim.x * blockDim.x) >> 3) {
        if (!__any_sync(RB_FULL_MASK, row < rows))  // a warp leaves as a whole: every lane takes part in the shuffles
            break;
        uint64_t h = 0;
        if (row < rows) {
            const uint4 v = __ldg(fp + row * 8 + sub);
            h = mix64(((uint64_t)v.y << 32 | v.x) + 0x9E3779B97F4A7C15ull * (2 * sub + 1));
            h ^= mix64(((uint64_t)v.w << 32 | v.z) + 0x9E3779B97F4A7C15ull * (2 * sub + 2));
        }
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 1);
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 2);
        h ^= __shfl_xor_sync(RB_FULL_MASK, h, 4);
        if (row < rows && sub == 0) {
            hashes[row] = mix64(h);
            ids[row] = (uint32_t)row;
        }
    }
}

// start[p] = p where a run of equal hashes begins, 0 elsewhere (an inclusive max-scan then spreads it over the run)
__global__ void __launch_bounds__(256) run_start_kernel(const uint64_t* __restrict__ sorted_hashes, uint64_t rows,
                                                        uint32_t* __restrict__ start) {
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < rows; p += (uint64_t)gridDim.x * blockDim.x)
        start[p] = (p && sorted_hashes[p] != sorted_hashes[p - 1]) ? (uint32_t)p : 0u;
}

// 8 lanes per sorted position: the first row of the run whose bytes equal this row's
__global__ void __launch_bounds__(256) representative_kernel(const uint4* __restrict__ fp, uint64_t rows,
                                                             const uint32_t* __restrict__ sorted_ids,
                                                             const uint32_t* __restrict__ run_start,
                                                             uint32_t* __restrict__ representative) {
    const uint32_t sub = threadIdx.x & 7;
    for (uint64_t p = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;; p += ((uint64_t)gridDim.x * blockDim.x) >> 3) {
        if (!__any_sync(RB_FULL_MASK, p < rows))
            break;
        const bool on = p < rows;
        const uint32_t me = on ? sorted_ids[p] : 0u;
        const uint4 mine = on ? __ldg(fp + (size_t)me * 8 + sub) : make_uint4(0, 0, 0, 0);
        uint64_t c = on ? run_start[p] : 0;
        uint32_t rep = me;
        bool done = !on;
        // the groups of a warp walk runs of different lengths: loop while any group still looks
        while (__any_sync(RB_FULL_MASK, !done)) {
            const bool look = !done && c < p;
            const uint32_t other = look ? sorted_ids[c] : me;
            const uint4 v = look ? __ldg(fp + (size_t)other * 8 + sub) : mine;
            bool same = v.x == mine.x && v.y == mine.y && v.z == mine.z && v.w == mine.w;
            uint32_t agree = __ballot_sync(RB_FULL_MASK, same);
            same = ((agree >> (threadIdx.x & 24)) & 0xFFu) == 0xFFu;  // all 8 lanes of this group
            if (!done) {
                if (!look) {
                    done = true;  // reached itself: the first of its kind
                } else if (same) {
                    rep = other;
                    done = true;
                } else {
                    ++c;
                }
            }
        }
        if (on && sub == 0)
            representative[me] = rep;
    }
}

}  // namespace rb
