// synth_kernels.cuh -- device twin of rad_b200/synth.py: counter-based clustered fingerprints.
// Bit b of row r depends only on (seed, r, b): h = splitmix64(seed * GOLD + (r << 10 | b)); a centre bit survives
// when (h & 0xFFFF) >= KEEP_T, a fresh bit appears when ((h >> 16) & 0xFFFF) < NOISE_T.  Bits are packed MSB-first
// inside each byte (numpy.packbits), bytes little-endian inside the 32-bit words the kernels read.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rb {

constexpr uint64_t kSynthGold = 0x9E3779B97F4A7C15ull;
constexpr uint32_t kSynthKeep = 19661;
constexpr uint32_t kSynthNoise = 983;

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x ^= x >> 30;
    x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}

// The generator's optional modes (rad_b200/synth.py: clustered_fingerprints): rows_per_centre > 0 = the analogue-series
// order (row r belongs to centre (r / rows_per_centre) mod C: the rows of a cluster are contiguous); dup_fold > 1 = runs
// of dup_fold exact duplicates inside [dup_start, dup_start + dup_rows).
struct SynthModes {
    uint64_t rows_per_centre, dup_fold, dup_start, dup_rows;
};

__device__ __forceinline__ uint32_t synth_word(uint64_t row, uint32_t w, uint64_t seed, const uint32_t* centres,
                                               uint32_t centre_count, SynthModes md) {
    const uint64_t base = seed * kSynthGold;
    if (md.dup_fold > 1 && row >= md.dup_start && row < md.dup_start + md.dup_rows)
        row = md.dup_start + (row - md.dup_start) / md.dup_fold * md.dup_fold;
    uint32_t assign = md.rows_per_centre ? (uint32_t)((row / md.rows_per_centre) % centre_count)
                                         : (uint32_t)(splitmix64(base + row + 0xA55Aull) % centre_count);
    uint32_t cw = centres[(size_t)assign * 32 + w];
    uint32_t word = 0;
#pragma unroll 4
    for (uint32_t t = 0; t < 32; ++t) {
        uint64_t h = splitmix64(base + ((row << 10) | (uint64_t)(w * 32 + t)));
        uint32_t keep = (uint32_t)(h & 0xFFFFu) >= kSynthKeep;
        uint32_t noise = (uint32_t)((h >> 16) & 0xFFFFu) < kSynthNoise;
        uint32_t shift = 8 * (t >> 3) + 7 - (t & 7);
        uint32_t bit = (((cw >> shift) & 1u) & keep) | noise;
        word |= bit << shift;
    }
    return word;
}

// one thread per 32-bit word: out is rows x 32 words
__global__ void synth_kernel(uint32_t* out, uint64_t first_row, uint64_t rows, uint64_t seed, const uint32_t* centres,
                             uint32_t centre_count, int edge_cases, SynthModes md) {
    uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= rows * 32)
        return;
    uint64_t row = first_row + gid / 32;
    uint32_t w = (uint32_t)(gid % 32);
    uint32_t word;
    if (edge_cases && row == 0)
        word = 0;
    else
        word = synth_word((edge_cases && row == 1) ? 2 : row, w, seed, centres, centre_count, md);
    out[gid] = word;
}

}  // namespace rb
