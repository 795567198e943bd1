// gather_microbench.cu -- what random 128-byte gathers can reach on this GPU (the access pattern of the HNSW build:
// one fingerprint = one 128-byte line at a random row of a 1.28 GB .. 12.8 GB array).  Development aid, not the bench.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_microbench scripts/gather_microbench.cu
//   ./gather_microbench [rows=10000000]
// Every group of 8 lanes fetches one line (LDG.128 per lane), UNROLL lines in flight per group, WARPS warps per SM.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16, x *= 0x7feb352du, x ^= x >> 15, x *= 0x846ca68bu, x ^= x >> 16;
    return x;
}

template <int UNROLL>
__global__ void gather_kernel(const uint4* __restrict__ fp, uint32_t rows, uint32_t rounds, uint32_t* sink) {
    const uint32_t lane = threadIdx.x & 31, sub = lane & 7;
    const uint32_t group = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    uint32_t acc = 0;
    for (uint32_t r = 0; r < rounds; ++r) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t row = mix(group * 0x9E3779B1u + r * UNROLL + u) % rows;
            v[u] = __ldg(fp + (size_t)row * 8 + sub);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            acc += __popc(v[u].x) + __popc(v[u].y) + __popc(v[u].z) + __popc(v[u].w);
    }
    if (acc == 0x12345678u)
        *sink = acc;
}

template <int UNROLL>
static void run(const uint4* fp, uint32_t rows, int warps_per_sm, int sms, uint32_t* sink) {
    const int threads = 128, blocks = sms * warps_per_sm / 4;
    const uint32_t rounds = 4096 / UNROLL;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    gather_kernel<UNROLL><<<blocks, threads>>>(fp, rows, rounds / 8, sink);
    cudaEventRecord(e0);
    gather_kernel<UNROLL><<<blocks, threads>>>(fp, rows, rounds, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double lines = (double)blocks * threads / 8 * rounds * UNROLL;
    printf("rows %u  warps/SM %2d  lines in flight per warp %2d  : %7.1f GB/s  (%.2f ms)\n", rows, warps_per_sm,
           4 * UNROLL, lines * 128 / ms / 1e6, ms);
}

int main(int argc, char** argv) {
    const uint32_t rows = argc > 1 ? (uint32_t)atoll(argv[1]) : 10000000u;
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    uint4* fp;
    uint32_t* sink;
    cudaMalloc(&fp, (size_t)rows * 128);
    cudaMalloc(&sink, 4);
    cudaMemset(fp, 0x5A, (size_t)rows * 128);
    for (int warps : {16, 28, 32, 48, 64}) {
        run<1>(fp, rows, warps, prop.multiProcessorCount, sink);
        run<2>(fp, rows, warps, prop.multiProcessorCount, sink);
        run<4>(fp, rows, warps, prop.multiProcessorCount, sink);
        run<8>(fp, rows, warps, prop.multiProcessorCount, sink);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
