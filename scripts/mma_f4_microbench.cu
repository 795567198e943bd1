// mma_f4_microbench.cu -- tcgen05.mma kind::mxf4 (block-scaled e2m1, K = 64 per instruction) on sm_100a as a BIT COUNTER:
// operands are 0 / 0.5 / 1.0 / 2.0 in e2m1, every scale factor is 1.0 (ue8m0 127), accumulators are f32 -- sums of
// up to 1024 products of {0, 0.25, 0.5, 1, 2, 4} are exact.  Part 1 checks the numbers against the CPU for the operand
// layout the exact-search kernel would use (K-major, SWIZZLE_NONE, 16-byte K pieces = 32 elements); part 2 measures
// the issue rate at N = 256 / 240 / 224 / 128 beside kind::i8.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/build/mma_f4_microbench scripts/mma_f4_microbench.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void mbar_wait(void* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
// instruction descriptor, block-scaled kinds (cute/arch/mma_sm100_desc.hpp: InstrDescriptorBlockScaled): a/b format
// E2M1 = 1 at [7,10) / [10,13), K-major, N >> 3 at [17,23), scale format UE8M0 = 1 at bit 23, M >> 4 at [24,29), K64
__host__ __device__ constexpr uint32_t idesc_f4(int n) {
    return (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (1u << 23) | ((uint32_t)(128 >> 4) << 24);
}
__host__ __device__ constexpr uint32_t idesc_i8(int n) { return (2u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24); }

__device__ __forceinline__ void mma_f4(uint32_t d, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t sfa, uint32_t sfb,
                                       uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc), "r"(sfa), "r"(sfb) : "memory");
}
__device__ __forceinline__ void mma_i8(uint32_t d, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
}

constexpr int kSfCol = 480;  // scale factors: TMEM columns [480, 512) of every lane hold 0x7F7F7F7F

// smem: A = 128 rows, B = 256 rows, `pieces` 16-byte K pieces per row (32 e2m1 elements each); layout
// address(row, piece) = (row / 8) * pieces * 128 + piece * 128 + (row % 8) * 16
// mode 0: correctness (results -> out_f[128 x n]); mode 1: rate f4; mode 2: rate i8
__global__ void __launch_bounds__(128, 1) bench(int mode, int n, int pieces, int iters, const uint8_t* a_in,
                                                const uint8_t* b_in, float* out_f, unsigned long long* out_cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tmem_base;
    __shared__ unsigned long long bar;
    unsigned char* a = smem;
    unsigned char* b = smem + 128 * pieces * 16;
    const int tid = threadIdx.x;
    for (int i = tid; i < 128 * pieces; i += blockDim.x) {
        int row = i / pieces, pc = i % pieces;
        uint4 v = a_in ? reinterpret_cast<const uint4*>(a_in)[i] : make_uint4(0x21212121u, 0x12121212u, 0x21212121u, 0x12121212u);
        *reinterpret_cast<uint4*>(a + (row >> 3) * pieces * 128 + pc * 128 + (row & 7) * 16) = v;
    }
    for (int i = tid; i < 256 * pieces; i += blockDim.x) {
        int row = i / pieces, pc = i % pieces;
        uint4 v = b_in ? reinterpret_cast<const uint4*>(b_in)[i] : make_uint4(0x12121212u, 0x21212121u, 0x12121212u, 0x21212121u);
        *reinterpret_cast<uint4*>(b + (row >> 3) * pieces * 128 + pc * 128 + (row & 7) * 16) = v;
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    {   // every scale factor = 1.0: thread = TMEM lane, 32 columns of 0x7F7F7F7F
        const uint32_t taddr = tm + ((uint32_t)((tid >> 5) * 32) << 16) + (uint32_t)kSfCol;
        const uint32_t one = 0x7F7F7F7Fu;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                     "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
                     "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(one) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint64_t ad0 = make_desc(smem_u32(a), 128, pieces * 128);
    const uint64_t bd0 = make_desc(smem_u32(b), 128, pieces * 128);
    const uint32_t sfa = tm + kSfCol, sfb = tm + kSfCol + 16;
    if (tid == 0) {
        long long t0 = clock64();
        if (mode == 0) {
            for (int k2 = 0; k2 < pieces / 2; ++k2)  // K = 64 elements = two 16-byte pieces per instruction
                mma_f4(tm, ad0 + (uint64_t)((k2 * 2 * 128) >> 4), bd0 + (uint64_t)((k2 * 2 * 128) >> 4), idesc_f4(n), sfa, sfb,
                       (uint32_t)(k2 != 0));
        } else {
            // the exact-search kernel's layout: A = 32 pieces per row (SBO 4096), B = stages of 8 pieces (SBO 1024, 32 KB)
            const uint64_t bs0 = make_desc(smem_u32(b), 128, 8 * 128);
            for (int i = 0; i < iters; i += 8) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int g = (i + j) >> 2, st = g & 3, c = g & 3, jj = j & 3;
                    const uint64_t ad = ad0 + (uint64_t)((((c * 8 + jj * 2) & 31) * 128) >> 4);  // (pieces == 32)
                    const uint64_t bd = bs0 + (uint64_t)((st * 32768 + jj * 2 * 128) >> 4);
                    if (mode == 1)
                        mma_f4(tm, ad, bd, idesc_f4(n), sfa, sfb, (uint32_t)((i | j) != 0));
                    else
                        mma_i8(tm, ad, bd, idesc_i8(n), (uint32_t)((i | j) != 0));
                }
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && out_cycles)
            out_cycles[0] = (unsigned long long)(t1 - t0);
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (mode == 0 && blockIdx.x == 0) {
        for (int c0 = 0; c0 < n; c0 += 8) {
            uint32_t r[8];
            const uint32_t taddr = tm + ((uint32_t)((tid >> 5) * 32) << 16) + (uint32_t)c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int j = 0; j < 8; ++j)
                out_f[(size_t)tid * n + c0 + j] = __uint_as_float(r[j]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

static float e2m1(uint8_t nib) {
    static const float mag[8] = {0.f, 0.5f, 1.f, 1.5f, 2.f, 3.f, 4.f, 6.f};
    float v = mag[nib & 7];
    return (nib & 8) ? -v : v;
}

int main() {
    const int pieces = 32;  // K = 1024 elements per row = one fingerprint
    const size_t a_bytes = 128 * pieces * 16, b_bytes = 256 * pieces * 16;
    const size_t smem = a_bytes + b_bytes + 1024;
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // ---- part 1: numbers.  Row-major host images [row][piece][16 bytes]; nibbles drawn from {0, 0.5, 1, 2}
    uint8_t *ha = (uint8_t*)malloc(a_bytes), *hb = (uint8_t*)malloc(b_bytes);
    const uint8_t pal[4] = {0x0, 0x1, 0x2, 0x4};
    uint32_t s = 12345;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return s >> 24; };
    for (size_t i = 0; i < a_bytes; ++i) ha[i] = (uint8_t)(pal[rnd() & 3] | (pal[rnd() & 3] << 4));
    for (size_t i = 0; i < b_bytes; ++i) hb[i] = (uint8_t)(pal[rnd() & 3] | (pal[rnd() & 3] << 4));
    uint8_t *da, *db;
    float* dout;
    unsigned long long* dcyc;
    cudaMalloc(&da, a_bytes), cudaMalloc(&db, b_bytes), cudaMalloc(&dout, 128 * 256 * 4), cudaMalloc(&dcyc, 8);
    cudaMemcpy(da, ha, a_bytes, cudaMemcpyHostToDevice);
    cudaMemcpy(db, hb, b_bytes, cudaMemcpyHostToDevice);
    for (int n : {256, 224}) {
        cudaMemset(dout, 0, 128 * 256 * 4);
        bench<<<1, 128, smem>>>(0, n, pieces, 0, da, db, dout, dcyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("correctness run N=%d: %s\n", n, cudaGetErrorString(e));
            return 1;
        }
        float* ho = (float*)malloc(128 * 256 * 4);
        cudaMemcpy(ho, dout, 128 * (size_t)n * 4, cudaMemcpyDeviceToHost);
        size_t bad = 0;
        double worst = 0;
        for (int r = 0; r < 128; ++r)
            for (int c = 0; c < n; ++c) {
                double want = 0;
                for (int k = 0; k < pieces * 16; ++k) {
                    uint8_t x = ha[(size_t)r * pieces * 16 + k], y = hb[(size_t)c * pieces * 16 + k];
                    want += (double)e2m1(x & 15) * e2m1(y & 15) + (double)e2m1(x >> 4) * e2m1(y >> 4);
                }
                double got = ho[(size_t)r * n + c];
                if (got != want) {
                    if (bad < 5)
                        printf("  mismatch N=%d (%d, %d): got %.3f want %.3f\n", n, r, c, got, want);
                    ++bad;
                }
                if (want > worst) worst = want;
            }
        printf("kind::mxf4 N=%d K=1024: %zu of %d sums differ from the CPU (largest sum %.2f)\n", n, bad, 128 * n, worst);
        free(ho);
    }
    // ---- part 2: issue rate, one CTA per SM, 4096 back-to-back instructions
    const int iters = 4096;
    for (int mode : {1, 2})
        for (int n : {256, 240, 224, 128}) {
            unsigned long long h = 0;
            for (int rep = 0; rep < 2; ++rep) {
                bench<<<148, 128, smem>>>(mode, n, pieces, iters, nullptr, nullptr, dout, dcyc);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) {
                    printf("rate run: %s\n", cudaGetErrorString(e));
                    return 1;
                }
            }
            cudaMemcpy(&h, dcyc, 8, cudaMemcpyDeviceToHost);
            const double cyc = (double)h / iters;
            const int k = mode == 1 ? 64 : 32;
            printf("%s M=128 N=%3d K=%d: %6.1f cycles per instruction = %6.1f T(multiply-add pairs)/s per 148 SMs at 1.965 GHz\n",
                   mode == 1 ? "kind::mxf4" : "kind::i8  ", n, k, cyc, 128.0 * n * k * 148 * 1.965e9 / cyc / 1e12);
        }
    return 0;
}
