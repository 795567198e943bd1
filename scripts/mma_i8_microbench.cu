// mma_i8_microbench.cu -- issue rate of tcgen05.mma kind::i8 on sm_100a with the operand placements the exact-search
// kernel could use: A from shared memory (SS) or from tensor memory (TS), N = 256 or 128, K = 32 per instruction.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ab/mma_bench.bin scripts/mma_i8_microbench.cu
// Run:   ab/mma_bench.bin            (one CTA per SM, 4096 back-to-back MMAs per measurement, cycles per MMA)
#include <cstdint>
#include <cstdio>
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

// mode: 0 = SS N256, 1 = SS N128, 2 = TS N128, 3 = TS N256;  stride = K pieces advanced per MMA (0: same operands)
__global__ void __launch_bounds__(128, 1) bench(int mode, int iters, int per_commit, unsigned long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tmem_base;
    __shared__ unsigned long long bar, scratch_bar;
    unsigned char* a = smem;                 // 128 rows x 64 K pieces (128 KB) like the kernel's query operand
    unsigned char* b = smem + 131072;        // 256 rows x 8 K pieces (32 KB) x 2 stages
    for (int i = threadIdx.x; i < (131072 + 65536) / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * ((i * 2654435761u) >> 31);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&scratch_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    const int n = (mode == 0 || mode == 3) ? 256 : 128;
    const uint32_t idesc = (2u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t ad0 = make_desc(smem_u32(a), 128, 64 * 128);
    const uint64_t bd0 = make_desc(smem_u32(b), 128, 8 * 128);
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        t0 = clock64();
        for (int i = 0; i < iters; i += 8) {
            const int chunk = (i >> 3) & 7, st = (i >> 3) & 1;  // like the kernel: 8 pieces per stage, 2 stages
#pragma unroll
            for (int j = 0; j < 8; ++j) {                       // 8 MMAs with compile-time offsets
                const int piece = (j & 3) * 2;
                const uint64_t ad = ad0 + (uint64_t)(((chunk * 8 + piece) * 128) >> 4);
                const uint64_t bd = bd0 + (uint64_t)((st * 32768 + piece * 128) >> 4);
                const uint32_t acc = (i | j) != 0;
                if (mode < 2) {
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(tm), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
                } else {
                    const uint32_t at = tm + 256 + (uint32_t)((chunk * 8 + piece) * 4);  // A in TMEM: 4 columns per piece
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                                 ::"r"(tm), "r"(at), "l"(bd), "r"(idesc), "r"(acc) : "memory");
                }
                if (per_commit == 4 && (j & 3) == 3)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                 ::"r"(smem_u32(&scratch_bar)) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                     ::"r"(smem_u32(&bar)) : "memory");
        mbar_wait(&bar, 0);
        t1 = clock64();
        if (blockIdx.x == 0)
            out[0] = (unsigned long long)(t1 - t0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

int main() {
    unsigned long long* d;
    cudaMalloc(&d, 8);
    const size_t smem = 131072 + 65536 + 1024;
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const char* names[4] = {"SS A smem, N=256", "SS A smem, N=128", "TS A tmem, N=128", "TS A tmem, N=256"};
    const int iters = 4096;
    for (int per_commit : {4096, 4}) {
        for (int mode = 0; mode < 4; ++mode) {
            unsigned long long h = 0;
            for (int rep = 0; rep < 2; ++rep) {
                bench<<<148, 128, smem>>>(mode, iters, per_commit, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) {
                    printf("%s: %s\n", names[mode], cudaGetErrorString(e));
                    return 1;
                }
            }
            cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
            printf("%-18s commit every %4d MMAs, one wait at the end: %7.1f cycles per MMA (M128 K32)\n", names[mode], per_commit,
                   (double)h / iters);
        }
    }
    return 0;
}
