// Issue-rate ceiling of tcgen05.mma.cta_group::1.kind::i8 on this GPU: one CTA per SM, no loads in the loop -- operands sit
// in shared memory (SS form) or tensor memory (TS form, what gemm_ts.cu uses for the expanded weights) and one elected
// thread issues back-to-back MMAs of shape 128 x N x 32 into two alternating TMEM accumulators, a tcgen05.commit every
// 64 groups of 8 MMAs.  Prints one JSON line: TOP/s per variant (burst: a few ms of MMAs; sustained: --seconds S of
// back-to-back launches).  This is the MEASURED int8 tensor ceiling the prefill GEMM's roofline fraction is quoted
// against (bench.py `gemm.int8_peak`).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o umma_i8_rate.bin umma_i8_rate.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t ph) {
    uint32_t d = 0;
    while (!d) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(d) : "r"(b), "r"(ph) : "memory");
}
// K-major SWIZZLE_128B descriptor: rows of 128 bytes, 8-row atoms 1024 bytes apart
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// kind::i8 instruction descriptor: D = s32, A = u8, B = s8, K-major both
__host__ __device__ constexpr uint32_t idesc_i8(int m, int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// TS: A operand from tensor memory (columns 2N.. hold 128 lanes x 32 B per K=32 slice)
template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) rate_kernel(int groups, unsigned long long* clocks) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    uint8_t* gb = raw + (base - smem_u32(raw));
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    // A: 128 rows x 128 B of codes 0..2, B: N rows x 128 B of arbitrary int8 (data toggling as in a real GEMM)
    uint32_t* w = reinterpret_cast<uint32_t*>(gb);
    for (int i = threadIdx.x; i < (128 + N) * 32; i += 128) {
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
        h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
        w[i] = (i < 128 * 32) ? ((h & 0x01010101u) + ((h >> 1) & 0x01010101u)) : h;
    }
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    if (TS) {
        uint32_t r[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = ((threadIdx.x * 32 + i) * 2246822519u >> 7) & 0x01010101u;
        const uint32_t ta = tm + ((uint32_t)(warp * 32) << 16) + 2 * N;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                     ::"r"(ta), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
                       "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
                       "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        const uint32_t sA = base, sB = base + 128 * 128;
        constexpr uint32_t idesc = idesc_i8(128, N);
        const long long t0 = clock64();
        uint32_t ph = 0;
        for (int g = 0; g < groups; ++g) {
            const uint32_t d = tm + (uint32_t)((g & 1) * N);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t ko = (uint32_t)(j & 3) * 32;                 // K slice inside the 128-byte swizzled row
                const uint64_t bd = desc_sw128(sB + ko);
                const uint32_t acc = (g >= 2 || j > 0) ? 1u : 0u;
                if (TS) {
                    const uint32_t ta = tm + 2 * N + (uint32_t)(j & 3) * 8;
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                                 ::"r"(d), "r"(ta), "l"(bd), "r"(idesc), "r"(acc) : "memory");
                } else {
                    const uint64_t ad = desc_sw128(sA + ko);
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
                }
            }
            if ((g & 63) == 63 || g == groups - 1) {                        // bound the number of MMAs in flight
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
                mbar_wait(smem_u32(&bar), ph); ph ^= 1u;
            }
        }
        const long long t1 = clock64();
        if (clocks) clocks[blockIdx.x] = (unsigned long long)(t1 - t0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

template <int N, bool TS>
static void run(const char* name, int sms, int groups, double seconds, bool first) {
    auto k = rate_kernel<N, TS>;
    const size_t smem = 1024 + (128 + N) * 128;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned long long* clk;
    cudaMalloc(&clk, sms * sizeof(unsigned long long));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) k<<<sms, 128, smem>>>(groups, clk);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < 5; ++i) {
        cudaEventRecord(e0);
        k<<<sms, 128, smem>>>(groups, clk);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    unsigned long long h[256];
    cudaMemcpy(h, clk, sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    double cmax = 0;
    for (int i = 0; i < sms && i < 256; ++i) if ((double)h[i] > cmax) cmax = (double)h[i];
    const double ops = 2.0 * 128 * N * 32 * 8.0 * groups * sms;
    double sustained = 0.0, sus_s = 0.0;
    if (seconds > 0) {
        const int launches = (int)(seconds * 1e3 / best) + 1;
        cudaEventRecord(e0);
        for (int i = 0; i < launches; ++i) k<<<sms, 128, smem>>>(groups, nullptr);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        sustained = ops * launches / (ms * 1e-3) / 1e12; sus_s = ms * 1e-3;
    }
    cudaError_t err = cudaGetLastError();
    printf("%s\"%s\": {\"tops_burst\": %.1f, \"ms\": %.4f, \"clk_per_mma\": %.2f, \"tops_sustained\": %.1f, \"sustained_s\": %.2f, \"err\": \"%s\"}",
           first ? "" : ", ", name, ops / (best * 1e-3) / 1e12, best, cmax / (8.0 * groups), sustained, sus_s, cudaGetErrorString(err));
    cudaFree(clk);
}

int main(int argc, char** argv) {
    double seconds = 0.0;
    int groups = 4096;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--seconds") && i + 1 < argc) seconds = atof(argv[++i]);
        if (!strcmp(argv[i], "--groups") && i + 1 < argc) groups = atoi(argv[++i]);
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { printf("{\"error\": \"no device\"}\n"); return 1; }
    const int sms = prop.multiProcessorCount;
    printf("{\"sms\": %d, \"mma\": \"tcgen05.mma.cta_group::1.kind::i8 128xNx32, u8 x s8 -> s32\", ", sms);
    run<256, false>("ss_n256", sms, groups, seconds, true);
    run<192, false>("ss_n192", sms, groups, 0, false);
    run<128, false>("ss_n128", sms, groups, 0, false);
    run<192, true>("ts_n192", sms, groups, seconds, false);
    run<128, true>("ts_n128", sms, groups, 0, false);
    printf("}\n");
    return 0;
}
