// Read-stream ceiling of one persistent CTA per SM fetching 16 KB tiles into a shared-memory ring (no compute).
// Variants: 0 = 4-D SWIZZLE_128B box (32 words, 16 rows, 8 segments) = the decode GEMV's tile (rows 1 KB apart);
//           1 = 1-D cp.async.bulk of the same contiguous 16 KB; 2 = 2-D box (128 B x 128 lines, contiguous);
//           3 = plain ld.global.nc.v4 grid-stride loop (no shared memory).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_stream.bin tma_stream.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t ph) {
    uint32_t d = 0;
    while (!d) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(d) : "r"(b), "r"(ph) : "memory");
}

template <int V, int TILE>
__global__ void __launch_bounds__(64, 1) stream_k(const __grid_constant__ CUtensorMap map, const uint8_t* src, long long ntiles, int stages,
                                                  int rows_per_layer, unsigned* sink) {
    extern __shared__ uint8_t sm_raw[];
    const uint32_t base = (smem_u32(sm_raw) + 1023u) & ~1023u;
    uint8_t* gb = sm_raw + (base - smem_u32(sm_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(gb + (size_t)stages * TILE);
    const uint32_t sb = smem_u32(bars);
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(sb + 8 * s, 1); mbar_init(sb + 8 * (32 + s), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
                mbar_wait(sb + 8 * (32 + s), ph ^ 1u);
                mbar_expect(sb + 8 * s, TILE);
                const uint32_t dst = base + s * TILE;
                if (V == 0) {
                    const int rb_per_layer = rows_per_layer / 16;
                    const int l = (int)(t / rb_per_layer), rb = (int)(t - (long long)l * rb_per_layer);
                    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                                 ::"r"(dst), "l"(&map), "r"(sb + 8 * s), "r"(0), "r"(rb * 16), "r"(0), "r"(l) : "memory");
                } else if (V == 1) {
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(dst), "l"(src + t * TILE), "r"(TILE), "r"(sb + 8 * s) : "memory");
                } else if (V == 2) {
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                                 ::"r"(dst), "l"(&map), "r"(sb + 8 * s), "r"(0), "r"((int)(t * 128)) : "memory");
                } else {
                    // V == 4 / 5: a TILE is (bw bytes x 128 rows) of a [rows][1024 B] matrix; consecutive tiles of a CTA walk
                    // the 1024/bw column blocks of one 128-row block (the tcgen05 GEMM's weight access pattern)
                    const int bw = TILE / 128, kbs = 1024 / bw;
                    const long long blk = t / kbs; const int kb = (int)(t - blk * kbs);
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                                 ::"r"(dst), "l"(&map), "r"(sb + 8 * s), "r"(kb * bw), "r"((int)(blk * 128)) : "memory");
                }
                if (++s == stages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        int s = 0; uint32_t ph = 0; unsigned acc = 0;
        for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
            mbar_wait(sb + 8 * s, ph);
            acc += reinterpret_cast<const unsigned*>(gb + s * TILE)[lane * (TILE / 128)];
            __syncwarp();
            if (lane == 0) mbar_arrive(sb + 8 * (32 + s));
            if (++s == stages) { s = 0; ph ^= 1u; }
        }
        if (acc == 0x12345678u) sink[0] = acc;
    }
}

__global__ void __launch_bounds__(512) ldg_k(const uint4* __restrict__ src, long long n16, unsigned* sink) {
    unsigned acc = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        uint4 a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
        acc += a.x ^ b.y ^ c.z ^ d.w;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                          const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int N = 11008, LD = 1024;
    const int L = 90;                                    // 90 x 11.27 MB = 1.01 GB
    const size_t bytes = (size_t)L * N * LD;
    uint8_t* buf; unsigned* sink;
    cudaMalloc(&buf, bytes); cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, bytes);
    void* fp; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    EncFn enc = (EncFn)fp;
    constexpr int TILE = 16384;
    const long long ntiles = (long long)bytes / TILE;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int promo = 0; promo < 2; ++promo) {
        CUtensorMapL2promotion pr = promo ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
        CUtensorMap m4, m2;
        { cuuint64_t d[4] = {32, (cuuint64_t)N, 8, (cuuint64_t)L}, st[3] = {(cuuint64_t)LD, 128, (cuuint64_t)N * LD};
          cuuint32_t bx[4] = {32, 16, 8, 1}, es[4] = {1, 1, 1, 1};
          CUresult r = enc(&m4, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, buf, d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          if (r) printf("enc4 %d\n", (int)r); }
        { cuuint64_t d[2] = {32, (cuuint64_t)(bytes / 128)}, st[1] = {128};
          cuuint32_t bx[2] = {32, 128}, es[2] = {1, 1};
          CUresult r = enc(&m2, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, buf, d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          if (r) printf("enc2 %d\n", (int)r); }
        for (int v = 0; v < 3; ++v) {
            if (v == 1 && promo) continue;
            for (int stages : {4, 8, 12}) {
                for (int grid : {148, 296}) {
                    const size_t smem = (size_t)stages * TILE + 1024 + 64 * 8 + 64;
                    if (grid == 296 && smem > 110 * 1024) continue;
                    auto launch = [&]() {
                        if (v == 0) { cudaFuncSetAttribute(stream_k<0, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); stream_k<0, 16384><<<grid, 64, smem>>>(m4, buf, ntiles, stages, N, sink); }
                        if (v == 1) { cudaFuncSetAttribute(stream_k<1, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); stream_k<1, 16384><<<grid, 64, smem>>>(m4, buf, ntiles, stages, N, sink); }
                        if (v == 2) { cudaFuncSetAttribute(stream_k<2, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); stream_k<2, 16384><<<grid, 64, smem>>>(m2, buf, ntiles, stages, N, sink); }
                    };
                    launch(); cudaDeviceSynchronize();
                    cudaEventRecord(e0); launch(); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1);
                    printf("variant %d promo %s stages %2d grid %3d: %7.1f GB/s  (%s)\n", v, promo ? "256B" : "128B", stages, grid,
                           2.0 * bytes / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
                }
            }
        }
    }
    {   // strided column-block boxes: 64 B x 128 rows (8 KB, SWIZZLE_64B) and 128 B x 128 rows (16 KB, SWIZZLE_128B)
        const long long rows = (long long)(bytes / 1024);
        CUtensorMap m64, m128;
        { cuuint64_t d[2] = {1024, (cuuint64_t)rows}, st[1] = {1024}; cuuint32_t bx[2] = {64, 128}, es[2] = {1, 1};
          enc(&m64, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        { cuuint64_t d[2] = {1024, (cuuint64_t)rows}, st[1] = {1024}; cuuint32_t bx[2] = {128, 128}, es[2] = {1, 1};
          enc(&m128, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        for (int v : {4, 5}) for (int stages : {8, 12, 20}) {
            const int tile = v == 4 ? 8192 : 16384;
            if (v == 5 && stages > 12) continue;
            const size_t smem = (size_t)stages * tile + 1024 + 64 * 8 + 64;
            const long long nt = (long long)bytes / tile;
            auto launch = [&]() {
                if (v == 4) { cudaFuncSetAttribute(stream_k<4, 8192>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); stream_k<4, 8192><<<148, 64, smem>>>(m64, buf, nt, stages, N, sink); }
                else { cudaFuncSetAttribute(stream_k<5, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); stream_k<5, 16384><<<148, 64, smem>>>(m128, buf, nt, stages, N, sink); }
            };
            launch(); cudaDeviceSynchronize();
            cudaEventRecord(e0); launch(); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            printf("variant %d (%3d B x 128 rows) stages %2d: %7.1f GB/s (%s)\n", v, tile / 128, stages, 2.0 * bytes / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
        }
    }
    for (int grid : {148 * 2, 148 * 4, 148 * 8}) {
        ldg_k<<<grid, 512>>>((const uint4*)buf, (long long)(bytes / 16), sink); cudaDeviceSynchronize();
        cudaEventRecord(e0); ldg_k<<<grid, 512>>>((const uint4*)buf, (long long)(bytes / 16), sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("variant 3 (ldg.v4) grid %4d: %7.1f GB/s (%s)\n", grid, bytes / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
