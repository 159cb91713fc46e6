// Isolates the consumer inner loop of the decode GEMV: per tile and warp 4 LDS.128 + 64 LOP3 + 16 IMMA (m16n8k32 u8.s8).
// Variants: 0 full, 1 no LOP3 (raw operands), 2 no IMMA (LOP3 + XOR), 3 full + mbarrier try_wait per tile,
//           4 full with 8 accumulator sets (chain distance 8)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o imma_loop imma_loop.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#ifndef TILES
#define TILES 64
#endif
__device__ __forceinline__ void imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <int V>
__global__ void __launch_bounds__(512) k(int* out, long long* cyc, uint32_t seed) {
    extern __shared__ uint4 sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    for (int i = threadIdx.x; i < 5 * 1024; i += blockDim.x) sm[i] = make_uint4(seed + i, seed * 3 + i, seed ^ i, i);
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)));
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
    __syncthreads();
    uint4 bq[2][4];
    for (int pp = 0; pp < 2; ++pp) for (int i = 0; i < 4; ++i) bq[pp][i] = sm[(pp * 4 + i) * 32 + lane];
    int acc[8][4] = {};
    const uint8_t* base = reinterpret_cast<const uint8_t*>(sm);
    long long t0 = clock64();
#pragma unroll 1
    for (int tile = 0; tile < TILES; ++tile) {
        if (V == 3) {
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(0u) : "memory");
        }
        const uint8_t* seg = base + (tile % 5) * 16384 + (warp & 7) * 2048 + g * 128;
        uint4 wlo[2], whi[2];
#pragma unroll
        for (int pp = 0; pp < 2; ++pp) {
            const int phys = ((2 * t + pp) ^ g) << 4;
            wlo[pp] = *reinterpret_cast<const uint4*>(seg + phys);
            whi[pp] = *reinterpret_cast<const uint4*>(seg + 1024 + phys);
        }
        if (V == 5 || V == 6) {
            // pinned schedule: all masks of a batch first (distinct registers), then the batch's IMMAs back to back
            auto vand = [](uint32_t a, uint32_t m) { uint32_t r; asm volatile("and.b32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(m)); return r; };
#pragma unroll
            for (int pp = 0; pp < 2; ++pp) {
                if (V == 5) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t m[4][4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint32_t mk = 0x03030303u << (2 * i);
                            m[i][0] = vand(h ? wlo[pp].z : wlo[pp].x, mk); m[i][1] = vand(h ? whi[pp].z : whi[pp].x, mk);
                            m[i][2] = vand(h ? wlo[pp].w : wlo[pp].y, mk); m[i][3] = vand(h ? whi[pp].w : whi[pp].y, mk);
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            imma(acc[i], m[i][0], m[i][1], m[i][2], m[i][3], h ? bq[pp][i].z : bq[pp][i].x, h ? bq[pp][i].w : bq[pp][i].y);
                    }
                } else {
                    uint32_t m[2][4][4];
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint32_t mk = 0x03030303u << (2 * i);
                            m[h][i][0] = vand(h ? wlo[pp].z : wlo[pp].x, mk); m[h][i][1] = vand(h ? whi[pp].z : whi[pp].x, mk);
                            m[h][i][2] = vand(h ? wlo[pp].w : wlo[pp].y, mk); m[h][i][3] = vand(h ? whi[pp].w : whi[pp].y, mk);
                        }
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            imma(acc[i + 4 * h], m[h][i][0], m[h][i][1], m[h][i][2], m[h][i][3], h ? bq[pp][i].z : bq[pp][i].x, h ? bq[pp][i].w : bq[pp][i].y);
                }
            }
            continue;
        }
#pragma unroll
        for (int pp = 0; pp < 2; ++pp) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    const uint32_t l0 = h ? wlo[pp].z : wlo[pp].x, l1 = h ? wlo[pp].w : wlo[pp].y;
                    const uint32_t h0 = h ? whi[pp].z : whi[pp].x, h1 = h ? whi[pp].w : whi[pp].y;
                    const uint32_t b0 = h ? bq[pp][i].z : bq[pp][i].x, b1 = h ? bq[pp][i].w : bq[pp][i].y;
                    const int a = (V == 4) ? i + 4 * h : i;
                    if (V == 1) imma(acc[a], l0, h0, l1, h1, b0, b1);
                    else if (V == 2) { acc[a][0] ^= l0 & mk; acc[a][1] ^= h0 & mk; acc[a][2] ^= l1 & mk; acc[a][3] ^= h1 & mk; }
                    else imma(acc[a], l0 & mk, h0 & mk, l1 & mk, h1 & mk, b0, b1);
                }
            }
        }
    }
    long long t1 = clock64();
    int s = 0;
    for (int a = 0; a < 8; ++a) for (int q = 0; q < 4; ++q) s += acc[a][q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int V> void run(const char* name, int threads) {
    int* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * 16384);
    for (int r = 0; r < 2; ++r) k<V><<<148, threads, 5 * 16384>>>(out, cyc, 12345u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const int warps = threads / 32;
    // a "tile" of the real kernel = 8 warps x 16 IMMA; here every warp does 16 IMMA per iteration
    printf("%-28s warps %2d: %7.1f clk per iteration (16 IMMA/warp) -> %6.1f clk per 128 IMMA, %.3f IMMA/clk/SM\n", name, warps,
           avg / TILES, avg / TILES * 8 / warps, 16.0 * warps * TILES / avg);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int th : {256, 384, 512}) {
        run<0>("full (LDS+LOP+IMMA)", th);
        run<1>("no LOP3", th);
        run<2>("no IMMA (LOP+XOR)", th);
        run<3>("full + try_wait", th);
        run<4>("full, 8 acc sets", th);
        run<5>("pinned batches of 4", th);
        run<6>("pinned batches of 8, 8 acc", th);
    }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
