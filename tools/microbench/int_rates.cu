// Measures per-SM issue rates of the integer dot-product candidates on sm_100a:
// IDP.4A (dp4a), LOP3, legacy IMMA (mma.sync m16n8k32 u8.s8), and a LOP3+IDP mix.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o int_rates int_rates.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096

__global__ void k_dp4a(int* out, uint32_t a, uint32_t b, long long* cyc) {
    int acc[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    uint32_t x = a + threadIdx.x, y = b;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(acc[j]) : "r"(x), "r"(y));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < 8; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_lop(int* out, uint32_t a, uint32_t b, long long* cyc) {
    uint32_t acc[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    uint32_t x = a + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(acc[j]) : "r"(x), "r"(b));
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int j = 0; j < 8; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_mix(int* out, uint32_t a, uint32_t b, long long* cyc) {
    int acc[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    uint32_t m[8];
    uint32_t x = a + threadIdx.x, y = b;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(m[j]) : "r"(x), "r"(0x03030303u << (2 * (j & 3))), "r"(acc[j]));
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(acc[j]) : "r"(m[j]), "r"(y));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < 8; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_imma(int* out, uint32_t a, uint32_t b, long long* cyc) {
    int c[4][4] = {};
    uint32_t a0 = a + threadIdx.x, a1 = a ^ 0x01010101u, a2 = a + 7, a3 = a + 9, b0 = b, b1 = b + 3;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < 4; ++j) for (int q = 0; q < 4; ++q) s += c[j][q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_hmma(int* out, uint32_t a, uint32_t b, long long* cyc) {
    float c[4][4] = {};
    uint32_t a0 = a + threadIdx.x, a1 = a ^ 0x01010101u, a2 = a + 7, a3 = a + 9, b0 = b, b1 = b + 3;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    long long t1 = clock64();
    float s = 0;
    for (int j = 0; j < 4; ++j) for (int q = 0; q < 4; ++q) s += c[j][q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (int)s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_imad(int* out, uint32_t a, uint32_t b, long long* cyc) {
    int acc[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    int x = a + threadIdx.x, y = b;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("mad.lo.s32 %0, %1, %2, %0;" : "+r"(acc[j]) : "r"(x), "r"(y));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < 8; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <typename F>
void run(const char* name, F kern, int threads, int per_iter, double work_per_instr) {
    int* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    kern<<<148, threads>>>(out, 0x01020102u, 0x7f01ff02u, cyc);
    kern<<<148, threads>>>(out, 0x01020102u, 0x7f01ff02u, cyc);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    double warp_instrs = (double)ITERS * per_iter * (threads / 32);
    printf("%-8s threads/SM %4d: %8.0f clk, %.3f warp-instr/clk/SM (%.1f lanes/clk), %.0f work/clk/SM\n", name, threads, avg,
           warp_instrs / avg, warp_instrs * 32 / avg, warp_instrs * work_per_instr / avg);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int th : {128, 256, 512, 1024}) {
        run("dp4a", k_dp4a, th, 8, 32 * 4);
        run("lop3", k_lop, th, 8, 32);
        run("imad", k_imad, th, 8, 32);
        run("lop+dp4a", k_mix, th, 16, 32 * 2);
        run("imma8", k_imma, th, 4, 16 * 8 * 32);
        run("hmma16", k_hmma, th, 4, 16 * 8 * 16);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
