// quant.cu — per-token symmetric int8 activation quantisation (stand-alone form).
// The decode GEMV fuses the same arithmetic into its prologue (gemv.cu); the tcgen05 GEMM
// consumes the output of this kernel with kperm=1.
//
//   s_x[m]  = max_k |x[m,k]| / 127          (1.0 when the row is all zero)
//   inv[m]  = 127 / max_k |x[m,k]|          (1.0 when the row is all zero)
//   q[m,k]  = clamp(rint(x[m,k] * inv[m]), -127, 127)     fp32 multiply, rint = round-half-even
//   sum[m]  = sum_k q[m,k]
// Output layouts: 0 natural K order; 1 "kperm" (tcgen05 GEMM operand order: within every group of
// 16, position 4i+b holds element 4b+i); 2 the GEMV's shared-memory image (kperm words, 16-byte
// chunks arranged [slab of 2048 k][word j of the packed chunk][lane]).
#include <type_traits>
#include "common.cuh"

namespace tb200 {

// rint + float->int via the 1.5*2^23 trick (no XU-pipe conversions); same bits as rintf + cast
__device__ __forceinline__ int quant1(float v, float inv) {
    float r = __fmul_rn(v, inv);
    r = fminf(fmaxf(r, -127.0f), 127.0f);
    return __float_as_int(__fadd_rn(r, 12582912.0f)) - 0x4B400000;
}

// 16 consecutive elements as fp32 (zero beyond K); 128-bit loads when the row is 16-byte aligned
template <typename T>
__device__ __forceinline__ void load16q(const T* __restrict__ xr, int64_t g, int64_t K, bool vec, float v[16]) {
    if (vec && 16 * g + 16 <= K) {
        if (sizeof(T) == 4) {
            const float4* q = reinterpret_cast<const float4*>(xr + 16 * g);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 f = q[j];
                v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
            }
        } else {
            const uint4* q = reinterpret_cast<const uint4*>(xr + 16 * g);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint4 u = q[j];
                const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
                for (int t = 0; t < 8; ++t) v[8 * j + t] = to_f32(e[t]);
            }
        }
    } else {
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            int64_t k = 16 * g + t;
            v[t] = k < K ? to_f32(xr[k]) : 0.0f;
        }
    }
}

__device__ __forceinline__ void store16q(int8_t* qr, int64_t g, const float (&v)[16], float inv, int layout, int& lsum) {
    int qv[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) {
        qv[t] = quant1(v[t], inv);
        lsum += qv[t];
    }
    const int kperm = layout != 0;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        // natural: word i = elements 4i..4i+3 ; kperm: word i = elements i, 4+i, 8+i, 12+i
        int e0 = kperm ? i : 4 * i, st = kperm ? 4 : 1;
        w[i] = (uint32_t)(qv[e0] & 0xff) | ((uint32_t)(qv[e0 + st] & 0xff) << 8) |
               ((uint32_t)(qv[e0 + 2 * st] & 0xff) << 16) | ((uint32_t)(qv[e0 + 3 * st] & 0xff) << 24);
    }
    int64_t pos = g;
    if (layout == 3) {      // field-split order: even fields in the first 128 bytes of every 256-element block, odd fields in the second
        int8_t* blk = qr + 256 * (g >> 4) + 8 * (g & 15);
        *reinterpret_cast<uint2*>(blk) = make_uint2(w[0], w[2]);
        *reinterpret_cast<uint2*>(blk + 128) = make_uint2(w[1], w[3]);
        return;
    }
    if (layout == 2) {
        const int64_t c = g >> 2, j = g & 3;
        pos = ((c >> 5) * 4 + j) * 32 + (c & 31);
    }
    *reinterpret_cast<uint4*>(qr + 16 * pos) = make_uint4(w[0], w[1], w[2], w[3]);
}

// One WARP per token row (no shared memory, no block barrier): lane owns 16-element groups lane, lane+32, ...
// Rows of up to 32*GC groups (K <= 4096 for GC=8) are held in registers between the absmax pass and the
// quantisation pass, so the activations are read from HBM once; longer rows are re-read (L2).
constexpr int QW = 4;     // warps (rows) per CTA
constexpr int GC_MAX = 8; // cached groups per lane (template parameter GC: 2, 4 or 8 by row length)
// register cap: 16 (GC=8) .. 32 (GC=2) resident warps per SM keep enough loads in flight for one-pass rows
template <typename T, int GC>
__global__ void __launch_bounds__(QW * 32, GC == 8 ? 4 : (GC == 4 ? 6 : 8))
quant_act_kernel(const T* __restrict__ x, int64_t M, int64_t K, int64_t ldx,
                 int8_t* __restrict__ q, int64_t ldq, float* __restrict__ s_x,
                 int32_t* __restrict__ sum_qx, int layout, bool vec_ok) {
    // a dependent kernel launched with programmatic stream serialisation (the grouped tcgen05 GEMM) may start its
    // prologue and weight stream now; it waits (griddepcontrol.wait) before it reads q / s_x / sum_qx
    asm volatile("griddepcontrol.launch_dependents;");
    const int lane = threadIdx.x & 31;
    const int64_t m = (int64_t)blockIdx.x * QW + (threadIdx.x >> 5);
    if (m >= M) return;
    const T* xr = x + m * ldx;
    int8_t* qr = q + m * ldq;
    const int64_t groups = ldq >> 4;          // ldq is a multiple of 16; groups beyond K quantise to zero
    const bool cached = groups <= 32 * GC;

    // Fast path (aligned rows, K a multiple of 16, whole row cached): ALL the row's 128-bit loads are issued before
    // the first use and the raw words stay in registers for both passes.  (The generic path below loads group by
    // group behind a bounds branch, which exposes one DRAM latency per group: 11 us for 800 rows of 4096 in ncu.)
    if (cached && vec_ok && (K & 15) == 0) {
        constexpr int NV = sizeof(T) == 4 ? 4 : 2;           // uint4 per 16 elements
        const int64_t kgroups = K >> 4;
        uint4 raw[GC][NV];
#pragma unroll
        for (int u = 0; u < GC; ++u) {
            const int64_t g = lane + 32 * u;
            const int64_t gg = g < kgroups ? g : kgroups - 1;    // clamped: always a valid address, never a branch
            const uint4* src = reinterpret_cast<const uint4*>(xr + 16 * gg);
#pragma unroll
            for (int j = 0; j < NV; ++j) raw[u][j] = __ldg(src + j);
        }
        auto unpack = [&](int u, float (&v)[16]) {
            const bool valid = lane + 32 * u < kgroups;
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                if (sizeof(T) == 4) {
                    v[4 * j] = __uint_as_float(raw[u][j].x); v[4 * j + 1] = __uint_as_float(raw[u][j].y);
                    v[4 * j + 2] = __uint_as_float(raw[u][j].z); v[4 * j + 3] = __uint_as_float(raw[u][j].w);
                } else {
                    const T* e = reinterpret_cast<const T*>(&raw[u][j]);
#pragma unroll
                    for (int t = 0; t < 8; ++t) v[8 * j + t] = to_f32(e[t]);
                }
            }
            if (!valid) {
#pragma unroll
                for (int t = 0; t < 16; ++t) v[t] = 0.0f;
            }
        };
        float amax = 0.0f;
#pragma unroll
        for (int u = 0; u < GC; ++u) {
            float v[16];
            unpack(u, v);
#pragma unroll
            for (int t = 0; t < 16; ++t) amax = fmaxf(amax, fabsf(v[t]));
        }
        amax = warp_max(amax);
        const float s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
        const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;
        int lsum = 0;
#pragma unroll
        for (int u = 0; u < GC; ++u) {
            const int64_t g = lane + 32 * u;
            if (g < groups) {
                float v[16];
                unpack(u, v);
                store16q(qr, g, v, inv, layout, lsum);
            }
        }
        lsum = warp_sum(lsum);
        if (lane == 0) {
            s_x[m] = s;
            if (sum_qx) sum_qx[m] = lsum;
        }
        return;
    }

    // generic path (unaligned rows, ragged K, or rows longer than the register cache): two passes, the second one
    // re-reads the row (L2)
    float amax = 0.0f;
    for (int64_t g = lane; g * 16 < K; g += 32) {
        float v[16];
        load16q<T>(xr, g, K, vec_ok, v);
#pragma unroll
        for (int t = 0; t < 16; ++t) amax = fmaxf(amax, fabsf(v[t]));
    }
    amax = warp_max(amax);
    const float s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;

    int lsum = 0;
    for (int64_t g = lane; g < groups; g += 32) {
        float v[16];
        load16q<T>(xr, g, K, vec_ok, v);
        store16q(qr, g, v, inv, layout, lsum);
    }
    lsum = warp_sum(lsum);
    if (lane == 0) {
        s_x[m] = s;
        if (sum_qx) sum_qx[m] = lsum;
    }
}

// Long rows (more than 32*GC groups): one CTA of 256 threads per row, up to 4 cached groups per thread
// (K <= 16384 stays a single HBM pass); absmax and sum meet in shared memory.
constexpr int GCB = 4;
template <typename T>
__global__ void __launch_bounds__(256)
quant_act_row_kernel(const T* __restrict__ x, int64_t K, int64_t ldx,
                     int8_t* __restrict__ q, int64_t ldq, float* __restrict__ s_x,
                     int32_t* __restrict__ sum_qx, int layout, bool vec_ok) {
    __shared__ float sh_max[8];
    __shared__ int sh_sum[8];
    asm volatile("griddepcontrol.launch_dependents;");      // see quant_act_kernel
    const int64_t m = blockIdx.x;
    const T* xr = x + m * ldx;
    int8_t* qr = q + m * ldq;
    const int64_t groups = ldq >> 4;
    const bool cached = groups <= 256 * GCB;
    float vc[GCB][16];
    float amax = 0.0f;
    if (cached) {
#pragma unroll
        for (int u = 0; u < GCB; ++u) {
            const int64_t g = threadIdx.x + 256 * u;
            if (g < groups) load16q<T>(xr, g, K, vec_ok, vc[u]);
        }
#pragma unroll
        for (int u = 0; u < GCB; ++u)
            if ((int64_t)threadIdx.x + 256 * u < groups) {
#pragma unroll
                for (int t = 0; t < 16; ++t) amax = fmaxf(amax, fabsf(vc[u][t]));
            }
    } else {
        for (int64_t g = threadIdx.x; g * 16 < K; g += 256) {
            float v[16];
            load16q<T>(xr, g, K, vec_ok, v);
#pragma unroll
            for (int t = 0; t < 16; ++t) amax = fmaxf(amax, fabsf(v[t]));
        }
    }
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0) sh_max[threadIdx.x >> 5] = amax;
    __syncthreads();
    amax = sh_max[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) amax = fmaxf(amax, sh_max[i]);
    const float s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;
    int lsum = 0;
    if (cached) {
#pragma unroll
        for (int u = 0; u < GCB; ++u) {
            const int64_t g = threadIdx.x + 256 * u;
            if (g < groups) store16q(qr, g, vc[u], inv, layout, lsum);
        }
    } else {
        for (int64_t g = threadIdx.x; g < groups; g += 256) {
            float v[16];
            load16q<T>(xr, g, K, vec_ok, v);
            store16q(qr, g, v, inv, layout, lsum);
        }
    }
    lsum = warp_sum(lsum);
    if ((threadIdx.x & 31) == 0) sh_sum[threadIdx.x >> 5] = lsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh_sum[i];
        s_x[m] = s;
        if (sum_qx) sum_qx[m] = t;
    }
}

// Specialised warp-per-row quantiser for the shapes the layers actually have: aligned rows of exactly K = 512 * GC
// elements, ldq == K, layout known at compile time.  No bounds predicates; ~8 instructions per element:
// the rounded value's low byte IS the int8 code (r + 1.5 * 2^23 keeps the two's-complement integer in the low mantissa
// bits), so packing is three PRMT per word and the row sum one dp4a per word.  Same bits as quant1 / store16q.
template <typename T, int GC, int LAYOUT>
__global__ void __launch_bounds__(QW * 32, GC == 8 ? 4 : (GC == 4 ? 6 : 8))
quant_act_fast_kernel(const T* __restrict__ x, int64_t M, int64_t ldx, int8_t* __restrict__ q,
                      float* __restrict__ s_x, int32_t* __restrict__ sum_qx) {
    asm volatile("griddepcontrol.launch_dependents;");      // see quant_act_kernel
    constexpr int K = 512 * GC;
    constexpr int NV = sizeof(T) == 4 ? 4 : 2;               // uint4 per 16 elements
    const int lane = threadIdx.x & 31;
    const int64_t m = (int64_t)blockIdx.x * QW + (threadIdx.x >> 5);
    if (m >= M) return;
    const uint4* xr = reinterpret_cast<const uint4*>(x + m * ldx);
    uint4 raw[GC][NV];
#pragma unroll
    for (int u = 0; u < GC; ++u)
#pragma unroll
        for (int j = 0; j < NV; ++j) raw[u][j] = __ldg(xr + (lane + 32 * u) * NV + j);
    auto unpack = [&](int u, float (&v)[16]) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const uint32_t w4[4] = {raw[u][j].x, raw[u][j].y, raw[u][j].z, raw[u][j].w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (sizeof(T) == 4) {
                    v[4 * j + c] = __uint_as_float(w4[c]);
                } else if (std::is_same<T, __nv_bfloat16>::value) {
                    v[8 * j + 2 * c] = __uint_as_float(w4[c] << 16);
                    v[8 * j + 2 * c + 1] = __uint_as_float(w4[c] & 0xffff0000u);
                } else {
                    const __half2 h = *reinterpret_cast<const __half2*>(&w4[c]);
                    const float2 f = __half22float2(h);
                    v[8 * j + 2 * c] = f.x; v[8 * j + 2 * c + 1] = f.y;
                }
            }
        }
    };
    float amax = 0.0f;
#pragma unroll
    for (int u = 0; u < GC; ++u) {
        float v[16];
        unpack(u, v);
#pragma unroll
        for (int t = 0; t < 16; ++t) amax = fmaxf(amax, fabsf(v[t]));
    }
    amax = warp_max(amax);
    const float s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;
    int lsum = 0;
    int8_t* qr = q + m * (int64_t)K;
#pragma unroll
    for (int u = 0; u < GC; ++u) {
        float v[16];
        unpack(u, v);
        uint32_t rb[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            float r = __fmul_rn(v[t], inv);
            r = fminf(fmaxf(r, -127.0f), 127.0f);
            rb[t] = __float_as_uint(__fadd_rn(r, 12582912.0f));
        }
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // natural: word i = elements 4i..4i+3 ; kperm: word i = elements i, 4+i, 8+i, 12+i
            const int e0 = LAYOUT != 0 ? i : 4 * i, st = LAYOUT != 0 ? 4 : 1;
            const uint32_t lo = __byte_perm(rb[e0], rb[e0 + st], 0x0040);
            const uint32_t hi = __byte_perm(rb[e0 + 2 * st], rb[e0 + 3 * st], 0x0040);
            w[i] = __byte_perm(lo, hi, 0x5410);
            lsum = __dp4a((int)w[i], 0x01010101, lsum);
        }
        const int g = lane + 32 * u;
        int pos = g;
        if (LAYOUT == 3) {
            int8_t* blk = qr + 256 * (g >> 4) + 8 * (g & 15);
            *reinterpret_cast<uint2*>(blk) = make_uint2(w[0], w[2]);
            *reinterpret_cast<uint2*>(blk + 128) = make_uint2(w[1], w[3]);
            continue;
        }
        if (LAYOUT == 2) {
            const int c = g >> 2, j = g & 3;
            pos = ((c >> 5) * 4 + j) * 32 + (c & 31);
        }
        *reinterpret_cast<uint4*>(qr + 16 * pos) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    lsum = warp_sum(lsum);
    if (lane == 0) {
        s_x[m] = s;
        if (sum_qx) sum_qx[m] = lsum;
    }
}

template <typename T, int GC>
static void launch_quant_fast(const T* x, int64_t M, int64_t ldx, int8_t* q, float* s_x, int32_t* sum_qx, int layout, cudaStream_t st) {
    const unsigned grid = (unsigned)((M + QW - 1) / QW);
    if (layout == 0) quant_act_fast_kernel<T, GC, 0><<<grid, QW * 32, 0, st>>>(x, M, ldx, q, s_x, sum_qx);
    else if (layout == 1) quant_act_fast_kernel<T, GC, 1><<<grid, QW * 32, 0, st>>>(x, M, ldx, q, s_x, sum_qx);
    else if (layout == 2) quant_act_fast_kernel<T, GC, 2><<<grid, QW * 32, 0, st>>>(x, M, ldx, q, s_x, sum_qx);
    else quant_act_fast_kernel<T, GC, 3><<<grid, QW * 32, 0, st>>>(x, M, ldx, q, s_x, sum_qx);
}

}  // namespace tb200

using namespace tb200;

extern "C" int tb200_quant_act_i8(const void* x, int dtype, int64_t M, int64_t K, int64_t ldx,
                                  int8_t* q_x, int64_t ldq, float* s_x, int32_t* sum_qx, int layout,
                                  void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0) return TB200_OK;
    TB_REQUIRE(x && q_x && s_x, "quant_act_i8: null pointer");
    TB_REQUIRE(M >= 0 && K > 0 && ldx >= K, "quant_act_i8: bad shape");
    TB_REQUIRE(ldq >= K && ldq % 16 == 0, "quant_act_i8: ldq=%lld must be >= K and a multiple of 16", (long long)ldq);
    TB_REQUIRE((uintptr_t)q_x % 16 == 0, "quant_act_i8: q_x must be 16-byte aligned");
    TB_REQUIRE(dtype_ok(dtype), "quant_act_i8: bad dtype %d", dtype);
    TB_REQUIRE(layout >= 0 && layout <= 3, "quant_act_i8: bad layout %d", layout);
    TB_REQUIRE(layout != 3 || ldq % 256 == 0, "quant_act_i8: layout 3 needs ldq %% 256 == 0");
    TB_REQUIRE(layout != 2 || ldq % 4096 == 0, "quant_act_i8: layout 2 needs ldq = 4096*ceil(K/4096)");
    const size_t es = dtype_size(dtype);
    const bool vec_ok = ((uintptr_t)x % 16 == 0) && ((ldx * es) % 16 == 0);
    // few rows: one CTA per row (the warp-per-row kernel would leave most SMs idle and is latency-bound per row)
    const int64_t warp_min_rows = tb_knob("TB200_QUANT_WARP_MIN_ROWS", 4 * (int64_t)sm_count());
    const int64_t groups = ldq >> 4;
    static const int fast_env = (int)tb_knob("TB200_QUANT_FAST", 1);
    if (fast_env && vec_ok && ldq == K && (K == 1024 || K == 2048 || K == 4096) && M >= warp_min_rows) {
        cudaStream_t st = (cudaStream_t)stream;
        if (K == 1024) { TB_DISPATCH_DTYPE(dtype, T, (launch_quant_fast<T, 2>((const T*)x, M, ldx, q_x, s_x, sum_qx, layout, st))); }
        else if (K == 2048) { TB_DISPATCH_DTYPE(dtype, T, (launch_quant_fast<T, 4>((const T*)x, M, ldx, q_x, s_x, sum_qx, layout, st))); }
        else { TB_DISPATCH_DTYPE(dtype, T, (launch_quant_fast<T, 8>((const T*)x, M, ldx, q_x, s_x, sum_qx, layout, st))); }
    } else if (groups <= 32 * GC_MAX && M >= warp_min_rows) {
        const unsigned grid = (unsigned)((M + QW - 1) / QW);
        cudaStream_t st = (cudaStream_t)stream;
        if (groups <= 64) {
            TB_DISPATCH_DTYPE(dtype, T, (quant_act_kernel<T, 2><<<grid, QW * 32, 0, st>>>((const T*)x, M, K, ldx, q_x, ldq, s_x, sum_qx, layout, vec_ok)));
        } else if (groups <= 128) {
            TB_DISPATCH_DTYPE(dtype, T, (quant_act_kernel<T, 4><<<grid, QW * 32, 0, st>>>((const T*)x, M, K, ldx, q_x, ldq, s_x, sum_qx, layout, vec_ok)));
        } else {
            TB_DISPATCH_DTYPE(dtype, T, (quant_act_kernel<T, 8><<<grid, QW * 32, 0, st>>>((const T*)x, M, K, ldx, q_x, ldq, s_x, sum_qx, layout, vec_ok)));
        }
    } else {
        TB_DISPATCH_DTYPE(dtype, T,
            (quant_act_row_kernel<T><<<(unsigned)M, 256, 0, (cudaStream_t)stream>>>(
                (const T*)x, K, ldx, q_x, ldq, s_x, sum_qx, layout, vec_ok)));
    }
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
