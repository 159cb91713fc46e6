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
#include "common.cuh"

namespace tb200 {

// rint + float->int via the 1.5*2^23 trick (no XU-pipe conversions); same bits as rintf + cast
__device__ __forceinline__ int quant1(float v, float inv) {
    float r = __fmul_rn(v, inv);
    r = fminf(fmaxf(r, -127.0f), 127.0f);
    return __float_as_int(__fadd_rn(r, 12582912.0f)) - 0x4B400000;
}

// one CTA per token row; thread owns 16-element groups (one 16-byte store each)
template <typename T>
__global__ void __launch_bounds__(256)
quant_act_kernel(const T* __restrict__ x, int64_t K, int64_t ldx,
                 int8_t* __restrict__ q, int64_t ldq, float* __restrict__ s_x,
                 int32_t* __restrict__ sum_qx, int layout, bool vec_ok) {
    const int kperm = layout != 0;
    __shared__ float sh_max[8];
    __shared__ int sh_sum[8];
    const int64_t m = blockIdx.x;
    const T* xr = x + m * ldx;
    int8_t* qr = q + m * ldq;
    const int64_t groups = ldq >> 4;          // ldq is a multiple of 16
    const int64_t full = K >> 4;              // groups entirely in range

    float amax = 0.0f;
    for (int64_t g = threadIdx.x; g < groups; g += blockDim.x) {
        if (g < full && vec_ok) {
            if (sizeof(T) == 4) {
                const float4* p = reinterpret_cast<const float4*>(xr + 16 * g);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float4 f = p[j];
                    amax = fmaxf(amax, fmaxf(fmaxf(fabsf(f.x), fabsf(f.y)), fmaxf(fabsf(f.z), fabsf(f.w))));
                }
            } else {
                const uint4* p = reinterpret_cast<const uint4*>(xr + 16 * g);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    uint4 u = p[j];
                    const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
                    for (int t = 0; t < 8; ++t) amax = fmaxf(amax, fabsf(to_f32(e[t])));
                }
            }
        } else {
            for (int t = 0; t < 16; ++t) {
                int64_t k = 16 * g + t;
                if (k < K) amax = fmaxf(amax, fabsf(to_f32(xr[k])));
            }
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
    for (int64_t g = threadIdx.x; g < groups; g += blockDim.x) {
        int qv[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            int64_t k = 16 * g + t;
            qv[t] = k < K ? quant1(to_f32(xr[k]), inv) : 0;
            lsum += qv[t];
        }
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // natural: word i = elements 4i..4i+3 ; kperm: word i = elements i, 4+i, 8+i, 12+i
            int e0 = kperm ? i : 4 * i, st = kperm ? 4 : 1;
            w[i] = (uint32_t)(qv[e0] & 0xff) | ((uint32_t)(qv[e0 + st] & 0xff) << 8) |
                   ((uint32_t)(qv[e0 + 2 * st] & 0xff) << 16) | ((uint32_t)(qv[e0 + 3 * st] & 0xff) << 24);
        }
        int64_t pos = g;
        if (layout == 2) {
            const int64_t c = g >> 2, j = g & 3;
            pos = ((c >> 5) * 4 + j) * 32 + (c & 31);
        }
        *reinterpret_cast<uint4*>(qr + 16 * pos) = make_uint4(w[0], w[1], w[2], w[3]);
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
    TB_REQUIRE(layout >= 0 && layout <= 2, "quant_act_i8: bad layout %d", layout);
    TB_REQUIRE(layout != 2 || ldq % 4096 == 0, "quant_act_i8: layout 2 needs ldq = 4096*ceil(K/4096)");
    const size_t es = dtype_size(dtype);
    const bool vec_ok = ((uintptr_t)x % 16 == 0) && ((ldx * es) % 16 == 0);
    TB_DISPATCH_DTYPE(dtype, T,
        (quant_act_kernel<T><<<(unsigned)M, 256, 0, (cudaStream_t)stream>>>(
            (const T*)x, K, ldx, q_x, ldq, s_x, sum_qx, layout, vec_ok)));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
