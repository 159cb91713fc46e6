// b1.cu — 1-bit packed weights for the binary layers.  ScaledBinaryLinear / ScaledBinary01Linear /
// SmoothlyBinarizedLinear (binary/modules.py:55-90, 112-146) have two-valued weights; the reference accounts them at one bit
// (binary/utils.py:42-45), and a decode GEMV is HBM-bound, so format "B1" stores them that way: row n, bit (k & 7) of byte
// k >> 3 = (w[n,k] >= 0) -- the decision rule of sign_ste and step_ste alike (-0.0 -> 1); the value is 2b - 1 (binary) or
// b (binary01).  Row pitch ld1 = multiple of 16 bytes >= ceil(K/8), padding bits 0 (they meet zero activations).
//
//   tb200_quantize_pack_b1 : latent weights -> B1                       (one pass, 16-byte loads, four in flight)
//   tb200_gemv_b1          : decode, M <= 8 tokens, on pre-quantised int8 activations (tb200_quant_act_i8, natural order):
//                            128-bit coalesced loads of packed rows, IN-PLACE bit masks (w & (0x01010101 << i): value
//                            bit * 2^i) against bit-plane ordered activations with dp4a -- eight accumulators per token,
//                            recombined exactly with shifts -- warp-shuffle reduction, fused epilogue
//                            ((2 acc - sum_qx) s_x scale + bias for binary, acc s_x scale + bias for binary01, clip);
//   tb200_expand_b1        : B1 -> the 2-bit format C (codes 0/2 or 1/2) for the tcgen05 GEMM (prefill) and the other
//                            2-bit kernels.
#include "common.cuh"

namespace tb200 {

constexpr int B1_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(B1_THREADS)
quantize_pack_b1_kernel(const T* __restrict__ w, int64_t K, int64_t ldw, uint8_t* __restrict__ packed, int64_t ld1, bool vec_ok) {
    constexpr int EPU = 16 / (int)sizeof(T);           // elements per 16-byte unit: 4 or 8
    const int64_t n = blockIdx.x;
    const T* wr = w + n * ldw;
    uint8_t* pr = packed + n * ld1;
    if (vec_ok) {
        // a thread owns one packed byte (8 weights = 2 or 1 units); four bytes per thread in flight
        const int64_t nb = K >> 3;
        for (int64_t b0 = threadIdx.x; b0 < nb; b0 += 4 * B1_THREADS) {
            uint4 v[4][8 / EPU];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t b = b0 + j * B1_THREADS;
#pragma unroll
                for (int u = 0; u < 8 / EPU; ++u)
                    v[j][u] = b < nb ? __ldg(reinterpret_cast<const uint4*>(wr + 8 * b) + u) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t b = b0 + j * B1_THREADS;
                if (b >= nb) continue;
                const T* e = reinterpret_cast<const T*>(&v[j][0]);
                uint32_t byte = 0;
#pragma unroll
                for (int i = 0; i < 8; ++i) byte |= (to_f32(e[i]) >= 0.0f ? 1u : 0u) << i;
                pr[b] = (uint8_t)byte;
            }
        }
        for (int64_t b = (K >> 3) + threadIdx.x; b < ld1; b += B1_THREADS) {     // ragged tail + padding
            uint32_t byte = 0;
            for (int i = 0; i < 8; ++i) { const int64_t k = 8 * b + i; if (k < K) byte |= (to_f32(wr[k]) >= 0.0f ? 1u : 0u) << i; }
            pr[b] = (uint8_t)byte;
        }
    } else {
        for (int64_t b = threadIdx.x; b < ld1; b += B1_THREADS) {
            uint32_t byte = 0;
            for (int i = 0; i < 8; ++i) { const int64_t k = 8 * b + i; if (k < K) byte |= (to_f32(wr[k]) >= 0.0f ? 1u : 0u) << i; }
            pr[b] = (uint8_t)byte;
        }
    }
}

// B1 -> format C (2 bits per weight, code = value + 1): one thread per B1 byte -> two format-C bytes
__global__ void __launch_bounds__(B1_THREADS)
expand_b1_kernel(const uint8_t* __restrict__ p1, int64_t ld1, int64_t N, int64_t K, int binary01, uint8_t* __restrict__ p2, int64_t ld2) {
    const int64_t per_row = ld2 >> 1, total = N * per_row;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = i / per_row, b = i - n * per_row;
        const uint32_t bits = b < ld1 ? p1[n * ld1 + b] : 0u;
        uint32_t out = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int64_t k = 8 * b + j;
            const uint32_t bit = (bits >> j) & 1u;
            const uint32_t code = k < K ? (binary01 ? 1u + bit : 2u * bit) : 1u;      // padding: code 1 (value 0)
            out |= code << (2 * j);
        }
        *reinterpret_cast<uint16_t*>(p2 + n * ld2 + 2 * b) = (uint16_t)out;
    }
}

__device__ __forceinline__ int dp4a_u8s8(uint32_t a_u8, uint32_t b_s8, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8), "r"(b_s8), "r"(c));
    return d;
}

// activation image in shared memory: for token m, 128-weight chunk c (= 16 bytes of a packed row), 32-bit word wd of the
// chunk and bit plane i, the word (q[k0 + i], q[k0 + 8 + i], q[k0 + 16 + i], q[k0 + 24 + i]), k0 = 128 c + 32 wd: the four
// activations that meet bit i of the four bytes of packed word wd.  Stored as uint4 (planes 4h .. 4h + 3) at
// ((m * 4 + wd) * 2 + h) * nchunks + c, so that consecutive lanes (chunks) read consecutive 16 bytes.
template <int M, int R>
__global__ void __launch_bounds__(B1_THREADS)
gemv_b1_kernel(const uint8_t* __restrict__ p1, int64_t ld1, int64_t N, int64_t K, const int8_t* __restrict__ q, int64_t ldq,
               const float* __restrict__ s_x, const int32_t* __restrict__ sum_qx, const float* __restrict__ scale,
               const float* __restrict__ bias, float clip_val, int binary01, void* __restrict__ out, int out_dtype, int64_t ldo,
               int pdl, const void* __restrict__ x, int x_dtype, int64_t ldx) {
    extern __shared__ uint4 img[];
    __shared__ float red_f[B1_THREADS / 32];
    __shared__ int red_i[B1_THREADS / 32];
    __shared__ float sx_sm[8];
    __shared__ int sq_sm[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunks = (int)((K + 127) >> 7);
    const int nj = (nchunks + 31) >> 5;                       // chunk iterations per row
    const int64_t total_warps = (int64_t)gridDim.x * (B1_THREADS / 32), gw = (int64_t)blockIdx.x * (B1_THREADS / 32) + warp;
    const int64_t ngroups = (N + R - 1) / R;                  // a warp takes R consecutive rows per step
    // the first rows' packed weights do not depend on the quantiser: requested before the grid dependency resolves
    uint4 wnext[R];
    {
        const int64_t r0 = gw * R;
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
            wnext[rr] = (gw < ngroups && r0 + rr < N && lane < nchunks) ? __ldg(reinterpret_cast<const uint4*>(p1 + (r0 + rr) * ld1) + lane)
                                                                         : make_uint4(0u, 0u, 0u, 0u);
    }
    if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (x) {
        // fused activation quantisation (every CTA quantises the M <= 2 rows for itself, the arithmetic of tb200_quant_act_i8:
        // absmax, 127 / amax, round-half-even), written into the bit-plane image
        uint8_t* img8 = reinterpret_cast<uint8_t*>(img);
        const bool fast = x_dtype != TB200_F32 && K % 16 == 0 && K <= 16 * B1_THREADS && (uintptr_t)x % 16 == 0 && (ldx * 2) % 16 == 0;
        auto ldx_f = [&](int m, int64_t k) -> float {
            const int64_t o = (int64_t)m * ldx + k;
            if (x_dtype == TB200_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[o]);
            if (x_dtype == TB200_F16) return __half2float(reinterpret_cast<const __half*>(x)[o]);
            return reinterpret_cast<const float*>(x)[o];
        };
        if (!fast) {
            for (int idx = threadIdx.x; idx < M * nchunks * 8; idx += B1_THREADS) img[idx] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
        }
#pragma unroll 1
        for (int m = 0; m < M; ++m) {
            // fast path (16-bit activations, K <= 4096, aligned): thread t owns elements 16 t .. 16 t + 15 = bytes b0, b0 + 1 of the
            // plane words of one 32-element group, kept in registers between the two passes; one 32-byte load per thread
            float v[16];
            const bool mine = fast && 16 * (int)threadIdx.x < K;
            float amax = 0.0f;
            if (fast) {
                if (mine) {
                    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(x) + (int64_t)m * ldx + 16 * threadIdx.x);
                    const uint4 a = src[0], b = src[1];
                    const uint32_t wv[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (x_dtype == TB200_BF16) { v[2 * j] = __uint_as_float(wv[j] << 16); v[2 * j + 1] = __uint_as_float(wv[j] & 0xffff0000u); }
                        else { const __half2 h = *reinterpret_cast<const __half2*>(&wv[j]); v[2 * j] = __low2float(h); v[2 * j + 1] = __high2float(h); }
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) amax = fmaxf(amax, fabsf(v[j]));
                }
            } else {
                for (int64_t k = threadIdx.x; k < K; k += B1_THREADS) amax = fmaxf(amax, fabsf(ldx_f(m, k)));
            }
            amax = warp_max(amax);
            if (lane == 0) red_f[warp] = amax;
            __syncthreads();
            amax = 0.0f;
#pragma unroll
            for (int i = 0; i < B1_THREADS / 32; ++i) amax = fmaxf(amax, red_f[i]);
            const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;
            int ls = 0;
            if (fast) {
                if (mine) {
                    uint32_t qb[16];                               // low byte = int8 code (rint through the 1.5 * 2^23 trick)
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float qf = __fmul_rn(v[j], inv);
                        qf = fminf(fmaxf(qf, -127.0f), 127.0f);
                        qb[j] = __float_as_uint(__fadd_rn(qf, 12582912.0f));
                        ls += (int)(int8_t)(qb[j] & 0xffu);
                    }
                    const int k0 = 16 * threadIdx.x;
                    const int c = k0 >> 7, wd = (k0 >> 5) & 3, bh = (k0 >> 4) & 1;        // bytes 2 bh, 2 bh + 1 of every plane word
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const uint32_t two = __byte_perm(qb[i], qb[8 + i], 0x0040);        // elements i, 8 + i: bytes b = 2 bh, 2 bh + 1
                        *reinterpret_cast<uint16_t*>(img8 + ((((m * 4 + wd) * 2 + (i >> 2)) * nchunks + c) << 4) + ((i & 3) << 2) + 2 * bh) = (uint16_t)two;
                    }
                } else if (16 * (int)threadIdx.x < 128 * nchunks) {                       // zero padding of the last chunk
                    const int k0 = 16 * threadIdx.x;
                    const int c = k0 >> 7, wd = (k0 >> 5) & 3, bh = (k0 >> 4) & 1;
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        *reinterpret_cast<uint16_t*>(img8 + ((((m * 4 + wd) * 2 + (i >> 2)) * nchunks + c) << 4) + ((i & 3) << 2) + 2 * bh) = 0;
                }
            } else {
                for (int64_t k = threadIdx.x; k < K; k += B1_THREADS) {
                    float qf = __fmul_rn(ldx_f(m, k), inv);
                    qf = fminf(fmaxf(qf, -127.0f), 127.0f);
                    const int qi = __float2int_rn(qf);
                    ls += qi;
                    const int c = (int)(k >> 7), wd = (int)(k >> 5) & 3, b = (int)(k >> 3) & 3, i = (int)k & 7;
                    img8[((((m * 4 + wd) * 2 + (i >> 2)) * nchunks + c) << 4) + ((i & 3) << 2) + b] = (uint8_t)qi;
                }
            }
            ls = warp_sum(ls);
            if (lane == 0) red_i[warp] = ls;
            __syncthreads();
            if (threadIdx.x == 0) {
                int t = 0;
                for (int i = 0; i < B1_THREADS / 32; ++i) t += red_i[i];
                sq_sm[m] = t;
                sx_sm[m] = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
            }
            __syncthreads();
        }
    }
    for (int idx = threadIdx.x; !x && idx < M * nchunks * 4; idx += B1_THREADS) {
        const int wd = idx & 3, c = (idx >> 2) % nchunks, m = (idx >> 2) / nchunks;
        const int64_t k0 = 128 * (int64_t)c + 32 * wd;
        uint32_t e[8];                                        // 32 natural-order int8 activations (zero beyond K)
        const int8_t* src = q + m * ldq + k0;
        if (k0 + 32 <= K) {
            const uint4 a = *reinterpret_cast<const uint4*>(src), b = *reinterpret_cast<const uint4*>(src + 16);
            e[0] = a.x; e[1] = a.y; e[2] = a.z; e[3] = a.w; e[4] = b.x; e[5] = b.y; e[6] = b.z; e[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                uint32_t v = 0;
#pragma unroll
                for (int t = 0; t < 4; ++t) { const int64_t k = k0 + 4 * j + t; if (k < K) v |= (uint32_t)(uint8_t)src[4 * j + t] << (8 * t); }
                e[j] = v;
            }
        }
        // e[2b], e[2b+1] hold elements 8b .. 8b+7: plane i takes byte (i & 3) of e[2b + (i >> 2)] for b = 0..3
        uint32_t pl[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int hw = i >> 2, by = i & 3;
            const uint32_t lo = __byte_perm(e[hw], e[2 + hw], by | ((4 + by) << 4));            // bytes b = 0, 1
            const uint32_t hi = __byte_perm(e[4 + hw], e[6 + hw], by | ((4 + by) << 4));        // bytes b = 2, 3
            pl[i] = __byte_perm(lo, hi, 0x5410);
        }
        img[((m * 4 + wd) * 2 + 0) * nchunks + c] = make_uint4(pl[0], pl[1], pl[2], pl[3]);
        img[((m * 4 + wd) * 2 + 1) * nchunks + c] = make_uint4(pl[4], pl[5], pl[6], pl[7]);
    }
    __syncthreads();
    float sx = 1.0f; int sq = 0;
    if (lane < M) { sx = x ? sx_sm[lane] : s_x[lane]; sq = x ? sq_sm[lane] : sum_qx[lane]; }
    for (int64_t grp = gw; grp < ngroups; grp += total_warps) {
        const int64_t r0 = grp * R;
        int acc[R][M][8];
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
            for (int m = 0; m < M; ++m)
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[rr][m][i] = 0;
        for (int j = 0; j < nj; ++j) {
            uint4 wv[R];
#pragma unroll
            for (int rr = 0; rr < R; ++rr) wv[rr] = wnext[rr];
            // next request: the following chunk iteration of these rows, or the first one of the warp's next rows
            {
                const bool same = j + 1 < nj;
                const int64_t rn = same ? r0 : r0 + total_warps * R;
                const int cn = (same ? 32 * (j + 1) : 0) + lane;
#pragma unroll
                for (int rr = 0; rr < R; ++rr)
                    wnext[rr] = (rn + rr < N && cn < nchunks) ? __ldg(reinterpret_cast<const uint4*>(p1 + (rn + rr) * ld1) + cn)
                                                              : make_uint4(0u, 0u, 0u, 0u);
            }
            const int c = 32 * j + lane;
            if (c < nchunks) {
#pragma unroll
                for (int wd = 0; wd < 4; ++wd) {
                    uint4 a[M][2];
#pragma unroll
                    for (int m = 0; m < M; ++m)
#pragma unroll
                        for (int h = 0; h < 2; ++h) a[m][h] = img[((m * 4 + wd) * 2 + h) * nchunks + c];
#pragma unroll
                    for (int rr = 0; rr < R; ++rr) {
                        const uint32_t word = wd == 0 ? wv[rr].x : wd == 1 ? wv[rr].y : wd == 2 ? wv[rr].z : wv[rr].w;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const uint32_t mk = word & (0x01010101u << i);          // bytes: bit * 2^i, in place
#pragma unroll
                            for (int m = 0; m < M; ++m) {
                                const uint4 av = a[m][i >> 2];
                                const uint32_t xw = (i & 3) == 0 ? av.x : (i & 3) == 1 ? av.y : (i & 3) == 2 ? av.z : av.w;
                                acc[rr][m][i] = dp4a_u8s8(mk, xw, acc[rr][m][i]);
                            }
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            const int64_t r = r0 + rr;
            int mine = 0;
#pragma unroll
            for (int m = 0; m < M; ++m) {
                int v = 0;
#pragma unroll
                for (int i = 0; i < 8; ++i) v += acc[rr][m][i] >> i;                 // exact: plane i is a multiple of 2^i
                v = warp_sum(v);
                if (lane == m) mine = v;
            }
            if (lane < M && r < N) {
                const int a = binary01 ? mine : 2 * mine - sq;                       // sum (2b - 1) q = 2 sum b q - sum q
                const float tv = __fmul_rn((float)a, sx);
                const float y = epilogue(tv, scale ? scale[r] : 1.0f, bias ? bias[r] : 0.0f, clip_val);
                const int64_t o = (int64_t)lane * ldo + r;
                if (out_dtype == TB200_BF16) reinterpret_cast<__nv_bfloat16*>(out)[o] = __float2bfloat16_rn(y);
                else if (out_dtype == TB200_F16) reinterpret_cast<__half*>(out)[o] = __float2half_rn(y);
                else reinterpret_cast<float*>(out)[o] = y;
            }
        }
    }
}

template <int M>
static int launch_gemv_b1(const uint8_t* p1, int64_t ld1, int64_t N, int64_t K, const int8_t* q, int64_t ldq, const float* s_x,
                          const int32_t* sum_qx, const float* scale, const float* bias, float clip_val, int binary01, void* out,
                          int out_dtype, int64_t ldo, int flags, cudaStream_t s, const void* x = nullptr, int x_dtype = 0, int64_t ldx = 0) {
    const int nchunks = (int)((K + 127) / 128);
    const size_t smem = (size_t)M * nchunks * 8 * sizeof(uint4);
    constexpr int R = M <= 2 ? 2 : 1;
    auto launch = [&](auto kern, SmemOptIn& optin) -> int {
        TB_CUDA(ensure_dynamic_smem(kern, optin, smem));
        const int64_t groups = (N + R - 1) / R;
        int64_t ctas = (groups + B1_THREADS / 32 - 1) / (B1_THREADS / 32);
        int per_sm = 1;
        TB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, B1_THREADS, smem));
        const int64_t cap = (int64_t)sm_count() * (per_sm < 1 ? 1 : per_sm);      // one resident wave: every warp strides over its rows
        if (ctas > cap) ctas = cap;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)ctas, 1, 1);
        cfg.blockDim = dim3(B1_THREADS, 1, 1);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = (flags & TB200_FLAG_PDL) ? 1 : 0;
        TB_CUDA(cudaLaunchKernelEx(&cfg, kern, p1, ld1, N, K, q, ldq, s_x, sum_qx, scale, bias, clip_val, binary01, out, out_dtype, ldo,
                                   (flags & TB200_FLAG_PDL) ? 1 : 0, x, x_dtype, ldx));
        TB_LAUNCH_CHECK();
        return TB200_OK;
    };
    // The lane's activation words come from shared memory on every use: keeping them in registers (32 per token) halves the
    // resident warps and with them the weight bytes in flight (measured: 6.9 us per 4096 x 11008 layer against 5.7 us).
    static SmemOptIn optin;
    return launch(gemv_b1_kernel<M, R>, optin);
}

}  // namespace tb200

using namespace tb200;

extern "C" int64_t tb200_b1_pitch(int64_t K) { return K <= 0 ? 16 : ((K + 127) / 128) * 16; }

extern "C" int tb200_quantize_pack_b1(const void* w, int dtype, int64_t N, int64_t K, int64_t ldw, uint8_t* packed, int64_t ld1,
                                      void* stream) {
    TB_DEVICE_OR_RETURN();
    if (N == 0) return TB200_OK;
    TB_REQUIRE(w && packed, "quantize_pack_b1: null pointer");
    TB_REQUIRE(N > 0 && K >= 0 && ldw >= K, "quantize_pack_b1: bad shape");
    TB_REQUIRE(ld1 >= tb200_b1_pitch(K) && ld1 % 16 == 0, "quantize_pack_b1: ld1 must be a multiple of 16 >= 16*ceil(K/128)");
    TB_REQUIRE(dtype_ok(dtype), "quantize_pack_b1: bad dtype %d", dtype);
    const size_t es = dtype_size(dtype);
    const bool vec_ok = (uintptr_t)w % 16 == 0 && (ldw * es) % 16 == 0;
    TB_DISPATCH_DTYPE(dtype, T,
        (quantize_pack_b1_kernel<T><<<(unsigned)N, B1_THREADS, 0, (cudaStream_t)stream>>>((const T*)w, K, ldw, packed, ld1, vec_ok)));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_expand_b1(const uint8_t* packed1, int64_t ld1, int64_t N, int64_t K, int binary01, uint8_t* packed2,
                               int64_t ld2, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (N == 0) return TB200_OK;
    TB_REQUIRE(packed1 && packed2, "expand_b1: null pointer");
    TB_REQUIRE(N > 0 && K > 0 && ld1 >= (K + 7) / 8 && ld2 >= (K + 3) / 4 && ld2 % 2 == 0, "expand_b1: bad shape / pitch");
    const int64_t total = N * (ld2 / 2);
    int64_t ctas = (total + B1_THREADS - 1) / B1_THREADS;
    if (ctas > (int64_t)sm_count() * 8) ctas = (int64_t)sm_count() * 8;
    expand_b1_kernel<<<(unsigned)ctas, B1_THREADS, 0, (cudaStream_t)stream>>>(packed1, ld1, N, K, binary01, packed2, ld2);
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

// x [M, ldx] (bf16 / f16 / f32), M <= 2: the activation quantisation runs in the kernel's prologue (one launch per layer)
extern "C" int tb200_gemv_b1_fused(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, const uint8_t* packed1, int64_t ld1,
                                   int64_t N, const float* scale, const float* bias, float clip_val, int binary01, void* out,
                                   int out_dtype, int64_t ldo, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(x && packed1 && out, "gemv_b1_fused: null pointer");
    TB_REQUIRE(M >= 1 && M <= 2, "gemv_b1_fused: M=%lld out of range [1,2]", (long long)M);
    TB_REQUIRE(K > 0 && ldx >= K && dtype_ok(x_dtype) && dtype_ok(out_dtype) && ldo >= N, "gemv_b1_fused: bad shape / dtype");
    TB_REQUIRE(ld1 >= tb200_b1_pitch(K) && ld1 % 16 == 0 && (uintptr_t)packed1 % 16 == 0, "gemv_b1_fused: bad packed layout");
    TB_REQUIRE((size_t)M * ((K + 127) / 128) * 128 <= 200 * 1024, "gemv_b1_fused: K too large for the shared-memory activation image");
    cudaStream_t s = (cudaStream_t)stream;
    if (M == 1) return launch_gemv_b1<1>(packed1, ld1, N, K, nullptr, 0, nullptr, nullptr, scale, bias, fabsf(clip_val), binary01, out, out_dtype, ldo, flags, s, x, x_dtype, ldx);
    return launch_gemv_b1<2>(packed1, ld1, N, K, nullptr, 0, nullptr, nullptr, scale, bias, fabsf(clip_val), binary01, out, out_dtype, ldo, flags, s, x, x_dtype, ldx);
}

extern "C" int tb200_gemv_b1(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx, int64_t M, int64_t K,
                             const uint8_t* packed1, int64_t ld1, int64_t N, const float* scale, const float* bias, float clip_val,
                             int binary01, void* out, int out_dtype, int64_t ldo, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(q_x && s_x && sum_qx && packed1 && out, "gemv_b1: null pointer");
    TB_REQUIRE(M >= 1 && M <= TB200_GEMV_FUSED_M, "gemv_b1: M=%lld out of range [1,%d]", (long long)M, TB200_GEMV_FUSED_M);
    TB_REQUIRE(K > 0 && ldq >= K && ldq % 16 == 0 && (uintptr_t)q_x % 16 == 0, "gemv_b1: q_x rows must be 16-byte aligned (ldq % 16 == 0)");
    TB_REQUIRE(ld1 >= tb200_b1_pitch(K) && ld1 % 16 == 0 && (uintptr_t)packed1 % 16 == 0, "gemv_b1: bad packed layout");
    TB_REQUIRE(ldo >= N && dtype_ok(out_dtype), "gemv_b1: bad output");
    TB_REQUIRE((size_t)M * ((K + 127) / 128) * 128 <= 200 * 1024, "gemv_b1: M*K too large for the shared-memory activation image");
    cudaStream_t s = (cudaStream_t)stream;
    const float cv = fabsf(clip_val);
#define B1_CASE(MM) case MM: return launch_gemv_b1<MM>(packed1, ld1, N, K, q_x, ldq, s_x, sum_qx, scale, bias, cv, binary01, out, out_dtype, ldo, flags, s)
    switch ((int)M) {
        B1_CASE(1); B1_CASE(2); B1_CASE(3); B1_CASE(4); B1_CASE(5); B1_CASE(6); B1_CASE(7);
        default: return launch_gemv_b1<8>(packed1, ld1, N, K, q_x, ldq, s_x, sum_qx, scale, bias, cv, binary01, out, out_dtype, ldo, flags, s);
    }
#undef B1_CASE
}
