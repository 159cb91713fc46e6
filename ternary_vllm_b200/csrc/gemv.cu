// gemv.cu — decode-shaped packed-ternary GEMV (M <= 32 tokens).  HBM-bound integer work:
//   * the packed weight rows are streamed once with 128-bit coalesced loads
//     (lane l of a warp owns 16-byte chunk l of every 512-byte slab of a row);
//   * the first slab of every row is requested BEFORE the activation prologue, so its DRAM
//     latency overlaps the per-token int8 quantisation (and, under programmatic dependent
//     launch, the tail of the previous layer's kernel);
//   * unpacking is 4 AND ops per 16 weights: field i of every byte is masked in place
//     (value code*4^i), multiplied with dp4a(u8 x s8) into its own accumulator, and the four
//     accumulators are recombined exactly with shifts at the end;
//   * codes are used unsigned (w+1 in {0,1,2}); sum_k q_x is subtracted once per token;
//   * warp-shuffle reduction, then the fused scale/bias/clip epilogue.
// Reference semantics: ScaledTernaryLinear.forward (binary/modules.py:101-109) and
// ops.ternary_matmul (impl/tf/moondream/ternary.py:279-281).
#include "common.cuh"

namespace tb200 {

constexpr int GEMV_WARPS = 4;
constexpr int GEMV_THREADS = GEMV_WARPS * 32;

struct GemvParams {
    const void* x; int64_t ldx; int64_t stride_x;      // activations [M,K]
    const uint8_t* packed; int64_t ld_packed; int64_t stride_packed;
    const float* scale; const float* bias; int64_t stride_sb;
    void* out; int64_t ldo; int64_t stride_out;
    int32_t* acc_out; float* sx_out; int8_t* qx_out;
    int64_t N, K;
    int M;            // tokens in this launch (<= MT)
    int cblocks;      // ceil(ceil(K/64)/32): 512-byte slabs per packed row
    int kc;           // ceil(K/64): valid 16-byte chunks per row
    float clip_val;
    int pdl;
    bool xvec;        // x rows are 16-byte aligned -> vector loads in the prologue
};

__device__ __forceinline__ int dp4a_us(uint32_t a_u8, uint32_t b_s8, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8), "r"(b_s8), "r"(c));
    return d;
}
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ int quant1g(float v, float s) {
    float r = rintf(__fdiv_rn(v, s));
    return (int)fminf(fmaxf(r, -127.0f), 127.0f);
}

// 16 consecutive activations of one token as fp32 (zero beyond K); vector loads when aligned
template <typename XT>
__device__ __forceinline__ void load16(const XT* __restrict__ xr, int64_t g, int64_t K, bool vec, float v[16]) {
    if (vec && 16 * g + 16 <= K) {
        if (sizeof(XT) == 4) {
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
                const XT* e = reinterpret_cast<const XT*>(&u);
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

template <typename XT, typename OT, int MT, int R>
__global__ void __launch_bounds__(GEMV_THREADS)
gemv_dp4a_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    // layout: xq[MT][cblocks*128] uint4 | sx[MT] | sumq[MT] | red_f[MT][WARPS] | red_i[MT][WARPS]
    uint4* xq = reinterpret_cast<uint4*>(smem_raw);
    const int slab_u4 = p.cblocks * 128;                      // uint4 per token
    float* sh_sx = reinterpret_cast<float*>(xq + (size_t)MT * slab_u4);
    int* sh_sumq = reinterpret_cast<int*>(sh_sx + MT);
    float* red_f = reinterpret_cast<float*>(sh_sumq + MT);
    int* red_i = reinterpret_cast<int*>(red_f + MT * GEMV_WARPS);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int layer = blockIdx.y;
    const XT* x = reinterpret_cast<const XT*>(p.x) + layer * p.stride_x;
    const uint8_t* packed = p.packed + layer * p.stride_packed;
    const float* scale = p.scale ? p.scale + layer * p.stride_sb : nullptr;
    const float* bias = p.bias ? p.bias + layer * p.stride_sb : nullptr;
    OT* out = reinterpret_cast<OT*>(p.out) + layer * p.stride_out;

    const int64_t row0 = ((int64_t)blockIdx.x * GEMV_WARPS + warp) * R;
    const uint4* wrow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int64_t row = row0 + r < p.N ? row0 + r : p.N - 1;   // clamp: loads stay in bounds, store is skipped
        wrow[r] = reinterpret_cast<const uint4*>(packed + row * p.ld_packed);
    }

    // ---- weight prefetch of slab 0: independent of the activations ---------------------
    uint4 wcur[R];
    {
        const bool valid = lane < p.kc;
#pragma unroll
        for (int r = 0; r < R; ++r) wcur[r] = valid ? ldg_stream(wrow[r] + lane) : make_uint4(0, 0, 0, 0);
    }
    if (p.pdl) {
        // let the next layer's kernel start its own weight prefetch, then wait for the
        // producer of x (the previous kernel in the stream) to finish
        asm volatile("griddepcontrol.launch_dependents;");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }

    // ---- prologue: per-token int8 quantisation of x into shared memory -----------------
    const int groups = p.cblocks * 128;          // 16-element groups incl. padding
    float amax[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) amax[m] = 0.0f;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        if (m < p.M) {
            const XT* xr = x + (int64_t)m * p.ldx;
            for (int g = threadIdx.x; (int64_t)g * 16 < p.K; g += GEMV_THREADS) {
                float v[16];
                load16<XT>(xr, g, p.K, p.xvec, v);
#pragma unroll
                for (int t = 0; t < 16; ++t) amax[m] = fmaxf(amax[m], fabsf(v[t]));
            }
        }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float v = warp_max(amax[m]);
        if (lane == 0) red_f[m * GEMV_WARPS + warp] = v;
    }
    __syncthreads();
    float sx[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float v = red_f[m * GEMV_WARPS];
#pragma unroll
        for (int w = 1; w < GEMV_WARPS; ++w) v = fmaxf(v, red_f[m * GEMV_WARPS + w]);
        sx[m] = v > 0.0f ? __fdiv_rn(v, 127.0f) : 1.0f;
    }
    int lsum[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) lsum[m] = 0;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const XT* xr = x + (int64_t)m * p.ldx;
        for (int g = threadIdx.x; g < groups; g += GEMV_THREADS) {
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            if (m < p.M && (int64_t)g * 16 < p.K) {
                float v[16];
                load16<XT>(xr, g, p.K, p.xvec, v);
                int qv[16];
#pragma unroll
                for (int t = 0; t < 16; ++t) {
                    qv[t] = quant1g(v[t], sx[m]);     // v is 0 beyond K -> q = 0
                    lsum[m] += qv[t];
                }
                if (p.qx_out && blockIdx.x == 0 && layer == 0) {
#pragma unroll
                    for (int t = 0; t < 16; ++t) {
                        int64_t k = 16 * (int64_t)g + t;
                        if (k < p.K) p.qx_out[(int64_t)m * p.K + k] = (int8_t)qv[t];
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)   // word i = elements i, 4+i, 8+i, 12+i (matches field i of bytes 0..3)
                    w[i] = (uint32_t)(qv[i] & 0xff) | ((uint32_t)(qv[4 + i] & 0xff) << 8) |
                           ((uint32_t)(qv[8 + i] & 0xff) << 16) | ((uint32_t)(qv[12 + i] & 0xff) << 24);
            }
            // chunk c = g/4 -> (slab cb = c/32, lane c%32); word j = g%4 of the chunk
            const int c = g >> 2, j = g & 3;
            xq[(size_t)m * slab_u4 + ((c >> 5) * 4 + j) * 32 + (c & 31)] = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        int v = warp_sum(lsum[m]);
        if (lane == 0) red_i[m * GEMV_WARPS + warp] = v;
    }
    __syncthreads();
    if (threadIdx.x < MT) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < GEMV_WARPS; ++w) t += red_i[threadIdx.x * GEMV_WARPS + w];
        sh_sumq[threadIdx.x] = t;
        sh_sx[threadIdx.x] = sx[0];   // overwritten just below with the right token's value
    }
#pragma unroll
    for (int m = 0; m < MT; ++m)
        if (threadIdx.x == m) {
            sh_sx[m] = sx[m];
            if (p.sx_out && blockIdx.x == 0 && layer == 0 && m < p.M) p.sx_out[m] = sx[m];
        }
    __syncthreads();

    // ---- main loop over 512-byte slabs ---------------------------------------------------
    int acc[R][MT][4];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[r][m][i] = 0;

    for (int cb = 0; cb < p.cblocks; ++cb) {
        uint4 wnext[R];
        {
            const int c = (cb + 1) * 32 + lane;
            const bool valid = (cb + 1 < p.cblocks) && (c < p.kc);
#pragma unroll
            for (int r = 0; r < R; ++r) wnext[r] = valid ? ldg_stream(wrow[r] + c) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint4 xv[MT];
#pragma unroll
            for (int m = 0; m < MT; ++m) xv[m] = xq[(size_t)m * slab_u4 + (cb * 4 + j) * 32 + lane];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t w = j == 0 ? wcur[r].x : j == 1 ? wcur[r].y : j == 2 ? wcur[r].z : wcur[r].w;
                const uint32_t f0 = w & 0x03030303u, f1 = w & 0x0c0c0c0cu;
                const uint32_t f2 = w & 0x30303030u, f3 = w & 0xc0c0c0c0u;
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    acc[r][m][0] = dp4a_us(f0, xv[m].x, acc[r][m][0]);
                    acc[r][m][1] = dp4a_us(f1, xv[m].y, acc[r][m][1]);
                    acc[r][m][2] = dp4a_us(f2, xv[m].z, acc[r][m][2]);
                    acc[r][m][3] = dp4a_us(f3, xv[m].w, acc[r][m][3]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) wcur[r] = wnext[r];
    }

    // ---- recombine fields (exact: acc_i is a multiple of 4^i), reduce, epilogue ----------
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            int v = acc[r][m][0] + (acc[r][m][1] >> 2) + (acc[r][m][2] >> 4) + (acc[r][m][3] >> 6);
            v = warp_sum(v);
            const int64_t row = row0 + r;
            if (lane == 0 && row < p.N && m < p.M) {
                const int a = v - sh_sumq[m];
                if (p.acc_out && layer == 0) p.acc_out[(int64_t)m * p.N + row] = a;
                const float t = __fmul_rn((float)a, sh_sx[m]);
                const float y = epilogue(t, scale ? scale[row] : 1.0f, bias ? bias[row] : 0.0f, p.clip_val);
                out[(int64_t)m * p.ldo + row] = from_f32<OT>(y);
            }
        }
    }
}

// ---- float-activation variant: the reference's exact semantics (no activation quantisation)
template <typename XT, typename OT, int MT>
__global__ void __launch_bounds__(GEMV_THREADS)
gemv_f32_kernel(const XT* __restrict__ x, int64_t ldx, int M, int64_t K,
                const uint8_t* __restrict__ packed, int64_t ld_packed, int64_t N,
                const float* __restrict__ scale, const float* __restrict__ bias, float clip_val,
                OT* __restrict__ out, int64_t ldo) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t)blockIdx.x * GEMV_WARPS + warp;
    if (row >= N) return;
    const uint8_t* pr = packed + row * ld_packed;
    float acc[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m] = 0.0f;
    const int64_t kb = (K + 3) >> 2;
    for (int64_t b = lane; b < kb; b += 32) {
        const uint32_t byte = pr[b];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t k = 4 * b + i;
            const float wv = (float)((int)((byte >> (2 * i)) & 3u) - 1);
            if (k < K) {
#pragma unroll
                for (int m = 0; m < MT; ++m)
                    if (m < M) acc[m] = fmaf(to_f32(x[(int64_t)m * ldx + k]), wv, acc[m]);
            }
        }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float v = warp_sum(acc[m]);
        if (lane == 0 && m < M) {
            // F.linear output is rounded to x.dtype before the in-place scale/bias (modules.py:104-107)
            float t = to_f32(from_f32<XT>(v));
            out[(int64_t)m * ldo + row] =
                from_f32<OT>(epilogue(t, scale ? scale[row] : 1.0f, bias ? bias[row] : 0.0f, clip_val));
        }
    }
}

template <typename XT, typename OT, int MT, int R>
static int launch_gemv(GemvParams p, int layers, cudaStream_t s) {
    const size_t smem = (size_t)MT * p.cblocks * 128 * sizeof(uint4) + MT * 8 + MT * GEMV_WARPS * 8 + 16;
    auto kern = gemv_dp4a_kernel<XT, OT, MT, R>;
    if (smem > 48 * 1024) {
        static size_t configured = 0;   // per instantiation
        if (smem > configured) {
            TB_REQUIRE(smem <= 227 * 1024, "gemv: K=%lld needs %zu bytes of shared memory for %d tokens",
                       (long long)p.K, smem, MT);
            TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured = smem;
        }
    }
    const int64_t rows_per_cta = (int64_t)GEMV_WARPS * R;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((p.N + rows_per_cta - 1) / rows_per_cta), (unsigned)layers, 1);
    cfg.blockDim = dim3(GEMV_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

template <typename XT, typename OT>
static int gemv_dispatch_mt(GemvParams p, int layers, cudaStream_t s) {
    if (p.M == 1) return launch_gemv<XT, OT, 1, 4>(p, layers, s);
    if (p.M == 2) return launch_gemv<XT, OT, 2, 4>(p, layers, s);
    if (p.M <= 4) return launch_gemv<XT, OT, 4, 2>(p, layers, s);
    return launch_gemv<XT, OT, 8, 2>(p, layers, s);
}

static int gemv_dispatch(int x_dtype, int out_dtype, GemvParams p, int layers, cudaStream_t s) {
#define TB_GEMV_CASE(XD, XT, OD, OT) \
    if (x_dtype == XD && out_dtype == OD) return gemv_dispatch_mt<XT, OT>(p, layers, s)
    TB_GEMV_CASE(TB200_BF16, __nv_bfloat16, TB200_BF16, __nv_bfloat16);
    TB_GEMV_CASE(TB200_BF16, __nv_bfloat16, TB200_F32, float);
    TB_GEMV_CASE(TB200_F32, float, TB200_F32, float);
    TB_GEMV_CASE(TB200_F16, __half, TB200_F16, __half);
    TB_GEMV_CASE(TB200_F16, __half, TB200_F32, float);
#undef TB_GEMV_CASE
    set_error("gemv: unsupported dtype pair x=%d out=%d (out must be the x dtype or f32)", x_dtype, out_dtype);
    return TB200_ENOTSUP;
}

static int gemv_common(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                       const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                       const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                       void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                       int32_t* acc_out, float* sx_out, int8_t* qx_out, int layers, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out, "gemv: null pointer");
    TB_REQUIRE(M >= 0 && M <= TB200_GEMV_MAX_M, "gemv: M=%lld out of range [0,%d]", (long long)M, TB200_GEMV_MAX_M);
    TB_REQUIRE(K > 0 && N >= 0 && ldx >= K && ldo >= N, "gemv: bad shape M=%lld K=%lld N=%lld", (long long)M, (long long)K, (long long)N);
    const int64_t kc = (K + 63) / 64;
    TB_REQUIRE(ld_packed % 16 == 0 && ld_packed >= kc * 16,
               "gemv: ld_packed=%lld must be a multiple of 16 and >= 16*ceil(K/64)=%lld", (long long)ld_packed, (long long)(kc * 16));
    TB_REQUIRE((uintptr_t)packed % 16 == 0 && stride_packed % 16 == 0, "gemv: packed weights must be 16-byte aligned");
    TB_REQUIRE(layers >= 1 && layers <= 65535, "gemv: bad group count %d", layers);
    if (M == 0 || N == 0) return TB200_OK;
    const size_t xs = dtype_size(x_dtype), os = dtype_size(out_dtype);
    GemvParams p;
    p.ldx = ldx; p.stride_x = stride_x;
    p.packed = packed; p.ld_packed = ld_packed; p.stride_packed = stride_packed;
    p.scale = scale; p.bias = bias; p.stride_sb = stride_sb;
    p.ldo = ldo; p.stride_out = stride_out;
    p.N = N; p.K = K; p.kc = (int)kc; p.cblocks = (int)((kc + 31) / 32);
    p.clip_val = fabsf(clip_val); p.pdl = flags & 1;
    p.xvec = ((uintptr_t)x % 16 == 0) && ((ldx * xs) % 16 == 0) && ((stride_x * xs) % 16 == 0);
    for (int64_t m0 = 0; m0 < M; m0 += 8) {   // passes of <= 8 tokens (quantisation is per token)
        p.M = (int)(M - m0 < 8 ? M - m0 : 8);
        p.x = (const char*)x + m0 * ldx * xs;
        p.out = (char*)out + m0 * ldo * os;
        p.acc_out = acc_out ? acc_out + m0 * N : nullptr;
        p.sx_out = sx_out ? sx_out + m0 : nullptr;
        p.qx_out = qx_out ? qx_out + m0 * K : nullptr;
        int rc = gemv_dispatch(x_dtype, out_dtype, p, layers, (cudaStream_t)stream);
        if (rc != TB200_OK) return rc;
    }
    return TB200_OK;
}

}  // namespace tb200

using namespace tb200;

extern "C" int tb200_gemv(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                          const uint8_t* packed, int64_t ld_packed, int64_t N,
                          const float* scale, const float* bias, float clip_val,
                          void* out, int out_dtype, int64_t ldo,
                          int32_t* acc_out, float* sx_out, int8_t* qx_out, int flags, void* stream) {
    return gemv_common(x, x_dtype, M, K, ldx, 0, packed, ld_packed, 0, N, scale, bias, 0, clip_val,
                       out, out_dtype, ldo, 0, acc_out, sx_out, qx_out, 1, flags, stream);
}

extern "C" int tb200_gemv_grouped(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                                  const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                                  const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                  void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                                  int count, int flags, void* stream) {
    return gemv_common(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N, scale, bias,
                       stride_sb, clip_val, out, out_dtype, ldo, stride_out, nullptr, nullptr, nullptr,
                       count, flags, stream);
}

extern "C" int tb200_gemv_f32(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                              const uint8_t* packed, int64_t ld_packed, int64_t N,
                              const float* scale, const float* bias, float clip_val,
                              void* out, int out_dtype, int64_t ldo, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out, "gemv_f32: null pointer");
    TB_REQUIRE(M >= 0 && M <= TB200_GEMV_MAX_M, "gemv_f32: M=%lld out of range", (long long)M);
    TB_REQUIRE(K > 0 && N >= 0 && ldx >= K && ldo >= N && ld_packed >= (K + 3) / 4, "gemv_f32: bad shape");
    TB_REQUIRE(x_dtype == out_dtype, "gemv_f32: out dtype must equal x dtype");
    if (M == 0 || N == 0) return TB200_OK;
    const size_t es = dtype_size(x_dtype);
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((unsigned)((N + GEMV_WARPS - 1) / GEMV_WARPS));
    for (int64_t m0 = 0; m0 < M; m0 += 8) {
        const int mt = (int)(M - m0 < 8 ? M - m0 : 8);
        const char* xp = (const char*)x + m0 * ldx * es;
        char* op = (char*)out + m0 * ldo * es;
        TB_DISPATCH_DTYPE(x_dtype, T,
            (gemv_f32_kernel<T, T, 8><<<grid, GEMV_THREADS, 0, s>>>((const T*)xp, ldx, mt, K, packed, ld_packed, N,
                                                                   scale, bias, fabsf(clip_val), (T*)op, ldo)));
        TB_LAUNCH_CHECK();
    }
    return TB200_OK;
}
