// init.cu — scale initialisation of linear_to_quantized (binary/modules.py:189-299).
// Reductions are accumulated in fp64 in a fixed order (deterministic); every elementwise
// decision reproduces the dtype arithmetic torch performs (see comments).
#include "common.cuh"

namespace tb200 {

struct InitWs {          // workspace layout for N rows
    double* sum;         // [N] row partial sums  (binary: positive part)
    double* sum2;        // [N] (binary: negative part)
    long long* cnt;      // [N] row partial counts
    float* m;            // [1] current tensor-wide mean ("uniform")
};
__host__ __device__ inline InitWs carve(void* ws, int64_t N) {
    InitWs r;
    char* p = (char*)ws;
    r.sum = (double*)p;  p += sizeof(double) * N;
    r.sum2 = (double*)p; p += sizeof(double) * N;
    r.cnt = (long long*)p; p += sizeof(long long) * N;
    r.m = (float*)p;
    return r;
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < nw; ++i) t += sh[i];   // fixed order -> deterministic
    return t;
}
__device__ __forceinline__ long long block_sum(long long v, long long* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    long long t = 0;
    for (int i = 0; i < nw; ++i) t += sh[i];
    return t;
}

// mask rule of modules.py:285-287.  `m_is_T`: first iteration, m still has the weight dtype
// (abs_w.mean() keeps dtype); afterwards m is fp32 and torch promotes the subtraction.
template <typename T>
__device__ __forceinline__ bool kmeans_mask(float a, float m, bool m_is_T) {
    float d = __fsub_rn(a, m);
    float thr = -m * 0.5f;
    if (m_is_T && sizeof(T) == 2) { d = to_f32(from_f32<T>(d)); thr = to_f32(from_f32<T>(thr)); }
    return d > thr;
}

// per-row variant: one CTA owns a row for all iterations.
template <typename T>
__global__ void __launch_bounds__(256)
ternary_init_row_kernel(const T* __restrict__ W, int64_t K, int iters,
                        float* __restrict__ latent, float* __restrict__ scale) {
    __shared__ double shd[8];
    __shared__ long long shl[8];
    const int64_t n = blockIdx.x;
    const T* wr = W + n * K;
    double s = 0.0;
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) s += (double)fabsf(to_f32(wr[k]));
    s = block_sum(s, shd);
    float m = to_f32(from_f32<T>((float)(s / (double)K)));       // mean keeps the dtype
    for (int it = 0; it < iters; ++it) {
        double ps = 0.0; long long pc = 0;
        for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
            float a = fabsf(to_f32(wr[k]));
            if (kmeans_mask<T>(a, m, it == 0)) { ps += (double)a; ++pc; }
        }
        ps = block_sum(ps, shd);
        pc = block_sum(pc, shl);
        m = __fdiv_rn((float)ps, (float)pc);                       // 0/0 -> NaN like the reference
    }
    const float inv = __fdiv_rn(1.0f, m);
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        float v = __fmul_rn(to_f32(wr[k]), inv);
        latent[n * K + k] = v != v ? v : fminf(fmaxf(v, -1.0f), 1.0f);
    }
    if (threadIdx.x == 0) scale[n] = m;
}

// tensor-wide ("uniform") variant: row partials + a single-CTA ordered finalise per iteration.
// phase 0: sum|w| ; phase 1: masked sum/count with m = *ws.m
template <typename T>
__global__ void __launch_bounds__(256)
ternary_partial_kernel(const T* __restrict__ W, int64_t K, void* wsp, int64_t N, int phase, int first) {
    __shared__ double shd[8];
    __shared__ long long shl[8];
    InitWs ws = carve(wsp, N);
    const int64_t n = blockIdx.x;
    const T* wr = W + n * K;
    const float m = phase ? *ws.m : 0.0f;
    double ps = 0.0; long long pc = 0;
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        float a = fabsf(to_f32(wr[k]));
        if (!phase || kmeans_mask<T>(a, m, first != 0)) { ps += (double)a; ++pc; }
    }
    ps = block_sum(ps, shd);
    pc = block_sum(pc, shl);
    if (threadIdx.x == 0) { ws.sum[n] = ps; ws.cnt[n] = pc; }
}

template <typename T>
__global__ void __launch_bounds__(256)
ternary_finalize_kernel(void* wsp, int64_t N, int phase) {
    __shared__ double shd[8];
    __shared__ long long shl[8];
    InitWs ws = carve(wsp, N);
    double s = 0.0; long long c = 0;
    for (int64_t i = threadIdx.x; i < N; i += blockDim.x) { s += ws.sum[i]; c += ws.cnt[i]; }
    s = block_sum(s, shd);
    c = block_sum(c, shl);
    if (threadIdx.x == 0) {
        if (!phase) *ws.m = to_f32(from_f32<T>((float)(s / (double)c)));
        else        *ws.m = __fdiv_rn((float)s, (float)c);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
ternary_latent_uniform_kernel(const T* __restrict__ W, int64_t N, int64_t K, const void* wsp,
                              float* __restrict__ latent, float* __restrict__ scale) {
    InitWs ws = carve(const_cast<void*>(wsp), N);
    const float m = *ws.m;
    const float inv = __fdiv_rn(1.0f, m);
    const int64_t total = N * K;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        float v = __fmul_rn(to_f32(W[i]), inv);
        latent[i] = v != v ? v : fminf(fmaxf(v, -1.0f), 1.0f);
        if (i < N) scale[i] = m;
    }
}

// binary / binary01 (modules.py:193-266): row partial sums of positive and negative parts
template <typename T>
__global__ void __launch_bounds__(256)
binary_partial_kernel(const T* __restrict__ W, int64_t K, void* wsp, int64_t N) {
    __shared__ double shd[8];
    InitWs ws = carve(wsp, N);
    const int64_t n = blockIdx.x;
    const T* wr = W + n * K;
    double pp = 0.0, pn = 0.0;
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        float v = to_f32(wr[k]);
        if (v > 0.0f) pp += (double)v;
        if (v < 0.0f) pn += (double)v;
    }
    pp = block_sum(pp, shd);
    pn = block_sum(pn, shd);
    if (threadIdx.x == 0) { ws.sum[n] = pp; ws.sum2[n] = pn; }
}

// per_row: scale[n] = (pos[n] - neg[n]) / K ; uniform: (sum pos - sum neg) / (N*K)
__global__ void __launch_bounds__(256)
binary_scale_kernel(void* wsp, int64_t N, int64_t K, int per_row, int positive_only, float* __restrict__ scale) {
    __shared__ double shd[8];
    InitWs ws = carve(wsp, N);
    if (per_row) {
        for (int64_t i = threadIdx.x; i < N; i += blockDim.x) {
            float mp = (float)ws.sum[i], mn = (float)ws.sum2[i];
            float num = positive_only ? mp : __fsub_rn(mp, mn);
            scale[i] = __fdiv_rn(num, (float)K);
        }
    } else {
        double sp = 0.0, sn = 0.0;
        for (int64_t i = threadIdx.x; i < N; i += blockDim.x) { sp += ws.sum[i]; sn += ws.sum2[i]; }
        sp = block_sum(sp, shd);
        sn = block_sum(sn, shd);
        float num = positive_only ? (float)sp : __fsub_rn((float)sp, (float)sn);
        float sc = __fdiv_rn(num, (float)(N * K));
        for (int64_t i = threadIdx.x; i < N; i += blockDim.x) scale[i] = sc;
    }
}

// latent = clip(W * prescale, +-1); natural: prescale = 1/scale[row] (fp32 promotion)
template <typename T>
__global__ void __launch_bounds__(256)
binary_latent_kernel(const T* __restrict__ W, int64_t N, int64_t K, float prescale, int natural,
                     const float* __restrict__ scale, float* __restrict__ latent) {
    const int64_t total = N * K;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        float w = to_f32(W[i]);
        float v;
        if (natural) v = __fmul_rn(w, __fdiv_rn(1.0f, scale[i / K]));
        else         v = to_f32(from_f32<T>(__fmul_rn(w, prescale)));   // python-float scale keeps dtype
        latent[i] = v != v ? v : fminf(fmaxf(v, -1.0f), 1.0f);
    }
}

}  // namespace tb200

using namespace tb200;

extern "C" size_t tb200_ternary_init_workspace(int64_t N) {
    if (N < 0) N = 0;
    return (size_t)N * (2 * sizeof(double) + sizeof(long long)) + 64;
}

extern "C" int tb200_ternary_init(const void* W, int dtype, int64_t N, int64_t K, int per_row, int iters,
                                  float* latent, float* scale, void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(W && latent && scale, "ternary_init: null pointer");
    TB_REQUIRE(N > 0 && K > 0, "ternary_init: bad shape N=%lld K=%lld", (long long)N, (long long)K);
    TB_REQUIRE(iters >= 0, "ternary_init: iters < 0");
    TB_REQUIRE(dtype_ok(dtype), "ternary_init: bad dtype %d", dtype);
    cudaStream_t s = (cudaStream_t)stream;
    if (per_row) {
        TB_DISPATCH_DTYPE(dtype, T,
            (ternary_init_row_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, iters, latent, scale)));
        TB_LAUNCH_CHECK();
        return TB200_OK;
    }
    TB_REQUIRE(workspace, "ternary_init: uniform mode needs a workspace");
    int64_t tot = N * K;
    int g = (int)((tot + 255) / 256 < (int64_t)sm_count() * 8 ? (tot + 255) / 256 : (int64_t)sm_count() * 8);
    TB_DISPATCH_DTYPE(dtype, T, {
        ternary_partial_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, workspace, N, 0, 0);
        TB_LAUNCH_CHECK();
        ternary_finalize_kernel<T><<<1, 256, 0, s>>>(workspace, N, 0);
        TB_LAUNCH_CHECK();
        for (int it = 0; it < iters; ++it) {
            ternary_partial_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, workspace, N, 1, it == 0);
            TB_LAUNCH_CHECK();
            ternary_finalize_kernel<T><<<1, 256, 0, s>>>(workspace, N, 1);
            TB_LAUNCH_CHECK();
        }
        ternary_latent_uniform_kernel<T><<<g, 256, 0, s>>>((const T*)W, N, K, workspace, latent, scale);
        TB_LAUNCH_CHECK();
    });
    return TB200_OK;
}

extern "C" int tb200_binary_init(const void* W, int dtype, int64_t N, int64_t K, int per_row, int positive_only,
                                 float prescale, int natural, float* latent, float* scale,
                                 void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(W && latent && scale && workspace, "binary_init: null pointer");
    TB_REQUIRE(N > 0 && K > 0, "binary_init: bad shape");
    TB_REQUIRE(dtype_ok(dtype), "binary_init: bad dtype %d", dtype);
    cudaStream_t s = (cudaStream_t)stream;
    int64_t tot = N * K;
    int g = (int)((tot + 255) / 256 < (int64_t)sm_count() * 8 ? (tot + 255) / 256 : (int64_t)sm_count() * 8);
    TB_DISPATCH_DTYPE(dtype, T, {
        binary_partial_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, workspace, N);
        TB_LAUNCH_CHECK();
        binary_scale_kernel<<<1, 256, 0, s>>>(workspace, N, K, per_row, positive_only, scale);
        TB_LAUNCH_CHECK();
        binary_latent_kernel<T><<<g, 256, 0, s>>>((const T*)W, N, K, prescale, natural, scale, latent);
        TB_LAUNCH_CHECK();
    });
    return TB200_OK;
}
