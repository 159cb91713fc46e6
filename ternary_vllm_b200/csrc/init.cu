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

// ---- fast paths (rows of at most 4096 elements, 4-element vectors aligned): the row lives in REGISTERS ----
// A thread owns up to four 4-element vectors (element index 4 * (t + 256 j) + e: coalesced 16 / 8 byte loads, all issued
// before the first use).  Per-thread partial sums are fp32 over <= 16 values added as a balanced tree (relative error
// <= 4 * 2^-24, far inside the 2e-6 contract of DESIGN.md section 2); everything ACROSS threads, rows and CTAs stays fp64 in a fixed order.
template <typename T>
__device__ __forceinline__ void load_vec4(const T* p, float (&v)[4]) {
    if (sizeof(T) == 4) {
        const float4 f = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
        const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = to_f32(e[i]);
    }
}
__device__ __forceinline__ void block_sum2(double& s, long long& c, double* shd, long long* shl) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();
    if (l == 0) { shd[w] = s; shl[w] = c; }
    __syncthreads();
    double ts = 0.0; long long tc = 0;
    for (int i = 0; i < nw; ++i) { ts += shd[i]; tc += shl[i]; }   // fixed order -> deterministic
    s = ts; c = tc;
}
// masked partial sum / count over the thread's register-resident vectors (balanced tree per vector, then across vectors)
template <typename T>
__device__ __forceinline__ void masked_partial(const float (&w)[4][4], int nv, float m, bool masked, bool first, float& ps, int& pc) {
    float vs[4]; pc = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float a[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float av = fabsf(w[j][e]);
            const bool on = j < nv && (!masked || kmeans_mask<T>(av, m, first));
            a[e] = on ? av : 0.0f;
            pc += on ? 1 : 0;
        }
        vs[j] = (a[0] + a[1]) + (a[2] + a[3]);
    }
    ps = (vs[0] + vs[1]) + (vs[2] + vs[3]);
}

template <typename T>
__global__ void __launch_bounds__(256)
ternary_init_row_fast_kernel(const T* __restrict__ W, int64_t K, int iters, float* __restrict__ latent, float* __restrict__ scale) {
    __shared__ double shd[8];
    __shared__ long long shl[8];
    const int64_t n = blockIdx.x;
    const T* wr = W + n * K;
    const int nvec = (int)(K >> 2);
    float w[4][4]; int nv = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int v = threadIdx.x + 256 * j;
        if (v < nvec) { load_vec4<T>(wr + 4 * v, w[j]); nv = j + 1; }
        else { w[j][0] = w[j][1] = w[j][2] = w[j][3] = 0.0f; }
    }
    float ps; int pc;
    masked_partial<T>(w, nv, 0.0f, false, false, ps, pc);
    double s = (double)ps; long long c = pc;
    block_sum2(s, c, shd, shl);
    float m = to_f32(from_f32<T>((float)(s / (double)K)));       // mean keeps the dtype
    for (int it = 0; it < iters; ++it) {
        masked_partial<T>(w, nv, m, true, it == 0, ps, pc);
        s = (double)ps; c = pc;
        block_sum2(s, c, shd, shl);
        m = __fdiv_rn((float)s, (float)c);                         // 0/0 -> NaN like the reference
    }
    const float inv = __fdiv_rn(1.0f, m);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int v = threadIdx.x + 256 * j;
        if (v < nvec) {
            float4 o; float* oe = reinterpret_cast<float*>(&o);
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float x = __fmul_rn(w[j][e], inv); oe[e] = x != x ? x : fminf(fmaxf(x, -1.0f), 1.0f); }
            *reinterpret_cast<float4*>(latent + n * K + 4 * v) = o;
        }
    }
    if (threadIdx.x == 0) scale[n] = m;
}

// uniform variant, one pass: row partial (phase 0: sum|w|, phase 1: masked sum / count with m = *ws.m) and, in the LAST CTA
// to finish (ticket counter behind ws.m), the ordered finalise over all rows -- one launch per pass instead of two
template <typename T>
__global__ void __launch_bounds__(256)
ternary_pass_fast_kernel(const T* __restrict__ W, int64_t K, void* wsp, int64_t N, int phase, int first) {
    __shared__ double shd[8];
    __shared__ long long shl[8];
    __shared__ unsigned last;
    InitWs ws = carve(wsp, N);
    unsigned* ticket = reinterpret_cast<unsigned*>(ws.m + 1);
    const int64_t n = blockIdx.x;
    const T* wr = W + n * K;
    const int nvec = (int)(K >> 2);
    float w[4][4]; int nv = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int v = threadIdx.x + 256 * j;
        if (v < nvec) { load_vec4<T>(wr + 4 * v, w[j]); nv = j + 1; }
        else { w[j][0] = w[j][1] = w[j][2] = w[j][3] = 0.0f; }
    }
    const float m = phase ? *ws.m : 0.0f;
    float ps; int pc;
    masked_partial<T>(w, nv, m, phase != 0, first != 0, ps, pc);
    double s = (double)ps; long long c = pc;
    block_sum2(s, c, shd, shl);
    if (threadIdx.x == 0) {
        ws.sum[n] = s; ws.cnt[n] = c;
        __threadfence();
        last = atomicAdd(ticket, 1u) == (unsigned)(N - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    s = 0.0; c = 0;
    for (int64_t i = threadIdx.x; i < N; i += blockDim.x) { s += __ldcg(ws.sum + i); c += __ldcg(ws.cnt + i); }
    block_sum2(s, c, shd, shl);
    if (threadIdx.x == 0) {
        if (!phase) *ws.m = to_f32(from_f32<T>((float)(s / (double)c)));
        else        *ws.m = __fdiv_rn((float)s, (float)c);
        *ticket = 0u;
    }
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

// 4 elements per thread and step (aligned, numel % 4 == 0): 16 / 8 byte loads, 16-byte stores
template <typename T>
__global__ void __launch_bounds__(256)
ternary_latent_uniform_vec_kernel(const T* __restrict__ W, int64_t N, int64_t K, const void* wsp,
                                  float* __restrict__ latent, float* __restrict__ scale) {
    InitWs ws = carve(const_cast<void*>(wsp), N);
    const float m = *ws.m;
    const float inv = __fdiv_rn(1.0f, m);
    const int64_t nvec = (N * K) >> 2, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float v[4]; load_vec4<T>(W + 4 * i, v);
        float4 o; float* oe = reinterpret_cast<float*>(&o);
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float x = __fmul_rn(v[e], inv); oe[e] = x != x ? x : fminf(fmaxf(x, -1.0f), 1.0f); }
        *reinterpret_cast<float4*>(latent + 4 * i) = o;
        if (i < N) scale[i] = m;
    }
    for (int64_t i = nvec + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += stride) scale[i] = m;
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
    return (size_t)N * (2 * sizeof(double) + sizeof(long long)) + 64;   // + m (float) and the last-CTA ticket behind it
}

extern "C" int tb200_ternary_init(const void* W, int dtype, int64_t N, int64_t K, int per_row, int iters,
                                  float* latent, float* scale, void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(W && latent && scale, "ternary_init: null pointer");
    TB_REQUIRE(N > 0 && K > 0, "ternary_init: bad shape N=%lld K=%lld", (long long)N, (long long)K);
    TB_REQUIRE(iters >= 0, "ternary_init: iters < 0");
    TB_REQUIRE(dtype_ok(dtype), "ternary_init: bad dtype %d", dtype);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t es = dtype_size(dtype);
    // register-resident fast paths: rows of <= 4096 elements, whole 4-element vectors, 16 / 8 byte aligned rows
    const bool fast = K <= 4096 && K % 4 == 0 && (uintptr_t)W % (4 * es) == 0 && (uintptr_t)latent % 16 == 0;
    if (per_row) {
        TB_DISPATCH_DTYPE(dtype, T, {
            if (fast) ternary_init_row_fast_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, iters, latent, scale);
            else ternary_init_row_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, iters, latent, scale);
        });
        TB_LAUNCH_CHECK();
        return TB200_OK;
    }
    TB_REQUIRE(workspace, "ternary_init: uniform mode needs a workspace");
    int64_t tot = N * K;
    int g = (int)((tot + 255) / 256 < (int64_t)sm_count() * 8 ? (tot + 255) / 256 : (int64_t)sm_count() * 8);
    if (fast) {
        InitWs ws = carve(workspace, N);
        TB_CUDA(cudaMemsetAsync(ws.m + 1, 0, sizeof(unsigned), s));          // the last-CTA ticket
        TB_DISPATCH_DTYPE(dtype, T, {
            ternary_pass_fast_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, workspace, N, 0, 0);
            TB_LAUNCH_CHECK();
            for (int it = 0; it < iters; ++it) {
                ternary_pass_fast_kernel<T><<<(unsigned)N, 256, 0, s>>>((const T*)W, K, workspace, N, 1, it == 0);
                TB_LAUNCH_CHECK();
            }
            const int64_t nvec = tot >> 2;
            const int gv = (int)((nvec + 255) / 256 < (int64_t)sm_count() * 8 ? (nvec + 255) / 256 : (int64_t)sm_count() * 8);
            ternary_latent_uniform_vec_kernel<T><<<gv, 256, 0, s>>>((const T*)W, N, K, workspace, latent, scale);
            TB_LAUNCH_CHECK();
        });
        return TB200_OK;
    }
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
