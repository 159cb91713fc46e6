// train.cu — optimiser-side helpers of the smooth-binary training loop on the GPU:
// compute_and_set_biobjective_gradient (binary/modules.py:149-176) as two kinds of launches per tensor instead of ~8 eager
// passes per parameter and two host synchronisations (`if norm_g > 0.0` on device scalars):
//   tb200_sumsq_accumulate  : acc += sum x^2   (task gradients)   or   acc += #nonzero(x)  (= sum sign(x)^2, the
//                             binarisation gradient -sign(w)); per-CTA fp64 partials, added in a fixed order by the
//                             last CTA to finish -> deterministic;
//   tb200_biobjective_apply : g <- the blended gradient, in place, reading both squared norms from device memory.
// HBM-bound elementwise / reduction work: 16-byte loads, four in flight per thread.
#include "common.cuh"

namespace tb200 {

constexpr int TR_THREADS = 256;
constexpr int TR_MAX_CTAS = 1024;

// workspace: double partial[TR_MAX_CTAS] | unsigned ticket (zero before the first call; the kernel resets it)
template <typename T, bool COUNT_NONZERO>
__global__ void __launch_bounds__(TR_THREADS)
sumsq_kernel(const T* __restrict__ x, int64_t numel, double* __restrict__ acc, double* __restrict__ partial, unsigned* __restrict__ ticket) {
    constexpr int EPV = 16 / (int)sizeof(T);
    __shared__ double sh[TR_THREADS / 32];
    __shared__ unsigned last;
    const int64_t nvec = ((uintptr_t)x % 16 == 0) ? numel / EPV : 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    double s = 0.0;
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    for (int64_t i0 = t0; i0 < nvec; i0 += 4 * stride) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { const int64_t i = i0 + j * stride; v[j] = i < nvec ? __ldg(xv + i) : make_uint4(0u, 0u, 0u, 0u); }
        float p = 0.0f;                                  // <= 32 squares per step in fp32, then fp64
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const T* e = reinterpret_cast<const T*>(&v[j]);
#pragma unroll
            for (int k = 0; k < EPV; ++k) {
                const float f = to_f32(e[k]);
                p += COUNT_NONZERO ? ((f > 0.0f || f < 0.0f) ? 1.0f : 0.0f) : f * f;     // sign(NaN)^2 is NaN in torch: not reproduced (weights are finite)
            }
        }
        s += (double)p;
    }
    for (int64_t i = nvec * EPV + t0; i < numel; i += stride) {
        const float f = to_f32(x[i]);
        s += COUNT_NONZERO ? ((f > 0.0f || f < 0.0f) ? 1.0 : 0.0) : (double)f * (double)f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < TR_THREADS / 32; ++i) t += sh[i];
        partial[blockIdx.x] = t;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last || threadIdx.x != 0) return;
    __threadfence();
    double t = 0.0;
    for (unsigned i = 0; i < gridDim.x; ++i) t += __ldcg(partial + i);      // fixed order
    *acc += t;
    *ticket = 0u;
}

// mode 0 (smooth-binary latent weights, w != null):  g <- ((g / n_g + (-sign(w)) / n_bg) * (n_g + n_bg) * 0.5) / n_g * (n_g + n_bg)
//   -- the reference first stores the blend into weights.grad and then rescales EVERY parameter gradient of the module,
//   the blended one included (modules.py:166-176); n_g == 0: the first quotient is g * 0 and the rescale is skipped;
// mode 1 (every other parameter):                    g <- (g / n_g) * (n_g + n_bg), untouched when n_g == 0.
// fp32 arithmetic in the reference's operation order (norms: fp64 accumulators rounded to fp32 once); 16-bit gradients are
// computed in fp32 and rounded once on the store.
template <typename T>
__global__ void __launch_bounds__(TR_THREADS)
biobjective_apply_kernel(T* __restrict__ g, const T* __restrict__ w, int64_t numel, const double* __restrict__ ng_sq,
                         const double* __restrict__ nbg_sq) {
    const float n_g = sqrtf((float)*ng_sq), n_bg = sqrtf((float)*nbg_sq);
    const float tot = __fadd_rn(n_g, n_bg);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < numel; i += stride) {
        float v = to_f32(g[i]);
        if (w) {
            const float wv = to_f32(w[i]);
            const float bg = wv > 0.0f ? -1.0f : (wv < 0.0f ? 1.0f : (wv == wv ? 0.0f : wv));   // -sign(w); NaN stays NaN
            const float a = n_g > 0.0f ? __fdiv_rn(v, n_g) : __fmul_rn(v, 0.0f);
            const float c = n_bg > 0.0f ? __fdiv_rn(bg, n_bg) : __fmul_rn(bg, 0.0f);
            v = __fmul_rn(__fmul_rn(__fadd_rn(a, c), tot), 0.5f);
        }
        if (n_g > 0.0f) v = __fmul_rn(__fdiv_rn(v, n_g), tot);
        g[i] = from_f32<T>(v);
    }
}

}  // namespace tb200

using namespace tb200;

extern "C" size_t tb200_train_workspace(void) { return sizeof(double) * TR_MAX_CTAS + 64; }

extern "C" int tb200_sumsq_accumulate(const void* x, int dtype, int64_t numel, int count_nonzero, double* acc,
                                      void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(acc && workspace && (x || numel == 0), "sumsq_accumulate: null pointer");
    TB_REQUIRE(dtype_ok(dtype), "sumsq_accumulate: bad dtype %d", dtype);
    TB_REQUIRE((uintptr_t)workspace % 8 == 0, "sumsq_accumulate: workspace must be 8-byte aligned");
    if (numel <= 0) return TB200_OK;
    int64_t ctas = (numel + TR_THREADS * 32 - 1) / (TR_THREADS * 32);
    const int64_t cap = (int64_t)sm_count() * 4 < TR_MAX_CTAS ? (int64_t)sm_count() * 4 : TR_MAX_CTAS;
    if (ctas > cap) ctas = cap;
    double* partial = (double*)workspace;
    unsigned* ticket = (unsigned*)(partial + TR_MAX_CTAS);
    cudaStream_t s = (cudaStream_t)stream;
    TB_DISPATCH_DTYPE(dtype, T, {
        if (count_nonzero) sumsq_kernel<T, true><<<(unsigned)ctas, TR_THREADS, 0, s>>>((const T*)x, numel, acc, partial, ticket);
        else sumsq_kernel<T, false><<<(unsigned)ctas, TR_THREADS, 0, s>>>((const T*)x, numel, acc, partial, ticket);
    });
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_biobjective_apply(void* grad, const void* weights, int dtype, int64_t numel, const double* norm_g_sq,
                                       const double* norm_bg_sq, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(norm_g_sq && norm_bg_sq && (grad || numel == 0), "biobjective_apply: null pointer");
    TB_REQUIRE(dtype_ok(dtype), "biobjective_apply: bad dtype %d", dtype);
    if (numel <= 0) return TB200_OK;
    int64_t ctas = (numel + TR_THREADS * 8 - 1) / (TR_THREADS * 8);
    const int64_t cap = (int64_t)sm_count() * 8;
    if (ctas > cap) ctas = cap;
    TB_DISPATCH_DTYPE(dtype, T,
        (biobjective_apply_kernel<T><<<(unsigned)ctas, TR_THREADS, 0, (cudaStream_t)stream>>>((T*)grad, (const T*)weights, numel, norm_g_sq, norm_bg_sq)));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
