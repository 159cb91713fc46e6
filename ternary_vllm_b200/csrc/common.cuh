// common.cuh — shared helpers for libternary_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <atomic>

#include "../../include/ternary_b200.h"

namespace tb200 {

// ---- error reporting (thread-local message, see tb200_last_error) ----------
void set_error(const char* fmt, ...);
// Device state is kept PER DEVICE ORDINAL (a process may drive several GPUs from several host threads):
int  current_device();               // cudaGetDevice, -1 on failure
int  check_device();                 // TB200_OK or TB200_ENODEV for the calling thread's current device
int  sm_count();                     // SM count of the calling thread's current device
extern std::atomic<uint64_t> g_launches;
constexpr int TB_MAX_DEVICES = 64;

// Tuning / ablation knobs.  The release library does not read the environment and carries no ablation branches in its
// kernels: `make ABLATE=1` (-DTB200_ABLATE) builds the experiment library that honours the TB200_* variables quoted in
// DESIGN.md / profiles/README.md.
#ifdef TB200_ABLATE
constexpr bool kAblate = true;
static inline long long tb_knob(const char* name, long long dflt) { const char* e = getenv(name); return e ? atoll(e) : dflt; }
#else
constexpr bool kAblate = false;
static inline long long tb_knob(const char*, long long dflt) { return dflt; }
#endif

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) attribute: remember what was requested
// for every device ordinal separately (a process-wide flag would skip the opt-in on the second GPU).  Races between host
// threads are benign (the attribute is set again with the same or a larger value).
struct SmemOptIn { std::atomic<size_t> bytes[TB_MAX_DEVICES]; };
template <typename Kern>
static inline cudaError_t ensure_dynamic_smem(Kern kern, SmemOptIn& st, size_t smem) {
    const int dev = current_device();
    if (dev >= 0 && dev < TB_MAX_DEVICES && smem <= st.bytes[dev].load(std::memory_order_acquire)) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess && dev >= 0 && dev < TB_MAX_DEVICES) st.bytes[dev].store(smem, std::memory_order_release);
    return e;
}

#define TB_REQUIRE(cond, ...)                                   \
    do {                                                        \
        if (!(cond)) {                                          \
            ::tb200::set_error(__VA_ARGS__);                    \
            return TB200_EINVAL;                                \
        }                                                       \
    } while (0)

#define TB_CUDA(expr)                                                            \
    do {                                                                         \
        cudaError_t _e = (expr);                                                 \
        if (_e != cudaSuccess) {                                                 \
            ::tb200::set_error("%s failed: %s (%s:%d)", #expr,                   \
                               cudaGetErrorString(_e), __FILE__, __LINE__);      \
            return TB200_ECUDA;                                                  \
        }                                                                        \
    } while (0)

#define TB_LAUNCH_CHECK()                                                        \
    do {                                                                         \
        ::tb200::g_launches.fetch_add(1, std::memory_order_relaxed);             \
        cudaError_t _e = cudaPeekAtLastError();                                  \
        if (_e != cudaSuccess) {                                                 \
            ::tb200::set_error("kernel launch failed: %s (%s:%d)",               \
                               cudaGetErrorString(_e), __FILE__, __LINE__);      \
            return TB200_ECUDA;                                                  \
        }                                                                        \
    } while (0)

#define TB_DEVICE_OR_RETURN()                                   \
    do {                                                        \
        int _d = ::tb200::check_device();                       \
        if (_d != TB200_OK) return _d;                          \
    } while (0)

// ---- dtype helpers ----------------------------------------------------------
template <int DT> struct elem;
template <> struct elem<TB200_F32>  { using type = float; };
template <> struct elem<TB200_BF16> { using type = __nv_bfloat16; };
template <> struct elem<TB200_F16>  { using type = __half; };

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

static inline size_t dtype_size(int dt) { return dt == TB200_F32 ? 4 : 2; }
static inline bool   dtype_ok(int dt) { return dt == TB200_F32 || dt == TB200_BF16 || dt == TB200_F16; }

// dispatch a lambda-like macro over the three float dtypes
#define TB_DISPATCH_DTYPE(dt, T, ...)                                           \
    switch (dt) {                                                               \
        case TB200_F32:  { using T = float;          __VA_ARGS__; break; }      \
        case TB200_BF16: { using T = __nv_bfloat16;  __VA_ARGS__; break; }      \
        case TB200_F16:  { using T = __half;         __VA_ARGS__; break; }      \
        default: ::tb200::set_error("bad dtype %d", dt); return TB200_EINVAL;   \
    }

// quantiser decision rules -> code = value + 1 in {0,1,2}
// (binary/modules.py:5-26; comparisons are done on the stored value, exactly as torch does)
template <int MODE>
__device__ __forceinline__ uint32_t quant_code(float w) {
    if (MODE == TB200_Q_TERNARY) return 1u + (w >= 0.5f ? 1u : 0u) - (w <= -0.5f ? 1u : 0u);
    if (MODE == TB200_Q_SIGN)    return (w >= 0.0f) ? 2u : 0u;   // -0.0 >= 0 is true; NaN -> -1 like torch
    return (w >= 0.0f) ? 2u : 1u;                                // step: {0,+1}
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// fused epilogue shared by GEMV and GEMM: the reference order is
//   y = acc (in x.dtype) ; y *= scale ; y += bias ; clip   (binary/modules.py:104-108)
__device__ __forceinline__ float epilogue(float acc, float scale, float bias, float clip_val) {
    float y = fmaf(acc, scale, bias);
    return fminf(fmaxf(y, -clip_val), clip_val);
}

}  // namespace tb200
