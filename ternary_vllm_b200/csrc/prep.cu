// prep.cu — weight preparation: quantise+pack, unpack, reference pack formats, clip.
// HBM-bound byte work: one pass over the latent weights, coalesced loads, no reshaping.
#include "common.cuh"

namespace tb200 {

// ---------------------------------------------------------------------------
// quantize_pack: one CTA per weight row.  Thread t produces packed byte b (4 weights).
// A warp reads 32 consecutive 4-element groups per step = 512 B (fp32) fully coalesced
// and writes 32 consecutive bytes.
// ---------------------------------------------------------------------------
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
quantize_pack_kernel(const T* __restrict__ w, int64_t K, int64_t ldw,
                     uint8_t* __restrict__ packed, int64_t ld_packed,
                     int8_t* __restrict__ codes_out, bool vec_ok) {
    const int64_t n = blockIdx.x;
    const T* wr = w + n * ldw;
    uint8_t* pr = packed + n * ld_packed;
    int8_t* cr = codes_out ? codes_out + n * K : nullptr;
    const int64_t kb_full = K >> 2;            // bytes made of 4 in-range weights
    for (int64_t b = threadIdx.x; b < ld_packed; b += blockDim.x) {
        uint32_t byte;
        if (b < kb_full) {
            float v0, v1, v2, v3;
            if (vec_ok) {
                if (sizeof(T) == 4) {
                    float4 f = *reinterpret_cast<const float4*>(wr + 4 * b);
                    v0 = f.x; v1 = f.y; v2 = f.z; v3 = f.w;
                } else {
                    uint2 u = *reinterpret_cast<const uint2*>(wr + 4 * b);
                    const T* e = reinterpret_cast<const T*>(&u);
                    v0 = to_f32(e[0]); v1 = to_f32(e[1]); v2 = to_f32(e[2]); v3 = to_f32(e[3]);
                }
            } else {
                v0 = to_f32(wr[4 * b]); v1 = to_f32(wr[4 * b + 1]);
                v2 = to_f32(wr[4 * b + 2]); v3 = to_f32(wr[4 * b + 3]);
            }
            uint32_t c0 = quant_code<MODE>(v0), c1 = quant_code<MODE>(v1);
            uint32_t c2 = quant_code<MODE>(v2), c3 = quant_code<MODE>(v3);
            byte = c0 | (c1 << 2) | (c2 << 4) | (c3 << 6);
            if (cr) {
                cr[4 * b] = (int8_t)c0 - 1; cr[4 * b + 1] = (int8_t)c1 - 1;
                cr[4 * b + 2] = (int8_t)c2 - 1; cr[4 * b + 3] = (int8_t)c3 - 1;
            }
        } else {
            byte = 0x55u;  // padding code 1 (value 0), ternary.py:380-386
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t k = 4 * b + i;
                if (k < K) {
                    uint32_t c = quant_code<MODE>(to_f32(wr[k]));
                    byte = (byte & ~(3u << (2 * i))) | (c << (2 * i));
                    if (cr) cr[k] = (int8_t)c - 1;
                }
            }
        }
        pr[b] = (uint8_t)byte;
    }
}

// Fast path (rows of whole 16-byte units, aligned): a thread turns 16 input bytes (4 fp32 weights -> 1 packed byte,
// 8 bf16 / f16 weights -> 2 packed bytes) per step and works on FOUR steps at once: the four 16-byte loads of a thread are
// issued before the first use (the single-load loop above leaves the memory system with one request per thread in flight:
// 61 % of the measured HBM rate for fp32 rows, 39 % for 16-bit rows).  Consecutive lanes read consecutive 16-byte units.
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
quantize_pack_vec_kernel(const T* __restrict__ w, int64_t K, int64_t ldw, uint8_t* __restrict__ packed, int64_t ld_packed) {
    constexpr int EPU = 16 / (int)sizeof(T);           // elements per unit: 4 or 8
    constexpr int BPU = EPU / 4;                        // packed bytes per unit: 1 or 2
    const int64_t n = blockIdx.x;
    const uint4* wr = reinterpret_cast<const uint4*>(w + n * ldw);
    uint8_t* pr = packed + n * ld_packed;
    const int units = (int)(K / EPU);
    for (int u0 = threadIdx.x; u0 < units; u0 += 4 * 256) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int u = u0 + 256 * j;
            v[j] = u < units ? __ldg(wr + u) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int u = u0 + 256 * j;
            if (u >= units) continue;
            const T* e = reinterpret_cast<const T*>(&v[j]);
            uint32_t word = 0;
#pragma unroll
            for (int i = 0; i < EPU; ++i) word |= quant_code<MODE>(to_f32(e[i])) << (2 * i);
            if (BPU == 1) pr[u] = (uint8_t)word;
            else *reinterpret_cast<uint16_t*>(pr + 2 * u) = (uint16_t)word;
        }
    }
    // row padding up to the pitch: code 1 (value 0)
    for (int64_t b = (K >> 2) + threadIdx.x; b < ld_packed; b += 256) pr[b] = 0x55u;
}

// unpack format C -> int8 values (code - 1); thread per packed byte
__global__ void __launch_bounds__(256)
unpack_kernel(const uint8_t* __restrict__ packed, int64_t ld_packed, int64_t N, int64_t K,
              int8_t* __restrict__ out) {
    const int64_t kb = (K + 3) >> 2;
    const int64_t total = N * kb;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t n = i / kb, b = i - n * kb;
        uint32_t byte = packed[n * ld_packed + b];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int64_t k = 4 * b + j;
            if (k < K) out[n * K + k] = (int8_t)((byte >> (2 * j)) & 3u) - 1;
        }
    }
}

// reference layouts A / B (planar over the flattened, zero-padded vector) and C (per row)
__device__ __forceinline__ uint32_t code_of(int v, int fmt) {
    if (fmt == TB200_FMT_A) return (v == 1 ? 1u : 0u) ^ (v == -1 ? 2u : 0u);   // ternary.py:72-74
    return (uint32_t)(v + 1) & 3u;                                             // ternary.py:128, :392
}
__device__ __forceinline__ int value_of(uint32_t f, int fmt) {
    if (fmt == TB200_FMT_A) return (int)(f & 1u) - (int)((f >> 1) & 1u);       // ternary.py:104-108
    return (int)f - 1;                                                         // ternary.py:150-153
}

__global__ void __launch_bounds__(256)
pack_format_kernel(const int8_t* __restrict__ codes, int64_t N, int64_t K, int fmt,
                   uint8_t* __restrict__ out) {
    if (fmt == TB200_FMT_C) {
        const int64_t kb = (K + 3) >> 2, total = N * kb;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
             i += (int64_t)gridDim.x * blockDim.x) {
            int64_t n = i / kb, b = i - n * kb;
            uint32_t byte = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int64_t k = 4 * b + j;
                int v = k < K ? codes[n * K + k] : 0;
                byte |= code_of(v, fmt) << (2 * j);
            }
            out[i] = (uint8_t)byte;
        }
    } else {
        const int64_t T = N * K, S = (T + 3) >> 2;
        for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < S;
             j += (int64_t)gridDim.x * blockDim.x) {
            uint32_t byte = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t t = i * S + j;
                int v = t < T ? codes[t] : 0;
                byte |= code_of(v, fmt) << (2 * i);
            }
            out[j] = (uint8_t)byte;
        }
    }
}

__global__ void __launch_bounds__(256)
unpack_format_kernel(const uint8_t* __restrict__ packed, int64_t N, int64_t K, int fmt,
                     int8_t* __restrict__ out) {
    if (fmt == TB200_FMT_C) {
        const int64_t kb = (K + 3) >> 2, total = N * kb;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
             i += (int64_t)gridDim.x * blockDim.x) {
            int64_t n = i / kb, b = i - n * kb;
            uint32_t byte = packed[i];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int64_t k = 4 * b + j;
                if (k < K) out[n * K + k] = (int8_t)value_of((byte >> (2 * j)) & 3u, fmt);
            }
        }
    } else {
        const int64_t T = N * K, S = (T + 3) >> 2;
        for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < S;
             j += (int64_t)gridDim.x * blockDim.x) {
            uint32_t byte = packed[j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                out[i * S + j] = (int8_t)value_of((byte >> (2 * i)) & 3u, fmt);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
clip_kernel(T* __restrict__ w, int64_t numel) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        float v = to_f32(w[i]);
        // torch.clip propagates NaN; fminf/fmaxf would not
        float c = v != v ? v : fminf(fmaxf(v, -1.0f), 1.0f);
        w[i] = from_f32<T>(c);
    }
}

// 16 bytes per thread and step, four steps in flight (aligned buffers; the tail goes through clip_kernel)
template <typename T>
__global__ void __launch_bounds__(256)
clip_vec_kernel(T* __restrict__ w, int64_t nvec) {
    constexpr int EPV = 16 / (int)sizeof(T);
    uint4* p = reinterpret_cast<uint4*>(w);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < nvec; i0 += 4 * stride) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { const int64_t i = i0 + j * stride; if (i < nvec) v[j] = p[i]; }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t i = i0 + j * stride;
            if (i >= nvec) continue;
            T* e = reinterpret_cast<T*>(&v[j]);
#pragma unroll
            for (int k = 0; k < EPV; ++k) {
                const float x = to_f32(e[k]);
                e[k] = from_f32<T>(x != x ? x : fminf(fmaxf(x, -1.0f), 1.0f));   // torch.clip propagates NaN
            }
            p[i] = v[j];
        }
    }
}

static inline int grid_for(int64_t total, int threads) {
    int64_t g = (total + threads - 1) / threads;
    int64_t cap = (int64_t)sm_count() * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace tb200

using namespace tb200;

extern "C" int tb200_quantize_pack(const void* w, int dtype, int64_t N, int64_t K, int64_t ldw,
                                   int mode, uint8_t* packed, int64_t ld_packed,
                                   int8_t* codes_out, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (N == 0) return TB200_OK;
    TB_REQUIRE(w && packed, "quantize_pack: null pointer");
    TB_REQUIRE(N >= 0 && K >= 0 && ldw >= K, "quantize_pack: bad shape N=%lld K=%lld ldw=%lld",
               (long long)N, (long long)K, (long long)ldw);
    TB_REQUIRE(ld_packed >= (K + 3) / 4, "quantize_pack: ld_packed %lld < ceil(K/4)", (long long)ld_packed);
    TB_REQUIRE(mode >= 0 && mode <= 2, "quantize_pack: bad mode %d", mode);
    TB_REQUIRE(dtype_ok(dtype), "quantize_pack: bad dtype %d", dtype);
    if (N == 0 || ld_packed == 0) return TB200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t es = dtype_size(dtype);
    // 4-element vector loads (16 B fp32 / 8 B bf16,fp16) need base and row pitch aligned to them
    const bool vec_ok = ((uintptr_t)w % (4 * es) == 0) && ((ldw * es) % (4 * es) == 0);
    dim3 grid((unsigned)N), block(256);
    // whole 16-byte units per row, 16-byte aligned rows, 2-byte aligned packed rows, no code dump: the vector kernel
    const bool fast = !codes_out && K > 0 && (K * es) % 16 == 0 && (uintptr_t)w % 16 == 0 && (ldw * es) % 16 == 0 &&
                      (uintptr_t)packed % 2 == 0 && ld_packed % 2 == 0;
#define LAUNCH_QP(T, MODE) \
    do { if (fast) quantize_pack_vec_kernel<T, MODE><<<grid, block, 0, s>>>((const T*)w, K, ldw, packed, ld_packed); \
         else quantize_pack_kernel<T, MODE><<<grid, block, 0, s>>>((const T*)w, K, ldw, packed, ld_packed, codes_out, vec_ok); } while (0)
    TB_DISPATCH_DTYPE(dtype, T,
        if (mode == TB200_Q_TERNARY) LAUNCH_QP(T, TB200_Q_TERNARY);
        else if (mode == TB200_Q_SIGN) LAUNCH_QP(T, TB200_Q_SIGN);
        else LAUNCH_QP(T, TB200_Q_STEP));
#undef LAUNCH_QP
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_unpack(const uint8_t* packed, int64_t ld_packed, int64_t N, int64_t K,
                            int8_t* out, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(packed && out, "unpack: null pointer");
    TB_REQUIRE(N >= 0 && K >= 0 && ld_packed >= (K + 3) / 4, "unpack: bad shape");
    if (N == 0 || K == 0) return TB200_OK;
    unpack_kernel<<<grid_for(N * ((K + 3) / 4), 256), 256, 0, (cudaStream_t)stream>>>(packed, ld_packed, N, K, out);
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_pack_format(const int8_t* codes, int64_t N, int64_t K, int fmt,
                                 uint8_t* out, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(codes && out, "pack_format: null pointer");
    TB_REQUIRE(fmt >= 0 && fmt <= 2, "pack_format: bad format %d", fmt);
    TB_REQUIRE(N >= 0 && K >= 0, "pack_format: bad shape");
    if (N * K == 0) return TB200_OK;
    int64_t total = fmt == TB200_FMT_C ? N * ((K + 3) / 4) : (N * K + 3) / 4;
    pack_format_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(codes, N, K, fmt, out);
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_unpack_format(const uint8_t* packed, int64_t N, int64_t K, int fmt,
                                   int8_t* out, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(packed && out, "unpack_format: null pointer");
    TB_REQUIRE(fmt >= 0 && fmt <= 2, "unpack_format: bad format %d", fmt);
    TB_REQUIRE(N >= 0 && K >= 0, "unpack_format: bad shape");
    if (N * K == 0) return TB200_OK;
    int64_t total = fmt == TB200_FMT_C ? N * ((K + 3) / 4) : (N * K + 3) / 4;
    unpack_format_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(packed, N, K, fmt, out);
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

extern "C" int tb200_clip_latent(void* w, int dtype, int64_t numel, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(w || numel == 0, "clip_latent: null pointer");
    TB_REQUIRE(dtype_ok(dtype), "clip_latent: bad dtype %d", dtype);
    if (numel <= 0) return TB200_OK;
    const size_t es = dtype_size(dtype);
    const int64_t epv = 16 / (int64_t)es;
    const int64_t nvec = ((uintptr_t)w % 16 == 0) ? numel / epv : 0;
    TB_DISPATCH_DTYPE(dtype, T, {
        if (nvec > 0) clip_vec_kernel<T><<<grid_for((nvec + 3) / 4, 256), 256, 0, (cudaStream_t)stream>>>((T*)w, nvec);
        if (numel - nvec * epv > 0)
            clip_kernel<T><<<grid_for(numel - nvec * epv, 256), 256, 0, (cudaStream_t)stream>>>((T*)w + nvec * epv, numel - nvec * epv);
    });
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
