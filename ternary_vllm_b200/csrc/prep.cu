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
#define LAUNCH_QP(T, MODE) \
    quantize_pack_kernel<T, MODE><<<grid, block, 0, s>>>((const T*)w, K, ldw, packed, ld_packed, codes_out, vec_ok)
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
    TB_DISPATCH_DTYPE(dtype, T,
        (clip_kernel<T><<<grid_for(numel, 256), 256, 0, (cudaStream_t)stream>>>((T*)w, numel)));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
