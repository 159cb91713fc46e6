// gemm_i8.cu — prefill-shaped ternary GEMM on the 5th-gen tensor cores (sm_100a).
//
//   out[m, n] = clip( (sum_k q_x[m,k] * w[n,k]) * s_x[m] * scale[n] + bias[n] )
//
// Orientation ("swap-AB"): the WEIGHT rows are the UMMA M dimension (128 rows per CTA tile,
// one TMEM lane each) and the TOKENS are the UMMA N dimension (BT columns).  Per k-block of
// 128:
//   * TMA brings the packed weight tile (128 rows x 32 bytes, 2 bits/weight) and the int8
//     activation tile (BT rows x 128 bytes, SWIZZLE_128B) into shared memory;
//   * expander warps turn every packed 32-bit word into one 16-byte chunk of u8 codes
//     {0,1,2} with 3 shifts + 4 ANDs and store it in the K-major SWIZZLE_128B layout the UMMA
//     descriptor expects (the activations were quantised in the matching K order, kperm=1);
//   * one thread issues 4 x tcgen05.mma.cta_group::1.kind::i8 (u8 x s8 -> s32, K=32 each)
//     into a TMEM accumulator; tcgen05.commit frees the stage;
//   * the epilogue warps read the accumulator with tcgen05.ld (double-buffered in TMEM so the
//     next tile's MMAs overlap), subtract sum_k q_x (codes are w+1), apply s_x*scale, bias,
//     clip and store bf16/f16/f32.
// Reference semantics: ops.ternary_matmul (impl/tf/moondream/ternary.py:279-281) /
// ScaledTernaryLinear.forward (binary/modules.py:101-109).
#include <stdlib.h>
#include "umma.cuh"

namespace tb200 {

constexpr int UM_ROWS = 128;          // weight rows per tile (UMMA M)
constexpr int UM_BK = 128;            // k per stage (bytes of int8 = one 128B swizzle row)
constexpr int UM_THREADS = 384;       // warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-7 epilogue, 8-11 expanders
constexpr int UM_EXP_WARPS = 4;
constexpr int UM_A_BYTES = UM_ROWS * UM_BK;        // 16 KB expanded weights
constexpr int UM_P_BYTES = UM_ROWS * (UM_BK / 4);  // 4 KB packed weights

template <int BT> struct UmmaCfg {
    static constexpr int B_BYTES = BT * UM_BK;
    static constexpr int STAGE_BYTES = UM_A_BYTES + B_BYTES + UM_P_BYTES;
    static constexpr int STAGES = (BT >= 256) ? 4 : 6;
    static constexpr int TMEM_COLS = (2 * BT < 32) ? 32 : 2 * BT;   // two accumulator buffers
    static constexpr int TAIL_BYTES = 1024;                         // barriers + tmem ptr
    static constexpr int SX_BYTES = 2 * BT * 8;                     // s_x and sum_qx, double-buffered
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + SX_BYTES + TAIL_BYTES + 1024 /*align slack*/;
};

template <typename OT, int BT>
__global__ void __launch_bounds__(UM_THREADS, 1)
umma_ternary_kernel(const __grid_constant__ CUtensorMap map_act, const __grid_constant__ CUtensorMap map_pw,
                    const UmmaParams p) {
    using Cfg = UmmaCfg<BT>;
    constexpr int NS = Cfg::STAGES;
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;      // SWIZZLE_128B atoms need 1024-B alignment
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    const uint32_t sA = base;                                          // [NS][16 KB] expanded weights
    const uint32_t sB = sA + NS * UM_A_BYTES;                          // [NS][BT*128] activations
    const uint32_t sP = sB + NS * Cfg::B_BYTES;                        // [NS][4 KB] packed weights
    const uint32_t sSX = sP + NS * UM_P_BYTES;                         // [2][BT] float s_x, [2][BT] int sum_qx
    const uint32_t sBar = sSX + Cfg::SX_BYTES;
    auto bar_full = [&](int s) { return sBar + 8u * s; };
    auto bar_aready = [&](int s) { return sBar + 8u * (NS + s); };
    auto bar_empty = [&](int s) { return sBar + 8u * (2 * NS + s); };
    auto bar_tfull = [&](int b) { return sBar + 8u * (3 * NS + b); };
    auto bar_tempty = [&](int b) { return sBar + 8u * (3 * NS + 2 + b); };
    const uint32_t sTmemPtr = sBar + 8u * (3 * NS + 4);
    float* sx_sm = reinterpret_cast<float*>(gbase + (sSX - base));
    int* sq_sm = reinterpret_cast<int*>(sx_sm + 2 * BT);
    volatile uint32_t* tmem_ptr_sm = reinterpret_cast<volatile uint32_t*>(gbase + (sTmemPtr - base));
    const uint8_t* gP = gbase + (sP - base);
    uint8_t* gA = gbase + (sA - base);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_tiles = p.m_tiles * p.n_tiles;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_act);
        tma_prefetch_desc(&map_pw);
        for (int s = 0; s < NS; ++s) {
            mbar_init(bar_full(s), 1);
            mbar_init(bar_aready(s), UM_EXP_WARPS);
            mbar_init(bar_empty(s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(bar_tfull(b), 1);
            mbar_init(bar_tempty(b), 4);
        }
        mbar_fence_init();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sTmemPtr), "n"(Cfg::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_sm;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                const int n_tile = t / p.m_tiles, m_tile = t - n_tile * p.m_tiles;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(bar_empty(s), ph ^ 1u);
                    mbar_arrive_expect_tx(bar_full(s), Cfg::B_BYTES + UM_P_BYTES);
                    tma_load_2d(sB + s * Cfg::B_BYTES, &map_act, bar_full(s), kb * UM_BK, m_tile * BT);
                    tma_load_2d(sP + s * UM_P_BYTES, &map_pw, bar_full(s), kb * (UM_BK / 4), n_tile * UM_ROWS);
                    if (++s == NS) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(UM_ROWS, BT);
            int s = 0; uint32_t ph = 0; int it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
                const int buf = it & 1;
                mbar_wait(bar_tempty(buf), (((uint32_t)it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BT);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(bar_full(s), ph);
                    mbar_wait(bar_aready(s), ph);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sA + s * UM_A_BYTES);
                    const uint64_t bdesc = umma_desc_sw128(sB + s * Cfg::B_BYTES);
#pragma unroll
                    for (int k = 0; k < UM_BK / 32; ++k)   // +32 bytes along K = +2 in the (addr>>4) field
                        umma_i8_ss(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) ? 1u : 0u);
                    umma_commit(bar_empty(s));
                    if (kb == p.k_blocks - 1) umma_commit(bar_tfull(buf));
                    if (++s == NS) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ===================== epilogue =====================
        const int q = warp & 3;                      // TMEM lane quarter this warp may access
        const int et = threadIdx.x - 128;            // 0..127
        OT* out = reinterpret_cast<OT*>(p.out);
        int it = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
            const int buf = it & 1;
            const int n_tile = t / p.m_tiles, m_tile = t - n_tile * p.m_tiles;
            const int64_t m0 = (int64_t)m_tile * BT;
            const int64_t n = (int64_t)n_tile * UM_ROWS + q * 32 + lane;
            // stage the per-token scale and code-offset for this tile
            for (int j = et; j < BT; j += 128) {
                const int64_t m = m0 + j;
                sx_sm[buf * BT + j] = m < p.M ? p.s_x[m] : 0.0f;
                sq_sm[buf * BT + j] = m < p.M ? p.sum_qx[m] : 0;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const float sc = (n < p.N && p.scale) ? p.scale[n] : 1.0f;
            const float bi = (n < p.N && p.bias) ? p.bias[n] : 0.0f;
            mbar_wait(bar_tfull(buf), ((uint32_t)it >> 1) & 1u);
            tc_fence_after();
            constexpr int CH = BT >= 32 ? 32 : 16;
#pragma unroll 1
            for (int c0 = 0; c0 < BT; c0 += CH) {
                uint32_t r[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0);
                if (CH == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r);
                if (n < p.N) {
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int64_t m = m0 + c0 + j;
                        if (m < p.M) {
                            const int a = (int)r[j] - sq_sm[buf * BT + c0 + j];
                            if (p.acc_out) p.acc_out[m * p.N + n] = a;
                            const float tv = __fmul_rn((float)a, sx_sm[buf * BT + c0 + j]);
                            out[m * p.ldo + n] = from_f32<OT>(epilogue(tv, sc, bi, p.clip_val));
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty(buf));
            asm volatile("bar.sync 1, 128;" ::: "memory");   // sx_sm[buf] may be refilled two tiles later
        }
    } else if (warp >= 8) {
        // ===================== expanders: packed 2-bit -> u8 codes in SWIZZLE_128B =====================
        const int xt = threadIdx.x - 256;            // 0..127
        const int c = xt & 7;                        // 16-byte chunk within the 128-byte row
        int s = 0; uint32_t ph = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                mbar_wait(bar_full(s), ph);
                const uint32_t* P = reinterpret_cast<const uint32_t*>(gP + s * UM_P_BYTES);
                uint8_t* A = gA + s * UM_A_BYTES;
                uint32_t w[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) w[i] = P[i * 128 + xt];            // row = i*16 + xt/8, word c
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = i * 16 + (xt >> 3);
                    uint4 o;
                    o.x = w[i] & 0x03030303u;
                    o.y = (w[i] >> 2) & 0x03030303u;
                    o.z = (w[i] >> 4) & 0x03030303u;
                    o.w = (w[i] >> 6) & 0x03030303u;
                    *reinterpret_cast<uint4*>(A + row * 128 + ((c ^ (row & 7)) << 4)) = o;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_aready(s));
                if (++s == NS) { s = 0; ph ^= 1u; }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(Cfg::TMEM_COLS) : "memory");
    }
}

// ---- host side ------------------------------------------------------------------------------
static int make_map_u8(CUtensorMap* map, const void* ptr, uint64_t inner, uint64_t outer, uint64_t pitch,
                       uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle swz) {
    uint64_t dims[2] = {inner, outer};
    uint64_t strides[1] = {pitch};
    uint32_t box[2] = {box_inner, box_outer};
    return make_tensor_map(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ptr, dims, strides, box, swz);
}

template <typename OT, int BT>
static int launch_umma(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed,
                       UmmaParams p, cudaStream_t s) {
    using Cfg = UmmaCfg<BT>;
    CUtensorMap map_act, map_pw;
    int rc = make_map_u8(&map_act, q_x, (uint64_t)ldq, (uint64_t)p.M, (uint64_t)ldq, UM_BK, BT, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    rc = make_map_u8(&map_pw, packed, (uint64_t)((p.K + 3) / 4), (uint64_t)p.N, (uint64_t)ld_packed, UM_BK / 4, UM_ROWS,
                     CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    auto kern = umma_ternary_kernel<OT, BT>;
    static bool configured = false;
    if (!configured) {
        TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        configured = true;
    }
    p.m_tiles = (int)((p.M + BT - 1) / BT);
    p.n_tiles = (int)((p.N + UM_ROWS - 1) / UM_ROWS);
    p.k_blocks = (int)((p.K + UM_BK - 1) / UM_BK);
    const int tiles = p.m_tiles * p.n_tiles;
    const int grid = tiles < sm_count() ? tiles : sm_count();
    kern<<<grid, UM_THREADS, Cfg::SMEM_BYTES, s>>>(map_act, map_pw, p);
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

template <typename OT>
static int umma_dispatch_bt(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed,
                            const UmmaParams& p, cudaStream_t s) {
    if (p.M <= 16) return launch_umma<OT, 16>(q_x, ldq, packed, ld_packed, p, s);
    if (p.M <= 32) return launch_umma<OT, 32>(q_x, ldq, packed, ld_packed, p, s);
    if (p.M <= 64) return launch_umma<OT, 64>(q_x, ldq, packed, ld_packed, p, s);
    if (p.M <= 128) return launch_umma<OT, 128>(q_x, ldq, packed, ld_packed, p, s);
    return launch_umma<OT, 256>(q_x, ldq, packed, ld_packed, p, s);
}

}  // namespace tb200

using namespace tb200;

extern "C" int tb200_gemm_i8(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx,
                             int64_t M, int64_t K,
                             const uint8_t* packed, int64_t ld_packed, int64_t N,
                             const float* scale, const float* bias, float clip_val,
                             void* out, int out_dtype, int64_t ldo,
                             int32_t* acc_out, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(q_x && s_x && sum_qx && packed && out, "gemm_i8: null pointer");
    TB_REQUIRE(M >= 0 && N >= 0 && K > 0 && ldo >= N, "gemm_i8: bad shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
    TB_REQUIRE(ldq >= K && ldq % 16 == 0, "gemm_i8: ldq=%lld must be >= K and a multiple of 16", (long long)ldq);
    TB_REQUIRE(ld_packed >= (K + 3) / 4 && ld_packed % 16 == 0, "gemm_i8: ld_packed=%lld must be >= ceil(K/4) and a multiple of 16", (long long)ld_packed);
    TB_REQUIRE((uintptr_t)q_x % 16 == 0 && (uintptr_t)packed % 16 == 0, "gemm_i8: q_x and packed must be 16-byte aligned");
    TB_REQUIRE(dtype_ok(out_dtype), "gemm_i8: bad out dtype %d", out_dtype);
    if (M == 0 || N == 0) return TB200_OK;
    UmmaParams p = {};
    p.s_x = s_x; p.sum_qx = sum_qx; p.scale = scale; p.bias = bias;
    p.out = out; p.ldo = ldo; p.acc_out = acc_out;
    p.M = M; p.N = N; p.K = K; p.clip_val = fabsf(clip_val);
    p.m_tiles = p.n_tiles = p.k_blocks = 0;
    p.layers = 1; p.stride_act_rows = 0; p.stride_sb = 0; p.stride_out = 0;
    { const char* e = getenv("TB200_GEMM_DBG"); p.dbg = e ? atoi(e) : 0; }
    cudaStream_t s = (cudaStream_t)stream;
    static int use_ts = -1;
    if (use_ts < 0) {
        const char* e = getenv("TB200_GEMM_PATH");      // "ss": expanded weights through shared memory; default "ts": through TMEM
        use_ts = (e && e[0] == 's' && e[1] == 's') ? 0 : 1;
    }
    if (use_ts) return gemm_ts_launch(q_x, ldq, packed, ld_packed, p, out_dtype, s);
    switch (out_dtype) {
        case TB200_F32:  return umma_dispatch_bt<float>(q_x, ldq, packed, ld_packed, p, s);
        case TB200_BF16: return umma_dispatch_bt<__nv_bfloat16>(q_x, ldq, packed, ld_packed, p, s);
        default:         return umma_dispatch_bt<__half>(q_x, ldq, packed, ld_packed, p, s);
    }
}
