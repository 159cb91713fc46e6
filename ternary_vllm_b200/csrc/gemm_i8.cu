// gemm_i8.cu — C-ABI entry of the prefill-shaped GEMM (tb200_gemm_i8): argument checks, then the tcgen05 kernel of
// gemm_ts.cu (weights expanded into TENSOR MEMORY as the UMMA A operand).  The first design, which wrote the expanded
// weights to shared memory and was bound by its bandwidth at 1.5 POPS (profiles/r1_gemm_cta1.*), has been removed.
#include "umma.cuh"

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
    p.dbg = (int)tb_knob("TB200_GEMM_DBG", 0);
    return gemm_ts_launch(q_x, ldq, packed, ld_packed, p, out_dtype, (cudaStream_t)stream);
}
