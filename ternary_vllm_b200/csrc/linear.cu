// linear.cu — whole-layer entry points: shape dispatch (GEMV vs tcgen05 GEMM) and the
// host-buffer end-to-end call.  Mirrors ScaledTernaryLinear.forward (binary/modules.py:101-109).
#include "common.cuh"

namespace tb200 {
static inline int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }
struct LinearWs { int8_t* q_x; int64_t ldq; float* s_x; int32_t* sum_qx; };
static LinearWs carve_ws(void* ws, int64_t M, int64_t K) {
    LinearWs r;
    r.ldq = round_up(K, 16);
    char* p = (char*)ws;
    r.q_x = (int8_t*)p;        p += round_up(M * r.ldq, 256);
    r.s_x = (float*)p;         p += round_up(M * 4, 256);
    r.sum_qx = (int32_t*)p;
    return r;
}
}  // namespace tb200
using namespace tb200;

extern "C" size_t tb200_linear_workspace(int64_t M, int64_t K) {
    if (M <= 2 || K <= 0) return 256;
    const int64_t ldq = round_up(K, 16);
    return (size_t)(round_up(M * ldq, 256) + 2 * round_up(M * 4, 256) + 256);
}

extern "C" int tb200_linear_forward(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                    const uint8_t* packed, int64_t ld_packed, int64_t N,
                                    const float* scale, const float* bias, float clip_val,
                                    void* out, int out_dtype, int64_t ldo,
                                    void* workspace, void* stream) {
    if (M <= 2)
        return tb200_gemv(x, x_dtype, M, K, ldx, packed, ld_packed, N, scale, bias, clip_val,
                          out, out_dtype, ldo, nullptr, nullptr, nullptr, 0, stream);
    TB_REQUIRE(workspace && (uintptr_t)workspace % 256 == 0, "linear_forward: workspace must be non-null and 256-byte aligned");
    LinearWs ws = carve_ws(workspace, M, K);
    if (M <= TB200_GEMV_MAX_M) {
        int rc = tb200_quant_act_i8(x, x_dtype, M, K, ldx, ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, TB200_QLAYOUT_GEMM, stream);
        if (rc) return rc;
        return tb200_gemv_q(ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, M, K, packed, ld_packed, N, scale, bias, clip_val,
                            out, out_dtype, ldo, nullptr, TB200_FLAG_PDL, stream);
    }
    int rc = tb200_quant_act_i8(x, x_dtype, M, K, ldx, ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, TB200_QLAYOUT_GEMM, stream);
    if (rc) return rc;
    return tb200_gemm_i8(ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, M, K, packed, ld_packed, N, scale, bias, clip_val,
                         out, out_dtype, ldo, nullptr, stream);
}

extern "C" int tb200_linear_forward_host(const void* x_host, int x_dtype, int64_t M, int64_t K,
                                         const uint8_t* packed, int64_t ld_packed, int64_t N,
                                         const float* scale, const float* bias, float clip_val,
                                         void* out_host, int out_dtype,
                                         void* dev_x, void* dev_out, void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(x_host && out_host && dev_x && dev_out, "linear_forward_host: null pointer");
    TB_REQUIRE(dtype_ok(x_dtype) && dtype_ok(out_dtype), "linear_forward_host: bad dtype");
    cudaStream_t s = (cudaStream_t)stream;
    TB_CUDA(cudaMemcpyAsync(dev_x, x_host, (size_t)(M * K) * dtype_size(x_dtype), cudaMemcpyHostToDevice, s));
    int rc = tb200_linear_forward(dev_x, x_dtype, M, K, K, packed, ld_packed, N, scale, bias, clip_val,
                                  dev_out, out_dtype, N, workspace, stream);
    if (rc) return rc;
    TB_CUDA(cudaMemcpyAsync(out_host, dev_out, (size_t)(M * N) * dtype_size(out_dtype), cudaMemcpyDeviceToHost, s));
    return TB200_OK;
}
