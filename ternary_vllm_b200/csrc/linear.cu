// linear.cu — whole-layer entry points: shape dispatch (GEMV vs tcgen05 GEMM) and the
// host-buffer end-to-end call.  Mirrors ScaledTernaryLinear.forward (binary/modules.py:101-109).
#include <cstdlib>
#include "umma.cuh"

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
namespace tb200 {
int gemv_fused(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
               const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
               const float* scale, const float* bias, int64_t stride_sb, float clip_val,
               void* out, int out_dtype, int64_t ldo, int64_t stride_out,
               int32_t* acc_out, int layers, int flags, cudaStream_t s,
               void* const* peers, int n_peers);
int gemv_q_impl(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx, int64_t M, int64_t K,
                const uint8_t* packed, int64_t ld_packed, int64_t N, const float* scale, const float* bias, float clip_val,
                void* out, int out_dtype, int64_t ldo, int32_t* acc_out, int flags, void* stream,
                void* const* peers, int n_peers);
}
using namespace tb200;

extern "C" size_t tb200_linear_workspace(int64_t M, int64_t K) {
    if (M <= TB200_GEMV_FUSED_M || K <= 0) return 256;
    const int64_t ldq = round_up(K, 16);
    return (size_t)(round_up(M * ldq, 256) + 2 * round_up(M * 4, 256) + 256);
}

extern "C" int tb200_linear_forward(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                    const uint8_t* packed, int64_t ld_packed, int64_t N,
                                    const float* scale, const float* bias, float clip_val,
                                    void* out, int out_dtype, int64_t ldo,
                                    void* workspace, void* stream) {
    if (M <= TB200_GEMV_FUSED_M)
        return tb200_gemv(x, x_dtype, M, K, ldx, packed, ld_packed, N, scale, bias, clip_val,
                          out, out_dtype, ldo, nullptr, nullptr, nullptr, 0, stream);
    TB_REQUIRE(workspace && (uintptr_t)workspace % 256 == 0, "linear_forward: workspace must be non-null and 256-byte aligned");
    LinearWs ws = carve_ws(workspace, M, K);
    // measured on B200 (4096 -> 11008): two 8-token GEMV passes beat the tcgen05 GEMM's fixed cost, four do not
    static const int gemv_max_env = (int)tb_knob("TB200_GEMV_MAX_M", TB200_LINEAR_GEMV_M);
    if (M <= gemv_max_env) {
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

// Tensor-parallel decode layer with the all-gather fused into the GEMV epilogue: this rank's [M, N_local]
// output block is stored straight into every rank's [M, ldo] buffer through peer-mapped (NVLink) pointers.
// out_peers[r] = address, in rank r's buffer, of element (0, first column of THIS rank's block).  The caller
// issues a cross-rank barrier afterwards (ternary_vllm_b200.parallel).  More than 32 tokens (prefill): activation
// quantiser + the tcgen05 GEMM, whose epilogue hands every 32-token x 128-channel output tile to ONE TMA tile store per
// rank (the all-gather of the [M, N_local] shards costs no extra pass over the output and no NCCL call).
extern "C" int tb200_linear_forward_tp(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                       const uint8_t* packed, int64_t ld_packed, int64_t N_local,
                                       const float* scale, const float* bias, float clip_val,
                                       void* const* out_peers, int n_peers, int out_dtype, int64_t ldo,
                                       void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N_local == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out_peers && ((n_peers >= 1 && n_peers <= 8) || n_peers == -1), "linear_forward_tp: bad peer list");
    TB_REQUIRE(n_peers != -1 || M > TB200_GEMV_MAX_M, "linear_forward_mc: the multicast form serves prefill (M > 32); use tb200_gemv_stream_mc for decode lists");
    TB_REQUIRE(K > 0 && ldx >= K && ldo >= N_local, "linear_forward_tp: bad shape");
    const int64_t kc = (K + 63) / 64;
    TB_REQUIRE(ld_packed % 16 == 0 && ld_packed >= kc * 16 && (uintptr_t)packed % 16 == 0, "linear_forward_tp: bad packed layout");
    if (M <= TB200_GEMV_FUSED_M)
        return gemv_fused(x, x_dtype, M, K, ldx, 0, packed, ld_packed, 0, N_local, scale, bias, 0, clip_val,
                          out_peers[0], out_dtype, ldo, 0, nullptr, 1, 0, (cudaStream_t)stream,
                          out_peers, n_peers);
    TB_REQUIRE(workspace && (uintptr_t)workspace % 256 == 0, "linear_forward_tp: workspace must be non-null and 256-byte aligned");
    LinearWs ws = carve_ws(workspace, M, K);
    int rc = tb200_quant_act_i8(x, x_dtype, M, K, ldx, ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, TB200_QLAYOUT_GEMM, stream);
    if (rc) return rc;
    const bool mc = n_peers < 0;                   // tb200_linear_forward_mc: out_peers[0] is a multicast address
    if (mc) n_peers = 1;
    if (M > TB200_GEMV_MAX_M) {
        TB_REQUIRE(dtype_ok(out_dtype), "linear_forward_tp: bad out dtype");
        TB_REQUIRE(!mc || N_local % (16 / (int64_t)dtype_size(out_dtype)) == 0, "linear_forward_mc: N_local must be a multiple of 8 (16-bit) / 4 (fp32)");
        UmmaParams p = {};
        p.mc = mc ? 1 : 0;
        p.s_x = ws.s_x; p.sum_qx = ws.sum_qx; p.scale = scale; p.bias = bias;
        p.out = out_peers[0]; p.ldo = ldo; p.acc_out = nullptr;
        p.M = M; p.N = N_local; p.K = K; p.clip_val = fabsf(clip_val);
        p.layers = 1; p.stride_act_rows = M; p.stride_sb = 0; p.stride_out = 0;
        p.n_peers = n_peers;
        for (int r = 0; r < n_peers; ++r) p.peers[r] = out_peers[r];
        p.pdl = 1;                                  // launched directly behind the quantiser
        return gemm_ts_launch(ws.q_x, ws.ldq, packed, ld_packed, p, out_dtype, (cudaStream_t)stream);
    }
    return gemv_q_impl(ws.q_x, ws.ldq, ws.s_x, ws.sum_qx, M, K, packed, ld_packed, N_local, scale, bias, clip_val,
                       out_peers[0], out_dtype, ldo, nullptr, TB200_FLAG_PDL, stream, out_peers, n_peers);
}

// Prefill tensor parallelism through an NVSwitch multicast mapping: as tb200_linear_forward_tp for M > 32, but every output tile
// leaves the GPU once (multimem stores from the tcgen05 GEMM's epilogue) and the switch replicates it into every rank's buffer.
extern "C" int tb200_linear_forward_mc(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                       const uint8_t* packed, int64_t ld_packed, int64_t N_local,
                                       const float* scale, const float* bias, float clip_val,
                                       void* out_multicast, int out_dtype, int64_t ldo, void* workspace, void* stream) {
    void* peers[1] = {out_multicast};
    return tb200_linear_forward_tp(x, x_dtype, M, K, ldx, packed, ld_packed, N_local, scale, bias, clip_val, peers, -1, out_dtype, ldo,
                                   workspace, stream);
}

// ---- allocation-free whole-layer call (no counterpart in the reference) ----
// Callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) have no
// workspace to hand to tb200_linear_forward.  This variant keeps the scratch of the activation quantiser (int8 rows,
// per-token scales and sums, read by the GEMV / tcgen05 GEMM that follows) in a file-scope device array: every GPU that
// loads the library holds TB200_STATIC_SCRATCH_BYTES until exit.  All calls share the array, on every stream: calls
// that overlap on one GPU race.  Nothing in it has to start at zero -- the quantiser writes every byte the matmul kernels
// read before they read it (the caller-provided workspace of tb200_linear_forward is uninitialised memory as well).
static __device__ __align__(256) unsigned char g_static_scratch[TB200_STATIC_SCRATCH_BYTES];

namespace tb200 {
// g_static_scratch is, in host code, a host-side stand-in for the symbol, not a device address: the device address is
// asked for once per device, with that device current.
constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
static std::atomic<void*> g_static_scratch_ptr[STATIC_SCRATCH_MAX_DEVICES];

static int static_scratch(void** ptr) {
    int dev = -1;
    TB_CUDA(cudaGetDevice(&dev));
    TB_REQUIRE(dev >= 0 && dev < STATIC_SCRATCH_MAX_DEVICES, "linear_forward_static: device ordinal %d out of range", dev);
    void* p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
    if (!p) {
        TB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));       // of the CURRENT device's copy of the array
        TB_REQUIRE(p && (uintptr_t)p % 256 == 0, "linear_forward_static: cudaGetSymbolAddress returned a bad address");
        g_static_scratch_ptr[dev].store(p, std::memory_order_release);   // racing first calls store the same value
    }
    *ptr = p;
    return TB200_OK;
}
}  // namespace tb200

extern "C" int tb200_linear_forward_static(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                           const uint8_t* packed, int64_t ld_packed, int64_t N,
                                           const float* scale, const float* bias, float clip_val,
                                           void* out, int out_dtype, int64_t ldo, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(M >= 0 && K >= 0, "linear_forward_static: bad shape");
    const size_t need = tb200_linear_workspace(M, K);
    TB_REQUIRE(need <= (size_t)TB200_STATIC_SCRATCH_BYTES,
               "linear_forward_static: M=%lld tokens x K=%lld needs %zu bytes of scratch, the static array holds %zu "
               "(TB200_STATIC_SCRATCH_BYTES); split the tokens or call tb200_linear_forward with a workspace",
               (long long)M, (long long)K, need, (size_t)TB200_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    int rc = static_scratch(&scratch);
    if (rc) return rc;
    return tb200_linear_forward(x, x_dtype, M, K, ldx, packed, ld_packed, N, scale, bias, clip_val, out, out_dtype, ldo, scratch, stream);
}
