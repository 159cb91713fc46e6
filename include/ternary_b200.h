/* ternary_b200.h — C ABI of libternary_b200.so (B200 / sm_100a only).
 *
 * This is the drop-in boundary for the quantised-Linear hot path of
 * BenCrulis/ternary-vllm.  Each entry point names the reference interface it
 * replaces (paths relative to the reference tree).  Conventions:
 *
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - weights are [N, K] = [out_features, in_features] row-major, as in the
 *     reference (binary/modules.py:96);
 *   - packed weights are "format C" of the reference
 *     (impl/tf/moondream/ternary.py:374-398): row n, byte j, bits [2i, 2i+1]
 *     hold code(W[n, 4j+i]) with code = value + 1 in {0,1,2}; padding code is 1
 *     (value 0).  `ld_packed` is the row pitch in bytes (>= ceil(K/4));
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*),
 *     allocates nothing unless stated, and returns 0 on success or a negative
 *     TB200_E* code; tb200_last_error() gives the message of the calling
 *     thread's last failure;
 *   - there is no CPU fallback: on a machine without an sm_100 GPU every
 *     compute entry returns TB200_ENODEV.
 */
#ifndef TERNARY_B200_H
#define TERNARY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB200_ABI_VERSION 1

/* error codes */
#define TB200_OK        0
#define TB200_EINVAL   -1 /* bad argument (shape, alignment, enum)            */
#define TB200_ECUDA    -2 /* CUDA runtime/driver error, see tb200_last_error  */
#define TB200_ENODEV   -3 /* no sm_100 device                                 */
#define TB200_ENOTSUP  -4 /* combination not implemented                      */

/* element types of float-like buffers */
#define TB200_F32  0
#define TB200_BF16 1
#define TB200_F16  2

/* quantiser decision rules (binary/modules.py:5-26) */
#define TB200_Q_TERNARY  0 /* tri_step_ste: +1 if w>=0.5, -1 if w<=-0.5, else 0 */
#define TB200_Q_SIGN     1 /* sign_ste:     +1 if w>=0 (incl. -0.0), else -1     */
#define TB200_Q_STEP     2 /* step_ste:     +1 if w>=0, else 0                   */

/* reference pack layouts (impl/tf/moondream/ternary.py) */
#define TB200_FMT_A 0 /* pack_ternary   :57-78   planar, bit0=(t==+1), bit1=(t==-1) */
#define TB200_FMT_B 1 /* pack_ternary2  :114-133 planar, code = t+1                 */
#define TB200_FMT_C 2 /* ScaledTernaryPackedMM.pack_fn :374-398 row-interleaved     */

int         tb200_abi_version(void);
const char* tb200_last_error(void);
/* number of kernels this library has launched in the calling process (for bench.py's gpu_launches) */
uint64_t    tb200_launch_count(void);

/* debug: when non-null, the decode GEMV writes 16 clock64() stamps per CTA (16*#CTAs int64) */
void        tb200_debug_set_trace(long long* dev_buf);

/* ---- weight preparation ------------------------------------------------ */

/* Replaces tri_step_ste / sign_ste / step_ste (binary/modules.py:5-26) followed by the
 * int8 cast + pack of utils/conversion.py:52-53,71-72 (ops.pack_fn).  One pass:
 * w [N,K] (dtype, row pitch ldw elements) -> packed [N, ld_packed] format C.
 * Bytes in [ceil(K/4), ld_packed) are written with the padding code 0x55.
 * `codes_out` (optional, int8 [N,K] dense) receives the unpacked codes {-1,0,1}. */
int tb200_quantize_pack(const void* w, int dtype, int64_t N, int64_t K, int64_t ldw,
                        int mode, uint8_t* packed, int64_t ld_packed,
                        int8_t* codes_out, void* stream);

/* Replaces ops.unpack_ternary(pw, orig_dim) (impl/tf/moondream/ternary.py:334):
 * packed format C -> int8 [N,K] values in {-1,0,1,(2)}. */
int tb200_unpack(const uint8_t* packed, int64_t ld_packed, int64_t N, int64_t K,
                 int8_t* out, void* stream);

/* int8 codes [T] (flattened) -> reference format A or B (planar, 1-D), or [N,K] -> C.
 * Replaces pack_ternary / pack_ternary2 / pack_fn.  out has ceil(T/4) bytes (A,B) or
 * N*ceil(K/4) bytes (C, tight pitch). For A/B pass N=1, K=T. */
int tb200_pack_format(const int8_t* codes, int64_t N, int64_t K, int fmt,
                      uint8_t* out, void* stream);
/* inverse: unpack_ternary_iterative / unpack_ternary_iterative2 / unpack (C). `out` has
 * 4*ceil(T/4) entries for A/B (the reference returns the padded length) and N*K for C. */
int tb200_unpack_format(const uint8_t* packed, int64_t N, int64_t K, int fmt,
                        int8_t* out, void* stream);

/* Replaces clip_binary_weights (binary/utils.py:74-77): in-place clip(w, -1, 1). */
int tb200_clip_latent(void* w, int dtype, int64_t numel, void* stream);

/* ---- 1-bit packed weights for the two-valued layers (ScaledBinaryLinear, ScaledBinary01Linear, SmoothlyBinarizedLinear:
 * binary/modules.py:55-90,112-146; one bit per weight as binary/utils.py:42-45 accounts them) ----
 * Format B1: [N, ld1] uint8, bit (k & 7) of byte k >> 3 of row n = (w[n,k] >= 0) (sign_ste and step_ste decide alike,
 * -0.0 -> 1); value = 2 bit - 1 (binary01 = 0) or bit (binary01 = 1); ld1 = tb200_b1_pitch(K) = 16 * ceil(K / 128), padding 0. */
int64_t tb200_b1_pitch(int64_t K);
int tb200_quantize_pack_b1(const void* w, int dtype, int64_t N, int64_t K, int64_t ldw, uint8_t* packed, int64_t ld1,
                           void* stream);
/* B1 -> the 2-bit format C (codes value + 1, pad code 1), row pitch ld2 (even, >= ceil(K/4)): feeds tb200_linear_forward /
 * tb200_gemm_i8 for more than 8 tokens. */
int tb200_expand_b1(const uint8_t* packed1, int64_t ld1, int64_t N, int64_t K, int binary01, uint8_t* packed2, int64_t ld2,
                    void* stream);
/* Decode GEMV on B1 weights, M <= TB200_GEMV_FUSED_M tokens, on the outputs of tb200_quant_act_i8(..., TB200_QLAYOUT_NATURAL)
 * (q_x [M, ldq] int8 with ldq % 16 == 0, s_x, sum_qx):  out = clip(((2 - binary01) * sum_k bit q - (1 - binary01) * sum_qx)
 * * s_x * scale + bias).  flags: TB200_FLAG_PDL when launched directly behind the quantiser. */
/* The same for M <= 2 tokens straight from x [M, ldx] (bf16 / f16 / f32): the activation quantisation runs in the kernel's
 * prologue, ONE launch per layer (the form ScaledBinaryLinear.forward takes at pack_bits == 1). */
int tb200_gemv_b1_fused(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, const uint8_t* packed1, int64_t ld1,
                        int64_t N, const float* scale, const float* bias, float clip_val, int binary01, void* out,
                        int out_dtype, int64_t ldo, int flags, void* stream);
int tb200_gemv_b1(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx, int64_t M, int64_t K,
                  const uint8_t* packed1, int64_t ld1, int64_t N, const float* scale, const float* bias, float clip_val,
                  int binary01, void* out, int out_dtype, int64_t ldo, int flags, void* stream);

/* ---- optimiser-side helpers of the smooth-binary training loop (binary/modules.py:149-176, train.py:296-297) ----
 * compute_and_set_biobjective_gradient(model) = for every parameter gradient tb200_sumsq_accumulate(grad, .., 0, norm_g_sq),
 * for every smooth-binary weight tensor tb200_sumsq_accumulate(weights, .., 1, norm_bg_sq)  (|-sign(w)|^2 = #nonzero),
 * then tb200_biobjective_apply per gradient.  The accumulators are device doubles the caller zeroes; nothing is read back
 * to the host.  `workspace`: tb200_train_workspace() bytes, 8-byte aligned, ZERO before the first call. */
size_t tb200_train_workspace(void);
int tb200_sumsq_accumulate(const void* x, int dtype, int64_t numel, int count_nonzero, double* acc,
                           void* workspace, void* stream);
/* In place on `grad` [numel].  weights != NULL: the layer is smooth-binary and `grad` is its weights.grad:
 *   grad <- (((grad / n_g + -sign(w) / n_bg) * (n_g + n_bg)) * 0.5 / n_g) * (n_g + n_bg)      (modules.py:166-176: the
 *   reference rescales the blended gradient once more in its per-parameter loop);  weights == NULL (any other parameter):
 *   grad <- (grad / n_g) * (n_g + n_bg).  n_g == 0 leaves the rescale out (`if norm_g > 0.0 else`), n_bg == 0 zeroes its term. */
int tb200_biobjective_apply(void* grad, const void* weights, int dtype, int64_t numel, const double* norm_g_sq,
                            const double* norm_bg_sq, void* stream);

/* Replaces the ternary branch of linear_to_quantized (binary/modules.py:275-298):
 * thresholded-mean ("k-means") scale, `iters` refinements, then latent = clip(W/m, +-1).
 * per_row=0: one scale for the tensor ("uniform"), broadcast into scale[N];
 * per_row=1: one scale per output row.  W: [N,K] dtype (dense); latent [N,K] and scale [N]
 * are fp32, like the parameters of the freshly built reference module (modules.py:277).
 * `workspace` must hold tb200_ternary_init_workspace(N) bytes. */
size_t tb200_ternary_init_workspace(int64_t N);
int tb200_ternary_init(const void* W, int dtype, int64_t N, int64_t K, int per_row, int iters,
                       float* latent, float* scale, void* workspace, void* stream);

/* Replaces the binary / binary01 branches (binary/modules.py:193-266) with
 * scale = (sum_{w>0} w - sum_{w<0} w)/count (binary01, positive_only=1: positive part
 * only), latent = clip(W * prescale, +-1).  prescale is a host float (1.0 for scaling
 * "none", 1/std for "std"); natural=1 uses 1/scale[row] instead (scaling "natural").
 * latent, scale: fp32.  workspace: tb200_ternary_init_workspace(N) bytes. */
int tb200_binary_init(const void* W, int dtype, int64_t N, int64_t K, int per_row, int positive_only,
                      float prescale, int natural, float* latent, float* scale,
                      void* workspace, void* stream);

/* ---- activations --------------------------------------------------------- */

/* Per-token symmetric int8 quantisation: s_x[m] = max|x[m,:]|/127, inv = 127/max|x[m,:]|
 * (both 1.0 if the row is all zero), q = clamp(rint(x * inv), -127, 127), fp32 multiply,
 * round-to-nearest-even.
 * x [M,K] dtype (pitch ldx elements) -> q_x int8 [M, ldq] ; s_x f32 [M] ; sum_qx i32 [M].
 * layout 0: natural K order.  layout 1: the K order the tcgen05 GEMM consumes (within every
 * group of 16, position 4i+b holds element 4b+i).  layout 2: the decode GEMV's shared-memory
 * image (ldq must be 4096*ceil(K/4096)).  Columns beyond K are zero-filled up to ldq. */
#define TB200_QLAYOUT_NATURAL 0
#define TB200_QLAYOUT_GEMM    1
#define TB200_QLAYOUT_GEMV    2
#define TB200_QLAYOUT_GEMM_SPLIT 3 /* library-internal operand order of the 32-token tcgen05 kernel of decode lists: inside every
                                     256-element block, element 16j + 4b + f (f = 2-bit field of packed byte b of word j) sits at
                                     128 (f & 1) + 8 j + 4 (f >> 1) + b; needs ldq % 256 == 0 (zero-filled beyond K) */
int tb200_quant_act_i8(const void* x, int dtype, int64_t M, int64_t K, int64_t ldx,
                       int8_t* q_x, int64_t ldq, float* s_x, int32_t* sum_qx, int layout,
                       void* stream);

/* ---- the fused op: ops.ternary_matmul(x, pw, scale, bias, clip_val) ------ */
/* (impl/tf/moondream/ternary.py:279-281; same maths as ScaledTernaryLinear.forward,
 *  binary/modules.py:101-109):  out = clip((x @ unpack(pw)^T) * scale + bias, +-clip_val)
 *
 * x      [M,K]  x_dtype, pitch ldx elements
 * packed [N, ld_packed] format C; scale [N] f32 or NULL (=1); bias [N] f32 or NULL (=0)
 * out    [M,N]  out_dtype, pitch ldo elements
 * acc_out (optional, i32 [M,N] dense): the exact integer accumulators sum_k q_x*q_w
 * sx_out  (optional, f32 [M]) / qx_out (optional, int8 [M,K] dense): the activation
 *         quantisation the kernel used, for parity tests.                             */

/* Decode shape, M <= TB200_GEMV_MAX_M.  HBM-bound: int8 activation quantisation in the
 * prologue (8 tokens per pass over the weights), TMA-staged packed rows, in-register unpack,
 * mma.sync.m16n8k32 (u8 x s8) integer dot products, fused epilogue.  K <= 20480. */
#define TB200_GEMV_MAX_M 32
#define TB200_GEMV_FUSED_M 8 /* tokens per pass over the weights; up to here tb200_linear_forward is ONE launch */
#define TB200_LINEAR_GEMV_M 16 /* tb200_linear_forward: largest M served by GEMV passes, above it the tcgen05 GEMM */
/* flags bit 0: programmatic dependent launch.  The kernel prefetches its packed weights
 * before waiting for the previous kernel in the stream (which normally produces x); the
 * caller asserts that the previous kernel does NOT write `packed`. */
#define TB200_FLAG_PDL 1
int tb200_gemv(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
               const uint8_t* packed, int64_t ld_packed, int64_t N,
               const float* scale, const float* bias, float clip_val,
               void* out, int out_dtype, int64_t ldo,
               int32_t* acc_out, float* sx_out, int8_t* qx_out, int flags, void* stream);

/* Decode GEMV on activations pre-quantised by tb200_quant_act_i8(layout 1): one pass over the
 * weights per 8 tokens.  tb200_gemv quantises in its prologue instead, which every CTA repeats
 * (each CTA reads all M*K activations from L2): preferable up to 8 tokens. */
int tb200_gemv_q(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx,
                 int64_t M, int64_t K,
                 const uint8_t* packed, int64_t ld_packed, int64_t N,
                 const float* scale, const float* bias, float clip_val,
                 void* out, int out_dtype, int64_t ldo, int32_t* acc_out, int flags, void* stream);

/* Same contract with float activations and fp32 accumulation (no activation
 * quantisation): the reference's exact semantics, for M <= TB200_GEMV_MAX_M. */
int tb200_gemv_f32(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                   const uint8_t* packed, int64_t ld_packed, int64_t N,
                   const float* scale, const float* bias, float clip_val,
                   void* out, int out_dtype, int64_t ldo, void* stream);

/* Grouped decode: `count` independent layers of identical shape in ONE launch (weights of
 * layer l at packed + l*stride_packed, activations at x + l*stride_x elements, ...). */
int tb200_gemv_grouped(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                       const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                       const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                       void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                       int count, int flags, void* stream);

/* Streaming decode over a LIST of `count` layers of identical shape (M <= 32 tokens each): a per-layer activation
 * quantiser launch followed by ONE persistent kernel that keeps the TMA weight stream running across layer boundaries
 * (csrc/gemv_stream.cu).  Same arguments as tb200_gemv_grouped plus acc_out (optional, i32 [M,N], layer 0 only) and
 * a 256-byte aligned workspace of tb200_gemv_stream_workspace(M, K, count) bytes.  count = 1 serves a single layer
 * with 9..32 tokens in one pass over the weights.  Returns TB200_ENOTSUP when K*M exceeds the shared-memory image. */
size_t tb200_gemv_stream_workspace(int64_t M, int64_t K, int count);
int tb200_gemv_stream(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                      const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                      const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                      void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                      int count, int32_t* acc_out, void* workspace, void* stream);

/* Tensor-parallel form of tb200_gemv_stream (the reference has no distributed code): `packed`,
 * `scale`, `bias` hold this rank's row block (N_local rows) of every layer; the all-gather is fused into the
 * epilogues: each output element is stored into EVERY rank's [count, M, ldo] buffer through peer-mapped (NVLink)
 * pointers.  out_peers[r] addresses element (layer 0, token 0, first column of THIS rank's block) of rank r's
 * buffer (own buffer included), 1 <= n_peers <= 8.  ONE cross-rank barrier by the caller after the call makes the
 * gathered outputs of the whole list consumable. */
int tb200_gemv_stream_tp(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                         const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N_local,
                         const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                         void* const* out_peers, int n_peers, int out_dtype, int64_t ldo, int64_t stride_out,
                         int count, void* workspace, void* stream);
/* The same through an NVSwitch MULTICAST mapping of the output buffers (e.g. torch symmetric memory, handle.multicast_ptr):
 * `out_multicast` = multicast address of element (layer 0, token 0, first column of THIS rank's block).  Every 16 bytes of output
 * leave the GPU once (multimem.st) and the switch writes them into every rank's buffer: egress 1x instead of n_ranks x.
 * N_local must be a multiple of 8 (fp32 outputs: 4).  The caller issues the cross-rank barrier afterwards. */
int tb200_gemv_stream_mc(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                         const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N_local,
                         const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                         void* out_multicast, int out_dtype, int64_t ldo, int64_t stride_out,
                         int count, void* workspace, void* stream);
/* Prefill (M > 32) tensor-parallel layer through a multicast mapping: quantiser + tcgen05 GEMM whose epilogue multicasts every
 * output tile (see tb200_linear_forward_tp for the unicast form and the argument meaning). */
int tb200_linear_forward_mc(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                            const uint8_t* packed, int64_t ld_packed, int64_t N_local,
                            const float* scale, const float* bias, float clip_val,
                            void* out_multicast, int out_dtype, int64_t ldo, void* workspace, void* stream);


/* tb200_gemv_stream with HOST activations x_host [count, M, K] and outputs out_host [count, M, N] (dense, pinned):
 * one H2D copy, the two launches, one D2H copy, in order on `stream`.  dev_x / dev_out: device staging buffers of the
 * same sizes.  The end-to-end form of a decode step over a list of layers (bench.py `e2e`). */
int tb200_gemv_stream_host(const void* x_host, int x_dtype, int64_t M, int64_t K,
                           const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                           const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                           void* out_host, int out_dtype, int count,
                           void* dev_x, void* dev_out, void* workspace, void* stream);

/* ---- a CHAIN of dependent layers in one launch (the fused MLP block) -----------------------
 * The reference wires the layer as fc1 -> gelu_new -> fc2 (impl/tf/moondream/layers.py:32-43) and passes the
 * activation INTO the ternary layer, applied after the clip (impl/tf/moondream/ternary.py:245-246,298-299;
 * utils/conversion.py:185-196):   h_{l+1} = act_l( clip((q(h_l) @ unpack(pw_l)^T) * s_x * scale_l + bias_l) ).
 * tb200_chain_forward runs n such layers, each consuming the previous one's output, in ONE persistent launch for
 * decode shapes (M <= TB200_CHAIN_MAX_M tokens): the weight stream never stops at a layer boundary, the boundary is
 * a flag in global memory, and every layer's input is quantised per token to int8 inside the kernel.  Activations
 * between layers are kept in x's dtype (the reference keeps x.dtype); the activation function is evaluated in fp32 on
 * that stored precision. */
#define TB200_ACT_NONE      0
#define TB200_ACT_GELU_TANH 1 /* gelu_new / keras gelu(approximate=True): 0.5x(1+tanh(sqrt(2/pi)(x+0.044715x^3))) */
#define TB200_ACT_RELU      2
#define TB200_CHAIN_MAX_LAYERS 128
#define TB200_CHAIN_MAX_M 8
#define TB200_CHAIN_CTRL_BYTES 16640 /* control block at the start of the workspace: 128 B + one 128-byte line per layer */
typedef struct tb200_chain_layer {
    const uint8_t* packed;   /* [N, ld_packed] format C, device */
    int64_t ld_packed, N, K; /* K must equal the previous layer's N */
    const float* scale;      /* [N] f32 or NULL, device */
    const float* bias;       /* [N] f32 or NULL, device */
    float clip_val;
    int32_t act;             /* TB200_ACT_*: applied after the clip */
} tb200_chain_layer;
/* A plan is a caller-owned blob: build it ONCE on the host (tb200_chain_plan_bytes(n) bytes; holds the layers' TMA
 * descriptors), copy it to 256-byte aligned device memory, pass both copies to tb200_chain_forward.  *workspace_bytes
 * receives the size of the device workspace (inter-layer activations + the control block); the workspace belongs
 * to this plan and its first TB200_CHAIN_CTRL_BYTES bytes must be ZERO before the first call (the kernel maintains them afterwards). */
size_t tb200_chain_plan_bytes(int n_layers);
int tb200_chain_plan_build(const tb200_chain_layer* layers, int n_layers, int64_t M_max, int act_elem_size,
                           void* plan_host, size_t* workspace_bytes);
/* x [M, ldx] x_dtype -> out [M, ldo] out_dtype (N of the last layer).  flags: TB200_FLAG_PDL as for tb200_gemv. */
int tb200_chain_forward(const void* plan_host, const void* plan_dev, const void* x, int x_dtype, int64_t M, int64_t ldx,
                        void* out, int out_dtype, int64_t ldo, void* workspace, int flags, void* stream);

/* Prefill shape: tcgen05.mma kind::i8, TMEM accumulators.  Takes PRE-QUANTISED activations
 * from tb200_quant_act_i8(kperm=1): q_x int8 [M, ldq], s_x [M], sum_qx [M].
 * Requires ldq % 16 == 0, ld_packed % 16 == 0, 16-byte aligned base pointers. */
int tb200_gemm_i8(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx,
                  int64_t M, int64_t K,
                  const uint8_t* packed, int64_t ld_packed, int64_t N,
                  const float* scale, const float* bias, float clip_val,
                  void* out, int out_dtype, int64_t ldo,
                  int32_t* acc_out, void* stream);

/* Whole layer by token count: M <= 8 tb200_gemv (fused quantisation, one launch); M <= 16 quantise
 * (layout 1) + tb200_gemv_q; larger M quantise (layout 1) + tb200_gemm_i8.
 * workspace: tb200_linear_workspace(M,K) bytes, 256-byte aligned. */
size_t tb200_linear_workspace(int64_t M, int64_t K);
int tb200_linear_forward(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                         const uint8_t* packed, int64_t ld_packed, int64_t N,
                         const float* scale, const float* bias, float clip_val,
                         void* out, int out_dtype, int64_t ldo,
                         void* workspace, void* stream);

/* Allocation-free variant of tb200_linear_forward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and results as tb200_linear_forward minus `workspace`: the activation quantiser's scratch
 * lives in a file-scope device array of TB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  Limit: returns TB200_EINVAL when
 * tb200_linear_workspace(M, K) > TB200_STATIC_SCRATCH_BYTES (8 MiB: e.g. 1024 tokens at K = 4096, 729 at K = 11008);
 * M <= 8 never touches the scratch.  All calls on a GPU share the array and it keeps the last call's contents (nothing in
 * it needs to start at zero: each call rewrites what it reads): calls that OVERLAP on one GPU -- two streams, or two
 * graphs replayed concurrently -- race; serialise them or give the others tb200_linear_forward with their own workspace. */
#define TB200_STATIC_SCRATCH_BYTES (8u << 20)
int tb200_linear_forward_static(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                                const uint8_t* packed, int64_t ld_packed, int64_t N,
                                const float* scale, const float* bias, float clip_val,
                                void* out, int out_dtype, int64_t ldo, void* stream);

/* Tensor-parallel decode layer (M <= 32) with the all-gather fused into the epilogue: this rank's
 * [M, N_local] output block is stored into EVERY rank's [M, ldo] buffer through peer-mapped (NVLink)
 * pointers; out_peers[r] addresses element (0, first column of this rank's block) of rank r's buffer
 * (own buffer included).  A cross-rank barrier by the caller makes the gathered rows consumable.
 * Replaces the NCCL all-gather of SURVEY.md §8e for decode shapes; the reference has no counterpart. */
int tb200_linear_forward_tp(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                            const uint8_t* packed, int64_t ld_packed, int64_t N_local,
                            const float* scale, const float* bias, float clip_val,
                            void* const* out_peers, int n_peers, int out_dtype, int64_t ldo,
                            void* workspace, void* stream);

/* End-to-end with HOST activations/outputs (weights stay resident on the device):
 * H2D x, forward (GEMV or GEMM by M), D2H out, all on `stream`; x_host/out_host should be
 * pinned.  dev_x / dev_out are caller-provided device staging buffers ([M,K] / [M,N]).    */
int tb200_linear_forward_host(const void* x_host, int x_dtype, int64_t M, int64_t K,
                              const uint8_t* packed, int64_t ld_packed, int64_t N,
                              const float* scale, const float* bias, float clip_val,
                              void* out_host, int out_dtype,
                              void* dev_x, void* dev_out, void* workspace, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TERNARY_B200_H */
