// gemv.cu — decode-shaped packed-ternary GEMV (M <= 32 tokens).  HBM-bound integer work.
//
// One persistent CTA per SM streams its share of the packed weight matrix through a ring of
// shared-memory stages filled by TMA (cp.async.bulk.tensor, SWIZZLE_128B): a stage is 16 rows x
// 1024 bytes (16 x 4096 weights) fetched as 8 boxes of 16 rows x 128 bytes by a producer warp.
//   * The first ring-full of tiles is requested BEFORE the kernel waits on the previous kernel
//     (programmatic dependent launch): the weights of layer i+1 stream in from HBM while layer i
//     is still reducing, so a chain of decode layers keeps HBM busy across kernel boundaries.
//   * Integer dot products run on the legacy warp-level tensor path: mma.sync.m16n8k32 (u8 x s8
//     -> s32) with the 16 weight rows of the tile as the M dimension and up to 8 tokens as N.
//     (dp4a was measured first: IDP.4A is a scoreboarded, reduced-rate instruction on sm_100 and
//     capped the kernel at ~2.8 TB/s; see profiles/r1_gemv_dp4a_grouped.*.)
//   * Unpacking is in-register and costs 4 AND per 16 weights: field i of every packed byte is
//     masked in place (value code*4^i) and fed to its own accumulator set; the four sets are
//     recombined exactly with arithmetic shifts at the end.  No shifts, no stores of expanded
//     weights.  Codes are unsigned (w+1 in {0,1,2}); sum_k q_x is subtracted once per token.
//   * Warp w owns the w-th 128-byte segment of every row of the tile (K split 8 ways); the
//     SWIZZLE_128B layout makes the per-lane 16-byte fragment loads conflict-free.  Partial
//     sums meet in shared memory once per 16-row block, followed by the fused epilogue.
//   * Per-token int8 quantisation of the activations is fused into the prologue (tb200_gemv), or
//     taken pre-quantised from tb200_quant_act_i8 layout 1 (tb200_gemv_q, used for > 2 tokens).
// Reference semantics: ScaledTernaryLinear.forward (binary/modules.py:101-109) and
// ops.ternary_matmul (impl/tf/moondream/ternary.py:279-281).
#include <mutex>
#include "tma.cuh"

namespace tb200 {

constexpr int GV_ROWS = 16;                       // rows per tile = MMA M
constexpr int GV_SLICE = 1024;                    // packed bytes per row per tile (4096 weights)
constexpr int GV_SEG = 128;                       // bytes per row per TMA box / per warp
constexpr int GV_TILE_BYTES = GV_ROWS * GV_SLICE; // 16 KB
constexpr int GV_CWARPS = GV_SLICE / GV_SEG;      // 8 consumer warps, one per segment
constexpr int GV_THREADS = (GV_CWARPS + 1) * 32;  // + producer warp
constexpr int GV_MAX_STAGES = 6;
constexpr int GV_TOK = 8;                         // tokens per MMA n-tile

struct GemvParams {
    const void* x; int64_t ldx; int64_t stride_x;            // raw activations (fused quantisation)
    const int8_t* qx; int64_t ldq; int64_t stride_q;         // or pre-quantised, layout 1 (kperm)
    const float* sx_in; const int32_t* sumq_in; int64_t stride_s;
    const float* scale; const float* bias; int64_t stride_sb;
    void* out; int64_t ldo; int64_t stride_out;
    int32_t* acc_out; float* sx_out; int8_t* qx_out;
    int64_t N, K;
    int M;            // tokens in this launch (<= 8*NT)
    int nks;          // k-slices of 4096 weights
    int row_blocks;   // ceil(N / 16)
    int nrb;          // row blocks per CTA = ceil(row_blocks / gridDim.x)
    int stages;       // ring depth
    float clip_val;
    int pdl;
    bool xvec;
    long long* trace;   // debug: per-CTA clock64 stamps (tb200_debug_set_trace), normally null
    // tensor-parallel gather fused into the epilogue: when n_peers > 0 the output element is stored into every
    // rank's [M, ldo] buffer (peer-mapped NVLink addresses, already offset to this rank's column block)
    void* peers[8]; int n_peers;
};
static long long* g_trace = nullptr;
#define GV_STAMP(slot) do { if (p.trace && threadIdx.x == 0 && blockIdx.y == 0) p.trace[blockIdx.x * 16 + (slot)] = clock64(); } while (0)

// q = clamp(rint(v * inv), +-127): the activation quantiser (ternary_b200.h, tb200_quant_act_i8)
// rint() and the float->int conversion are done by adding 1.5*2^23 (the FADD rounds to nearest-even
// integer, the low mantissa bits are the result): identical bits, but no XU-pipe conversions.
__device__ __forceinline__ int quant_mul(float v, float inv) {
    float r = __fmul_rn(v, inv);
    r = fminf(fmaxf(r, -127.0f), 127.0f);
    return __float_as_int(__fadd_rn(r, 12582912.0f)) - 0x4B400000;
}

// 16 consecutive activations of one token as fp32 (zero beyond K); vector loads when aligned
template <typename XT>
__device__ __forceinline__ void load16(const XT* __restrict__ xr, int64_t g, int64_t K, bool vec, float v[16]) {
    if (vec && 16 * g + 16 <= K) {
        if (sizeof(XT) == 4) {
            const float4* q = reinterpret_cast<const float4*>(xr + 16 * g);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 f = q[j];
                v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
            }
        } else {
            const uint4* q = reinterpret_cast<const uint4*>(xr + 16 * g);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint4 u = q[j];
                const XT* e = reinterpret_cast<const XT*>(&u);
#pragma unroll
                for (int t = 0; t < 8; ++t) v[8 * j + t] = to_f32(e[t]);
            }
        }
    } else {
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            int64_t k = 16 * g + t;
            v[t] = k < K ? to_f32(xr[k]) : 0.0f;
        }
    }
}

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GV_CWARPS * 32) : "memory"); }
// the 8 consumer warps work as two groups of 4 that alternate row blocks (one reduces while the other issues MMAs)
constexpr int GV_GWARPS = GV_CWARPS / 2;
__device__ __forceinline__ void group_bar(int grp) { asm volatile("bar.sync %0, %1;" ::"r"(3 + grp), "n"(GV_GWARPS * 32) : "memory"); }

// D(16x8,s32) += A(16x32,u8 row) * B(32x8,s8 col)
__device__ __forceinline__ void imma_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                           uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Shared-memory image of the quantised activations, in 32-bit words.  For token n (0..7 of an
// n-tile), 16-byte chunk c of the packed row (64 weights), packed word j (16 weights) and field
// i, the word holds bytes b=0..3 = q[64c + 16j + 4b + i].  Lane (g = lane>>2, t = lane&3) of
// warp `seg` reads, for its chunk pair p, the 4 words j=0..3 of (c = 8*seg + 2t + p, i, n = g)
// with one LDS.128; lane l reads the l-th 16 bytes of a 512-byte run (conflict-free).  Token
// slots >= M are never initialised: column n of the MMA only feeds output column n.
__device__ __forceinline__ int xq_word(int c, int j, int i, int n) {
    const int sg = c >> 3, lc = c & 7, t = lc >> 1, pp = lc & 1;
    return ((((sg * 2 + pp) * 4 + i) * GV_TOK + n) * 4 + t) * 4 + j;     // uint4 index = ... + lane (lane = 4n + t)
}

template <typename XT, typename OT, int NT, bool FUSED>
__global__ void __launch_bounds__(GV_THREADS, 2)
gemv_mma_kernel(const __grid_constant__ CUtensorMap map_w, const GemvParams p) {
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;         // SWIZZLE_128B boxes: 1024-B aligned
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    // layout: ring[stages][16 KB] | xq[NT][8 tokens][nks*4096] | slots[2][NT][8 warps][32] uint4 | sx | sumq | red | barriers
    const uint32_t sRing = base;
    const int xq_words_nt = p.nks * 4096 * GV_TOK / 4;
    uint32_t* xq = reinterpret_cast<uint32_t*>(gbase + p.stages * GV_TILE_BYTES);
    uint4* slots = reinterpret_cast<uint4*>(xq + (size_t)NT * xq_words_nt);
    float* sh_sx = reinterpret_cast<float*>(slots + 2 * NT * GV_CWARPS * 32);   // slots: [2 groups][2 bufs][NT][4 warps][32]
    int* sh_sumq = reinterpret_cast<int*>(sh_sx + NT * GV_TOK);
    float* red_f = reinterpret_cast<float*>(sh_sumq + NT * GV_TOK);       // [8 tokens][8 warps]
    int* red_i = reinterpret_cast<int*>(red_f + GV_TOK * GV_CWARPS);
    uint64_t* bars = reinterpret_cast<uint64_t*>(red_i + GV_TOK * GV_CWARPS);
    float* sh_scale = reinterpret_cast<float*>(bars + 2 * GV_MAX_STAGES); // [nrb][16] this CTA's rows
    float* sh_bias = sh_scale + p.nrb * GV_ROWS;
    const uint32_t sBar = smem_u32(bars);
    auto bar_full = [&](int s) { return sBar + 8u * s; };
    auto bar_empty = [&](int s) { return sBar + 8u * (GV_MAX_STAGES + s); };

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int layer = blockIdx.y;
    GV_STAMP(0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(bar_full(s), 1);
            mbar_init(bar_empty(s), GV_GWARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    GV_STAMP(10);    // barriers initialised
    if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");

    if (warp == GV_CWARPS) {
        // ===================== producer: weights do not depend on the previous kernel =====================
        // Without PDL the activations are requested at the same time as the weights and would queue
        // behind a ring-full of weight traffic: let the consumers issue their x loads first.
        if (!p.pdl) asm volatile("bar.sync 2, %0;" ::"n"(GV_THREADS) : "memory");
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int rb = blockIdx.x; rb < p.row_blocks; rb += gridDim.x) {
                for (int ks = 0; ks < p.nks; ++ks) {
                    mbar_wait(bar_empty(s), ph ^ 1u);
                    mbar_arrive_expect_tx(bar_full(s), GV_TILE_BYTES);
#pragma unroll
                    for (int sg = 0; sg < GV_CWARPS; ++sg)
                        tma_load_3d(sRing + s * GV_TILE_BYTES + sg * (GV_ROWS * GV_SEG), &map_w, bar_full(s),
                                    ks * (GV_SLICE / 4) + sg * (GV_SEG / 4), rb * GV_ROWS, layer);
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
        return;
    }

    // ===================== consumers =====================
    const int ct = threadIdx.x;                       // 0..255
    const int g = lane >> 2, t = lane & 3;
    // this CTA's per-row scale and bias (weights: independent of the previous kernel): loads are issued
    // now, the values are parked in shared memory at the end of the prologue
    float sb_scale[2], sb_bias[2];
    {
        const float* scale_g = p.scale ? p.scale + layer * p.stride_sb : nullptr;
        const float* bias_g = p.bias ? p.bias + layer * p.stride_sb : nullptr;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int i = ct + u * GV_CWARPS * 32;
            const int64_t row = ((int64_t)blockIdx.x + (int64_t)(i >> 4) * gridDim.x) * GV_ROWS + (i & 15);
            const bool ok = i < p.nrb * GV_ROWS && row < p.N;
            sb_scale[u] = (scale_g && ok) ? scale_g[row] : 1.0f;
            sb_bias[u] = (bias_g && ok) ? bias_g[row] : 0.0f;
        }
    }
    if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");   // x comes from the previous kernel
    GV_STAMP(14);    // previous kernel complete (PDL) / consumers start

    if (FUSED) {
        const XT* x = reinterpret_cast<const XT*>(p.x) + layer * p.stride_x;
        const int groups = p.nks * 256;                      // 16-element groups incl. zero padding up to the slice
        constexpr int GC = 1;                                // groups cached in registers per thread (K <= 4096)
        const bool cached = groups <= GC * GV_CWARPS * 32;
        if (!cached && !p.pdl) asm volatile("bar.arrive 2, %0;" ::"n"(GV_THREADS) : "memory");
        for (int m = 0; m < p.M; ++m) {
            const XT* xr = x + (int64_t)m * p.ldx;
            float vc[GC][16];
            float amax = 0.0f;
            if (cached) {
#pragma unroll
                for (int u = 0; u < GC; ++u) {
                    const int gi = ct + u * GV_CWARPS * 32;
                    if (gi < groups) load16<XT>(xr, gi, p.K, p.xvec, vc[u]);
                }
                // the activation loads are in flight: now let the producer start the weight stream
                if (m == 0 && !p.pdl) asm volatile("bar.arrive 2, %0;" ::"n"(GV_THREADS) : "memory");
#pragma unroll
                for (int u = 0; u < GC; ++u) {
                    const int gi = ct + u * GV_CWARPS * 32;
                    if (gi < groups) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) amax = fmaxf(amax, fabsf(vc[u][e]));
                    }
                }
            } else {
                for (int gi = ct; (int64_t)gi * 16 < p.K; gi += GV_CWARPS * 32) {
                    float v[16];
                    load16<XT>(xr, gi, p.K, p.xvec, v);
#pragma unroll
                    for (int e = 0; e < 16; ++e) amax = fmaxf(amax, fabsf(v[e]));
                }
            }
            amax = warp_max(amax);
            if (lane == 0) red_f[m * GV_CWARPS + warp] = amax;
            GV_STAMP(11);    // activations arrived, amax of this warp known
            consumer_bar();
            GV_STAMP(12);
            float v = red_f[m * GV_CWARPS];
#pragma unroll
            for (int w = 1; w < GV_CWARPS; ++w) v = fmaxf(v, red_f[m * GV_CWARPS + w]);
            const float sxm = v > 0.0f ? __fdiv_rn(v, 127.0f) : 1.0f;
            const float inv = v > 0.0f ? __fdiv_rn(127.0f, v) : 1.0f;
            if (ct == 0) {
                sh_sx[m] = sxm;
                if (p.sx_out && blockIdx.x == 0 && layer == 0) p.sx_out[m] = sxm;
            }
            int lsum = 0;
            auto emit = [&](int gi, const float (&v16)[16]) {
                int qv[16];
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    qv[e] = quant_mul(v16[e], inv);          // v is 0 beyond K -> q = 0
                    lsum += qv[e];
                }
                if (p.qx_out && blockIdx.x == 0 && layer == 0) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        int64_t k = 16 * (int64_t)gi + e;
                        if (k < p.K) p.qx_out[(int64_t)m * p.K + k] = (int8_t)qv[e];
                    }
                }
                const int c = gi >> 2, j = gi & 3;
#pragma unroll
                for (int i = 0; i < 4; ++i)   // word i = elements i, 4+i, 8+i, 12+i (field i of bytes 0..3)
                    xq[xq_word(c, j, i, m)] = (uint32_t)(qv[i] & 0xff) | ((uint32_t)(qv[4 + i] & 0xff) << 8) |
                                              ((uint32_t)(qv[8 + i] & 0xff) << 16) | ((uint32_t)(qv[12 + i] & 0xff) << 24);
            };
            if (cached) {
#pragma unroll
                for (int u = 0; u < GC; ++u) {
                    const int gi = ct + u * GV_CWARPS * 32;
                    if (gi < groups) emit(gi, vc[u]);
                }
            } else {
                for (int gi = ct; gi < groups; gi += GV_CWARPS * 32) {
                    float v16[16];
                    load16<XT>(xr, gi, p.K, p.xvec, v16);      // zero beyond K
                    emit(gi, v16);
                }
            }
            lsum = warp_sum(lsum);
            if (lane == 0) red_i[m * GV_CWARPS + warp] = lsum;
            GV_STAMP(13);    // quantised image written
        }
    } else {
        // pre-quantised activations (tb200_quant_act_i8 layout 1): rows of kperm words
        const int8_t* q = p.qx + layer * p.stride_q;
        const int groups = p.nks * 256, have = (int)(p.ldq >> 4);   // zero padding up to the slice
        const int total = p.M * groups;
        bool arrived = p.pdl != 0;
        for (int base_i = 0; base_i < total; base_i += 8 * GV_CWARPS * 32) {
            uint4 u[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int idx = base_i + r * GV_CWARPS * 32 + ct;
                const int m = idx / groups, gi = idx - m * groups;
                u[r] = (idx < total && gi < have) ? *reinterpret_cast<const uint4*>(q + (int64_t)m * p.ldq + 16 * (int64_t)gi)
                                                   : make_uint4(0u, 0u, 0u, 0u);
            }
            if (!arrived) {   // first batch of activation loads is in flight: release the weight stream
                asm volatile("bar.arrive 2, %0;" ::"n"(GV_THREADS) : "memory");
                arrived = true;
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int idx = base_i + r * GV_CWARPS * 32 + ct;
                if (idx < total) {
                    const int m = idx / groups, gi = idx - m * groups;
                    const int c = gi >> 2, j = gi & 3;
                    uint32_t* dst = xq + (size_t)(m >> 3) * xq_words_nt;
                    dst[xq_word(c, j, 0, m & 7)] = u[r].x;
                    dst[xq_word(c, j, 1, m & 7)] = u[r].y;
                    dst[xq_word(c, j, 2, m & 7)] = u[r].z;
                    dst[xq_word(c, j, 3, m & 7)] = u[r].w;
                }
            }
        }
        if (!arrived) asm volatile("bar.arrive 2, %0;" ::"n"(GV_THREADS) : "memory");
        if (ct < p.M) {
            sh_sx[ct] = p.sx_in[layer * p.stride_s + ct];
            sh_sumq[ct] = p.sumq_in[layer * p.stride_s + ct];
        }
    }

    OT* out = reinterpret_cast<OT*>(p.out) + layer * p.stride_out;
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int i = ct + u * GV_CWARPS * 32;
        if (i < p.nrb * GV_ROWS) { sh_scale[i] = sb_scale[u]; sh_bias[i] = sb_bias[u]; }
    }
    for (int i = ct + 2 * GV_CWARPS * 32; i < p.nrb * GV_ROWS; i += GV_CWARPS * 32) {   // > 32 row blocks per CTA
        const int64_t row = ((int64_t)blockIdx.x + (int64_t)(i >> 4) * gridDim.x) * GV_ROWS + (i & 15);
        sh_scale[i] = (p.scale && row < p.N) ? p.scale[layer * p.stride_sb + row] : 1.0f;
        sh_bias[i] = (p.bias && row < p.N) ? p.bias[layer * p.stride_sb + row] : 0.0f;
    }
    consumer_bar();      // activation image, scale/bias and the per-warp sum_q partials are visible
    if (FUSED && ct < p.M) {   // consumed only by the epilogue, which sits behind another barrier
        int tsum = 0;
#pragma unroll
        for (int w = 0; w < GV_CWARPS; ++w) tsum += red_i[ct * GV_CWARPS + w];
        sh_sumq[ct] = tsum;
    }
    if (FUSED) consumer_bar();
    GV_STAMP(1);     // prologue done

    const int grp = warp >> 2, wl = warp & 3, gt = ct & (GV_GWARPS * 32 - 1);
    bool wrote_peer = false;
    int s = 0; uint32_t ph = 0; int buf = 0; int tile_no = 0; int rbi = 0;
    for (int rb = blockIdx.x; rb < p.row_blocks; rb += gridDim.x, ++rbi) {
        if ((rbi & 1) != grp) {                       // the other group's row block: just track the ring
            s += p.nks;
            while (s >= p.stages) { s -= p.stages; ph ^= 1u; }
            continue;
        }
        int acc[NT][4][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[nt][i][q] = 0;

        for (int ks = 0; ks < p.nks; ++ks) {
            mbar_wait(bar_full(s), ph);
            if (tile_no < 4 && grp == 0) GV_STAMP(2 + 2 * tile_no);          // tile arrived
#pragma unroll
            for (int sl2 = 0; sl2 < 2; ++sl2) {
                const int sgi = wl * 2 + sl2;                               // 128-byte segment of the row
                const uint8_t* seg = gbase + s * GV_TILE_BYTES + sgi * (GV_ROWS * GV_SEG);
#pragma unroll
                for (int pp = 0; pp < 2; ++pp) {
                    const int phys = ((2 * t + pp) ^ g) << 4;           // SWIZZLE_128B: chunk ^= row & 7
                    const uint4 wlo = *reinterpret_cast<const uint4*>(seg + g * GV_SEG + phys);
                    const uint4 whi = *reinterpret_cast<const uint4*>(seg + (g + 8) * GV_SEG + phys);
                    const int sgp = (ks * GV_CWARPS + sgi) * 2 + pp;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint32_t mk = 0x03030303u << (2 * i);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            // only the lanes whose token (n = lane>>2) exists load their B fragment: column n of the
                            // MMA feeds output column n only, and the shared-memory wavefronts scale with M, not 8
                            uint4 xv = make_uint4(0u, 0u, 0u, 0u);
                            if (g < p.M - nt * GV_TOK)
                                xv = *reinterpret_cast<const uint4*>(
                                    xq + (size_t)nt * xq_words_nt + ((sgp * 4 + i) * 32 + lane) * 4);
                            imma_16832(acc[nt][i], wlo.x & mk, whi.x & mk, wlo.y & mk, whi.y & mk, xv.x, xv.y);
                            imma_16832(acc[nt][i], wlo.z & mk, whi.z & mk, wlo.w & mk, whi.w & mk, xv.z, xv.w);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty(s));
            if (tile_no < 4 && grp == 0) GV_STAMP(3 + 2 * tile_no);          // tile consumed (warp 0)
            ++tile_no;
            if (++s == p.stages) { s = 0; ph ^= 1u; }
        }

        // recombine fields (exact: set i is a multiple of 4^i) and publish this warp's K-partials
        uint4* gslots = slots + (size_t)(grp * 2 + buf) * NT * GV_GWARPS * 32;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            uint4 v;
            v.x = (uint32_t)(acc[nt][0][0] + (acc[nt][1][0] >> 2) + (acc[nt][2][0] >> 4) + (acc[nt][3][0] >> 6));
            v.y = (uint32_t)(acc[nt][0][1] + (acc[nt][1][1] >> 2) + (acc[nt][2][1] >> 4) + (acc[nt][3][1] >> 6));
            v.z = (uint32_t)(acc[nt][0][2] + (acc[nt][1][2] >> 2) + (acc[nt][2][2] >> 4) + (acc[nt][3][2] >> 6));
            v.w = (uint32_t)(acc[nt][0][3] + (acc[nt][1][3] >> 2) + (acc[nt][2][3] >> 4) + (acc[nt][3][3] >> 6));
            gslots[(nt * GV_GWARPS + wl) * 32 + lane] = v;
        }
        group_bar(grp);
        // reduce over the group's 4 K-quarters and run the fused epilogue: one thread per (row, token)
        const int* sl = reinterpret_cast<const int*>(gslots);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int lq = gt;                                   // lane*4 + q of the accumulator fragment
            int v = 0;
#pragma unroll
            for (int w = 0; w < GV_GWARPS; ++w) v += sl[(nt * GV_GWARPS + w) * 128 + lq];
            const int l = lq >> 2, q = lq & 3;
            const int rl = (l >> 2) + 8 * (q >> 1);
            const int64_t row = (int64_t)rb * GV_ROWS + rl;
            const int tok = nt * GV_TOK + 2 * (l & 3) + (q & 1);
            if (row < p.N && tok < p.M) {
                const int a = v - sh_sumq[tok];
                if (p.acc_out && layer == 0) p.acc_out[(int64_t)tok * p.N + row] = a;
                const float tv = __fmul_rn((float)a, sh_sx[tok]);
                const float y = epilogue(tv, sh_scale[rbi * GV_ROWS + rl], sh_bias[rbi * GV_ROWS + rl], p.clip_val);
                const OT yv = from_f32<OT>(y);
                if (p.n_peers == 0) {
                    out[(int64_t)tok * p.ldo + row] = yv;
                } else {
#pragma unroll 1
                    for (int r = 0; r < p.n_peers; ++r)     // one small store per peer over NVLink (P2P-mapped memory)
                        reinterpret_cast<OT*>(p.peers[r])[(int64_t)tok * p.ldo + row] = yv;
                    wrote_peer = true;
                }
            }
        }
        buf ^= 1;
    }
    if (wrote_peer) __threadfence_system();      // peer stores are visible system-wide before the grid completes
    GV_STAMP(15);
}

// ---- float-activation variant: the reference's exact semantics (no activation quantisation)
template <typename XT, typename OT, int MT>
__global__ void __launch_bounds__(128)
gemv_f32_kernel(const XT* __restrict__ x, int64_t ldx, int M, int64_t K,
                const uint8_t* __restrict__ packed, int64_t ld_packed, int64_t N,
                const float* __restrict__ scale, const float* __restrict__ bias, float clip_val,
                OT* __restrict__ out, int64_t ldo) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t)blockIdx.x * 4 + warp;
    if (row >= N) return;
    const uint8_t* pr = packed + row * ld_packed;
    float acc[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m] = 0.0f;
    const int64_t kb = (K + 3) >> 2;
    for (int64_t b = lane; b < kb; b += 32) {
        const uint32_t byte = pr[b];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t k = 4 * b + i;
            const float wv = (float)((int)((byte >> (2 * i)) & 3u) - 1);
            if (k < K) {
#pragma unroll
                for (int m = 0; m < MT; ++m)
                    if (m < M) acc[m] = fmaf(to_f32(x[(int64_t)m * ldx + k]), wv, acc[m]);
            }
        }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float v = warp_sum(acc[m]);
        if (lane == 0 && m < M) {
            // F.linear output is rounded to x.dtype before the in-place scale/bias (modules.py:104-107)
            float t = to_f32(from_f32<XT>(v));
            out[(int64_t)m * ldo + row] =
                from_f32<OT>(epilogue(t, scale ? scale[row] : 1.0f, bias ? bias[row] : 0.0f, clip_val));
        }
    }
}

// ---- host side ------------------------------------------------------------------------------
// tensor maps are cached per (pointer, geometry): encoding costs ~1 us of host time per call
struct MapKey { const void* ptr; int64_t ld, N, stride; int layers; };
struct MapEntry { MapKey key; CUtensorMap map; bool used; };
static std::mutex g_map_mu;
static MapEntry g_maps[64];
static int g_map_next = 0;

static int get_weight_map(const uint8_t* packed, int64_t ld, int64_t N, int64_t stride, int layers, CUtensorMap* out) {
    std::lock_guard<std::mutex> lk(g_map_mu);
    for (auto& e : g_maps)
        if (e.used && e.key.ptr == packed && e.key.ld == ld && e.key.N == N && e.key.stride == stride && e.key.layers == layers) {
            *out = e.map;
            return TB200_OK;
        }
    uint64_t dims[3] = {(uint64_t)(ld / 4), (uint64_t)N, (uint64_t)layers};
    uint64_t strides[2] = {(uint64_t)ld, (uint64_t)(layers > 1 ? stride : ld * N)};
    uint32_t box[3] = {GV_SEG / 4, GV_ROWS, 1};
    MapEntry& e = g_maps[g_map_next];
    int rc = make_tensor_map(&e.map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, packed, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) { e.used = false; return rc; }
    e.key = MapKey{packed, ld, N, stride, layers};
    e.used = true;
    g_map_next = (g_map_next + 1) % 64;
    *out = e.map;
    return TB200_OK;
}

int gemv_nks(int64_t K) { return (int)((K + 4095) / 4096); }

static size_t gemv_smem(int nt, int nks, int stages, int nrb = 8) {
    return 1024 + (size_t)stages * GV_TILE_BYTES + (size_t)nt * GV_TOK * nks * 4096 + (size_t)2 * nt * GV_CWARPS * 32 * 16 +
           (size_t)nt * GV_TOK * 8 + GV_TOK * GV_CWARPS * 8 + 2 * GV_MAX_STAGES * 8 + (size_t)nrb * GV_ROWS * 8 + 64;
}
// deepest ring (<= 6, >= 2 stages) that fits next to the activation image; 0 if nothing fits
static int gemv_pick_stages(int nt, int nks, int nrb = 8) {
    // two CTAs per SM (this layer's and, under programmatic dependent launch, the next layer's)
    // when a >= 3-deep ring still fits in half of the SM's shared memory ...
    for (int s = GV_MAX_STAGES - 1; s >= 3; --s)
        if (gemv_smem(nt, nks, s, nrb) <= 113 * 1024) return s;
    // ... else the deepest ring that fits at one CTA per SM
    for (int s = GV_MAX_STAGES - 1; s >= 2; --s)
        if (gemv_smem(nt, nks, s, nrb) <= 227 * 1024) return s;
    return 0;
}

template <typename XT, typename OT, int NT, bool FUSED>
static int launch_gemv(const CUtensorMap& map, GemvParams p, int layers, cudaStream_t s) {
    const int ctas = p.row_blocks < sm_count() ? p.row_blocks : sm_count();
    p.nrb = (p.row_blocks + ctas - 1) / ctas;
    p.stages = gemv_pick_stages(NT, p.nks, p.nrb);
    TB_REQUIRE(p.stages >= 2, "gemv: K=%lld N=%lld is too large for the shared-memory activation image", (long long)p.K, (long long)p.N);
    const size_t smem = gemv_smem(NT, p.nks, p.stages, p.nrb);
    auto kern = gemv_mma_kernel<XT, OT, NT, FUSED>;
    static size_t configured = 0;   // per instantiation
    if (smem > configured) {
        TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas, (unsigned)layers, 1);
    cfg.blockDim = dim3(GV_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, map, p));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

template <int NT, bool FUSED>
static int gemv_dispatch(int x_dtype, int out_dtype, const CUtensorMap& map, const GemvParams& p, int layers, cudaStream_t s) {
#define TB_GEMV_CASE(XD, XT, OD, OT) \
    if (x_dtype == XD && out_dtype == OD) return launch_gemv<XT, OT, NT, FUSED>(map, p, layers, s)
    TB_GEMV_CASE(TB200_BF16, __nv_bfloat16, TB200_BF16, __nv_bfloat16);
    TB_GEMV_CASE(TB200_BF16, __nv_bfloat16, TB200_F32, float);
    TB_GEMV_CASE(TB200_F32, float, TB200_F32, float);
    TB_GEMV_CASE(TB200_F16, __half, TB200_F16, __half);
    TB_GEMV_CASE(TB200_F16, __half, TB200_F32, float);
#undef TB_GEMV_CASE
    set_error("gemv: unsupported dtype pair x=%d out=%d (out must be the x dtype or f32)", x_dtype, out_dtype);
    return TB200_ENOTSUP;
}

static int gemv_check(int64_t M, int64_t K, int64_t N, int64_t ldo, const uint8_t* packed, int64_t ld_packed,
                      int64_t stride_packed, int layers) {
    TB_REQUIRE(M >= 0 && M <= TB200_GEMV_MAX_M, "gemv: M=%lld out of range [0,%d]", (long long)M, TB200_GEMV_MAX_M);
    TB_REQUIRE(K > 0 && N >= 0 && ldo >= N, "gemv: bad shape M=%lld K=%lld N=%lld", (long long)M, (long long)K, (long long)N);
    const int64_t kc = (K + 63) / 64;
    TB_REQUIRE(ld_packed % 16 == 0 && ld_packed >= kc * 16,
               "gemv: ld_packed=%lld must be a multiple of 16 and >= 16*ceil(K/64)=%lld", (long long)ld_packed, (long long)(kc * 16));
    TB_REQUIRE((uintptr_t)packed % 16 == 0 && stride_packed % 16 == 0, "gemv: packed weights must be 16-byte aligned");
    TB_REQUIRE(layers >= 1 && layers <= 65535, "gemv: bad group count %d", layers);
    return TB200_OK;
}

static void fill_common(GemvParams& p, int64_t K, int64_t N, const float* scale, const float* bias,
                        int64_t stride_sb, float clip_val, int64_t ldo, int64_t stride_out, int flags) {
    p.scale = scale; p.bias = bias; p.stride_sb = stride_sb;
    p.ldo = ldo; p.stride_out = stride_out;
    p.N = N; p.K = K; p.nks = gemv_nks(K);
    p.row_blocks = (int)((N + GV_ROWS - 1) / GV_ROWS);
    p.clip_val = fabsf(clip_val); p.pdl = flags & 1;
    p.trace = g_trace;
}

// fused-quantisation entry (raw activations), <= 8 tokens per pass over the weights
static void set_peers(GemvParams& p, void* const* peers, int n_peers, int64_t row_offset_bytes) {
    p.n_peers = peers ? n_peers : 0;
    for (int r = 0; r < p.n_peers; ++r) p.peers[r] = (char*)peers[r] + row_offset_bytes;
}

int gemv_fused(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                      const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                      const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                      void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                      int32_t* acc_out, float* sx_out, int8_t* qx_out, int layers, int flags, cudaStream_t s,
                      void* const* peers = nullptr, int n_peers = 0) {
    CUtensorMap map;
    int rc = get_weight_map(packed, ld_packed, N, stride_packed, layers, &map);
    if (rc) return rc;
    const size_t xs = dtype_size(x_dtype), os = dtype_size(out_dtype);
    GemvParams p = {};
    fill_common(p, K, N, scale, bias, stride_sb, clip_val, ldo, stride_out, flags);
    p.ldx = ldx; p.stride_x = stride_x;
    p.xvec = ((uintptr_t)x % 16 == 0) && ((ldx * xs) % 16 == 0) && ((stride_x * xs) % 16 == 0);
    for (int64_t m0 = 0; m0 < M; m0 += GV_TOK) {
        p.M = (int)(M - m0 < GV_TOK ? M - m0 : GV_TOK);
        p.x = (const char*)x + m0 * ldx * xs;
        p.out = (char*)out + m0 * ldo * os;
        p.acc_out = acc_out ? acc_out + m0 * N : nullptr;
        p.sx_out = sx_out ? sx_out + m0 : nullptr;
        p.qx_out = qx_out ? qx_out + m0 * K : nullptr;
        set_peers(p, peers, n_peers, m0 * ldo * (int64_t)os);
        rc = gemv_dispatch<1, true>(x_dtype, out_dtype, map, p, layers, s);
        if (rc) return rc;
    }
    return TB200_OK;
}

}  // namespace tb200

using namespace tb200;

extern "C" void tb200_debug_set_trace(long long* dev_buf) { g_trace = dev_buf; }

namespace tb200 {
int gemv_q_impl(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx,
                            int64_t M, int64_t K,
                            const uint8_t* packed, int64_t ld_packed, int64_t N,
                            const float* scale, const float* bias, float clip_val,
                            void* out, int out_dtype, int64_t ldo, int32_t* acc_out, int flags, void* stream,
                void* const* peers, int n_peers) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(q_x && s_x && sum_qx && packed && out, "gemv_q: null pointer");
    int rc = gemv_check(M, K, N, ldo, packed, ld_packed, 0, 1);
    if (rc) return rc;
    TB_REQUIRE(ldq >= K && ldq % 16 == 0 && (uintptr_t)q_x % 16 == 0,
               "gemv_q: q_x must come from tb200_quant_act_i8(layout 1): ldq >= K, multiple of 16, 16-byte aligned");
    TB_REQUIRE(dtype_ok(out_dtype), "gemv_q: bad out dtype");
    CUtensorMap map;
    rc = get_weight_map(packed, ld_packed, N, 0, 1, &map);
    if (rc) return rc;
    const size_t os = dtype_size(out_dtype);
    GemvParams p = {};
    fill_common(p, K, N, scale, bias, 0, clip_val, ldo, 0, flags);
    p.ldq = ldq;
    // 16 tokens per pass when the activation image fits next to a >= 4-deep ring, else 8
    const bool two = M > GV_TOK && gemv_pick_stages(2, p.nks) >= 4;
    const int64_t step = two ? 2 * GV_TOK : GV_TOK;
    for (int64_t m0 = 0; m0 < M; m0 += step) {
        p.M = (int)(M - m0 < step ? M - m0 : step);
        p.qx = q_x + m0 * ldq;
        p.sx_in = s_x + m0; p.sumq_in = sum_qx + m0;
        p.out = (char*)out + m0 * ldo * os;
        p.acc_out = acc_out ? acc_out + m0 * N : nullptr;
        set_peers(p, peers, n_peers, m0 * ldo * (int64_t)os);
        // activations are int8 already: the XT template parameter is unused on this path
        if (two && p.M > GV_TOK) rc = gemv_dispatch<2, false>(out_dtype, out_dtype, map, p, 1, (cudaStream_t)stream);
        else                     rc = gemv_dispatch<1, false>(out_dtype, out_dtype, map, p, 1, (cudaStream_t)stream);
        if (rc) return rc;
    }
    return TB200_OK;
}
}  // namespace tb200

extern "C" int tb200_gemv_q(const int8_t* q_x, int64_t ldq, const float* s_x, const int32_t* sum_qx,
                            int64_t M, int64_t K,
                            const uint8_t* packed, int64_t ld_packed, int64_t N,
                            const float* scale, const float* bias, float clip_val,
                            void* out, int out_dtype, int64_t ldo, int32_t* acc_out, int flags, void* stream) {
    return gemv_q_impl(q_x, ldq, s_x, sum_qx, M, K, packed, ld_packed, N, scale, bias, clip_val, out, out_dtype, ldo,
                       acc_out, flags, stream, nullptr, 0);
}

extern "C" int tb200_gemv(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                          const uint8_t* packed, int64_t ld_packed, int64_t N,
                          const float* scale, const float* bias, float clip_val,
                          void* out, int out_dtype, int64_t ldo,
                          int32_t* acc_out, float* sx_out, int8_t* qx_out, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out, "gemv: null pointer");
    TB_REQUIRE(ldx >= K, "gemv: ldx < K");
    int rc = gemv_check(M, K, N, ldo, packed, ld_packed, 0, 1);
    if (rc) return rc;
    return gemv_fused(x, x_dtype, M, K, ldx, 0, packed, ld_packed, 0, N, scale, bias, 0, clip_val,
                      out, out_dtype, ldo, 0, acc_out, sx_out, qx_out, 1, flags, (cudaStream_t)stream);
}

extern "C" int tb200_gemv_grouped(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                                  const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                                  const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                  void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                                  int count, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0 || count == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out, "gemv_grouped: null pointer");
    TB_REQUIRE(ldx >= K, "gemv_grouped: ldx < K");
    int rc = gemv_check(M, K, N, ldo, packed, ld_packed, stride_packed, count);
    if (rc) return rc;
    return gemv_fused(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N, scale, bias,
                      stride_sb, clip_val, out, out_dtype, ldo, stride_out, nullptr, nullptr, nullptr,
                      count, flags, (cudaStream_t)stream);
}

extern "C" int tb200_gemv_f32(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx,
                              const uint8_t* packed, int64_t ld_packed, int64_t N,
                              const float* scale, const float* bias, float clip_val,
                              void* out, int out_dtype, int64_t ldo, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out, "gemv_f32: null pointer");
    TB_REQUIRE(M >= 0 && M <= TB200_GEMV_MAX_M, "gemv_f32: M=%lld out of range", (long long)M);
    TB_REQUIRE(K > 0 && N >= 0 && ldx >= K && ldo >= N && ld_packed >= (K + 3) / 4, "gemv_f32: bad shape");
    TB_REQUIRE(x_dtype == out_dtype, "gemv_f32: out dtype must equal x dtype");
    const size_t es = dtype_size(x_dtype);
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((unsigned)((N + 3) / 4));
    for (int64_t m0 = 0; m0 < M; m0 += 8) {
        const int mt = (int)(M - m0 < 8 ? M - m0 : 8);
        const char* xp = (const char*)x + m0 * ldx * es;
        char* op = (char*)out + m0 * ldo * es;
        TB_DISPATCH_DTYPE(x_dtype, T,
            (gemv_f32_kernel<T, T, 8><<<grid, 128, 0, s>>>((const T*)xp, ldx, mt, K, packed, ld_packed, N,
                                                          scale, bias, fabsf(clip_val), (T*)op, ldo)));
        TB_LAUNCH_CHECK();
    }
    return TB200_OK;
}
