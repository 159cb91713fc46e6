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
#include <cstdlib>
#include "tma.cuh"

namespace tb200 {

constexpr int GV_ROWS = 16;                       // rows per tile = MMA M
constexpr int GV_SLICE = 1024;                    // packed bytes per row per tile (4096 weights)
constexpr int GV_SEG = 128;                       // bytes per row per TMA box / per consumer warp
constexpr int GV_TILE_BYTES = GV_ROWS * GV_SLICE; // 16 KB
constexpr int GV_CWARPS = GV_SLICE / GV_SEG;      // 8 consumer warps, one per 128-byte K segment
constexpr int GV_CTHREADS = GV_CWARPS * 32;
constexpr int GV_WARP_PRODUCER = GV_CWARPS;       // warp 8: TMA producer
constexpr int GV_THREADS = (GV_CWARPS + 1) * 32;
constexpr int GV_MAX_STAGES = 12;
constexpr int GV_MAX_SLOTS = 8;                   // row blocks whose K-partials (4 KB each) are parked before a reduction round
constexpr int GV_TOK = 8;                         // tokens per pass = MMA N

struct GemvParams {
    const void* x; int64_t ldx; int64_t stride_x;            // raw activations (fused quantisation)
    const int8_t* qx; int64_t ldq; int64_t stride_q;         // or pre-quantised, layout 1 (kperm)
    const float* sx_in; const int32_t* sumq_in; int64_t stride_s;
    const float* scale; const float* bias; int64_t stride_sb;
    void* out; int64_t ldo; int64_t stride_out;
    int32_t* acc_out;
    int64_t N, K;
    int M;            // tokens in this launch (<= 8)
    int nks;          // k-slices of 4096 weights
    int row_blocks;   // ceil(N / 16)
    int nrb;          // row blocks per CTA = ceil(row_blocks / gridDim.x)
    int stages;       // ring depth
    int nslots;       // row blocks per reduction round (partial-sum slots in shared memory)
    float clip_val;
    int pdl;
    int one_box;      // the tensor map is 4-D (word, row, segment, layer): one 16 KB box per tile
    bool xvec;
    long long* trace;   // debug: per-CTA globaltimer stamps (tb200_debug_set_trace), normally null
    // tensor-parallel gather fused into the epilogue: when n_peers > 0 the output element is stored into every
    // rank's [M, ldo] buffer (peer-mapped NVLink addresses, already offset to this rank's column block)
    void* peers[8]; int n_peers;
};
static long long* g_trace = nullptr;
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define GV_STAMP(slot) do { if (p.trace && blockIdx.y == 0) { if ((slot) == 0 || (slot) == 15 || (slot) == 1) p.trace[blockIdx.x * 64 + (slot)] = (long long)gtimer(); p.trace[blockIdx.x * 64 + 16 + (slot)] = clock64(); } } while (0)
#define GV_CLK(slot) do { if (p.trace && blockIdx.y == 0) p.trace[blockIdx.x * 64 + 32 + (slot)] = clock64(); } while (0)

// q = clamp(rint(v * inv), +-127): the activation quantiser (ternary_b200.h, tb200_quant_act_i8)
// rint() and the float->int conversion are done by adding 1.5*2^23 (the FADD rounds to nearest-even
// integer, the low mantissa bits are the result): identical bits, but no XU-pipe conversions.
__device__ __forceinline__ int quant_mul(float v, float inv) {
    float r = __fmul_rn(v, inv);
    r = fminf(fmaxf(r, -127.0f), 127.0f);
    return __float_as_int(__fadd_rn(r, 12582912.0f)) - 0x4B400000;
}

// 16 consecutive activations of one token in their storage type (zero beyond K): NV 16-byte vectors.
// VEC: rows are 16-byte aligned and K % 16 == 0 (the hot instantiation: vector loads only, small code).
template <typename XT> struct Raw16 { static constexpr int NV = sizeof(XT); uint4 v[NV]; };   // 2 (16-bit) or 4 (fp32) vectors
template <typename XT, bool VEC>
__device__ __forceinline__ void load_raw16(const XT* __restrict__ xr, int64_t g, int64_t K, Raw16<XT>& r) {
    if (VEC) {
        if (16 * g < K) {
            const uint4* q = reinterpret_cast<const uint4*>(xr + 16 * g);
#pragma unroll
            for (int j = 0; j < Raw16<XT>::NV; ++j) r.v[j] = q[j];
        } else {
#pragma unroll
            for (int j = 0; j < Raw16<XT>::NV; ++j) r.v[j] = make_uint4(0u, 0u, 0u, 0u);
        }
    } else {
        XT* e = reinterpret_cast<XT*>(r.v);
#pragma unroll 1
        for (int t = 0; t < 16; ++t) {
            const int64_t k = 16 * g + t;
            e[t] = k < K ? xr[k] : from_f32<XT>(0.0f);
        }
    }
}
template <typename XT>
__device__ __forceinline__ float raw_get(const Raw16<XT>& r, int e) { return to_f32(reinterpret_cast<const XT*>(r.v)[e]); }

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GV_CTHREADS) : "memory"); }

// D(16x8,s32) += A(16x32,u8 row) * B(32x8,s8 col)
__device__ __forceinline__ void imma_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                           uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Shared-memory image of the quantised activations of the M tokens of this pass, in 32-bit words.  For
// 16-byte chunk c of the packed row (64 weights; c counts across k-slices), packed word j (16 weights),
// field i and token n, the word holds bytes b=0..3 = q[n][64c + 16j + 4b + i]: the four activations that
// meet field i of the four bytes of packed word j.  The words j=0..3 form one uint4, which is exactly
// the B fragments {b0,b1} of two m16n8k32 MMAs for lane (g = n, t) with chunk c = 8*segment + 2t + p.
// uint4 index = (((segment*2 + p)*4 + i) * M + n) * 4 + t.
__device__ __forceinline__ int xq_vec(int sg, int pp, int i, int n, int t, int M) {
    return ((((sg * 2 + pp) * 4 + i) * M + n) << 2) + t;
}

template <typename XT, typename OT, bool FUSED, bool VEC>
__global__ void __launch_bounds__(GV_THREADS, 2)
gemv_mma_kernel(const __grid_constant__ CUtensorMap map_w, const GemvParams p) {
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;         // SWIZZLE_128B boxes: 1024-B aligned
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    // layout: ring[stages][16 KB] | image[nks*4096*M] | slots[nslots][8 warps][32] uint4 | sx | sumq | red_f | red_i | barriers | scale | bias
    const uint32_t sRing = base;
    uint32_t* xq = reinterpret_cast<uint32_t*>(gbase + p.stages * GV_TILE_BYTES);
    uint4* slots = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(xq) + (size_t)p.nks * 4096 * p.M);
    float* sh_sx = reinterpret_cast<float*>(slots + (size_t)p.nslots * GV_CTHREADS);
    int* sh_sumq = reinterpret_cast<int*>(sh_sx + GV_TOK);
    float* red_f = reinterpret_cast<float*>(sh_sumq + GV_TOK);             // [8 tokens][8 warps]
    int* red_i = reinterpret_cast<int*>(red_f + GV_TOK * GV_CWARPS);      // [8 tokens][8 warps]
    uint64_t* bars = reinterpret_cast<uint64_t*>(red_i + GV_TOK * GV_CWARPS);
    float* sh_scale = reinterpret_cast<float*>(bars + 2 * GV_MAX_STAGES);   // [nrb][16] this CTA's rows
    float* sh_bias = sh_scale + p.nrb * GV_ROWS;
    const uint32_t sBar = smem_u32(bars);
    auto bar_full = [&](int s) { return sBar + 8u * s; };
    auto bar_empty = [&](int s) { return sBar + 8u * (GV_MAX_STAGES + s); };

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int layer = blockIdx.y;
    const int M = p.M;
    if (threadIdx.x == 0) {
        GV_STAMP(0);
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(bar_full(s), 1);
            mbar_init(bar_empty(s), GV_CWARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");

    if (warp == GV_WARP_PRODUCER) {
        // ===================== producer: weights do not depend on the previous kernel =====================
        // Without PDL the activations are requested at the same time as the weights and would queue
        // behind a ring-full of weight traffic: let the consumers issue their x loads first.
        if (!p.pdl) asm volatile("bar.sync 2, %0;" ::"n"(GV_CTHREADS + 32) : "memory");
        if (lane == 0) {
            int s = 0; uint32_t ph = 0; int idx = 0;
            for (int rb = blockIdx.x; rb < p.row_blocks; rb += gridDim.x) {
                for (int ks = 0; ks < p.nks; ++ks, ++idx) {
                    if (idx >= p.stages) mbar_wait(bar_empty(s), ph ^ 1u);     // first pass over the ring: stages are free
                    mbar_arrive_expect_tx(bar_full(s), GV_TILE_BYTES);
                    if (p.one_box) {
                        tma_load_4d(sRing + s * GV_TILE_BYTES, &map_w, bar_full(s), 0, rb * GV_ROWS, ks * GV_CWARPS, layer);
                    } else {
#pragma unroll
                        for (int sg = 0; sg < GV_CWARPS; ++sg)
                            tma_load_3d(sRing + s * GV_TILE_BYTES + sg * (GV_ROWS * GV_SEG), &map_w, bar_full(s),
                                        ks * (GV_SLICE / 4) + sg * (GV_SEG / 4), rb * GV_ROWS, layer);
                    }
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
        return;
    }

    // ===================== consumers =====================
    const int ct = threadIdx.x;                       // 0..255
    const int g = lane >> 2, t = lane & 3;
    // this CTA's per-row scale and bias (weights: independent of the previous kernel): the loads are issued
    // now, the values are parked in shared memory once the activations have been requested
    float sb_scale[2], sb_bias[2];
    {
        const float* scale_g = p.scale ? p.scale + layer * p.stride_sb : nullptr;
        const float* bias_g = p.bias ? p.bias + layer * p.stride_sb : nullptr;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int i = ct + u * GV_CTHREADS;
            const int64_t row = ((int64_t)blockIdx.x + (int64_t)(i >> 4) * gridDim.x) * GV_ROWS + (i & 15);
            const bool ok = i < p.nrb * GV_ROWS && row < p.N;
            sb_scale[u] = (scale_g && ok) ? scale_g[row] : 1.0f;
            sb_bias[u] = (bias_g && ok) ? bias_g[row] : 0.0f;
        }
    }
    // x comes from the previous kernel
    if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (ct == 0) GV_STAMP(1);
    bool released = p.pdl != 0;                       // producer already running under PDL
    auto release_producer = [&]() {
        if (!released) { asm volatile("bar.arrive 2, %0;" ::"n"(GV_CTHREADS + 32) : "memory"); released = true; }
    };

    int lsum_tok[GV_TOK];                             // this thread's sum of codes per token (reduced after the image barrier)
#pragma unroll
    for (int m = 0; m < GV_TOK; ++m) lsum_tok[m] = 0;
    if (FUSED) {
        const XT* x = reinterpret_cast<const XT*>(p.x) + layer * p.stride_x;
        constexpr int TB = sizeof(XT) == 4 ? 2 : 4;           // tokens whose loads are in flight together (32 registers)
#pragma unroll
        for (int mb = 0; mb < GV_TOK / TB; ++mb) {
            const int m0 = mb * TB;
            if (m0 >= M) break;
            Raw16<XT> raw[TB];
            float amax[TB];
#pragma unroll
            for (int tb = 0; tb < TB; ++tb) amax[tb] = 0.0f;
            // pass A: row maxima.  Thread ct owns 16-element group ct of every k-slice; the last slice stays in registers.
#pragma unroll 1
            for (int u = 0; u < p.nks; ++u) {
#pragma unroll
                for (int tb = 0; tb < TB; ++tb)
                    if (m0 + tb < M) load_raw16<XT, VEC>(x + (int64_t)(m0 + tb) * p.ldx, u * GV_CTHREADS + ct, p.K, raw[tb]);
                release_producer();
#pragma unroll
                for (int tb = 0; tb < TB; ++tb)
                    if (m0 + tb < M) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) amax[tb] = fmaxf(amax[tb], fabsf(raw_get<XT>(raw[tb], e)));
                    }
            }
#pragma unroll
            for (int tb = 0; tb < TB; ++tb) {
                if (m0 + tb < M) {
                    // non-negative floats order like their bit patterns: one REDUX instead of five shuffles
                    const uint32_t a = __reduce_max_sync(0xffffffffu, __float_as_uint(amax[tb]));
                    if (lane == 0) red_f[(m0 + tb) * GV_CWARPS + warp] = __uint_as_float(a);
                }
            }
            consumer_bar();
            float inv[TB];
#pragma unroll
            for (int tb = 0; tb < TB; ++tb) {
                inv[tb] = 1.0f;
                if (m0 + tb < M) {
                    const int m = m0 + tb;
                    const float4 r0 = *reinterpret_cast<const float4*>(red_f + m * GV_CWARPS);
                    const float4 r1 = *reinterpret_cast<const float4*>(red_f + m * GV_CWARPS + 4);
                    const float v = fmaxf(fmaxf(fmaxf(r0.x, r0.y), fmaxf(r0.z, r0.w)), fmaxf(fmaxf(r1.x, r1.y), fmaxf(r1.z, r1.w)));
                    inv[tb] = v > 0.0f ? __fdiv_rn(127.0f, v) : 1.0f;
                    if (ct == 0) sh_sx[m] = v > 0.0f ? __fdiv_rn(v, 127.0f) : 1.0f;
                }
            }
            // pass B: quantise, last slice first (its values are still in registers).  |x*inv| < 127.5 by construction
            // (|x| <= amax, two roundings), so the clamp of the definition never binds and is not evaluated.
#pragma unroll 1
            for (int u = p.nks - 1; u >= 0; --u) {
                const int gi = u * GV_CTHREADS + ct;
                const int c = gi >> 2, j = gi & 3, lc = c & 7;
#pragma unroll
                for (int tb = 0; tb < TB; ++tb) {
                    if (m0 + tb < M) {
                        if (u != p.nks - 1) load_raw16<XT, VEC>(x + (int64_t)(m0 + tb) * p.ldx, gi, p.K, raw[tb]);
                        uint32_t qb[16];      // low byte = two's-complement code (rint via the 1.5*2^23 trick)
#pragma unroll
                        for (int e = 0; e < 16; ++e)
                            qb[e] = __float_as_uint(__fadd_rn(__fmul_rn(raw_get<XT>(raw[tb], e), inv[tb]), 12582912.0f));
                        int ls = lsum_tok[m0 + tb];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {  // word i = elements i, 4+i, 8+i, 12+i (field i of bytes 0..3)
                            const uint32_t w = __byte_perm(__byte_perm(qb[i], qb[4 + i], 0x0040), __byte_perm(qb[8 + i], qb[12 + i], 0x0040), 0x5410);
                            ls = __dp4a((int)w, 0x01010101, ls);
                            xq[xq_vec(c >> 3, lc & 1, i, m0 + tb, lc >> 1, M) * 4 + j] = w;
                        }
                        lsum_tok[m0 + tb] = ls;
                    }
                }
            }
        }
    } else {
        // pre-quantised activations (tb200_quant_act_i8 layout 1): rows of kperm words
        const int8_t* q = p.qx + layer * p.stride_q;
        const int groups = p.nks * 256, have = (int)(p.ldq >> 4);   // zero padding up to the slice
        const int total = M * groups;
        for (int base_i = 0; base_i < total; base_i += 8 * GV_CTHREADS) {
            uint4 u[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int idx = base_i + r * GV_CTHREADS + ct;
                const int m = idx / groups, gi = idx - m * groups;
                u[r] = (idx < total && gi < have) ? *reinterpret_cast<const uint4*>(q + (int64_t)m * p.ldq + 16 * (int64_t)gi)
                                                   : make_uint4(0u, 0u, 0u, 0u);
            }
            release_producer();
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int idx = base_i + r * GV_CTHREADS + ct;
                if (idx < total) {
                    const int m = idx / groups, gi = idx - m * groups;
                    const int c = gi >> 2, j = gi & 3, lc = c & 7;
                    const uint32_t w4[4] = {u[r].x, u[r].y, u[r].z, u[r].w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) xq[xq_vec(c >> 3, lc & 1, i, m, lc >> 1, M) * 4 + j] = w4[i];
                }
            }
        }
    }
    release_producer();
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int i = ct + u * GV_CTHREADS;
        if (i < p.nrb * GV_ROWS) { sh_scale[i] = sb_scale[u]; sh_bias[i] = sb_bias[u]; }
    }
    for (int i = ct + 2 * GV_CTHREADS; i < p.nrb * GV_ROWS; i += GV_CTHREADS) {   // > 32 row blocks per CTA
        const int64_t row = ((int64_t)blockIdx.x + (int64_t)(i >> 4) * gridDim.x) * GV_ROWS + (i & 15);
        sh_scale[i] = (p.scale && row < p.N) ? p.scale[layer * p.stride_sb + row] : 1.0f;
        sh_bias[i] = (p.bias && row < p.N) ? p.bias[layer * p.stride_sb + row] : 0.0f;
    }
    consumer_bar();               // activation image (and sh_sx) complete
    if (ct == 0) GV_STAMP(3);

    // B fragments of this warp's K segment: 32 registers, loaded once when K <= 4096
    const uint4* xqv = reinterpret_cast<const uint4*>(xq);
    uint4 bq[2][4];
    auto load_b = [&](int ks) {
#pragma unroll
        for (int pp = 0; pp < 2; ++pp)
#pragma unroll
            for (int i = 0; i < 4; ++i)
                bq[pp][i] = g < M ? xqv[xq_vec(ks * GV_CWARPS + warp, pp, i, g, t, M)] : make_uint4(0u, 0u, 0u, 0u);
    };
    load_b(0);
    if (FUSED) {                  // per-token sum of codes: needed by the epilogue only, off the critical path
#pragma unroll
        for (int m = 0; m < GV_TOK; ++m)
            if (m < M) {
                const int v = __reduce_add_sync(0xffffffffu, lsum_tok[m]);
                if (lane == 0) red_i[m * GV_CWARPS + warp] = v;
            }
    }

    // Main loop, software-pipelined: the packed rows of tile i+1 are fetched into registers (and its ring stage
    // released) between the two halves of the MMAs of tile i.
    const uint32_t sSeg = warp * (GV_ROWS * GV_SEG) + g * GV_SEG;
    const uint32_t phys0 = ((2 * t) ^ g) << 4, phys1 = ((2 * t + 1) ^ g) << 4;   // SWIZZLE_128B: chunk ^= row & 7
    // loop-invariant launch parameters pinned in registers (otherwise every use is a constant-bank load on the
    // critical path of a loop that runs only a handful of times)
    auto pin = [](int v) { int r; asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v)); return r; };
    const int nks_r = pin(p.nks), stages_r = pin(p.stages), nslots_r = pin(p.nslots);
    const int my_rbs = pin((p.row_blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x);
    const int ntiles = pin(nks_r * my_rbs);
    const bool recycle = ntiles > stages_r;           // stages are handed back to the producer only when the ring wraps
    int s = 0; uint32_t ph = 0;
    uint4 wlo[2], whi[2], nlo[2], nhi[2];
    // shared-memory addresses of the loop as 32-bit values the compiler cannot rematerialise: left to itself it rebuilt the
    // ring and barrier bases from SR_CgaCtaId (S2UR + a chain of uniform-datapath instructions) on every tile
    uint32_t a_seg0 = base + sSeg + phys0, a_seg1 = base + sSeg + phys1, a_bar = sBar;
    asm volatile("" : "+r"(a_seg0), "+r"(a_seg1), "+r"(a_bar));
    auto lds128 = [](uint32_t a) {
        uint4 v;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
        return v;
    };
    auto fetch = [&](uint4 (&lo)[2], uint4 (&hi)[2]) {
        mbar_wait(a_bar + 8u * s, ph);
        const uint32_t o = (uint32_t)s * GV_TILE_BYTES;
        lo[0] = lds128(a_seg0 + o);
        hi[0] = lds128(a_seg0 + o + 8 * GV_SEG);
        lo[1] = lds128(a_seg1 + o);
        hi[1] = lds128(a_seg1 + o + 8 * GV_SEG);
        if (recycle) {
            __syncwarp();
            if (lane == 0) mbar_arrive(a_bar + 8u * (GV_MAX_STAGES + s));            // the tile lives in registers now
        }
        if (++s == stages_r) { s = 0; ph ^= 1u; }
    };
    auto mma_half = [&](int (&acc)[4][4], const uint4& lo, const uint4& hi, const uint4 (&b)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t mk = 0x03030303u << (2 * i);
            imma_16832(acc[i], lo.x & mk, hi.x & mk, lo.y & mk, hi.y & mk, b[i].x, b[i].y);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t mk = 0x03030303u << (2 * i);
            imma_16832(acc[i], lo.z & mk, hi.z & mk, lo.w & mk, hi.w & mk, b[i].z, b[i].w);
        }
    };
    // Reduction round: the K-partials of up to nslots row blocks are parked in shared memory; after a barrier warp w
    // sums the eight partials of row blocks w, w+8, ... of the round and runs the fused epilogue for them.
    OT* out = reinterpret_cast<OT*>(p.out) + layer * p.stride_out;
    bool wrote_peer = false;
    auto reduce_round = [&](int rbi0, int count) {
        consumer_bar();
        int sumq[2]; float sxv[2];                // this lane's two tokens (2t, 2t+1)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int tok = 2 * t + e;
            sumq[e] = 0; sxv[e] = 0.0f;
            if (tok < M) {
                sxv[e] = sh_sx[tok];
                if (FUSED) {
                    const int4 r0 = *reinterpret_cast<const int4*>(red_i + tok * GV_CWARPS);
                    const int4 r1 = *reinterpret_cast<const int4*>(red_i + tok * GV_CWARPS + 4);
                    sumq[e] = r0.x + r0.y + r0.z + r0.w + r1.x + r1.y + r1.z + r1.w;
                } else {
                    sumq[e] = sh_sumq[tok];
                }
            }
        }
        for (int j = warp; j < count; j += GV_CWARPS) {
            const uint4* sl = slots + (size_t)j * GV_CTHREADS + lane;
            int vv[4] = {0, 0, 0, 0};
#pragma unroll
            for (int w = 0; w < GV_CWARPS; ++w) {
                const uint4 u = sl[w * 32];
                vv[0] += (int)u.x; vv[1] += (int)u.y; vv[2] += (int)u.z; vv[3] += (int)u.w;
            }
            // accumulator fragment: vv0 (row g, tok 2t), vv1 (row g, tok 2t+1), vv2 (row g+8, tok 2t), vv3 (row g+8, tok 2t+1)
            const int rbi = rbi0 + j;
            const int64_t rb = (int64_t)blockIdx.x + (int64_t)rbi * gridDim.x;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int rl = g + 8 * (q >> 1), tok = 2 * t + (q & 1);
                const int64_t row = rb * GV_ROWS + rl;
                if (row < p.N && tok < M) {
                    const int a = vv[q] - sumq[q & 1];
                    if (p.acc_out && layer == 0) p.acc_out[(int64_t)tok * p.N + row] = a;
                    const float tv = __fmul_rn((float)a, sxv[q & 1]);
                    const float y = epilogue(tv, sh_scale[rbi * GV_ROWS + rl], sh_bias[rbi * GV_ROWS + rl], p.clip_val);
                    const OT yv = from_f32<OT>(y);
                    if (p.n_peers == 0) {
                        out[(int64_t)tok * p.ldo + row] = yv;
                    } else {
#pragma unroll
                        for (int r = 0; r < 8; ++r)             // one small store per peer over NVLink (P2P-mapped memory);
                            if (r < p.n_peers)                  // constant indices keep the parameter block out of local memory
                                reinterpret_cast<OT*>(p.peers[r])[(int64_t)tok * p.ldo + row] = yv;
                        wrote_peer = true;
                    }
                }
            }
        }
    };
    if (!FUSED && ct < M) {       // pre-quantised activations: scale and sum of codes come from the quantiser kernel
        sh_sx[ct] = p.sx_in[layer * p.stride_s + ct];
        sh_sumq[ct] = p.sumq_in[layer * p.stride_s + ct];
    }

    int acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[i][q] = 0;
    if (ntiles > 0) fetch(wlo, whi);
    int ks = 0, rbi = 0, round0 = 0;
#pragma unroll 1
    for (int it = 0; it < ntiles; ++it) {
        mma_half(acc, wlo[0], whi[0], bq[0]);
        if (it + 1 < ntiles) fetch(nlo, nhi);
        mma_half(acc, wlo[1], whi[1], bq[1]);
#pragma unroll
        for (int pp = 0; pp < 2; ++pp) { wlo[pp] = nlo[pp]; whi[pp] = nhi[pp]; }
        if (++ks == nks_r) {
            // recombine fields (exact: set i is a multiple of 4^i) and park this warp's K-partials
            uint4 v;
            v.x = (uint32_t)(acc[0][0] + (acc[1][0] >> 2) + (acc[2][0] >> 4) + (acc[3][0] >> 6));
            v.y = (uint32_t)(acc[0][1] + (acc[1][1] >> 2) + (acc[2][1] >> 4) + (acc[3][1] >> 6));
            v.z = (uint32_t)(acc[0][2] + (acc[1][2] >> 2) + (acc[2][2] >> 4) + (acc[3][2] >> 6));
            v.w = (uint32_t)(acc[0][3] + (acc[1][3] >> 2) + (acc[2][3] >> 4) + (acc[3][3] >> 6));
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[i][q] = 0;
            slots[(size_t)(rbi - round0) * GV_CTHREADS + warp * 32 + lane] = v;
            ks = 0; ++rbi;
            if (rbi - round0 == nslots_r || rbi == my_rbs) {
                if (ct == 0 && rbi == my_rbs) GV_STAMP(5);
                reduce_round(round0, rbi - round0);
                round0 = rbi;
                if (rbi < my_rbs) consumer_bar();     // the slots are rewritten by the next round
            }
        }
        if (nks_r > 1) load_b(ks);
    }
    if (wrote_peer) __threadfence_system();      // peer stores are visible system-wide before the grid completes
    if (ct == 0) GV_STAMP(15);
}

// ---- parity aid: the activation quantisation of the fused prologue, written out densely (sx_out / qx_out of tb200_gemv)
template <typename XT>
__global__ void __launch_bounds__(256)
gemv_debug_quant_kernel(const XT* __restrict__ x, int64_t ldx, int64_t K, float* __restrict__ sx_out, int8_t* __restrict__ qx_out) {
    __shared__ float red[8];
    const XT* xr = x + (int64_t)blockIdx.x * ldx;
    float amax = 0.0f;
    for (int64_t k = threadIdx.x; k < K; k += 256) amax = fmaxf(amax, fabsf(to_f32(xr[k])));
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
    __syncthreads();
    float v = red[0];
    for (int w = 1; w < 8; ++w) v = fmaxf(v, red[w]);
    const float inv = v > 0.0f ? __fdiv_rn(127.0f, v) : 1.0f;
    if (threadIdx.x == 0 && sx_out) sx_out[blockIdx.x] = v > 0.0f ? __fdiv_rn(v, 127.0f) : 1.0f;
    if (qx_out)
        for (int64_t k = threadIdx.x; k < K; k += 256) qx_out[(int64_t)blockIdx.x * K + k] = (int8_t)quant_mul(to_f32(xr[k]), inv);
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
struct MapKey { const void* ptr; int64_t ld, N, stride; int layers; };   // one map per weight tensor
struct MapEntry { MapKey key; CUtensorMap map; bool used; bool one_box; };
static std::mutex g_map_mu;
static MapEntry g_maps[64];
static int g_map_next = 0;

int get_weight_map(const uint8_t* packed, int64_t ld, int64_t N, int64_t stride, int layers, CUtensorMap* out, int* one_box) {
    std::lock_guard<std::mutex> lk(g_map_mu);
    for (auto& e : g_maps)
        if (e.used && e.key.ptr == packed && e.key.ld == ld && e.key.N == N && e.key.stride == stride && e.key.layers == layers) {
            *out = e.map;
            *one_box = e.one_box;
            return TB200_OK;
        }
    MapEntry& e = g_maps[g_map_next];
    static const int no4d = (int)tb_knob("TB200_GEMV_NO4D", 0);
    int rc = TB200_ECUDA;
    e.one_box = false;
    if (ld % GV_SEG == 0 && !no4d) {
        // 4-D view (word in segment, row, 128-byte segment of the row, layer): a whole 16-row x 1024-byte tile is ONE box
        // that lands as [segment][row][128 B], the layout the consumers read (SWIZZLE_128B on the row index).
        uint64_t dims4[4] = {GV_SEG / 4, (uint64_t)N, (uint64_t)(ld / GV_SEG), (uint64_t)layers};
        uint64_t strides4[3] = {(uint64_t)ld, (uint64_t)GV_SEG, (uint64_t)(layers > 1 ? stride : ld * N)};
        uint32_t box4[4] = {GV_SEG / 4, GV_ROWS, GV_CWARPS, 1};
        rc = make_tensor_map(&e.map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, packed, dims4, strides4, box4, CU_TENSOR_MAP_SWIZZLE_128B);
        e.one_box = rc == TB200_OK;
    }
    uint64_t dims[3] = {(uint64_t)(ld / 4), (uint64_t)N, (uint64_t)layers};
    uint64_t strides[2] = {(uint64_t)ld, (uint64_t)(layers > 1 ? stride : ld * N)};
    uint32_t box[3] = {GV_SEG / 4, GV_ROWS, 1};
    if (rc) rc = make_tensor_map(&e.map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, packed, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) { e.used = false; return rc; }
    e.key = MapKey{packed, ld, N, stride, layers};
    e.used = true;
    g_map_next = (g_map_next + 1) % 64;
    *out = e.map;
    *one_box = e.one_box;
    return TB200_OK;
}

int gemv_nks(int64_t K) { return (int)((K + 4095) / 4096); }

static size_t gemv_smem(int m, int nks, int stages, int nslots, int nrb) {
    return 1024 + (size_t)stages * GV_TILE_BYTES + (size_t)nks * 4096 * m + (size_t)nslots * GV_CTHREADS * 16 +
           2 * GV_TOK * 4 + 2 * GV_TOK * GV_CWARPS * 4 + 2 * GV_MAX_STAGES * 8 + (size_t)nrb * GV_ROWS * 8 + 64;
}
// Ring depth and reduction slots.  Preferred: the CTA's whole share of the weights in the ring (then every load is
// issued before the kernel waits for its predecessor) and one slot per row block (one reduction round).  Two CTAs per
// SM must fit (this layer's and, under programmatic dependent launch, the next layer's): 2 x (smem + 1 KB reserved)
// <= 228 KB; large K*M falls back to one CTA per SM.
static bool gemv_pick(int m, int nks, int nrb, int* stages, int* nslots) {
    const int tiles = nrb * nks;
    const int want = tiles < GV_MAX_STAGES ? tiles : GV_MAX_STAGES;
    const int smax = nrb < GV_MAX_SLOTS ? nrb : GV_MAX_SLOTS;
    const size_t limits[2] = {113 * 1024, 227 * 1024};
    for (int l = 0; l < 2; ++l)
        for (int st = want; st >= (want < 2 ? want : 2); --st)
            for (int ns = smax; ns >= (smax < 2 ? smax : 2); --ns)
                if (gemv_smem(m, nks, st, ns, nrb) <= limits[l]) { *stages = st; *nslots = ns; return true; }
    if (gemv_smem(m, nks, 1, 1, nrb) <= limits[1]) { *stages = 1; *nslots = 1; return true; }
    return false;
}

template <typename XT, typename OT, bool FUSED, bool VEC>
static int launch_gemv(const CUtensorMap& map, GemvParams p, int layers, cudaStream_t s) {
    const int ctas = p.row_blocks < sm_count() ? p.row_blocks : sm_count();
    p.nrb = (p.row_blocks + ctas - 1) / ctas;
    TB_REQUIRE(gemv_pick(p.M, p.nks, p.nrb, &p.stages, &p.nslots),
               "gemv: K=%lld N=%lld is too large for the shared-memory activation image", (long long)p.K, (long long)p.N);
    size_t smem = gemv_smem(p.M, p.nks, p.stages, p.nslots, p.nrb);
    static const int solo_env = (int)tb_knob("TB200_GEMV_SOLO", 0);
    if (solo_env && smem < 120 * 1024) smem = 120 * 1024;     // experiment: one CTA per SM
    auto kern = gemv_mma_kernel<XT, OT, FUSED, VEC>;
    static SmemOptIn optin;         // per instantiation and device
    TB_CUDA(ensure_dynamic_smem(kern, optin, smem));
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

template <bool FUSED>
static int gemv_dispatch(int x_dtype, int out_dtype, const CUtensorMap& map, const GemvParams& p, int layers, cudaStream_t s) {
#define TB_GEMV_CASE(XD, XT, OD, OT) \
    if (x_dtype == XD && out_dtype == OD) \
        return (!FUSED || p.xvec) ? launch_gemv<XT, OT, FUSED, true>(map, p, layers, s) \
                                  : launch_gemv<XT, OT, FUSED, false>(map, p, layers, s)
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
                      int32_t* acc_out, int layers, int flags, cudaStream_t s,
                      void* const* peers = nullptr, int n_peers = 0) {
    CUtensorMap map;
    int one_box = 0;
    int rc = get_weight_map(packed, ld_packed, N, stride_packed, layers, &map, &one_box);
    if (rc) return rc;
    const size_t xs = dtype_size(x_dtype), os = dtype_size(out_dtype);
    GemvParams p = {};
    fill_common(p, K, N, scale, bias, stride_sb, clip_val, ldo, stride_out, flags);
    p.ldx = ldx; p.stride_x = stride_x; p.one_box = one_box;
    p.xvec = ((uintptr_t)x % 16 == 0) && ((ldx * xs) % 16 == 0) && ((stride_x * xs) % 16 == 0) && (K % 16 == 0);
    for (int64_t m0 = 0; m0 < M; m0 += GV_TOK) {
        p.M = (int)(M - m0 < GV_TOK ? M - m0 : GV_TOK);
        p.x = (const char*)x + m0 * ldx * xs;
        p.out = (char*)out + m0 * ldo * os;
        p.acc_out = acc_out ? acc_out + m0 * N : nullptr;
        set_peers(p, peers, n_peers, m0 * ldo * (int64_t)os);
        rc = gemv_dispatch<true>(x_dtype, out_dtype, map, p, layers, s);
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
    int one_box = 0;
    rc = get_weight_map(packed, ld_packed, N, 0, 1, &map, &one_box);
    if (rc) return rc;
    const size_t os = dtype_size(out_dtype);
    GemvParams p = {};
    fill_common(p, K, N, scale, bias, 0, clip_val, ldo, 0, flags);
    p.ldq = ldq; p.one_box = one_box;
    for (int64_t m0 = 0; m0 < M; m0 += GV_TOK) {     // one pass over the weights per 8 tokens
        p.M = (int)(M - m0 < GV_TOK ? M - m0 : GV_TOK);
        p.qx = q_x + m0 * ldq;
        p.sx_in = s_x + m0; p.sumq_in = sum_qx + m0;
        p.out = (char*)out + m0 * ldo * os;
        p.acc_out = acc_out ? acc_out + m0 * N : nullptr;
        set_peers(p, peers, n_peers, m0 * ldo * (int64_t)os);
        // activations are int8 already: the XT template parameter is unused on this path
        rc = gemv_dispatch<false>(out_dtype, out_dtype, map, p, 1, (cudaStream_t)stream);
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
    if (sx_out || qx_out) {
        TB_DISPATCH_DTYPE(x_dtype, T,
            (gemv_debug_quant_kernel<T><<<(unsigned)M, 256, 0, (cudaStream_t)stream>>>((const T*)x, ldx, K, sx_out, qx_out)));
        TB_LAUNCH_CHECK();
    }
    return gemv_fused(x, x_dtype, M, K, ldx, 0, packed, ld_packed, 0, N, scale, bias, 0, clip_val,
                      out, out_dtype, ldo, 0, acc_out, 1, flags, (cudaStream_t)stream);
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
                      stride_sb, clip_val, out, out_dtype, ldo, stride_out, nullptr,
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
