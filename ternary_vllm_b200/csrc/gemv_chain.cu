// gemv_chain.cu — a CHAIN of dependent quantised Linear layers in ONE persistent launch (decode, M <= 8 tokens).
//
//   h_0 = x ;  h_{l+1} = act_l( clip( (q(h_l) @ codes_l^T) * s_x * scale_l + bias_l ) ) ;  out = h_L
//
// This is what a model does with the layer (fc1 -> gelu_new -> fc2, impl/tf/moondream/layers.py:32-43 with the
// activation applied after the clip, impl/tf/moondream/ternary.py:245-246,298-299; out_proj after Wqkv's attention):
// every layer consumes the previous layer's output.  A launch per layer is latency-bound (~4.6 us of
// serialised prologue -> dot products -> reduction per layer against 1.7 us of HBM time for 4096 x 11008), so here
//   * ONE warp per CTA keeps the TMA weight stream running across ALL layers of the chain: weights do not depend on
//     activations, so while the CTAs wait for a layer's input their shared-memory rings already hold the next layer's
//     tiles (148 x 8 x 16 KB = 19 MB in flight, more than a whole 4096 x 11008 layer);
//   * the layer boundary is a counting flag in global memory, not a kernel boundary: every finished 16-row block adds 1
//     to its layer's counter (release), consumers poll it (acquire), then read the previous layer's output from L2,
//     quantise it per token to int8 (same arithmetic as tb200_quant_act_i8: absmax, 127/amax, round-half-even) straight
//     into the kperm activation image in shared memory -- no quantiser launch, no int8 round trip through HBM;
//   * dot products are the decode GEMV's (gemv_stream.cu): in-place 2-bit field masks + mma.sync.m16n8k32.u8.s8, two
//     teams of four warps on alternate row blocks, K-partials through shared-memory slots, and two aux warps that
//     reduce them and run the fused epilogue (-sum_qx, s_x*scale, bias, clip, ACTIVATION, store) off the MMA warps' path.
// Integer accumulators are bit-identical to the single-layer kernels'.  Inter-layer activations are stored in the
// activation dtype (the reference keeps x.dtype between layers, binary/modules.py:102-108); gelu_new is evaluated in
// fp32 on the stored-precision value and rounded once.
#include <math.h>
#include <string.h>
#include "tma.cuh"

namespace tb200 {

constexpr int CH_ROWS = 16, CH_SEG = 128, CH_TILE = 16 * 1024;
constexpr int CH_DOT_WARPS = 16, CH_TEAM = 8, CH_AUX = 3;   // 20 warps = 5 warpgroups: four of dot-product warps, one of producer + aux warps
constexpr int CH_WARP_PRODUCER = CH_DOT_WARPS, CH_WARP_AUX0 = CH_DOT_WARPS + 1;
constexpr int CH_THREADS = (CH_DOT_WARPS + 1 + CH_AUX) * 32;
constexpr int CH_DOT_THREADS = CH_DOT_WARPS * 32;
constexpr int CH_MAX_STAGES = 12, CH_MAX_SLOTS = 8;
constexpr int CH_CACHE_WORDS = 16;                 // raw 32-bit words of activations a thread keeps in registers between its two passes
constexpr uint32_t CH_MAGIC = 0x7b200c43u;

// device-side layer descriptor inside the plan blob (192 bytes, 64-byte aligned: the tensor map is used in place)
struct ChainLayer {
    CUtensorMap map;                               // 128 B
    const float* scale; const float* bias;
    int32_t N, K, nks, row_blocks;
    float clip_val; int32_t act, one_box, h_pitch; // h_pitch: ELEMENTS per token of this layer's output buffer (N rounded up to 16)
    int64_t h_off;                                 // byte offset of that buffer (two sets) in the workspace (last layer: unused)
    int64_t h_set_stride;                          // bytes between the two sets (even / odd generation)
};
static_assert(sizeof(ChainLayer) == 192, "plan layout");
struct ChainHeader {
    uint32_t magic, n_layers;
    int32_t max_nks, max_rb, M_max, esz;           // esz: bytes per inter-layer activation element
    int64_t ws_bytes;
    char pad[32];
};
static_assert(sizeof(ChainHeader) == 64, "plan layout");
// control block at the start of the workspace (zero before the first call; the kernel maintains it)
// every layer's counter sits on its own 128-byte line (pollers of one layer do not queue behind arrivals of another)
constexpr int CH_CNT_STRIDE = 32;
struct ChainCtrl { uint32_t gen, done, last_M, pad[29]; uint32_t counter[TB200_CHAIN_MAX_LAYERS * CH_CNT_STRIDE]; };
constexpr size_t CH_CTRL_BYTES = TB200_CHAIN_CTRL_BYTES;
static_assert(sizeof(ChainCtrl) <= CH_CTRL_BYTES, "control block");

struct ChainParams {
    const void* x; int64_t ldx;
    void* out; int64_t ldo; int out_dtype;
    uint8_t* ws;
    int stages, nslots, n_layers;
    uint32_t img_bytes;
    int pdl;
    int gate;             // 1: one thread per CTA waits for the layer's row-block counter before the data words are read
    int dbg;              // experiment build only: 1 no IMMAs, 2 no LOP3 (unmasked operands), 4 no weight LDS
    long long* trace;     // experiment build only: per CTA and layer, time stamps (tools/chain_trace.py)
};
__device__ __forceinline__ long long chain_gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
constexpr int CH_TR = 32;   // trace slots per (CTA, layer)
#define CH_SLOT(l, k) p.trace[((int64_t)b * TB200_CHAIN_MAX_LAYERS + (l)) * CH_TR + (k)]
#define CH_STAMP(l, k) do { if (kAblate && p.trace && ct == 0) CH_SLOT(l, k) = chain_gtimer(); } while (0)
#define CH_CLK(l, k) do { if (kAblate && p.trace && ct == 0) CH_SLOT(l, k) = clock64(); } while (0)
#define CH_STAMP_AUX(l, k) do { if (kAblate && p.trace && lane == 0) CH_SLOT(l, k) = chain_gtimer(); } while (0)
#define CH_CLK_AUX(l, k) do { if (kAblate && p.trace && lane == 0) CH_SLOT(l, k) = clock64(); } while (0)

// per-layer facts the warps need, copied to shared memory once
struct LayerInfo {
    const float* scale; const float* bias;
    uint8_t* hout;                                 // this layer's output buffer, set 0 (null for the last layer)
    int64_t set_stride;                            // bytes to set 1
    int N, K, nks, RB;
    float clip; int act, pitch, one_box;
};

__device__ __forceinline__ void imma_u8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// shared-memory accesses through 32-bit addresses kept in ordinary registers.  The addresses are made OPAQUE to the compiler
// (keep_reg) once per kernel: otherwise it rematerialises every shared-memory base from scratch at each use -- an
// S2UR SR_CgaCtaId + ~10 dependent uniform-datapath instructions per mbarrier / LDS address inside the per-tile loop, which
// measured ~600 cycles per 16 KB tile with the dot products compiled out (tools/chain_ablate.sh).
__device__ __forceinline__ void keep_reg(uint32_t& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void keep_reg(int& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ int lds32(uint32_t a) {
    int v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(uint32_t a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// 32 bytes in one request (sm_100: 256-bit global loads), from L2, never from a stale L1 line
__device__ __forceinline__ void ld_relaxed_256(const void* p, uint32_t (&r)[8]) {
    asm volatile("ld.relaxed.gpu.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_relaxed_u16(void* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u16 [%0], %1;" ::"l"(p), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ void st_relaxed_u32(void* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_add(uint32_t* p, uint32_t v) {
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ uint4 ld_cg_u4(const void* p) {
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
template <typename T> __device__ __forceinline__ T ld_cg_elem(const T* p);
template <> __device__ __forceinline__ float ld_cg_elem<float>(const float* p) { float v; asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p)); return v; }
template <> __device__ __forceinline__ __nv_bfloat16 ld_cg_elem<__nv_bfloat16>(const __nv_bfloat16* p) {
    unsigned short v; asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(v) : "l"(p)); return *reinterpret_cast<__nv_bfloat16*>(&v);
}
template <> __device__ __forceinline__ __half ld_cg_elem<__half>(const __half* p) {
    unsigned short v; asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(v) : "l"(p)); return *reinterpret_cast<__half*>(&v);
}

// gelu_new / keras gelu(approximate=True): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))  (layers.py:21-27)
__device__ __forceinline__ float gelu_tanh(float x) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.0f + tanhf(u));
}
template <typename T> __device__ __forceinline__ float round_to(float v) { return to_f32(from_f32<T>(v)); }

// Inter-layer activations are stored INVERTED (~bits) in the activation dtype, and an all-zero element means "not
// written yet": the buffers start as zeros, the value whose bits are all ones (a NaN no conversion produces) is never
// stored, so data and validity travel in the same single-copy-atomic element and no fence is needed between a
// producer's stores and its counter increment.
template <typename T> __device__ __forceinline__ uint32_t chain_encode(float y);
template <> __device__ __forceinline__ uint32_t chain_encode<float>(float y) {
    const uint32_t bits = __float_as_uint(y);
    return ~(bits == 0xffffffffu ? 0x7fffffffu : bits);
}
template <> __device__ __forceinline__ uint32_t chain_encode<__nv_bfloat16>(float y) {
    const __nv_bfloat16 h = __float2bfloat16_rn(y);
    const uint32_t bits = *reinterpret_cast<const unsigned short*>(&h);
    return (~(bits == 0xffffu ? 0x7fffu : bits)) & 0xffffu;
}
template <> __device__ __forceinline__ uint32_t chain_encode<__half>(float y) {
    const __half h = __float2half_rn(y);
    const uint32_t bits = *reinterpret_cast<const unsigned short*>(&h);
    return (~(bits == 0xffffu ? 0x7fffu : bits)) & 0xffffu;
}

// position of 16-element group gi in the activation image (uint4 units): inside every 32-group segment the (chunk, word)
// index is transposed so that the eight chunk lanes of a warp read eight consecutive uint4 (no bank conflict)
__device__ __forceinline__ int chain_img_pos(int gi) { return (gi & ~31) | ((gi & 3) << 3) | ((gi >> 2) & 7); }

// walk of a CTA through ITS row blocks of all layers: (l, rb) with rb = b, b + G, ... < RB[l]; n counts them
struct Walk {
    int l, rb, n;
    __device__ __forceinline__ void settle(const LayerInfo* lay, int L, int b) {
        while (l < L && rb >= lay[l].RB) { ++l; rb = b; }
    }
    __device__ __forceinline__ void start(const LayerInfo* lay, int L, int b) { l = 0; rb = b; n = 0; settle(lay, L, b); }
    __device__ __forceinline__ void next(const LayerInfo* lay, int L, int b, int G) {
        rb += G; ++n;
        if (l < L && rb >= lay[l].RB) { ++l; rb = b; settle(lay, L, b); }
    }
};

template <typename XT, int M>
__global__ void __launch_bounds__(CH_THREADS, 1)
gemv_chain_kernel(const ChainHeader* __restrict__ plan, const ChainParams p) {
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    // ring[stages][16 KB] | image[M][groups] uint4 | slots[nslots][16 warps][16 rows][MP] int | red_f[2][M][16] | red_i[2][M][16] |
    // sx[2][8] | layers | barriers
    constexpr int MP = M | 1;                                   // odd token pitch of a slot row: conflict-free reads
    constexpr int SLOT_INTS = CH_TEAM * CH_ROWS * MP;
    const uint32_t sRing = base;
    uint4* img = reinterpret_cast<uint4*>(gbase + (size_t)p.stages * CH_TILE);
    int* slots = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(img) + ((p.img_bytes + 127u) & ~127u));
    float* red_f = reinterpret_cast<float*>(slots + (size_t)p.nslots * SLOT_INTS);
    int* red_i = reinterpret_cast<int*>(red_f + 2 * M * CH_DOT_WARPS);
    float* sx_sm = reinterpret_cast<float*>(red_i + 2 * M * CH_DOT_WARPS);
    LayerInfo* lay = reinterpret_cast<LayerInfo*>(sx_sm + 16);
    const int L = p.n_layers;
    uint64_t* bars = reinterpret_cast<uint64_t*>(lay + L);
    const uint32_t sBar = smem_u32(bars);
    auto bar_full = [&](int s) { return sBar + 8u * s; };
    auto bar_empty = [&](int s) { return sBar + 8u * (CH_MAX_STAGES + s); };
    auto bar_sfull = [&](int q) { return sBar + 8u * (2 * CH_MAX_STAGES + q); };
    auto bar_sempty = [&](int q) { return sBar + 8u * (2 * CH_MAX_STAGES + CH_MAX_SLOTS + q); };

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = gridDim.x, b = blockIdx.x;
    const ChainLayer* layers = reinterpret_cast<const ChainLayer*>(plan + 1);
    ChainCtrl* ctrl = reinterpret_cast<ChainCtrl*>(p.ws);

    for (int l = threadIdx.x; l < L; l += CH_THREADS) {
        const ChainLayer& cl = layers[l];
        LayerInfo li;
        li.scale = cl.scale; li.bias = cl.bias;
        li.hout = (l + 1 < L) ? p.ws + cl.h_off : nullptr;
        li.set_stride = cl.h_set_stride;
        li.N = cl.N; li.K = cl.K; li.nks = cl.nks; li.RB = cl.row_blocks;
        li.clip = cl.clip_val; li.act = cl.act; li.pitch = cl.h_pitch; li.one_box = cl.one_box;
        lay[l] = li;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), CH_TEAM); }
        for (int q = 0; q < p.nslots; ++q) { mbar_init(bar_sfull(q), CH_TEAM); mbar_init(bar_sempty(q), 1); }
        mbar_fence_init();
    }
    __syncthreads();
    if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");

    if (warp >= CH_DOT_WARPS) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");   // 4 x 32 x (96 - 64) = exactly what 16 dot-product warps need to go from 96 to 104
        if (warp == CH_WARP_PRODUCER) {
            // ===================== weight tiles of ALL layers: independent of every activation =====================
            if (lane == 0) {
                int s = 0; uint32_t ph = 0;
                Walk w; w.start(lay, L, b);
                while (w.l < L) {
                    const CUtensorMap* map = &layers[w.l].map;
                    const int nks = lay[w.l].nks, one_box = lay[w.l].one_box;
                    for (int ks = 0; ks < nks; ++ks) {
                        mbar_wait_relaxed(bar_empty(s), ph ^ 1u);
                        mbar_arrive_expect_tx(bar_full(s), CH_TILE);
                        if (one_box) {
                            tma_load_4d(sRing + s * CH_TILE, map, bar_full(s), 0, w.rb * CH_ROWS, ks * 8, 0);
                        } else {
#pragma unroll
                            for (int sg = 0; sg < 8; ++sg)
                                tma_load_3d(sRing + s * CH_TILE + sg * (CH_ROWS * CH_SEG), map, bar_full(s), ks * 256 + sg * 32, w.rb * CH_ROWS, 0);
                        }
                        if (++s == p.stages) { s = 0; ph ^= 1u; }
                    }
                    w.next(lay, L, b, G);
                }
            }
            return;
        }
        // ===================== reduction of the K-partials + fused epilogue (+ activation) =====================
        // lane -> row (lane & 15); the two half-warps add four warps' partials each, then take alternate tokens
        const int a = warp - CH_WARP_AUX0;
        const int row_l = lane & 15, half = lane >> 4;
        uint32_t a_bar = sBar;                                                                        // see keep_reg
        uint32_t a_sl = smem_u32(slots) + 4u * ((half * (CH_TEAM / 2) * CH_ROWS + row_l) * MP);  // this lane's partials in slot 0
        int nslots_a = p.nslots;
        keep_reg(a_bar); keep_reg(a_sl); keep_reg(nslots_a);
        if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // earlier kernels may still read out / the h buffers
        const uint32_t gen = ld_acquire_u32(&ctrl->gen);
        const int set = (int)(gen & 1u);
        const int prevM = (int)ld_relaxed_u32(&ctrl->last_M);
        Walk w; w.start(lay, L, b);
        while (w.l < L) {
            if ((w.n % CH_AUX) == a) {
                const LayerInfo& li = lay[w.l];
                const int l = w.l, par = l & 1;
                const int sq_i = w.n % nslots_a;
                const uint32_t spar = (uint32_t)(w.n / nslots_a) & 1u;
                const int64_t row = (int64_t)w.rb * CH_ROWS + row_l;
                float sc = 1.0f, bi = 0.0f;                               // requested before the wait for the partial sums
                if (li.scale && row < li.N) sc = __ldg(li.scale + row);
                if (li.bias && row < li.N) bi = __ldg(li.bias + row);
                // scale / bias are read once per call (cold in L2) and their loads would queue behind the weight stream right
                // when the layer's last row blocks wait for them: ask L2 for the same rows of the next two layers now
                if (lane < 4) {
                    const int l2 = l + 1 + (lane >> 1);
                    if (l2 < L && w.rb < lay[l2].RB) {
                        const float* ptr = (lane & 1) ? lay[l2].bias : lay[l2].scale;
                        if (ptr) prefetch_l2(ptr + (int64_t)w.rb * CH_ROWS);
                    }
                }
                CH_CLK_AUX(l, 17);
                mbar_wait(a_bar + 8u * (2 * CH_MAX_STAGES + sq_i), spar);
                CH_CLK_AUX(l, 18);
                const uint32_t sl = a_sl + 4u * (sq_i * SLOT_INTS);
                int tot[M]; float sxv[M]; int sumq[M];
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    int v = 0;
#pragma unroll
                    for (int wv = 0; wv < CH_TEAM / 2; ++wv) v += lds32(sl + 4u * (wv * CH_ROWS * MP + m));
                    tot[m] = v;
                    sxv[m] = sx_sm[par * 8 + m];
                    int sq = 0;
#pragma unroll
                    for (int j = 0; j < CH_DOT_WARPS / 4; ++j) {
                        const int4 a0 = *reinterpret_cast<const int4*>(red_i + (par * M + m) * CH_DOT_WARPS + 4 * j);
                        sq += a0.x + a0.y + a0.z + a0.w;
                    }
                    sumq[m] = sq;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(a_bar + 8u * (2 * CH_MAX_STAGES + CH_MAX_SLOTS + sq_i));   // partial sums and token scalars are in registers
                float yv[M];
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const int av = tot[m] + __shfl_xor_sync(0xffffffffu, tot[m], 16) - sumq[m];
                    const float tv = __fmul_rn((float)av, sxv[m]);
                    float y = epilogue(tv, sc, bi, li.clip);
                    // the activation function sees the value in the STORED (activation dtype) precision
                    if (li.act == TB200_ACT_GELU_TANH) y = gelu_tanh(round_to<XT>(y));
                    else if (li.act == TB200_ACT_RELU) y = fmaxf(y, 0.0f);
                    yv[m] = row < li.N ? y : 0.0f;                           // rows of the padding: zeros for the next layer's image
                }
                CH_CLK_AUX(l, 19);
                if (li.hout) {
                    // this call's set gets the values (every row up to the pitch: the consumers read whole 16-element groups);
                    // the other set, read by the previous call, is cleared for the next one
                    uint8_t* hs = li.hout + (int64_t)set * li.set_stride;
                    uint8_t* ho = li.hout + (int64_t)(set ^ 1) * li.set_stride;
#pragma unroll
                    for (int m = 0; m < M; ++m)
                        if (half == (m & 1)) {
                            const int64_t o = ((int64_t)m * li.pitch + row) * sizeof(XT);
                            if (sizeof(XT) == 2) st_relaxed_u16(hs + o, chain_encode<XT>(yv[m]));
                            else st_relaxed_u32(hs + o, chain_encode<XT>(yv[m]));
                        }
#pragma unroll
                    for (int m = 0; m < TB200_CHAIN_MAX_M; ++m)
                        if (m < prevM && half == (m & 1)) {
                            const int64_t o = ((int64_t)m * li.pitch + row) * sizeof(XT);
                            if (sizeof(XT) == 2) st_relaxed_u16(ho + o, 0u);
                            else st_relaxed_u32(ho + o, 0u);
                        }
                    __syncwarp();
                    if (lane == 0) red_relaxed_add(&ctrl->counter[l * CH_CNT_STRIDE], 1u);     // hint for the pollers
                } else {
#pragma unroll
                    for (int m = 0; m < M; ++m)
                        if (half == (m & 1) && row < li.N) {
                            const int64_t o = (int64_t)m * p.ldo + row;
                            if (p.out_dtype == TB200_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(yv[m]);
                            else if (p.out_dtype == TB200_F16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(yv[m]);
                            else reinterpret_cast<float*>(p.out)[o] = yv[m];
                        }
                }
                CH_CLK_AUX(l, 20);
            }
            w.next(lay, L, b, G);
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
        // ===================== activation image of every layer + dot products =====================
        // Two teams of eight warps take alternate row blocks, so that one team's per-row-block latencies (barrier waits, the
        // recombination, the slot hand-off) overlap the other team's MMAs.  In a team, warp wq owns 128-byte K segment wq of
        // every tile (512 weights of each of the 16 rows).  Lane (g, t) holds the mma.sync.m16n8k32 A fragments of rows g and
        // g + 8: 2-bit field i of every packed byte is masked IN PLACE (value code * 4^i) and multiplied with the int8
        // activations that meet it -- four accumulator sets, recombined exactly with shifts at the end of the row block.  The
        // B fragments (token = g, zero for g >= M) of a whole k-slice stay in registers.  A tile costs a warp 4 LDS.128 + 64 LOP3 + 16 IMMA whatever M is.
        const int ct = threadIdx.x;                          // 0..511
        const int g = lane >> 2, t = lane & 3;
        const int team = warp >> 3, wq = warp & 7;
        int stages = p.stages, nslots = p.nslots;
        keep_reg(stages); keep_reg(nslots);
        // shared-memory addresses of this thread, computed ONCE (see keep_reg)
        uint32_t a_bar = sBar;                                                                   // mbarriers
        uint32_t a_w = base + wq * (CH_ROWS * CH_SEG) + g * CH_SEG;                               // row g of this warp's segment in stage 0
        uint32_t a_slot = smem_u32(slots) + 4u * ((wq * CH_ROWS + g) * MP + 2 * t);               // partial sums in slot 0
        uint32_t a_img = smem_u32(img) + 16u * (wq * 32 + 2 * t);                                 // image words of token 0, k-slice 0
        keep_reg(a_bar); keep_reg(a_w); keep_reg(a_slot); keep_reg(a_img);
        const uint32_t sw0 = (uint32_t)((2 * t) ^ g) << 4, sw1 = (uint32_t)((2 * t + 1) ^ g) << 4;   // SWIZZLE_128B: chunk ^ (row & 7)
        if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // x (and the control block) come from earlier kernels
        const uint32_t gen = ld_acquire_u32(&ctrl->gen);
        const int set = (int)(gen & 1u);
        int sq = 0; uint32_t spar = 0;                       // slot ring (every row block of the CTA takes the next slot)
        uint4 bq[2][4];                                      // image words of packed words 0..3 of this lane's two chunks: .x/.y/.z/.w = field 0..3
        int cur = -1, ngroups = 0;
        long long tr_full = 0, tr_slot = 0;                  // experiment build: cycles this thread waited for tiles / for a slot
        Walk w; w.start(lay, L, b);
        auto load_b = [&](int ks) {
#pragma unroll
            for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    bq[pp][j] = g < M ? lds128(a_img + 16u * (g * ngroups + ks * 256 + j * 8 + pp)) : make_uint4(0u, 0u, 0u, 0u);
        };
        int s = 0; uint32_t ph = 0;                          // weight ring (global tile sequence of this CTA)
        constexpr int RW = 4 * (int)sizeof(XT);              // 32-bit words per 16-element group (8 or 16)
        while (w.l < L) {
            const LayerInfo& li = lay[w.l];
            if (w.l != cur) {
                // ---------- entering layer w.l: its input must be complete, then quantise it into the image ----------
                cur = w.l;
                const int l = cur, par = l & 1;
                CH_STAMP(l, 0); CH_CLK(l, 8);
                const int K = li.K, nks = li.nks;
                ngroups = nks * 256;
                const int nu = (nks + 1) >> 1;               // 512-group slices of the input: group gi = u * 512 + ct
                const XT* xin = reinterpret_cast<const XT*>(p.x);
                const uint8_t* hin = nullptr; int64_t pitch_b = 0;
                const bool vec0 = ((reinterpret_cast<uintptr_t>(p.x) | (uintptr_t)(p.ldx * sizeof(XT))) & 15) == 0;
                if (l > 0) {
                    const LayerInfo& pl = lay[l - 1];
                    hin = pl.hout + (int64_t)set * pl.set_stride; pitch_b = (int64_t)pl.pitch * sizeof(XT);
                    if (p.gate) {
                        if (ct == 0) {
                            // coarse wait: ONE poller per CTA on the layer's row-block counter (relaxed loads; the counter is only
                            // a hint that the data is on its way -- the elements themselves say whether they have been written)
                            const uint32_t need = (gen + 1u) * (uint32_t)pl.RB;
                            const uint32_t* cnt = &ctrl->counter[(l - 1) * CH_CNT_STRIDE];
                            uint32_t spins = 0;
                            while ((int)(ld_relaxed_u32(cnt) - need) < 0) {
                                if (++spins > 50000000u) __trap();          // a protocol bug traps instead of hanging the GPU
                            }
                        }
                        asm volatile("bar.sync 1, %0;" ::"n"(CH_DOT_THREADS) : "memory");
                    }
                }
                CH_CLK(l, 9);
                // one 16-element group of token m as PLAIN bits of the activation dtype (zero beyond K).  Layer 0 reads the
                // caller's activations; later layers read what the previous layer's epilogue stored (inverted bits, zero = not
                // written yet): a group with a missing element is simply fetched again.
                auto fetch = [&](int m, int gi, uint32_t (&r)[RW]) -> bool {
                    if (16 * gi >= K) {
#pragma unroll
                        for (int j = 0; j < RW; ++j) r[j] = 0u;
                        return true;
                    }
                    if (l == 0) {
                        const XT* src = xin + (int64_t)m * p.ldx + 16 * (int64_t)gi;
                        if (vec0 && 16 * gi + 16 <= K) {
#pragma unroll
                            for (int j = 0; j < RW / 4; ++j) {
                                const uint4 q4 = ld_cg_u4(reinterpret_cast<const uint4*>(src) + j);
                                r[4 * j] = q4.x; r[4 * j + 1] = q4.y; r[4 * j + 2] = q4.z; r[4 * j + 3] = q4.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < RW; ++j) r[j] = 0u;
#pragma unroll
                            for (int tt = 0; tt < 16; ++tt)
                                if (16 * gi + tt < K) {
                                    const XT e = ld_cg_elem<XT>(src + tt);
                                    if (sizeof(XT) == 4) r[tt * RW / 16] = *reinterpret_cast<const uint32_t*>(&e);
                                    else r[tt * RW / 16] |= (uint32_t)(*reinterpret_cast<const unsigned short*>(&e)) << (16 * (tt & 1));
                                }
                        }
                        return true;
                    }
                    const uint8_t* src = hin + (int64_t)m * pitch_b + (int64_t)gi * (16 * sizeof(XT));
                    uint32_t miss = 0;
#pragma unroll
                    for (int h = 0; h < RW / 8; ++h) {
                        uint32_t q8[8];
                        ld_relaxed_256(src + 32 * h, q8);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (sizeof(XT) == 2) miss |= (q8[j] - 0x00010001u) & ~q8[j] & 0x80008000u;      // a 16-bit lane that is zero
                            else miss |= (uint32_t)(q8[j] == 0u);
                            r[8 * h + j] = ~q8[j];
                        }
                    }
                    return miss == 0u;
                };
                auto to_f32v = [&](const uint32_t (&r)[RW], float (&v)[16]) {
#pragma unroll
                    for (int tt = 0; tt < 16; ++tt) {
                        if (sizeof(XT) == 4) v[tt] = __uint_as_float(r[tt * RW / 16]);
                        else {
                            const unsigned short hb = (unsigned short)(r[tt * RW / 16] >> (16 * (tt & 1)));
                            v[tt] = to_f32(*reinterpret_cast<const XT*>(&hb));
                        }
                    }
                };
                auto group_amax = [&](const uint32_t (&r)[RW]) -> float {
                    float v[16]; to_f32v(r, v);
                    float am = 0.0f;
#pragma unroll
                    for (int tt = 0; tt < 16; ++tt) am = fmaxf(am, fabsf(v[tt]));
                    return am;
                };
                auto quantise_group = [&](const uint32_t (&r)[RW], float inv, int m, int gi, int& ls) {
                    if (gi >= ngroups) return;    // beyond the image (odd number of 256-group slices)
                    float v[16]; to_f32v(r, v);
                    uint32_t qb[16];              // low byte = two's-complement int8 code (rint via the 1.5 * 2^23 trick)
#pragma unroll
                    for (int tt = 0; tt < 16; ++tt) {
                        float q = __fmul_rn(v[tt], inv);
                        q = fminf(fmaxf(q, -127.0f), 127.0f);
                        qb[tt] = __float_as_uint(__fadd_rn(q, 12582912.0f));
                    }
                    uint32_t wd[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {   // word i = elements i, 4+i, 8+i, 12+i: the bytes that meet field i of packed word gi
                        wd[i] = __byte_perm(__byte_perm(qb[i], qb[4 + i], 0x0040), __byte_perm(qb[8 + i], qb[12 + i], 0x0040), 0x5410);
                        ls = __dp4a((int)wd[i], 0x01010101, ls);
                    }
                    img[m * ngroups + chain_img_pos(gi)] = make_uint4(wd[0], wd[1], wd[2], wd[3]);
                };
                auto token_scale = [&](int m) -> float {     // after the barrier: absmax of token m -> 127/amax; s_x to shared
                    float vmax = 0.0f;
#pragma unroll
                    for (int j = 0; j < CH_DOT_WARPS / 4; ++j) {
                        const float4 r0 = *reinterpret_cast<const float4*>(red_f + (par * M + m) * CH_DOT_WARPS + 4 * j);
                        vmax = fmaxf(vmax, fmaxf(fmaxf(r0.x, r0.y), fmaxf(r0.z, r0.w)));
                    }
                    if (ct == 0) sx_sm[par * 8 + m] = vmax > 0.0f ? __fdiv_rn(vmax, 127.0f) : 1.0f;
                    return vmax > 0.0f ? __fdiv_rn(127.0f, vmax) : 1.0f;
                };
                // NOTE: register arrays below are indexed with compile-time constants only.  With ~225 KB of the SM carved out as
                // shared memory there is practically no L1 left, so a local-memory (stack) array costs an L2 round trip per access.
                constexpr int NKC = CH_CACHE_WORDS / (M * RW) > 4 ? 4 : CH_CACHE_WORDS / (M * RW);   // 512-group slices per token whose raw words fit the register cache
                bool cached = false;
                if constexpr (NKC > 0) {
                    if (nu <= NKC) {
                        cached = true;
                        uint32_t rc[M][NKC][RW];
                        // all loads of the thread are issued before the first use (ONE L2 round trip), and again until every element is there
                        uint32_t spins = 0;
                        bool ok;
                        do {
                            ok = true;
#pragma unroll
                            for (int m = 0; m < M; ++m)
#pragma unroll
                                for (int u = 0; u < NKC; ++u)
                                    if (u < nu) ok = fetch(m, u * CH_DOT_THREADS + ct, rc[m][u]) && ok;
                            if (!ok && ++spins > 20000000u) __trap();
                        } while (!ok);
                        CH_CLK(l, 10);
                        if (kAblate && p.trace && ct == 0) CH_SLOT(l, 7) = spins;
#pragma unroll
                        for (int m = 0; m < M; ++m) {            // pass A: per-token absmax
                            float amax = 0.0f;
#pragma unroll
                            for (int u = 0; u < NKC; ++u)
                                if (u < nu) amax = fmaxf(amax, group_amax(rc[m][u]));
                            const uint32_t am = __reduce_max_sync(0xffffffffu, __float_as_uint(amax));
                            if (lane == 0) red_f[(par * M + m) * CH_DOT_WARPS + warp] = __uint_as_float(am);
                        }
                        asm volatile("bar.sync 1, %0;" ::"n"(CH_DOT_THREADS) : "memory");
                        CH_CLK(l, 11);
#pragma unroll
                        for (int m = 0; m < M; ++m) {            // pass B: quantise into the image (zero beyond K), code sums
                            const float inv = token_scale(m);
                            int ls = 0;
#pragma unroll
                            for (int u = 0; u < NKC; ++u)
                                if (u < nu) quantise_group(rc[m][u], inv, m, u * CH_DOT_THREADS + ct, ls);
                            ls = __reduce_add_sync(0xffffffffu, ls);
                            if (lane == 0) red_i[(par * M + m) * CH_DOT_WARPS + warp] = ls;
                        }
                    }
                }
                if (!cached) {
#pragma unroll 1
                    for (int m = 0; m < M; ++m) {            // pass A (the data is fetched again in pass B: L2)
                        float amax = 0.0f;
                        for (int u = 0; u < nu; ++u) {
                            uint32_t r[RW];
                            uint32_t spins = 0;
                            while (!fetch(m, u * CH_DOT_THREADS + ct, r)) { if (++spins > 20000000u) __trap(); }
                            amax = fmaxf(amax, group_amax(r));
                        }
                        const uint32_t am = __reduce_max_sync(0xffffffffu, __float_as_uint(amax));
                        if (lane == 0) red_f[(par * M + m) * CH_DOT_WARPS + warp] = __uint_as_float(am);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(CH_DOT_THREADS) : "memory");
                    CH_CLK(l, 11);
#pragma unroll 1
                    for (int m = 0; m < M; ++m) {
                        const float inv = token_scale(m);
                        int ls = 0;
                        for (int u = 0; u < nu; ++u) {
                            uint32_t r[RW];
                            (void)fetch(m, u * CH_DOT_THREADS + ct, r);           // verified in pass A
                            quantise_group(r, inv, m, u * CH_DOT_THREADS + ct, ls);
                        }
                        ls = __reduce_add_sync(0xffffffffu, ls);
                        if (lane == 0) red_i[(par * M + m) * CH_DOT_WARPS + warp] = ls;
                    }
                }
                CH_CLK(l, 14);
                // image complete for every warp; sx_sm and red_i (read by the aux warps) are ordered by this barrier + the first
                // slot hand-off of the layer
                asm volatile("bar.sync 1, %0;" ::"n"(CH_DOT_THREADS) : "memory");
                load_b(0);
                CH_CLK(l, 12);
                tr_full = 0; tr_slot = 0;
            }
            // ---------- one row block: team (n & 1) computes it (one K segment of every k-slice per warp) ----------
            const int nks = li.nks;
            if ((w.n & 1) == team) {
                const int rbi = (w.rb - b) / G;                      // index of this row block among the CTA's blocks of the layer
                const int tsl = rbi < 3 ? (rbi < 2 ? 1 + 3 * rbi : 21) : -1;
                if (kAblate && p.trace && ct == 0 && tsl >= 0) CH_SLOT(w.l, tsl) = clock64();
                int acc[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[i][e] = 0;
#pragma unroll 1
                for (int k2 = 0; k2 < nks; ++k2) {
                    const long long tw0 = kAblate && p.trace ? clock64() : 0;
                    mbar_wait(a_bar + 8u * s, ph);
                    if (kAblate && p.trace) tr_full += clock64() - tw0;
                    const uint32_t aw = a_w + s * CH_TILE;
                    uint4 lo0 = make_uint4(aw, s, ph, 3u), hi0 = lo0, lo1 = lo0, hi1 = lo0;
                    if (!(kAblate && (p.dbg & 4))) {
                        lo0 = lds128(aw + sw0); hi0 = lds128(aw + 8 * CH_SEG + sw0);
                        lo1 = lds128(aw + sw1); hi1 = lds128(aw + 8 * CH_SEG + sw1);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(a_bar + 8u * (CH_MAX_STAGES + s));   // this warp's part of the tile lives in registers now
                    if (++s == stages) { s = 0; ph ^= 1u; }
                    if (nks > 1) load_b(k2);           // K > 4096: the activation words of this k-slice (a single slice stays in registers)
                    if (kAblate && (p.dbg & 1)) {
                        acc[0][0] += (int)(lo0.x ^ hi0.y ^ lo1.z ^ hi1.w);
                    } else {
#pragma unroll
                        for (int pp = 0; pp < 2; ++pp) {
                            const uint4 lo = pp ? lo1 : lo0, hi = pp ? hi1 : hi0;
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const uint32_t mk = 0x03030303u << (2 * i);
                                const uint32_t b0 = i == 0 ? bq[pp][0].x : i == 1 ? bq[pp][0].y : i == 2 ? bq[pp][0].z : bq[pp][0].w;
                                const uint32_t b1 = i == 0 ? bq[pp][1].x : i == 1 ? bq[pp][1].y : i == 2 ? bq[pp][1].z : bq[pp][1].w;
                                imma_u8s8(acc[i], lo.x & mk, hi.x & mk, lo.y & mk, hi.y & mk, b0, b1);
                            }
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const uint32_t mk = 0x03030303u << (2 * i);
                                const uint32_t b2 = i == 0 ? bq[pp][2].x : i == 1 ? bq[pp][2].y : i == 2 ? bq[pp][2].z : bq[pp][2].w;
                                const uint32_t b3 = i == 0 ? bq[pp][3].x : i == 1 ? bq[pp][3].y : i == 2 ? bq[pp][3].z : bq[pp][3].w;
                                imma_u8s8(acc[i], lo.z & mk, hi.z & mk, lo.w & mk, hi.w & mk, b2, b3);
                            }
                        }
                    }
                }
                // exact recombination: accumulator set i is a multiple of 4^i.  Lane (g, t) holds rows g / g + 8, tokens 2t / 2t + 1
                int fin[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) fin[e] = acc[0][e] + (acc[1][e] >> 2) + (acc[2][e] >> 4) + (acc[3][e] >> 6);
                if (kAblate && p.trace && ct == 0 && tsl >= 0) CH_SLOT(w.l, tsl + 1) = clock64() + (fin[0] & 0);
                const long long tw1 = kAblate && p.trace ? clock64() : 0;
                mbar_wait(a_bar + 8u * (2 * CH_MAX_STAGES + CH_MAX_SLOTS + sq), spar ^ 1u);
                if (kAblate && p.trace) tr_slot += clock64() - tw1;
                const uint32_t dst = a_slot + 4u * (sq * SLOT_INTS);
                if (2 * t < M) { sts32(dst, fin[0]); sts32(dst + 4u * (8 * MP), fin[2]); }
                if (2 * t + 1 < M) { sts32(dst + 4u, fin[1]); sts32(dst + 4u * (8 * MP + 1), fin[3]); }
                __syncwarp();
                if (lane == 0) mbar_arrive(a_bar + 8u * (2 * CH_MAX_STAGES + sq));
                CH_CLK(w.l, 13);
                if (kAblate && p.trace && ct == 0 && tsl >= 0) CH_SLOT(w.l, tsl + 2) = clock64();
                if (kAblate && p.trace && ct == 0) { CH_SLOT(w.l, 15) = tr_full; CH_SLOT(w.l, 16) = tr_slot; }
            } else {
                s += nks;                                                // the other team's tiles
                while (s >= stages) { s -= stages; ph ^= 1u; }
            }
            if (++sq == nslots) { sq = 0; spar ^= 1u; }
            w.next(lay, L, b, G);
        }
    }
    // ---- the last CTA to finish advances the generation (counters are cumulative across calls, never reset) ----
    asm volatile("bar.sync 2, %0;" ::"n"(CH_THREADS - 32) : "memory");     // every warp but the producer
    if (threadIdx.x == 0) {
        __threadfence();
        const uint32_t old = atomicAdd(&ctrl->done, 1u);
        if (old == (uint32_t)G - 1u) {
            ctrl->done = 0u;
            ctrl->last_M = (uint32_t)M;
            __threadfence();
            atomicAdd(&ctrl->gen, 1u);
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------------
static inline int64_t ch_round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }
static inline int ch_nks(int64_t K) { return (int)((K + 4095) / 4096); }

static size_t chain_smem(int n_layers, int stages, int nslots, uint32_t img_bytes, int M) {
    return 1024 + (size_t)stages * CH_TILE + ((img_bytes + 127u) & ~127u) + (size_t)nslots * CH_TEAM * CH_ROWS * (M | 1) * 4 +
           2 * (size_t)M * CH_DOT_WARPS * 4 * 2 + 16 * 4 + (size_t)n_layers * sizeof(LayerInfo) + (2 * CH_MAX_STAGES + 2 * CH_MAX_SLOTS) * 8 + 64;
}

}  // namespace tb200

using namespace tb200;

extern "C" size_t tb200_chain_plan_bytes(int n_layers) {
    if (n_layers < 1 || n_layers > TB200_CHAIN_MAX_LAYERS) return 0;
    return sizeof(ChainHeader) + (size_t)n_layers * sizeof(ChainLayer);
}

// Fills `plan_host` (tb200_chain_plan_bytes(n) bytes) from the caller's layer list; the caller copies the blob to
// 256-byte aligned device memory and passes that copy to tb200_chain_forward.  act_elem_size: bytes per element of the
// activations between layers (= element size of x: 2 for bf16 / f16, 4 for f32).
extern "C" int tb200_chain_plan_build(const tb200_chain_layer* layers_in, int n_layers, int64_t M_max, int act_elem_size,
                                      void* plan_host, size_t* workspace_bytes) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(layers_in && plan_host && workspace_bytes, "chain_plan_build: null pointer");
    TB_REQUIRE(n_layers >= 1 && n_layers <= TB200_CHAIN_MAX_LAYERS, "chain_plan_build: n_layers=%d out of range [1,%d]", n_layers, TB200_CHAIN_MAX_LAYERS);
    TB_REQUIRE(M_max >= 1 && M_max <= TB200_CHAIN_MAX_M, "chain_plan_build: M_max=%lld out of range [1,%d]", (long long)M_max, TB200_CHAIN_MAX_M);
    TB_REQUIRE(act_elem_size == 2 || act_elem_size == 4, "chain_plan_build: act_elem_size must be 2 or 4");
    ChainHeader* hd = (ChainHeader*)plan_host;
    ChainLayer* out = (ChainLayer*)(hd + 1);
    memset(plan_host, 0, tb200_chain_plan_bytes(n_layers));
    int64_t off = (int64_t)CH_CTRL_BYTES;
    int max_nks = 1, max_rb = 1;
    for (int l = 0; l < n_layers; ++l) {
        const tb200_chain_layer& in = layers_in[l];
        TB_REQUIRE(in.packed && in.N > 0 && in.K > 0, "chain_plan_build: layer %d: bad shape / null weights", l);
        TB_REQUIRE(l == 0 || in.K == layers_in[l - 1].N, "chain_plan_build: layer %d consumes K=%lld but layer %d produces N=%lld",
                   l, (long long)in.K, l - 1, (long long)layers_in[l - 1].N);
        const int64_t kc = (in.K + 63) / 64;
        TB_REQUIRE(in.ld_packed % 16 == 0 && in.ld_packed >= kc * 16 && (uintptr_t)in.packed % 16 == 0,
                   "chain_plan_build: layer %d: ld_packed must be a multiple of 16, >= 16*ceil(K/64), base 16-byte aligned", l);
        TB_REQUIRE(in.act >= 0 && in.act <= TB200_ACT_RELU, "chain_plan_build: layer %d: unknown activation %d", l, in.act);
        ChainLayer& o = out[l];
        o.one_box = 0;
        int rc = TB200_ECUDA;
        if (in.ld_packed % CH_SEG == 0) {
            uint64_t dims4[4] = {CH_SEG / 4, (uint64_t)in.N, (uint64_t)(in.ld_packed / CH_SEG), 1};
            uint64_t str4[3] = {(uint64_t)in.ld_packed, (uint64_t)CH_SEG, (uint64_t)(in.ld_packed * in.N)};
            uint32_t box4[4] = {CH_SEG / 4, CH_ROWS, 8, 1};
            rc = make_tensor_map(&o.map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, in.packed, dims4, str4, box4, CU_TENSOR_MAP_SWIZZLE_128B);
            o.one_box = rc == TB200_OK;
        }
        if (rc) {
            uint64_t dims[3] = {(uint64_t)(in.ld_packed / 4), (uint64_t)in.N, 1};
            uint64_t str[2] = {(uint64_t)in.ld_packed, (uint64_t)(in.ld_packed * in.N)};
            uint32_t box[3] = {CH_SEG / 4, CH_ROWS, 1};
            rc = make_tensor_map(&o.map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, in.packed, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
            if (rc) return rc;
        }
        o.scale = in.scale; o.bias = in.bias;
        o.N = (int32_t)in.N; o.K = (int32_t)in.K; o.nks = ch_nks(in.K);
        o.row_blocks = (int32_t)((in.N + CH_ROWS - 1) / CH_ROWS);
        o.clip_val = fabsf(in.clip_val); o.act = in.act;
        // inter-layer buffer: two sets (even / odd call) of M_max rows of N rounded up to whole 16-element groups
        o.h_pitch = (int32_t)ch_round_up(in.N, CH_ROWS);
        o.h_off = off;
        o.h_set_stride = ch_round_up((int64_t)M_max * o.h_pitch * act_elem_size, 256);
        if (l + 1 < n_layers) off += o.h_set_stride * 2;
        if (o.nks > max_nks) max_nks = o.nks;
        if (o.row_blocks > max_rb) max_rb = o.row_blocks;
    }
    hd->magic = CH_MAGIC; hd->n_layers = (uint32_t)n_layers;
    hd->max_nks = max_nks; hd->max_rb = max_rb; hd->M_max = (int32_t)M_max; hd->esz = act_elem_size;
    hd->ws_bytes = off;
    *workspace_bytes = (size_t)off;
    return TB200_OK;
}

template <typename XT, int M>
static int launch_chain(const ChainHeader* plan_dev, const ChainParams& p, int grid, size_t smem, cudaStream_t s) {
    auto kern = gemv_chain_kernel<XT, M>;
    static SmemOptIn optin;
    TB_CUDA(ensure_dynamic_smem(kern, optin, smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3(CH_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, plan_dev, p));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}
template <typename XT>
static int launch_chain_m(const ChainHeader* plan_dev, const ChainParams& p, int M, int grid, size_t smem, cudaStream_t s) {
    switch (M) {
        case 1: return launch_chain<XT, 1>(plan_dev, p, grid, smem, s);
        case 2: return launch_chain<XT, 2>(plan_dev, p, grid, smem, s);
        case 3: return launch_chain<XT, 3>(plan_dev, p, grid, smem, s);
        case 4: return launch_chain<XT, 4>(plan_dev, p, grid, smem, s);
        case 5: return launch_chain<XT, 5>(plan_dev, p, grid, smem, s);
        case 6: return launch_chain<XT, 6>(plan_dev, p, grid, smem, s);
        case 7: return launch_chain<XT, 7>(plan_dev, p, grid, smem, s);
        default: return launch_chain<XT, 8>(plan_dev, p, grid, smem, s);
    }
}

// plan_host: the blob tb200_chain_plan_build filled (read on the host for shapes); plan_dev: its device copy.
extern "C" int tb200_chain_forward(const void* plan_host, const void* plan_dev, const void* x, int x_dtype, int64_t M, int64_t ldx,
                                   void* out, int out_dtype, int64_t ldo, void* workspace, int flags, void* stream) {
    TB_DEVICE_OR_RETURN();
    TB_REQUIRE(plan_host && plan_dev && x && out && workspace, "chain_forward: null pointer");
    const ChainHeader* hd = (const ChainHeader*)plan_host;
    TB_REQUIRE(hd->magic == CH_MAGIC, "chain_forward: plan_host is not a plan built by tb200_chain_plan_build");
    if (M == 0) return TB200_OK;
    TB_REQUIRE(M >= 1 && M <= hd->M_max, "chain_forward: M=%lld exceeds the plan's M_max=%d", (long long)M, hd->M_max);
    TB_REQUIRE(dtype_ok(x_dtype) && dtype_ok(out_dtype), "chain_forward: bad dtype");
    TB_REQUIRE((int)dtype_size(x_dtype) == hd->esz, "chain_forward: x dtype size %d differs from the plan's activation element size %d",
               (int)dtype_size(x_dtype), hd->esz);
    const ChainLayer* layers = (const ChainLayer*)(hd + 1);
    const int L = (int)hd->n_layers;
    TB_REQUIRE(ldx >= layers[0].K && ldo >= layers[L - 1].N, "chain_forward: ldx / ldo smaller than the first K / last N");
    TB_REQUIRE((uintptr_t)plan_dev % 64 == 0 && (uintptr_t)workspace % 256 == 0, "chain_forward: plan_dev must be 64-byte and workspace 256-byte aligned");
    ChainParams p = {};
    p.x = x; p.ldx = ldx; p.out = out; p.ldo = ldo; p.out_dtype = out_dtype;
    p.ws = (uint8_t*)workspace;
    p.img_bytes = (uint32_t)(hd->max_nks * 4096 * (int)M);
    p.pdl = flags & TB200_FLAG_PDL;
    p.trace = kAblate ? (long long*)tb_knob("TB200_CHAIN_TRACE_PTR", 0) : nullptr;
    p.gate = (int)tb_knob("TB200_CHAIN_GATE", 1);
    p.dbg = kAblate ? (int)tb_knob("TB200_CHAIN_DBG", 0) : 0;
    // ring depth: as many 16 KB stages as fit beside the image and the slots (bytes in flight per SM)
    p.n_layers = L;
    p.nslots = M <= 4 ? 2 * CH_AUX : CH_AUX;     // a multiple of the aux-warp count: slot q is always served by aux warp q % CH_AUX
    int st = CH_MAX_STAGES;
    while (st > 2 && chain_smem(L, st, p.nslots, p.img_bytes, (int)M) > 227 * 1024) --st;
    if (chain_smem(L, st, p.nslots, p.img_bytes, (int)M) > 227 * 1024) {
        set_error("chain_forward: M=%lld with K up to %d does not fit the shared-memory activation image", (long long)M, hd->max_nks * 4096);
        return TB200_ENOTSUP;
    }
    p.stages = st;
    const int grid = hd->max_rb < sm_count() ? hd->max_rb : sm_count();
    const size_t smem = chain_smem(L, p.stages, p.nslots, p.img_bytes, (int)M);
    switch (x_dtype) {
        case TB200_F32:  return launch_chain_m<float>((const ChainHeader*)plan_dev, p, (int)M, grid, smem, (cudaStream_t)stream);
        case TB200_BF16: return launch_chain_m<__nv_bfloat16>((const ChainHeader*)plan_dev, p, (int)M, grid, smem, (cudaStream_t)stream);
        default:         return launch_chain_m<__half>((const ChainHeader*)plan_dev, p, (int)M, grid, smem, (cudaStream_t)stream);
    }
}
