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
constexpr int CH_MMA_WARPS = 8, CH_AUX = 3;     // 12 warps = 3 warpgroups: two of MMA warps, one of producer + aux warps
constexpr int CH_WARP_PRODUCER = CH_MMA_WARPS, CH_WARP_AUX0 = CH_MMA_WARPS + 1;
constexpr int CH_THREADS = (CH_MMA_WARPS + 1 + CH_AUX) * 32;
constexpr int CH_MMA_THREADS = CH_MMA_WARPS * 32;
constexpr int CH_MAX_STAGES = 12, CH_MAX_SLOTS = 8;
constexpr int CH_SLOT_U4 = 4 * 32;                 // four K-partials (one team) x 32 lanes, uint4 each
constexpr uint32_t CH_MAGIC = 0x7b200c42u;
constexpr int CH_REPLICAS = 16;                    // copies of every inter-layer activation buffer

// device-side layer descriptor inside the plan blob (192 bytes, 64-byte aligned: the tensor map is used in place)
struct ChainLayer {
    CUtensorMap map;                               // 128 B
    const float* scale; const float* bias;
    int32_t N, K, nks, row_blocks;
    float clip_val; int32_t act, one_box, h_pitch; // h_pitch: flagged 8-byte WORDS per token of this layer's output buffer
    int64_t h_off;                                 // byte offset of that buffer in the workspace (last layer: unused)
    int64_t h_rep_stride;                          // words between the replicas of that buffer
};
static_assert(sizeof(ChainLayer) == 192, "plan layout");
struct ChainHeader {
    uint32_t magic, n_layers;
    int32_t max_nks, max_rb, M_max, esz;           // esz: bytes per inter-layer activation element
    int64_t ws_bytes;
    int32_t replicas;
    char pad[28];
};
static_assert(sizeof(ChainHeader) == 64, "plan layout");
// control block at the start of the workspace (zero before the first call; the kernel maintains it)
// every layer's counter sits on its own 128-byte line (pollers of one layer do not queue behind arrivals of another)
constexpr int CH_CNT_STRIDE = 32;
struct ChainCtrl { uint32_t gen, done, pad[30]; uint32_t counter[TB200_CHAIN_MAX_LAYERS * CH_CNT_STRIDE]; };
constexpr size_t CH_CTRL_BYTES = TB200_CHAIN_CTRL_BYTES;
static_assert(sizeof(ChainCtrl) <= CH_CTRL_BYTES, "control block");

struct ChainParams {
    const void* x; int64_t ldx;
    void* out; int64_t ldo; int out_dtype;
    uint8_t* ws;
    int M, stages, nslots;
    uint32_t img_bytes;
    int pdl;
    int inflight;         // tiles of one SM allowed to be outstanding in the memory system (0: the whole ring)
    int replicas;         // copies of every inter-layer buffer (see the consumer side)
    long long* trace;     // experiment build only: per CTA and layer, globaltimer stamps {enter, input ready, image built, last tile}
};
__device__ __forceinline__ long long chain_gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define CH_STAMP(l, k) do { if (kAblate && p.trace && ct == 0) p.trace[((int64_t)b * TB200_CHAIN_MAX_LAYERS + (l)) * 8 + (k)] = chain_gtimer(); } while (0)
#define CH_STAMP_AUX(l, k) do { if (kAblate && p.trace && lane == 0) p.trace[((int64_t)b * TB200_CHAIN_MAX_LAYERS + (l)) * 8 + (k)] = chain_gtimer(); } while (0)

// per-layer facts the warps need, copied to shared memory once
struct LayerInfo {
    const float* scale; const float* bias;
    uint8_t* hout;                                 // this layer's output rows (inter-layer buffer), null for the last layer
    int64_t rep_stride;                            // words between its replicas
    int N, K, nks, RB;
    float clip; int act, pitch, one_box;
};

__device__ __forceinline__ void imma_c(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
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
// one fence + one relaxed reduction by ONE lane after a __syncwarp(): the barrier orders the other lanes' stores
// before the fence (cumulativity), as in a cooperative grid barrier.  (__threadfence() is fence.sc here and
// red.release adds a second MEMBAR: 2 x MEMBAR + CCTL.IVALL per row block in the first version.)
__device__ __forceinline__ void publish_add(uint32_t* p, uint32_t v) {
    asm volatile("fence.acq_rel.gpu;\n\tred.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ ulonglong2 ld_relaxed_u64x2(const uint64_t* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_add(uint32_t* p, uint32_t v) {
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// 16 bytes from L2 (never from a possibly stale L1 line: the buffers are rewritten by other SMs during the kernel)
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

__device__ __forceinline__ int chain_img_vec(int sg, int pp, int i, int n, int t, int Mg) {
    return ((((sg * 2 + pp) * 4 + i) * Mg + n) << 2) + t;
}

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

template <typename XT>
__global__ void __launch_bounds__(CH_THREADS, 1)
gemv_chain_kernel(const ChainHeader* __restrict__ plan, const ChainParams p) {
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    // ring[stages][16 KB] | image[img_bytes] | slots[nslots][128] uint4 | red_f[64] | red_i[64] | sx[16] | layers | barriers
    const uint32_t sRing = base;
    uint32_t* xq = reinterpret_cast<uint32_t*>(gbase + (size_t)p.stages * CH_TILE);
    uint4* slots = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(xq) + ((p.img_bytes + 127u) & ~127u));
    float* red_f = reinterpret_cast<float*>(slots + (size_t)p.nslots * CH_SLOT_U4);
    int* red_i = reinterpret_cast<int*>(red_f + 64);
    float* sx_sm = reinterpret_cast<float*>(red_i + 64);
    LayerInfo* lay = reinterpret_cast<LayerInfo*>(sx_sm + 16);
    const int L = (int)plan->n_layers;
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
        li.rep_stride = cl.h_rep_stride;
        li.N = cl.N; li.K = cl.K; li.nks = cl.nks; li.RB = cl.row_blocks;
        li.clip = cl.clip_val; li.act = cl.act; li.pitch = cl.h_pitch; li.one_box = cl.one_box;
        lay[l] = li;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), 4); }
        for (int q = 0; q < p.nslots; ++q) { mbar_init(bar_sfull(q), 4); mbar_init(bar_sempty(q), 1); }
        mbar_fence_init();
    }
    __syncthreads();
    if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");
    // register budget by role (per warpgroup): the MMA warps hold 64 registers of B fragments, 32 of accumulators and 32 of
    // A fragments; the producer and the aux warps need few
    if (warp < CH_MMA_WARPS) asm volatile("setmaxnreg.inc.sync.aligned.u32 200;");
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 104;");

    if (warp == CH_WARP_PRODUCER) {
        // ===================== weight tiles of ALL layers: independent of every activation =====================
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            int sl = 0; uint32_t phl = 0; int issued = 0;       // trailing position: the tile `inflight` issues back
            Walk w; w.start(lay, L, b);
            while (w.l < L) {
                const CUtensorMap* map = &layers[w.l].map;
                const int nks = lay[w.l].nks, one_box = lay[w.l].one_box;
                for (int ks = 0; ks < nks; ++ks) {
                    mbar_wait_relaxed(bar_empty(s), ph ^ 1u);
                    // at most `inflight` tiles of this SM outstanding in the memory system: a deeper queue adds no bandwidth
                    // (the ring still fills up with LANDED tiles) but delays every other load and store of the kernel -- the
                    // flag and activation traffic of the layer boundaries -- by the time it takes to drain
                    if (p.inflight > 0 && issued >= p.inflight) {
                        mbar_wait_relaxed(bar_full(sl), phl);
                        if (++sl == p.stages) { sl = 0; phl ^= 1u; }
                    }
                    ++issued;
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
    } else if (warp < CH_MMA_WARPS) {
        // ===================== activation image of every layer + dot products =====================
        const int ct = threadIdx.x;                          // 0..255
        const int g = lane >> 2, t = lane & 3;
        const int team = warp >> 2, wq = warp & 3, wseg = 2 * wq;
        const uint32_t sSeg = wseg * (CH_ROWS * CH_SEG) + g * CH_SEG;
        const uint32_t phys0 = ((2 * t) ^ g) << 4, phys1 = ((2 * t + 1) ^ g) << 4;
        const int M = p.M, stages = p.stages, nslots = p.nslots;
        if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // x (and the control block) come from earlier kernels
        const uint32_t gen = ld_acquire_u32(&ctrl->gen);
        int s = 0; uint32_t ph = 0;                          // weight ring (global tile sequence of this CTA)
        int sq = 0; uint32_t spar = 0;                       // slot ring
        int acc[2][4][4];                                    // [segment][field][fragment]: eight independent IMMA chains
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[j][i][e] = 0;
        uint4 b2[2][2][4];                                   // B fragments of this warp's two segments for the current k-slice
        const uint4* xqv = reinterpret_cast<const uint4*>(xq);
        auto load_b = [&](int ks) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        b2[j][pp][i] = g < M ? xqv[chain_img_vec(ks * 8 + wseg + j, pp, i, g, t, M)] : make_uint4(0u, 0u, 0u, 0u);
        };
        int cur = -1;
        Walk w; w.start(lay, L, b);
        while (w.l < L) {
            const LayerInfo& li = lay[w.l];
            if (w.l != cur) {
                // ---------- entering layer w.l: its input must be complete, then quantise it into the image ----------
                cur = w.l;
                const int l = cur;
                CH_STAMP(l, 0);
                const int K = li.K, nks = li.nks;
                const uint32_t want = gen + 1u;              // flag value of this call's inter-layer words
                const XT* xin = reinterpret_cast<const XT*>(p.x);
                const uint64_t* hw = nullptr; int64_t pitch_w = 0;
                const bool vec0 = ((reinterpret_cast<uintptr_t>(p.x) | (uintptr_t)(p.ldx * sizeof(XT))) & 15) == 0;
                if (l > 0) {
                    const LayerInfo& pl = lay[l - 1];
                    // every CTA reads the WHOLE previous output at the same moment: 148 readers of the same L2 lines are served one
                    // after the other (measured: 7 KB/us per SM), so the producers store `replicas` copies and CTA b reads copy b % R
                    hw = reinterpret_cast<const uint64_t*>(pl.hout) + (int64_t)(b % p.replicas) * pl.rep_stride; pitch_w = pl.pitch;
                    if (ct == 0) {
                        // coarse wait: ONE poller per CTA on the layer's row-block counter (relaxed loads; the counter is only
                        // a hint that the data is on its way -- the flags inside the data words are what is verified below)
                        const uint32_t need = want * (uint32_t)pl.RB;
                        const uint32_t* cnt = &ctrl->counter[(l - 1) * CH_CNT_STRIDE];
                        uint32_t spins = 0;
                        while ((int)(ld_relaxed_u32(cnt) - need) < 0) {
                            if (++spins > 50000000u) __trap();          // a protocol bug traps instead of hanging the GPU
                        }
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(CH_MMA_THREADS) : "memory");
                }
                CH_STAMP(l, 1);
                // one 16-element group of token m as fp32 (zero beyond K).  Layer 0 reads the caller's plain activations;
                // later layers read the flagged words the previous layer's epilogue stored: {value(s), flag} per 8 bytes, so
                // no fence is needed between the producer's data and its counter increment -- a word whose flag is not
                // `want` yet simply makes the group invalid and it is fetched again.
                constexpr int EPW = sizeof(XT) == 4 ? 1 : 2;           // elements per flagged word
                auto fetch = [&](int m, int gi, float (&v)[16]) -> bool {
                    if (l == 0) {
                        const XT* src = xin + (int64_t)m * p.ldx + 16 * (int64_t)gi;
                        if (vec0 && 16 * gi + 16 <= K) {
                            constexpr int NV = sizeof(XT) == 4 ? 4 : 2;
#pragma unroll
                            for (int j = 0; j < NV; ++j) {
                                const uint4 r = ld_cg_u4(reinterpret_cast<const uint4*>(src) + j);
                                const XT* e = reinterpret_cast<const XT*>(&r);
#pragma unroll
                                for (int tt = 0; tt < 16 / NV; ++tt) v[j * (16 / NV) + tt] = to_f32(e[tt]);
                            }
                        } else {
#pragma unroll
                            for (int tt = 0; tt < 16; ++tt) v[tt] = (16 * gi + tt < K) ? to_f32(ld_cg_elem<XT>(src + tt)) : 0.0f;
                        }
                        return true;
                    }
                    const uint64_t* src = hw + (int64_t)m * pitch_w + (16 / EPW) * (int64_t)gi;
                    bool ok = true;
                    if (16 * gi + 16 <= K) {
#pragma unroll
                        for (int j = 0; j < 8 / EPW; ++j) {             // 16-byte loads of two flagged words
                            const ulonglong2 r = ld_relaxed_u64x2(src + 2 * j);
                            ok = ok && (uint32_t)(r.x >> 32) == want && (uint32_t)(r.y >> 32) == want;
                            if (EPW == 2) {
                                const uint32_t w0 = (uint32_t)r.x, w1 = (uint32_t)r.y;
                                const XT* e0 = reinterpret_cast<const XT*>(&w0);
                                const XT* e1 = reinterpret_cast<const XT*>(&w1);
                                v[4 * j] = to_f32(e0[0]); v[4 * j + 1] = to_f32(e0[1]); v[4 * j + 2] = to_f32(e1[0]); v[4 * j + 3] = to_f32(e1[1]);
                            } else {
                                v[2 * j] = __uint_as_float((uint32_t)r.x); v[2 * j + 1] = __uint_as_float((uint32_t)r.y);
                            }
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 16 / EPW; ++j) {            // ragged tail: word by word
                            const bool in_k = 16 * gi + EPW * j < K;
                            const uint64_t r = in_k ? ld_relaxed_u64(src + j) : ((uint64_t)want << 32);
                            ok = ok && (uint32_t)(r >> 32) == want;
                            const uint32_t w0 = (uint32_t)r;
                            if (EPW == 2) {
                                const XT* e0 = reinterpret_cast<const XT*>(&w0);
                                v[2 * j] = to_f32(e0[0]);
                                v[2 * j + 1] = (16 * gi + 2 * j + 1 < K) ? to_f32(e0[1]) : 0.0f;
                            } else {
                                v[j] = __uint_as_float(w0);
                            }
                        }
                    }
                    return ok;
                };
                constexpr int CACHE = 3;                     // groups per thread kept in registers between the two passes
                const int tot = M * nks;                     // groups of this thread
                const bool cached = tot <= CACHE;
                // NOTE: everything below indexes its register arrays with compile-time constants only.  With ~225 KB of the
                // SM carved out as shared memory there is practically no L1 left, so a local-memory (stack) array costs an L2
                // round trip per access: the first version kept the cached groups in a dynamically indexed array and spent
                // 2.3 us (K = 4096) / 5.6 us (K = 11008) per layer right here.
                auto quantise_group = [&](const float (&v)[16], float inv, int m, int gi, int& ls) {
                    uint32_t qb[16];              // low byte = two's-complement int8 code (rint via the 1.5 * 2^23 trick)
#pragma unroll
                    for (int tt = 0; tt < 16; ++tt) {
                        float q = __fmul_rn(v[tt], inv);
                        q = fminf(fmaxf(q, -127.0f), 127.0f);
                        qb[tt] = __float_as_uint(__fadd_rn(q, 12582912.0f));
                    }
                    const int c = gi >> 2, jw = gi & 3, lc = c & 7;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {   // word i = elements i, 4+i, 8+i, 12+i (field i of the 4 bytes of packed word gi)
                        const uint32_t wd = __byte_perm(__byte_perm(qb[i], qb[4 + i], 0x0040), __byte_perm(qb[8 + i], qb[12 + i], 0x0040), 0x5410);
                        ls = __dp4a((int)wd, 0x01010101, ls);
                        xq[chain_img_vec(c >> 3, lc & 1, i, m, lc >> 1, M) * 4 + jw] = wd;
                    }
                };
                auto token_scale = [&](int m) -> float {     // after the barrier: absmax of token m -> 127/amax; s_x to shared
                    const float4 r0 = *reinterpret_cast<const float4*>(red_f + m * 8);
                    const float4 r1 = *reinterpret_cast<const float4*>(red_f + m * 8 + 4);
                    const float vmax = fmaxf(fmaxf(fmaxf(r0.x, r0.y), fmaxf(r0.z, r0.w)), fmaxf(fmaxf(r1.x, r1.y), fmaxf(r1.z, r1.w)));
                    if (ct == 0) sx_sm[m] = vmax > 0.0f ? __fdiv_rn(vmax, 127.0f) : 1.0f;
                    return vmax > 0.0f ? __fdiv_rn(127.0f, vmax) : 1.0f;
                };
                if (cached) {
                    float vc[CACHE][16];
                    // all loads of the thread are issued before the first use (ONE L2 round trip), and again until every flag is set
                    uint32_t spins = 0;
                    bool ok;
                    do {
                        ok = true;
#pragma unroll
                        for (int c = 0; c < CACHE; ++c)
                            if (c < tot) { const int m = c / nks, u = c - m * nks; ok = fetch(m, u * 256 + ct, vc[c]) && ok; }
                        if (!ok && ++spins > 20000000u) __trap();
                    } while (!ok);
                    CH_STAMP(l, 6);
                    if (kAblate && p.trace && ct == 0) p.trace[((int64_t)b * TB200_CHAIN_MAX_LAYERS + l) * 8 + 7] = spins;
                    for (int m = 0; m < M; ++m) {            // pass A: per-token absmax
                        float amax = 0.0f;
#pragma unroll
                        for (int c = 0; c < CACHE; ++c)
                            if (c < tot && c / nks == m) {
#pragma unroll
                                for (int tt = 0; tt < 16; ++tt) amax = fmaxf(amax, fabsf(vc[c][tt]));
                            }
                        const uint32_t am = __reduce_max_sync(0xffffffffu, __float_as_uint(amax));
                        if (lane == 0) red_f[m * 8 + warp] = __uint_as_float(am);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(CH_MMA_THREADS) : "memory");
                    for (int m = 0; m < M; ++m) {            // pass B: quantise into the kperm image (zero beyond K), code sums
                        const float inv = token_scale(m);
                        int ls = 0;
#pragma unroll
                        for (int c = 0; c < CACHE; ++c)
                            if (c < tot && c / nks == m) quantise_group(vc[c], inv, m, (c - m * nks) * 256 + ct, ls);
                        ls = __reduce_add_sync(0xffffffffu, ls);
                        if (lane == 0) red_i[m * 8 + warp] = ls;
                    }
                } else {
                    for (int m = 0; m < M; ++m) {            // pass A (the data is fetched again in pass B: L2)
                        float amax = 0.0f;
                        for (int u = 0; u < nks; ++u) {
                            float v[16];
                            uint32_t spins = 0;
                            while (!fetch(m, u * 256 + ct, v)) { if (++spins > 20000000u) __trap(); }
#pragma unroll
                            for (int tt = 0; tt < 16; ++tt) amax = fmaxf(amax, fabsf(v[tt]));
                        }
                        const uint32_t am = __reduce_max_sync(0xffffffffu, __float_as_uint(amax));
                        if (lane == 0) red_f[m * 8 + warp] = __uint_as_float(am);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(CH_MMA_THREADS) : "memory");
                    for (int m = 0; m < M; ++m) {
                        const float inv = token_scale(m);
                        int ls = 0;
                        for (int u = 0; u < nks; ++u) {
                            float v[16];
                            (void)fetch(m, u * 256 + ct, v);           // verified in pass A
                            quantise_group(v, inv, m, u * 256 + ct, ls);
                        }
                        ls = __reduce_add_sync(0xffffffffu, ls);
                        if (lane == 0) red_i[m * 8 + warp] = ls;
                    }
                }
                // image, sx_sm and red_i (per-warp code sums, added up by the aux warps) are complete: every MMA warp passes
                // this barrier before its first slot hand-off of the layer, which is what the aux warps wait for
                asm volatile("bar.sync 1, %0;" ::"n"(CH_MMA_THREADS) : "memory");
                load_b(0);
                CH_STAMP(l, 2);
            }
            // ---------- one row block: team (n & 1) computes it, the other team only advances its ring position ----------
            const int nks = li.nks;
            if ((w.n & 1) == team) {
#pragma unroll 1
                for (int k2 = 0; k2 < nks; ++k2) {
                    if (nks > 1) load_b(k2);          // K > 4096: the fragments of this k-slice (a single slice stays in registers)
                    mbar_wait(bar_full(s), ph);
                    const uint8_t* seg = gbase + s * CH_TILE + sSeg;
                    uint4 alo[2][2], ahi[2][2];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const uint8_t* sj = seg + j * (CH_ROWS * CH_SEG);
                        alo[j][0] = *reinterpret_cast<const uint4*>(sj + phys0);
                        ahi[j][0] = *reinterpret_cast<const uint4*>(sj + 8 * CH_SEG + phys0);
                        alo[j][1] = *reinterpret_cast<const uint4*>(sj + phys1);
                        ahi[j][1] = *reinterpret_cast<const uint4*>(sj + 8 * CH_SEG + phys1);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_empty(s));            // the tile lives in registers now
                    if (++s == stages) { s = 0; ph ^= 1u; }
#pragma unroll
                    for (int j = 0; j < 2; ++j)
#pragma unroll
                        for (int pp = 0; pp < 2; ++pp) {
                            const uint4 lo = alo[j][pp], hi = ahi[j][pp];
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const uint32_t mk = 0x03030303u << (2 * i);
                                imma_c(acc[j][i], lo.x & mk, hi.x & mk, lo.y & mk, hi.y & mk, b2[j][pp][i].x, b2[j][pp][i].y);
                            }
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const uint32_t mk = 0x03030303u << (2 * i);
                                imma_c(acc[j][i], lo.z & mk, hi.z & mk, lo.w & mk, hi.w & mk, b2[j][pp][i].z, b2[j][pp][i].w);
                            }
                        }
                }
                // recombine the four fields (exact: set i is a multiple of 4^i) and park this warp's K-partial
                mbar_wait(bar_sempty(sq), spar ^ 1u);
                uint32_t vr[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    vr[e] = (uint32_t)((acc[0][0][e] + acc[1][0][e]) + ((acc[0][1][e] + acc[1][1][e]) >> 2) +
                                       ((acc[0][2][e] + acc[1][2][e]) >> 4) + ((acc[0][3][e] + acc[1][3][e]) >> 6));
                }
                slots[(size_t)sq * CH_SLOT_U4 + wq * 32 + lane] = make_uint4(vr[0], vr[1], vr[2], vr[3]);
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[j][i][e] = 0;
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_sfull(sq));
                CH_STAMP(w.l, 3);
            } else {
                s += nks;                                                // the other team's tiles
                while (s >= stages) { s -= stages; ph ^= 1u; }
            }
            if (++sq == nslots) { sq = 0; spar ^= 1u; }
            w.next(lay, L, b, G);
        }
    } else {
        // ===================== reduction of the K-partials + fused epilogue (+ activation) =====================
        const int a = warp - CH_WARP_AUX0;
        const int g = lane >> 2, t = lane & 3;
        if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // earlier kernels may still read out / the h buffers
        const uint32_t gen = ld_acquire_u32(&ctrl->gen);
        Walk w; w.start(lay, L, b);
        while (w.l < L) {
            if ((w.n % CH_AUX) == a) {
                const LayerInfo& li = lay[w.l];
                const int l = w.l;
                const int sq_i = w.n % p.nslots;
                const uint32_t spar = (uint32_t)(w.n / p.nslots) & 1u;
                const int64_t row0 = (int64_t)w.rb * CH_ROWS + g, row1 = row0 + 8;
                float sc0 = 1.0f, sc1 = 1.0f, bi0 = 0.0f, bi1 = 0.0f;         // requested before the wait for the partial sums
                if (li.scale) { if (row0 < li.N) sc0 = __ldg(li.scale + row0); if (row1 < li.N) sc1 = __ldg(li.scale + row1); }
                if (li.bias) { if (row0 < li.N) bi0 = __ldg(li.bias + row0); if (row1 < li.N) bi1 = __ldg(li.bias + row1); }
                // scale / bias are read once per call (cold in L2) and their loads would queue behind the weight stream right
                // when the layer's last row blocks wait for them: ask L2 for the same rows of the next two layers now
                if (lane < 4) {
                    const int l2 = l + 1 + (lane >> 1);
                    if (l2 < L && w.rb < lay[l2].RB) {
                        const float* ptr = (lane & 1) ? lay[l2].bias : lay[l2].scale;
                        if (ptr) prefetch_l2(ptr + (int64_t)w.rb * CH_ROWS);
                    }
                }
                mbar_wait(bar_sfull(sq_i), spar);
                int vv[4] = {0, 0, 0, 0};
                const uint4* sl = slots + (size_t)sq_i * CH_SLOT_U4 + lane;
#pragma unroll
                for (int wv = 0; wv < 4; ++wv) {
                    const uint4 v = sl[wv * 32];
                    vv[0] += (int)v.x; vv[1] += (int)v.y; vv[2] += (int)v.z; vv[3] += (int)v.w;
                }
                float sxv[2]; int sumq[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int tok = 2 * t + e;
                    sxv[e] = 0.0f; sumq[e] = 0;
                    if (tok < p.M) {
                        sxv[e] = sx_sm[tok];
                        const int4 a0 = *reinterpret_cast<const int4*>(red_i + tok * 8);
                        const int4 a1 = *reinterpret_cast<const int4*>(red_i + tok * 8 + 4);
                        sumq[e] = a0.x + a0.y + a0.z + a0.w + a1.x + a1.y + a1.z + a1.w;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_sempty(sq_i));                // partial sums and token scalars are in registers
                // fragment: vv0 (row g, tok 2t), vv1 (row g, tok 2t+1), vv2 (row g+8, tok 2t), vv3 (row g+8, tok 2t+1)
                float yv[4];
#pragma unroll
                for (int e4 = 0; e4 < 4; ++e4) {
                    const int e = e4 & 1, hi = e4 >> 1;
                    const int av = vv[e4] - sumq[e];
                    const float tv = __fmul_rn((float)av, sxv[e]);
                    float y = epilogue(tv, hi ? sc1 : sc0, hi ? bi1 : bi0, li.clip);
                    // the activation function sees the value in the STORED (activation dtype) precision
                    if (li.act == TB200_ACT_GELU_TANH) y = gelu_tanh(round_to<XT>(y));
                    else if (li.act == TB200_ACT_RELU) y = fmaxf(y, 0.0f);
                    yv[e4] = y;
                }
                CH_STAMP_AUX(l, 4);
                if (li.hout) {
                    // inter-layer activations as flagged 8-byte words {value(s) in the activation dtype, flag = generation}:
                    // a single-copy-atomic store carries data and validity together, so no fence orders data before the counter
                    const uint32_t want = gen + 1u;
                    uint64_t* hw = reinterpret_cast<uint64_t*>(li.hout);
                    if (sizeof(XT) == 4) {
#pragma unroll
                        for (int e4 = 0; e4 < 4; ++e4) {
                            const int tok = 2 * t + (e4 & 1);
                            const int64_t row = (e4 >> 1) ? row1 : row0;
                            if (row < li.N && tok < p.M) {
                                uint64_t* dst = hw + (int64_t)tok * li.pitch + row;
                                for (int r = 0; r < p.replicas; ++r) st_relaxed_u64(dst + r * li.rep_stride, ((uint64_t)want << 32) | __float_as_uint(yv[e4]));
                            }
                        }
                    } else {
                        // a word holds two consecutive rows of one token: lanes g and g^1 exchange halves.  Even g stores
                        // rows (g, g+1), odd g stores rows (g+7, g+8), for tokens 2t and 2t+1.
                        const bool odd = g & 1;
                        const float s0 = __shfl_xor_sync(0xffffffffu, odd ? yv[0] : yv[2], 4);
                        const float s1 = __shfl_xor_sync(0xffffffffu, odd ? yv[1] : yv[3], 4);
                        const int64_t rlo = odd ? row0 + 7 : row0;             // first (even) row of my word
                        const float lo[2] = {odd ? s0 : yv[0], odd ? s1 : yv[1]};
                        const float hi2[2] = {odd ? yv[2] : s0, odd ? yv[3] : s1};
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int tok = 2 * t + e;
                            if (rlo < li.N && tok < p.M) {
                                const XT v0 = from_f32<XT>(lo[e]);
                                const XT v1 = from_f32<XT>(rlo + 1 < li.N ? hi2[e] : 0.0f);
                                const uint32_t pair = (uint32_t)*reinterpret_cast<const unsigned short*>(&v0) |
                                                      ((uint32_t)*reinterpret_cast<const unsigned short*>(&v1) << 16);
                                uint64_t* dst = hw + (int64_t)tok * li.pitch + (rlo >> 1);
                                for (int r = 0; r < p.replicas; ++r) st_relaxed_u64(dst + r * li.rep_stride, ((uint64_t)want << 32) | pair);
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) red_relaxed_add(&ctrl->counter[l * CH_CNT_STRIDE], 1u);     // hint for the pollers
                } else {
#pragma unroll
                    for (int e4 = 0; e4 < 4; ++e4) {
                        const int tok = 2 * t + (e4 & 1);
                        const int64_t row = (e4 >> 1) ? row1 : row0;
                        if (row < li.N && tok < p.M) {
                            const int64_t o = (int64_t)tok * p.ldo + row;
                            if (p.out_dtype == TB200_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(yv[e4]);
                            else if (p.out_dtype == TB200_F16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(yv[e4]);
                            else reinterpret_cast<float*>(p.out)[o] = yv[e4];
                        }
                    }
                }
                CH_STAMP_AUX(l, 5);
            }
            w.next(lay, L, b, G);
        }
    }
    // ---- the last CTA to finish advances the generation (counters are cumulative across calls, never reset) ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const uint32_t old = atomicAdd(&ctrl->done, 1u);
        if (old == (uint32_t)G - 1u) {
            ctrl->done = 0u;
            __threadfence();
            atomicAdd(&ctrl->gen, 1u);
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------------
static inline int64_t ch_round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }
static inline int ch_nks(int64_t K) { return (int)((K + 4095) / 4096); }

static size_t chain_smem(int n_layers, int stages, int nslots, uint32_t img_bytes) {
    return 1024 + (size_t)stages * CH_TILE + ((img_bytes + 127u) & ~127u) + (size_t)nslots * CH_SLOT_U4 * 16 +
           64 * 4 * 2 + 8 * 4 * 2 + (size_t)n_layers * sizeof(LayerInfo) + (2 * CH_MAX_STAGES + 2 * CH_MAX_SLOTS) * 8 + 64;
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
        // inter-layer buffer: per token ceil(N / elements-per-word) flagged words {value(s), flag}, rows 16-byte aligned
        o.h_pitch = (int32_t)ch_round_up((in.N + (act_elem_size == 4 ? 0 : 1)) / (act_elem_size == 4 ? 1 : 2), 2);
        o.h_off = off;
        o.h_rep_stride = ch_round_up((int64_t)M_max * o.h_pitch * 8, 256) / 8;
        if (l + 1 < n_layers) off += o.h_rep_stride * 8 * CH_REPLICAS;
        if (o.nks > max_nks) max_nks = o.nks;
        if (o.row_blocks > max_rb) max_rb = o.row_blocks;
    }
    hd->magic = CH_MAGIC; hd->n_layers = (uint32_t)n_layers;
    hd->max_nks = max_nks; hd->max_rb = max_rb; hd->M_max = (int32_t)M_max; hd->esz = act_elem_size;
    hd->ws_bytes = off;
    hd->replicas = CH_REPLICAS;
    *workspace_bytes = (size_t)off;
    return TB200_OK;
}

template <typename XT>
static int launch_chain(const ChainHeader* plan_dev, const ChainParams& p, int n_layers, int grid, size_t smem, cudaStream_t s) {
    auto kern = gemv_chain_kernel<XT>;
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
    p.ws = (uint8_t*)workspace; p.M = (int)M;
    p.img_bytes = (uint32_t)(hd->max_nks * 4096 * (int)M);
    p.pdl = flags & TB200_FLAG_PDL;
    p.trace = kAblate ? (long long*)tb_knob("TB200_CHAIN_TRACE_PTR", 0) : nullptr;
    p.inflight = (int)tb_knob("TB200_CHAIN_INFLIGHT", 4);
    p.replicas = (int)tb_knob("TB200_CHAIN_REPLICAS", hd->replicas);
    if (p.replicas < 1 || p.replicas > hd->replicas) p.replicas = hd->replicas;
    // ring depth: as many 16 KB stages as fit beside the image and the slots (bytes in flight per SM)
    p.nslots = 2 * CH_AUX;     // a multiple of the aux-warp count: slot q is always served by aux warp q % CH_AUX
    int st = CH_MAX_STAGES;
    while (st > 2 && chain_smem(L, st, p.nslots, p.img_bytes) > 227 * 1024) --st;
    if (chain_smem(L, st, p.nslots, p.img_bytes) > 227 * 1024) {
        set_error("chain_forward: M=%lld with K up to %d does not fit the shared-memory activation image", (long long)M, hd->max_nks * 4096);
        return TB200_ENOTSUP;
    }
    p.stages = st;
    const int grid = hd->max_rb < sm_count() ? hd->max_rb : sm_count();
    const size_t smem = chain_smem(L, p.stages, p.nslots, p.img_bytes);
    switch (x_dtype) {
        case TB200_F32:  return launch_chain<float>((const ChainHeader*)plan_dev, p, L, grid, smem, (cudaStream_t)stream);
        case TB200_BF16: return launch_chain<__nv_bfloat16>((const ChainHeader*)plan_dev, p, L, grid, smem, (cudaStream_t)stream);
        default:         return launch_chain<__half>((const ChainHeader*)plan_dev, p, L, grid, smem, (cudaStream_t)stream);
    }
}
