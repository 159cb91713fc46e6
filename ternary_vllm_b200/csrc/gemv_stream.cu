// gemv_stream.cu — persistent, warp-specialised decode GEMV over a LIST of layers (M <= 32 tokens).
//
// The single-layer kernel (gemv.cu) is latency-bound: activation prologue -> dot products -> reduction run one
// after the other inside every CTA, and a 4096 x 11008 layer is only 1.7 us of HBM time.  This kernel keeps HBM
// busy across layer boundaries instead.  One CTA per SM walks the global list of 16-row blocks of ALL layers
// (block u = layer * row_blocks + rb, CTA b takes u = b, b + grid, ...: balanced to one block over the whole
// list) with specialised warps that never wait for each other's phase:
//   * warp 8  (1 thread): streams the packed weight tiles (16 rows x 1024 B, SWIZZLE_128B) through a ring of
//     shared-memory stages with TMA, continuously across layers;
//   * warp 9  (1 thread): brings the int8 activation image of the next layer (written once per layer by
//     quant_image_kernel below, not by every CTA) into shared memory with cp.async.bulk while the MMA warps are
//     still on the previous layer;
//   * warps 0-7: warp w owns the w-th 128-byte K segment of every row: 4 LDS.128 of packed weights per tile, in-place
//     field masks (4 LOP3 per 16 weights) and mma.sync.m16n8k32.u8.s8 against up to four groups of 8 tokens; the
//     K-partials of a row block are parked in a ring of shared-memory slots;
//   * warps 10-11: sum the eight partials of a slot and run the fused epilogue (-sum_qx, s_x * scale, bias, clip,
//     store) off the critical path of the MMA warps.
// Integer accumulators are the same as the single-layer kernel's (bit-exact); the epilogue arithmetic is identical.
// Reference semantics: ScaledTernaryLinear.forward (binary/modules.py:101-109), ops.ternary_matmul
// (impl/tf/moondream/ternary.py:279-281).
#include <cstdlib>
#include "umma.cuh"

namespace tb200 {

int get_weight_map(const uint8_t* packed, int64_t ld, int64_t N, int64_t stride, int layers, CUtensorMap* out, int* one_box);
int gemv_nks(int64_t K);

constexpr int SV_ROWS = 16;
constexpr int SV_SEG = 128;
constexpr int SV_TILE_BYTES = 16 * 1024;
constexpr int SV_MMA_WARPS = 8;
// warp roles: [0, 8) MMA | 8 tile producer | 9 image producer | 10.. aux
constexpr int sv_threads(int aux) { return (8 + 2 + aux) * 32; }
// TEAMS == 2 with FOUR aux warps serves the tensor-parallel epilogue (its staging + peer / multicast stores do not fit the two aux
// warps' time budget: tools/peer_cost.py, 212 -> 170 us with one local destination against 147 us without the gather).  14 warps at
// 128 registers; a 16-warp form re-balanced with setmaxnreg (MMA warps 160, the rest 96) spilled in the aux loop and was slower.
constexpr int sv_threads2(int teams, int aux) { return sv_threads(aux); }
constexpr int SV_MAX_STAGES = 12;
constexpr int SV_MAX_SLOTS = 8;
constexpr int SV_MAX_IMG = 4;
constexpr int SV_PEER_BUFS = 4;     // staging tiles per aux warp (tensor-parallel epilogue)
constexpr int SV_TAIL = 256;           // blob tail: s_x[32] f32 | sum_qx[32] i32

struct StreamParams {
    const uint8_t* blob; int64_t blob_stride;      // per layer: activation image | s_x[32] | sum_qx[32]
    const float* scale; const float* bias; int64_t stride_sb;
    void* out; int64_t ldo; int64_t stride_out; int out_dtype;
    int32_t* acc_out;
    // tensor-parallel gather fused into the epilogue: when n_peers > 0 every [tokens x 16 rows] output tile is staged in
    // shared memory and stored with ONE TMA tile store per rank into that rank's [layers, M, ldo] buffer (peer-mapped
    // NVLink addresses; tensor maps in PeerMaps, already offset to this rank's column block; TMA clips rows >= N, tokens >= M)
    int n_peers; uint32_t peer_stage_off;     // staging tiles of the aux warps, byte offset in dynamic shared memory
    void* peer_ptrs[8]; int direct;           // direct != 0: 16-byte st.global per piece and rank straight from the staging tile (no TMA)
    void* mc_out;                             // non-null: NVSwitch multicast address of this rank's column block -> multimem stores instead of per-rank TMA stores
    int64_t N;
    int M, Mg, nks, row_blocks, layers;
    int stages, nslots, nimg;
    uint32_t img_bytes, grp_bytes;
    float clip_val;
    int one_box;
    int contig;       // 1: a CTA takes a contiguous chunk of the global row-block list, 0: the strided list b, b + grid, ...
    int dbg;          // experiments: bit 0 skips the MMAs (memory-stream ceiling of the kernel)
};

__device__ __forceinline__ void imma_s(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// uint4 index inside one token group's image: see gemv.cu (xq_vec); Mg = token stride
__device__ __forceinline__ int img_vec(int sg, int pp, int i, int n, int t, int Mg) {
    return ((((sg * 2 + pp) * 4 + i) * Mg + n) << 2) + t;
}

// NG = groups of 8 tokens; BREG = the B fragments of the whole layer live in registers (NG == 1, K <= 4096)
// TEAMS = 2 (NG == 1 only): two teams of 4 MMA warps take alternate row blocks, every warp owning TWO 128-byte K
// segments: the per-row-block work of a warp (barriers, recombination, slot hand-off) is amortised over twice the MMAs
// and only four K-partials per output meet in shared memory.
// PEER: tensor-parallel variant (fused gather); a separate instantiation so that the single-GPU kernel keeps its exact
// code (the kernel is sensitive to the size of the aux warps' loop)
struct NoPeerMaps { int unused; };
template <bool PEER> struct PeerArg { using type = NoPeerMaps; };
template <> struct PeerArg<true> { using type = PeerMaps; };
template <int NG, bool BREG, int TEAMS, int AUX, bool PEER>
__global__ void __launch_bounds__(sv_threads2(TEAMS, AUX), 1)
gemv_stream_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ typename PeerArg<PEER>::type pm, const StreamParams p) {
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    // ring[stages][16 KB] | image[nimg][img_bytes] | slots[nslots][NG][8 warps][32] uint4 | barriers
    const uint32_t sRing = base;
    const uint32_t img_pitch = (p.img_bytes + 127u) & ~127u;
    uint8_t* img0 = gbase + (size_t)p.stages * SV_TILE_BYTES;
    uint4* slots = reinterpret_cast<uint4*>(img0 + (size_t)p.nimg * img_pitch);
    uint64_t* bars = reinterpret_cast<uint64_t*>(slots + (size_t)p.nslots * NG * 256);
    const uint32_t sBar = smem_u32(bars);
    auto bar_full = [&](int s) { return sBar + 8u * s; };
    auto bar_empty = [&](int s) { return sBar + 8u * (SV_MAX_STAGES + s); };
    auto bar_ifull = [&](int b) { return sBar + 8u * (2 * SV_MAX_STAGES + b); };
    auto bar_iempty = [&](int b) { return sBar + 8u * (2 * SV_MAX_STAGES + SV_MAX_IMG + b); };
    auto bar_sfull = [&](int q) { return sBar + 8u * (2 * SV_MAX_STAGES + 2 * SV_MAX_IMG + q); };
    auto bar_sempty = [&](int q) { return sBar + 8u * (2 * SV_MAX_STAGES + 2 * SV_MAX_IMG + SV_MAX_SLOTS + q); };

    constexpr int WPT = SV_MMA_WARPS / TEAMS;              // MMA warps per tile (team size)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = gridDim.x, b = blockIdx.x;
    const int RB = p.row_blocks;
    const int total = p.layers * RB;                       // row blocks of all layers
    // which blocks of the global list this CTA takes: a CONTIGUOUS chunk [u0, u0 + my_groups) (round 2: consecutive blocks of a
    // CTA then belong to the same layer, so the per-layer switch -- wait for the image, 8 LDS.128 of B fragments per lane -- is
    // paid once per ~RB blocks instead of once per block when a layer has fewer row blocks than the grid has CTAs), or the
    // strided list b, b + G, ... of round 1 (p.contig == 0)
    const int chunk = (total + G - 1) / G;
    const int u0 = p.contig ? b * chunk : b, ustep = p.contig ? 1 : G;
    const int my_groups = p.contig ? (total > u0 ? (total - u0 < chunk ? total - u0 : chunk) : 0)
                                   : (total > b ? (total - b + G - 1) / G : 0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), WPT); }
        for (int i = 0; i < p.nimg; ++i) { mbar_init(bar_ifull(i), 1); mbar_init(bar_iempty(i), SV_MMA_WARPS); }
        for (int q = 0; q < p.nslots; ++q) { mbar_init(bar_sfull(q), WPT); mbar_init(bar_sempty(q), 1); }
        mbar_fence_init();
    }
    __syncthreads();
    asm volatile("griddepcontrol.launch_dependents;");

    constexpr int SV_WARP_TILES = 8, SV_WARP_IMAGE = 9, SV_WARP_AUX0 = 10;
    constexpr bool REBAL = false;
    if constexpr (REBAL) {
        if (warp < SV_MMA_WARPS) asm volatile("setmaxnreg.inc.sync.aligned.u32 160;");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 96;");
        if (warp >= SV_WARP_AUX0 + AUX) return;              // two idle warps complete the fourth warpgroup
    }
    if (warp == SV_WARP_TILES) {
        // ===================== weight tiles: independent of any previous kernel =====================
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int gi = 0, u = u0; gi < my_groups; ++gi, u += ustep) {
                const int l = u / RB, rb = u - l * RB;
                for (int ks = 0; ks < p.nks; ++ks) {
                    mbar_wait_relaxed(bar_empty(s), ph ^ 1u);
                    mbar_arrive_expect_tx(bar_full(s), SV_TILE_BYTES);
                    if (p.one_box) {
                        tma_load_4d(sRing + s * SV_TILE_BYTES, &map_w, bar_full(s), 0, rb * SV_ROWS, ks * 8, l);
                    } else {
#pragma unroll
                        for (int sg = 0; sg < 8; ++sg)
                            tma_load_3d(sRing + s * SV_TILE_BYTES + sg * (SV_ROWS * SV_SEG), &map_w, bar_full(s),
                                        ks * 256 + sg * 32, rb * SV_ROWS, l);
                    }
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == SV_WARP_IMAGE) {
        // ===================== activation images (written by the preceding quantiser kernel) =====================
        if (lane == 0) {
            asm volatile("griddepcontrol.wait;" ::: "memory");
            int cur = -1, j = 0;
            for (int gi = 0, u = u0; gi < my_groups; ++gi, u += ustep) {
                const int l = u / RB;
                if (l == cur) continue;
                cur = l;
                const int ib = j % p.nimg;
                const uint32_t par = (uint32_t)(j / p.nimg) & 1u;
                mbar_wait_relaxed(bar_iempty(ib), par ^ 1u);
                mbar_arrive_expect_tx(bar_ifull(ib), p.img_bytes);
                const uint8_t* src = p.blob + (int64_t)l * p.blob_stride;
                const uint32_t dst = smem_u32(img0) + ib * img_pitch;
                // chunks of <= 32 KB keep each bulk request small enough to interleave with the weight stream
                for (uint32_t off = 0; off < p.img_bytes; off += 32768u) {
                    const uint32_t n = p.img_bytes - off < 32768u ? p.img_bytes - off : 32768u;
                    bulk_g2s(dst + off, src + off, n, bar_ifull(ib));
                }
                ++j;
            }
        }
    } else if (warp < SV_MMA_WARPS) {
        // ===================== dot products =====================
        const int g = lane >> 2, t = lane & 3;
        const int team = warp / WPT, wq = warp % WPT;       // TEAMS == 2: segments 2*wq, 2*wq+1
        const int wseg = TEAMS == 1 ? warp : 2 * wq;        // (first) K segment of this warp
        const uint32_t sSeg = wseg * (SV_ROWS * SV_SEG) + g * SV_SEG;
        const uint32_t phys0 = ((2 * t) ^ g) << 4, phys1 = ((2 * t + 1) ^ g) << 4;
        const int nks = p.nks, stages = p.stages, nslots = p.nslots, nimg = p.nimg, Mg = p.Mg;
        int ntok[NG];                                   // tokens of each group
#pragma unroll
        for (int q = 0; q < NG; ++q) { const int r = p.M - 8 * q; ntok[q] = r < 0 ? 0 : (r > 8 ? 8 : r); }
        int s = 0; uint32_t ph = 0;
        uint4 wlo[2], whi[2], nlo[2], nhi[2];
        auto fetch = [&](uint4 (&lo)[2], uint4 (&hi)[2], uint32_t ready) {
            if (!ready) mbar_wait(bar_full(s), ph);
            const uint8_t* seg = gbase + s * SV_TILE_BYTES + sSeg;
            lo[0] = *reinterpret_cast<const uint4*>(seg + phys0);
            hi[0] = *reinterpret_cast<const uint4*>(seg + 8 * SV_SEG + phys0);
            lo[1] = *reinterpret_cast<const uint4*>(seg + phys1);
            hi[1] = *reinterpret_cast<const uint4*>(seg + 8 * SV_SEG + phys1);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty(s));            // the tile lives in registers now
            if (++s == stages) { s = 0; ph ^= 1u; }
        };
        int acc[NG][4][4];
#pragma unroll
        for (int q = 0; q < NG; ++q)
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[q][i][e] = 0;
        uint4 bq[2][4];                                  // BREG: B fragments of the layer
        const uint4* img = nullptr;
        const int ntiles = my_groups * nks;
        if (TEAMS == 1 && ntiles > 0) fetch(wlo, whi, 0u);
        // position in the global row-block list, advanced without divisions: u = l * RB + rb, step G = Gq * RB + Gr
        const int Gq = ustep / RB, Gr = ustep - Gq * RB;
        int l = u0 / RB, rb = u0 - l * RB;
        int cur = -1, ks = 0, it = 0;
        int ib = 0, ib_cur = 0; uint32_t ipar = 0;       // image buffer ring
        int sq = 0; uint32_t spar = 0;                   // partial-sum slot ring
        // one half (pp) of a tile: 8 (BREG / NG == 1) or 8 * NG MMAs
        auto half = [&](const uint4& lo, const uint4& hi, int pp, int ksl) {
#ifdef TB200_STREAM_NOMMA
            acc[0][0][0] += (int)(lo.x ^ hi.y); return;
#endif
            if (BREG) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    imma_s(acc[0][i], lo.x & mk, hi.x & mk, lo.y & mk, hi.y & mk, bq[pp][i].x, bq[pp][i].y);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    imma_s(acc[0][i], lo.z & mk, hi.z & mk, lo.w & mk, hi.w & mk, bq[pp][i].z, bq[pp][i].w);
                }
            } else if (NG <= 2) {
                uint4 bv[NG][4];
#pragma unroll
                for (int q = 0; q < NG; ++q)
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        bv[q][i] = g < ntok[q] ? img[q * (p.grp_bytes >> 4) + img_vec(ksl * 8 + wseg, pp, i, g, t, Mg)]
                                               : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    const uint32_t a0 = lo.x & mk, a1 = hi.x & mk, a2 = lo.y & mk, a3 = hi.y & mk;
#pragma unroll
                    for (int q = 0; q < NG; ++q) imma_s(acc[q][i], a0, a1, a2, a3, bv[q][i].x, bv[q][i].y);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    const uint32_t a0 = lo.z & mk, a1 = hi.z & mk, a2 = lo.w & mk, a3 = hi.w & mk;
#pragma unroll
                    for (int q = 0; q < NG; ++q) imma_s(acc[q][i], a0, a1, a2, a3, bv[q][i].z, bv[q][i].w);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 bv[NG];
#pragma unroll
                    for (int q = 0; q < NG; ++q)
                        bv[q] = g < ntok[q] ? img[q * (p.grp_bytes >> 4) + img_vec(ksl * 8 + wseg, pp, i, g, t, Mg)]
                                            : make_uint4(0u, 0u, 0u, 0u);
                    const uint32_t mk = 0x03030303u << (2 * i);
                    {
                        const uint32_t a0 = lo.x & mk, a1 = hi.x & mk, a2 = lo.y & mk, a3 = hi.y & mk;
#pragma unroll
                        for (int q = 0; q < NG; ++q) imma_s(acc[q][i], a0, a1, a2, a3, bv[q].x, bv[q].y);
                    }
                    {
                        const uint32_t a0 = lo.z & mk, a1 = hi.z & mk, a2 = lo.w & mk, a3 = hi.w & mk;
#pragma unroll
                        for (int q = 0; q < NG; ++q) imma_s(acc[q][i], a0, a1, a2, a3, bv[q].z, bv[q].w);
                    }
                }
            }
        };
        // one tile: A fragments in (lo, hi); the next tile's fragments are fetched into (nl, nh) between the halves
        auto step = [&](uint4 (&lo)[2], uint4 (&hi)[2], uint4 (&nl)[2], uint4 (&nh)[2]) {
            if (ks == 0 && l != cur) {                   // first row block of a new layer: its image must have landed
                cur = l;
                ib_cur = ib;
                mbar_wait(bar_ifull(ib), ipar);
                img = reinterpret_cast<const uint4*>(img0 + ib * img_pitch);
                if (BREG) {
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            bq[pp][i] = g < ntok[0] ? img[img_vec(wseg, pp, i, g, t, Mg)] : make_uint4(0u, 0u, 0u, 0u);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_iempty(ib));      // the image buffer can take the next layer
                }
                if (++ib == nimg) { ib = 0; ipar ^= 1u; }
            }
            const uint32_t ready = mbar_test(bar_full(s), ph);      // next tile's stage, tested ahead of its use
            half(lo[0], hi[0], 0, ks);
            if (++it < ntiles) fetch(nl, nh, ready);
            half(lo[1], hi[1], 1, ks);
            if (++ks == nks) {
                // recombine the four fields (exact: set i is a multiple of 4^i) and park this warp's K-partials
                ks = 0;
                mbar_wait(bar_sempty(sq), spar ^ 1u);
#pragma unroll
                for (int q = 0; q < NG; ++q) {
                    uint4 v;
                    v.x = (uint32_t)(acc[q][0][0] + (acc[q][1][0] >> 2) + (acc[q][2][0] >> 4) + (acc[q][3][0] >> 6));
                    v.y = (uint32_t)(acc[q][0][1] + (acc[q][1][1] >> 2) + (acc[q][2][1] >> 4) + (acc[q][3][1] >> 6));
                    v.z = (uint32_t)(acc[q][0][2] + (acc[q][1][2] >> 2) + (acc[q][2][2] >> 4) + (acc[q][3][2] >> 6));
                    v.w = (uint32_t)(acc[q][0][3] + (acc[q][1][3] >> 2) + (acc[q][2][3] >> 4) + (acc[q][3][3] >> 6));
                    slots[(size_t)(sq * NG + q) * 256 + wseg * 32 + lane] = v;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[q][i][e] = 0;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_sfull(sq));
                if (++sq == nslots) { sq = 0; spar ^= 1u; }
                l += Gq; rb += Gr;
                if (rb >= RB) { rb -= RB; ++l; }
                if (!BREG && (it >= ntiles || l != cur)) {              // last row block of this layer for this CTA
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_iempty(ib_cur));
                }
            }
        };
        if constexpr (TEAMS == 1) {
#pragma unroll 1
            while (it < ntiles) {
                step(wlo, whi, nlo, nhi);
                if (it >= ntiles) break;
                step(nlo, nhi, wlo, whi);
            }
        } else {
            // Two teams of four warps: every team walks the whole row-block list (layer changes, ring and slot positions)
            // but computes only its own row blocks (gi & 1 == team).
            (void)step;
            uint32_t ready = 0;                          // early test of the stage of this team's next tile
            uint4 b2[2][2][4];                           // BREG: B fragments of both segments
            auto seg_half = [&](const uint4& lo, const uint4& hi, int j, int pp, int ksl) {
#ifdef TB200_STREAM_NOMMA
                acc[0][0][0] += (int)(lo.x ^ hi.y); return;
#endif
                uint4 bv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (BREG) bv[i] = b2[j][pp][i];
                    else bv[i] = g < ntok[0] ? img[img_vec(ksl * 8 + wseg + j, pp, i, g, t, Mg)] : make_uint4(0u, 0u, 0u, 0u);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    imma_s(acc[0][i], lo.x & mk, hi.x & mk, lo.y & mk, hi.y & mk, bv[i].x, bv[i].y);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t mk = 0x03030303u << (2 * i);
                    imma_s(acc[0][i], lo.z & mk, hi.z & mk, lo.w & mk, hi.w & mk, bv[i].z, bv[i].w);
                }
            };
#pragma unroll 1
            for (int gi = 0; gi < my_groups; ++gi) {
                if (l != cur) {
                    cur = l;
                    ib_cur = ib;
                    mbar_wait(bar_ifull(ib), ipar);
                    img = reinterpret_cast<const uint4*>(img0 + ib * img_pitch);
                    if (BREG) {
#pragma unroll
                        for (int j = 0; j < 2; ++j)
#pragma unroll
                            for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    b2[j][pp][i] = g < ntok[0] ? img[img_vec(wseg + j, pp, i, g, t, Mg)] : make_uint4(0u, 0u, 0u, 0u);
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_iempty(ib));
                    }
                    if (++ib == nimg) { ib = 0; ipar ^= 1u; }
                }
                if ((gi & 1) == team) {
                    const uint32_t sready = mbar_test(bar_sempty(sq), spar ^ 1u);     // consumed after the MMAs
#pragma unroll 1
                    for (int k2 = 0; k2 < nks; ++k2) {
                        if (!ready) mbar_wait(bar_full(s), ph);
                        const uint8_t* seg = gbase + s * SV_TILE_BYTES + sSeg;
                        uint4 alo[2][2], ahi[2][2];
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const uint8_t* sj = seg + j * (SV_ROWS * SV_SEG);
                            alo[j][0] = *reinterpret_cast<const uint4*>(sj + phys0);
                            ahi[j][0] = *reinterpret_cast<const uint4*>(sj + 8 * SV_SEG + phys0);
                            alo[j][1] = *reinterpret_cast<const uint4*>(sj + phys1);
                            ahi[j][1] = *reinterpret_cast<const uint4*>(sj + 8 * SV_SEG + phys1);
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_empty(s));
                        if (++s == stages) { s = 0; ph ^= 1u; }
                        {   // test the stage of this team's NEXT tile now, use the answer after the MMAs
                            int s2 = s; uint32_t ph2 = ph;
                            if (k2 + 1 == nks) { s2 += nks; while (s2 >= stages) { s2 -= stages; ph2 ^= 1u; } }
                            ready = mbar_test(bar_full(s2), ph2);
                        }
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            seg_half(alo[j][0], ahi[j][0], j, 0, k2);
                            seg_half(alo[j][1], ahi[j][1], j, 1, k2);
                        }
                    }
                    if (!sready) mbar_wait(bar_sempty(sq), spar ^ 1u);
                    uint4 v;
                    v.x = (uint32_t)(acc[0][0][0] + (acc[0][1][0] >> 2) + (acc[0][2][0] >> 4) + (acc[0][3][0] >> 6));
                    v.y = (uint32_t)(acc[0][0][1] + (acc[0][1][1] >> 2) + (acc[0][2][1] >> 4) + (acc[0][3][1] >> 6));
                    v.z = (uint32_t)(acc[0][0][2] + (acc[0][1][2] >> 2) + (acc[0][2][2] >> 4) + (acc[0][3][2] >> 6));
                    v.w = (uint32_t)(acc[0][0][3] + (acc[0][1][3] >> 2) + (acc[0][2][3] >> 4) + (acc[0][3][3] >> 6));
                    slots[(size_t)(sq * NG) * 256 + wq * 32 + lane] = v;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[0][i][e] = 0;
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_sfull(sq));
                } else {
                    s += nks;                                            // the other team's tiles
                    while (s >= stages) { s -= stages; ph ^= 1u; }
                }
                if (++sq == nslots) { sq = 0; spar ^= 1u; }
                l += Gq; rb += Gr;
                if (rb >= RB) { rb -= RB; ++l; }
                if (!BREG && (gi + 1 == my_groups || l != cur)) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_iempty(ib_cur));
                }
            }
        }
    } else {
        // ===================== reduction of the K-partials + fused epilogue =====================
        // Aux warp a takes row blocks a, a + AUX, ...; nslots is a multiple of AUX, so slot buffer q is only ever waited
        // on by warp q % AUX, in phase order (a parity wait must never be more than one phase ahead of its barrier).
        const int a = warp - SV_WARP_AUX0;
        const int g = lane >> 2, t = lane & 3;
        asm volatile("griddepcontrol.wait;" ::: "memory");
        // position of this warp's current row block and of the one it prefetches for, advanced without divisions
        const int step = AUX * ustep;
        const int Sq = step / RB, Sr = step - Sq * RB;
        struct Pos { int l, rb; };
        auto advance = [&](Pos& q) { q.l += Sq; q.rb += Sr; if (q.rb >= RB) { q.rb -= RB; ++q.l; } };
        Pos cp, np;
        { const int u = u0 + a * ustep; cp.l = u / RB; cp.rb = u - cp.l * RB; }
        np = cp;
        // scale/bias of a row block are read once per launch (cold in L2): the loads for this warp's NEXT row block are
        // issued before the current one is reduced
        auto load_sb = [&](const Pos& q, bool valid, float (&v)[4]) {
            v[0] = v[1] = 1.0f; v[2] = v[3] = 0.0f;
            if (valid) {
                const int64_t row0 = (int64_t)q.rb * SV_ROWS + g, row1 = row0 + 8;
                if (p.scale) {
                    const float* sc = p.scale + (int64_t)q.l * p.stride_sb;
                    if (row0 < p.N) v[0] = sc[row0];
                    if (row1 < p.N) v[1] = sc[row1];
                }
                if (p.bias) {
                    const float* bi = p.bias + (int64_t)q.l * p.stride_sb;
                    if (row0 < p.N) v[2] = bi[row0];
                    if (row1 < p.N) v[3] = bi[row1];
                }
            }
        };
        int cur = -1;
        float sxv[NG][2]; int sumq[NG][2];
#pragma unroll
        for (int q = 0; q < NG; ++q) { sxv[q][0] = sxv[q][1] = 0.0f; sumq[q][0] = sumq[q][1] = 0; }
        // Under a saturated read stream every load waits behind ~20 MB of queued tile traffic (microseconds), so the
        // scale/bias values are requested SV_FIFO row blocks of this warp ahead and kept in a register FIFO.
        constexpr int SV_FIFO = 4;
        float sbf[SV_FIFO][4];
#pragma unroll
        for (int d = 0; d < SV_FIFO; ++d) {
            load_sb(np, a + d * AUX < my_groups, sbf[d]);
            advance(np);
        }
        int sq_i = a % p.nslots; uint32_t spar = (uint32_t)(a / p.nslots) & 1u;     // AUX <= nslots: a / nslots == 0
        const bool store_acc = p.acc_out != nullptr;
        if constexpr (PEER) {
            if (lane == 0) for (int r = 0; r < p.n_peers; ++r) tma_prefetch_desc(&pm.m[r]);
        }
        [[maybe_unused]] int pcnt = 0;                   // staging tiles written by this warp (peer gather)
        [[maybe_unused]] int64_t prev_off = 0; [[maybe_unused]] int prev_rb = 0;
        // per-lane constants of the direct / multicast store of a staged tile: lane -> (token, 16-byte piece of its 16 rows)
        const uint32_t esz_c = p.out_dtype == TB200_F32 ? 4u : 2u;
        const int ppr_c = (int)(SV_ROWS * esz_c / 16), tok_c = lane / ppr_c, cpi_c = lane - tok_c * ppr_c;
        const uint32_t stage0 = smem_u32(gbase) + p.peer_stage_off + (uint32_t)(a * SV_PEER_BUFS * (NG * 8 * SV_ROWS)) * esz_c;
        auto flush_tile = [&](int buf, int64_t obase_prev, int rb_prev, bool valid) {
            if constexpr (PEER) {
                if (!valid) return;
                for (int pc = lane, tok = tok_c; pc < p.M * ppr_c; pc += 32, tok += 32 / ppr_c) {
                    const int64_t row = (int64_t)rb_prev * SV_ROWS + cpi_c * (16 / (int)esz_c);
                    if (row < p.N) {
                        uint4 v;
                        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                     : "r"(stage0 + (uint32_t)(buf * (NG * 8 * SV_ROWS)) * esz_c + (uint32_t)(pc * 16)));
                        const int64_t off = (obase_prev + (int64_t)tok * p.ldo + row) * esz_c;
                        if (p.mc_out) {
                            multimem_st_128(reinterpret_cast<uint8_t*>(p.mc_out) + off, v);
                        } else {
#pragma unroll
                            for (int r = 0; r < 8; ++r)
                                if (r < p.n_peers) *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.peer_ptrs[r]) + off) = v;
                        }
                    }
                }
            }
        };
#pragma unroll 1
        for (int sn0 = a; sn0 < my_groups; sn0 += SV_FIFO * AUX) {
#pragma unroll
          for (int d = 0; d < SV_FIFO; ++d) {
            const int sn = sn0 + d * AUX;
            if (sn >= my_groups) break;
            const int l = cp.l;
            const int64_t row0 = (int64_t)cp.rb * SV_ROWS + g, row1 = row0 + 8;
            const float sc0 = sbf[d][0], sc1 = sbf[d][1], bi0 = sbf[d][2], bi1 = sbf[d][3];
            load_sb(np, sn + SV_FIFO * AUX < my_groups, sbf[d]);
            advance(np);
            if (l != cur) {
                cur = l;
                const uint8_t* tail = p.blob + (int64_t)l * p.blob_stride + p.img_bytes;
                const float* sx = reinterpret_cast<const float*>(tail);
                const int* sq = reinterpret_cast<const int*>(tail + 128);
#pragma unroll
                for (int q = 0; q < NG; ++q)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int tok = 8 * q + 2 * t + e;
                        sxv[q][e] = tok < p.M ? sx[tok] : 0.0f;
                        sumq[q][e] = tok < p.M ? sq[tok] : 0;
                    }
            }
            mbar_wait_relaxed(bar_sfull(sq_i), spar);
            int vv[NG][4];
#pragma unroll
            for (int q = 0; q < NG; ++q) {
                const uint4* sl = slots + (size_t)(sq_i * NG + q) * 256 + lane;
                vv[q][0] = vv[q][1] = vv[q][2] = vv[q][3] = 0;
#pragma unroll
                for (int w = 0; w < WPT; ++w) {
                    const uint4 v = sl[w * 32];
                    vv[q][0] += (int)v.x; vv[q][1] += (int)v.y; vv[q][2] += (int)v.z; vv[q][3] += (int)v.w;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_sempty(sq_i));      // the partial sums live in registers: the slot is free again
            sq_i += AUX;
            if (sq_i >= p.nslots) { sq_i -= p.nslots; spar ^= 1u; }
            // fragment: vv0 (row g, tok 2t), vv1 (row g, tok 2t+1), vv2 (row g+8, tok 2t), vv3 (row g+8, tok 2t+1)
            const int64_t obase = (int64_t)l * p.stride_out;
            float yv[PEER ? NG : 1][4];
#pragma unroll
            for (int q = 0; q < NG; ++q) {
#pragma unroll
                for (int e4 = 0; e4 < 4; ++e4) {
                    if constexpr (PEER) yv[q][e4] = 0.0f;
                    const int e = e4 & 1, hi = e4 >> 1;
                    const int tok = 8 * q + 2 * t + e;
                    const int64_t row = hi ? row1 : row0;
                    if (row < p.N && tok < p.M) {
                        const int av = vv[q][e4] - sumq[q][e];
                        if (store_acc && l == 0) p.acc_out[(int64_t)tok * p.N + row] = av;
                        const float tv = __fmul_rn((float)av, sxv[q][e]);
                        const float y = epilogue(tv, hi ? sc1 : sc0, hi ? bi1 : bi0, p.clip_val);
                        const int64_t o = obase + (int64_t)tok * p.ldo + row;
                        if constexpr (!PEER) {
                            if (p.out_dtype == TB200_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(y);
                            else if (p.out_dtype == TB200_F16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(y);
                            else reinterpret_cast<float*>(p.out)[o] = y;
                        } else {
                            yv[q][e4] = y;
                        }
                    }
                }
            }
            if (PEER && kAblate && (p.dbg & 2)) {                  // experiment: the PEER instantiation without its stores
                if (yv[0][0] == 123.456f) p.acc_out[0] = 1;
            } else if constexpr (PEER) {
                // tile [8*NG tokens][16 rows] -> staging (two buffers per aux warp) -> one TMA store per rank
                const uint32_t esz = p.out_dtype == TB200_F32 ? 4u : 2u;
                const uint32_t stg = smem_u32(gbase) + p.peer_stage_off + (uint32_t)((a * SV_PEER_BUFS + (pcnt & (SV_PEER_BUFS - 1))) * (NG * 8 * SV_ROWS)) * esz;
                // the stores that last read `stg` are done (a tile store queues behind the weight tiles in flight, so
                // several staging buffers per warp are needed to keep the warp from waiting for it)
                if (!(p.direct || p.mc_out)) {
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(SV_PEER_BUFS - 1) : "memory");
                    __syncwarp();
                }
                if (kAblate && (p.dbg & 4)) { if (yv[0][0] + yv[0][1] + yv[0][2] + yv[0][3] == 123.456f) p.acc_out[0] = 1; } else
#pragma unroll
                for (int q = 0; q < NG; ++q)
#pragma unroll
                    for (int e4 = 0; e4 < 4; ++e4) {
                        const uint32_t el = (uint32_t)((8 * q + 2 * t + (e4 & 1)) * SV_ROWS + g + 8 * (e4 >> 1));
                        if (p.out_dtype == TB200_F32) {
                            asm volatile("st.shared.f32 [%0], %1;" ::"r"(stg + el * 4u), "f"(yv[q][e4]));
                        } else {
                            unsigned short h;
                            if (p.out_dtype == TB200_BF16) { const __nv_bfloat16 v = __float2bfloat16_rn(yv[q][e4]); h = *reinterpret_cast<const unsigned short*>(&v); }
                            else { const __half v = __float2half_rn(yv[q][e4]); h = *reinterpret_cast<const unsigned short*>(&v); }
                            asm volatile("st.shared.u16 [%0], %1;" ::"r"(stg + el * 2u), "h"(h));
                        }
                    }
                if (p.direct || p.mc_out) {
                    // 16-byte stores straight from the staging tile (per rank, or ONE multimem store that the switch replicates),
                    // software-pipelined: the tile staged for the PREVIOUS row block leaves now, so the st.shared -> ld.shared
                    // round trip of the current one is never waited for.  (TMA tile stores queue behind the ~128 KB of weight
                    // tiles in flight per SM; the aux warps have no slack: tools/peer_cost.py.)
                    if (!(kAblate && (p.dbg & 8))) flush_tile((pcnt - 1) & (SV_PEER_BUFS - 1), prev_off, prev_rb, pcnt > 0);
                    prev_off = obase; prev_rb = cp.rb;
                } else {
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
#pragma unroll 1
                    for (int r = 0; r < p.n_peers; ++r) tma_store_3d(&pm.m[r], stg, cp.rb * SV_ROWS, 0, l);
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                }
                ++pcnt;
            }
            advance(cp);
          }
        }
        if constexpr (PEER) {
            if (p.direct || p.mc_out) { __syncwarp(); flush_tile((pcnt - 1) & (SV_PEER_BUFS - 1), prev_off, prev_rb, pcnt > 0); }
        }
        if (PEER && lane == 0) {
            // all tile stores of this warp have completed; grid completion orders them before the caller's barrier kernel
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        }
    }
}

// ---- activation quantiser writing the per-layer blob: image (token groups of 8, see img_vec) | s_x | sum_qx ----
// One CTA per (token, layer).  Same arithmetic as tb200_quant_act_i8 (quant.cu).
template <typename T>
__global__ void __launch_bounds__(256)
quant_image_kernel(const T* __restrict__ x, int64_t ldx, int64_t stride_x, int64_t K, int M, int Mg, int nks,
                   uint8_t* __restrict__ blob, int64_t blob_stride, uint32_t img_bytes, uint32_t grp_bytes, bool vec_ok) {
    __shared__ float sh_max[8];
    __shared__ int sh_sum[8];
    asm volatile("griddepcontrol.launch_dependents;");            // the GEMV may start streaming weights now
    asm volatile("griddepcontrol.wait;" ::: "memory");            // the previous kernel may still be reading the blob
    const int m = blockIdx.x, layer = blockIdx.y;
    const T* xr = x + (int64_t)layer * stride_x + (int64_t)m * ldx;
    uint8_t* bl = blob + (int64_t)layer * blob_stride;
    uint32_t* image = reinterpret_cast<uint32_t*>(bl + (size_t)(m >> 3) * grp_bytes);
    const int n = m & 7;
    const int groups = nks * 256;
    constexpr int QC = 2;                                  // groups per thread kept in registers between the passes (K <= 8192)
    const bool cached = nks <= QC;
    float vc[QC][16];
    auto load16 = [&](int gi, float (&v)[16]) {
        if (vec_ok && 16 * (int64_t)gi + 16 <= K) {
            const uint4* q = reinterpret_cast<const uint4*>(xr + 16 * (int64_t)gi);
#pragma unroll
            for (int jv = 0; jv < (int)sizeof(T); ++jv) {
                const uint4 u4 = q[jv];
                const T* e = reinterpret_cast<const T*>(&u4);
#pragma unroll
                for (int tt = 0; tt < 16 / (int)sizeof(T); ++tt) v[jv * (16 / (int)sizeof(T)) + tt] = to_f32(e[tt]);
            }
        } else {
#pragma unroll
            for (int tt = 0; tt < 16; ++tt) {
                const int64_t k = 16 * (int64_t)gi + tt;
                v[tt] = k < K ? to_f32(xr[k]) : 0.0f;
            }
        }
    };
    float amax = 0.0f;
    if (cached) {
#pragma unroll
        for (int c = 0; c < QC; ++c)
            if (c < nks) load16(threadIdx.x + 256 * c, vc[c]);
#pragma unroll
        for (int c = 0; c < QC; ++c)
            if (c < nks) {
#pragma unroll
                for (int tt = 0; tt < 16; ++tt) amax = fmaxf(amax, fabsf(vc[c][tt]));
            }
    } else {
        for (int gi = threadIdx.x; 16 * (int64_t)gi < K; gi += 256) {
            float v[16];
            load16(gi, v);
#pragma unroll
            for (int tt = 0; tt < 16; ++tt) amax = fmaxf(amax, fabsf(v[tt]));
        }
    }
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0) sh_max[threadIdx.x >> 5] = amax;
    __syncthreads();
    amax = sh_max[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) amax = fmaxf(amax, sh_max[i]);
    const float s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    const float inv = amax > 0.0f ? __fdiv_rn(127.0f, amax) : 1.0f;
    int lsum = 0;
    auto emit = [&](int gi, const float (&v)[16]) {
        int qv[16];
#pragma unroll
        for (int tt = 0; tt < 16; ++tt) {
            float r = __fmul_rn(v[tt], inv);
            r = fminf(fmaxf(r, -127.0f), 127.0f);
            qv[tt] = __float_as_int(__fadd_rn(r, 12582912.0f)) - 0x4B400000;
            lsum += qv[tt];
        }
        const int c = gi >> 2, jw = gi & 3, lc = c & 7;
#pragma unroll
        for (int i = 0; i < 4; ++i) {     // word i = elements i, 4+i, 8+i, 12+i (field i of bytes 0..3 of packed word gi)
            const uint32_t w = (uint32_t)(qv[i] & 0xff) | ((uint32_t)(qv[4 + i] & 0xff) << 8) |
                               ((uint32_t)(qv[8 + i] & 0xff) << 16) | ((uint32_t)(qv[12 + i] & 0xff) << 24);
            image[img_vec(c >> 3, lc & 1, i, n, lc >> 1, Mg) * 4 + jw] = w;
        }
    };
    if (cached) {
#pragma unroll
        for (int c = 0; c < QC; ++c)
            if (c < nks) emit(threadIdx.x + 256 * c, vc[c]);
    } else {
        for (int gi = threadIdx.x; gi < groups; gi += 256) {
            float v[16];
            load16(gi, v);
            emit(gi, v);
        }
    }
    lsum = warp_sum(lsum);
    if ((threadIdx.x & 31) == 0) sh_sum[threadIdx.x >> 5] = lsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tsum = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) tsum += sh_sum[i];
        reinterpret_cast<float*>(bl + img_bytes)[m] = s;
        reinterpret_cast<int*>(bl + img_bytes + 128)[m] = tsum;
    }
}

// ---- host side ------------------------------------------------------------------------------
struct StreamGeom { int Mg, ng, nks; uint32_t grp_bytes, img_bytes; int64_t blob_stride; };
static StreamGeom stream_geom(int64_t M, int64_t K) {
    StreamGeom gm;
    gm.Mg = (int)(M < 8 ? M : 8);
    gm.ng = (int)((M + 7) / 8);
    gm.nks = gemv_nks(K);
    gm.grp_bytes = (uint32_t)(gm.nks * 4096 * gm.Mg);
    gm.img_bytes = gm.grp_bytes * (uint32_t)gm.ng;
    gm.blob_stride = ((int64_t)gm.img_bytes + SV_TAIL + 255) / 256 * 256;
    return gm;
}
// byte offset (from the 1024-aligned base) of the peer staging tiles = end of the barrier block, 128-byte aligned
static size_t stream_stage_off(const StreamGeom& gm, int stages, int nslots, int nimg) {
    const size_t img_pitch = ((size_t)gm.img_bytes + 127) & ~(size_t)127;
    const size_t end = (size_t)stages * SV_TILE_BYTES + nimg * img_pitch + (size_t)nslots * gm.ng * 4096 +
                       (2 * SV_MAX_STAGES + 2 * SV_MAX_IMG + 2 * SV_MAX_SLOTS) * 8;
    return (end + 127) & ~(size_t)127;
}
// staging tiles of the tensor-parallel epilogue: aux warps x SV_PEER_BUFS buffers x [8 * ng tokens][16 rows] x element size
static size_t stream_stage_bytes(const StreamGeom& gm, bool peers, int aux, size_t esz) { return peers ? (size_t)aux * SV_PEER_BUFS * gm.ng * 8 * SV_ROWS * esz : 0; }
static size_t stream_smem(const StreamGeom& gm, int stages, int nslots, int nimg, size_t peer_esz = 0) {
    static const int teams = (int)tb_knob("TB200_STREAM_TEAMS", 2);
    static const int aux4 = (int)tb_knob("TB200_STREAM_AUX4", -1);
    const int aux = (gm.ng == 1 && (teams != 2 || aux4 > 0 || (aux4 < 0 && peer_esz > 0))) ? 4 : 2;         // as dispatched in gemv_stream_impl (staging: an upper bound is fine)
    return 1024 + stream_stage_off(gm, stages, nslots, nimg) + stream_stage_bytes(gm, peer_esz > 0, aux, peer_esz) + 64;
}
// ring depth first (bytes in flight), then a second image buffer, then more slots
static bool stream_pick(const StreamGeom& gm, bool breg, int* stages, int* nslots, int* nimg, size_t peer_esz = 0) {
    const size_t limit = 227 * 1024;
    static const int st_env = (int)tb_knob("TB200_STREAM_STAGES", 0);
    const int want_st = st_env > 0 ? st_env : 8;
    const int min_st = gm.nks < 3 ? 3 : (gm.nks < SV_MAX_STAGES ? gm.nks : SV_MAX_STAGES);
    // images: with the B fragments in registers (breg) a buffer is free again as soon as the layer's fragments are loaded,
    // but the NEXT layer's image must already be in flight then (its load queues behind the weight ring), so small
    // images get up to four buffers
    static const int img_env = (int)tb_knob("TB200_STREAM_NIMG", 0);
    const int want_img = img_env > 0 ? (img_env > SV_MAX_IMG ? SV_MAX_IMG : img_env) : (breg ? (gm.img_bytes <= 16384 ? 4 : 2) : 2);
    for (int img = want_img; img >= 1; --img)
        for (int st = want_st; st >= min_st; --st)
            for (int ns = (gm.ng > 2 ? 2 : gm.ng == 2 ? 4 : 8); ns >= (gm.ng == 1 ? 4 : 2); ns -= (gm.ng == 1 ? 4 : 2))   // multiples of the aux-warp count
                if (stream_smem(gm, st, ns, img, peer_esz) <= limit) { *stages = st; *nslots = ns; *nimg = img; return true; }
    return false;
}

bool gemv_stream_fits(int64_t M, int64_t K) {
    if (M < 1 || M > TB200_GEMV_MAX_M) return false;
    const StreamGeom gm = stream_geom(M, K);
    int a, b2, c;
    return stream_pick(gm, gm.ng == 1 && gm.nks == 1, &a, &b2, &c);
}

template <int NG, bool BREG, int TEAMS, int AUX, bool PEER>
static int launch_stream_v(const CUtensorMap& map, const typename PeerArg<PEER>::type& pm, const StreamParams& p, size_t smem, int ctas, bool pdl, cudaStream_t s) {
    auto kern = gemv_stream_kernel<NG, BREG, TEAMS, AUX, PEER>;
    static SmemOptIn optin;
    TB_CUDA(ensure_dynamic_smem(kern, optin, smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas, 1, 1);
    cfg.blockDim = dim3(sv_threads2(TEAMS, AUX), 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, map, pm, p));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

template <int NG, bool BREG, int TEAMS, int AUX>
static int launch_stream(const CUtensorMap& map, const PeerMaps* pm, const StreamParams& p, size_t smem, int ctas, bool pdl, cudaStream_t s) {
    if (pm) return launch_stream_v<NG, BREG, TEAMS, AUX, true>(map, *pm, p, smem, ctas, pdl, s);
    return launch_stream_v<NG, BREG, TEAMS, AUX, false>(map, NoPeerMaps{0}, p, smem, ctas, pdl, s);
}

int gemv_stream_impl(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                     const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                     const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                     void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                     int32_t* acc_out, int layers, void* workspace, cudaStream_t s,
                     void* const* peers = nullptr, int n_peers = 0) {
    const StreamGeom gm = stream_geom(M, K);
    const bool breg = gm.ng == 1 && gm.nks == 1;
    StreamParams p = {};
    const bool mc = peers && n_peers < 0;          // peers[0] is an NVSwitch multicast address
    if (mc) n_peers = 1;
    const bool tp = peers && n_peers > 0;
    const size_t peer_esz = tp ? dtype_size(out_dtype) : 0;
    if (!stream_pick(gm, breg, &p.stages, &p.nslots, &p.nimg, peer_esz)) {
        set_error("gemv_stream: M=%lld K=%lld does not fit the shared-memory activation image", (long long)M, (long long)K);
        return TB200_ENOTSUP;
    }
    CUtensorMap map;
    int one_box = 0;
    int rc = get_weight_map(packed, ld_packed, N, stride_packed, layers, &map, &one_box);
    if (rc) return rc;
    // 1. activations -> int8 image per layer (one CTA per token and layer)
    uint8_t* blob = (uint8_t*)workspace;
    const size_t xs = dtype_size(x_dtype);
    const bool vec_ok = ((uintptr_t)x % 16 == 0) && ((ldx * xs) % 16 == 0) && ((stride_x * xs) % 16 == 0);
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)M, (unsigned)layers, 1);
        cfg.blockDim = dim3(256, 1, 1);
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        TB_DISPATCH_DTYPE(x_dtype, T,
            TB_CUDA(cudaLaunchKernelEx(&cfg, quant_image_kernel<T>, (const T*)x, ldx, stride_x, K, (int)M, gm.Mg, gm.nks,
                                       blob, gm.blob_stride, gm.img_bytes, gm.grp_bytes, vec_ok)));
        TB_LAUNCH_CHECK();
    }
    // 2. the persistent GEMV
    p.blob = blob; p.blob_stride = gm.blob_stride;
    p.scale = scale; p.bias = bias; p.stride_sb = stride_sb;
    p.out = out; p.ldo = ldo; p.stride_out = stride_out; p.out_dtype = out_dtype;
    p.acc_out = acc_out;
    PeerMaps peer_maps;
    const PeerMaps* pm = nullptr;
    p.n_peers = 0;
    if (tp) {
        // one output tensor map per rank: [layers][M][N] elements at (stride_out, ldo) pitches, box = 8 * ng tokens x 16 rows
        const size_t os = dtype_size(out_dtype);
        if ((ldo * os) % 16 || (stride_out * os) % 16) {
            set_error("gemv_stream_tp: ldo and stride_out must be multiples of 16 bytes");
            return TB200_EINVAL;
        }
        const CUtensorMapDataType dt = out_dtype == TB200_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                     : out_dtype == TB200_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
        uint64_t dims_o[3] = {(uint64_t)N, (uint64_t)M, (uint64_t)layers};
        uint64_t str_o[2] = {(uint64_t)(ldo * os), (uint64_t)((layers > 1 ? stride_out : ldo * M) * os)};
        uint32_t box_o[3] = {SV_ROWS, (uint32_t)(8 * gm.ng), 1};
        for (int r = 0; r < n_peers; ++r) {
            if ((uintptr_t)peers[r] % 16) { set_error("gemv_stream_tp: peer pointer %d is not 16-byte aligned", r); return TB200_EINVAL; }
            int rc2 = make_tensor_map(&peer_maps.m[r], dt, 3, peers[r], dims_o, str_o, box_o, CU_TENSOR_MAP_SWIZZLE_NONE);
            if (rc2) return rc2;
        }
        for (int r = n_peers; r < 8; ++r) peer_maps.m[r] = peer_maps.m[0];
        pm = &peer_maps;
        p.n_peers = n_peers;
        for (int r = 0; r < n_peers; ++r) p.peer_ptrs[r] = peers[r];
        static const int direct_env = (int)tb_knob("TB200_STREAM_PEER_DIRECT", 1);
        p.direct = (direct_env && N % (16 / (int64_t)os) == 0) ? 1 : 0;     // whole 16-byte pieces only (else TMA clips the ragged edge)
        if (mc) {
            if (N % (16 / (int64_t)os)) { set_error("gemv_stream_mc: N_local must be a multiple of %d", (int)(16 / os)); return TB200_EINVAL; }
            p.mc_out = peers[0];
        }
        p.peer_stage_off = (uint32_t)stream_stage_off(gm, p.stages, p.nslots, p.nimg);
    }
    p.N = N; p.M = (int)M; p.Mg = gm.Mg; p.nks = gm.nks;
    p.row_blocks = (int)((N + SV_ROWS - 1) / SV_ROWS);
    p.layers = layers;
    p.img_bytes = gm.img_bytes; p.grp_bytes = gm.grp_bytes;
    p.clip_val = fabsf(clip_val);
    p.one_box = one_box;
    static const int dbg_env = (int)tb_knob("TB200_STREAM_DBG", 0);
    p.dbg = dbg_env;
    static const int contig_env = (int)tb_knob("TB200_STREAM_CONTIG", 1);
    p.contig = contig_env;
    const int64_t total = (int64_t)p.row_blocks * layers;
    const int ctas = (int)(total < sm_count() ? total : sm_count());
    const size_t smem = stream_smem(gm, p.stages, p.nslots, p.nimg, peer_esz);
    switch (gm.ng) {
        case 1: {
            static const int teams = (int)tb_knob("TB200_STREAM_TEAMS", 2);
            static const int aux4 = (int)tb_knob("TB200_STREAM_AUX4", -1);     // -1: four aux warps for the tensor-parallel epilogue only
            if (teams == 2 && breg && (aux4 > 0 || (aux4 < 0 && pm))) return launch_stream<1, true, 2, 4>(map, pm, p, smem, ctas, true, s);
            if (teams == 2)
                return breg ? launch_stream<1, true, 2, 2>(map, pm, p, smem, ctas, true, s) : launch_stream<1, false, 2, 2>(map, pm, p, smem, ctas, true, s);
            return breg ? launch_stream<1, true, 1, 4>(map, pm, p, smem, ctas, true, s) : launch_stream<1, false, 1, 4>(map, pm, p, smem, ctas, true, s);
        }
        case 2: return launch_stream<2, false, 1, 2>(map, pm, p, smem, ctas, true, s);
        case 3: return launch_stream<3, false, 1, 2>(map, pm, p, smem, ctas, true, s);
        default: return launch_stream<4, false, 1, 2>(map, pm, p, smem, ctas, true, s);
    }
}

}  // namespace tb200

using namespace tb200;

// tokens from which the list is served by the grouped tcgen05 kernel (gemm_ts.cu) instead of mma.sync: above 8 tokens
// the legacy IMMA pipe, not HBM, bounds the mma.sync kernel
static int stream_umma_min_m() {
    static const int v = (int)tb_knob("TB200_STREAM_UMMA_MIN_M", 9);
    return v;
}
static inline int64_t ru(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

extern "C" size_t tb200_gemv_stream_workspace(int64_t M, int64_t K, int count) {
    if (M < 1 || K < 1 || count < 1) return 256;
    const StreamGeom gm = stream_geom(M, K);
    const size_t image = (size_t)gm.blob_stride * (size_t)count + 256;
    const int64_t rows = M * count;
    const size_t umma = (size_t)(ru(rows * ru(K, 256), 256) + 2 * ru(rows * 4, 256) + 256);     // field-split operand order: 256-element blocks
    return image > umma ? image : umma;
}

// > 8 tokens: per-token int8 quantisation of all layers in one launch (layout 1), then the grouped tcgen05 GEMM
static int stream_umma(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                       const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                       const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                       void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                       int32_t* acc_out, int layers, void* workspace, cudaStream_t s,
                       void* const* peers = nullptr, int n_peers = 0) {
    // <= 32 tokens: the field-split form of the tcgen05 kernel (in-place masks of the odd 2-bit fields, gemm_ts.cu TsCfg NF == 2)
    const int qlayout = M <= 32 ? TB200_QLAYOUT_GEMM_SPLIT : TB200_QLAYOUT_GEMM;
    const int64_t ldq = qlayout == TB200_QLAYOUT_GEMM_SPLIT ? ru(K, 256) : ru(K, 16), rows = M * layers;
    char* w = (char*)workspace;
    int8_t* q_x = (int8_t*)w;      w += ru(rows * ldq, 256);
    float* s_x = (float*)w;        w += ru(rows * 4, 256);
    int32_t* sum_qx = (int32_t*)w;
    const size_t xs = dtype_size(x_dtype);
    if (layers == 1 || stride_x == M * ldx) {
        int rc = tb200_quant_act_i8(x, x_dtype, rows, K, ldx, q_x, ldq, s_x, sum_qx, qlayout, s);
        if (rc) return rc;
    } else {
        for (int l = 0; l < layers; ++l) {
            int rc = tb200_quant_act_i8((const char*)x + (size_t)l * stride_x * xs, x_dtype, M, K, ldx, q_x + (int64_t)l * M * ldq, ldq,
                                        s_x + (int64_t)l * M, sum_qx + (int64_t)l * M, qlayout, s);
            if (rc) return rc;
        }
    }
    UmmaParams p = {};
    p.qlayout = qlayout;
    p.s_x = s_x; p.sum_qx = sum_qx; p.scale = scale; p.bias = bias;
    p.out = out; p.ldo = ldo; p.acc_out = acc_out;
    p.M = M; p.N = N; p.K = K; p.clip_val = fabsf(clip_val);
    p.layers = layers; p.stride_act_rows = M; p.stride_sb = stride_sb; p.stride_out = stride_out;
    p.mc = (peers && n_peers < 0) ? 1 : 0;
    if (p.mc) n_peers = 1;
    p.n_peers = peers ? n_peers : 0;
    for (int r = 0; r < p.n_peers; ++r) p.peers[r] = peers[r];
    if (p.mc && N % (16 / (int64_t)dtype_size(out_dtype))) { set_error("gemv_stream_mc: N_local must be a multiple of 8 (16-bit) / 4 (fp32)"); return TB200_EINVAL; }
    p.dbg = (int)tb_knob("TB200_GEMM_DBG", 0);
    // the GEMM's prologue and launch latency overlap the quantiser (single quantiser launch directly before it)
    static const int pdl_env = (int)tb_knob("TB200_STREAM_UMMA_PDL", 1);
    p.pdl = (pdl_env && (layers == 1 || stride_x == M * ldx)) ? 1 : 0;
    return gemm_ts_launch(q_x, ldq, packed, ld_packed, p, out_dtype, s, M, stride_packed);
}

static int gemv_stream_entry(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                             const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                             const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                             void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                             int count, int32_t* acc_out, void* workspace, void* stream,
                             void* const* peers, int n_peers) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0 || count == 0) return TB200_OK;
    TB_REQUIRE(x && packed && out && workspace, "gemv_stream: null pointer");
    TB_REQUIRE(M >= 1 && M <= TB200_GEMV_MAX_M, "gemv_stream: M=%lld out of range [1,%d]", (long long)M, TB200_GEMV_MAX_M);
    TB_REQUIRE(K > 0 && N > 0 && ldx >= K && ldo >= N, "gemv_stream: bad shape M=%lld K=%lld N=%lld", (long long)M, (long long)K, (long long)N);
    TB_REQUIRE(dtype_ok(x_dtype) && dtype_ok(out_dtype), "gemv_stream: bad dtype");
    const int64_t kc = (K + 63) / 64;
    TB_REQUIRE(ld_packed % 16 == 0 && ld_packed >= kc * 16, "gemv_stream: ld_packed=%lld must be a multiple of 16 and >= %lld",
               (long long)ld_packed, (long long)(kc * 16));
    TB_REQUIRE((uintptr_t)packed % 16 == 0 && stride_packed % 16 == 0, "gemv_stream: packed weights must be 16-byte aligned");
    TB_REQUIRE((uintptr_t)workspace % 256 == 0, "gemv_stream: workspace must be 256-byte aligned");
    TB_REQUIRE(count >= 1 && count <= 65535, "gemv_stream: bad layer count %d", count);
    if (M >= stream_umma_min_m())
        return stream_umma(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N, scale, bias, stride_sb,
                           clip_val, out, out_dtype, ldo, stride_out, acc_out, count, workspace, (cudaStream_t)stream, peers, n_peers);
    return gemv_stream_impl(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N, scale, bias, stride_sb,
                            clip_val, out, out_dtype, ldo, stride_out, acc_out, count, workspace, (cudaStream_t)stream, peers, n_peers);
}

extern "C" int tb200_gemv_stream(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                                 const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                                 const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                 void* out, int out_dtype, int64_t ldo, int64_t stride_out,
                                 int count, int32_t* acc_out, void* workspace, void* stream) {
    return gemv_stream_entry(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N, scale, bias, stride_sb,
                             clip_val, out, out_dtype, ldo, stride_out, count, acc_out, workspace, stream, nullptr, 0);
}

// Tensor-parallel form: this rank's row block (N_local rows) of every layer in the list; the epilogues store each
// output element into EVERY rank's [count, M, ldo] buffer through peer-mapped (NVLink) pointers, so the all-gather of
// the whole list costs no extra pass: ONE cross-rank barrier by the caller after the call makes all layers consumable.
extern "C" int tb200_gemv_stream_tp(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                                    const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N_local,
                                    const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                    void* const* out_peers, int n_peers, int out_dtype, int64_t ldo, int64_t stride_out,
                                    int count, void* workspace, void* stream) {
    if (!(out_peers && n_peers >= 1 && n_peers <= 8)) {
        set_error("gemv_stream_tp: bad peer list (n_peers=%d, 1..8 supported)", n_peers);
        return TB200_EINVAL;
    }
    for (int r = 0; r < n_peers; ++r)
        if (!out_peers[r]) { set_error("gemv_stream_tp: null peer pointer %d", r); return TB200_EINVAL; }
    return gemv_stream_entry(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N_local, scale, bias, stride_sb,
                             clip_val, out_peers[0], out_dtype, ldo, stride_out, count, nullptr, workspace, stream, out_peers, n_peers);
}

// The same through an NVSwitch MULTICAST mapping (torch symmetric memory: handle.multicast_ptr): `out_multicast` is the multicast
// address of element (layer 0, token 0, first column of THIS rank's block); every 16 bytes of output leave the GPU ONCE and the
// switch writes them into every rank's buffer (egress 1x instead of n_ranks x).  N_local must be a multiple of 8 (fp32: 4).
extern "C" int tb200_gemv_stream_mc(const void* x, int x_dtype, int64_t M, int64_t K, int64_t ldx, int64_t stride_x,
                                    const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N_local,
                                    const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                    void* out_multicast, int out_dtype, int64_t ldo, int64_t stride_out,
                                    int count, void* workspace, void* stream) {
    if (!out_multicast) { set_error("gemv_stream_mc: null multicast pointer"); return TB200_EINVAL; }
    void* peers[1] = {out_multicast};
    return gemv_stream_entry(x, x_dtype, M, K, ldx, stride_x, packed, ld_packed, stride_packed, N_local, scale, bias, stride_sb,
                             clip_val, out_multicast, out_dtype, ldo, stride_out, count, nullptr, workspace, stream, peers, -1);
}

// End-to-end with HOST activations / outputs for a list of layers (weights resident on the device): one H2D copy of
// all activations, the streaming GEMV, one D2H copy of all outputs, in order on `stream`.  x_host [count, M, K] and
// out_host [count, M, N] are dense and should be pinned; dev_x / dev_out are device staging buffers of the same sizes.
extern "C" int tb200_gemv_stream_host(const void* x_host, int x_dtype, int64_t M, int64_t K,
                                      const uint8_t* packed, int64_t ld_packed, int64_t stride_packed, int64_t N,
                                      const float* scale, const float* bias, int64_t stride_sb, float clip_val,
                                      void* out_host, int out_dtype, int count,
                                      void* dev_x, void* dev_out, void* workspace, void* stream) {
    TB_DEVICE_OR_RETURN();
    if (M == 0 || N == 0 || count == 0) return TB200_OK;
    TB_REQUIRE(x_host && out_host && dev_x && dev_out, "gemv_stream_host: null pointer");
    TB_REQUIRE(dtype_ok(x_dtype) && dtype_ok(out_dtype), "gemv_stream_host: bad dtype");
    cudaStream_t s = (cudaStream_t)stream;
    TB_CUDA(cudaMemcpyAsync(dev_x, x_host, (size_t)count * M * K * dtype_size(x_dtype), cudaMemcpyHostToDevice, s));
    int rc = tb200_gemv_stream(dev_x, x_dtype, M, K, K, M * K, packed, ld_packed, stride_packed, N, scale, bias, stride_sb,
                               clip_val, dev_out, out_dtype, N, M * N, count, nullptr, workspace, stream);
    if (rc) return rc;
    TB_CUDA(cudaMemcpyAsync(out_host, dev_out, (size_t)count * M * N * dtype_size(out_dtype), cudaMemcpyDeviceToHost, s));
    return TB200_OK;
}
