// gemm_ts.cu — prefill-shaped ternary GEMM, weights as the TMEM A operand.
//
//   out[m, n] = clip( (sum_k q_x[m,k] * w[n,k]) * s_x[m] * scale[n] + bias[n] )
//
// Orientation ("swap-AB"): 128 weight rows = UMMA M / TMEM lanes, BT tokens = UMMA N.  Per 256-wide
// k-block (one pipeline stage):
//   * TMA brings the packed weight tile (128 rows x 64 bytes, SWIZZLE_64B) and the int8 activation tile
//     (two BT x 128-byte SWIZZLE_128B sub-tiles);
//   * expander warp q (TMEM lane quarter q; two warp sets alternate k-blocks) reads the 64 packed bytes of
//     its rows (4 conflict-free LDS.128 per thread), turns the 16 words into 64 registers of u8 codes
//     (3 shifts + 4 ANDs per word) and writes them with two tcgen05.st.32x32b.x32 into one of two A
//     slots in tensor memory (64 columns each): the expanded weights never touch shared memory;
//   * the MMA warp issues 8 x tcgen05.mma.cta_group::1.kind::i8 (A from TMEM, B from shared memory) after
//     ONE mbarrier wait (ready[s] = activation bytes landed + A slot written) and signals ONE
//     tcgen05.commit (done[s] frees the shared-memory stage and the A slot) -- all nine under ONE election
//     in one asm block (umma_i8_ts_kblock_warp): the issuing loop's latency, not the tensor pipe, bounded the
//     first versions (profiles/README.md, ablation table) and still sets how soon an A slot is free again;
//   * the TMA producer, like the MMA issuer, is a whole warp running warp-uniform code with one elected lane
//     issuing: a single producer THREAD moved every TMA operand through an R2UR broadcast loop, which was
//     co-critical for 32-token tiles (three TMAs per ~600-clock k-block);
//   * 32-token tiles of decode lists use the field-split form (TsCfg NF == 2): odd 2-bit fields masked in
//     place (value 4c) into their own accumulators, activations in TB200_QLAYOUT_GEMM_SPLIT order;
//   * accumulators are double-buffered in TMEM (2 x BT columns); the epilogue warps read them with
//     tcgen05.ld, apply -sum_qx, s_x*scale, bias, clip, and hand 32-token x 128-channel chunks to TMA
//     stores through a double-buffered shared-memory staging tile (2-byte scattered stores cost 60 us).
// TMEM budget at BT=192: 2*192 (D) + 2*64 (A) = 512 columns; at BT<=64: 2*BT (D) + 6*64 (A).
#include <stdlib.h>
#include <type_traits>
#include "umma.cuh"

namespace tb200 {

constexpr int TS_ROWS = 128;
constexpr int TS_BK = 256;           // int8 k per stage = two 128-byte swizzle atoms per token row
constexpr int TS_KS = TS_BK / 128;   // activation sub-tiles per stage
constexpr int TS_EXP_SETS = 2;       // expander warp sets (4 warps each) taking k-blocks round-robin
constexpr int TS_THREADS = 256 + 128 * TS_EXP_SETS;   // warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-7 epilogue, 8.. expanders
constexpr int TS_P_BYTES = TS_ROWS * (TS_BK / 4);   // 8 KB packed weights per stage (64 bytes per row)
constexpr int TS_A_COLS = TS_BK / 4; // 64 columns per A tile
// experiment switches per warp role (TB200_GEMM_DBG bits in the experiment library; -DTS_DBG_x=1 keeps one role's test in a
// release build: the presence of these branches changes the compiler's schedule of the role's loop)
#ifndef TS_DBG_PROD
#define TS_DBG_PROD kAblate
#endif
#ifndef TS_DBG_MMA
#define TS_DBG_MMA kAblate
#endif
#ifndef TS_DBG_EPI
#define TS_DBG_EPI kAblate
#endif
#ifndef TS_DBG_EXP
#define TS_DBG_EXP kAblate
#endif
constexpr bool kDbgProd = TS_DBG_PROD, kDbgMma = TS_DBG_MMA, kDbgEpi = TS_DBG_EPI, kDbgExp = TS_DBG_EXP;

template <int BT, int OSZ = 4, int NF = 1> struct TsCfg {
    static constexpr int B_BYTES = BT * TS_BK;
    static constexpr int STAGE_BYTES = B_BYTES + TS_P_BYTES;
    static constexpr int STAGES = BT >= 192 ? 3 : BT >= 128 ? 4 : BT >= 64 ? 6 : BT >= 32 ? 12 : 8;   // 160 -> 4 stages of 48 KB
    // skinny token tiles: TWO warps issue the MMAs of alternate k-blocks into their own accumulator (the issuing warp's
    // instruction chain, not the tensor pipe, bounds 32-token tiles); the epilogue adds the two partial sums
    static constexpr int DUAL = BT <= 32 ? 2 : 1;
    // NF == 2 ("field-split", 32-token tiles of decode lists, activations in TB200_QLAYOUT_GEMM_SPLIT order): the expanders mask
    // the odd 2-bit fields IN PLACE (w & 0x0c0c0c0c: value 4c) instead of shifting them down -- 5 instead of 7 alu-pipe
    // instructions per packed word, and that pipe bounds skinny token tiles -- so even and odd fields go to separate
    // accumulators (lo, hi) and the epilogue forms lo + (hi >> 2), exactly.
    static constexpr int D_COLS = 2 * BT * DUAL * NF;
    // A tiles in TMEM: the expand -> MMA -> retire -> reuse loop of a slot is ~2000 clocks long, so skinny token tiles
    // (which finish a k-block's MMAs quickly and have TMEM columns to spare) keep six of them in flight
    // (160-token tiles leave room for a third slot: 24.6 -> 21.5 us on the ViT projection M=8224 K=1024 N=1024 in the experiment
    // build, no change in the release build; a third slot for 128-token tiles changes nothing, 128-token tiles for large M lose 15 %)
    static constexpr int NSA = NF == 2 ? 4 : BT <= 32 ? 6 : BT <= 64 ? 4 : BT == 160 ? 3 : 2;
    static constexpr int NEED = D_COLS + NSA * TS_A_COLS;
    static constexpr int TMEM_COLS = NEED <= 32 ? 32 : NEED <= 64 ? 64 : NEED <= 128 ? 128 : NEED <= 256 ? 256 : 512;
    static_assert(NEED <= 512, "TMEM budget");
    static constexpr int SX_BYTES = 2 * BT * 8;
    static constexpr int OUT_STAGE_BYTES = 32 * TS_ROWS * OSZ;        // one 32-token x 128-channel output chunk
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + SX_BYTES + 2 * OUT_STAGE_BYTES + 1024 + 1024;
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// CL = thread-block cluster size: the CL CTAs of a cluster work on CL adjacent weight tiles and the SAME
// token tile; each loads 1/CL of the activation tile and TMA-multicasts it to all of them, which divides
// the L2 -> SM operand traffic that bounds the single-CTA kernel (profiles/README.md, ablation table).
template <typename OT, int BT, int CL, int NF = 1>
__global__ void __launch_bounds__(TS_THREADS, 1)
umma_ts_kernel(const __grid_constant__ CUtensorMap map_act, const __grid_constant__ CUtensorMap map_pw,
               const __grid_constant__ CUtensorMap map_out, const __grid_constant__ PeerMaps pm, const UmmaParams p) {
    using Cfg = TsCfg<BT, (int)sizeof(OT), NF>;
    static_assert(NF == 1 || (CL == 1 && BT == 32), "field-split accumulators: 32-token single-CTA form only");
    constexpr int NS = Cfg::STAGES;
    constexpr int TS_NSA = Cfg::NSA;
    static_assert(NS > TS_NSA, "a stage's done barrier must not wrap before its A slot is reused");
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    uint8_t* gbase = smem_dyn + (base - smem_u32(smem_dyn));
    const uint32_t sB = base;                                     // [NS][BT*128] activations (SWIZZLE_128B)
    const uint32_t sP = sB + NS * Cfg::B_BYTES;                   // [NS][4 KB] packed weights (SWIZZLE_32B)
    const uint32_t sSX = sP + NS * TS_P_BYTES;
    const uint32_t sOut = sSX + Cfg::SX_BYTES;                    // [2][32 tokens][128 channels] output staging for TMA stores
    const uint32_t sBar = sOut + 2 * Cfg::OUT_STAGE_BYTES;
    // One wait and one commit per k-block for the single MMA-issuing thread (its loop latency, not the
    // tensor pipe, bounded the first versions): ready[s] completes when the activation bytes have landed
    // AND the stage's A tile is in TMEM; done[s] frees both the shared-memory stage and the A slot.
    auto bar_pfull = [&](int s) { return sBar + 8u * s; };                       // packed weights landed
    auto bar_ready = [&](int s) { return sBar + 8u * (NS + s); };                // activations landed + A expanded
    auto bar_done = [&](int s) { return sBar + 8u * (2 * NS + s); };             // MMAs of the stage retired
    auto bar_tfull = [&](int b) { return sBar + 8u * (3 * NS + b); };
    auto bar_tempty = [&](int b) { return sBar + 8u * (3 * NS + 2 + b); };
    const uint32_t sTmemPtr = sBar + 8u * (3 * NS + 4);
    float* sx_sm = reinterpret_cast<float*>(gbase + (sSX - base));
    int* sq_sm = reinterpret_cast<int*>(sx_sm + 2 * BT);
    volatile uint32_t* tmem_ptr_sm = reinterpret_cast<volatile uint32_t*>(gbase + (sTmemPtr - base));
    const uint8_t* gP = gbase + (sP - base);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t cta_rank = 0;
    if (CL > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    const int cluster_id = blockIdx.x / CL, num_clusters = gridDim.x / CL;
    const int n_groups = (p.n_tiles + CL - 1) / CL;
    const int tiles_per_layer = p.m_tiles * n_groups;       // super-tiles: CL weight tiles x one token tile
    const int total_tiles = tiles_per_layer * p.layers;
    constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);
    const int t_begin = cluster_id, t_end = total_tiles, t_step = num_clusters;     // tiles of this CTA (cluster): round-robin

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_act);
        tma_prefetch_desc(&map_pw);
        for (int s = 0; s < NS; ++s) {
            mbar_init(bar_pfull(s), 1);
            mbar_init(bar_ready(s), 1 + 4);                  // producer (with the activation bytes) + 4 expander warps
            mbar_init(bar_done(s), CL);                      // every CTA of the cluster must have retired the stage
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(bar_tfull(b), (CL == 1) ? Cfg::DUAL : 1);
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
    if (CL > 1) {   // peers' barriers must be initialised before the first multicast load / remote arrive
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_sm;
    const uint32_t tmem_a0 = tmem_base + (uint32_t)Cfg::D_COLS;     // first A tile column

    if (warp == 0) {
        // ===================== TMA producer =====================
        // CL == 1: the WHOLE warp runs the loop and one elected lane issues (as the MMA warps do): the loop is warp-uniform, so
        // barrier addresses and tile coordinates stay in uniform registers.  In the single-thread form every TMA operand went
        // through an R2UR broadcast loop and the thread's ~3 TMAs per k-block were co-critical for skinny token tiles
        // (a k-block is ~600 clocks there; tools/stream_probe.py: 207 -> 190 us on 64 x 4096 -> 11008 at 32 tokens).
        if (CL == 1) {
            int s = 0; uint32_t ph = 0;
            if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");     // the quantised activations are complete
            for (int t = t_begin; t < t_end; t += t_step) {
                const int layer = t / tiles_per_layer, tl = t - layer * tiles_per_layer;
                const int n_tile = tl / p.m_tiles, m_tile = tl - n_tile * p.m_tiles;
                if (!(kDbgProd && (p.dbg & 8))) {
                    for (int kb = 0; kb < p.k_blocks; ++kb) {
                        mbar_wait_uniform(bar_done(s), ph ^ 1u);
                        mbar_arrive_expect_tx_warp(bar_pfull(s), TS_P_BYTES);
                        tma_load_3d_warp(sP + s * TS_P_BYTES, &map_pw, bar_pfull(s), kb * (TS_BK / 4), n_tile * TS_ROWS, layer);
                        mbar_arrive_expect_tx_warp(bar_ready(s), Cfg::B_BYTES);
#pragma unroll
                        for (int j = 0; j < TS_KS; ++j)
                            tma_load_3d_warp(sB + s * Cfg::B_BYTES + j * (BT * 128), &map_act, bar_ready(s),
                                             kb * TS_BK + j * 128, m_tile * BT, layer);
                        if (++s == NS) { s = 0; ph ^= 1u; }
                    }
                } else {                                  // experiment: no activation traffic
                    for (int kb = 0; kb < p.k_blocks; ++kb) {
                        mbar_wait_uniform(bar_done(s), ph ^ 1u);
                        mbar_arrive_expect_tx_warp(bar_pfull(s), TS_P_BYTES);
                        tma_load_3d_warp(sP + s * TS_P_BYTES, &map_pw, bar_pfull(s), kb * (TS_BK / 4), n_tile * TS_ROWS, layer);
                        mbar_arrive_warp(bar_ready(s));
                        if (++s == NS) { s = 0; ph ^= 1u; }
                    }
                }
            }
        } else if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");     // the quantised activations are complete
            for (int t = t_begin; t < t_end; t += t_step) {
                const int layer = t / tiles_per_layer, tl = t - layer * tiles_per_layer;
                const int n_group = tl / p.m_tiles, m_tile = tl - n_group * p.m_tiles;
                const int n_tile = n_group * CL + (int)cta_rank;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(bar_done(s), ph ^ 1u);
                    mbar_arrive_expect_tx(bar_pfull(s), TS_P_BYTES);
                    tma_load_3d(sP + s * TS_P_BYTES, &map_pw, bar_pfull(s), kb * (TS_BK / 4), n_tile * TS_ROWS, layer);
                    if (kDbgProd && (p.dbg & 8)) { mbar_arrive(bar_ready(s)); if (++s == NS) { s = 0; ph ^= 1u; } continue; }   // experiment: no activation traffic
                    mbar_arrive_expect_tx(bar_ready(s), Cfg::B_BYTES);
#pragma unroll
                    for (int j = 0; j < TS_KS; ++j) {
                        const uint32_t dst = sB + s * Cfg::B_BYTES + j * (BT * 128);
                        if (CL == 1) {
                            tma_load_3d(dst, &map_act, bar_ready(s), kb * TS_BK + j * 128, m_tile * BT, layer);
                        } else {   // my 1/CL slice of the token tile, delivered to every CTA of the cluster
                            constexpr int SL = BT / CL;
                            tma_load_3d_mc(dst + cta_rank * (SL * 128), &map_act, bar_ready(s),
                                           kb * TS_BK + j * 128, m_tile * BT + (int)cta_rank * SL, layer, kMask);
                        }
                    }
                    if (++s == NS) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1 || (warp == 3 && CL == 1 && Cfg::DUAL == 2)) {
        // ===================== MMA issuer(s) =====================
        // CL == 1: the whole warp runs the loop (uniform code, see umma_i8_ts_warp); clusters keep the single-thread form
        if (CL == 1) {
            constexpr uint32_t idesc = umma_idesc_i8(TS_ROWS, BT);
            int s = 0; uint32_t ph = 0; int a = 0; int it = 0;
            for (int t = t_begin; t < t_end; t += t_step, ++it) {
                const int buf = it & 1;
                mbar_wait_uniform(bar_tempty(buf), (((uint32_t)it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                constexpr int ND = Cfg::DUAL;
                const int me = ND == 2 ? (warp == 3 ? 1 : 0) : 0;             // which alternate k-blocks / accumulator
                const uint32_t d_tmem = tmem_base + (uint32_t)((buf * ND + me) * NF * BT);     // NF == 2: lo | hi
                int mine_done = 0;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    if (ND == 1 || (kb & 1) == me) {
                        mbar_wait_uniform(bar_ready(s), ph);
                        tc_fence_after();
                        const uint32_t a_tmem = tmem_a0 + (uint32_t)(a * TS_A_COLS);
                        static_assert(TS_KS == 2, "umma_i8_ts_kblock_warp issues two sub-tiles");
                        if (!(kDbgMma && (p.dbg & 4)))
                            umma_i8_ts_kblock_warp<NF == 2>(d_tmem, d_tmem + (uint32_t)(NF == 2 ? BT : 0), a_tmem,
                                                            umma_desc_sw128(sB + s * Cfg::B_BYTES), umma_desc_sw128(sB + s * Cfg::B_BYTES + BT * 128),
                                                            idesc, (uint32_t)mine_done, bar_done(s));
                        else umma_commit_warp(bar_done(s));
                        mine_done = 1;
                        if (kb + ND >= p.k_blocks) umma_commit_warp(bar_tfull(buf));   // my last k-block of the tile
                    }
                    if (++s == NS) { s = 0; ph ^= 1u; }
                    if (++a == TS_NSA) a = 0;
                }
                if (ND == 2 && me >= p.k_blocks) {      // a single k-block: the second issuer has nothing to add
                    if (lane == 0) mbar_arrive(bar_tfull(buf));
                }
            }
        } else if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(TS_ROWS, BT);
            int s = 0; uint32_t ph = 0; int a = 0; int it = 0;
            for (int t = t_begin; t < t_end; t += t_step, ++it) {
                const int buf = it & 1;
                mbar_wait(bar_tempty(buf), (((uint32_t)it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BT);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(bar_ready(s), ph);
                    tc_fence_after();
                    const uint32_t a_tmem = tmem_a0 + (uint32_t)(a * TS_A_COLS);
                    if (!(kDbgMma && (p.dbg & 4))) {
#pragma unroll
                        for (int j = 0; j < TS_KS; ++j) {
                            const uint64_t bdesc = umma_desc_sw128(sB + s * Cfg::B_BYTES + j * (BT * 128));
#pragma unroll
                            for (int k = 0; k < 4; ++k)   // K=32 int8 per MMA = 8 TMEM columns of A, 32 bytes of B
                                umma_i8_ts(d_tmem, a_tmem + (uint32_t)(j * 32 + k * 8), bdesc + 2u * k, idesc,
                                           (kb | j | k) ? 1u : 0u);
                        }
                    }
                    umma_commit_mc(bar_done(s), kMask);   // frees the stage in every CTA of the cluster
                    if (kb == p.k_blocks - 1) umma_commit(bar_tfull(buf));
                    if (++s == NS) { s = 0; ph ^= 1u; }
                    if (++a == TS_NSA) a = 0;
                }
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ===================== epilogue =====================
        const int q = warp & 3;
        const int et = threadIdx.x - 128;
        OT* out_base = reinterpret_cast<OT*>(p.out);
        int it = 0, chunk_no = 0;
        if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");         // s_x / sum_qx come from the quantiser
        if (et == 0 && p.tma_out) {
            tma_prefetch_desc(&map_out);
            for (int r = 0; r < p.n_peers; ++r) tma_prefetch_desc(&pm.m[r]);
        }
        for (int t = t_begin; t < t_end; t += t_step, ++it) {
            const int buf = it & 1;
            const int layer = t / tiles_per_layer, tl = t - layer * tiles_per_layer;
            const int n_group = tl / p.m_tiles, m_tile = tl - n_group * p.m_tiles;
            const int n_tile = n_group * CL + (int)cta_rank;
            const int64_t m0 = (int64_t)m_tile * BT;
            const int64_t n = (int64_t)n_tile * TS_ROWS + q * 32 + lane;
            const float* s_x_l = p.s_x + layer * p.stride_act_rows;
            const int32_t* sum_qx_l = p.sum_qx + layer * p.stride_act_rows;
            OT* out = out_base + layer * p.stride_out;
            const bool acc_on = p.acc_out != nullptr && layer == 0;
            for (int j = et; j < BT; j += 128) {
                const int64_t m = m0 + j;
                sx_sm[buf * BT + j] = m < p.M ? s_x_l[m] : 0.0f;
                sq_sm[buf * BT + j] = m < p.M ? sum_qx_l[m] : 0;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const float sc = (n < p.N && p.scale) ? p.scale[layer * p.stride_sb + n] : 1.0f;
            const float bi = (n < p.N && p.bias) ? p.bias[layer * p.stride_sb + n] : 0.0f;
            // skinny token tiles are bound by the expanders' instruction stream on the same schedulers: poll with a back-off
            if (BT <= 64 && !(kDbgEpi && (p.dbg & 32))) mbar_wait_relaxed(bar_tfull(buf), ((uint32_t)it >> 1) & 1u);
            else mbar_wait(bar_tfull(buf), ((uint32_t)it >> 1) & 1u);
            tc_fence_after();
            constexpr int CH = BT >= 32 ? 32 : 16;
            const uint32_t sx_s = sSX + (uint32_t)(buf * BT) * 4u;                 // shared-space addresses: explicit
            const uint32_t sq_s = sSX + (uint32_t)(2 * BT + buf * BT) * 4u;       // ld.shared keeps them off the generic path
#pragma unroll 1
            for (int c0 = 0; c0 < ((kDbgEpi && (p.dbg & 1)) ? 0 : BT); c0 += CH) {
                // per-token scale and code offset of the chunk first (registers), so that the 32 output
                // chains below are independent of each other and of the global stores
                float sxv[CH]; int sqv[CH];
                uint32_t r[32];
                constexpr int ND = (CL == 1) ? Cfg::DUAL : 1;
                constexpr bool SPLIT_LD = (CH == 32 && ND == 1);      // TMEM load issued first, waited for after the barrier below
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * ND * NF * BT + c0);
                if (SPLIT_LD) tmem_ld32_issue(taddr, r);
                if constexpr (NF == 2) {          // lo + (hi >> 2), both issuers' partial sums; before sxv / sqv take their registers
                    uint32_t rh[32], r2[32];
                    tmem_ld32(taddr, r);
                    tmem_ld32(taddr + (uint32_t)BT, rh);
                    if (ND == 2 && p.k_blocks > 1) {
                        tmem_ld32(taddr + (uint32_t)(2 * BT), r2);
#pragma unroll
                        for (int j = 0; j < 32; ++j) r[j] += r2[j];
                        tmem_ld32(taddr + (uint32_t)(3 * BT), r2);
#pragma unroll
                        for (int j = 0; j < 32; ++j) rh[j] += r2[j];
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += (uint32_t)((int)rh[j] >> 2);      // hi is a multiple of 4
                }
#pragma unroll
                for (int j = 0; j < CH; j += 4) {
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(sxv[j]), "=f"(sxv[j + 1]), "=f"(sxv[j + 2]), "=f"(sxv[j + 3]) : "r"(sx_s + (c0 + j) * 4));
                    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(sqv[j]), "=r"(sqv[j + 1]), "=r"(sqv[j + 2]), "=r"(sqv[j + 3]) : "r"(sq_s + (c0 + j) * 4));
                }
                if (!SPLIT_LD && NF == 1) { if (CH == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r); }
                if (NF == 1 && ND == 2 && p.k_blocks > 1) {
                    uint32_t r2[32];
                    if (CH == 32) tmem_ld32(taddr + (uint32_t)BT, r2); else tmem_ld16(taddr + (uint32_t)BT, r2);
#pragma unroll
                    for (int j = 0; j < CH; ++j) r[j] += r2[j];
                }
                if (p.tma_out && CH == 32) {
                    // registers -> shared staging [token][channel] -> one TMA store of 32 tokens x 128 channels
                    // (256-byte rows, asynchronous); the staging buffers alternate per chunk
                    const uint32_t stg = sOut + (uint32_t)(chunk_no & 1) * Cfg::OUT_STAGE_BYTES;
                    if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store that last read `stg` is done
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (SPLIT_LD) tmem_ld32_wait(r);
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int a = (int)r[j] - sqv[j];
                        const float tv = __fmul_rn((float)a, sxv[j]);
                        const OT yv = from_f32<OT>(epilogue(tv, sc, bi, p.clip_val));
                        const uint32_t addr = stg + (uint32_t)((j * TS_ROWS + q * 32 + lane) * sizeof(OT));
                        if (sizeof(OT) == 4) asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(to_f32(yv)) : "memory");
                        else asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<const unsigned short*>(&yv)) : "memory");
                    }
                    if (acc_on) {                   // parity output (debug): a separate, warp-uniformly skipped block keeps the
#pragma unroll                                     // 32 predicated global stores out of the hot loop above
                        for (int j = 0; j < CH; ++j)
                            if (n < p.N && m0 + c0 + j < p.M) p.acc_out[(m0 + c0 + j) * p.N + n] = (int)r[j] - sqv[j];
                    }
                    if (p.mc) {
                        // multicast gather: the staged [32 tokens][128 channels] tile leaves as 16-byte multimem stores, four per
                        // thread (the switch writes them into every rank's buffer)
                        asm volatile("bar.sync 1, 128;" ::: "memory");
                        constexpr int PPR = TS_ROWS * (int)sizeof(OT) / 16;          // 16-byte pieces per token row of the tile
                        constexpr int CPP = 16 / (int)sizeof(OT);                    // channels per piece
                        OT* mcb = reinterpret_cast<OT*>(p.peers[0]) + layer * p.stride_out;
#pragma unroll
                        for (int i = 0; i < 32 * PPR / 128; ++i) {
                            const int pc = et + 128 * i, j = pc / PPR, cp = pc - j * PPR;
                            const int64_t tok = m0 + c0 + j, ch = (int64_t)n_tile * TS_ROWS + cp * CPP;
                            if (tok < p.M && ch < p.N) {
                                uint4 v;
                                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                             : "r"(stg + (uint32_t)(pc * 16)) : "memory");
                                multimem_st_128(mcb + tok * p.ldo + ch, v);
                            }
                        }
                        ++chunk_no;
                        continue;
                    }
                    fence_proxy_async();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (et == 0) {
                        if (p.n_peers == 0) {
                            tma_store_3d(&map_out, stg, n_tile * TS_ROWS, (int)(m0 + c0), layer);
                        } else {
#pragma unroll 1
                            for (int r = 0; r < p.n_peers; ++r)     // tensor-parallel gather: the tile goes to every rank over NVLink
                                tma_store_3d(&pm.m[r], stg, n_tile * TS_ROWS, (int)(m0 + c0), layer);
                        }
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                    ++chunk_no;
                } else if (SPLIT_LD ? (tmem_ld32_wait(r), n < p.N) : (n < p.N)) {     // the wait is warp-wide: before the row test
                    OT* orow = out + m0 * p.ldo + n + (int64_t)c0 * p.ldo;
                    int valid = (int)((p.M - m0 - c0) < CH ? (p.M - m0 - c0) : CH);            // tokens of the chunk in range
                    if (kDbgEpi && (p.dbg & 256)) valid = (sc == 12345.678f) ? valid : 0;
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int a = (int)r[j] - sqv[j];
                        const float tv = __fmul_rn((float)a, sxv[j]);
                        const OT yv = from_f32<OT>(epilogue(tv, sc, bi, p.clip_val));
                        if (j < valid) {
                            if (p.n_peers == 0) orow[(int64_t)j * p.ldo] = yv;
                            else {
#pragma unroll 1
                                for (int r = 0; r < p.n_peers; ++r)     // unaligned peer buffers: 64-byte warp segments per rank
                                    reinterpret_cast<OT*>(p.peers[r])[(orow - out_base) + (int64_t)j * p.ldo] = yv;
                            }
                        }
                    }
                    if (acc_on) {
#pragma unroll
                        for (int j = 0; j < CH; ++j)
                            if (j < valid) p.acc_out[(m0 + c0 + j) * p.N + n] = (int)r[j] - sqv[j];
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty(buf));
            asm volatile("bar.sync 1, 128;" ::: "memory");
        }
        if (et == 0 && p.tma_out) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // staging must outlive the reads
    } else if (warp >= 8) {
        // ===================== expanders: packed 2-bit -> u8 codes -> TMEM A tile =====================
        const int q = warp & 3;                       // TMEM lane quarter; thread = weight row q*32 + lane
        const int row = q * 32 + lane;
        const int set = (warp - 8) >> 2;              // this warp set expands k-blocks set, set+SETS, ...
        // This set's k-blocks are kbc = set, set + SETS, ... of the CTA's whole k-block sequence (tiles back to back): the ring
        // positions are stepped by SETS per OWN k-block (walking every k-block and skipping the other set's cost ~20 alu-pipe
        // instructions per k-block and warp, 16 % of the pipe that bounds skinny token tiles).
        const long long my_tiles = t_begin < t_end ? (t_end - t_begin + t_step - 1) / t_step : 0;
        const long long total_kb = my_tiles * p.k_blocks;
        static_assert(NS >= TS_EXP_SETS && TS_NSA >= TS_EXP_SETS, "ring steps assume at most one wrap");
        int s = set; uint32_t ph = 0;                 // ring position of the current k-block (kbc % NS)
        int a = set % TS_NSA;                         // A slot (kbc % TS_NSA)
        // ring position of the k-block TS_NSA earlier (A-slot reuse), valid from the first own kbc >= TS_NSA
        int s2 = ((set - TS_NSA) % TS_EXP_SETS + TS_EXP_SETS) % TS_EXP_SETS; uint32_t ph2 = 0;
        for (long long kbc = set; kbc < total_kb; kbc += TS_EXP_SETS) {
            mbar_wait(bar_pfull(s), ph);
            // SWIZZLE_64B: 16-byte chunk index ^= address bits [7:8] = (row >> 1) & 3 at a 64-byte row pitch
            const uint8_t* P = gP + s * TS_P_BYTES + row * 64;
            const int sw = (row >> 1) & 3;
            uint32_t w[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint4 v = *reinterpret_cast<const uint4*>(P + ((c ^ sw) << 4));
                w[4 * c] = v.x; w[4 * c + 1] = v.y; w[4 * c + 2] = v.z; w[4 * c + 3] = v.w;
            }
            uint32_t r0[32], r1[32];
            if (NF == 2 && !(kDbgExp && (p.dbg & 2))) {
                // columns 0..31: even fields (values c) -> accumulator lo; columns 32..63: odd fields in place (values 4c) -> hi
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const uint32_t w4 = w[j] >> 4;
                    r0[2 * j + 0] = w[j] & 0x03030303u;
                    r0[2 * j + 1] = w4 & 0x03030303u;
                    r1[2 * j + 0] = w[j] & 0x0c0c0c0cu;
                    r1[2 * j + 1] = w4 & 0x0c0c0c0cu;
                }
            } else if (!(kDbgExp && (p.dbg & 2))) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    r0[4 * j + 0] = w[j] & 0x03030303u;
                    r0[4 * j + 1] = (w[j] >> 2) & 0x03030303u;
                    r0[4 * j + 2] = (w[j] >> 4) & 0x03030303u;
                    r0[4 * j + 3] = (w[j] >> 6) & 0x03030303u;
                    r1[4 * j + 0] = w[8 + j] & 0x03030303u;
                    r1[4 * j + 1] = (w[8 + j] >> 2) & 0x03030303u;
                    r1[4 * j + 2] = (w[8 + j] >> 4) & 0x03030303u;
                    r1[4 * j + 3] = (w[8 + j] >> 6) & 0x03030303u;
                }
            }
            // pin the 64 expanded words HERE: without it the compiler sinks the whole expansion below the wait for the A slot
            // (to shorten the live ranges), and the slot ring stops hiding the 112 alu-pipe instructions behind that wait
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                asm volatile("" : "+r"(r0[j]), "+r"(r0[j + 1]), "+r"(r0[j + 2]), "+r"(r0[j + 3]), "+r"(r0[j + 4]), "+r"(r0[j + 5]), "+r"(r0[j + 6]), "+r"(r0[j + 7]));
                asm volatile("" : "+r"(r1[j]), "+r"(r1[j + 1]), "+r"(r1[j + 2]), "+r"(r1[j + 3]), "+r"(r1[j + 4]), "+r"(r1[j + 5]), "+r"(r1[j + 6]), "+r"(r1[j + 7]));
            }
            const bool reuse = kbc >= TS_NSA;
            if (reuse) mbar_wait(bar_done(s2), ph2);   // the MMAs that read this A slot have retired
            tc_fence_after();
            if (!(kDbgExp && (p.dbg & 2))) {
                const uint32_t ta = tmem_a0 + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * TS_A_COLS);
                tmem_st32(ta, r0);
                tmem_st32(ta + 32u, r1);
                tmem_wait_st();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_ready(s));
            s += TS_EXP_SETS; if (s >= NS) { s -= NS; ph ^= 1u; }
            if (reuse) { s2 += TS_EXP_SETS; if (s2 >= NS) { s2 -= NS; ph2 ^= 1u; } }
            a += TS_EXP_SETS; if (a >= TS_NSA) a -= TS_NSA;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CL > 1) {   // no CTA may exit while a peer can still multicast into its shared memory or arrive on its barriers
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(Cfg::TMEM_COLS) : "memory");
    }
}

template <typename OT, int BT, int CL, int NF = 1>
static int launch_ts(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed, UmmaParams p, cudaStream_t s,
                     int64_t stride_q_rows, int64_t stride_packed) {
    using Cfg = TsCfg<BT, (int)sizeof(OT), NF>;
    static_assert(BT % (8 * CL) == 0, "token-tile slices must be whole 8-row swizzle atoms");
    CUtensorMap map_act, map_pw;
    const uint64_t L = (uint64_t)p.layers;
    uint64_t dims_a[3] = {(uint64_t)ldq, (uint64_t)p.M, L};
    uint64_t str_a[2] = {(uint64_t)ldq, (uint64_t)(L > 1 ? stride_q_rows * ldq : p.M * ldq)};
    uint32_t box_a[3] = {128, BT / CL, 1};
    int rc = make_tensor_map(&map_act, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, q_x, dims_a, str_a, box_a, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    uint64_t dims_p[3] = {(uint64_t)((p.K + 3) / 4), (uint64_t)p.N, L};
    uint64_t str_p[2] = {(uint64_t)ld_packed, (uint64_t)(L > 1 ? stride_packed : ld_packed * p.N)};
    uint32_t box_p[3] = {TS_BK / 4, TS_ROWS, 1};
    rc = make_tensor_map(&map_pw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, packed, dims_p, str_p, box_p, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
    // output tile stores through TMA need a 16-byte aligned base and row pitch
    CUtensorMap map_out = map_pw;
    PeerMaps peer_maps;
    for (int r = 0; r < 8; ++r) peer_maps.m[r] = map_pw;
    p.tma_out = (BT >= 32) && ((uintptr_t)p.out % 16 == 0) && ((p.ldo * sizeof(OT)) % 16 == 0) &&
                ((p.stride_out * sizeof(OT)) % 16 == 0) && !(kAblate && (p.dbg & 512));
    for (int r = 0; r < p.n_peers; ++r) p.tma_out = p.tma_out && ((uintptr_t)p.peers[r] % 16 == 0);
    if (p.tma_out) {
        uint64_t dims_o[3] = {(uint64_t)p.N, (uint64_t)p.M, L};
        uint64_t str_o[2] = {(uint64_t)(p.ldo * sizeof(OT)), (uint64_t)((L > 1 ? p.stride_out : p.ldo * p.M) * sizeof(OT))};
        uint32_t box_o[3] = {TS_ROWS, 32, 1};
        rc = make_tensor_map(&map_out, sizeof(OT) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (std::is_same<OT, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16),
                             3, p.out, dims_o, str_o, box_o, CU_TENSOR_MAP_SWIZZLE_NONE);
        if (rc) return rc;
        for (int r = 0; r < p.n_peers; ++r) {          // one output map per tensor-parallel rank
            rc = make_tensor_map(&peer_maps.m[r], sizeof(OT) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (std::is_same<OT, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16),
                                 3, p.peers[r], dims_o, str_o, box_o, CU_TENSOR_MAP_SWIZZLE_NONE);
            if (rc) return rc;
        }
        for (int r = p.n_peers; r < 8; ++r) peer_maps.m[r] = map_out;
    }
    auto kern = umma_ts_kernel<OT, BT, CL, NF>;
    static SmemOptIn optin;
    TB_CUDA(ensure_dynamic_smem(kern, optin, Cfg::SMEM_BYTES));
    p.m_tiles = (int)((p.M + BT - 1) / BT);
    p.n_tiles = (int)((p.N + TS_ROWS - 1) / TS_ROWS);
    p.k_blocks = (int)((p.K + TS_BK - 1) / TS_BK);
    const int super_tiles = p.m_tiles * ((p.n_tiles + CL - 1) / CL) * p.layers;
    int clusters = sm_count() / CL;
    if (super_tiles < clusters) clusters = super_tiles;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * CL), 1, 1);
    cfg.blockDim = dim3(TS_THREADS, 1, 1);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (CL > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = CL; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
        ++na;
    }
    if (p.pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, map_act, map_pw, map_out, peer_maps, p));
    TB_LAUNCH_CHECK();
    return TB200_OK;
}

static int ts_cluster_size() {
    // measured: multicast clusters do not pay (the kernel is not L2-bound); an option of the experiment build only
    static const int cl = [] { const int v = (int)tb_knob("TB200_GEMM_CLUSTER", 1); return (v == 2 || v == 4) ? v : 1; }();
    return cl;
}

template <typename OT>
static int ts_dispatch_bt(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed,
                          const UmmaParams& p, cudaStream_t s, int64_t sq, int64_t sp) {
    if (p.M <= 32) {
        // measured and dropped for these tiles: one multicast copy of the activation tile per cluster of 2 / 4 CTAs (the lock-step of
        // the ring across the cluster costs 40 %), and activations resident in shared memory for a whole layer (no gain once the
        // producer warp issued its TMAs from uniform code: the re-reads from L2 are not the bound)
        if (p.qlayout == TB200_QLAYOUT_GEMM_SPLIT) return launch_ts<OT, 32, 1, 2>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
        return launch_ts<OT, 32, 1>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    }
    if (p.qlayout == TB200_QLAYOUT_GEMM_SPLIT) { set_error("gemm_ts: the field-split activation layout serves at most 32 tokens"); return TB200_EINVAL; }
    if (p.M <= 64) return launch_ts<OT, 64, 1>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    if (p.M <= 128) return launch_ts<OT, 128, 1>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    const int n_tiles = (int)((p.N + TS_ROWS - 1) / TS_ROWS);
    const int cl = n_tiles >= 8 ? ts_cluster_size() : 1;
    static const int bt_env = (int)tb_knob("TB200_GEMM_BT", 0);
    int bt = bt_env;
    if (bt == 0) {
        // tiles go round-robin to one CTA per SM: pick the token tile (192 or 160) with the smaller rounds x tile cost
        // (wave quantisation: M=8224, N=1024 is 3 rounds either way, so the smaller tile wins; ties go to 192)
        const int64_t layers = p.layers < 1 ? 1 : p.layers;
        auto cost = [&](int64_t b) { const int64_t tiles = (int64_t)n_tiles * ((p.M + b - 1) / b) * layers; return (tiles + sm_count() - 1) / sm_count() * b; };
        bt = (cl == 1 && cost(160) < cost(192)) ? 160 : 192;
    }
    if constexpr (sizeof(OT) == 2) { if (bt == 160) return launch_ts<OT, 160, 1>(q_x, ldq, packed, ld_packed, p, s, sq, sp); }
    if (cl == 4) return launch_ts<OT, 192, 4>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    if (cl == 2) return launch_ts<OT, 192, 2>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    return launch_ts<OT, 192, 1>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
}

int gemm_ts_launch(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed, UmmaParams p,
                   int out_dtype, cudaStream_t s, int64_t sq, int64_t sp) {
    if (p.layers < 1) p.layers = 1;
    switch (out_dtype) {
        case TB200_F32:  return ts_dispatch_bt<float>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
        case TB200_BF16: return ts_dispatch_bt<__nv_bfloat16>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
        default:         return ts_dispatch_bt<__half>(q_x, ldq, packed, ld_packed, p, s, sq, sp);
    }
}

}  // namespace tb200
