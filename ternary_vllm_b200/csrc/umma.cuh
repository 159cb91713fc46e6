// umma.cuh — tcgen05 (UMMA / TMEM) PTX wrappers shared by the tensor-core GEMM kernels.
#pragma once
#include "tma.cuh"

namespace tb200 {

// ---- tcgen05 PTX wrappers (mbarrier / TMA wrappers live in tma.cuh) ------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (rows of 128 bytes, 8-row atoms of
// 1024 bytes): start>>4 | LBO(16B, ignored for swizzled K-major)<<16 | SBO(1024B)>>4<<32 |
// version 1 <<46 | layout SWIZZLE_128B(2) <<61
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// instruction descriptor, kind::i8: D=s32, A=u8 (codes), B=s8 (activations), both K-major
__host__ __device__ constexpr uint32_t umma_idesc_i8(int m, int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | (0u << 15) | (0u << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_i8_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// arrive on the mbarrier at this shared-memory offset in every CTA of cta_mask when the MMAs retire
__device__ __forceinline__ void umma_commit_mc(uint32_t bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(cta_mask) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// split form: issue the TMEM load, overlap its latency with other work, then wait; the wait names the destination
// registers as in/out operands so that no use of them can be scheduled above it
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32_wait(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
        : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
          "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
          "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
          "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
        :: "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
// Whole-warp forms: every lane executes the (warp-uniform) address arithmetic, so the compiler keeps it in uniform
// registers, and one elected lane issues.  The single-thread forms above make the issuing loop divergent code whose
// operands reach the uniform datapath through ELECT + R2UR.BROADCAST chains: ~13 dependent instructions per MMA, which
// bounds skinny token tiles (the MMAs themselves are short).
__device__ __forceinline__ void umma_i8_ts_warp(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
// One k-block of the TS kernel (256 int8 of K = two 128-byte activation sub-tiles x four K=32 slices) issued under ONE election,
// followed by the commit that frees the stage: 8 MMAs + commit = 10 issued instructions instead of 8 x (setp + elect + mma) + elect
// + commit with their operand moves.  A columns a .. a + 63; sub-tile 0 accumulates into d0, sub-tile 1 into d1 (the same
// accumulator, or the "hi" one of the field-split form, whose first MMA then also takes the `accum` flag).
template <bool SPLIT>
__device__ __forceinline__ void umma_i8_ts_kblock_warp(uint32_t d0, uint32_t d1, uint32_t a, uint64_t bdesc0, uint64_t bdesc1,
                                                       uint32_t idesc, uint32_t accum, uint32_t bar_done) {
    asm volatile(
        "{\n\t.reg .pred p, p1, q, t;\n\t"
        ".reg .b32 a1, a2, a3, a4, a5, a6, a7;\n\t"
        ".reg .b64 b01, b02, b03, b11, b12, b13;\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "setp.eq.b32 t, %6, %6;\n\t"
        "setp.ne.or.b32 p1, %8, 0, p;\n\t"
        "add.u32 a1, %1, 8;\n\tadd.u32 a2, %1, 16;\n\tadd.u32 a3, %1, 24;\n\tadd.u32 a4, %1, 32;\n\t"
        "add.u32 a5, %1, 40;\n\tadd.u32 a6, %1, 48;\n\tadd.u32 a7, %1, 56;\n\t"
        "add.u64 b01, %2, 2;\n\tadd.u64 b02, %2, 4;\n\tadd.u64 b03, %2, 6;\n\t"
        "add.u64 b11, %3, 2;\n\tadd.u64 b12, %3, 4;\n\tadd.u64 b13, %3, 6;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %5, p;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], [a1], b01, %5, t;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], [a2], b02, %5, t;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], [a3], b03, %5, t;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%4], [a4], %3, %5, p1;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%4], [a5], b11, %5, t;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%4], [a6], b12, %5, t;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%4], [a7], b13, %5, t;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%7];\n\t}"
        ::"r"(d0), "r"(a), "l"(bdesc0), "l"(bdesc1), "r"(d1), "r"(idesc), "r"(accum), "r"(bar_done), "r"(SPLIT ? 0u : 1u) : "memory");
}
__device__ __forceinline__ void umma_commit_warp(uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}
// 32 lanes x 32 columns (one 32-bit register per column) from registers into TMEM
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
        "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct UmmaParams {
    const float* s_x; const int32_t* sum_qx;
    const float* scale; const float* bias;
    void* out; int64_t ldo; int32_t* acc_out;
    int64_t M, N, K;
    int m_tiles, n_tiles, k_blocks;
    float clip_val;
    int tma_out;      // epilogue stores through TMA (aligned output)
    // grouped form (gemm_ts only): `layers` problems of identical shape in one persistent launch
    int layers;                       // >= 1
    int64_t stride_act_rows;          // s_x / sum_qx elements per layer (= M)
    int64_t stride_sb;                // scale / bias elements per layer
    int64_t stride_out;               // out elements per layer
    void* peers[8]; int n_peers;      // tensor-parallel gather: store every element into each rank's buffer (see gemv.cu)
    int mc;           // n_peers == 1 and peers[0] is an NVSwitch multicast address: ONE multimem store per 16 bytes reaches every rank
    int pdl;          // launched with programmatic stream serialisation: the activation quantiser may still be running
    int qlayout;      // 0 / TB200_QLAYOUT_GEMM: activations in layout 1; TB200_QLAYOUT_GEMM_SPLIT (M <= 32, ldq % 256 == 0): field-split kernel
    int dbg;          // TB200_GEMM_DBG experiments: 1 skip the epilogue body, 2 skip the expansion, 4 skip the MMAs
};

// shared epilogue of the tensor-core kernels: TMEM accumulator tile -> out (lane = output channel)
int gemm_ts_launch(const int8_t* q_x, int64_t ldq, const uint8_t* packed, int64_t ld_packed, UmmaParams p,
                   int out_dtype, cudaStream_t s, int64_t stride_q_rows = 0, int64_t stride_packed = 0);

}  // namespace tb200
