// Tensor Memory (TMEM, 256 KB per SM on sm_100a: 128 lanes x 512 columns x 32 bit) used as
// per-thread private storage.  tcgen05.ld / tcgen05.st in the 32x32b shape move N consecutive
// columns of ONE lane per thread; warp w may touch lanes 32*(w%4) .. 32*(w%4)+31, so thread
// (w, lane) owns TMEM row 32*(w%4)+lane and the two warps that share a lane quarter (w, w+4) use
// different column ranges.  No tensor-core instruction is involved: the jet kernel parks its per-layer
// stash there instead of round-tripping it through L2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace tdq {

// one warp allocates all 512 columns for the CTA and publishes the base address through shared memory
__device__ __forceinline__ void tmem_alloc_512(uint32_t* smem_slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
        (uint32_t)__cvta_generic_to_shared(smem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_512(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 16 consecutive columns of this thread's lane  <->  16 registers
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// the caller must tmem_wait_ld() before reading v

// 8-column variants (every per-layer stash size C*TP*4 is a multiple of 8)
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                 "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
                 "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// 4-column variants (16-point tiles: 4 points per thread and channel)
__device__ __forceinline__ void tmem_st4(uint32_t taddr, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(__float_as_uint(v[0])),
                 "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3]))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}

// TMEM address of column `col` in this thread's row: lane quarter of the warp in bits 31:16
__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, int warp, int col) {
    return base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)col;
}

// ---- tcgen05.mma (kind::tf32, fp32 accumulate in TMEM), operands in shared memory -----------------------
// Shared-memory matrix descriptor of a K-major operand in the SWIZZLE_128B canonical layout: one row of the
// matrix = one 128-byte line (32 fp32 along K), 16-byte chunk index XOR (row & 7), groups of 8 rows 1024 bytes
// apart; `addr` (shared-window byte address) is the 1024-byte aligned base of the plane, plus 32 bytes per K step
// of 8.  Field positions checked against the instruction by tdq_ffma_probe modes 42 / 43.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                          // leading byte offset: not used by swizzled K-major layouts
    d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;     // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                          // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                          // SWIZZLE_128B
    return d;
}
// instruction descriptor: tf32 x tf32 -> f32, both operands K-major, M x N output tile
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, K = 8; issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, int accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a), "l"(b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Lean forms for the issue loops (one thread issues every MMA of the CTA, so each extra instruction there costs
// tensor-core time): the accumulate flag is an immediate.  Call them under `if (leader)` from a warp whose other
// lanes run the same (warp-uniform) control flow, so that the descriptors are computed in uniform registers.
template <bool ACC>
__device__ __forceinline__ void umma_tf32_i(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc) {
    if (ACC) asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 1;" ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc) : "memory");
    else asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 0;" ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc) : "memory");
}
// Two consecutive MMAs with the SAME A operand (W x a_lo then W x a_hi; a_hi x zbar_lo then a_hi x zbar_hi): the first
// keeps A in the tensor core's collector buffer (.collector::a::fill, SASS A_KEEP), the second takes it from there
// (.collector::a::lastuse, SASS A_REUSE) instead of reading its 4 KB from shared memory again.
template <bool ACC>
__device__ __forceinline__ void umma_tf32_keep_i(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc) {
    if (ACC) asm volatile("tcgen05.mma.cta_group::1.kind::tf32.collector::a::fill [%0], %1, %2, %3, 1;" ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc) : "memory");
    else asm volatile("tcgen05.mma.cta_group::1.kind::tf32.collector::a::fill [%0], %1, %2, %3, 0;" ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void umma_tf32_reuse_i(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc) {
    asm volatile("tcgen05.mma.cta_group::1.kind::tf32.collector::a::lastuse [%0], %1, %2, %3, 1;" ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc) : "memory");
}
// A operand in Tensor Memory: lanes = the M rows, 8 consecutive 32-bit columns at a_tmem = K
template <bool ACC>
__device__ __forceinline__ void umma_tf32_ts_i(uint32_t d_tmem, uint32_t a_tmem, uint64_t b, uint32_t idesc) {
    if (ACC) asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, 1;" ::"r"(d_tmem), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
    else asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, 0;" ::"r"(d_tmem), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
}
// kind::f16 with bf16 inputs, fp32 accumulate, K = 16 per instruction; the A operand in Tensor Memory holds two bf16 per
// 32-bit column (even K in the low half): 8 consecutive columns at a_tmem.  Layouts pinned by tools/experiments/bf16_probe.cu.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
template <bool ACC>
__device__ __forceinline__ void umma_bf16_ts_i(uint32_t d_tmem, uint32_t a_tmem, uint64_t b, uint32_t idesc) {
    if (ACC) asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 1;" ::"r"(d_tmem), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
    else asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 0;" ::"r"(d_tmem), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
}
// the mbarrier gets one arrival when every tcgen05.mma this thread issued so far has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
        (uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads, TMA)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace tdq
