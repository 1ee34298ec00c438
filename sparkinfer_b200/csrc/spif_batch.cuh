// spif_batch.cuh -- device helpers for the batched (n_tok > 1) sparse FFN: tcgen05 tensor-core MMA with accumulators in
// tensor memory, cp.async gathers of weight rows into 128-byte-swizzled shared-memory operand tiles.
//
// Encodings follow the PTX ISA "tcgen05 matrix descriptor" / "instruction descriptor" tables:
//   shared-memory descriptor (64 bit): [0,14) start address >> 4 | [16,30) leading byte offset >> 4 |
//       [32,46) stride byte offset >> 4 | [46,48) version = 1 | [49,52) base offset = 0 | [61,64) swizzle (2 = 128 B)
//   instruction descriptor (32 bit, kind::f16): [4,6) D format (1 = F32) | [7,10) A format | [10,13) B format
//       (0 = F16, 1 = BF16) | bit 15 A major | bit 16 B major (0 = K, 1 = MN) | [17,23) N >> 3 | [24,29) M >> 4
#pragma once
#include "spif_common.cuh"

__device__ __forceinline__ void cp_async16(uint32_t dst, const void * src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// src_bytes == 0 writes 16 zero bytes without touching global memory
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void * src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// asynchronous arrive on `bar` once all cp.async issued so far by this thread have landed; does not change the barrier's
// pending count, so the expected count has to include this arrival
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint64_t * bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// generic-proxy writes to shared memory (st.shared, cp.async) -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void tc_fence_before_sync() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after_sync() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// whole warp: allocate `cols` (power of two >= 32) TMEM columns, base address written to *slot (shared memory)
__device__ __forceinline__ void tmem_alloc(uint32_t * slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t base, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(cols) : "memory");
}

__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout = 2 /* SWIZZLE_128B; 6 = SWIZZLE_32B */) {
    uint64_t d = 0;
    d |= (uint64_t) ((saddr >> 4) & 0x3FFFu);
    d |= (uint64_t) ((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t) ((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t) 1 << 46;  // descriptor version (Blackwell)
    d |= (uint64_t) layout << 61;
    return d;
}

__host__ __device__ constexpr uint32_t umma_idesc_f16(int fmt /*0 F16, 1 BF16*/, int a_mn_major, int b_mn_major, int M, int N) {
    return (1u << 4) | ((uint32_t) fmt << 7) | ((uint32_t) fmt << 10) | ((uint32_t) a_mn_major << 15) |
           ((uint32_t) b_mn_major << 16) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (M >> 4) << 24);
}

// one thread: D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// one thread: arrive on `bar` when all MMAs issued so far by this thread have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t * bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// warp-collective: lane l reads TMEM lane (taddr.lane + l), 16 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        v[i] = __uint_as_float(r[i]);
    }
}

// byte offset of 16-byte chunk `c` (0..7) of row `r` inside a K-major SWIZZLE_128B tile (rows of 128 B = 64 f16 along K,
// 8-row atoms of 1024 B stacked along M/N)
__device__ __forceinline__ uint32_t swz_kmajor(int r, int c) {
    return (uint32_t) ((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

// K-major SWIZZLE_32B tile: rows of 32 B (one K = 16 MMA step), 8-row atoms of 256 B; 16-byte half `h` of row `r`
__device__ __forceinline__ uint32_t swz32_kmajor(int r, int h) {
    return (uint32_t) ((r >> 3) * 256 + (r & 7) * 32 + ((h ^ ((r >> 2) & 1)) << 4));
}

template <typename WT> struct wt_fmt;
template <> struct wt_fmt<spif_f16> { static constexpr int id = 0; };
template <> struct wt_fmt<spif_bf16> { static constexpr int id = 1; };

template <typename WT> __device__ __forceinline__ uint16_t f32_to_wt_bits(float v);
template <> __device__ __forceinline__ uint16_t f32_to_wt_bits<spif_f16>(float v) {
    return __half_as_ushort(__float2half_rn(v));
}
template <> __device__ __forceinline__ uint16_t f32_to_wt_bits<spif_bf16>(float v) {
    return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
