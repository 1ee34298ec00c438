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
// hidden values of the union GEMM travel in the weight float type: F16 would turn |h| > 65504 into inf (and y into inf / NaN)
// where the reference CUDA kernels, which keep hidden in F32 (axpy-sparse.cu:44), stay finite -> saturate; NaN stays NaN
template <typename WT> __device__ __forceinline__ float saturate_to_wt(float v) { return v; }
template <> __device__ __forceinline__ float saturate_to_wt<spif_f16>(float v) {
    return (v != v) ? v : fminf(fmaxf(v, -65504.0f), 65504.0f);
}

// ---------------------------------------------------------------------------------------------------------------
// Q8_0 / Q4_0 block PAIRS (64 consecutive weights) -> 64 F16 values = 8 x 16-byte operand chunks.
// Layouts: block_q8_0 {half d; int8 qs[32]} (34 B), block_q4_0 {half d; uint8 qs[16]} (18 B, element j = low nibble of
// qs[j], element j + 16 = high nibble; value = (nibble - 8) * d)   -- ggml/src/ggml-common.h:170-174,211-215.
// A pair starts 4-byte aligned (68 / 36 bytes), the second block of a pair does not: bytes are re-aligned with prmt.
// Conversion without int->float instructions: byte u -> half bits 0x6400 | u == 1024 + u exactly; subtracting the bias
// (1152 for int8 ^ 0x80, 1032 for nibbles) is exact in F16, the product with d rounds once: w16 = RN_f16(q * d).
template <int QT> struct qpair;  // QT = ggml type id

__device__ __forceinline__ uint32_t h2_sub_mul(uint32_t biased, uint32_t bias2, uint32_t d2) {
    __half2 v = *reinterpret_cast<__half2 *>(&biased);
    v         = __hmul2(__hsub2(v, *reinterpret_cast<__half2 *>(&bias2)), *reinterpret_cast<__half2 *>(&d2));
    return *reinterpret_cast<uint32_t *>(&v);
}

template <> struct qpair<SPIF_TYPE_Q8_0> {
    static constexpr int BS = 34, NW = 17;
    // 4 int8 -> two half2
    static __device__ __forceinline__ void cvt4(uint32_t q, uint32_t d2, uint32_t & lo, uint32_t & hi) {
        const uint32_t x = q ^ 0x80808080u;
        lo               = h2_sub_mul(__byte_perm(x, 0x64646464u, 0x4140), 0x64806480u /* 1152 */, d2);
        hi               = h2_sub_mul(__byte_perm(x, 0x64646464u, 0x4342), 0x64806480u, d2);
    }
    static __device__ __forceinline__ void dequant(const uint32_t (&w)[NW], uint4 (&out)[8]) {
        const uint32_t d0 = (w[0] & 0xffffu) * 0x10001u, d1 = (w[8] >> 16) * 0x10001u;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t qa = __byte_perm(w[2 * c], w[2 * c + 1], 0x5432), qb = __byte_perm(w[2 * c + 1], w[2 * c + 2], 0x5432);
            cvt4(qa, d0, out[c].x, out[c].y);
            cvt4(qb, d0, out[c].z, out[c].w);
            cvt4(w[9 + 2 * c], d1, out[4 + c].x, out[4 + c].y);
            cvt4(w[10 + 2 * c], d1, out[4 + c].z, out[4 + c].w);
        }
    }
};

template <> struct qpair<SPIF_TYPE_Q4_0> {
    static constexpr int BS = 18, NW = 9;
    // one 4-byte q word -> low nibbles of its 4 bytes (lo0, lo1) and high nibbles (hi0, hi1), each a half2
    static __device__ __forceinline__ void cvt8(uint32_t q, uint32_t d2, uint32_t & lo0, uint32_t & lo1, uint32_t & hi0, uint32_t & hi1) {
        const uint32_t t01 = __byte_perm(q, 0u, 0x4140), t23 = __byte_perm(q, 0u, 0x4342);
        const uint32_t bias = 0x64086408u;  // 1032
        lo0                 = h2_sub_mul((t01 & 0x000F000Fu) | 0x64006400u, bias, d2);
        lo1                 = h2_sub_mul((t23 & 0x000F000Fu) | 0x64006400u, bias, d2);
        hi0                 = h2_sub_mul(((t01 >> 4) & 0x000F000Fu) | 0x64006400u, bias, d2);
        hi1                 = h2_sub_mul(((t23 >> 4) & 0x000F000Fu) | 0x64006400u, bias, d2);
    }
    static __device__ __forceinline__ void block(const uint32_t (&q)[4], uint32_t d2, uint4 * out) {
        // chunk 0 = elements 0..7 (low nibbles of q words 0,1), chunk 1 = 8..15, chunk 2 = 16..23 (high nibbles), chunk 3
        cvt8(q[0], d2, out[0].x, out[0].y, out[2].x, out[2].y);
        cvt8(q[1], d2, out[0].z, out[0].w, out[2].z, out[2].w);
        cvt8(q[2], d2, out[1].x, out[1].y, out[3].x, out[3].y);
        cvt8(q[3], d2, out[1].z, out[1].w, out[3].z, out[3].w);
    }
    static __device__ __forceinline__ void dequant(const uint32_t (&w)[NW], uint4 (&out)[8]) {
        const uint32_t d0 = (w[0] & 0xffffu) * 0x10001u, d1 = (w[4] >> 16) * 0x10001u;
        uint32_t       q0[4], q1[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            q0[k] = __byte_perm(w[k], w[k + 1], 0x5432);
            q1[k] = w[5 + k];
        }
        block(q0, d0, out);
        block(q1, d1, out + 4);
    }
};
