// spif_common.cuh -- shared device helpers for the sm_100a sparse-FFN kernels.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/spif_b200.h"

#define SPIF_WARP 32
#define SPIF_QK   32 /* QK8_0 == QK4_0, ggml-common.h */

// ---------------------------------------------------------------------------------------------------------------
// host-side error plumbing: C-ABI functions never throw and never abort; they return codes.
#define SPIF_CUDA_TRY(expr)                 \
    do {                                    \
        cudaError_t _e = (expr);            \
        if (_e != cudaSuccess) {            \
            return (int) _e;                \
        }                                   \
    } while (0)

#define SPIF_REQUIRE(cond, code) \
    do {                         \
        if (!(cond)) {           \
            return (code);       \
        }                        \
    } while (0)

// ---------------------------------------------------------------------------------------------------------------
// mbarrier + TMA 1-D bulk copy (cp.async.bulk -> SASS UBLKCP), sm_90+/sm_100a

__device__ __forceinline__ uint32_t smem_u32(const void * p) {
    return (uint32_t) __cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

// make barrier inits visible to the async proxy before any TMA targets them
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t * bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

// Blocking wait for a phase.  The suspend-time hint matters: without it try_wait returns after a short system-dependent
// limit and the loop becomes a hot spin -- 20 spinning consumer warps took 5/6 of the issue slots of the scheduler that
// also runs the producer warp, which then needed ~1 us per ring item (profiles/r2_layer_trace.txt: the "22 GB/s per SM
// cap" of every mbarrier ring in this repo).  With the hint the thread sleeps in hardware until the phase completes.
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity) {
    const uint32_t addr  = smem_u32(bar);
    const uint32_t ticks = 0x989680u;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "SPIF_WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
        "@P1 bra SPIF_WAIT_DONE;\n\t"
        "bra SPIF_WAIT_LOOP;\n\t"
        "SPIF_WAIT_DONE:\n\t"
        "}" ::"r"(addr),
        "r"(parity), "r"(ticks)
        : "memory");
}

// the same without the suspend-time hint (default, short, system-dependent limit): for single waiting warps
__device__ __forceinline__ void mbar_wait_spin(uint64_t * bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "SPIF_WAITS_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra SPIF_WAITS_DONE;\n\t"
        "bra SPIF_WAITS_LOOP;\n\t"
        "SPIF_WAITS_DONE:\n\t"
        "}" ::"r"(addr),
        "r"(parity)
        : "memory");
}

// non-blocking probe of a phase: true if the phase with this parity has completed
__device__ __forceinline__ bool mbar_test(uint64_t * bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

// L2 policy for streamed-once weights
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned; completion on `bar` (complete_tx)
__device__ __forceinline__ void tma_load_1d(void * smem_dst, const void * gmem_src, uint32_t bytes, uint64_t * bar,
                                            uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}

__device__ __forceinline__ void tma_load_1d_nohint(void * smem_dst, const void * gmem_src, uint32_t bytes, uint64_t * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// named barrier among a subset of the CTA's warps
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// programmatic dependent launch hooks (no-ops when the launch did not opt in)
__device__ __forceinline__ void pdl_wait_prior_grid() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// streaming 128-bit global load that does not allocate in L1
__device__ __forceinline__ uint4 ldg_stream_v4(const void * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------------------------------------
// element decode: 8 weights packed in one 16-byte vector -> 8 floats

struct spif_f16 {};
struct spif_bf16 {};

template <typename WT> __device__ __forceinline__ void unpack8(const uint4 & v, float (&f)[8]);

template <> __device__ __forceinline__ void unpack8<spif_f16>(const uint4 & v, float (&f)[8]) {
    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 h  = *reinterpret_cast<const __half2 *>(&w[i]);
        const float2  t  = __half22float2(h);
        f[2 * i]         = t.x;
        f[2 * i + 1]     = t.y;
    }
}

template <> __device__ __forceinline__ void unpack8<spif_bf16>(const uint4 & v, float (&f)[8]) {
    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i]     = __uint_as_float(w[i] << 16);
        f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}

// round an F32 through the weight's float type (SPIF_FLAG_CPU_ROUNDING), RNE like ggml-impl.h:401,550
template <typename WT> __device__ __forceinline__ float round_through(float v);
template <> __device__ __forceinline__ float round_through<spif_f16>(float v) {
    return __half2float(__float2half_rn(v));
}
template <> __device__ __forceinline__ float round_through<spif_bf16>(float v) {
    return __bfloat162float(__float2bfloat16_rn(v));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    return v;
}

// the mask convention: active <=> !(p < thr); NaN is active (mm-sparse.cu:22)
__device__ __forceinline__ bool spif_is_active(float p, float thr) {
    return !(p < thr);
}

// activation recipe, src/llama-graph.cpp:1063-1094 (+ silu for the Llama-3 config)
__device__ __forceinline__ float spif_activation(int arch, float g, float u) {
    switch (arch) {
        case SPIF_ARCH_PROSPARSE:
            return ((g > 0.01f) ? g : 0.0f) * u;
        case SPIF_ARCH_RELU_GATED:
            return ((g > 0.f) ? g : 0.f) * ((u > 0.f) ? u : 0.f);
        case SPIF_ARCH_RELU_UP:
            return (u > 0.f) ? u : 0.f;
        default: /* SPIF_ARCH_SILU_GATED */
            return (g / (1.0f + expf(-g))) * u;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// internal launchers (defined in the .cu files, called from spif_api.cu)

struct spif_launch_ctx {
    int          device;
    int          sm_count;
    int          max_smem_optin;
    cudaStream_t stream;
    const spif_shard_comm * comm;  // non-NULL: the fused layer's partial-sum kernel also sums over the GPU shards
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel instantiation, device): `done` is a per-instantiation
// bitmask; setting the attribute twice is harmless, so no locking is needed.
template <typename K> static inline int spif_set_max_smem_once(K kern, size_t smem, int device, unsigned long long & done) {
    const unsigned long long bit = 1ull << (device & 63);
    if (!(done & bit) || smem > 48 * 1024) {
        if (!(done & bit)) {
            SPIF_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            done |= bit;
        }
    }
    return 0;
}

int spif_launch_ffn(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                    size_t ws_bytes, const spif_launch_ctx & ctx);

enum { SPIF_MODE_FUSED = 0, SPIF_MODE_GEMV = 1, SPIF_MODE_AXPY = 2 };

// workspace regions (spif_workspace_size reserves them for 320 CTAs): [partials + stats][predictor hidden][...]
static inline size_t spif_ws_partials_bytes(int n_embd, int n_tok) {
    size_t b = sizeof(float) * (size_t) n_tok * 320 * n_embd + sizeof(int32_t) * (size_t) n_tok * 320 * 2;
    return (b + 255) & ~(size_t) 255;
}


int spif_launch_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                          const float * x, float * sparse_idx_out, float * logits_out, int n_embd, int r, int n_ff,
                          int n_tok, int flags, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx);
int spif_launch_mask_compact(const float * sparse_idx, const int32_t * neuron_idx, int m, int n_ff, int n_tok, float thr,
                             int flags, int32_t * idx_out, int32_t * count_out, int32_t * union_idx, int32_t * union_count,
                             cudaStream_t stream);
int spif_launch_topk_mask(const float * score, int n_ff, int n_tok, int k, float * out, cudaStream_t stream);
int spif_launch_unary(int op, const float * x, float * y, int64_t n, float t, cudaStream_t stream);
int spif_launch_dfr_update(const float * sparse_idx, int n_ff, int n_tok, int g, float thr, float lambda, int ema,
                           float * dfr, cudaStream_t stream);
