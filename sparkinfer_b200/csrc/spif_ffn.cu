// spif_ffn.cu -- the sparse-FFN hot path for F16/BF16 weights on sm_100a.
//
// One persistent kernel per layer (SPIF_MODE_FUSED) replaces the reference's
//   memset + mul_mat_vec_sparse(up) + memset + mul_mat_vec_sparse(gate) + fatrelu + mul + memset +
//   mul_mat_axpy_sparse_rowwise        (mm-sparse.cu:10-102,397; unary.cu:571; axpy-sparse.cu:16-86)
// and the same kernel template, in SPIF_MODE_GEMV / SPIF_MODE_AXPY, is the drop-in for the two single ops.
//
// Design (B200: 148 SMs, HBM-bound, ~10 active neurons per SM per layer at 10% activity):
//   * grid = one CTA per SM.  Every CTA scans the float mask itself (warp ballot + prefix scan, ascending order) and
//     keeps the active rows whose global rank r satisfies r % gridDim.x == blockIdx.x -> static, balanced to +-1 row,
//     deterministic, no extra launch and no global index list on the critical path.
//   * a producer warp streams each active neuron's gate and up rows (contiguous, 16-byte aligned) into a shared-memory
//     ring with TMA 1-D bulk copies (cp.async.bulk + mbarrier complete_tx); with 8 stages x 20 KB practically all of a
//     CTA's rows are in flight from the first microsecond (Little: 6.5 TB/s x ~1 us / 148 SMs ~= 44 KB per SM).
//   * consumer threads own fixed 16-byte column slices: x lives in registers, the gate/up dots are FP32, reduced with
//     shuffles + one named barrier per batch of neurons; the activation is applied in-register, and the neuron's
//     W_down^T row (contiguous because the reference stores ffn_down transposed) is fetched by TMA only if hid != 0
//     (axpy-sparse.cu:53) and FMA'd into per-thread FP32 accumulators -> no grid-wide dependency between the GEMV
//     and the axpy, no dense hidden vector, no float atomics.
//   * each CTA writes one partial y; a tiny second kernel (launched with programmatic dependent launch) adds the
//     partials in a fixed order -> run-to-run bit-stable, unlike axpy-sparse.cu:84.
#include "spif_common.cuh"

#include <cstdlib>

#define SPIF_NB        2   /* neurons per reduction batch */
#define SPIF_MAX_WARPS 17  /* 16 consumer warps + 1 producer */

struct FfnKParams {
    const unsigned char * w_up;
    const unsigned char * w_gate;
    const unsigned char * w_down;
    const float *         b_up;
    const float *         b_gate;
    const float *         x;          // [n_tok][n_embd]
    const float *         mask;       // [n_tok][mask_len]
    const float *         hid_in;     // MODE_AXPY: [n_tok][n_ff]
    const int32_t *       neuron_idx; // nullable [m]
    float *               partials;   // [n_tok][P][n_embd]
    float *               dense_out;  // GEMV: dst [n_tok][n_ff]; FUSED: hidden_out or NULL
    int32_t *             cta_stats;  // [n_tok][P][2]
    int                   n_embd, n_ff, m, mask_len;
    int                   V;          // 16-byte vectors per row
    int                   row_bytes;
    int                   nc;         // consumer threads (multiple of 32)
    int                   S, D;       // ring depths
    int                   list_cap;
    float                 thr;
    int                   arch, flags;
    int                   zero_fill;  // GEMV with identity mapping: zero-fill inactive dst entries in-kernel
    unsigned int          magic;      // floor(2^32 / P) + 1: rank / P == __umulhi(rank, magic) for rank < 2^20
    unsigned long long *  trace;      // debug: [P][8] globaltimer stamps (spif_debug_set_trace), normally NULL
};

static unsigned long long * g_spif_trace = nullptr;  // debug only; see tools/trace_ffn.py
extern "C" void spif_debug_set_trace(void * dev_ptr) {
    g_spif_trace = (unsigned long long *) dev_ptr;
}

__device__ __forceinline__ unsigned long long gtime_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define SPIF_STAMP(k)                                                         \
    do {                                                                      \
        if (p.trace && threadIdx.x == 0 && blockIdx.y == 0) {                 \
            p.trace[(size_t) blockIdx.x * 8 + (k)] = gtime_ns();              \
        }                                                                     \
    } while (0)

// ---------------------------------------------------------------------------------------------------------------
// mask -> this CTA's share of the ascending active list.  All threads of the CTA take part.
template <int MODE, typename WT>
__device__ __forceinline__ void build_my_list(const FfnKParams & p, const float * mask, const float * hid, float * dense,
                                              int * warp_cnt, int * list_row, int * list_neu, float * list_alpha,
                                              int * s_counts) {
    const int tid   = threadIdx.x;
    const int lane  = tid & 31;
    const int warp  = tid >> 5;
    const int nwarp = blockDim.x >> 5;
    const int P     = gridDim.x;
    const int c     = blockIdx.x;
    if (p.flags & SPIF_INTERNAL_DENSE) {  // dense GEMV (predictor up projection): every row, round-robin over CTAs
        const int cnt = (p.m > c) ? (p.m - c + P - 1) / P : 0;
        for (int k = tid; k < cnt; k += blockDim.x) {
            list_row[k] = c + k * P;
            list_neu[k] = c + k * P;
        }
        if (tid == 0) {
            s_counts[0] = cnt;
            s_counts[1] = p.m;
        }
        __syncthreads();
        return;
    }
    // Each warp owns a contiguous chunk of rows (a multiple of 128 long), so global ranks are ordered by
    // (warp, position in chunk).  Two passes: count, then rank = warp base + running popcount.  All loads of a batch
    // are unconditional (clamped addresses, no branches) so ptxas issues them back to back: a pass costs one memory
    // round trip per batch, not one per 32 rows.  Pass 2 re-reads the mask from L1.
    const int  chunk    = ((p.m + nwarp * 128 - 1) / (nwarp * 128)) * 128;
    const int  r0       = min(warp * chunk, p.m);
    const int  r1       = min(r0 + chunk, p.m);
    const bool indirect = (p.neuron_idx != nullptr);
    const bool by_row   = (p.flags & SPIF_FLAG_MASK_BY_ROW) != 0;
    const bool rounding = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
    const bool vec_ok   = !indirect && (p.m % 4 == 0) && ((reinterpret_cast<uintptr_t>(mask) & 15) == 0) &&
                        (MODE != SPIF_MODE_AXPY || (reinterpret_cast<uintptr_t>(hid) & 15) == 0);

    auto alpha_of = [&](float h) -> float { return rounding ? round_through<WT>(h) : h; };
    int  cnt = 0, before = 0, total = 0;

    if (vec_ok) {
        // ---- direct mask: 128-bit loads, lane l holds rows base + 4l .. 4l+3 of each 128-row group
        constexpr int  VB   = (MODE == SPIF_MODE_AXPY) ? 5 : 10;  // 128 * VB rows per warp and batch
        const float4 * m4   = reinterpret_cast<const float4 *>(mask);
        const float4 * h4   = reinterpret_cast<const float4 *>(hid);
        const int      last = p.m / 4 - 1;
        auto bits_of = [&](const float4 & v, const float4 & h, int row) -> unsigned {
            unsigned b = 0;
            b |= (row + 0 < r1 && spif_is_active(v.x, p.thr)) ? 1u : 0u;
            b |= (row + 1 < r1 && spif_is_active(v.y, p.thr)) ? 2u : 0u;
            b |= (row + 2 < r1 && spif_is_active(v.z, p.thr)) ? 4u : 0u;
            b |= (row + 3 < r1 && spif_is_active(v.w, p.thr)) ? 8u : 0u;
            if (MODE == SPIF_MODE_AXPY) {  // axpy-sparse.cu:53, ggml-cpu.c:2208: alpha == 0 rows are skipped
                b &= (alpha_of(h.x) == 0.0f ? 0u : 1u) | (alpha_of(h.y) == 0.0f ? 0u : 2u) |
                     (alpha_of(h.z) == 0.0f ? 0u : 4u) | (alpha_of(h.w) == 0.0f ? 0u : 8u);
            }
            return b;
        };
        for (int base = r0; base < r1; base += 128 * VB) {
            float4 v[VB], h[VB];
#pragma unroll
            for (int j = 0; j < VB; ++j) {
                const int f = min((base + 128 * j) / 4 + lane, last);
                v[j]        = __ldg(m4 + f);
                h[j]        = (MODE == SPIF_MODE_AXPY) ? __ldg(h4 + f) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int j = 0; j < VB; ++j) {
                cnt += __popc(bits_of(v[j], h[j], base + 128 * j + 4 * lane));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        }
        if (lane == 0) {
            warp_cnt[warp] = cnt;
        }
        __syncthreads();
        for (int w = 0; w < nwarp; ++w) {
            const int x = warp_cnt[w];
            before += (w < warp) ? x : 0;
            total += x;
        }
        int run = before;
        for (int base = r0; base < r1; base += 128 * VB) {
            float4 v[VB], h[VB];
#pragma unroll
            for (int j = 0; j < VB; ++j) {
                const int f = min((base + 128 * j) / 4 + lane, last);
                v[j]        = __ldg(m4 + f);
                h[j]        = (MODE == SPIF_MODE_AXPY) ? __ldg(h4 + f) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int j = 0; j < VB; ++j) {
                const int      row0 = base + 128 * j + 4 * lane;
                const unsigned b    = bits_of(v[j], h[j], row0);
                const int      mine = __popc(b);
                int            incl = mine;  // inclusive prefix of `mine` over lanes
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, o);
                    incl += (lane >= o) ? t : 0;
                }
                int         rank  = run + incl - mine;
                const float hv[4] = { h[j].x, h[j].y, h[j].z, h[j].w };
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = row0 + e;
                    if (b & (1u << e)) {
                        const int k = (int) __umulhi((unsigned) rank, p.magic);  // rank / P without a divide
                        if (rank - k * P == c) {
                            list_row[k] = row;
                            list_neu[k] = row;
                            if (MODE == SPIF_MODE_AXPY) {
                                list_alpha[k] = alpha_of(hv[e]);
                            }
                        }
                        ++rank;
                    } else if (MODE == SPIF_MODE_GEMV && p.zero_fill && row < r1 && (row - (int) __umulhi((unsigned) row, p.magic) * P) == c) {
                        dense[row] = 0.0f;  // dst[neu] of inactive neurons (mm-sparse.cu:397)
                    }
                }
                run += __shfl_sync(0xffffffffu, incl, 31);
            }
        }
    } else {
        // ---- gathered mask (GPU cache / shard rows -> neuron ids) or unaligned: scalar loads, batches of 8
        constexpr int SB    = 8;
        const int     lastr = max(r1 - 1, 0);
        for (int pass = 0; pass < 2; ++pass) {
            int run = before;
            for (int base = r0; base < r1; base += 32 * SB) {
                int   neu[SB];
                float v[SB], al[SB];
#pragma unroll
                for (int j = 0; j < SB; ++j) {
                    const int row = min(base + j * 32 + lane, lastr);
                    neu[j]        = indirect ? __ldg(p.neuron_idx + row) : row;
                }
#pragma unroll
                for (int j = 0; j < SB; ++j) {
                    const int row = min(base + j * 32 + lane, lastr);
                    v[j]          = __ldg(mask + (by_row ? row : neu[j]));
                    al[j]         = (MODE == SPIF_MODE_AXPY) ? alpha_of(__ldg(hid + neu[j])) : 1.0f;
                }
#pragma unroll
                for (int j = 0; j < SB; ++j) {
                    const int      row  = base + j * 32 + lane;
                    const bool     a    = (row < r1) && spif_is_active(v[j], p.thr) && !(al[j] == 0.0f);
                    const unsigned ball = __ballot_sync(0xffffffffu, a);
                    if (pass == 0) {
                        cnt += __popc(ball);
                    } else {
                        if (a) {
                            const int rank = run + __popc(ball & ((1u << lane) - 1u));
                            const int k = (int) __umulhi((unsigned) rank, p.magic);
                            if (rank - k * P == c) {
                                list_row[k] = row;
                                list_neu[k] = neu[j];
                                if (MODE == SPIF_MODE_AXPY) {
                                    list_alpha[k] = al[j];
                                }
                            }
                        } else if (MODE == SPIF_MODE_GEMV && p.zero_fill && row < r1 && (row - (int) __umulhi((unsigned) row, p.magic) * P) == c) {
                            dense[row] = 0.0f;
                        }
                        run += __popc(ball);
                    }
                }
            }
            if (pass == 0) {
                if (lane == 0) {
                    warp_cnt[warp] = cnt;
                }
                __syncthreads();
                for (int w = 0; w < nwarp; ++w) {
                    const int x = warp_cnt[w];
                    before += (w < warp) ? x : 0;
                    total += x;
                }
            }
        }
    }
    if (tid == 0) {
        s_counts[0] = (total > c) ? (total - c + P - 1) / P : 0;  // my share
        s_counts[1] = total;
    }
    __syncthreads();
}

template <typename WT, int VPT>
__device__ __forceinline__ void axpy_row(const uint4 * row, float alpha, float (&acc)[VPT * 8], int ctid, int nc, int V) {
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = ctid + k * nc;
        if (v < V) {
            float w[8];
            unpack8<WT>(row[v], w);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                acc[k * 8 + j] = fmaf(w[j], alpha, acc[k * 8 + j]);
            }
        }
    }
}

template <typename WT, int VPT, int MODE>
__global__ void __launch_bounds__(SPIF_MAX_WARPS * 32, 1) spif_ffn_kernel(const FfnKParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int nc   = p.nc;
    const int nwc  = nc >> 5;  // consumer warps; warp index nwc is the producer
    const int P    = gridDim.x;
    const int c    = blockIdx.x;
    const int tok  = blockIdx.y;
    const int V    = p.V;

    const bool gated       = (MODE == SPIF_MODE_FUSED) && (p.arch != SPIF_ARCH_RELU_UP);
    const int  nmat        = gated ? 2 : 1;
    const int  stage_bytes = nmat * p.row_bytes;

    // ---- shared memory carve-up
    unsigned char * ring  = smem;
    unsigned char * dring = ring + (size_t) p.S * stage_bytes;
    unsigned char * tail  = dring + (size_t) ((MODE == SPIF_MODE_FUSED) ? p.D : 0) * p.row_bytes;
    uint64_t *      full  = reinterpret_cast<uint64_t *>(tail);
    uint64_t *      empty = full + p.S;
    uint64_t *      dfull = empty + p.S;
    float *         red   = reinterpret_cast<float *>(dfull + p.D);      // [2][16][NB][2]
    int *           warp_cnt   = reinterpret_cast<int *>(red + 2 * 16 * SPIF_NB * 2);  // [32]
    int *           s_counts   = warp_cnt + 32;                          // [4]
    int *           list_row   = s_counts + 4;
    int *           list_neu   = list_row + p.list_cap;
    float *         list_alpha = reinterpret_cast<float *>(list_neu + p.list_cap);

    if (tid == 0) {
        for (int s = 0; s < p.S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], nwc);
        }
        for (int d = 0; d < p.D; ++d) {
            mbar_init(&dfull[d], 1);
        }
        mbar_fence_init();
    }
    SPIF_STAMP(0);
    // everything above overlaps the previous kernel's tail under programmatic dependent launch
    pdl_wait_prior_grid();
    SPIF_STAMP(1);

    const float * mask  = p.mask + (size_t) tok * p.mask_len;
    const float * xg    = p.x ? p.x + (size_t) tok * p.n_embd : nullptr;
    const float * hid   = p.hid_in ? p.hid_in + (size_t) tok * p.n_ff : nullptr;
    float *       dense = p.dense_out ? p.dense_out + (size_t) tok * p.n_ff : nullptr;

    build_my_list<MODE, WT>(p, mask, hid, dense, warp_cnt, list_row, list_neu, list_alpha, s_counts);  // ends with __syncthreads
    const int my_count = s_counts[0];
    pdl_launch_dependents();
    SPIF_STAMP(2);

    if (warp == nwc) {
        // ================= producer warp: stream weight rows of my active neurons =================
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            for (int i = 0; i < my_count; ++i) {
                const int s = i % p.S;
                if (i >= p.S) {
                    mbar_wait(&empty[s], ((i / p.S) - 1) & 1);
                }
                const size_t off = (size_t) list_row[i] * p.row_bytes;
                mbar_arrive_expect_tx(&full[s], stage_bytes);
                unsigned char * dst = ring + (size_t) s * stage_bytes;
                if (MODE == SPIF_MODE_AXPY) {
                    tma_load_1d(dst, p.w_down + off, p.row_bytes, &full[s], pol);
                } else if (MODE == SPIF_MODE_GEMV) {
                    tma_load_1d(dst, p.w_up + off, p.row_bytes, &full[s], pol);
                } else {
                    tma_load_1d(dst, p.w_up + off, p.row_bytes, &full[s], pol);
                    if (gated) {
                        tma_load_1d(dst + p.row_bytes, p.w_gate + off, p.row_bytes, &full[s], pol);
                    }
                }
            }
        }
        return;
    }
    if (warp > nwc) {
        return;
    }

    // ================= consumer warps =================
    const int ctid = tid;  // consumer thread id == tid (< nc)
    float     acc[VPT * 8];
#pragma unroll
    for (int j = 0; j < VPT * 8; ++j) {
        acc[j] = 0.f;
    }
    int n_nonzero = 0;

    if (MODE == SPIF_MODE_AXPY) {
        for (int i = 0; i < my_count; ++i) {
            const int s = i % p.S;
            mbar_wait(&full[s], (i / p.S) & 1);
            axpy_row<WT, VPT>(reinterpret_cast<const uint4 *>(ring + (size_t) s * stage_bytes), list_alpha[i], acc, ctid,
                              nc, V);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&empty[s]);
            }
        }
        n_nonzero = my_count;
    } else {
        // x slice in registers
        float xr[VPT * 8];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xr[k * 8 + j] = 0.f;
            }
            if (v < V) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v + 1);
                xr[k * 8 + 0] = a.x, xr[k * 8 + 1] = a.y, xr[k * 8 + 2] = a.z, xr[k * 8 + 3] = a.w;
                xr[k * 8 + 4] = b.x, xr[k * 8 + 5] = b.y, xr[k * 8 + 6] = b.z, xr[k * 8 + 7] = b.w;
                if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        xr[k * 8 + j] = round_through<WT>(xr[k * 8 + j]);
                    }
                }
            }
        }

        const uint64_t pol = l2_policy_evict_first();
        int   q_issued = 0, q_done = 0;  // down-row queue (FUSED): sequence numbers, identical in every thread
        float pend_h[SPIF_NB];
        int   n_pend = 0;

        const int nbatch = (my_count + SPIF_NB - 1) / SPIF_NB;
        for (int b = 0; b < nbatch; ++b) {
            const int nb = min(SPIF_NB, my_count - b * SPIF_NB);
            float     pu[SPIF_NB], pg[SPIF_NB];
#pragma unroll
            for (int j = 0; j < SPIF_NB; ++j) {
                pu[j] = 0.f, pg[j] = 0.f;
                if (j < nb) {
                    const int i = b * SPIF_NB + j;
                    const int s = i % p.S;
                    mbar_wait(&full[s], (i / p.S) & 1);
                    if (i == 0) {
                        SPIF_STAMP(3);
                    }
                    const uint4 * ur = reinterpret_cast<const uint4 *>(ring + (size_t) s * stage_bytes);
                    const uint4 * gr = reinterpret_cast<const uint4 *>(ring + (size_t) s * stage_bytes + p.row_bytes);
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        const int v = ctid + k * nc;
                        if (v < V) {
                            float w[8];
                            unpack8<WT>(ur[v], w);
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                pu[j] = fmaf(w[e], xr[k * 8 + e], pu[j]);
                            }
                            if (gated) {
                                unpack8<WT>(gr[v], w);
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    pg[j] = fmaf(w[e], xr[k * 8 + e], pg[j]);
                                }
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&empty[s]);  // stage free for the producer
                    }
                }
            }
            // cross-warp reduction of the batch
            float * rb = red + (size_t) (b & 1) * 16 * SPIF_NB * 2;
#pragma unroll
            for (int j = 0; j < SPIF_NB; ++j) {
                pu[j] = warp_sum(pu[j]);
                if (gated) {
                    pg[j] = warp_sum(pg[j]);
                }
            }
            if (lane == 0) {
#pragma unroll
                for (int j = 0; j < SPIF_NB; ++j) {
                    rb[(warp * SPIF_NB + j) * 2 + 0] = pu[j];
                    rb[(warp * SPIF_NB + j) * 2 + 1] = pg[j];
                }
            }
            named_bar_sync(1, nc);
            float hcur[SPIF_NB];
#pragma unroll
            for (int j = 0; j < SPIF_NB; ++j) {
                hcur[j] = 0.f;
                if (j < nb) {
                    float u = 0.f, g = 0.f;
                    for (int w = 0; w < nwc; ++w) {
                        u += rb[(w * SPIF_NB + j) * 2 + 0];
                        g += rb[(w * SPIF_NB + j) * 2 + 1];
                    }
                    const int i   = b * SPIF_NB + j;
                    const int neu = list_neu[i];
                    if (MODE == SPIF_MODE_GEMV) {
                        if (p.flags & SPIF_INTERNAL_DENSE) {  // predictor up: + bias, relu (llama-graph.cpp:871-880)
                            if (p.b_up) {
                                u += __ldg(p.b_up + neu);
                            }
                            u = (u > 0.f) ? u : 0.f;
                        }
                        if (ctid == 0) {
                            dense[neu] = u;
                        }
                    } else {
                        if (p.b_up) {
                            u += __ldg(p.b_up + neu);
                        }
                        if (gated && p.b_gate) {
                            g += __ldg(p.b_gate + neu);
                        }
                        float h = spif_activation(p.arch, g, u);
                        if (ctid == 0 && dense) {
                            dense[neu] = h;
                        }
                        if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
                            h = round_through<WT>(h);
                        }
                        hcur[j] = h;
                    }
                }
            }
            if (MODE == SPIF_MODE_FUSED) {
                // issue this batch's down rows (only where hid != 0), then consume the previous batch's
                float hq[SPIF_NB];
                int   nq = 0;
#pragma unroll
                for (int j = 0; j < SPIF_NB; ++j) {
                    if (j < nb && !(hcur[j] == 0.0f)) {
                        if (ctid == 0) {
                            const int    slot = q_issued % p.D;
                            const size_t off  = (size_t) list_row[b * SPIF_NB + j] * p.row_bytes;
                            mbar_arrive_expect_tx(&dfull[slot], p.row_bytes);
                            tma_load_1d(dring + (size_t) slot * p.row_bytes, p.w_down + off, p.row_bytes, &dfull[slot], pol);
                        }
                        ++q_issued;
                        hq[nq++] = hcur[j];
                    }
                }
                for (int j = 0; j < n_pend; ++j) {
                    const int slot = q_done % p.D;
                    mbar_wait(&dfull[slot], (q_done / p.D) & 1);
                    axpy_row<WT, VPT>(reinterpret_cast<const uint4 *>(dring + (size_t) slot * p.row_bytes), pend_h[j], acc,
                                      ctid, nc, V);
                    ++q_done;
                }
                n_pend = nq;
#pragma unroll
                for (int j = 0; j < SPIF_NB; ++j) {
                    pend_h[j] = hq[j];
                }
                n_nonzero += nq;
            }
        }
        SPIF_STAMP(4);
        if (MODE == SPIF_MODE_FUSED) {
            for (int j = 0; j < n_pend; ++j) {
                const int slot = q_done % p.D;
                mbar_wait(&dfull[slot], (q_done / p.D) & 1);
                axpy_row<WT, VPT>(reinterpret_cast<const uint4 *>(dring + (size_t) slot * p.row_bytes), pend_h[j], acc, ctid,
                                  nc, V);
                ++q_done;
            }
        }
    }

    if (MODE != SPIF_MODE_GEMV) {
        float * part = p.partials + ((size_t) tok * P + c) * p.n_embd;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
            if (v < V) {
                float4 * o = reinterpret_cast<float4 *>(part) + 2 * v;
                o[0]       = make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]);
                o[1]       = make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]);
            }
        }
        if (ctid == 0 && p.cta_stats) {
            int32_t * st = p.cta_stats + ((size_t) tok * P + c) * 2;
            st[0]        = s_counts[1];
            st[1]        = n_nonzero;
        }
    }
    SPIF_STAMP(5);
    if (p.trace && threadIdx.x == 0 && blockIdx.y == 0) {
        p.trace[(size_t) blockIdx.x * 8 + 6] = (unsigned long long) my_count;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// y[t][col] = sum_p partials[t][p][col] (+ b_down[col]) in a fixed order.  block (32 column-vectors, 16 p-chunks):
// every thread has all of its <= SPIF_RED_PER loads in flight at once (one L2 round trip), then a fixed-order smem sum.
#define SPIF_RED_CHUNKS 16
#define SPIF_RED_PER    20 /* ceil(P / 16) for P <= 320 CTAs */
__global__ void __launch_bounds__(32 * SPIF_RED_CHUNKS) spif_reduce_partials_kernel(const float * __restrict__ partials,
                                                                                     const float * __restrict__ b_down,
                                                                                     float * __restrict__ y,
                                                                                     const int32_t * __restrict__ cta_stats,
                                                                                     int32_t * __restrict__ stats_out,
                                                                                     int n_embd, int P) {
    __shared__ float4 sm[SPIF_RED_CHUNKS][32];
    const int         tok  = blockIdx.y;
    const int         col4 = blockIdx.x * 32 + threadIdx.x;
    const int         ch   = threadIdx.y;
    const int         n4   = n_embd >> 2;
    const int         per  = (P + SPIF_RED_CHUNKS - 1) / SPIF_RED_CHUNKS;
    pdl_wait_prior_grid();
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (col4 < n4) {
        const float4 * base = reinterpret_cast<const float4 *>(partials + (size_t) tok * P * n_embd) + col4;
        const int      p0   = ch * per;
        float4         v[SPIF_RED_PER];
#pragma unroll
        for (int q = 0; q < SPIF_RED_PER; ++q) {
            v[q] = (q < per && p0 + q < P) ? base[(size_t) (p0 + q) * n4] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int q = 0; q < SPIF_RED_PER; ++q) {
            s.x += v[q].x, s.y += v[q].y, s.z += v[q].z, s.w += v[q].w;
        }
    }
    sm[ch][threadIdx.x] = s;
    __syncthreads();
    if (ch == 0 && col4 < n4) {
        float4 t = sm[0][threadIdx.x];
#pragma unroll
        for (int k = 1; k < SPIF_RED_CHUNKS; ++k) {
            const float4 v = sm[k][threadIdx.x];
            t.x += v.x, t.y += v.y, t.z += v.z, t.w += v.w;
        }
        if (b_down) {
            const float4 bb = reinterpret_cast<const float4 *>(b_down)[col4];
            t.x += bb.x, t.y += bb.y, t.z += bb.z, t.w += bb.w;
        }
        reinterpret_cast<float4 *>(y + (size_t) tok * n_embd)[col4] = t;
    }
    if (stats_out && blockIdx.x == 0 && threadIdx.x == 0 && ch == 0) {
        int nz = 0;
        for (int q = 0; q < P; ++q) {
            nz += cta_stats[((size_t) tok * P + q) * 2 + 1];
        }
        stats_out[tok * 4 + 0] = cta_stats[((size_t) tok * P) * 2 + 0];
        stats_out[tok * 4 + 1] = nz;
        stats_out[tok * 4 + 2] = 0;
        stats_out[tok * 4 + 3] = 0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host-side launcher

template <typename WT, int VPT, int MODE>
static int launch_t(const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    auto                      kern = spif_ffn_kernel<WT, VPT, MODE>;
    static unsigned long long done = 0;
    int                       rc0  = spif_set_max_smem_once(kern, smem, device, done);
    if (rc0) {
        return rc0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = grid;
    cfg.blockDim           = dim3(threads);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, kp));
    return 0;
}

template <typename WT, int MODE>
static int launch_v(int vpt, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    switch (vpt) {
        case 1:
            return launch_t<WT, 1, MODE>(kp, grid, threads, smem, stream, device);
        case 2:
            return launch_t<WT, 2, MODE>(kp, grid, threads, smem, stream, device);
        default:
            return launch_t<WT, 4, MODE>(kp, grid, threads, smem, stream, device);
    }
}

template <typename WT>
static int launch_m(int mode, int vpt, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    switch (mode) {
        case SPIF_MODE_FUSED:
            return launch_v<WT, SPIF_MODE_FUSED>(vpt, kp, grid, threads, smem, stream, device);
        case SPIF_MODE_GEMV:
            return launch_v<WT, SPIF_MODE_GEMV>(vpt, kp, grid, threads, smem, stream, device);
        default:
            return launch_v<WT, SPIF_MODE_AXPY>(vpt, kp, grid, threads, smem, stream, device);
    }
}

// workspace layout shared with spif_api.cu: [partials n_tok*P*n_embd f32][cta_stats n_tok*P*2 i32]
int spif_launch_ffn(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws, size_t ws_bytes,
                    const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(a.type == SPIF_TYPE_F16 || a.type == SPIF_TYPE_BF16, SPIF_E_TYPE);
    SPIF_REQUIRE(a.n_embd > 0 && a.n_embd % 8 == 0 && a.n_ff > 0 && a.m > 0 && a.n_tok > 0, SPIF_E_SHAPE);
    SPIF_REQUIRE(a.n_embd <= 16384, SPIF_E_SHAPE);
    SPIF_REQUIRE(a.neuron_idx != nullptr || a.m == a.n_ff, SPIF_E_SHAPE);

    // tuning knobs (read once): CTAs per SM and ring depth.  Defaults chosen from profiles/ (see DESIGN.md).
    static const int k_ctas_per_sm = [] { const char * e = getenv("SPIF_CTAS_PER_SM"); int v = e ? atoi(e) : 1; return v < 1 ? 1 : (v > 2 ? 2 : v); }();
    static const int k_max_stages  = [] { const char * e = getenv("SPIF_STAGES"); int v = e ? atoi(e) : 8; return v < 2 ? 2 : (v > 8 ? 8 : v); }();
    const int P = ctx.sm_count * k_ctas_per_sm;
    SPIF_REQUIRE(P >= 1 && P <= SPIF_RED_CHUNKS * SPIF_RED_PER, SPIF_E_UNSUPPORTED);
    FfnKParams kp = {};
    kp.w_up       = (const unsigned char *) a.w_up;
    kp.w_gate     = (const unsigned char *) a.w_gate;
    kp.w_down     = (const unsigned char *) a.w_down_t;
    kp.b_up       = a.b_up;
    kp.b_gate     = a.b_gate;
    kp.x          = a.x;
    kp.mask       = a.sparse_idx;
    kp.hid_in     = hid_in;
    kp.neuron_idx = a.neuron_idx;
    kp.n_embd     = a.n_embd;
    kp.n_ff       = a.n_ff;
    kp.m          = a.m;
    kp.mask_len   = (a.flags & SPIF_FLAG_MASK_BY_ROW) ? a.m : a.n_ff;
    kp.V          = a.n_embd / 8;
    kp.row_bytes  = a.n_embd * 2;
    kp.thr        = a.thr;
    kp.arch       = a.arch;
    kp.flags      = a.flags;
    kp.dense_out  = (mode == SPIF_MODE_GEMV) ? dst_dense : a.hidden_out;
    kp.trace      = g_spif_trace;
    kp.magic      = (unsigned int) ((1ull << 32) / (unsigned) P) + 1u;
    const bool dense_list = (a.flags & SPIF_INTERNAL_DENSE) != 0;
    SPIF_REQUIRE(dense_list || a.sparse_idx != nullptr, SPIF_E_NULL);
    kp.zero_fill  = (mode == SPIF_MODE_GEMV && a.neuron_idx == nullptr && !dense_list) ? 1 : 0;

    // thread shape: VPT 16-byte vectors per consumer thread
    const int V   = kp.V;
    const int vpt = (V <= 128) ? 1 : (V <= 768 ? 2 : 4);
    kp.nc         = ((V + vpt - 1) / vpt + 31) / 32 * 32;
    SPIF_REQUIRE(kp.nc <= 512, SPIF_E_SHAPE);
    const int threads = kp.nc + 32;

    const bool gated = (mode == SPIF_MODE_FUSED) && (a.arch != SPIF_ARCH_RELU_UP);
    const int  nmat  = gated ? 2 : 1;
    kp.list_cap      = (a.m + P - 1) / P + 1;
    kp.D             = (mode == SPIF_MODE_FUSED) ? 2 * SPIF_NB : 0;
    const size_t fixed =
        (size_t) 8 * (2 * 8 + kp.D) + sizeof(float) * 2 * 16 * SPIF_NB * 2 + sizeof(int) * (32 + 4) + (size_t) 12 * kp.list_cap + 256;
    const size_t budget = ((size_t) ctx.max_smem_optin - 1024) / k_ctas_per_sm - (k_ctas_per_sm > 1 ? 1024 : 0);
    const size_t dbytes = (size_t) kp.D * kp.row_bytes;
    SPIF_REQUIRE(budget > fixed + dbytes + 2 * (size_t) nmat * kp.row_bytes, SPIF_E_SHAPE);
    int S = (int) ((budget - fixed - dbytes) / ((size_t) nmat * kp.row_bytes));
    S     = S > k_max_stages ? k_max_stages : S;
    kp.S  = S;
    const size_t smem = (size_t) S * nmat * kp.row_bytes + dbytes + fixed;

    // workspace
    const size_t part_bytes = sizeof(float) * (size_t) a.n_tok * P * a.n_embd;
    const size_t stat_bytes = sizeof(int32_t) * (size_t) a.n_tok * P * 2;
    if (mode != SPIF_MODE_GEMV) {
        SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0 && ws_bytes >= part_bytes + stat_bytes, SPIF_E_WORKSPACE);
        kp.partials  = (float *) ws;
        kp.cta_stats = (int32_t *) ((char *) ws + part_bytes);
    }
    if (mode == SPIF_MODE_GEMV && !kp.zero_fill && !dense_list) {
        SPIF_CUDA_TRY(cudaMemsetAsync(dst_dense, 0, sizeof(float) * (size_t) a.n_ff * a.n_tok, ctx.stream));
    }
    if (mode == SPIF_MODE_FUSED && a.hidden_out) {
        SPIF_CUDA_TRY(cudaMemsetAsync(a.hidden_out, 0, sizeof(float) * (size_t) a.n_ff * a.n_tok, ctx.stream));
    }

    const dim3 grid(P, a.n_tok);
    int        rc = (a.type == SPIF_TYPE_F16) ? launch_m<spif_f16>(mode, vpt, kp, grid, threads, smem, ctx.stream, ctx.device) :
                                                launch_m<spif_bf16>(mode, vpt, kp, grid, threads, smem, ctx.stream, ctx.device);
    if (rc) {
        return rc;
    }
    if (mode != SPIF_MODE_GEMV) {
        float *            y  = (mode == SPIF_MODE_AXPY) ? dst_dense : a.y;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim            = dim3((a.n_embd / 4 + 31) / 32, a.n_tok);
        cfg.blockDim           = dim3(32, SPIF_RED_CHUNKS);
        cfg.stream             = ctx.stream;
        cudaLaunchAttribute at[1];
        at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs                                        = at;
        cfg.numAttrs                                     = 1;
        SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_reduce_partials_kernel, (const float *) kp.partials,
                                         (mode == SPIF_MODE_FUSED) ? a.b_down : (const float *) nullptr, y,
                                         (const int32_t *) kp.cta_stats, a.stats_out, a.n_embd, P));
    }
    return 0;
}
