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
// mask -> this CTA's share of the active rows.  All threads of the CTA take part; every CTA scans the whole mask
// (n_ff floats, L2-resident) and keeps the rows whose rank r satisfies r % gridDim.x == blockIdx.x.
//
// The rank order only has to be the SAME in every CTA (it decides ownership and the fixed summation order), it does
// not have to be ascending in the row id.  Lane-major order (warp, lane, position in the lane's own rows) needs ONE
// warp scan over per-lane totals instead of one per 32 rows: each warp owns a contiguous chunk, a lane reads the
// chunk with coalesced 128-bit loads (float4 j of the lane = chunk + 32 j + lane) and packs its activity bits into
// registers (4 bits per load).  ~200 instructions per warp at n_ff = 13824 instead of ~1500 for the ballot-per-row form.
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
    const bool indirect = (p.neuron_idx != nullptr);
    const bool by_row   = (p.flags & SPIF_FLAG_MASK_BY_ROW) != 0;
    const bool rounding = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
    const bool vec_ok   = !indirect && (p.m % 4 == 0) && ((reinterpret_cast<uintptr_t>(mask) & 15) == 0) &&
                        (MODE != SPIF_MODE_AXPY || (reinterpret_cast<uintptr_t>(hid) & 15) == 0);
    auto alpha_of = [&](float h) -> float { return rounding ? round_through<WT>(h) : h; };

    unsigned long long bits[2] = { 0ull, 0ull };
    int                base    = 0;  // first float4 (vec) / first row (scalar) of my warp's chunk
    if (vec_ok) {
        const int      n4   = p.m >> 2;
        const int      q4   = (n4 + nwarp * 32 - 1) / (nwarp * 32);  // float4 per lane, <= 32 (checked by the launcher)
        const float4 * m4   = reinterpret_cast<const float4 *>(mask);
        const float4 * h4   = reinterpret_cast<const float4 *>(hid);
        const int      last = n4 - 1;
        base                = warp * q4 * 32;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int jn = min(16, q4 - 16 * h);
#pragma unroll 8
            for (int j = 0; j < jn; ++j) {
                const int    f  = base + (16 * h + j) * 32 + lane;
                const float4 v  = __ldg(m4 + min(f, last));
                unsigned     nb = (spif_is_active(v.x, p.thr) ? 1u : 0u) | (spif_is_active(v.y, p.thr) ? 2u : 0u) |
                              (spif_is_active(v.z, p.thr) ? 4u : 0u) | (spif_is_active(v.w, p.thr) ? 8u : 0u);
                if (MODE == SPIF_MODE_AXPY) {  // axpy-sparse.cu:53, ggml-cpu.c:2208: alpha == 0 rows are skipped
                    const float4 a = __ldg(h4 + min(f, last));
                    nb &= (alpha_of(a.x) == 0.0f ? 0u : 1u) | (alpha_of(a.y) == 0.0f ? 0u : 2u) |
                          (alpha_of(a.z) == 0.0f ? 0u : 4u) | (alpha_of(a.w) == 0.0f ? 0u : 8u);
                }
                nb = (f <= last) ? nb : 0u;
                bits[h] |= (unsigned long long) nb << (4 * j);
            }
        }
    } else {
        // gathered mask (GPU cache / shard rows -> neuron ids) or unaligned: scalar loads, one bit per load
        const int q    = (p.m + nwarp * 32 - 1) / (nwarp * 32);  // rows per lane, <= 128
        const int last = p.m - 1;
        base           = warp * q * 32;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int jn = min(64, q - 64 * h);
#pragma unroll 8
            for (int j = 0; j < jn; ++j) {
                const int   row = base + (64 * h + j) * 32 + lane;
                const int   rc  = min(row, last);
                const int   neu = indirect ? __ldg(p.neuron_idx + rc) : rc;
                const float v   = __ldg(mask + (by_row ? rc : neu));
                bool        a   = (row <= last) && spif_is_active(v, p.thr);
                if (MODE == SPIF_MODE_AXPY) {
                    a = a && !(alpha_of(__ldg(hid + neu)) == 0.0f);
                }
                bits[h] |= (unsigned long long) (a ? 1u : 0u) << j;
            }
        }
    }
    const int cnt  = __popcll(bits[0]) + __popcll(bits[1]);
    int       incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        incl += (lane >= o) ? t : 0;
    }
    if (lane == 31) {
        warp_cnt[warp] = incl;
    }
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < nwarp; ++w) {
        const int x = warp_cnt[w];
        before += (w < warp) ? x : 0;
        total += x;
    }
    int rank = before + incl - cnt;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        unsigned long long w = bits[h];
        while (w) {
            const int b = __ffsll((long long) w) - 1;
            w &= w - 1;
            const int k = (int) __umulhi((unsigned) rank, p.magic);  // rank / P without a divide
            if (rank - k * P == c) {
                const int row = vec_ok ? ((base + (16 * h + (b >> 2)) * 32 + lane) * 4 + (b & 3)) :
                                         (base + (64 * h + b) * 32 + lane);
                const int neu = indirect ? __ldg(p.neuron_idx + row) : row;
                list_row[k]   = row;
                list_neu[k]   = neu;
                if (MODE == SPIF_MODE_AXPY) {
                    list_alpha[k] = alpha_of(__ldg(hid + neu));
                }
            }
            ++rank;
        }
    }
    if (MODE == SPIF_MODE_GEMV && p.zero_fill) {
        // dst[neu] = 0 for inactive neurons (mm-sparse.cu:397): this CTA clears its contiguous slice of dst; active
        // entries are written by the CTA that owns the row, so the two never touch the same element
        const int lo = (int) ((long long) c * p.m / P), hi = (int) ((long long) (c + 1) * p.m / P);
        for (int r = lo + tid; r < hi; r += blockDim.x) {
            if (!spif_is_active(__ldg(mask + r), p.thr)) {
                dense[r] = 0.0f;
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

// ring cursor: slot index + phase parity, advanced without integer division
struct RingPos {
    int      slot;
    unsigned phase;
    __device__ __forceinline__ void advance(int depth) {
        if (++slot == depth) {
            slot = 0;
            phase ^= 1u;
        }
    }
};

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
    unsigned char * ring   = smem;
    unsigned char * dring  = ring + (size_t) p.S * stage_bytes;
    unsigned char * tail   = dring + (size_t) p.D * p.row_bytes;
    uint64_t *      full   = reinterpret_cast<uint64_t *>(tail);
    uint64_t *      empty  = full + p.S;
    uint64_t *      dfull  = empty + p.S;
    uint64_t *      dempty = dfull + p.D;
    float *         dalpha = reinterpret_cast<float *>(dempty + p.D);    // [D] hidden value of the row in each down slot
    float *         red    = dalpha + p.D;                               // [2][16 warps][NB][2]
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
            mbar_init(&dempty[d], nwc);
        }
        mbar_fence_init();
    }
    SPIF_STAMP(0);
    // everything above overlaps the previous kernel's tail under programmatic dependent launch; with
    // SPIF_FLAG_MASK_STABLE so do the mask scan and the first ring-fill of weight rows (only x has to wait)
    const bool early = (p.flags & SPIF_FLAG_MASK_STABLE) != 0 && MODE == SPIF_MODE_FUSED;  // the single ops write dst early
    if (!early) {
        pdl_wait_prior_grid();
    }
    SPIF_STAMP(1);

    const float * mask  = p.mask + (size_t) tok * p.mask_len;
    const float * xg    = p.x ? p.x + (size_t) tok * p.n_embd : nullptr;
    const float * hid   = p.hid_in ? p.hid_in + (size_t) tok * p.n_ff : nullptr;
    float *       dense = p.dense_out ? p.dense_out + (size_t) tok * p.n_ff : nullptr;

    build_my_list<MODE, WT>(p, mask, hid, dense, warp_cnt, list_row, list_neu, list_alpha, s_counts);  // ends with __syncthreads
    const int my_count = s_counts[0];
    SPIF_STAMP(2);

    if (warp == nwc) {
        // ================= producer warp: stream weight rows of my active neurons =================
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            RingPos        rp  = { 0, 1u };  // phase of the "empty" barrier to wait for (first lap: never waits)
            for (int i = 0; i < my_count; ++i) {
                if (i >= p.S) {
                    mbar_wait(&empty[rp.slot], rp.phase);
                }
                const size_t off = (size_t) list_row[i] * p.row_bytes;
                mbar_arrive_expect_tx(&full[rp.slot], stage_bytes);
                unsigned char * dst = ring + (size_t) rp.slot * stage_bytes;
                if (MODE == SPIF_MODE_AXPY) {
                    tma_load_1d(dst, p.w_down + off, p.row_bytes, &full[rp.slot], pol);
                } else {
                    tma_load_1d(dst, p.w_up + off, p.row_bytes, &full[rp.slot], pol);
                    if (gated) {
                        tma_load_1d(dst + p.row_bytes, p.w_gate + off, p.row_bytes, &full[rp.slot], pol);
                    }
                }
                rp.advance(p.S);
            }
        }
        return;
    }
    if (warp > nwc) {
        return;
    }
    if (early) {
        pdl_wait_prior_grid();  // x, y and the partials workspace belong to the previous kernels until here
    }
    pdl_launch_dependents();    // always after this CTA's wait: keeps "my start => my predecessor's predecessor is done"

    // ================= consumer warps =================
    const int ctid = tid;  // consumer thread id == tid (< nc)
    float     acc[VPT * 8];
#pragma unroll
    for (int j = 0; j < VPT * 8; ++j) {
        acc[j] = 0.f;
    }
    int     n_nonzero = 0;
    RingPos sp        = { 0, 0u };  // gate/up (or axpy) ring, consumer side

    if (MODE == SPIF_MODE_AXPY) {
        for (int i = 0; i < my_count; ++i) {
            mbar_wait(&full[sp.slot], sp.phase);
            axpy_row<WT, VPT>(reinterpret_cast<const uint4 *>(ring + (size_t) sp.slot * stage_bytes), list_alpha[i], acc,
                              ctid, nc, V);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&empty[sp.slot]);
            }
            sp.advance(p.S);
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
        // down-row queue (FUSED): a ring of D slots filled by TMA as soon as a neuron's hidden value is known and
        // drained lazily, always in issue order (-> fixed summation order).  The number of rows issued is identical in
        // every consumer thread; each warp drains at its own pace, slot reuse is gated by dempty (one arrival per warp).
        int     q_pending = 0;            // issued - drained (this warp's view)
        int     q_total   = 0;            // issued so far
        RingPos dq        = { 0, 0u };    // next slot to drain
        RingPos iq        = { 0, 1u };    // next slot to issue into (+ phase of dempty to wait for once q_total >= D)
        auto    drain_one = [&]() {
            axpy_row<WT, VPT>(reinterpret_cast<const uint4 *>(dring + (size_t) dq.slot * p.row_bytes), dalpha[dq.slot], acc,
                              ctid, nc, V);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&dempty[dq.slot]);
            }
            dq.advance(p.D);
            --q_pending;
        };

        int rbuf = 0;
        for (int i0 = 0; i0 < my_count; i0 += SPIF_NB) {
            const int nb = min(SPIF_NB, my_count - i0);
            float     pu[SPIF_NB], pg[SPIF_NB];
#pragma unroll
            for (int j = 0; j < SPIF_NB; ++j) {
                pu[j] = 0.f, pg[j] = 0.f;
                if (j < nb) {
                    mbar_wait(&full[sp.slot], sp.phase);
                    if (i0 == 0 && j == 0) {
                        SPIF_STAMP(3);
                    }
                    const uint4 * ur = reinterpret_cast<const uint4 *>(ring + (size_t) sp.slot * stage_bytes);
                    const uint4 * gr = reinterpret_cast<const uint4 *>(ring + (size_t) sp.slot * stage_bytes + p.row_bytes);
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
                        mbar_arrive(&empty[sp.slot]);  // stage free for the producer
                    }
                    sp.advance(p.S);
                }
            }
            // cross-warp reduction of the batch: warp partials -> smem -> every warp re-sums them, one value per lane
            float * rb = red + rbuf * (16 * SPIF_NB * 2);
            rbuf ^= 1;
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
                    rb[(j * 2 + 0) * 16 + warp] = pu[j];
                    rb[(j * 2 + 1) * 16 + warp] = pg[j];
                }
            }
            named_bar_sync(1, nc);
            // lane q < 2*NB sums column q of the [2*NB][16] table in warp order (fixed order), then broadcast
            float tot = 0.f;
            if (lane < 2 * SPIF_NB) {
                for (int w = 0; w < nwc; ++w) {
                    tot += rb[lane * 16 + w];
                }
            }
            float hcur[SPIF_NB];
            int   nq = 0;
#pragma unroll
            for (int j = 0; j < SPIF_NB; ++j) {
                float u = __shfl_sync(0xffffffffu, tot, 2 * j);
                float g = __shfl_sync(0xffffffffu, tot, 2 * j + 1);
                hcur[j] = 0.f;
                if (j < nb) {
                    const int neu = list_neu[i0 + j];
                    if (MODE == SPIF_MODE_GEMV) {
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
                        nq += !(h == 0.0f) ? 1 : 0;  // axpy-sparse.cu:53: rows with hid == 0 are never read
                    }
                }
            }
            if (MODE == SPIF_MODE_FUSED) {
                // make room: every warp drains (blocking) until this batch's rows fit -- the same rule in every warp, so
                // the issuing thread's wait on dempty below never depends on a warp that is parked at the next barrier
                while (q_pending + nq > p.D) {
                    mbar_wait(&dfull[dq.slot], dq.phase);
                    drain_one();
                }
#pragma unroll
                for (int j = 0; j < SPIF_NB; ++j) {
                    if (j < nb && !(hcur[j] == 0.0f)) {
                        if (ctid == 0) {
                            if (q_total >= p.D) {
                                mbar_wait(&dempty[iq.slot], iq.phase);
                            }
                            dalpha[iq.slot]  = hcur[j];
                            const size_t off = (size_t) list_row[i0 + j] * p.row_bytes;
                            mbar_arrive_expect_tx(&dfull[iq.slot], p.row_bytes);
                            tma_load_1d(dring + (size_t) iq.slot * p.row_bytes, p.w_down + off, p.row_bytes, &dfull[iq.slot], pol);
                        }
                        iq.advance(p.D);
                        ++q_total;
                        ++q_pending;
                    }
                }
                // opportunistic: rows that have already landed are folded in now instead of in the tail
                while (q_pending > 0) {
                    unsigned ok = (lane == 0 && mbar_test(&dfull[dq.slot], dq.phase)) ? 1u : 0u;
                    ok          = __shfl_sync(0xffffffffu, ok, 0);  // warp-uniform decision
                    if (!ok) {
                        break;
                    }
                    mbar_wait(&dfull[dq.slot], dq.phase);  // returns at once; every lane gets its own acquire on the data
                    drain_one();
                }
            }
        }
        SPIF_STAMP(4);
        if (MODE == SPIF_MODE_FUSED) {
            n_nonzero = q_total;
            while (q_pending > 0) {
                mbar_wait(&dfull[dq.slot], dq.phase);
                drain_one();
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
// y[t][col] = sum_p partials[t][p][col] (+ b_down[col]) in a fixed order.  Block = 8 column-vectors (128 B of every
// partial row) x 32 partial lanes: thread (cx, py) loads partials py, py+32, ... (<= SPIF_RED_PER, all in flight at
// once -> one L2 round trip), then a fixed-shape two-level sum through shared memory.  n_embd/32 CTAs (160 at 13B).
#define SPIF_RED_COLS  8
#define SPIF_RED_LANES 32
#define SPIF_RED_PER   10 /* ceil(P / 32) for P <= 320 CTAs */
__global__ void __launch_bounds__(SPIF_RED_COLS * SPIF_RED_LANES) spif_reduce_partials_kernel(
    const float * __restrict__ partials, const float * __restrict__ b_down, float * __restrict__ y,
    const int32_t * __restrict__ cta_stats, int32_t * __restrict__ stats_out, int n_embd, int P) {
    __shared__ float4 sm[SPIF_RED_LANES][SPIF_RED_COLS];
    __shared__ float4 sm2[4][SPIF_RED_COLS];
    const int         tok  = blockIdx.y;
    const int         cx   = threadIdx.x & (SPIF_RED_COLS - 1);
    const int         py   = threadIdx.x / SPIF_RED_COLS;
    const int         col4 = blockIdx.x * SPIF_RED_COLS + cx;
    const int         n4   = n_embd >> 2;
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (col4 < n4) {
        const float4 * base = reinterpret_cast<const float4 *>(partials + (size_t) tok * P * n_embd) + col4;
        float4         v[SPIF_RED_PER];
#pragma unroll
        for (int q = 0; q < SPIF_RED_PER; ++q) {
            const int pp = py + q * SPIF_RED_LANES;
            v[q]         = (pp < P) ? base[(size_t) pp * n4] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int q = 0; q < SPIF_RED_PER; ++q) {
            s.x += v[q].x, s.y += v[q].y, s.z += v[q].z, s.w += v[q].w;
        }
    }
    sm[py][cx] = s;
    __syncthreads();
    if (py < 4) {  // 4 x 8 threads: each sums 8 lanes in order
        float4 t = sm[py * 8][cx];
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            const float4 v = sm[py * 8 + k][cx];
            t.x += v.x, t.y += v.y, t.z += v.z, t.w += v.w;
        }
        sm2[py][cx] = t;
    }
    __syncthreads();
    if (py == 0 && col4 < n4) {
        float4 t = sm2[0][cx];
#pragma unroll
        for (int k = 1; k < 4; ++k) {
            const float4 v = sm2[k][cx];
            t.x += v.x, t.y += v.y, t.z += v.z, t.w += v.w;
        }
        if (b_down) {
            const float4 bb = reinterpret_cast<const float4 *>(b_down)[col4];
            t.x += bb.x, t.y += bb.y, t.z += bb.z, t.w += bb.w;
        }
        reinterpret_cast<float4 *>(y + (size_t) tok * n_embd)[col4] = t;
    }
    if (stats_out && blockIdx.x == 0 && threadIdx.x < 32) {
        int nz = 0;
        for (int q = threadIdx.x; q < P; q += 32) {
            nz += cta_stats[((size_t) tok * P + q) * 2 + 1];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nz += __shfl_xor_sync(0xffffffffu, nz, o);
        }
        if (threadIdx.x == 0) {
            stats_out[tok * 4 + 0] = cta_stats[((size_t) tok * P) * 2 + 0];
            stats_out[tok * 4 + 1] = nz;
            stats_out[tok * 4 + 2] = 0;
            stats_out[tok * 4 + 3] = 0;
        }
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
    SPIF_REQUIRE(P >= 1 && P <= SPIF_RED_LANES * SPIF_RED_PER, SPIF_E_UNSUPPORTED);
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
    SPIF_REQUIRE(a.sparse_idx != nullptr, SPIF_E_NULL);
    kp.zero_fill  = (mode == SPIF_MODE_GEMV && a.neuron_idx == nullptr && !(a.flags & SPIF_FLAG_MASK_BY_ROW)) ? 1 : 0;

    // thread shape: VPT 16-byte vectors per consumer thread
    const int V   = kp.V;
    const int vpt = (V <= 128) ? 1 : (V <= 768 ? 2 : 4);
    kp.nc         = ((V + vpt - 1) / vpt + 31) / 32 * 32;
    SPIF_REQUIRE(kp.nc <= 512, SPIF_E_SHAPE);
    const int threads = kp.nc + 32;

    const bool gated = (mode == SPIF_MODE_FUSED) && (a.arch != SPIF_ARCH_RELU_UP);
    const int  nmat  = gated ? 2 : 1;
    kp.list_cap      = (a.m + P - 1) / P + 1;
    static const int k_dslots = [] { const char * e = getenv("SPIF_DSLOTS"); int v = e ? atoi(e) : 0; return v < 0 ? 0 : (v > 16 ? 16 : v); }();
    const int    nwarp_all  = threads / 32;
    SPIF_REQUIRE(a.m <= 4096 * nwarp_all, SPIF_E_SHAPE);  // activity bits kept in registers by build_my_list
    const size_t stage_b    = (size_t) nmat * kp.row_bytes;
    const size_t fixed      = (size_t) 8 * (2 * 8 + 2 * 16) + 4 * 16 + sizeof(float) * 2 * 16 * SPIF_NB * 2 + sizeof(int) * (32 + 4) +
                         (size_t) 12 * kp.list_cap + 256;
    // Half of the SM's shared memory per CTA (SPIF_SMEM_SPLIT, default 2): the NEXT layer's CTA can then be co-resident
    // and fill its ring under this layer's tail (SPIF_FLAG_MASK_STABLE), and the register-only predictor kernel fits too.
    static const int k_split = [] { const char * e = getenv("SPIF_SMEM_SPLIT"); int v = e ? atoi(e) : 2; return v < 1 ? 1 : (v > 2 ? 2 : v); }();
    const int        n_share = k_ctas_per_sm > k_split ? k_ctas_per_sm : k_split;
    size_t           budget  = ((size_t) ctx.max_smem_optin - 1024) / n_share - (n_share > 1 ? 1024 : 0);
    if (budget < fixed + 2 * stage_b + (mode == SPIF_MODE_FUSED ? 2 * (size_t) kp.row_bytes : 0) + 1) {
        budget = (size_t) ctx.max_smem_optin - 1024;  // rows too long to share the SM: take all of it
    }
    SPIF_REQUIRE(budget > fixed + 2 * stage_b + (mode == SPIF_MODE_FUSED ? 2 * (size_t) kp.row_bytes : 0), SPIF_E_SHAPE);
    const size_t ring_budget = budget - fixed;
    int          S, D = 0;
    if (mode == SPIF_MODE_FUSED) {
        // roughly half of the ring for gate/up stages (enough bytes in flight to cover HBM latency), the rest for down
        // rows so that their fetch never has to wait for a slot at ~10% activity
        S = (int) ((ring_budget * 45 / 100) / stage_b);
        S = S < 2 ? 2 : (S > k_max_stages ? k_max_stages : S);
        D = (int) ((ring_budget - (size_t) S * stage_b) / kp.row_bytes);
        D = D > 16 ? 16 : D;
        if (k_dslots >= 2 && k_dslots < D) {
            D = k_dslots;
        }
        while (D < 2 && S > 2) {
            --S;
            D = (int) ((ring_budget - (size_t) S * stage_b) / kp.row_bytes);
        }
        SPIF_REQUIRE(D >= 2, SPIF_E_SHAPE);
    } else {
        S = (int) (ring_budget / stage_b);
        S = S > k_max_stages ? k_max_stages : S;
    }
    kp.S = S;
    kp.D = D;
    const size_t smem = (size_t) S * stage_b + (size_t) D * kp.row_bytes + fixed;

    // workspace
    const size_t part_bytes = sizeof(float) * (size_t) a.n_tok * P * a.n_embd;
    const size_t stat_bytes = sizeof(int32_t) * (size_t) a.n_tok * P * 2;
    if (mode != SPIF_MODE_GEMV) {
        SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0 && ws_bytes >= part_bytes + stat_bytes, SPIF_E_WORKSPACE);
        kp.partials  = (float *) ws;
        kp.cta_stats = (int32_t *) ((char *) ws + part_bytes);
    }
    if (mode == SPIF_MODE_GEMV && !kp.zero_fill) {
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
        cfg.gridDim            = dim3((a.n_embd / 4 + SPIF_RED_COLS - 1) / SPIF_RED_COLS, a.n_tok);
        cfg.blockDim           = dim3(SPIF_RED_COLS * SPIF_RED_LANES);
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
