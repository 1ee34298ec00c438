// spif_ffn_kernel.cuh -- the persistent sparse-FFN kernel template shared by the float (spif_ffn.cu) and quantised
// (spif_ffn_q.cu) weight formats.  See spif_ffn.cu for the design notes.
#pragma once
#include "spif_common.cuh"

#include <cstdlib>

#define SPIF_NB        2   /* neurons per reduction batch */
#define SPIF_MAX_WARPS 17  /* 16 consumer warps + 1 producer */

#define SPIF_KFLAG_PDL_LATE (1 << 30) /* kernel-internal: launch the dependent grid after the row loop */

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

extern unsigned long long * g_spif_trace;  // debug only (spif_ffn.cu); see tools/trace_ffn.py

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
// WIDE: all (up to 16) mask loads of a lane are issued before the first one is used -- one L2 round trip instead of one per
// unrolled group of 8 plus one per load of the remainder loop (2.6 us -> ~1 us at n_ff = 13824 with 10 warps); costs 64
// registers at this point of the kernel, so only kernels whose register budget is set elsewhere ask for it.
template <int MODE, typename C, bool WIDE = false>
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
    auto alpha_of = [&](float h) -> float { return rounding ? C::round_alpha(h) : h; };

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
            if (WIDE) {
                if (jn > 0) {
                    float4 v[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int f = base + (16 * h + j) * 32 + lane;
                        v[j]        = (j < jn) ? __ldg(m4 + min(f, last)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int f  = base + (16 * h + j) * 32 + lane;
                        unsigned  nb = (spif_is_active(v[j].x, p.thr) ? 1u : 0u) | (spif_is_active(v[j].y, p.thr) ? 2u : 0u) |
                                      (spif_is_active(v[j].z, p.thr) ? 4u : 0u) | (spif_is_active(v[j].w, p.thr) ? 8u : 0u);
                        if (MODE == SPIF_MODE_AXPY && j < jn) {
                            const float4 a = __ldg(h4 + min(f, last));
                            nb &= (alpha_of(a.x) == 0.0f ? 0u : 1u) | (alpha_of(a.y) == 0.0f ? 0u : 2u) |
                                  (alpha_of(a.z) == 0.0f ? 0u : 4u) | (alpha_of(a.w) == 0.0f ? 0u : 8u);
                        }
                        nb = (j < jn && f <= last) ? nb : 0u;
                        bits[h] |= (unsigned long long) nb << (4 * j);
                    }
                }
                continue;
            }
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


// ---------------------------------------------------------------------------------------------------------------
// Codec interface (what spif_ffn_kernel needs from a weight format):
//   NACC                    FP32 accumulators per consumer thread for the down axpy
//   XS                      the thread's slice of x in registers
//   load_x / dot / axpy / store / round_alpha
// FloatCodec<WT, VPT>: F16 / BF16 rows; consumer thread t owns the 16-byte vectors t, t + nc, ... (VPT of them).
template <typename WT, int VPT> struct FloatCodec {
    static constexpr int NACC = VPT * 8;
    struct XS {
        float xr[VPT * 8];
    };
    __device__ static __forceinline__ float round_alpha(float h) { return round_through<WT>(h); }
    __device__ static __forceinline__ void  load_x(XS & xs, const float * xg, int ctid, int nc, int n_embd, int flags) {
        const int V = n_embd >> 3;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xs.xr[k * 8 + j] = 0.f;
            }
            if (v < V) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v + 1);
                xs.xr[k * 8 + 0] = a.x, xs.xr[k * 8 + 1] = a.y, xs.xr[k * 8 + 2] = a.z, xs.xr[k * 8 + 3] = a.w;
                xs.xr[k * 8 + 4] = b.x, xs.xr[k * 8 + 5] = b.y, xs.xr[k * 8 + 6] = b.z, xs.xr[k * 8 + 7] = b.w;
                if (flags & SPIF_FLAG_CPU_ROUNDING) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        xs.xr[k * 8 + j] = round_through<WT>(xs.xr[k * 8 + j]);
                    }
                }
            }
        }
    }
    __device__ static __forceinline__ float dot(const unsigned char * row, const XS & xs, int ctid, int nc, int n_embd) {
        const int     V = n_embd >> 3;
        const uint4 * r = reinterpret_cast<const uint4 *>(row);
        float         s = 0.f;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
            if (v < V) {
                float w[8];
                unpack8<WT>(r[v], w);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    s = fmaf(w[e], xs.xr[k * 8 + e], s);
                }
            }
        }
        return s;
    }
    __device__ static __forceinline__ void axpy(const unsigned char * row, float alpha, float (&acc)[NACC], int ctid, int nc,
                                                int n_embd) {
        const int     V = n_embd >> 3;
        const uint4 * r = reinterpret_cast<const uint4 *>(row);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
            if (v < V) {
                float w[8];
                unpack8<WT>(r[v], w);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc[k * 8 + j] = fmaf(w[j], alpha, acc[k * 8 + j]);
                }
            }
        }
    }
    __device__ static __forceinline__ void store(float * part, const float (&acc)[NACC], int ctid, int nc, int n_embd) {
        const int V = n_embd >> 3;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = ctid + k * nc;
            if (v < V) {
                float4 * o = reinterpret_cast<float4 *>(part) + 2 * v;
                o[0]       = make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]);
                o[1]       = make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]);
            }
        }
    }
};

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

template <typename C, int MODE>
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

    build_my_list<MODE, C>(p, mask, hid, dense, warp_cnt, list_row, list_neu, list_alpha, s_counts);  // ends with __syncthreads
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
    float     acc[C::NACC];
#pragma unroll
    for (int j = 0; j < C::NACC; ++j) {
        acc[j] = 0.f;
    }
    int     n_nonzero = 0;
    RingPos sp        = { 0, 0u };  // gate/up (or axpy) ring, consumer side

    if (MODE == SPIF_MODE_AXPY) {
        for (int i = 0; i < my_count; ++i) {
            mbar_wait(&full[sp.slot], sp.phase);
            C::axpy(ring + (size_t) sp.slot * stage_bytes, list_alpha[i], acc, ctid, nc, p.n_embd);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&empty[sp.slot]);
            }
            sp.advance(p.S);
        }
        n_nonzero = my_count;
    } else {
        // this thread's slice of x (registers), converted as the codec needs it
        typename C::XS xs;
        C::load_x(xs, xg, ctid, nc, p.n_embd, p.flags);

        const uint64_t pol = l2_policy_evict_first();
        // down-row queue (FUSED): a ring of D slots filled by TMA as soon as a neuron's hidden value is known and
        // drained lazily, always in issue order (-> fixed summation order).  The number of rows issued is identical in
        // every consumer thread; each warp drains at its own pace, slot reuse is gated by dempty (one arrival per warp).
        int     q_pending = 0;            // issued - drained (this warp's view)
        int     q_total   = 0;            // issued so far
        RingPos dq        = { 0, 0u };    // next slot to drain
        RingPos iq        = { 0, 1u };    // next slot to issue into (+ phase of dempty to wait for once q_total >= D)
        auto    drain_one = [&]() {
            C::axpy(dring + (size_t) dq.slot * p.row_bytes, dalpha[dq.slot], acc, ctid, nc, p.n_embd);
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
                    const unsigned char * ur = ring + (size_t) sp.slot * stage_bytes;
                    pu[j]                    = C::dot(ur, xs, ctid, nc, p.n_embd);
                    if (gated) {
                        pg[j] = C::dot(ur + p.row_bytes, xs, ctid, nc, p.n_embd);
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
                            h = C::round_alpha(h);
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
        C::store(p.partials + ((size_t) tok * P + c) * p.n_embd, acc, ctid, nc, p.n_embd);
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
// launch plumbing shared by the float and the quantised instantiations
template <typename C, int MODE>
static int spif_ffn_launch_kernel(const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    auto                      kern = spif_ffn_kernel<C, MODE>;
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

// y = fixed-order sum of the per-CTA partials (+ b_down) [+ the shard exchange]: spif_reduce_partials_kernel (spif_ffn.cu)
int spif_launch_reduce_partials(const FfnKParams & kp, const float * b_down, float * y, int32_t * stats_out, int n_embd, int n_tok,
                                int P, cudaStream_t stream, const spif_shard_comm * comm);

typedef int (*spif_ffn_launch_fn)(int type, int mode, int variant, const FfnKParams & kp, dim3 grid, int threads, size_t smem,
                                  cudaStream_t stream, int device);
int spif_launch_ffn_common(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                           size_t ws_bytes, const spif_launch_ctx & ctx, int row_bytes, int nc, int variant,
                           spif_ffn_launch_fn launch);
