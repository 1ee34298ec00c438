// spif_ffn_qwarp.cuh -- Q8_0 / Q4_0 sparse FFN at one token per CTA row: warp-per-neuron gate/up dots, CTA-wide column-owned
// down axpy.  Replaces the shared ring + "one block pair per consumer thread" schedule of round 1 (spif_ffn_kernel<QuantCodec>)
// which kept 80 of 224 threads busy with serial 64-element dots behind a CTA barrier per two neurons (35 us per 13B layer).
//
// Reference arithmetic: mul_mat_vec_sparse_q8_0_q8_1 + quantize_q8_1 (mmq-sparse.cu:12-163, quantize.cu:5-36) and
// mul_mat_axpy_sparse_rowwise_q (axpyq-sparse.cu:26-105); CPU: ggml_vec_dot_q8_0_q8_0 / q4_0_q8_0 (arch/x86/quants.c:543-583,
// 1011-1045), ggml_axpy_avx_q8_0_alphaf32 (ggml-cpu.c:2060-2146).
//
// A quantised 13B layer at 10 % activity is 9-18 MB (1.4-2.7 us of HBM time), 9 neurons per SM: the layer is a LATENCY chain
// (list -> gate/up rows -> hidden -> down row -> y), so the kernel is built to put every byte in flight at once and to cross
// as few synchronisation points as possible:
//   * slot s of shared memory = {up row, gate row, W_down^T row} of ONE neuron and belongs to warp s.  Lane 0 of the warp hands
//     the neuron's up and gate rows to the TMA unit (cp.async.bulk, completion on the slot's mbarrier) the moment the CTA's
//     active-row list exists; with <= NSLOT neurons per CTA (the normal case) every row of the CTA is requested in one go;
//   * x is quantised to 8 bit per 32-block ONCE per CTA into shared memory (both quantiser variants of the reference), laid
//     out so that a lane's 16-byte reads are conflict-free;
//   * the warp waits for ITS rows only, lanes take block pairs lane, lane+32, ... (36 / 68 bytes, 4-byte aligned, payload
//     realigned with funnel shifts), integer dp4a dots with d_w * d_x per block, one shuffle tree -> hidden value; if it is
//     non-zero (axpy-sparse.cu:53) the same lane requests the W_down^T row; no CTA barrier so far;
//   * one CTA barrier, then every thread owns 8 columns (IPT x) of y and adds the landed down rows in LIST ORDER:
//     acc = fma(q, d * alpha, acc) -- fixed order, bit-stable; per-CTA partial y -> spif_reduce_partials_kernel as before.
// More neurons than slots: rounds of NSLOT; a warp requests its next neuron's gate/up rows as soon as its dots are done.
#pragma once
#include "spif_ffn_kernel.cuh"

struct spif_q8_0 {};
struct spif_q4_0 {};

template <typename QT> struct QuantTraits;
template <> struct QuantTraits<spif_q8_0> {
    static constexpr int BLOCK_BYTES = 34;
    static constexpr int PAIR_WORDS  = 17;
};
template <> struct QuantTraits<spif_q4_0> {
    static constexpr int BLOCK_BYTES = 18;
    static constexpr int PAIR_WORDS  = 9;
};

__device__ __forceinline__ float half_bits_to_float(unsigned h) {
    return __half2float(__ushort_as_half((unsigned short) (h & 0xffffu)));
}

struct QwarpAlpha {  // build_my_list's codec hook: alpha stays F32 for quantised rows (ggml-cpu.c:2215-2218)
    __device__ static __forceinline__ float round_alpha(float h) { return h; }
};

// shared-memory layout, computed identically on host and device
struct QwarpSmem {
    uint32_t warp_cnt, counts, list_row, list_neu, list_alpha, hs, bars, xd, xs, xq, slots, total;
};
static inline __host__ __device__ QwarpSmem qwarp_smem_layout(int list_cap, int nslot, int n_embd, int slot_bytes) {
    QwarpSmem L;
    uint32_t  o   = 0;
    const int np  = n_embd / 64 + ((n_embd % 64) ? 1 : 0);
    L.warp_cnt    = o, o += 32 * 4;
    L.counts      = o, o += 4 * 4;
    L.list_row    = o, o += 4 * list_cap;
    L.list_neu    = o, o += 4 * list_cap;
    L.list_alpha  = o, o += 4 * list_cap;
    L.hs          = o, o += 4 * nslot;
    o             = (o + 7) & ~7u;
    L.bars        = o, o += 16 * nslot;  // {gate/up, down} mbarrier per slot
    L.xd          = o, o += 8 * np;      // float2 per pair: activation scale of its two blocks
    L.xs          = o, o += 8 * np;      // int2 per pair: sum of the 8-bit activations per block (Q4_0's -8 offset)
    o             = (o + 15) & ~15u;
    L.xq          = o, o += 64 * np;     // [4][np] uint4: quad k of pair p = 16 activations
    L.slots       = o, o += (uint32_t) nslot * slot_bytes;
    L.total       = o;
    return L;
}

template <typename QT, int MODE, int IPT, int MAXT>
__global__ void __launch_bounds__(MAXT, (MAXT <= 320 ? 2 : 1)) spif_ffn_qwarp_kernel(const FfnKParams p) {
    using T = QuantTraits<QT>;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NT = blockDim.x;
    const int P = gridDim.x, c = blockIdx.x, tok = blockIdx.y;
    const int NSLOT = p.S;
    const bool gated = (MODE == SPIF_MODE_FUSED) && (p.arch != SPIF_ARCH_RELU_UP);
    const int  rows_per_slot = (MODE == SPIF_MODE_FUSED) ? (gated ? 3 : 2) : 1;
    const int  slot_bytes    = rows_per_slot * p.row_bytes;
    const int  nblk = p.n_embd >> 5, np = (nblk + 1) >> 1;

    const QwarpSmem L = qwarp_smem_layout(p.list_cap, NSLOT, p.n_embd, slot_bytes);
    int *      warp_cnt   = reinterpret_cast<int *>(smem + L.warp_cnt);
    int *      s_counts   = reinterpret_cast<int *>(smem + L.counts);
    int *      list_row   = reinterpret_cast<int *>(smem + L.list_row);
    int *      list_neu   = reinterpret_cast<int *>(smem + L.list_neu);
    float *    list_alpha = reinterpret_cast<float *>(smem + L.list_alpha);
    float *    hs         = reinterpret_cast<float *>(smem + L.hs);
    uint64_t * bars       = reinterpret_cast<uint64_t *>(smem + L.bars);
    float2 *   xd         = reinterpret_cast<float2 *>(smem + L.xd);
    int2 *     xs         = reinterpret_cast<int2 *>(smem + L.xs);
    uint4 *    xq         = reinterpret_cast<uint4 *>(smem + L.xq);
    unsigned char * slots = smem + L.slots;

    const bool early = (p.flags & SPIF_FLAG_MASK_STABLE) != 0;
    SPIF_STAMP(0);
    if (tid == 0) {
        for (int s = 0; s < 2 * NSLOT; ++s) {
            mbar_init(bars + s, 1);
        }
        mbar_fence_init();
    }
    if (!early) {
        pdl_wait_prior_grid();
    }
    SPIF_STAMP(1);
    // this thread's part of x (8 lanes per 32-block, one float4 each; XQ = 4 covers every launch shape: threads >= n_embd / 16).
    // Without MASK_STABLE x is already valid here: request it BEFORE the mask scan, whose L2 round trips then cover its latency.
    constexpr int XQ = 4;
    float4        xv[XQ];
    auto load_x = [&]() {
        const int wi = lane & 7, bstep = NT >> 3;
#pragma unroll
        for (int u = 0; u < XQ; ++u) {
            const int b = (tid >> 3) + u * bstep;
            xv[u]       = (MODE != SPIF_MODE_AXPY && b < nblk) ?
                              __ldg(reinterpret_cast<const float4 *>(p.x + (size_t) blockIdx.y * p.n_embd + (size_t) b * 32) + wi) :
                              make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    if (!early) {
        load_x();
    }
    const float * mask  = p.mask + (size_t) tok * p.mask_len;
    const float * xg    = p.x + (size_t) tok * p.n_embd;
    const float * hid   = (MODE == SPIF_MODE_AXPY) ? p.hid_in + (size_t) tok * p.n_ff : nullptr;
    float *       dense = p.dense_out ? p.dense_out + (size_t) tok * p.n_ff : nullptr;
    build_my_list<MODE, QwarpAlpha, true>(p, mask, hid, dense, warp_cnt, list_row, list_neu, list_alpha, s_counts);  // syncs the CTA
    const int n = s_counts[0];
    SPIF_STAMP(2);

    const uint64_t pol = l2_policy_evict_first();
    auto slot_ptr = [&](int s) { return slots + (size_t) s * slot_bytes; };
    auto issue_gu = [&](int s, int i) {  // one lane: up (+ gate) rows of list entry i -> slot s
        const size_t off = (size_t) list_row[i] * p.row_bytes;
        mbar_arrive_expect_tx(bars + 2 * s, (uint32_t) ((gated ? 2 : 1) * p.row_bytes));
        tma_load_1d(slot_ptr(s), p.w_up + off, (uint32_t) p.row_bytes, bars + 2 * s, pol);
        if (gated) {
            tma_load_1d(slot_ptr(s) + p.row_bytes, p.w_gate + off, (uint32_t) p.row_bytes, bars + 2 * s, pol);
        }
    };
    auto issue_dn = [&](int s, int i) {  // one lane: W_down^T row of list entry i -> slot s
        const size_t off = (size_t) list_row[i] * p.row_bytes;
        mbar_arrive_expect_tx(bars + 2 * s + 1, (uint32_t) p.row_bytes);
        tma_load_1d(slot_ptr(s) + (size_t) (rows_per_slot - 1) * p.row_bytes, p.w_down + off, (uint32_t) p.row_bytes,
                    bars + 2 * s + 1, pol);
    };
    if (warp < NSLOT && warp < n && lane == 0) {
        if (MODE == SPIF_MODE_AXPY) {
            issue_dn(warp, warp);
        } else {
            issue_gu(warp, warp);
        }
    }
    if (early) {
        pdl_wait_prior_grid();  // x, y and the partials workspace belong to the previous kernels until here
        load_x();
    }
    pdl_launch_dependents();

    // ---- x -> 8 bit per 32-block, once per CTA: 8 lanes per block, one float4 each (quantize.cu:19-31 / quants.c:290-356).
    // All of a thread's loads are issued first (XQ = 4 covers every launch shape: threads >= n_embd / 16): the first cut did a
    // dependent L2 round trip per block group, 3 us of the layer's 10.
    if (MODE != SPIF_MODE_AXPY) {
        const bool    cpu = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
        const int     wi  = lane & 7, bstep = NT >> 3;
        {
            const int b0 = tid >> 3;
#pragma unroll
            for (int u = 0; u < XQ; ++u) {
                const int    b = b0 + u * bstep;  // (b < nblk) is uniform over the 8 lanes of a block and over whole warps' shuffles
                const float4 v = xv[u];
                float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) {
                    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                }
                const float d  = amax / 127.f;
                const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
                const float e4[4] = { v.x, v.y, v.z, v.w };
                int         sm = 0, w = 0;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    // default = the CUDA reference's roundf(x / d) (quantize.cu:29).  t = x * id is within 1.1e-5 of x / d for
                    // |x / d| <= 127, and round-to-nearest-even of t equals roundf(x / d) unless t is that close to a half-integer
                    // (where ties-away and a last-bit difference matter): only then divide and roundf.  Bit-identical to
                    // roundf(x / d); the IEEE division + roundf for every element was a third of the x stage.
                    int q;
                    if (cpu) {
                        q = __float2int_rn(e4[e] * id);
                    } else {
                        const float t = e4[e] * id;
                        q             = __float2int_rn(t);  // amax == 0: id == 0 -> t == 0 -> q == 0
                        if (fabsf(t - (float) q) > 0.5f - 1e-4f) {
                            q = (int) roundf(e4[e] / d);
                        }
                    }
                    sm += q;
                    w |= (q & 0xff) << (8 * e);
                }
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) {
                    sm += __shfl_xor_sync(0xffffffffu, sm, o);
                }
                if (b < nblk) {
                    const int pr = b >> 1, bb = b & 1;
                    reinterpret_cast<int *>(xq)[((size_t) (bb * 2 + (wi >> 2)) * np + pr) * 4 + (wi & 3)] = w;
                    if (wi == 0) {
                        reinterpret_cast<float *>(xd)[2 * pr + bb] = __half2float(__float2half_rn(d));
                        reinterpret_cast<int *>(xs)[2 * pr + bb]   = sm;
                    }
                }
            }
        }
        if ((nblk & 1) && tid < 8) {  // odd block count: the missing half of the last pair contributes nothing
            reinterpret_cast<int *>(xq)[((size_t) (2 + (tid >> 2)) * np + (np - 1)) * 4 + (tid & 3)] = 0;
            if (tid == 0) {
                reinterpret_cast<float *>(xd)[2 * (np - 1) + 1] = 0.f;
                reinterpret_cast<int *>(xs)[2 * (np - 1) + 1]   = 0;
            }
        }
    }
    __syncthreads();
    SPIF_STAMP(3);

    float acc[IPT][8];
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            acc[k][e] = 0.f;
        }
    }
    const int n_items   = p.n_embd >> 3;  // 8 columns each: (block, quarter)
    int       n_nonzero = 0;

    for (int base = 0, round = 0; base < n; base += NSLOT, ++round) {
        const uint32_t par = (uint32_t) round & 1u;
        const int      i   = base + warp;
        // ---- phase A: warp s owns list entry base + s
        if (warp < NSLOT) {
            if (MODE == SPIF_MODE_AXPY) {
                if (lane == 0) {
                    hs[warp] = (i < n) ? list_alpha[i] : 0.f;
                }
            } else if (i < n) {
                const unsigned char * up = slot_ptr(warp);
                const unsigned char * gt = up + p.row_bytes;
                mbar_wait_spin(bars + 2 * warp, par);
#ifdef SPIF_QW_TRACE_LANDED
                if (round == 0 && warp == 0) {
                    SPIF_STAMP(5);
                }
#endif
                float su = 0.f, sg = 0.f;
                for (int pr = lane; pr < np; pr += 32) {
                    const uint4  q0 = xq[pr], q1 = xq[np + pr], q2 = xq[2 * np + pr], q3 = xq[3 * np + pr];
                    const float2 dx = xd[pr];
                    const int2   sx = xs[pr];
                    const int    qx[16] = { (int) q0.x, (int) q0.y, (int) q0.z, (int) q0.w, (int) q1.x, (int) q1.y, (int) q1.z,
                                            (int) q1.w, (int) q2.x, (int) q2.y, (int) q2.z, (int) q2.w, (int) q3.x, (int) q3.y,
                                            (int) q3.z, (int) q3.w };
                    const bool two = (2 * pr + 1 < nblk);
#pragma unroll
                    for (int mtx = 0; mtx < 2; ++mtx) {
                        if (mtx == 1 && !gated) {
                            break;
                        }
                        const unsigned * wp = reinterpret_cast<const unsigned *>((mtx ? gt : up) + (size_t) pr * (2 * T::BLOCK_BYTES));
                        unsigned         w[T::PAIR_WORDS];
#pragma unroll
                        for (int k = 0; k < T::PAIR_WORDS; ++k) {
                            // the second block of an odd row's last pair lies outside the row: its activations are zero,
                            // the words only have to be readable (they are: the next row of the slot or the slot's tail pad)
                            w[k] = wp[k];
                        }
                        float out;
                        if constexpr (T::BLOCK_BYTES == 34) {
                            // [d0 | q0 x32 | d1 | q1 x32] = 68 B: q0 starts at byte 2 (funnel shift by 16), q1 at byte 36 (aligned)
                            int s0 = 0, s1 = 0;
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                s0 = __dp4a((int) __funnelshift_r(w[k], w[k + 1], 16), qx[k], s0);
                                s1 = __dp4a((int) w[9 + k], qx[8 + k], s1);
                            }
                            out = half_bits_to_float(w[0]) * dx.x * (float) s0;
                            if (two) {
                                out += half_bits_to_float(w[8] >> 16) * dx.y * (float) s1;
                            }
                        } else {
                            // [d0 | n0 x16 | d1 | n1 x16] = 36 B; byte j of a block = elements j (low nibble) and j+16 (high)
                            int s0 = 0, s1 = 0;
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const unsigned v0 = __funnelshift_r(w[k], w[k + 1], 16);
                                const unsigned v1 = w[5 + k];
                                s0 = __dp4a((int) (v0 & 0x0F0F0F0Fu), qx[k], s0);
                                s0 = __dp4a((int) ((v0 >> 4) & 0x0F0F0F0Fu), qx[4 + k], s0);
                                s1 = __dp4a((int) (v1 & 0x0F0F0F0Fu), qx[8 + k], s1);
                                s1 = __dp4a((int) ((v1 >> 4) & 0x0F0F0F0Fu), qx[12 + k], s1);
                            }
                            out = half_bits_to_float(w[0]) * dx.x * (float) (s0 - 8 * sx.x);
                            if (two) {
                                out += half_bits_to_float(w[4] >> 16) * dx.y * (float) (s1 - 8 * sx.y);
                            }
                        }
                        if (mtx) {
                            sg += out;
                        } else {
                            su += out;
                        }
                    }
                }
                su = warp_sum(su);
                sg = warp_sum(sg);
                const int neu = list_neu[i];
                if (p.b_up) {
                    su += __ldg(p.b_up + neu);
                }
                if (gated && p.b_gate) {
                    sg += __ldg(p.b_gate + neu);
                }
                if (MODE == SPIF_MODE_GEMV) {
                    if (lane == 0) {
                        dense[neu] = su;
                    }
                } else {
                    const float h = spif_activation(p.arch, sg, su);
                    if (lane == 0) {
                        if (dense) {
                            dense[neu] = h;
                        }
                        hs[warp] = h;
                        if (!(h == 0.0f)) {
                            issue_dn(warp, i);
                        } else {
                            mbar_arrive(bars + 2 * warp + 1);  // keep the slot's down barrier in step with the rounds
                        }
                    }
                }
                __syncwarp();
                if (lane == 0 && i + NSLOT < n) {
                    issue_gu(warp, i + NSLOT);  // the slot's gate/up rows are consumed: next round's go in flight now
                }
            } else if (lane == 0) {
                hs[warp] = 0.f;
            }
        }
        if (MODE == SPIF_MODE_GEMV) {
            continue;
        }
        __syncthreads();
        if (round == 0) {
            SPIF_STAMP(7);
        }
        // ---- phase B: every thread adds the landed W_down^T rows into its columns, in list order
        const int cnt = min(NSLOT, n - base);
        for (int s = 0; s < cnt; ++s) {
            const float alpha = hs[s];
            if (alpha == 0.0f) {
                continue;  // uniform over the CTA
            }
            ++n_nonzero;
            mbar_wait_spin(bars + 2 * s + 1, par);
            const unsigned char * dn = slot_ptr(s) + (size_t) (rows_per_slot - 1) * p.row_bytes;
#pragma unroll
            for (int k = 0; k < IPT; ++k) {
                const int it = tid + k * NT;
                if (it < n_items) {
                    const int              b  = it >> 2, q = it & 3;
                    const unsigned short * bp = reinterpret_cast<const unsigned short *>(dn + (size_t) b * T::BLOCK_BYTES);
                    const float            sc = half_bits_to_float(bp[0]) * alpha;  // scale = d * alpha (axpyq-sparse.cu:60)
                    if constexpr (T::BLOCK_BYTES == 34) {
                        const unsigned lo = (unsigned) bp[1 + 4 * q] | ((unsigned) bp[2 + 4 * q] << 16);
                        const unsigned hi = (unsigned) bp[3 + 4 * q] | ((unsigned) bp[4 + 4 * q] << 16);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            acc[k][e]     = fmaf((float) (int) (signed char) (lo >> (8 * e)), sc, acc[k][e]);
                            acc[k][4 + e] = fmaf((float) (int) (signed char) (hi >> (8 * e)), sc, acc[k][4 + e]);
                        }
                    } else {
                        const unsigned v = (unsigned) bp[1 + 2 * q] | ((unsigned) bp[2 + 2 * q] << 16);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int by  = (v >> (8 * e)) & 0xff;
                            acc[k][e]     = fmaf((float) ((by & 0xF) - 8), sc, acc[k][e]);
                            acc[k][4 + e] = fmaf((float) ((by >> 4) - 8), sc, acc[k][4 + e]);
                        }
                    }
                }
            }
        }
        if (base + NSLOT < n) {
            __syncthreads();  // hs[] and the down rows of this round are consumed
            if (MODE == SPIF_MODE_AXPY && warp < NSLOT && lane == 0 && i + NSLOT < n) {
                issue_dn(warp, i + NSLOT);
            }
        }
    }
    SPIF_STAMP(4);
    if (MODE != SPIF_MODE_GEMV) {
        float * part = p.partials + ((size_t) tok * P + c) * p.n_embd;
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const int it = tid + k * NT;
            if (it < n_items) {
                const int b = it >> 2, q = it & 3;
                if constexpr (T::BLOCK_BYTES == 34) {
                    float4 * o = reinterpret_cast<float4 *>(part + (size_t) b * 32 + 8 * q);
                    o[0]       = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                    o[1]       = make_float4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]);
                } else {
                    *reinterpret_cast<float4 *>(part + (size_t) b * 32 + 4 * q)      = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                    *reinterpret_cast<float4 *>(part + (size_t) b * 32 + 16 + 4 * q) = make_float4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]);
                }
            }
        }
        if (tid == 0 && p.cta_stats) {
            int32_t * st = p.cta_stats + ((size_t) tok * P + c) * 2;
            st[0]        = s_counts[1];
            st[1]        = n_nonzero;
        }
    }
#ifndef SPIF_QW_TRACE_LANDED
    SPIF_STAMP(5);
#endif
    if (p.trace && tid == 0 && blockIdx.y == 0) {
        p.trace[(size_t) blockIdx.x * 8 + 6] = (unsigned long long) n;
    }
}
