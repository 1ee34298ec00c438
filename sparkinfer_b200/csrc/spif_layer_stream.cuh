// spif_layer_stream.cuh -- the look-ahead layer kernel, version 5: a STATIC per-CTA schedule of register-direct row batches.
// Included by spif_layer.cu (shares LayerParams, layer_sum_partials and the launch code).
//
// What versions 2-4 measured (profiles/r2_layer_trace_*.txt, profiles/r2_ncu_layer_*.txt):
//   * TMA ring (v2/v3): ONE issuing thread hands the ring a 40 KB item per ~1.0 us whatever the item is (even items without
//     copies), i.e. one item in flight: 40 GB/s per SM at best, and three roles in series per item.
//   * cp.async staging (v4): 22 M warp instructions per launch (address selection per 16 bytes, polling): issue-bound, 33 us.
//   * tools/ubench/stream_rows.cu: plain 128-bit loads straight into registers, 8 rows in flight per thread, warps running
//     free of each other, reach 6.5 TB/s -- the best of all load paths, with no staging at all.
// So: no ring, no issuer, no dynamic queue.  The work of a CTA is a fixed sequence of ITEMS of 4 row slices per thread,
// two items in flight per thread (8 x 16 B, 80 KB per SM at 13B): the loads of item t+1 are issued before item t is used,
// so a short item (the last rows of a phase) costs its bytes, not a memory round trip:
//
//     Wpu rows -> [row list ready] -> gate/up rows -> first Wpd item -> [hidden values known] -> W_down^T rows of the
//     non-zero hidden values -> partial y stored -> remaining Wpd rows -> column-sliced sum of all CTAs' partial y
//
// Every consumer warp walks this sequence on its own (it owns the 16-byte column slices [32w, 32w+32) of every row): issue
// the 4 loads of the next item, multiply the current one, one packed shuffle tree for its 4 partial dots, 4 floats into a
// small ring in shared memory, one mbarrier arrival.  Warps never wait for each other inside an item type; the three points where the sequence
// depends on other warps / CTAs (row list, hidden values, the predictor's hidden vector) sit behind other streaming work.
// The epilogue warp retires the partial dots in the canonical order Wpu -> gate/up -> Wpd (bias, activation, sigmoid,
// compaction); helper warps do everything that waits on other CTAs (row list, h1 gather, partial-y publication).
// Determinism: fixed shuffle tree, fixed warp order, W_down^T rows in ascending neuron order, partials in CTA order.
#pragma once

#define SPIF_L3_DEPTH 4 /* items in flight per consumer thread (cp.async groups) */
#define SPIF_L3_K 16 /* depth of the partial-dot ring (reduce items between the consumer warps and the epilogue warp) */

struct Layer3Smem {
    uint32_t done, part, ctl, list_row, list_neu, list_bu, list_bg, tab_a, tab_b, tab_c, tab_d, bpu_s, bpd_s, lout_s, h1_s, blk_pre,
        red, wdesc, stage, total;
};

__host__ __device__ static inline Layer3Smem layer3_smem_layout(int cap_ffn, int cap_pd, int cap_pu, int r, int nblk, int nw) {
    Layer3Smem L;
    uint32_t   o = 0;
    auto       take = [&](uint32_t bytes, uint32_t align) {
        o               = (o + align - 1) / align * align;
        const uint32_t r = o;
        o += bytes;
        return r;
    };
    L.done       = take(8u * SPIF_L3_K, 8);
    L.part       = take(4u * 128 * SPIF_L3_K, 16);  // [K][4][32]
    L.ctl        = take(96u, 16);  // int[8] {pend_tail, -, b_end, total_active, items retired, ...}, 4 mbarriers at +32
    L.list_row   = take(4u * cap_ffn, 4);
    L.list_neu   = take(4u * cap_ffn, 4);
    L.list_bu    = take(4u * cap_ffn, 4);
    L.list_bg    = take(4u * cap_ffn, 4);
    L.tab_a      = take(16u * (cap_pu + 4), 16);       // Wpu rows
    L.tab_b      = take(16u * (2 * cap_ffn + 4), 16);  // gate/up rows
    L.tab_c      = take(16u * (cap_pd + 4), 16);       // Wpd virtual rows (at most one per Wpd row)
    L.tab_d      = take(16u * (cap_ffn + 4), 16);      // W_down^T rows
    L.bpu_s      = take(4u * cap_pu, 4);
    L.bpd_s      = take(4u * cap_pd, 4);
    L.lout_s     = take(4u * cap_pd, 4);
    L.h1_s       = take(4u * r, 16);
    L.blk_pre    = take(4u * (nblk + 1), 4);
    L.red        = take(16u * (2 * nw + 4) * 16, 16);  // scratch of the column-sliced sum
    L.wdesc      = take(8u * SPIF_L3_DEPTH * nw, 16);  // per warp: {type, table address} of its items in flight
    L.stage      = take(16u * SPIF_L3_DEPTH * 4 * nw * 32, 128);  // [DEPTH][4 vectors per thread][NC] x 16 B
    L.total      = (o + 127) / 128 * 128;
    return L;
}

// ---- exact f16 x f16 (bf16 x bf16) products accumulated in F32 by ONE instruction per element (fma.rn.f32.f16 -> FHFMA):
// with SPIF_FLAG_CPU_ROUNDING x / the hidden values are values of the weight type (ggml-cpu.c:1832-1856), so the 8
// conversions per 16 bytes of weights disappear and the results are bit-identical to convert + FFMA (both round once).
template <typename WT> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<spif_f16>(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<spif_bf16>(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}
template <typename WT> __device__ __forceinline__ void fhfma2(float & s0, float & s1, uint32_t w, uint32_t x);
template <> __device__ __forceinline__ void fhfma2<spif_f16>(float & s0, float & s1, uint32_t w, uint32_t x) {
    asm("{\n\t.reg .b16 a0, a1, b0, b1;\n\tmov.b32 {a0, a1}, %2;\n\tmov.b32 {b0, b1}, %3;\n\t"
        "fma.rn.f32.f16 %0, a0, b0, %0;\n\tfma.rn.f32.f16 %1, a1, b1, %1;\n\t}"
        : "+f"(s0), "+f"(s1) : "r"(w), "r"(x));
}
template <> __device__ __forceinline__ void fhfma2<spif_bf16>(float & s0, float & s1, uint32_t w, uint32_t x) {
    asm("{\n\t.reg .b16 a0, a1, b0, b1;\n\tmov.b32 {a0, a1}, %2;\n\tmov.b32 {b0, b1}, %3;\n\t"
        "fma.rn.f32.bf16 %0, a0, b0, %0;\n\tfma.rn.f32.bf16 %1, a1, b1, %1;\n\t}"
        : "+f"(s0), "+f"(s1) : "r"(w), "r"(x));
}
// same order of additions as dot8: even elements in one chain, odd elements in the other
template <typename WT> __device__ __forceinline__ float dot8h(const uint4 & wv, const uint32_t * xh) {
    float s0 = 0.f, s1 = 0.f;
    fhfma2<WT>(s0, s1, wv.x, xh[0]);
    fhfma2<WT>(s0, s1, wv.y, xh[1]);
    fhfma2<WT>(s0, s1, wv.z, xh[2]);
    fhfma2<WT>(s0, s1, wv.w, xh[3]);
    return s0 + s1;
}

// A row of the schedule: {global address lo, hi, valid bytes of the row (0: no row), alpha of a W_down^T row}.  Tables in shared
// memory, one per source: Wpu rows and Wpd "virtual rows" (RP adjacent Wpd rows = row_bytes) are filled at kernel entry, the
// gate/up table by the helper warp that fetches the active-row list, the W_down^T table by the epilogue warp as it retires
// the hidden values.  Every table is padded with empty rows to whole items.
enum { T_DOT = 0, T_DOTH = 1, T_AXPY = 2, T_FLUSH = 3, T_NONE = 4, T_END = 5 };

// Warp roles: [0, NW) consumers, NW = epilogue, NW + 1 .. NW + 2 = helpers (row list, h1 gather, partial-y publication).
// An item is RPI = 4 / VPT rows (4 x 16 bytes per consumer thread); a thread keeps two items in flight.
template <typename WT, int VPT, int MAXT, bool TRACE, bool RND>
__global__ void __launch_bounds__(MAXT, 1) spif_layer3_kernel(const __grid_constant__ LayerParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int RPI   = 4 / VPT;  // rows per item: 4 x 16 bytes per thread
    constexpr int DEPTH = SPIF_L3_DEPTH;
    constexpr int K   = SPIF_L3_K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NW = p.nw, NC = NW * 32;
    const int P = gridDim.x, c = blockIdx.x;

    const Layer3Smem L = layer3_smem_layout(p.cap_ffn, p.cap_pd, p.cap_pu, p.has_pred ? p.r : 0, p.lin_nblk, NW);
    uint64_t * done       = reinterpret_cast<uint64_t *>(smem + L.done);
    float *    part       = reinterpret_cast<float *>(smem + L.part);  // [K][4][32]
    int *      ctl        = reinterpret_cast<int *>(smem + L.ctl);
    uint64_t * h1bar      = reinterpret_cast<uint64_t *>(smem + L.ctl + 32);
    uint64_t * flushbar   = reinterpret_cast<uint64_t *>(smem + L.ctl + 40);
    uint64_t * listbar    = reinterpret_cast<uint64_t *>(smem + L.ctl + 48);
    int *      list_row   = reinterpret_cast<int *>(smem + L.list_row);
    int *      list_neu   = reinterpret_cast<int *>(smem + L.list_neu);
    float *    list_bu    = reinterpret_cast<float *>(smem + L.list_bu);
    float *    list_bg    = reinterpret_cast<float *>(smem + L.list_bg);
    uint4 *    tab_a      = reinterpret_cast<uint4 *>(smem + L.tab_a);
    uint4 *    tab_b      = reinterpret_cast<uint4 *>(smem + L.tab_b);
    uint4 *    tab_c      = reinterpret_cast<uint4 *>(smem + L.tab_c);
    uint4 *    tab_d      = reinterpret_cast<uint4 *>(smem + L.tab_d);
    float *    bpu_s      = reinterpret_cast<float *>(smem + L.bpu_s);
    float *    bpd_s      = reinterpret_cast<float *>(smem + L.bpd_s);
    int *      lout_s     = reinterpret_cast<int *>(smem + L.lout_s);
    float *    h1_s       = reinterpret_cast<float *>(smem + L.h1_s);
    int *      blk_pre    = reinterpret_cast<int *>(smem + L.blk_pre);
    float4 *   red        = reinterpret_cast<float4 *>(smem + L.red);

    if (tid == 0) {
        LAYER_STAMP(7);
        if (p.trace) {
            p.trace[(size_t) blockIdx.x * 512 + 15] = (unsigned long long) clock64();
        }
    }
    if (tid < K + 3) {  // one barrier per thread: 19 serial mbarrier.init cost 0.3 us of every launch
        if (tid < K) {
            mbar_init(&done[tid], NW);
        } else if (tid == K) {
            mbar_init(h1bar, 64);
        } else if (tid == K + 1) {
            mbar_init(flushbar, NW + 1);
        } else {
            mbar_init(listbar, 1);
        }
        mbar_fence_init();
    }
    if (tid >= 32 && tid < 40) {
        ctl[tid - 32] = 0;
    }

    const bool gated    = (p.arch != SPIF_ARCH_RELU_UP);
    const bool rounding = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
    const int  world    = p.comm.world;
    const int  h1_off   = p.n_embd;  // h1 words live behind the y words of a slot
    uint32_t   seq      = 0;

    // my shares of the dense predictor matrices
    int ra = 0, rb = 0, pa = 0, pb = 0;
    if (p.has_pred) {
        const int rw = p.r / world, R0 = p.comm.rank * rw;
        ra = R0 + share(c, rw, P);
        rb = R0 + share(c + 1, rw, P);
        pa = share(c, p.pm, P);
        pb = share(c + 1, p.pm, P);
    }
    const int npg   = gated ? RPI / 2 : RPI;                 // neurons per gate/up item
    const int n_wpu = (rb - ra + RPI - 1) / RPI;             // items of Wpu rows
    const int wpi   = RPI * p.RP;                            // Wpd rows per item (RP of them = one "virtual row" of row_bytes)
    const int n_wpd = (pb - pa + wpi - 1) / wpi;

    // ---- the two tables that are known at entry (weights only: the first Wpu items are issued before the PDL wait)
    const uint32_t row_bytes = (uint32_t) p.row_bytes, wrow_bytes = (uint32_t) p.r * 2;
    const uint32_t tabA_s = smem_u32(tab_a), tabB_s = smem_u32(tab_b), tabC_s = smem_u32(tab_c), tabD_s = smem_u32(tab_d);
    for (int i = tid; i < n_wpu * RPI; i += blockDim.x) {
        const unsigned long long a = (unsigned long long) (p.wpu + (size_t) (ra + i) * row_bytes);
        tab_a[i] = make_uint4((uint32_t) a, (uint32_t) (a >> 32), (ra + i < rb) ? row_bytes : 0u, 0u);
    }
    for (int i = tid; i < n_wpd * RPI; i += blockDim.x) {
        const int                row = pa + i * p.RP;  // first Wpd row of virtual row i
        const unsigned long long a   = (unsigned long long) (p.wpd + (size_t) row * wrow_bytes);
        tab_c[i] = make_uint4((uint32_t) a, (uint32_t) (a >> 32), (uint32_t) max(0, min(p.RP, pb - row)) * wrow_bytes, 0u);
    }
    __syncthreads();
    if (tid == 0) {
        LAYER_STAMP(0);
    }

    if (warp < NW) {
        // ================================================ consumer warps ================================================
        // vr: x while Wpu / gate/up items are consumed, then (x is dead) my 8 (16) values of the predictor's hidden vector
        // (RND: both are values of the weight type, kept as packed pairs in vh -- see fhfma2)
        float    vr[VPT * 8], acc[VPT * 8];
        uint32_t vh[VPT * 4];
        uint32_t toff[VPT];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            toff[k] = (uint32_t) (tid + k * NC) * 16u;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                vr[k * 8 + e] = 0.f, acc[k * 8 + e] = 0.f;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                vh[k * 4 + e] = 0u;
            }
        }
        // my 16 bytes of (slot, row j, vector k): consecutive threads are consecutive in shared memory (conflict-free both ways);
        // a thread only ever reads back what it copied itself, so the staging needs no barrier -- cp.async groups are the
        // whole pipeline: DEPTH items in flight per thread, completion in order (cp.async.wait_group)
        const uint32_t stage_s = smem_u32(smem + L.stage) + (uint32_t) tid * 16u;
        const uint32_t jstride = (uint32_t) NC * 16u, sstride = jstride * (RPI * VPT);
        const uint32_t desc_s  = smem_u32(smem + L.wdesc) + (uint32_t) warp * (DEPTH * 8u);  // {type | name << 8, table address}

        // ---- which item next?  Priority: Wpu rows (the predictor's hidden vector has the longest way to go: every CTA of every
        // shard) > W_down^T rows whose hidden values the epilogue warp has retired (4 table rows at a time, so the partial y is
        // complete -- and on its way to the other CTAs -- right after the last gate/up item) > gate/up rows > Wpd rows (dense,
        // the cover for everything that waits on other CTAs).  Nothing here blocks: when no table is ready the slot gets an
        // empty item and the items already in flight are consumed meanwhile.
        int      a_left = n_wpu, b_left = p.has_ffn ? -1 : 0, c_left = n_wpd, d_next = 0;
        bool     flushed = !p.has_ffn;
        uint32_t a_ptr = tabA_s, b_ptr = tabB_s, c_ptr = tabC_s;
        auto pick = [&](int & ty, uint32_t & pt) {
            if (a_left > 0) {
                ty = T_DOT | (1 << 8), pt = a_ptr;
                --a_left, a_ptr += RPI * 16u;
                return;
            }
            if (b_left < 0) {  // my share of the active-row list: a helper warp builds the gate/up table
                if (!mbar_test(listbar, 0)) {
                    return;
                }
                b_left = ctl[5];
            }
            if (!flushed) {
                const int final_rows = ld_acquire_cta_s32(ctl + 1);  // > 0: every gate/up item retired, rows incl. padding + 1
                const int rows       = final_rows > 0 ? final_rows - 1 : ld_acquire_cta_s32(ctl + 0);
                if ((d_next + 1) * RPI <= rows) {
                    ty = T_AXPY | (3 << 8), pt = tabD_s + (uint32_t) d_next * (RPI * 16u);
                    ++d_next;
                    return;
                }
                if (final_rows > 0) {  // the W_down^T rows are all issued: my slice of the partial y follows them
                    ty      = T_FLUSH | (5 << 8);
                    flushed = true;
                    return;
                }
            }
            if (b_left > 0) {
                ty = T_DOT | (2 << 8), pt = b_ptr;
                --b_left, b_ptr += RPI * 16u;
                return;
            }
            if (c_left > 0) {
                ty = T_DOTH | (4 << 8), pt = c_ptr;
                --c_left, c_ptr += RPI * 16u;
                return;
            }
            if (flushed) {
                ty = T_END | (6 << 8);
            }
        };

        int  qi = 0, qc = 0, kseq = 0;
        auto issue = [&](int slot) {
            int      ty = T_NONE | (7 << 8);
            uint32_t pt = 0;
            pick(ty, pt);
            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(desc_s + (uint32_t) slot * 8u), "r"(ty), "r"(pt) : "memory");
            if (TRACE && warp == 0 && (ty & 0xff) != T_END) {
                ITEM_STAMP(qi, 0, (unsigned long long) (ty >> 8));
                ITEM_STAMP(qi, 1, 0);
                ++qi;
            }
            if ((ty & 0xff) <= T_AXPY) {  // an item with rows: table entries {address, valid bytes, alpha}
                const uint32_t dst = stage_s + (uint32_t) slot * sstride;
#pragma unroll
                for (int j = 0; j < RPI; ++j) {
                    const uint4                 e   = lds128(pt + j * 16u);
                    const unsigned char * const src = reinterpret_cast<const unsigned char *>(((unsigned long long) e.y << 32) | e.x);
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {  // rows that are not there: 0 bytes are read (the address is a valid one)
                        cp_async_16(dst + (uint32_t) (j * VPT + k) * jstride, src + toff[k], toff[k] < e.z ? 16 : 0);
                    }
                }
            }
            cp_async_commit();  // also for items without rows: one group per item keeps wait_group in step
        };

        bool h1_ready = false;
        auto consume = [&](int slot) -> bool {  // false: the END item
            int      ty;
            uint32_t pt;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(ty), "=r"(pt) : "r"(desc_s + (uint32_t) slot * 8u) : "memory");
            ty &= 0xff;
            if (ty == T_END) {
                return false;
            }
            if (TRACE && warp == 0) {
                ITEM_STAMP(qc, 2, 0);
            }
            if (ty <= T_AXPY) {
                uint4 w[RPI][VPT];
#pragma unroll
                for (int j = 0; j < RPI; ++j) {
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        w[j][k] = lds128(stage_s + (uint32_t) slot * sstride + (uint32_t) (j * VPT + k) * jstride);
                    }
                }
                if (ty == T_AXPY) {
                    // table order = ascending neuron order: a fixed FMA chain per element; missing rows are zeros times zero
#pragma unroll
                    for (int j = 0; j < RPI; ++j) {
                        float alpha;
                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(alpha) : "r"(pt + j * 16u + 12u));
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            if (RND) {  // alpha is a value of the weight type (the epilogue warp rounded it)
                                const uint32_t a2 = pack2<WT>(alpha, alpha);
                                fhfma2<WT>(acc[k * 8 + 0], acc[k * 8 + 1], w[j][k].x, a2);
                                fhfma2<WT>(acc[k * 8 + 2], acc[k * 8 + 3], w[j][k].y, a2);
                                fhfma2<WT>(acc[k * 8 + 4], acc[k * 8 + 5], w[j][k].z, a2);
                                fhfma2<WT>(acc[k * 8 + 6], acc[k * 8 + 7], w[j][k].w, a2);
                            } else {
                                float f[8];
                                unpack8<WT>(w[j][k], f);
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    acc[k * 8 + e] = fmaf(f[e], alpha, acc[k * 8 + e]);
                                }
                            }
                        }
                    }
                } else {
                    if (ty == T_DOTH && !h1_ready) {
                        // the hidden vector of the predictor, gathered by the helper warps while the rows above streamed
                        h1_ready = true;
                        mbar_wait(h1bar, 0);
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            const int    v    = tid + k * NC;
                            const int    hcol = (v < p.RP * p.Vr) ? v % p.Vr : 0;
                            const float4 h0   = *reinterpret_cast<const float4 *>(h1_s + hcol * 8);
                            const float4 h1v  = *reinterpret_cast<const float4 *>(h1_s + hcol * 8 + 4);
                            if (RND) {
                                vh[k * 4 + 0] = pack2<WT>(h0.x, h0.y), vh[k * 4 + 1] = pack2<WT>(h0.z, h0.w);
                                vh[k * 4 + 2] = pack2<WT>(h1v.x, h1v.y), vh[k * 4 + 3] = pack2<WT>(h1v.z, h1v.w);
                                continue;
                            }
                            vr[k * 8 + 0] = h0.x, vr[k * 8 + 1] = h0.y, vr[k * 8 + 2] = h0.z, vr[k * 8 + 3] = h0.w;
                            vr[k * 8 + 4] = h1v.x, vr[k * 8 + 5] = h1v.y, vr[k * 8 + 6] = h1v.z, vr[k * 8 + 7] = h1v.w;
                            if (p.pflags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    vr[k * 8 + e] = round_through<WT>(vr[k * 8 + e]);
                                }
                            }
                        }
                    }
                    // RPI x VPT = 4 partial dots -> one packed shuffle tree -> slot kseq % K of the ring -> one arrival
                    float v0, v1, v2, v3;
                    auto dotv = [&](const uint4 & wv, int k) -> float {
                        return RND ? dot8h<WT>(wv, vh + k * 4) : dot8<WT>(wv, vr + k * 8);
                    };
                    if (VPT == 1) {
                        v0 = dotv(w[0][0], 0), v1 = dotv(w[1 % RPI][0], 0);
                        v2 = dotv(w[2 % RPI][0], 0), v3 = dotv(w[3 % RPI][0], 0);
                    } else {
                        v0 = dotv(w[0][0], 0), v1 = dotv(w[0][VPT - 1], VPT - 1);
                        v2 = dotv(w[1 % RPI][0], 0), v3 = dotv(w[1 % RPI][VPT - 1], VPT - 1);
                    }
                    const float s = reduce4(v0, v1, v2, v3, lane);  // lanes [8q, 8q+8): partial q = row * VPT + vector
                    if (kseq >= K) {  // the epilogue warp must be done with item kseq - K (it trails by one or two items)
                        while (kseq - ld_volatile_s32(ctl + 4) >= K) {
                            __nanosleep(20);
                        }
                    }
                    const int ks = kseq & (K - 1);
                    if ((lane & 7) == 0) {
                        part[ks * 128 + (lane >> 3) * 32 + warp] = s;
                    }
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&done[ks]);
                    }
                    ++kseq;
                }
            } else if (ty == T_FLUSH) {
                // my slice of this CTA's partial y; a helper thread publishes it (fence + counter) off this path
                float * prt = p.partials + (size_t) c * p.n_embd;
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    if (toff[k] < row_bytes) {
                        float4 * o = reinterpret_cast<float4 *>(prt) + 2 * (tid + k * NC);
                        __stcg(o, make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]));
                        __stcg(o + 1, make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]));
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(flushbar);
                }
            } else {
                __nanosleep(100);  // an empty item: a table is not ready yet
            }
            if (TRACE && warp == 0) {
                ITEM_STAMP(qc, 3, 0);
                ++qc;
            }
            return true;
        };

        // ---- DEPTH - 1 items are in flight before the first one is used.  Wpu items are weights only: they are on their way
        // before the programmatic-dependent-launch wait; x is read after it.
        int n_issued = 0;
        while (n_issued < DEPTH - 1 && n_issued < n_wpu) {
            issue(n_issued);
            ++n_issued;
        }
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        if (warp == 0) {
            LAYER_STAMP(1);
        }
        seq = spif_comm_call_seq(p.comm);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            if (toff[k] < row_bytes) {
                const int    v = tid + k * NC;
                const float4 a = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v + 1);
                if (RND) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
                    vh[k * 4 + 0] = pack2<WT>(a.x, a.y), vh[k * 4 + 1] = pack2<WT>(a.z, a.w);
                    vh[k * 4 + 2] = pack2<WT>(b.x, b.y), vh[k * 4 + 3] = pack2<WT>(b.z, b.w);
                    continue;
                }
                vr[k * 8 + 0] = a.x, vr[k * 8 + 1] = a.y, vr[k * 8 + 2] = a.z, vr[k * 8 + 3] = a.w;
                vr[k * 8 + 4] = b.x, vr[k * 8 + 5] = b.y, vr[k * 8 + 6] = b.z, vr[k * 8 + 7] = b.w;
                if (rounding) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        vr[k * 8 + e] = round_through<WT>(vr[k * 8 + e]);
                    }
                }
            }
        }
        if (warp == 0) {
            LAYER_STAMP(8);
        }
        while (n_issued < DEPTH - 1) {
            issue(n_issued);
            ++n_issued;
        }
        for (int slot = 0, islot = DEPTH - 1;;) {
            issue(islot);
            cp_async_wait<DEPTH - 1>();  // the oldest item in flight has landed
            if (!consume(slot)) {
                break;
            }
            slot  = (slot + 1 == DEPTH) ? 0 : slot + 1;
            islot = (islot + 1 == DEPTH) ? 0 : islot + 1;
        }
        cp_async_wait<0>();
        if (warp == 0) {
            LAYER_STAMP(10);
        }
        // ---- y[columns of this CTA] = sum over all CTAs of their partial y, fixed order (+ b_down) [+ sum over the shards]
        if (p.has_ffn) {
            layer_sum_partials<true>(p, tid, NC, 1, red, seq, 2 * NW);
        }
    } else if (warp == NW) {
        // ================================================== epilogue warp ==================================================
        for (int i = lane; i < rb - ra; i += 32) {
            bpu_s[i] = p.bpu ? __ldg(p.bpu + ra + i) : 0.f;
        }
        for (int i = lane; i < pb - pa; i += 32) {
            bpd_s[i] = p.bpd ? __ldg(p.bpd + pa + i) : 0.f;
        }
        __syncwarp();
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq                = spif_comm_call_seq(p.comm);
        const uint32_t par = seq & 1u;
        if (p.has_pred && rb == ra && lane == 0) {
            atomicAdd(p.comm.seq + 4, 1u);  // no Wpu rows here: nothing to wait for from this CTA
        }
        constexpr int LPR = 32 / RPI;  // lanes per row sum
        const int     erow = lane / LPR, esub = lane % LPR;
        int  pend_tail = 0, n_nonzero = 0, cnt_out = 0, n_gu = 0, b_end = 0;
        bool have_list = !p.has_ffn;
        int  total = n_wpu + n_wpd;  // + n_gu once the list is known
        for (int k = 0;; ++k) {
            if (!have_list && k >= n_wpu) {  // the gate/up items follow the Wpu items; their number comes with the row list
                mbar_wait(listbar, 0);
                have_list = true;
                b_end     = ctl[2];
                n_gu      = (b_end + npg - 1) / npg;
                total += n_gu;
                if (n_gu == 0 && lane == 0) {
                    p.cta_stats[c * 2 + 0] = ctl[3];
                    p.cta_stats[c * 2 + 1] = 0;
                    st_release_cta_s32(ctl + 1, 1);
                    mbar_arrive(flushbar);
                }
            }
            if (k >= total) {
                break;
            }
            mbar_wait_spin(&done[k % K], (unsigned) (k / K) & 1u);
            const float * ps = part + (k % K) * 128;
            if (k < n_wpu + n_gu) {
                // row sum = sum over (vector k', warp w) of the partial dots, in a fixed order: LPR lanes per row, then a tree
                float s = 0.f;
                for (int e = esub; e < VPT * NW; e += LPR) {
                    const int kk = e / NW, ww = e - kk * NW;
                    s += ps[(erow * VPT + kk) * 32 + ww];
                }
#pragma unroll
                for (int o = LPR / 2; o > 0; o >>= 1) {
                    s += __shfl_xor_sync(0xffffffffu, s, o);
                }
                if (k < n_wpu) {
                    // h1[row] = relu(Wpu[row] . x + bpu[row])  (llama-graph.cpp:871-880), published to every shard
                    const int a = ra + k * RPI, n = min(RPI, rb - a);
                    if (esub == 0 && erow < n) {
                        const int row = a + erow;
                        float     z   = s + bpu_s[row - ra];
                        z             = (z > 0.f) ? z : 0.f;
                        const size_t off = ((size_t) par * world + p.comm.rank) * p.comm.cap + h1_off + row;
                        for (int r = 0; r < world; ++r) {
                            st_word_sys(p.comm.data[r] + off, __float_as_uint(z), seq);
                        }
                    }
                    if (a + n >= rb) {  // hint for the pollers (not a release: every word carries its own sequence number)
                        __syncwarp();
                        if (lane == 0) {
                            atomicAdd(p.comm.seq + 4, 1u);
                        }
                        LAYER_STAMP(13);
                    }
                } else {
                    // gated: rows (up_0, gate_0, up_1, gate_1, ...); ungated: (up_0 .. up_{RPI-1})
                    const int i0 = (k - n_wpu) * npg, n = min(npg, b_end - i0);
                    for (int j = 0; j < n; ++j) {
                        const int   i = i0 + j;
                        const float u = __shfl_sync(0xffffffffu, s, (gated ? 2 * j : j) * LPR) + list_bu[i];
                        const float g = gated ? __shfl_sync(0xffffffffu, s, (2 * j + 1) * LPR) + list_bg[i] : 0.f;
                        float       h = spif_activation(p.arch, g, u);
                        if (lane == 0 && p.hidden_out) {
                            p.hidden_out[list_neu[i]] = h;
                        }
                        if (rounding) {
                            h = round_through<WT>(h);
                        }
                        if (!(h == 0.0f)) {  // axpy-sparse.cu:53: rows with hid == 0 are never read
                            if (lane == 0) {
                                const unsigned long long a = (unsigned long long) (p.w_down + (size_t) list_row[i] * row_bytes);
                                tab_d[pend_tail] = make_uint4((uint32_t) a, (uint32_t) (a >> 32), row_bytes, __float_as_uint(h));
                            }
                            ++pend_tail;
                            ++n_nonzero;
                        }
                    }
                    if (lane == 0) {
                        if (k - n_wpu + 1 < n_gu) {
                            st_release_cta_s32(ctl + 0, pend_tail);  // after the table rows
                        } else {  // the W_down^T table is final: pad it to whole items
                            const int                n_dn = (pend_tail + RPI - 1) / RPI;
                            const unsigned long long a0   = (unsigned long long) p.w_down;  // a valid address, no bytes, alpha 0
                            for (int q = pend_tail; q < n_dn * RPI; ++q) {
                                tab_d[q] = make_uint4((uint32_t) a0, (uint32_t) (a0 >> 32), 0u, 0u);
                            }
                            p.cta_stats[c * 2 + 0] = ctl[3];
                            p.cta_stats[c * 2 + 1] = n_nonzero;
                            st_release_cta_s32(ctl + 1, n_dn * RPI + 1);
                            mbar_arrive(flushbar);
                        }
                    }
                }
            } else {
                const int row0 = pa + (k - n_wpu - n_gu) * wpi, n = min(wpi, pb - row0);
                for (int base = 0; base < n; base += 32) {
                    const int  idx   = base + lane;
                    const bool valid = idx < n;
                    float      z     = 0.f;
                    if (valid) {
                        // Wpd row idx of the item lives in virtual row j = idx / RP; its 32-vector groups are g0 .. g0 + WR - 1
                        // of that virtual row; group g was reduced by consumer warp g % NW for its vector kv = g / NW
                        const int j = idx / p.RP, g0 = (idx - j * p.RP) * p.WR;
                        for (int t = 0; t < p.WR; ++t) {  // fixed order
                            const int g = g0 + t, kv = g / NW, ww = g - kv * NW;
                            z += ps[(j * VPT + kv) * 32 + ww];
                        }
                        z += bpd_s[row0 - pa + idx];
                    }
                    const float pr = 1.f / (1.f + expf(-z));  // unary-ops.cpp:31-33 / unary.cu:48-50
                    if (valid) {
                        if (p.p_out) {
                            p.p_out[row0 + idx] = pr;
                        }
                        if (p.logits_out) {
                            p.logits_out[row0 + idx] = z;
                        }
                    }
                    const bool     act  = valid && spif_is_active(pr, p.thr_next);
                    const unsigned ball = __ballot_sync(0xffffffffu, act);
                    if (act) {
                        lout_s[cnt_out + __popc(ball & ((1u << lane) - 1u))] = row0 + idx;
                    }
                    cnt_out += __popc(ball);
                }
            }
            __syncwarp();
            if (lane == 0) {
                st_volatile_s32(ctl + 4, k + 1);  // slot k % K of the partial ring is free again
            }
        }
        // ---- the next layer's active-row list: this block's rows in ascending order + its count
        if (p.has_pred && p.lout_cnt) {
            __syncwarp();
            for (int i = lane; i < cnt_out; i += 32) {
                p.lout_rows[pa + i] = lout_s[i];
            }
            if (lane == 0) {
                p.lout_cnt[c] = cnt_out;
                if (c == 0) {
                    p.lout_hdr[0] = P;
                    p.lout_hdr[1] = p.pm;
                }
            }
        }
        LAYER_STAMP(6);
    } else {
        // ================================================== helper warps ==================================================
        const int ht = tid - (NW + 1) * 32;  // 0 .. 63
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq = spif_comm_call_seq(p.comm);
        // ---- my share of this layer's active rows: ranks [c*A/P, (c+1)*A/P) of the concatenated block lists
        if (p.has_ffn && ht >= 32) {
            // lane L owns blocks [L*kb, (L+1)*kb): all counts are requested at once (one L2 round trip), one warp scan, the
            // exclusive prefix of every block goes to shared memory; then one lane per list entry finds its block by binary
            // search and fetches the row (second round trip, all entries in parallel)
            const int nb = p.lin_nblk, kb = (nb + 31) / 32;  // kb <= 8 (checked by the launcher: nblk <= 256)
            const int b0 = min(lane * kb, nb), b1 = min(b0 + kb, nb);
            int       cb[8];
            int       local = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                cb[j] = (b0 + j < b1) ? __ldcg(p.lin_cnt + b0 + j) : 0;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                local += cb[j];
            }
            int incl = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                incl += (lane >= o) ? t : 0;
            }
            const int total_active = __shfl_sync(0xffffffffu, incl, 31);
            int       pbk          = incl - local;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (b0 + j < b1) {
                    blk_pre[b0 + j] = pbk;
                }
                pbk += cb[j];
            }
            if (lane == 31) {
                blk_pre[nb] = total_active;
            }
            __syncwarp();
            const int lo = share(c, total_active, P), hi = share(c + 1, total_active, P);
            const int b_end = hi - lo;
            for (int j = lane; j < b_end; j += 32) {
                const int rank = lo + j;
                int       bl = 0, bh = nb;  // blk_pre[bl] <= rank < blk_pre[bh]
                while (bh - bl > 1) {
                    const int mid = (bl + bh) >> 1;
                    if (blk_pre[mid] <= rank) {
                        bl = mid;
                    } else {
                        bh = mid;
                    }
                }
                const int row = __ldcg(p.lin_rows + share(bl, p.m, nb) + (rank - blk_pre[bl]));
                const int neu = p.neuron_idx ? __ldg(p.neuron_idx + row) : row;
                list_row[j]   = row;
                list_neu[j]   = neu;
                list_bu[j]    = p.b_up ? __ldg(p.b_up + neu) : 0.f;
                list_bg[j]    = p.b_gate ? __ldg(p.b_gate + neu) : 0.f;
                // the rows of this neuron in the gate/up table: gated (up, gate), else (up)
                const unsigned long long au = (unsigned long long) (p.w_up + (size_t) row * row_bytes);
                if (gated) {
                    const unsigned long long ag = (unsigned long long) (p.w_gate + (size_t) row * row_bytes);
                    tab_b[2 * j]     = make_uint4((uint32_t) au, (uint32_t) (au >> 32), row_bytes, 0u);
                    tab_b[2 * j + 1] = make_uint4((uint32_t) ag, (uint32_t) (ag >> 32), row_bytes, 0u);
                } else {
                    tab_b[j] = make_uint4((uint32_t) au, (uint32_t) (au >> 32), row_bytes, 0u);
                }
            }
            const int n_rows_b = gated ? 2 * b_end : b_end, n_gu_items = (n_rows_b + RPI - 1) / RPI;
            const unsigned long long a0 = (unsigned long long) p.w_up;  // a valid address, no bytes
            for (int q = n_rows_b + lane; q < n_gu_items * RPI; q += 32) {
                tab_b[q] = make_uint4((uint32_t) a0, (uint32_t) (a0 >> 32), 0u, 0u);
            }
            __syncwarp();
            if (lane == 0) {
                ctl[2] = b_end;
                ctl[5] = n_gu_items;
                ctl[3] = total_active;
                mbar_arrive(listbar);  // release: test_wait / try_wait on the other side acquire the list
            }
            LAYER_STAMP(2);
        }
        if (p.has_pred) {
            // the hidden vector of the predictor: {value, seq} words, complete once every CTA (of every shard) has retired its
            // Wpu rows.  One word is polled until every CTA of THIS shard has sent its rows (148 pollers x 8 KB of words per round
            // used a quarter of the L2's request rate: 2.7 us per layer); the words themselves are then read, normally once.
            const uint32_t par = seq & 1u;
            const int      rw  = p.r / world;  // rows per owner rank
            unsigned       spins = 0;
            while (ld_u32_relaxed_gpu(p.comm.seq + 4) < (unsigned) P && ++spins < SPIF_COMM_SPIN_LIMIT) {
                __nanosleep(100);
            }
            if (ht < 32) {
                LAYER_STAMP(14);
            }
            for (int row = ht * 16; row < p.r; row += 1024) {  // 16 words per thread per pass (r % 256 == 0: whole threads)
                const int     owner = row / rw;
                const uint2 * src   = p.comm.data[p.comm.rank] + ((size_t) par * world + owner) * p.comm.cap + h1_off + row;
                uint2         w16[16];
                bool          ok = false;
                while (true) {
#pragma unroll
                    for (int e = 0; e < 16; e += 2) {
                        ld_word2_sys(src + e, w16[e], w16[e + 1]);
                    }
                    ok = true;
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        ok = ok && (w16[e].y == seq);
                    }
                    if (ok || ++spins > SPIF_COMM_SPIN_LIMIT) {
                        break;
                    }
                    __nanosleep(100);
                }
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    h1_s[row + e] = ok ? __uint_as_float(w16[e].x) : __uint_as_float(0x7fc00000u);
                }
                if (!ok) {
                    atomicOr(p.comm.seq + 2, SPIF_COMM_ERR_TIMEOUT);
                }
            }
            mbar_arrive(h1bar);  // release: the consumers' wait acquires h1_s
            if (ht < 32) {
                LAYER_STAMP(9);
            }
        }
        if (p.has_ffn && ht == 0) {
            mbar_wait(flushbar, 0);  // every consumer warp has stored its slice of the partial y, the epilogue warp the counts
            __threadfence();
            atomicAdd(p.comm.seq + 3, 1u);
            LAYER_STAMP(4);
        }
    }
    __syncthreads();
    if (tid == 0) {
        spif_comm_block_done(p.comm, seq, (unsigned) P);
    }
}
