// spif_layer_stream.cuh -- the look-ahead layer kernel, version 5: a STATIC per-CTA schedule of register-direct row batches.
// Included by spif_layer.cu (shares LayerParams, layer_sum_partials and the launch code).
//
// What versions 2-4 measured (profiles/r2_layer_trace_*.txt, profiles/r2_ncu_layer_*.txt):
//   * TMA ring (v2/v3): ONE issuing thread hands the ring a 40 KB item per ~1.0 us whatever the item is (even items without
//     copies), i.e. one item in flight: 40 GB/s per SM at best, and three roles in series per item.
//   * cp.async staging (v4): 22 M warp instructions per launch (address selection per 16 bytes, polling): issue-bound, 33 us.
//   * tools/ubench/stream_rows.cu: plain 128-bit loads straight into registers, 8 rows in flight per thread, warps running
//     free of each other, reach 6.5 TB/s -- the best of all load paths, with no staging at all.
// So: no ring, no issuer, no dynamic queue.  The work of a CTA is a fixed sequence of ITEMS of 8 row slices per thread
// (8 x 16 B in flight per thread, 80 KB per SM at 13B):
//
//     Wpu rows -> [row list ready] -> gate/up rows -> first Wpd item -> [hidden values known] -> W_down^T rows of the
//     non-zero hidden values -> partial y stored -> remaining Wpd rows -> column-sliced sum of all CTAs' partial y
//
// Every consumer warp walks this sequence on its own (it owns the 16-byte column slices [32w, 32w+32) of every row): issue
// the 8 loads of an item, multiply, one packed shuffle tree for the 8 partial dots, 8 floats into a small ring in shared
// memory, one mbarrier arrival.  Warps never wait for each other inside an item type; the three points where the sequence
// depends on other warps / CTAs (row list, hidden values, the predictor's hidden vector) sit behind other streaming work.
// The epilogue warp retires the partial dots in the canonical order Wpu -> gate/up -> Wpd (bias, activation, sigmoid,
// compaction); helper warps do everything that waits on other CTAs (row list, h1 gather, partial-y publication).
// Determinism: fixed shuffle tree, fixed warp order, W_down^T rows in ascending neuron order, partials in CTA order.
#pragma once

#define SPIF_L3_K 8 /* depth of the partial-dot ring (reduce items between the consumer warps and the epilogue warp) */

struct Layer3Smem {
    uint32_t done, part, ctl, list_row, list_neu, list_bu, list_bg, pend_row, pend_alpha, bpu_s, bpd_s, lout_s, h1_s, blk_pre,
        red, total;
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
    L.part       = take(4u * 256 * SPIF_L3_K, 16);  // [K][8][32]
    L.ctl        = take(96u, 16);  // int[8] {pend_tail, -, b_end, total_active, items retired, ...}, 4 mbarriers at +32
    L.list_row   = take(4u * cap_ffn, 4);
    L.list_neu   = take(4u * cap_ffn, 4);
    L.list_bu    = take(4u * cap_ffn, 4);
    L.list_bg    = take(4u * cap_ffn, 4);
    L.pend_row   = take(4u * cap_ffn, 4);
    L.pend_alpha = take(4u * cap_ffn, 4);
    L.bpu_s      = take(4u * cap_pu, 4);
    L.bpd_s      = take(4u * cap_pd, 4);
    L.lout_s     = take(4u * cap_pd, 4);
    L.h1_s       = take(4u * r, 16);
    L.blk_pre    = take(4u * (nblk + 1), 4);
    L.red        = take(16u * (2 * nw + 4) * 16, 16);  // scratch of the column-sliced sum
    L.total      = (o + 127) / 128 * 128;
    return L;
}

// eight warp reductions with nine shuffles: every lane of group g = lane >> 2 returns the warp sum of v[g]
__device__ __forceinline__ float reduce8(const float (&v)[8], int lane) {
    const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
    float      a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        a[i] = (b4 ? v[4 + i] : v[i]) + __shfl_xor_sync(0xffffffffu, b4 ? v[i] : v[4 + i], 16);
    }
    float b[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        b[i] = (b3 ? a[2 + i] : a[i]) + __shfl_xor_sync(0xffffffffu, b3 ? a[i] : a[2 + i], 8);
    }
    float c = (b2 ? b[1] : b[0]) + __shfl_xor_sync(0xffffffffu, b2 ? b[0] : b[1], 4);
    c += __shfl_xor_sync(0xffffffffu, c, 2);
    c += __shfl_xor_sync(0xffffffffu, c, 1);
    return c;
}

// Warp roles: [0, NW) consumers, NW = epilogue, NW + 1 .. NW + 2 = helpers (row list, h1 gather, partial-y publication).
// An item is RPI = 8 / VPT rows: every consumer thread holds 8 x 16 bytes of it.
template <typename WT, int VPT, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) spif_layer3_kernel(const __grid_constant__ LayerParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int RPI = 8 / VPT;
    constexpr int K   = SPIF_L3_K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NW = p.nw, NC = NW * 32;
    const int P = gridDim.x, c = blockIdx.x;

    const Layer3Smem L = layer3_smem_layout(p.cap_ffn, p.cap_pd, p.cap_pu, p.has_pred ? p.r : 0, p.lin_nblk, NW);
    uint64_t * done       = reinterpret_cast<uint64_t *>(smem + L.done);
    float *    part       = reinterpret_cast<float *>(smem + L.part);  // [K][8][32]
    int *      ctl        = reinterpret_cast<int *>(smem + L.ctl);
    uint64_t * h1bar      = reinterpret_cast<uint64_t *>(smem + L.ctl + 32);
    uint64_t * flushbar   = reinterpret_cast<uint64_t *>(smem + L.ctl + 40);
    uint64_t * listbar    = reinterpret_cast<uint64_t *>(smem + L.ctl + 48);
    uint64_t * dnbar      = reinterpret_cast<uint64_t *>(smem + L.ctl + 56);
    int *      list_row   = reinterpret_cast<int *>(smem + L.list_row);
    int *      list_neu   = reinterpret_cast<int *>(smem + L.list_neu);
    float *    list_bu    = reinterpret_cast<float *>(smem + L.list_bu);
    float *    list_bg    = reinterpret_cast<float *>(smem + L.list_bg);
    int *      pend_row   = reinterpret_cast<int *>(smem + L.pend_row);
    float *    pend_alpha = reinterpret_cast<float *>(smem + L.pend_alpha);
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
        for (int s = 0; s < K; ++s) {
            mbar_init(&done[s], NW);
        }
        mbar_init(h1bar, 64);
        mbar_init(flushbar, NW + 1);
        mbar_init(listbar, 1);
        mbar_init(dnbar, 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            ctl[i] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();

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
    const int n_wpd1 = p.has_ffn ? min(n_wpd, 1) : n_wpd;    // Wpd items streamed while the hidden values are being retired

    if (warp < NW) {
        // ================================================ consumer warps ================================================
        float    xr[VPT * 8], acc[VPT * 8];
        uint32_t toff[VPT];
        bool     vok[VPT], pok[VPT];
        int      wrow[VPT];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = tid + k * NC;
            vok[k]      = v < p.V;
            toff[k]     = (uint32_t) v * 16u;
            wrow[k]     = v / p.Vr;                  // Wpd: which of the RP rows of a virtual row my vector belongs to
            pok[k]      = vok[k] && wrow[k] < p.RP;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                xr[k * 8 + e] = 0.f, acc[k * 8 + e] = 0.f;
            }
        }
        const uint32_t row_bytes = (uint32_t) p.row_bytes, wrow_bytes = (uint32_t) p.r * 2;
        const uint4    zero4     = make_uint4(0u, 0u, 0u, 0u);
        uint4          w[RPI][VPT];
        int            kseq = 0, qstamp = 0;

        auto load_wpu = [&](int i) {
            const int a = ra + i * RPI, n = min(RPI, rb - a);
#pragma unroll
            for (int j = 0; j < RPI; ++j) {
                const unsigned char * src = p.wpu + (size_t) (a + j) * row_bytes;
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    w[j][k] = (j < n && vok[k]) ? ldg_stream_v4(src + toff[k]) : zero4;
                }
            }
        };
        // RPI x VPT partial dots of the item against vec (x or the predictor's hidden vector) -> slot kidx % K of the ring
        auto dot_deposit = [&](const float * vec, int kidx) {
            float v[8];
#pragma unroll
            for (int j = 0; j < RPI; ++j) {
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    v[j * VPT + k] = dot8<WT>(w[j][k], vec + k * 8);
                }
            }
            const float s = reduce8(v, lane);
            if (kidx >= K) {  // the epilogue warp must be done with item kidx - K (it trails by one or two items)
                while (kidx - ld_volatile_s32(ctl + 4) >= K) {
                    __nanosleep(20);
                }
            }
            if ((lane & 3) == 0) {
                part[(kidx % K) * 256 + (lane >> 2) * 32 + warp] = s;
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&done[kidx % K]);
            }
        };
#define L3_STAMP_BEGIN(ty)                                        \
    if (warp == 0) {                                              \
        ITEM_STAMP(qstamp, 0, (unsigned long long) (ty));         \
        ITEM_STAMP(qstamp, 1, 0);                                 \
    }
#define L3_STAMP_END()               \
    if (warp == 0) {                 \
        ITEM_STAMP(qstamp, 3, 0);    \
        ++qstamp;                    \
    }

        // ---- Wpu rows: the first item is on its way before the programmatic-dependent-launch wait (weights only)
        L3_STAMP_BEGIN(D_WPU);
        if (n_wpu > 0) {
            load_wpu(0);
        }
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        if (warp == 0) {
            LAYER_STAMP(1);
        }
        seq = spif_comm_call_seq(p.comm);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            if (vok[k]) {
                const int    v = tid + k * NC;
                const float4 a = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v + 1);
                xr[k * 8 + 0] = a.x, xr[k * 8 + 1] = a.y, xr[k * 8 + 2] = a.z, xr[k * 8 + 3] = a.w;
                xr[k * 8 + 4] = b.x, xr[k * 8 + 5] = b.y, xr[k * 8 + 6] = b.z, xr[k * 8 + 7] = b.w;
                if (rounding) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        xr[k * 8 + e] = round_through<WT>(xr[k * 8 + e]);
                    }
                }
            }
        }
        for (int i = 0; i < n_wpu; ++i) {
            if (i > 0) {
                L3_STAMP_BEGIN(D_WPU);
                load_wpu(i);
            }
            dot_deposit(xr, kseq++);
            L3_STAMP_END();
        }
        if (n_wpu == 0) {
            L3_STAMP_END();
        }

        // ---- gate/up rows of my share of the active-row list (the helper warp has fetched it meanwhile)
        if (p.has_ffn) {
            mbar_wait(listbar, 0);
            const int b_end = ctl[2];
            for (int i0 = 0; i0 < b_end; i0 += npg) {
                const int n = min(npg, b_end - i0);
                L3_STAMP_BEGIN(D_GU);
#pragma unroll
                for (int j = 0; j < RPI; ++j) {
                    const int  jj = gated ? (j >> 1) : j;
                    const bool ok = jj < n;
                    const unsigned char * src = p.w_up;
                    if (ok) {
                        src = ((gated && (j & 1)) ? p.w_gate : p.w_up) + (size_t) list_row[i0 + jj] * row_bytes;
                    }
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        w[j][k] = (ok && vok[k]) ? ldg_stream_v4(src + toff[k]) : zero4;
                    }
                }
                dot_deposit(xr, kseq++);
                L3_STAMP_END();
            }
        }

        // ---- Wpd rows: dense, known at entry; they need the predictor's hidden vector (gathered by the helper warps while the
        // rows above streamed).  x is dead from here on: its registers hold my 8 (16) values of h1.
        float hr[VPT * 8];
        bool  h1_ready = false;
        auto wpd_item = [&](int i) {
            L3_STAMP_BEGIN(D_WPD);
            const int row0 = pa + i * wpi;
#pragma unroll
            for (int j = 0; j < RPI; ++j) {
                const int             rj  = row0 + j * p.RP;  // first Wpd row of virtual row j
                const unsigned char * src = p.wpd + (size_t) rj * wrow_bytes;
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    w[j][k] = (pok[k] && rj + wrow[k] < pb) ? ldg_stream_v4(src + toff[k]) : zero4;
                }
            }
            if (!h1_ready) {
                h1_ready = true;
                mbar_wait(h1bar, 0);
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    const int    hcol = (tid + k * NC) - wrow[k] * p.Vr;
                    const float4 h0   = *reinterpret_cast<const float4 *>(h1_s + (pok[k] ? hcol : 0) * 8);
                    const float4 h1v  = *reinterpret_cast<const float4 *>(h1_s + (pok[k] ? hcol : 0) * 8 + 4);
                    hr[k * 8 + 0] = h0.x, hr[k * 8 + 1] = h0.y, hr[k * 8 + 2] = h0.z, hr[k * 8 + 3] = h0.w;
                    hr[k * 8 + 4] = h1v.x, hr[k * 8 + 5] = h1v.y, hr[k * 8 + 6] = h1v.z, hr[k * 8 + 7] = h1v.w;
                    if (p.pflags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            hr[k * 8 + e] = round_through<WT>(hr[k * 8 + e]);
                        }
                    }
                }
            }
            dot_deposit(hr, kseq++);
            L3_STAMP_END();
        };
        for (int i = 0; i < n_wpd1; ++i) {
            wpd_item(i);
        }

        // ---- W_down^T rows of the non-zero hidden values (axpy-sparse.cu:53), ascending: a fixed FMA chain per element
        if (p.has_ffn) {
            mbar_wait(dnbar, 0);  // the epilogue warp has retired every gate/up item
            const int n_pend = ctl[0];
            for (int i0 = 0; i0 < n_pend; i0 += RPI) {
                const int n = min(RPI, n_pend - i0);
                L3_STAMP_BEGIN(D_DN);
#pragma unroll
                for (int j = 0; j < RPI; ++j) {
                    const bool            ok  = j < n;
                    const unsigned char * src = p.w_down + (size_t) (ok ? pend_row[i0 + j] : 0) * row_bytes;
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        w[j][k] = (ok && vok[k]) ? ldg_stream_v4(src + toff[k]) : zero4;
                    }
                }
#pragma unroll
                for (int j = 0; j < RPI; ++j) {
                    if (j < n) {  // uniform
                        const float alpha = pend_alpha[i0 + j];
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            float f[8];
                            unpack8<WT>(w[j][k], f);
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                acc[k * 8 + e] = fmaf(f[e], alpha, acc[k * 8 + e]);
                            }
                        }
                    }
                }
                L3_STAMP_END();
            }
            // my slice of this CTA's partial y; a helper thread publishes it (fence + counter) off this path
            float * prt = p.partials + (size_t) c * p.n_embd;
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                if (vok[k]) {
                    float4 * o = reinterpret_cast<float4 *>(prt) + 2 * (tid + k * NC);
                    __stcg(o, make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]));
                    __stcg(o + 1, make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]));
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(flushbar);
            }
        }

        // ---- the remaining Wpd rows cover the time until every CTA has published its partial y
        for (int i = n_wpd1; i < n_wpd; ++i) {
            wpd_item(i);
        }
        if (warp == 0) {
            LAYER_STAMP(10);
        }
        // ---- y[columns of this CTA] = sum over all CTAs of their partial y, fixed order (+ b_down) [+ sum over the shards]
        if (p.has_ffn) {
            layer_sum_partials<true>(p, tid, NC, 1, red, seq, 2 * NW);
        }
#undef L3_STAMP_BEGIN
#undef L3_STAMP_END
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
                    mbar_arrive(dnbar);
                    mbar_arrive(flushbar);
                }
            }
            if (k >= total) {
                break;
            }
            mbar_wait_spin(&done[k % K], (unsigned) (k / K) & 1u);
            const float * ps = part + (k % K) * 256;
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
                                pend_row[pend_tail]   = list_row[i];
                                pend_alpha[pend_tail] = h;
                            }
                            ++pend_tail;
                            ++n_nonzero;
                        }
                    }
                    if (k - n_wpu + 1 == n_gu && lane == 0) {  // the W_down^T list is final
                        ctl[0]                 = pend_tail;
                        p.cta_stats[c * 2 + 0] = ctl[3];
                        p.cta_stats[c * 2 + 1] = n_nonzero;
                        mbar_arrive(dnbar);     // release: the consumers' wait acquires pend_row / pend_alpha / ctl[0]
                        mbar_arrive(flushbar);
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
            }
            __syncwarp();
            if (lane == 0) {
                ctl[2] = b_end;
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
