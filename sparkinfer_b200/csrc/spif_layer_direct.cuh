// spif_layer_direct.cuh -- the look-ahead layer kernel, version 4: every consumer warp streams its own column slice.
// Included by spif_layer.cu (shares LayerParams, the packed shuffle trees, layer_sum_partials and the launch code).
//
// What the TMA-ring versions taught (profiles/r2_layer_trace.txt):
//   * one issuing warp needs ~0.9 us per 40 KB ring item (≈250 dependent instructions at ~5 cycles each: policy, descriptors,
//     uniform-register shuffling for UBLKCP) -- 43 GB/s per SM, exactly the rate every ring version of this repo got stuck at;
//   * ring handshakes (full / done / free) put three roles in series for every item.
// Here there is no issuer and no ring.  Consumer warp w owns the 16-byte column slices [32w, 32w+32) of every weight row of
// the CTA's work (as before), and loads them itself: cp.async (LDGSTS) 16 B per lane into the thread's OWN 16 bytes of a
// staging slot, PIPE items of RPS rows in flight per thread (13B: 4 x 40 KB per SM), completion by cp.async.wait_group --
// a thread only ever reads back what it copied itself, so the staging needs no barrier at all.  (A first cut with plain
// ld.global into registers, two items in flight, ran at one item per 2 us: the loads of the next item and the use of the
// current one ended up behind the same scoreboard.)  A warp chooses its next item by itself:
//   W_down^T rows whose hidden value is known (pending list, filled by the epilogue warp)  >  Wpu rows  >  gate/up rows of
//   the active-row list  >  store my slice of the partial y  >  Wpd rows
// The only cross-warp traffic is the per-item partial dot (4 floats per warp) into a small ring in shared memory + one
// mbarrier arrival; the epilogue warp retires those in the canonical order Wpu -> gate/up -> Wpd.  W_down^T rows are summed
// by every warp in pending-list order (= ascending neuron order), so y stays bit-stable however the item types interleave.
#pragma once

#define SPIF_L2_K 8 /* depth of the partial-dot ring (items between a consumer warp and the epilogue warp) */
#define SPIF_L2_PIPE 4 /* items in flight per consumer thread */

enum { D_NONE = 0, D_WPU = 1, D_GU = 2, D_DN = 3, D_WPD = 4, D_FLUSH = 5, D_END = 6 };

struct Layer2Smem {
    uint32_t done, part, ctl, list_row, list_neu, list_bu, list_bg, pend_row, pend_alpha, bpu_s, bpd_s, lout_s, h1_s, blk_pre,
        red, wdesc, stage, total;
};

__host__ __device__ static inline Layer2Smem layer2_smem_layout(int cap_ffn, int cap_pd, int cap_pu, int r, int nblk, int nw,
                                                                int item_vecs) {
    Layer2Smem L;
    uint32_t   o = 0;
    auto       take = [&](uint32_t bytes, uint32_t align) {
        o               = (o + align - 1) / align * align;
        const uint32_t r = o;
        o += bytes;
        return r;
    };
    L.done       = take(8u * SPIF_L2_K, 8);
    L.part       = take(4u * 128 * SPIF_L2_K, 16);
    L.ctl        = take(64u, 16);  // int[8] {pend_tail, gu retired, b_end, total_active, items retired, -, -, -}, 3 mbarriers
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
    L.wdesc      = take(16u * SPIF_L2_PIPE * nw, 16);                  // per warp: descriptors of its items in flight
    L.stage      = take(16u * SPIF_L2_PIPE * item_vecs * nw * 32, 128);  // [PIPE][RPS * VPT][NC] x 16 B
    L.total      = (o + 127) / 128 * 128;
    return L;
}

// 16-byte asynchronous copy global -> shared (LDGSTS, L1 bypassed); src_bytes == 0 writes zeros without reading
__device__ __forceinline__ void cp_async_16(uint32_t dst_smem, const void * src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ int ld_acquire_cta_s32(const int * p) {
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_cta_s32(int * p, int v) {
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

struct L2Desc {
    int type, a, n, k;  // a: first row / list index / pending index / first Wpd row; n: rows (neurons for gate/up); k: reduce seq
};

// Warp roles: [0, NW) consumers, NW = epilogue, NW + 1 .. NW + 2 = helpers (row list, h1 polling, partial-y publication).
template <typename WT, int VPT, int RPS, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) spif_layer2_kernel(const __grid_constant__ LayerParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NW = p.nw, NC = NW * 32;
    const int P = gridDim.x, c = blockIdx.x;
    constexpr int K = SPIF_L2_K;

    const Layer2Smem L = layer2_smem_layout(p.cap_ffn, p.cap_pd, p.cap_pu, p.has_pred ? p.r : 0, p.lin_nblk, NW, RPS * VPT);
    uint64_t * done       = reinterpret_cast<uint64_t *>(smem + L.done);
    float *    part       = reinterpret_cast<float *>(smem + L.part);  // [K][4][32]
    int *      ctl        = reinterpret_cast<int *>(smem + L.ctl);     // see layer2_smem_layout
    uint64_t * h1bar      = reinterpret_cast<uint64_t *>(smem + L.ctl + 32);
    uint64_t * flushbar   = reinterpret_cast<uint64_t *>(smem + L.ctl + 40);
    uint64_t * listbar    = reinterpret_cast<uint64_t *>(smem + L.ctl + 48);
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
    int4 *     wdesc      = reinterpret_cast<int4 *>(smem + L.wdesc);

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

    int ra = 0, rb = 0, pa = 0, pb = 0;
    if (p.has_pred) {
        const int rw = p.r / world, R0 = p.comm.rank * rw;
        ra = R0 + share(c, rw, P);
        rb = R0 + share(c + 1, rw, P);
        pa = share(c, p.pm, P);
        pb = share(c + 1, p.pm, P);
    }
    const int neu_per_gu   = gated ? RPS / 2 : RPS;
    const int n_wpu        = (rb - ra + RPS - 1) / RPS;
    const int wpd_per_item = RPS * p.RP;
    const int n_wpd        = (pb - pa + wpd_per_item - 1) / wpd_per_item;

    if (warp < NW) {
        // ================================================ consumer warps ================================================
        float xr[VPT * 8], acc[VPT * 8];
        bool  vok[VPT];
        int   wrow[VPT], hcol[VPT];  // WPD: which of the RP Wpd rows of a virtual row my vector belongs to, and its h1 vector
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = tid + k * NC;
            vok[k]      = v < p.V;
            wrow[k]     = v / p.Vr;
            hcol[k]     = v - wrow[k] * p.Vr;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                xr[k * 8 + e] = 0.f, acc[k * 8 + e] = 0.f;
            }
        }
        const uint32_t row_bytes = (uint32_t) p.row_bytes, wrow_bytes = (uint32_t) p.r * 2;
        int  wpu_i = 0, gu_i = 0, wpd_i = 0, dn_i = 0, kseq = 0, n_gu = 0, b_end = 0;
        bool list_ready = !p.has_ffn, flushed = !p.has_ffn, h1_ready = false;

        auto next_item = [&](bool prewait) -> L2Desc {
            L2Desc d = { D_NONE, 0, 0, 0 };
            if (wpu_i < n_wpu) {
                d.type = D_WPU, d.a = ra + wpu_i * RPS, d.n = min(RPS, rb - d.a), d.k = kseq++;
                ++wpu_i;
                return d;
            }
            if (prewait) {
                return d;
            }
            if (p.has_ffn && !flushed) {
                if (!list_ready) {
                    if (!mbar_test(listbar, 0)) {
                        return d;
                    }
                    list_ready = true;
                    b_end      = ctl[2];
                    n_gu       = (b_end + neu_per_gu - 1) / neu_per_gu;
                }
                const int  gu_ret  = ld_acquire_cta_s32(ctl + 1);
                const int  avail   = ld_acquire_cta_s32(ctl + 0) - dn_i;  // read AFTER gu_ret: final once gu_ret == n_gu
                const bool gu_left = gu_i < n_gu;
                if (avail >= RPS || (avail > 0 && !gu_left)) {
                    d.type = D_DN, d.a = dn_i, d.n = min(RPS, avail);
                    dn_i += d.n;
                    return d;
                }
                if (gu_left) {
                    d.type = D_GU, d.a = gu_i * neu_per_gu, d.n = min(neu_per_gu, b_end - d.a), d.k = kseq++;
                    ++gu_i;
                    return d;
                }
                if (gu_ret == n_gu && avail == 0) {
                    d.type  = D_FLUSH;
                    flushed = true;
                    return d;
                }
            }
            if (wpd_i < n_wpd) {
                d.type = D_WPD, d.a = pa + wpd_i * wpd_per_item, d.n = min(wpd_per_item, pb - d.a), d.k = kseq++;
                ++wpd_i;
                return d;
            }
            if (flushed) {
                d.type = D_END;
            }
            return d;  // D_NONE: gate/up items are still with the epilogue warp, their W_down^T rows may follow
        };

        // my 16 bytes of (slot, row j): consecutive threads are consecutive in shared memory (conflict-free both ways)
        constexpr int  PIPE    = SPIF_L2_PIPE;
        const uint32_t stage_s = smem_u32(smem + L.stage) + (uint32_t) tid * 16u;
        const uint32_t jstride = (uint32_t) NC * 16u, sstride = jstride * (RPS * VPT);
        auto load_item = [&](const L2Desc & d, int slot) {
            if (d.type != D_NONE && d.type < D_FLUSH) {
#pragma unroll
                for (int r = 0; r < RPS; ++r) {
                    const unsigned char * src = reinterpret_cast<const unsigned char *>(p.x);  // any valid address (0 bytes read)
                    bool                  okr = false;
                    int                   lim = 0;  // WPD: Wpd rows of this virtual row that exist
                    if (d.type == D_WPU) {
                        okr = r < d.n;
                        src = p.wpu + (size_t) (d.a + r) * row_bytes;
                    } else if (d.type == D_GU) {
                        const int j = gated ? (r >> 1) : r;
                        okr         = j < d.n;
                        if (okr) {
                            const size_t off = (size_t) list_row[d.a + j] * row_bytes;
                            src              = ((gated && (r & 1)) ? p.w_gate : p.w_up) + off;
                        }
                    } else if (d.type == D_DN) {
                        okr = r < d.n;
                        if (okr) {
                            src = p.w_down + (size_t) pend_row[d.a + r] * row_bytes;
                        }
                    } else {  // D_WPD: virtual row r = RP consecutive Wpd rows
                        lim = d.n - r * p.RP;
                        okr = lim > 0;
                        src = p.wpd + (size_t) (d.a + r * p.RP) * wrow_bytes;
                    }
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        bool ok = okr && vok[k];
                        if (d.type == D_WPD) {
                            ok = ok && wrow[k] < min(lim, p.RP);
                        }
                        const unsigned char * sp = ok ? src + (size_t) (tid + k * NC) * 16 : reinterpret_cast<const unsigned char *>(p.x);
                        cp_async_16(stage_s + (uint32_t) slot * sstride + (uint32_t) (r * VPT + k) * jstride, sp, ok ? 16 : 0);
                    }
                }
            }
            cp_async_commit();  // also for items without rows: one group per item keeps wait_group in step
        };

        int qstamp = 0;
        auto compute_item = [&](const L2Desc & d, int slot) {
            if (d.type == D_NONE || d.type == D_END) {
                return;
            }
            uint4 w[RPS * VPT];
            if (d.type < D_FLUSH) {
#pragma unroll
                for (int j = 0; j < RPS * VPT; ++j) {
                    w[j] = lds128(stage_s + (uint32_t) slot * sstride + (uint32_t) j * jstride);
                }
            }
            if (warp == 0) {
                ITEM_STAMP(qstamp, 0, (unsigned long long) d.type);
                ITEM_STAMP(qstamp, 1, 0);
            }
            if (d.type == D_DN) {
#pragma unroll
                for (int r = 0; r < RPS; ++r) {
                    if (r < d.n) {  // uniform; pending order = ascending neuron order: a fixed FMA chain per element
                        const float alpha = pend_alpha[d.a + r];
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            float f[8];
                            unpack8<WT>(w[r * VPT + k], f);
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                acc[k * 8 + e] = fmaf(f[e], alpha, acc[k * 8 + e]);
                            }
                        }
                    }
                }
            } else if (d.type == D_FLUSH) {
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
                    mbar_arrive(flushbar);  // release; a helper thread fences and counts this CTA in (off this path)
                }
            } else {
                // a reduce item: 4 partial dots per warp into slot k % K of the ring; the epilogue warp must be done with
                // item k - K (it almost always is: it trails the consumers by one or two items)
                if (d.k >= K) {
                    while (d.k - ld_volatile_s32(ctl + 4) >= K) {
                        __nanosleep(20);
                    }
                }
                float * pw = part + (d.k % K) * 128 + warp;
                float   k2;
                if (d.type == D_WPD) {
                    if (!h1_ready) {
                        mbar_wait(h1bar, 0);  // the helper warps have collected the predictor's hidden vector
                        h1_ready = true;
                    }
                    float v[RPS * VPT];
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        const float4 h0 = *reinterpret_cast<const float4 *>(h1_s + hcol[k] * 8);
                        const float4 h1v = *reinterpret_cast<const float4 *>(h1_s + hcol[k] * 8 + 4);
                        float        hv[8] = { h0.x, h0.y, h0.z, h0.w, h1v.x, h1v.y, h1v.z, h1v.w };
                        if (p.pflags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                hv[e] = round_through<WT>(hv[e]);
                            }
                        }
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            v[r * VPT + k] = dot8<WT>(w[r * VPT + k], hv);
                        }
                    }
                    k2 = reduce4(v[0], v[1], v[2], v[3], lane);  // RPS * VPT == 4
                    if ((lane & 7) == 0) {
                        pw[(lane >> 3) * 32] = k2;
                    }
                } else {
                    float v[RPS];
#pragma unroll
                    for (int r = 0; r < RPS; ++r) {
                        v[r] = 0.f;
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            v[r] += dot8<WT>(w[r * VPT + k], xr + k * 8);
                        }
                    }
                    if (RPS == 4) {
                        k2 = reduce4(v[0], v[1], v[RPS - 2], v[RPS - 1], lane);
                        if ((lane & 7) == 0) {
                            pw[(lane >> 3) * 32] = k2;
                        }
                    } else {
                        k2 = reduce2(v[0], v[1], lane);
                        if ((lane & 15) == 0) {
                            pw[(lane >> 4) * 32] = k2;
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&done[d.k % K]);
                }
            }
            if (warp == 0) {
                ITEM_STAMP(qstamp, 3, 0);
                ++qstamp;
            }
        };

        // ---- before the programmatic-dependent-launch wait: the first predictor up-projection rows are on their way
        int4 * mydesc = wdesc + warp * PIPE;  // every lane writes the same value and reads back its own write: no sync needed
#pragma unroll
        for (int i = 0; i < PIPE; ++i) {
            const L2Desc d = next_item(true);
            mydesc[i]      = make_int4(d.type, d.a, d.n, d.k);
            load_item(d, i);
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
        bool ended = false;
        for (int slot = 0;;) {
            cp_async_wait<PIPE - 1>();  // the oldest item in flight has landed
            const int4   dd = mydesc[slot];
            const L2Desc d  = { dd.x, dd.y, dd.z, dd.w };
            compute_item(d, slot);
            if (d.type == D_END) {
                break;
            }
            L2Desc nd = { D_END, 0, 0, 0 };
            if (!ended) {
                nd    = next_item(false);
                ended = nd.type == D_END;
            }
            mydesc[slot] = make_int4(nd.type, nd.a, nd.n, nd.k);
            load_item(nd, slot);
            if (d.type == D_NONE && nd.type == D_NONE) {
                __nanosleep(40);
            }
            if (++slot == PIPE) {
                slot = 0;
            }
        }
        cp_async_wait<0>();
        if (warp == 0) {
            LAYER_STAMP(10);
        }
        // ---- D: y[columns of this CTA] = sum over all CTAs of their partial y, fixed order (+ b_down) [+ sum over shards]
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
        int  pend_tail = 0, n_nonzero = 0, cnt_out = 0, n_gu = 0, b_end = 0;
        bool have_list = !p.has_ffn;
        int  total = n_wpu + n_wpd;  // + n_gu once the list is known
        for (int k = 0;; ++k) {
            if (!have_list && k >= n_wpu) {  // the gate/up items follow the Wpu items; their number comes with the row list
                mbar_wait(listbar, 0);
                have_list = true;
                b_end     = ctl[2];
                n_gu      = (b_end + neu_per_gu - 1) / neu_per_gu;
                total += n_gu;
                if (n_gu == 0) {
                    if (lane == 0) {
                        p.cta_stats[c * 2 + 0] = ctl[3];
                        p.cta_stats[c * 2 + 1] = 0;
                        mbar_arrive(flushbar);
                    }
                }
            }
            if (k >= total) {
                break;
            }
            mbar_wait_spin(&done[k % K], (unsigned) (k / K) & 1u);
            const float * ps = part + (k % K) * 128;
            if (k < n_wpu + n_gu) {
                float kk;
                if (RPS == 4) {
                    const float a0 = (lane < NW) ? ps[lane] : 0.f, a1 = (lane < NW) ? ps[32 + lane] : 0.f;
                    const float a2 = (lane < NW) ? ps[64 + lane] : 0.f, a3 = (lane < NW) ? ps[96 + lane] : 0.f;
                    kk             = reduce4(a0, a1, a2, a3, lane);  // lanes [8j, 8j+8): row j
                } else {
                    const float a0 = (lane < NW) ? ps[lane] : 0.f, a1 = (lane < NW) ? ps[32 + lane] : 0.f;
                    kk             = reduce2(a0, a1, lane);          // lanes [16j, 16j+16): row j
                }
                constexpr int LPR = 32 / RPS;  // lanes per row sum
                if (k < n_wpu) {
                    // h1[row] = relu(Wpu[row] . x + bpu[row])  (llama-graph.cpp:871-880), published to every shard
                    const int a = ra + k * RPS, n = min(RPS, rb - a);
                    const int j = lane / LPR;
                    if ((lane % LPR) == 0 && j < n) {
                        const int row = a + j;
                        float     z   = kk + bpu_s[row - ra];
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
                    // gated: rows (up_0, gate_0, up_1, gate_1); ungated: (up_0 .. up_{RPS-1})
                    const int i0 = (k - n_wpu) * neu_per_gu, n = min(neu_per_gu, b_end - i0);
                    for (int j = 0; j < n; ++j) {
                        const int   i = i0 + j;
                        const float u = __shfl_sync(0xffffffffu, kk, (gated ? 2 * j : j) * LPR) + list_bu[i];
                        const float g = gated ? __shfl_sync(0xffffffffu, kk, (2 * j + 1) * LPR) + list_bg[i] : 0.f;
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
                    if (lane == 0) {
                        st_release_cta_s32(ctl + 0, pend_tail);            // after the pending entries
                        st_release_cta_s32(ctl + 1, k - n_wpu + 1);        // gate/up items retired, after pend_tail
                        if (k - n_wpu + 1 == n_gu) {
                            p.cta_stats[c * 2 + 0] = ctl[3];
                            p.cta_stats[c * 2 + 1] = n_nonzero;
                            mbar_arrive(flushbar);
                        }
                    }
                }
            } else {
                const int row0 = pa + (k - n_wpu - n_gu) * wpd_per_item, n = min(wpd_per_item, pb - row0);
                for (int base = 0; base < n; base += 32) {
                    const int  idx   = base + lane;
                    const bool valid = idx < n;
                    float      z     = 0.f;
                    if (valid) {
                        // Wpd row idx of the item lives in virtual row j = idx / RP; its 32-vector groups are g0 .. g0 + WR - 1
                        // of that virtual row; group g was reduced by consumer warp g % NW for its vector kv = g / NW
                        const int j = idx / p.RP, g0 = (idx - j * p.RP) * p.WR;
                        for (int t = 0; t < p.WR; ++t) {  // fixed order
                            const int g = g0 + t, kv = g / NW, w = g - kv * NW;
                            z += ps[(j * VPT + kv) * 32 + w];
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
