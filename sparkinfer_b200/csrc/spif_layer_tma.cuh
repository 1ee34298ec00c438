// spif_layer_tma.cuh -- the look-ahead layer kernel, version 6: row tables + ONE producer lane + the TMA unit.
// Included by spif_layer.cu (shares LayerParams, the packed shuffle trees, layer_sum_partials and the launch code).
//
// Version 5 (spif_layer_stream.cuh) proved the schedule -- tables of rows in shared memory, a non-blocking priority pick, the
// partial y complete right after the last gate/up item -- but its consumers issued their own loads: a table lookup, 64-bit
// address arithmetic, a predicate and an LDGSTS per 16 bytes (which ptxas pads with three dummy LDS on sm_100a), and read the
// staging back: ~190 warp instructions per 4-row item, 91 K per SM and launch -- issue-bound at 23 us per 13B layer whatever
// the arithmetic cost (profiles/r2_ncu_layer_stream.txt).  Here one lane of a producer warp hands whole rows to the TMA
// unit (cp.async.bulk, 10 KB per instruction, address and length straight from the table) into a ring of 4-row slots;
// completion by complete_tx on the slot's `full` barrier, recycling by the consumer warps' arrivals on its `empty` barrier
// as soon as their 16-byte slices are in registers.  A consumer's item is: wait, 4 LDS.128, 32 FHFMA (or 64 convert + FFMA),
// one packed shuffle tree, 4 floats into the partial ring, two barrier arrivals.
#pragma once

__device__ __forceinline__ int ld_acquire_cta_s32(const int * p) {
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_cta_s32(int * p, int v) {
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
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

#define SPIF_L4_K 16 /* depth of the partial-dot ring (reduce items between the consumer warps and the epilogue warp) */

// extra clock64 stamps of the traced instantiation: [500] epilogue sees the last gate/up item complete, [501] epilogue has
// published the final W_down^T table, [502] producer first sees it, [504] helper: every warp has stored its partial y
#define CLK_STAMP(idx)                                                                              \
    do {                                                                                            \
        if (TRACE && p.trace && lane == 0) {                                                        \
            p.trace[(size_t) blockIdx.x * 512 + (idx)] = (unsigned long long) clock64();            \
        }                                                                                           \
    } while (0)

struct Layer4Smem {
    uint32_t ring, full, empty, sdesc, done, part, ctl, list_row, list_neu, list_bu, list_bg, tab_a, tab_b, tab_c, tab_d, bpu_s,
        bpd_s, lout_s, h1_s, blk_pre, nsum, total;
};

__host__ __device__ static inline Layer4Smem layer4_smem_layout(int nslots, int slot_bytes, int cap_ffn, int cap_pd, int cap_pu,
                                                                int r, int nblk) {
    Layer4Smem L;
    uint32_t   o = 0;
    auto       take = [&](uint32_t bytes, uint32_t align) {
        o               = (o + align - 1) / align * align;
        const uint32_t r = o;
        o += bytes;
        return r;
    };
    L.ring     = take((uint32_t) nslots * slot_bytes, 128);  // also the scratch of the final column-sliced sum
    L.full     = take(8u * nslots, 8);
    L.empty    = take(8u * nslots, 8);
    L.sdesc    = take(16u * nslots, 16);
    L.done     = take(8u * SPIF_L4_K, 8);
    L.part     = take(4u * 128 * SPIF_L4_K, 16);  // [K][4][32]
    L.ctl      = take(96u, 16);  // int[8] {W_down^T rows so far, final rows + 1, b_end, total_active, items retired, gate/up items}, 3 mbarriers at +32
    L.list_row = take(4u * cap_ffn, 4);
    L.list_neu = take(4u * cap_ffn, 4);
    L.list_bu  = take(4u * cap_ffn, 4);
    L.list_bg  = take(4u * cap_ffn, 4);
    L.tab_a    = take(16u * (cap_pu + 4), 16);       // Wpu rows
    L.tab_b    = take(16u * (2 * cap_ffn + 4), 16);  // gate/up rows
    L.tab_c    = take(16u * (cap_pd + 4), 16);       // Wpd virtual rows (at most one per Wpd row)
    L.tab_d    = take(16u * (cap_ffn + 4), 16);      // W_down^T rows
    L.bpu_s    = take(4u * cap_pu, 4);
    L.bpd_s    = take(4u * cap_pd, 4);
    L.lout_s   = take(4u * cap_pd, 4);
    L.h1_s     = take(4u * r, 16);
    L.blk_pre  = take(4u * (nblk + 1), 4);
    L.nsum     = take(8u * 32, 8);  // per consumer warp: sum of squares of its slice of x (fused RMSNorm)
    L.total    = (o + 127) / 128 * 128;
    return L;
}

// 1 / sqrtf(mean(x^2) + eps) from every consumer thread's double partial sum of squares (out of line: its doubles and the
// barrier must not touch the register allocation of the streaming loop)
__device__ __noinline__ float layer_rms_scale(double ss, double * nsum, int warp, int lane, int NW, int NC, int n_embd, float eps) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
    }
    if (lane == 0) {
        nsum[warp] = ss;
    }
    named_bar_sync(2, NC);
    double tot = 0.0;
    for (int w = 0; w < NW; ++w) {
        tot += nsum[w];
    }
    const float mean = (float) (tot / (double) n_embd);
    return __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
}

// Warp roles: [0, NW) consumers, NW = epilogue, NW + 1 .. NW + 2 = helpers (row list, h1 gather, partial-y publication),
// NW + 3 = producer.
// An item is RPI = 4 / VPT rows (4 x 16 bytes per consumer thread) = one slot of the ring.
template <typename WT, int VPT, int MAXT, bool TRACE, bool RND>
__global__ void __launch_bounds__(MAXT, 1) spif_layer4_kernel(const __grid_constant__ LayerParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int RPI   = 4 / VPT;  // rows per item: 4 x 16 bytes per thread
    constexpr int K   = SPIF_L4_K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NW = p.nw, NC = NW * 32;
    const int P = gridDim.x, c = blockIdx.x;

    {   // the parameters come through the constant cache, cold at every launch: touch all of its lines at once instead of
        // one miss after the other along the prologue's dependency chain
        const int * pw = reinterpret_cast<const int *>(&p);
#pragma unroll
        for (int i = 0; i < (int) (sizeof(LayerParams) / sizeof(int)); i += 16) {
            asm volatile("" ::"r"(pw[i]));
        }
    }
    const int S = p.nslots, slot_bytes = RPI * p.row_bytes;

    const Layer4Smem L = layer4_smem_layout(S, slot_bytes, p.cap_ffn, p.cap_pd, p.cap_pu, p.has_pred ? p.r : 0, p.lin_nblk);
    unsigned char * ring  = smem + L.ring;
    uint64_t * full       = reinterpret_cast<uint64_t *>(smem + L.full);
    uint64_t * empty      = reinterpret_cast<uint64_t *>(smem + L.empty);
    int4 *     sdesc      = reinterpret_cast<int4 *>(smem + L.sdesc);
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
    double *   nsum       = reinterpret_cast<double *>(smem + L.nsum);

    if (tid == 0) {
        LAYER_STAMP(7);
        if (p.trace) {
            p.trace[(size_t) blockIdx.x * 512 + 15] = (unsigned long long) clock64();
        }
    }
    if (tid < K + 3 + 2 * S) {  // one barrier per thread: serial mbarrier.init cost 0.3 us of every launch
        if (tid < K) {
            mbar_init(&done[tid], NW);
        } else if (tid == K) {
            mbar_init(h1bar, 64);
        } else if (tid == K + 1) {
            mbar_init(flushbar, NW + 1);
        } else if (tid == K + 2) {
            mbar_init(listbar, 1);
        } else if (tid < K + 3 + S) {
            mbar_init(&full[tid - (K + 3)], 1);
        } else {
            mbar_init(&empty[tid - (K + 3 + S)], NW);
        }
        mbar_fence_init();
    }
    if (tid >= 32 && tid < 40) {
        ctl[tid - 32] = 0;
    }

    const bool gated    = (p.arch != SPIF_ARCH_RELU_UP);
    const bool rounding = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
    // (rejected experiments, profiles/r2_layer_experiments.txt: consumer warps in step per gate/up and W_down^T item +0.6 us;
    // every warp publishing its own slice of the partial y with its own fence +0.4 us -- a __threadfence under this load costs
    // ~1.1 us wherever it stands, so one per CTA it stays)
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
                vr[k * 8 + 0] = a.x, vr[k * 8 + 1] = a.y, vr[k * 8 + 2] = a.z, vr[k * 8 + 3] = a.w;
                vr[k * 8 + 4] = b.x, vr[k * 8 + 5] = b.y, vr[k * 8 + 6] = b.z, vr[k * 8 + 7] = b.w;
            }
        }
        if (p.norm_w) {
            // build_norm(ffn_inp, ffn_norm, LLM_NORM_RMS) fused (spif_layer_io): the arithmetic of
            // ggml_compute_forward_rms_norm_f32 (ops.cpp:3756-3775) + ggml_mul -- squares rounded to F32, summed in double,
            // mean and scale in F32, y = (x * scale) * w with two roundings.  Every CTA normalises the whole vector itself (its
            // consumer threads hold all of x): one shuffle tree in double + one named barrier among the consumer warps.
            double ss = 0.0;
#pragma unroll
            for (int e = 0; e < VPT * 8; ++e) {
                ss += (double) __fmul_rn(vr[e], vr[e]);  // threads beyond the row hold zeros
            }
            const float scale = layer_rms_scale(ss, nsum, warp, lane, NW, NC, p.n_embd, p.norm_eps);
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                if (toff[k] < row_bytes) {
                    const int    v = tid + k * NC;
                    const float4 a = __ldg(reinterpret_cast<const float4 *>(p.norm_w) + 2 * v);
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.norm_w) + 2 * v + 1);
                    const float  wv[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        vr[k * 8 + e] = __fmul_rn(__fmul_rn(vr[k * 8 + e], scale), wv[e]);
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            if (RND) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
                vh[k * 4 + 0] = pack2<WT>(vr[k * 8 + 0], vr[k * 8 + 1]), vh[k * 4 + 1] = pack2<WT>(vr[k * 8 + 2], vr[k * 8 + 3]);
                vh[k * 4 + 2] = pack2<WT>(vr[k * 8 + 4], vr[k * 8 + 5]), vh[k * 4 + 3] = pack2<WT>(vr[k * 8 + 6], vr[k * 8 + 7]);
            } else if (rounding) {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    vr[k * 8 + e] = round_through<WT>(vr[k * 8 + e]);
                }
            }
        }
        if (warp == 0) {
            LAYER_STAMP(8);
        }
        // my vector(s) of slot 0, row 0.  Threads beyond the end of a row (n_embd / 8 not a multiple of 32, e.g. 1376 or 4544)
        // hold x == 0 / unsummed h1 lanes and re-read the row's LAST vector: bytes beyond the row belong to the next row, the next
        // slot or stale shared memory, and a NaN bit pattern there times their zero made the whole dot NaN -- which the ReLU-type
        // activations then turned into a silent hidden == 0 (seen as a flaky parity failure at 1376 x 2048 in round 2)
        const uint32_t ring_s = smem_u32(ring) + min((uint32_t) tid * 16u, row_bytes - 16u);
        const uint32_t koff   = min((uint32_t) (tid + NC) * 16u, row_bytes - 16u) - min((uint32_t) tid * 16u, row_bytes - 16u);  // my next vector (VPT == 2)
        bool           h1_ready = false;
        int            kseq = 0, qc = 0;
        uint32_t       slot_s = ring_s;
        for (int s = 0, ph = 0;;) {
            mbar_wait(&full[s], (unsigned) ph);
            const int4 d  = sdesc[s];  // {type, table address of the item's rows, 1: some rows are short or missing, name}
            const int  ty = d.x;
            if (ty == T_END) {
                break;
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
                        w[j][k] = lds128(slot_s + (uint32_t) j * row_bytes + (uint32_t) k * koff);
                    }
                }
                if (d.z) {  // rare (the last item of a table): bytes beyond a row's valid length are stale
#pragma unroll
                    for (int j = 0; j < RPI; ++j) {
                        uint32_t lim;
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lim) : "r"((uint32_t) d.y + j * 16u + 8u));
#pragma unroll
                        for (int k = 0; k < VPT; ++k) {
                            if (!(toff[k] < lim)) {
                                w[j][k] = make_uint4(0u, 0u, 0u, 0u);
                            }
                        }
                    }
                }
                if (ty == T_AXPY) {
                    // table order = ascending neuron order: a fixed FMA chain per element; missing rows are zeros times zero
#pragma unroll
                    for (int j = 0; j < RPI; ++j) {
                        float alpha;
                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(alpha) : "r"((uint32_t) d.y + j * 16u + 12u));
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
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&empty[s]);  // my rows are in registers / consumed: the slot can be refilled
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
                    auto dotv = [&](const uint4 & wv, int k) -> float {
                        return RND ? dot8h<WT>(wv, vh + k * 4) : dot8<WT>(wv, vr + k * 8);
                    };
                    float v0, v1, v2, v3;
                    if (VPT == 1) {
                        v0 = dotv(w[0][0], 0), v1 = dotv(w[1 % RPI][0], 0);
                        v2 = dotv(w[2 % RPI][0], 0), v3 = dotv(w[3 % RPI][0], 0);
                    } else {
                        v0 = dotv(w[0][0], 0), v1 = dotv(w[0][VPT - 1], VPT - 1);
                        v2 = dotv(w[1 % RPI][0], 0), v3 = dotv(w[1 % RPI][VPT - 1], VPT - 1);
                    }
                    const float sred = reduce4(v0, v1, v2, v3, lane);  // lanes [8q, 8q+8): partial q = row * VPT + vector
                    if (lane == 0) {
                        mbar_arrive(&empty[s]);  // the shuffles above needed every lane's dots: all rows are consumed
                    }
                    if (kseq >= K) {  // the epilogue warp must be done with item kseq - K (it trails by one or two items)
                        while (kseq - ld_volatile_s32(ctl + 4) >= K) {
                            __nanosleep(20);
                        }
                    }
                    const int ks = kseq & (K - 1);
                    if ((lane & 7) == 0) {
                        part[ks * 128 + (lane >> 3) * 32 + warp] = sred;
                    }
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&done[ks]);
                    }
                    ++kseq;
                }
            } else {  // T_FLUSH
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
                    mbar_arrive(&empty[s]);
                }
            }
            if (TRACE && warp == 0) {
                ITEM_STAMP(qc, 3, 0);
                ++qc;
            }
            slot_s += slot_bytes;
            if (++s == S) {
                s = 0, ph ^= 1, slot_s = ring_s;
            }
        }
        if (warp == 0) {
            LAYER_STAMP(10);
        }
        // ---- y[columns of this CTA] = sum over all CTAs of their partial y, fixed order (+ b_down) [+ sum over the shards].
        // Its scratch is the ring: every consumer warp is past the loop when the first barrier inside releases.
        if (p.has_ffn) {
            layer_sum_partials<true>(p, tid, NC, 1, reinterpret_cast<float4 *>(ring), seq, 2 * NW);
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
            if (have_list && k == n_wpu + n_gu - 1) {
                CLK_STAMP(500);
            }
            const float * ps = part + (k % K) * 128;
            if (k < n_wpu + n_gu) {
                // row sum = sum over (vector k', warp w) of the partial dots, in a fixed order: LPR lanes per row, then a tree
                float s = 0.f;
                if (VPT == 1) {
                    for (int e = esub; e < NW; e += LPR) {
                        s += ps[erow * 32 + e];
                    }
                } else {
                    for (int e = esub; e < VPT * NW; e += LPR) {
                        const int kk = e / NW, ww = e - kk * NW;
                        s += ps[(erow * VPT + kk) * 32 + ww];
                    }
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
                        if (p.comm.mc) {  // one multicast store instead of one per shard
                            st_word_sys(p.comm.mc + off, __float_as_uint(z), seq);
                        } else {
                            for (int r = 0; r < world; ++r) {
                                st_word_sys(p.comm.data[r] + off, __float_as_uint(z), seq);
                            }
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
                            st_release_cta_s32(ctl + 1, n_dn * RPI + 1);  // first: the producer is waiting for exactly this
                            CLK_STAMP(501);
                            p.cta_stats[c * 2 + 0] = ctl[3];
                            p.cta_stats[c * 2 + 1] = n_nonzero;
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
    } else if (warp == NW + 3) {
        // ================================================== producer warp ==================================================
        // ONE lane decides what streams next and hands it to the TMA unit: a bulk copy per row (address and valid bytes from
        // the row tables) into the next free slot of the ring, completion by complete_tx on the slot's `full` barrier.  The
        // consumers never compute an address and never issue a load.
        //
        // Priority: Wpu rows (the predictor's hidden vector has the longest way to go: every CTA of every shard) > W_down^T rows
        // whose hidden values the epilogue warp has retired (4 table rows at a time, so the partial y is complete -- and on
        // its way to the other CTAs -- right after the last gate/up item) > gate/up rows > Wpd rows (dense, the cover for
        // everything that waits on other CTAs).
        // The whole warp runs the (uniform) decision; lane j < RPI looks up row j and issues its copy, so an item costs one
        // pass instead of RPI serial ones -- a single lane needed ~0.6 us per item on a busy SM, more than the 0.9 us the HBM
        // needs for it (profiles/r2_layer_trace_tma.txt, first cut).
        {
            const uint64_t pol  = l2_policy_evict_first();
            const bool     hint = p.l2_hint != 0;
            int      a_left = n_wpu, b_left = p.has_ffn ? -1 : 0, c_left = n_wpd, d_next = 0;
            bool     flushed = !p.has_ffn;
            uint32_t a_ptr = tabA_s, b_ptr = tabB_s, c_ptr = tabC_s;
            int      s = 0, q = 0, qi = 0, b_done = 0;
            bool     seen_final = false;
            unsigned eph = 0;  // parity of empty[s] to wait for once the ring has wrapped
            // 1: an item was issued, 0: nothing can be issued right now, -1: END was issued
            auto put = [&](bool prewait) -> int {
                int      ty = T_NONE, name = 7;
                uint32_t pt = 0;
                if (a_left > 0) {
                    ty = T_DOT, name = 1, pt = a_ptr;
                    --a_left, a_ptr += RPI * 16u;
                } else if (prewait) {
                    return 0;
                } else {
                    if (b_left < 0 && mbar_test(listbar, 0)) {  // my share of the active-row list is complete
                        b_left = ctl[5] - b_done;
                    }
                    if (!flushed && (b_left >= 0 || b_done > 0)) {
                        const int final_rows = ld_acquire_cta_s32(ctl + 1);  // > 0: all gate/up items retired; rows incl. padding + 1
                        const int rows       = final_rows > 0 ? final_rows - 1 : ld_acquire_cta_s32(ctl + 0);
                        if (TRACE && final_rows > 0 && !seen_final) {
                            seen_final = true;
                            CLK_STAMP(502);
                        }
                        if ((d_next + 1) * RPI <= rows) {
                            ty = T_AXPY, name = 3, pt = tabD_s + (uint32_t) d_next * (RPI * 16u);
                            ++d_next;
                        } else if (final_rows > 0) {  // every W_down^T row is issued: the partial y follows them
                            ty = T_FLUSH, name = 5;
                            flushed = true;
                        }
                    }
                    if (ty == T_NONE) {
                        // gate/up rows: the first few (my own block of the list) are there one round trip after the wait
                        const bool b_ready = b_left > 0 || (b_left < 0 && (b_done + 1) * RPI <= ld_acquire_cta_s32(ctl + 6));
                        if (b_ready) {
                            ty = T_DOT, name = 2, pt = b_ptr;
                            b_ptr += RPI * 16u;
                            ++b_done;
                            if (b_left > 0) {
                                --b_left;
                            }
                        } else if (b_left == 0 && c_left > 0) {
                            // Wpd rows.  While the last hidden values are still with the epilogue warp only three unconsumed items
                            // may sit in the ring: the W_down^T rows must not queue behind a ring full of Wpd rows (the partial
                            // y has to reach the other CTAs while Wpd rows are still left to cover the wait)
                            if (!flushed && q >= 3) {
                                const int qq = q - 3, sl = qq % S;
                                if (!mbar_test(&empty[sl], (unsigned) (qq / S) & 1u)) {
                                    return 0;
                                }
                            }
                            ty = T_DOTH, name = 4, pt = c_ptr;
                            --c_left, c_ptr += RPI * 16u;
                        } else if (b_left == 0 && flushed) {
                            ty = T_END, name = 6;
                        } else {
                            return 0;  // the row list or the last hidden values are not there yet
                        }
                    }
                }
                if (q >= S) {
                    mbar_wait_spin(&empty[s], eph);
                }
                if (TRACE && ty != T_END) {
                    ITEM_STAMP(qi, 0, (unsigned long long) name);
                    ITEM_STAMP(qi, 1, 0);
                    ++qi;
                }
                uint4 e   = make_uint4(0u, 0u, 0u, 0u);
                bool  cut = false;
                if (ty <= T_AXPY && lane < RPI) {
                    e   = lds128(pt + (uint32_t) lane * 16u);
                    cut = e.z != row_bytes;
                    if (e.z > 0) {  // before the arrival below: the phase cannot complete under the copies
                        asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(e.z) : "memory");
                    }
                }
                const unsigned any_cut = __ballot_sync(0xffffffffu, cut);
                if (lane == 0) {
                    sdesc[s] = make_int4(ty, (int) pt, any_cut != 0u, name);
                    mbar_arrive(&full[s]);  // release: the consumers' wait acquires the descriptor
                }
                if (e.z > 0) {
                    unsigned char * dst = ring + (size_t) s * slot_bytes + (size_t) lane * row_bytes;
                    const void *    src = reinterpret_cast<const void *>(((unsigned long long) e.y << 32) | e.x);
                    if (hint) {
                        tma_load_1d(dst, src, e.z, &full[s], pol);
                    } else {
                        tma_load_1d_nohint(dst, src, e.z, &full[s]);
                    }
                }
                ++q;
                if (++s == S) {
                    s = 0;
                    eph ^= (q > S) ? 1u : 0u;
                }
                return ty == T_END ? -1 : 1;
            };
            // ---- before the programmatic-dependent-launch wait: Wpu rows are weights only, they are on their way already
            while (q < S && put(true) > 0) {
            }
            pdl_wait_prior_grid();
            pdl_launch_dependents();
            for (;;) {
                const int rc = put(false);
                if (rc < 0) {
                    break;
                }
                if (rc == 0 && (flushed || b_left != 0)) {
                    __nanosleep(40);  // (no sleep while only the last hidden values are missing: the W_down^T rows are the critical path)
                }
            }
        }
    } else {
        // ================================================== helper warps ==================================================
        const int ht = tid - (NW + 1) * 32;  // 0 .. 63
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq = spif_comm_call_seq(p.comm);
        // ---- my share of this layer's active rows, in two stages so that the first gate/up rows are on their way one memory
        // round trip after the wait above instead of three:
        //   1. the first T1 active rows of MY OWN block of the list (block c was written by CTA c of the previous launch):
        //      its count and its first rows are requested together, nothing else is needed;
        //   2. the rest -- every block's rows beyond its first T1 -- is split evenly: ranks [c*A'/P, (c+1)*A'/P) of the
        //      concatenated remainders (all counts, one warp scan, one more round trip for the rows).
        // The order of a CTA's neurons (own first rows, then its share of the rest, both ascending) is fixed by the list
        // alone: the sums stay bit-stable.
        if (p.has_ffn && ht >= 32) {
            constexpr int T1 = 4;
            const int nb = p.lin_nblk, kb = (nb + 31) / 32;  // kb <= 8 (checked by the launcher: nblk <= 256); nb == gridDim.x
            const int b0 = min(lane * kb, nb), b1 = min(b0 + kb, nb);
            const int off_c = share(c, p.m, nb), cap_c = share(c + 1, p.m, nb) - off_c;
            const int cnt_c = __ldcg(p.lin_cnt + c);
            const int row1  = (lane < T1 && lane < cap_c) ? __ldcg(p.lin_rows + off_c + lane) : 0;
            int       cb[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                cb[j] = (b0 + j < b1) ? __ldcg(p.lin_cnt + b0 + j) : 0;
            }
            const unsigned long long a0 = (unsigned long long) p.w_up;  // a valid address for rows that are not there
            auto entry = [&](int j, int row) {  // neuron j of this CTA's list
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
            };
            const int k1 = min(cnt_c, T1);
            if (lane < k1) {
                entry(lane, row1);
            }
            __syncwarp();
            if (lane == 0) {
                st_release_cta_s32(ctl + 6, gated ? 2 * k1 : k1);  // table rows the producer may issue so far
            }
            LAYER_STAMP(2);
            int local = 0, local_all = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                local_all += cb[j];
                cb[j] = max(cb[j] - T1, 0);
                local += cb[j];
            }
            int incl = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                incl += (lane >= o) ? t : 0;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                local_all += __shfl_xor_sync(0xffffffffu, local_all, o);
            }
            const int total_rest = __shfl_sync(0xffffffffu, incl, 31), total_active = local_all;
            int       pbk        = incl - local;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (b0 + j < b1) {
                    blk_pre[b0 + j] = pbk;
                }
                pbk += cb[j];
            }
            if (lane == 31) {
                blk_pre[nb] = total_rest;
            }
            __syncwarp();
            const int lo = share(c, total_rest, P), hi = share(c + 1, total_rest, P);
            const int b_end = k1 + hi - lo;
            for (int j = lane; j < hi - lo; j += 32) {
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
                entry(k1 + j, __ldcg(p.lin_rows + share(bl, p.m, nb) + T1 + (rank - blk_pre[bl])));
            }
            const int n_rows_b = gated ? 2 * b_end : b_end, n_gu_items = (n_rows_b + RPI - 1) / RPI;
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
            if (TRACE && p.trace) {
                p.trace[(size_t) blockIdx.x * 512 + 504] = (unsigned long long) clock64();
            }
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
