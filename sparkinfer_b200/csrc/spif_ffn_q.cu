// spif_ffn_q.cu -- Q8_0 / Q4_0 variants of the sparse-FFN hot path: dequantisation fused into the dot products.
//
// Reference: Q8_0 GPU path mul_mat_vec_sparse_q8_0_q8_1 + quantize_q8_1 (mmq-sparse.cu:12-163, quantize.cu:5-36) and
// mul_mat_axpy_sparse_rowwise_q (axpyq-sparse.cu:26-105); CPU path ggml_vec_dot_q8_0_q8_0 / q4_0_q8_0 on a Q8_0
// activation (ggml-cpu.c:1832-1856, arch/x86/quants.c:543-583,1011-1045) and ggml_axpy_avx_q8_0_alphaf32
// (ggml-cpu.c:2060-2146).  Q4_0 sparse ops are claimed by supports_op (ggml-cuda.cu:4594-4663) but never implemented
// on the GPU; the CPU backend has the Q4_0 dot only (the Q4_0 axpy aborts, ggml-cpu.c:2225-2226).
//
// Same persistent kernel as the float formats (spif_ffn_kernel.cuh: in-kernel mask scan, TMA row ring, lazy down ring,
// fixed-order partial sum); only the codec differs:
//   * a consumer thread owns one PAIR of 32-element blocks (68 B for Q8_0, 36 B for Q4_0: pairs are 4-byte aligned
//     although single blocks are not), reads it from the shared-memory ring with 32-bit loads and realigns the first
//     block's payload with funnel shifts -- instead of the reference's per-byte assembly (mmq-sparse.cu:20-25);
//   * x is quantised to 8 bit per 32-block ONCE per CTA, in registers, by the thread that owns the pair (the reference
//     launches a separate quantize kernel and re-reads its output from global memory in every CTA);
//   * gate/up dots are integer (dp4a) with d_w * d_x applied per block, exactly the reference arithmetic; the down
//     axpy is acc = fma(q, d * alpha, acc) in FP32 like axpyq-sparse.cu / ggml_axpy_avx_q8_0_alphaf32.
// Activation quantiser: default = the CUDA reference's (d = amax/127, q = roundf(x/d), quantize.cu:19-31);
// SPIF_FLAG_CPU_ROUNDING = the CPU backend's (id = 127/amax, round-half-even, arch/x86/quants.c:290-356).
#include "spif_ffn_qwarp.cuh"

template <typename QT> struct QuantCodec {
    using T                   = QuantTraits<QT>;
    static constexpr int NACC = 64;  // the two blocks of my pair
    struct XS {
        int   qx[16];  // my 64 activations, 8 bit
        float dx[2];   // per-block activation scale (rounded through F16 like block_q8_0.d / block_q8_1.ds)
        int   sx[2];   // per-block sum of qx (Q4_0's -8 offset)
        int   nb;      // blocks of my pair that exist (0, 1, 2)
    };
    __device__ static __forceinline__ float round_alpha(float h) { return h; }  // alpha stays F32, ggml-cpu.c:2215-2218

    __device__ static __forceinline__ void load_x(XS & xs, const float * xg, int ctid, int nc, int n_embd, int flags) {
        (void) nc;
        const int nblk = n_embd >> 5;
        xs.nb          = max(0, min(2, nblk - 2 * ctid));
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            xs.dx[b] = 0.f;
            xs.sx[b] = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                xs.qx[b * 8 + k] = 0;
            }
            if (b < xs.nb) {
                const float4 * x4 = reinterpret_cast<const float4 *>(xg + (size_t) (2 * ctid + b) * 32);
                float          v[32];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float4 t = __ldg(x4 + k);
                    v[4 * k] = t.x, v[4 * k + 1] = t.y, v[4 * k + 2] = t.z, v[4 * k + 3] = t.w;
                }
                float amax = 0.f;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    amax = fmaxf(amax, fabsf(v[j]));
                }
                const float d   = amax / 127.f;
                const bool  cpu = (flags & SPIF_FLAG_CPU_ROUNDING) != 0;
                const float id  = (amax != 0.0f) ? 127.f / amax : 0.0f;
                int         s   = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    int w = 0;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float xv = v[4 * k + e];
                        const int   q  = cpu ? __float2int_rn(xv * id) : ((amax == 0.0f) ? 0 : (int) roundf(xv / d));
                        s += q;
                        w |= (q & 0xff) << (8 * e);
                    }
                    xs.qx[b * 8 + k] = w;
                }
                xs.sx[b] = s;
                xs.dx[b] = __half2float(__float2half_rn(d));
            }
        }
    }

    // the pair's words from shared memory; first block's payload realigned
    __device__ static __forceinline__ void load_pair(const unsigned char * row, int ctid, unsigned (&w)[T::PAIR_WORDS]) {
        const unsigned * p = reinterpret_cast<const unsigned *>(row + (size_t) ctid * (2 * T::BLOCK_BYTES));
#pragma unroll
        for (int k = 0; k < T::PAIR_WORDS; ++k) {
            w[k] = p[k];
        }
    }

    __device__ static __forceinline__ float dot(const unsigned char * row, const XS & xs, int ctid, int nc, int n_embd) {
        (void) nc;
        (void) n_embd;
        if (xs.nb == 0) {
            return 0.f;
        }
        unsigned w[T::PAIR_WORDS];
        load_pair(row, ctid, w);
        float out = 0.f;
        if constexpr (T::BLOCK_BYTES == 34) {
            // [d0 | q0 x32 | d1 | q1 x32] = 68 B: q0 starts at byte 2 (funnel shift by 16), q1 at byte 36 (aligned)
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                s0 = __dp4a((int) __funnelshift_r(w[k], w[k + 1], 16), xs.qx[k], s0);
                s1 = __dp4a((int) w[9 + k], xs.qx[8 + k], s1);
            }
            out = half_bits_to_float(w[0]) * xs.dx[0] * (float) s0;
            if (xs.nb > 1) {
                out += half_bits_to_float(w[8] >> 16) * xs.dx[1] * (float) s1;
            }
        } else {
            // [d0 | n0 x16 | d1 | n1 x16] = 36 B; byte j of a block = elements j (low nibble) and j+16 (high nibble)
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned v0 = __funnelshift_r(w[k], w[k + 1], 16);
                const unsigned v1 = w[5 + k];
                s0 = __dp4a((int) (v0 & 0x0F0F0F0Fu), xs.qx[k], s0);
                s0 = __dp4a((int) ((v0 >> 4) & 0x0F0F0F0Fu), xs.qx[4 + k], s0);
                s1 = __dp4a((int) (v1 & 0x0F0F0F0Fu), xs.qx[8 + k], s1);
                s1 = __dp4a((int) ((v1 >> 4) & 0x0F0F0F0Fu), xs.qx[12 + k], s1);
            }
            out = half_bits_to_float(w[0]) * xs.dx[0] * (float) (s0 - 8 * xs.sx[0]);
            if (xs.nb > 1) {
                out += half_bits_to_float(w[4] >> 16) * xs.dx[1] * (float) (s1 - 8 * xs.sx[1]);
            }
        }
        return out;
    }

    __device__ static __forceinline__ void axpy(const unsigned char * row, float alpha, float (&acc)[NACC], int ctid, int nc,
                                                int n_embd) {
        (void) nc;
        const int nb = max(0, min(2, (n_embd >> 5) - 2 * ctid));
        if (nb == 0) {
            return;
        }
        unsigned w[T::PAIR_WORDS];
        load_pair(row, ctid, w);
        if constexpr (T::BLOCK_BYTES == 34) {
            const float sc0 = half_bits_to_float(w[0]) * alpha;                      // scale = d * alpha
            const float sc1 = (nb > 1) ? half_bits_to_float(w[8] >> 16) * alpha : 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const unsigned q0 = __funnelshift_r(w[k], w[k + 1], 16);
                const unsigned q1 = (nb > 1) ? w[9 + k] : 0u;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float f0      = (float) (int) (signed char) (q0 >> (8 * e));
                    const float f1      = (float) (int) (signed char) (q1 >> (8 * e));
                    acc[4 * k + e]      = fmaf(f0, sc0, acc[4 * k + e]);         // fma(q, scale, y)
                    acc[32 + 4 * k + e] = fmaf(f1, sc1, acc[32 + 4 * k + e]);
                }
            }
        } else {
            const float sc0 = half_bits_to_float(w[0]) * alpha;
            const float sc1 = (nb > 1) ? half_bits_to_float(w[4] >> 16) * alpha : 0.f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned v0 = __funnelshift_r(w[k], w[k + 1], 16);
                const unsigned v1 = (nb > 1) ? w[5 + k] : 0x88888888u;  // nibble 8 == value 0
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int b0 = (v0 >> (8 * e)) & 0xff, b1 = (v1 >> (8 * e)) & 0xff;
                    const int j  = 4 * k + e;
                    acc[j]           = fmaf((float) ((b0 & 0xF) - 8), sc0, acc[j]);
                    acc[j + 16]      = fmaf((float) ((b0 >> 4) - 8), sc0, acc[j + 16]);
                    acc[32 + j]      = fmaf((float) ((b1 & 0xF) - 8), sc1, acc[32 + j]);
                    acc[32 + j + 16] = fmaf((float) ((b1 >> 4) - 8), sc1, acc[32 + j + 16]);
                }
            }
        }
    }

    __device__ static __forceinline__ void store(float * part, const float (&acc)[NACC], int ctid, int nc, int n_embd) {
        (void) nc;
        const int nb = max(0, min(2, (n_embd >> 5) - 2 * ctid));
        float4 *  o  = reinterpret_cast<float4 *>(part + (size_t) ctid * 64);
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < nb * 8) {
                o[k] = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
            }
        }
    }
};

template <typename C>
static int launch_qm(int mode, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    switch (mode) {
        case SPIF_MODE_FUSED:
            return spif_ffn_launch_kernel<C, SPIF_MODE_FUSED>(kp, grid, threads, smem, stream, device);
        case SPIF_MODE_GEMV:
            return spif_ffn_launch_kernel<C, SPIF_MODE_GEMV>(kp, grid, threads, smem, stream, device);
        default:
            return spif_ffn_launch_kernel<C, SPIF_MODE_AXPY>(kp, grid, threads, smem, stream, device);
    }
}

static int launch_quant(int type, int mode, int variant, const FfnKParams & kp, dim3 grid, int threads, size_t smem,
                        cudaStream_t stream, int device) {
    (void) variant;
    return (type == SPIF_TYPE_Q8_0) ? launch_qm<QuantCodec<spif_q8_0>>(mode, kp, grid, threads, smem, stream, device) :
                                      launch_qm<QuantCodec<spif_q4_0>>(mode, kp, grid, threads, smem, stream, device);
}

// ---------------------------------------------------------------------------------------------------------------
// warp-per-neuron kernel (spif_ffn_qwarp.cuh): the default for every mode
template <typename QT, int IPT, int MAXT>
static int launch_qwarp_m(int mode, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    auto go = [&](auto kern, unsigned long long & done) -> int {
        const int rc0 = spif_set_max_smem_once(kern, smem, device, done);
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
    };
    static unsigned long long d0 = 0, d1 = 0, d2 = 0;
    switch (mode) {
        case SPIF_MODE_FUSED:
            return go(spif_ffn_qwarp_kernel<QT, SPIF_MODE_FUSED, IPT, MAXT>, d0);
        case SPIF_MODE_GEMV:
            return go(spif_ffn_qwarp_kernel<QT, SPIF_MODE_GEMV, IPT, MAXT>, d1);
        default:
            return go(spif_ffn_qwarp_kernel<QT, SPIF_MODE_AXPY, IPT, MAXT>, d2);
    }
}

// returns SPIF_E_UNSUPPORTED when the shape does not qualify (the ring kernel handles it then)
static int spif_launch_ffn_qwarp(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                                 size_t ws_bytes, const spif_launch_ctx & ctx) {
    const int nblk      = a.n_embd / 32;
    const int row_bytes = nblk * (a.type == SPIF_TYPE_Q8_0 ? 34 : 18);
    const int P         = ctx.sm_count;
    SPIF_REQUIRE(P >= 1 && P <= 320, SPIF_E_UNSUPPORTED);  // spif_reduce_partials_kernel: 32 lanes x 10 partials
    SPIF_REQUIRE(a.n_embd % 256 == 0 && a.n_embd <= 16384, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(a.neuron_idx != nullptr || a.m == a.n_ff, SPIF_E_SHAPE);
    SPIF_REQUIRE(a.sparse_idx != nullptr, SPIF_E_NULL);
    const int n_items = a.n_embd / 8;
    // SPIF_QW_HALF=1 (default): half-size CTAs (two column items per thread) whenever that still leaves >= 8 warps, so that
    // the next launch of a PDL chain fits on the SM beside this one (registers and shared memory) and scans its mask / requests
    // its first rows under this launch's tail
    static const int k_half = [] { const char * e = getenv("SPIF_QW_HALF"); return e ? atoi(e) : 1; }();
    int threads = (n_items + 31) / 32 * 32;
    if (k_half && threads >= 512) {
        threads = (n_items / 2 + 31) / 32 * 32;
    }
    threads = threads < 256 ? 256 : (threads > 1024 ? 1024 : threads);
    while (a.m > 4096 * (threads / 32) && threads < 1024) {  // activity bits kept in registers by build_my_list
        threads += 32;
    }
    SPIF_REQUIRE(a.m <= 4096 * (threads / 32), SPIF_E_UNSUPPORTED);
    const int ipt = (n_items + threads - 1) / threads;
    SPIF_REQUIRE(ipt <= 2, SPIF_E_UNSUPPORTED);

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
    kp.row_bytes  = row_bytes;
    kp.thr        = a.thr;
    kp.arch       = a.arch;
    kp.flags      = a.flags;
    kp.dense_out  = (mode == SPIF_MODE_GEMV) ? dst_dense : a.hidden_out;
    kp.trace      = g_spif_trace;
    kp.magic      = (unsigned int) ((1ull << 32) / (unsigned) P) + 1u;
    kp.zero_fill  = (mode == SPIF_MODE_GEMV && a.neuron_idx == nullptr && !(a.flags & SPIF_FLAG_MASK_BY_ROW)) ? 1 : 0;
    kp.list_cap   = (a.m + P - 1) / P + 1;
    kp.nc         = threads;
    if (mode == SPIF_MODE_AXPY) {  // the axpy-only op reads its rows from the down pointer
        kp.w_down = (const unsigned char *) a.w_down_t;
    }

    const bool gated         = (mode == SPIF_MODE_FUSED) && (a.arch != SPIF_ARCH_RELU_UP);
    const int  rows_per_slot = (mode == SPIF_MODE_FUSED) ? (gated ? 3 : 2) : 1;
    const int  slot_bytes    = rows_per_slot * row_bytes;
    // slots: one per warp at most; half of the SM's shared memory when that still gives 12 slots (the next launch of a PDL
    // chain can then become resident under this one's tail), else all of it
    const int nwarp = threads / 32;
    auto      fit   = [&](size_t budget) {
        int ns = nwarp;
        while (ns > 1 && qwarp_smem_layout(kp.list_cap, ns, a.n_embd, slot_bytes).total + 64 > budget) {
            --ns;
        }
        return ns;
    };
    const size_t full_b = (size_t) ctx.max_smem_optin - 1024;
    int          nslot  = fit(full_b / 2 - 1024);
    if (nslot < 12 && nslot < nwarp) {
        nslot = fit(full_b);
    }
    const QwarpSmem L = qwarp_smem_layout(kp.list_cap, nslot, a.n_embd, slot_bytes);
    SPIF_REQUIRE(L.total + 64 <= full_b, SPIF_E_UNSUPPORTED);
    kp.S              = nslot;
    const size_t smem = L.total + 64;  // + readable pad behind the last slot (odd block counts read one block past a row)

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
    int        rc;
#define SPIF_QW_GO(QT_)                                                                                          \
    (ipt == 2 ? (threads <= 320 ? launch_qwarp_m<QT_, 2, 320>(mode, kp, grid, threads, smem, ctx.stream, ctx.device) :  \
                 threads <= 512 ? launch_qwarp_m<QT_, 2, 512>(mode, kp, grid, threads, smem, ctx.stream, ctx.device) :  \
                                  launch_qwarp_m<QT_, 2, 1024>(mode, kp, grid, threads, smem, ctx.stream, ctx.device)) : \
     threads <= 768 ? launch_qwarp_m<QT_, 1, 768>(mode, kp, grid, threads, smem, ctx.stream, ctx.device) :       \
                      launch_qwarp_m<QT_, 1, 1024>(mode, kp, grid, threads, smem, ctx.stream, ctx.device))
    rc = (a.type == SPIF_TYPE_Q8_0) ? SPIF_QW_GO(spif_q8_0) : SPIF_QW_GO(spif_q4_0);
#undef SPIF_QW_GO
    if (rc) {
        return rc;
    }
    if (mode != SPIF_MODE_GEMV) {
        return spif_launch_reduce_partials(kp, (mode == SPIF_MODE_FUSED) ? a.b_down : (const float *) nullptr,
                                           (mode == SPIF_MODE_AXPY) ? dst_dense : a.y, a.stats_out, a.n_embd, a.n_tok, P,
                                           ctx.stream, (mode == SPIF_MODE_FUSED) ? ctx.comm : nullptr);
    }
    return 0;
}

int spif_launch_ffn_q(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                      size_t ws_bytes, const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(a.type == SPIF_TYPE_Q8_0 || a.type == SPIF_TYPE_Q4_0, SPIF_E_TYPE);
    // rows are streamed with TMA bulk copies: 16-byte multiples <=> n_embd % 256 == 0 for both block sizes (the
    // reference asserts % 32, mmq-sparse.cu:177; every architecture it supports has n_embd % 256 == 0)
    SPIF_REQUIRE(a.n_embd > 0 && a.n_embd % 256 == 0, SPIF_E_SHAPE);
    static const int k_ring = [] { const char * e = getenv("SPIF_Q_RING"); return e ? atoi(e) : 0; }();  // 1: round-1 ring kernel
    if (!k_ring) {
        const int rc = spif_launch_ffn_qwarp(a, mode, hid_in, dst_dense, ws, ws_bytes, ctx);
        if (rc != SPIF_E_UNSUPPORTED) {
            return rc;
        }
    }
    const int nblk      = a.n_embd / 32;
    const int row_bytes = nblk * (a.type == SPIF_TYPE_Q8_0 ? 34 : 18);
    const int npairs    = (nblk + 1) / 2;
    int       nc        = (npairs + 31) / 32 * 32;
    // at least 7 consumer warps: the whole CTA scans the mask (4096 rows per warp at most) and idle warps cost nothing
    const int need_warps = (a.m + 4095) / 4096;
    nc                   = nc < 224 ? 224 : nc;
    nc                   = nc < need_warps * 32 ? need_warps * 32 : nc;
    return spif_launch_ffn_common(a, mode, hid_in, dst_dense, ws, ws_bytes, ctx, row_bytes, nc, 0, launch_quant);
}
