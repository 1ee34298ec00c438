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
#include "spif_ffn_kernel.cuh"
#include "spif_ffn_direct.cuh"
#include "spif_comm.cuh"

unsigned long long * g_spif_trace = nullptr;  // debug only; see tools/trace_ffn.py
extern "C" void spif_debug_set_trace(void * dev_ptr) {
    g_spif_trace = (unsigned long long *) dev_ptr;
}

// ---------------------------------------------------------------------------------------------------------------
// y[t][col] = sum_p partials[t][p][col] (+ b_down[col]) in a fixed order.  Block = 8 column-vectors (128 B of every
// partial row) x 32 partial lanes: thread (cx, py) loads partials py, py+32, ... (<= SPIF_RED_PER, all in flight at
// once -> one L2 round trip), then a fixed-shape two-level sum through shared memory.  n_embd/32 CTAs (160 at 13B).
#define SPIF_RED_COLS  8
#define SPIF_RED_LANES 32
#define SPIF_RED_PER   10 /* ceil(P / 32) for P <= 320 CTAs */
// COMM: the block's 32 finished columns are not written to y but pushed into every shard's exchange buffer; the same block
// then waits for the peers' flags and writes the sum over shards (spif_comm.cuh) -- partial sum + allreduce in ONE kernel.
template <bool COMM>
__global__ void __launch_bounds__(SPIF_RED_COLS * SPIF_RED_LANES) spif_reduce_partials_kernel(
    const float * __restrict__ partials, const float * __restrict__ b_down, float * __restrict__ y,
    const int32_t * __restrict__ cta_stats, int32_t * __restrict__ stats_out, int n_embd, int P, const spif_comm_dev comm,
    const int launch_first) {
    __shared__ float4 sm[SPIF_RED_LANES][SPIF_RED_COLS];
    __shared__ float4 sm2[4][SPIF_RED_COLS];
    const int         tok  = blockIdx.y;
    const int         cx   = threadIdx.x & (SPIF_RED_COLS - 1);
    const int         py   = threadIdx.x / SPIF_RED_COLS;
    const int         col4 = blockIdx.x * SPIF_RED_COLS + cx;
    const int         n4   = n_embd >> 2;
    if (launch_first) {
        // the next layer's kernel may start (mask scan, first weight rows) while this grid still waits for the partials;
        // it touches x / y / the workspace only after ITS griddepcontrol.wait, i.e. after this grid has completed
        pdl_launch_dependents();
    }
    pdl_wait_prior_grid();
    if (!launch_first) {
        pdl_launch_dependents();
    }
    const uint32_t seq = COMM ? spif_comm_call_seq(comm) : 0u;  // after the wait: the previous call on this comm is complete
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
    float4     t   = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool own = (py == 0 && col4 < n4);
    if (own) {
        t = sm2[0][cx];
#pragma unroll
        for (int k = 1; k < 4; ++k) {
            const float4 v = sm2[k][cx];
            t.x += v.x, t.y += v.y, t.z += v.z, t.w += v.w;
        }
        if (b_down) {
            const float4 bb = reinterpret_cast<const float4 *>(b_down)[col4];
            t.x += bb.x, t.y += bb.y, t.z += bb.z, t.w += bb.w;
        }
    }
    if (COMM) {
        t = spif_peer_allreduce_f4(comm, seq, t, tok * n4 + col4, own);
    }
    if (own) {
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
    if (COMM) {
        __syncthreads();
        if (threadIdx.x == 0) {
            spif_comm_block_done(comm, seq, gridDim.x * gridDim.y);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host-side launcher

template <typename C, int MODE>
static int launch_t(const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    return spif_ffn_launch_kernel<C, MODE>(kp, grid, threads, smem, stream, device);
}

template <typename C>
static int launch_m(int mode, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    switch (mode) {
        case SPIF_MODE_FUSED:
            return launch_t<C, SPIF_MODE_FUSED>(kp, grid, threads, smem, stream, device);
        case SPIF_MODE_GEMV:
            return launch_t<C, SPIF_MODE_GEMV>(kp, grid, threads, smem, stream, device);
        default:
            return launch_t<C, SPIF_MODE_AXPY>(kp, grid, threads, smem, stream, device);
    }
}

template <typename WT>
static int launch_v(int mode, int vpt, const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream, int device) {
    switch (vpt) {
        case 1:
            return launch_m<FloatCodec<WT, 1>>(mode, kp, grid, threads, smem, stream, device);
        case 2:
            return launch_m<FloatCodec<WT, 2>>(mode, kp, grid, threads, smem, stream, device);
        default:
            return launch_m<FloatCodec<WT, 4>>(mode, kp, grid, threads, smem, stream, device);
    }
}

static int launch_float(int type, int mode, int vpt, const FfnKParams & kp, dim3 grid, int threads, size_t smem,
                        cudaStream_t stream, int device) {
    return (type == SPIF_TYPE_F16) ? launch_v<spif_f16>(mode, vpt, kp, grid, threads, smem, stream, device) :
                                     launch_v<spif_bf16>(mode, vpt, kp, grid, threads, smem, stream, device);
}

static int launch_reduce(const FfnKParams & kp, const float * b_down, float * y, int32_t * stats_out, int n_embd, int n_tok,
                         int P, cudaStream_t stream, const spif_shard_comm * comm = nullptr) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3((n_embd / 4 + SPIF_RED_COLS - 1) / SPIF_RED_COLS, n_tok);
    cfg.blockDim           = dim3(SPIF_RED_COLS * SPIF_RED_LANES);
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    static const int k_red_first = [] { const char * e = getenv("SPIF_RED_FIRST"); return e ? atoi(e) : 1; }();
    spif_comm_dev cd = {};
    if (comm) {
        int rc = spif_comm_to_dev(comm, cd);
        if (rc) {
            return rc;
        }
        SPIF_REQUIRE((int64_t) n_embd * n_tok <= comm->cap_floats, SPIF_E_WORKSPACE);
        SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_reduce_partials_kernel<true>, (const float *) kp.partials, b_down, y,
                                         (const int32_t *) kp.cta_stats, stats_out, n_embd, P, cd, k_red_first));
        return 0;
    }
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_reduce_partials_kernel<false>, (const float *) kp.partials, b_down, y,
                                     (const int32_t *) kp.cta_stats, stats_out, n_embd, P, cd, k_red_first));
    return 0;
}

int spif_launch_reduce_partials(const FfnKParams & kp, const float * b_down, float * y, int32_t * stats_out, int n_embd, int n_tok,
                                int P, cudaStream_t stream, const spif_shard_comm * comm) {
    return launch_reduce(kp, b_down, y, stats_out, n_embd, n_tok, P, stream, comm);
}

// workspace layout shared with spif_api.cu: [partials n_tok*P*n_embd f32][cta_stats n_tok*P*2 i32]
// `row_bytes` / `nc` (consumer threads) / `variant` come from the weight format; `launch` instantiates the kernel.
int spif_launch_ffn_common(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                           size_t ws_bytes, const spif_launch_ctx & ctx, int row_bytes, int nc, int variant,
                           spif_ffn_launch_fn launch) {
    SPIF_REQUIRE(a.n_embd > 0 && a.n_embd % 8 == 0 && a.n_ff > 0 && a.m > 0 && a.n_tok > 0, SPIF_E_SHAPE);
    SPIF_REQUIRE(a.n_embd <= 16384 && row_bytes % 16 == 0, SPIF_E_SHAPE);
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
    kp.row_bytes  = row_bytes;
    kp.thr        = a.thr;
    kp.arch       = a.arch;
    kp.flags      = a.flags;
    kp.dense_out  = (mode == SPIF_MODE_GEMV) ? dst_dense : a.hidden_out;
    kp.trace      = g_spif_trace;
    kp.magic      = (unsigned int) ((1ull << 32) / (unsigned) P) + 1u;
    SPIF_REQUIRE(a.sparse_idx != nullptr, SPIF_E_NULL);
    kp.zero_fill  = (mode == SPIF_MODE_GEMV && a.neuron_idx == nullptr && !(a.flags & SPIF_FLAG_MASK_BY_ROW)) ? 1 : 0;

    kp.nc = nc;
    SPIF_REQUIRE(kp.nc >= 32 && kp.nc <= 512 && kp.nc % 32 == 0, SPIF_E_SHAPE);
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
        static const int k_gu_pct = [] { const char * e = getenv("SPIF_GU_PCT"); int v = e ? atoi(e) : 45; return v < 10 ? 10 : (v > 90 ? 90 : v); }();
        S = (int) ((ring_budget * k_gu_pct / 100) / stage_b);
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
    int        rc = launch(a.type, mode, variant, kp, grid, threads, smem, ctx.stream, ctx.device);
    if (rc) {
        return rc;
    }
    if (mode != SPIF_MODE_GEMV) {
        return launch_reduce(kp, (mode == SPIF_MODE_FUSED) ? a.b_down : (const float *) nullptr,
                             (mode == SPIF_MODE_AXPY) ? dst_dense : a.y, a.stats_out, a.n_embd, a.n_tok, P, ctx.stream,
                             (mode == SPIF_MODE_FUSED) ? ctx.comm : nullptr);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// register-streaming fused layer (spif_ffn_direct.cuh)
template <typename WT, int VPT, int NB, int MAXT>
static int launch_direct_t(const FfnKParams & kp, dim3 grid, int threads, size_t smem, cudaStream_t stream) {
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
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_ffn_direct_kernel<WT, VPT, NB, MAXT>, kp));
    return 0;
}

// returns SPIF_E_UNSUPPORTED when the shape does not qualify (the ring kernel handles it then)
static int spif_launch_ffn_direct(const spif_ffn_args & a, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx) {
    const int V = a.n_embd / 8;
    SPIF_REQUIRE(V <= 1024, SPIF_E_UNSUPPORTED);
    // half-size CTAs (<= 320 threads x 2 vectors, 41 K registers) leave room for a co-resident predictor CTA of the next
    // layer (spif_launch_predictor caps its CTAs to match); SPIF_DIRECT_SMALL=0 restores the 640-thread shape
    static const int k_small = [] { const char * e = getenv("SPIF_DIRECT_SMALL"); return e ? atoi(e) : 1; }();
    const bool small   = k_small && V <= 640 && V > 320;
    const int  vpt     = (V <= 640 && !small) ? 1 : 2;
    const int  threads = ((V + vpt - 1) / vpt + 31) / 32 * 32;
    SPIF_REQUIRE(a.m <= 4096 * (threads / 32), SPIF_E_UNSUPPORTED);  // activity bits kept in registers by build_my_list
    // SPIF_FFN_CTAS < SM count leaves whole SMs to the look-ahead predictor grid (experiment knob; 0 = one CTA per SM)
    static const int k_ffn_ctas = [] { const char * e = getenv("SPIF_FFN_CTAS"); return e ? atoi(e) : 0; }();
    const int        P          = (k_ffn_ctas > 0 && k_ffn_ctas < ctx.sm_count) ? k_ffn_ctas : ctx.sm_count;
    SPIF_REQUIRE(P >= 1 && P <= SPIF_RED_LANES * SPIF_RED_PER, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(a.neuron_idx != nullptr || a.m == a.n_ff, SPIF_E_SHAPE);
    SPIF_REQUIRE(a.sparse_idx != nullptr, SPIF_E_NULL);
    FfnKParams kp = {};
    kp.w_up       = (const unsigned char *) a.w_up;
    kp.w_gate     = (const unsigned char *) a.w_gate;
    kp.w_down     = (const unsigned char *) a.w_down_t;
    kp.b_up       = a.b_up;
    kp.b_gate     = a.b_gate;
    kp.x          = a.x;
    kp.mask       = a.sparse_idx;
    kp.neuron_idx = a.neuron_idx;
    kp.n_embd     = a.n_embd;
    kp.n_ff       = a.n_ff;
    kp.m          = a.m;
    kp.mask_len   = (a.flags & SPIF_FLAG_MASK_BY_ROW) ? a.m : a.n_ff;
    kp.V          = V;
    kp.row_bytes  = a.n_embd * 2;
    kp.thr        = a.thr;
    kp.arch       = a.arch;
    static const int k_pdl_late = [] { const char * e = getenv("SPIF_PDL_LATE"); return e ? atoi(e) : 0; }();
    kp.flags      = a.flags | (k_pdl_late ? SPIF_KFLAG_PDL_LATE : 0);
    kp.dense_out  = a.hidden_out;
    kp.trace      = g_spif_trace;
    kp.magic      = (unsigned int) ((1ull << 32) / (unsigned) P) + 1u;
    kp.list_cap   = (a.m + P - 1) / P + 1;
    const size_t part_bytes = sizeof(float) * (size_t) a.n_tok * P * a.n_embd;
    const size_t stat_bytes = sizeof(int32_t) * (size_t) a.n_tok * P * 2;
    SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0 && ws_bytes >= part_bytes + stat_bytes, SPIF_E_WORKSPACE);
    kp.partials  = (float *) ws;
    kp.cta_stats = (int32_t *) ((char *) ws + part_bytes);
    const size_t smem = sizeof(float) * 2 * 2 * 2 * 32 + sizeof(int) * 36 + (size_t) 12 * kp.list_cap + 64;
    SPIF_REQUIRE(smem <= 48 * 1024, SPIF_E_UNSUPPORTED);
    if (a.hidden_out) {
        SPIF_CUDA_TRY(cudaMemsetAsync(a.hidden_out, 0, sizeof(float) * (size_t) a.n_ff * a.n_tok, ctx.stream));
    }
    const dim3 grid(P, a.n_tok);
    const bool f16 = a.type == SPIF_TYPE_F16;
    int        rc;
    if (vpt == 1) {
        rc = f16 ? launch_direct_t<spif_f16, 1, 2, 640>(kp, grid, threads, smem, ctx.stream) :
                   launch_direct_t<spif_bf16, 1, 2, 640>(kp, grid, threads, smem, ctx.stream);
    } else if (small) {
        rc = f16 ? launch_direct_t<spif_f16, 2, 1, 320>(kp, grid, threads, smem, ctx.stream) :
                   launch_direct_t<spif_bf16, 2, 1, 320>(kp, grid, threads, smem, ctx.stream);
    } else {
        rc = f16 ? launch_direct_t<spif_f16, 2, 1, 512>(kp, grid, threads, smem, ctx.stream) :
                   launch_direct_t<spif_bf16, 2, 1, 512>(kp, grid, threads, smem, ctx.stream);
    }
    if (rc) {
        return rc;
    }
    return launch_reduce(kp, a.b_down, a.y, a.stats_out, a.n_embd, a.n_tok, P, ctx.stream, ctx.comm);
}

// F16 / BF16 entry: thread shape = VPT 16-byte vectors per consumer thread
int spif_launch_ffn(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws, size_t ws_bytes,
                    const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(a.type == SPIF_TYPE_F16 || a.type == SPIF_TYPE_BF16, SPIF_E_TYPE);
    SPIF_REQUIRE(a.n_embd > 0 && a.n_embd % 8 == 0, SPIF_E_SHAPE);
    static const int k_direct = [] { const char * e = getenv("SPIF_DIRECT"); return e ? atoi(e) : 1; }();
    if (mode == SPIF_MODE_FUSED && k_direct) {
        const int rc = spif_launch_ffn_direct(a, ws, ws_bytes, ctx);
        if (rc != SPIF_E_UNSUPPORTED) {
            return rc;
        }
    }
    const int V   = a.n_embd / 8;
    const int vpt = (V <= 128) ? 1 : (V <= 768 ? 2 : 4);
    const int nc  = ((V + vpt - 1) / vpt + 31) / 32 * 32;
    return spif_launch_ffn_common(a, mode, hid_in, dst_dense, ws, ws_bytes, ctx, a.n_embd * 2, nc, vpt, launch_float);
}
