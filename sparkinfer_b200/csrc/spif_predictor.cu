// spif_predictor.cu -- the per-layer activation predictor  p = sigmoid(Wpd . relu(Wpu . x + bpu) + bpd).
//
// Reference: llm_graph_context::build_predictor (src/llama-graph.cpp:865-894) = six stock ggml nodes
// (MUL_MAT, ADD, RELU, MUL_MAT, ADD, SIGMOID; on CUDA bs=1: mul_mat_vec_f mmvf.cu:6, unary.cu:44-50).
// Here: two kernels.
//   up   : rows of Wpu are n_embd long -> the cooperative K-slice GEMV of spif_ffn.cu (SPIF_MODE_GEMV, dense list,
//          bias + relu epilogue), x in registers, rows streamed by TMA.
//   down : rows of Wpd are short (r elements, 2 KB at r = 1024) -> this file: chunks of 8 consecutive rows are one
//          contiguous byte range, streamed by TMA bulk copies into a shared-memory ring; one warp per row, the hidden
//          vector in registers, bias + sigmoid fused, logits optionally kept for flip accounting.
#include "spif_common.cuh"

#define PD_ROWS   8   /* rows per chunk == consumer warps */
#define PD_STAGES 6

struct PredDownParams {
    const unsigned char * wpd;   // [n_ff][r]
    const float *         bpd;   // nullable [n_ff]
    const float *         h1;    // [n_tok][r]
    float *               out;   // [n_tok][n_ff] sigmoid
    float *               logits;// nullable [n_tok][n_ff]
    int                   n_ff, r, row_bytes, flags;
};

template <typename WT, int VL>
__global__ void __launch_bounds__((PD_ROWS + 1) * 32, 1) spif_pred_down_kernel(const PredDownParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int P = gridDim.x, c = blockIdx.x, tok = blockIdx.y;
    const int chunk_bytes = PD_ROWS * p.row_bytes;

    unsigned char * ring  = smem;
    uint64_t *      full  = reinterpret_cast<uint64_t *>(ring + (size_t) PD_STAGES * chunk_bytes);
    uint64_t *      empty = full + PD_STAGES;

    if (tid == 0) {
        for (int s = 0; s < PD_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], PD_ROWS);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const int n_chunks = (p.n_ff + PD_ROWS - 1) / PD_ROWS;
    const int my_n     = (n_chunks > c) ? (n_chunks - c + P - 1) / P : 0;

    if (warp == PD_ROWS) {
        // producer: weights do not depend on the previous kernel -> start streaming before the PDL wait
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            for (int i = 0; i < my_n; ++i) {
                const int s = i % PD_STAGES;
                if (i >= PD_STAGES) {
                    mbar_wait(&empty[s], ((i / PD_STAGES) - 1) & 1);
                }
                const int      ch    = c + i * P;
                const int      rows  = min(PD_ROWS, p.n_ff - ch * PD_ROWS);
                const uint32_t bytes = (uint32_t) rows * p.row_bytes;
                mbar_arrive_expect_tx(&full[s], bytes);
                tma_load_1d(ring + (size_t) s * chunk_bytes, p.wpd + (size_t) ch * chunk_bytes, bytes, &full[s], pol);
            }
        }
        return;
    }

    pdl_wait_prior_grid();  // h1 comes from the up kernel
    pdl_launch_dependents();

    // hidden vector slice in registers: vectors v = lane + 32*k
    const int     Vr = p.r >> 3;
    const float * h  = p.h1 + (size_t) tok * p.r;
    float         hr[VL * 8];
#pragma unroll
    for (int k = 0; k < VL; ++k) {
        const int v = lane + 32 * k;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            hr[k * 8 + j] = 0.f;
        }
        if (v < Vr) {
            const float4 a = __ldg(reinterpret_cast<const float4 *>(h) + 2 * v);
            const float4 b = __ldg(reinterpret_cast<const float4 *>(h) + 2 * v + 1);
            hr[k * 8 + 0] = a.x, hr[k * 8 + 1] = a.y, hr[k * 8 + 2] = a.z, hr[k * 8 + 3] = a.w;
            hr[k * 8 + 4] = b.x, hr[k * 8 + 5] = b.y, hr[k * 8 + 6] = b.z, hr[k * 8 + 7] = b.w;
            if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    hr[k * 8 + j] = round_through<WT>(hr[k * 8 + j]);
                }
            }
        }
    }

    for (int i = 0; i < my_n; ++i) {
        const int s = i % PD_STAGES;
        mbar_wait(&full[s], (i / PD_STAGES) & 1);
        const int ch  = c + i * P;
        const int row = ch * PD_ROWS + warp;
        float     z   = 0.f;
        if (row < p.n_ff) {
            const uint4 * wr = reinterpret_cast<const uint4 *>(ring + (size_t) s * chunk_bytes + (size_t) warp * p.row_bytes);
#pragma unroll
            for (int k = 0; k < VL; ++k) {
                const int v = lane + 32 * k;
                if (v < Vr) {
                    float w[8];
                    unpack8<WT>(wr[v], w);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        z = fmaf(w[e], hr[k * 8 + e], z);
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(&empty[s]);
        }
        z = warp_sum(z);
        if (lane == 0 && row < p.n_ff) {
            if (p.bpd) {
                z += __ldg(p.bpd + row);
            }
            if (p.logits) {
                p.logits[(size_t) tok * p.n_ff + row] = z;
            }
            p.out[(size_t) tok * p.n_ff + row] = 1.f / (1.f + expf(-z));  // unary-ops.cpp:31-33 / unary.cu:48-50
        }
    }
}

template <typename WT, int VL>
static int launch_pd(const PredDownParams & p, dim3 grid, size_t smem, cudaStream_t stream, int device) {
    auto                      kern = spif_pred_down_kernel<WT, VL>;
    static unsigned long long done = 0;
    int                       rc0  = spif_set_max_smem_once(kern, smem, device, done);
    if (rc0) {
        return rc0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = grid;
    cfg.blockDim           = dim3((PD_ROWS + 1) * 32);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, p));
    return 0;
}

template <typename WT> static int launch_pd_v(int vl, const PredDownParams & p, dim3 grid, size_t smem, cudaStream_t stream, int device) {
    switch (vl) {
        case 1:
            return launch_pd<WT, 1>(p, grid, smem, stream, device);
        case 2:
            return launch_pd<WT, 2>(p, grid, smem, stream, device);
        case 4:
            return launch_pd<WT, 4>(p, grid, smem, stream, device);
        default:
            return launch_pd<WT, 8>(p, grid, smem, stream, device);
    }
}

int spif_launch_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                          const float * x, float * sparse_idx_out, float * logits_out, int n_embd, int r, int n_ff,
                          int n_tok, int flags, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(ptype == SPIF_TYPE_F16 || ptype == SPIF_TYPE_BF16, SPIF_E_TYPE);
    SPIF_REQUIRE(n_embd > 0 && n_embd % 8 == 0 && r > 0 && r % 8 == 0 && r <= 2048 && n_ff > 0 && n_tok > 0, SPIF_E_SHAPE);
    // the hidden vector lives after the FFN's partials region so a predictor may run concurrently with an FFN
    // kernel that shares the workspace (look-ahead overlap, llama-graph.cpp:939-946)
    const size_t h1_off   = spif_ws_partials_bytes(n_embd, n_tok);
    const size_t h1_bytes = sizeof(float) * (size_t) n_tok * r;
    SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0 && ws_bytes >= h1_off + h1_bytes, SPIF_E_WORKSPACE);
    float * h1 = (float *) ((char *) ws + h1_off);

    // up: dense GEMV + bias + relu through the FFN kernel template
    spif_ffn_args a = {};
    a.arch          = SPIF_ARCH_RELU_UP;
    a.type          = ptype;
    a.n_embd        = n_embd;
    a.n_ff          = r;
    a.m             = r;
    a.n_tok         = n_tok;
    a.flags         = (flags & SPIF_FLAG_CPU_ROUNDING) | (1 << 30);  // bit 30: dense list + bias/relu epilogue
    a.thr           = 0.5f;
    a.w_up          = wpu;
    a.b_up          = bpu;
    a.x             = x;
    a.sparse_idx    = nullptr;
    int rc          = spif_launch_ffn(a, SPIF_MODE_GEMV, nullptr, h1, nullptr, 0, ctx);
    if (rc) {
        return rc;
    }

    PredDownParams p = {};
    p.wpd            = (const unsigned char *) wpd;
    p.bpd            = bpd;
    p.h1             = h1;
    p.out            = sparse_idx_out;
    p.logits         = logits_out;
    p.n_ff           = n_ff;
    p.r              = r;
    p.row_bytes      = r * 2;
    p.flags          = flags;
    const int    Vr  = r / 8;
    const int    vl  = Vr <= 32 ? 1 : (Vr <= 64 ? 2 : (Vr <= 128 ? 4 : 8));
    const size_t smem = (size_t) PD_STAGES * PD_ROWS * p.row_bytes + 16 * PD_STAGES + 64;
    const dim3   grid(ctx.sm_count, n_tok);
    return (ptype == SPIF_TYPE_F16) ? launch_pd_v<spif_f16>(vl, p, grid, smem, ctx.stream, ctx.device) :
                                      launch_pd_v<spif_bf16>(vl, p, grid, smem, ctx.stream, ctx.device);
}
