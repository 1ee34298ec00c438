// spif_predictor.cu -- the per-layer activation predictor  p = sigmoid(Wpd . relu(Wpu . x + bpu) + bpd).
//
// Reference: llm_graph_context::build_predictor (src/llama-graph.cpp:865-894) = six stock ggml nodes
// (MUL_MAT, ADD, RELU, MUL_MAT, ADD, SIGMOID; on CUDA bs=1: mul_mat_vec_f mmvf.cu:6, unary.cu:44-50).
// Here: two launches of ONE dense streaming GEMV kernel with fused epilogues
//   up   : h1 = relu(Wpu . x + bpu)                 rows = r,    K = n_embd
//   down : z = Wpd . h1 + bpd ; p = sigmoid(z)      rows = n_ff, K = r
//
// Both matrices are read exactly once per token and nothing about their addresses depends on data, so the kernel is
// a pure HBM stream: every warp owns a fixed K-slice (its slice of the input vector stays in registers) and walks the
// rows of its CTA's contiguous row block with U rows x VL 128-bit loads in flight per lane (ld.global.nc, L1 no-allocate).
// No shared-memory staging: the kernel is deliberately light on registers and uses (almost) no shared memory so that it
// can be co-resident on an SM with the TMA-ring sparse-FFN kernel of the previous layer (the look-ahead predictor of
// layer l+1 only depends on x_l, llama-graph.cpp:939-946, so the two run on parallel graph branches).
// The first batch of weight loads is issued BEFORE griddepcontrol.wait: weights never depend on the previous kernel.
#include "spif_common.cuh"

#define DG_MAX_WARPS 12
#define DG_U         2 /* rows in flight per warp */
#define DG_DEPTH     4 /* ASYNC variant: batches of DG_U rows in flight per warp */

// ASYNC variant (spif_dense_gemv_kernel<..., true>): the same walk, but the rows travel through a thread-PRIVATE shared-memory
// ring filled by 16-byte cp.async (LDGSTS): DG_DEPTH batches are in flight per lane and hold no registers while they fly.
// Every lane reads back only the bytes it requested itself, so cp.async.wait_group is the only synchronisation.  Purpose:
// next to a resident sparse-FFN CTA only ONE small predictor CTA fits by registers (profiles/r1d_coresidency.txt); with
// register-held loads that CTA keeps 24 KB in flight, with the ring 4x as much -- shared memory is what the FFN kernel
// leaves free.
__device__ __forceinline__ void dg_cp_async16(uint32_t dst, const void * src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void dg_cp_async_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void dg_cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

enum { DG_EPI_RELU = 0, DG_EPI_SIGMOID = 1 };

struct DenseGemvParams {
    const unsigned char * w;       // [rows][K] weight type
    const float *         bias;    // nullable [rows]
    const float *         x;       // [n_tok][K]
    float *               out;     // [n_tok][rows]
    float *               logits;  // nullable [n_tok][rows] (pre-activation, SIGMOID epilogue only)
    int                   rows, K, V;  // V = K / 8 vectors per row
    int                   ns, nrw;     // K-slices per row, row lanes per CTA
    int                   rpc;         // rows per CTA (contiguous block)
    int                   flags;
};

template <typename WT, int VL, int EPI, bool ASYNC>
__global__ void __launch_bounds__(DG_MAX_WARPS * 32, 2) spif_dense_gemv_kernel(const DenseGemvParams p) {
    extern __shared__ __align__(16) float part[];  // [rpc][ns] partial dots (only when ns > 1), then the ASYNC ring
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tok  = blockIdx.y;
    const int sl   = warp % p.ns;  // my K-slice
    const int rl   = warp / p.ns;  // my row lane
    const int r0   = blockIdx.x * p.rpc;
    const int r1   = min(r0 + p.rpc, p.rows);
    const int v0   = sl * (VL * 32);  // first vector of my slice
    const size_t row_bytes = (size_t) p.K * 2;

    // vector k of lane: v0 + lane + 32 k   (coalesced 512 B per warp instruction)
    uint4 wv[DG_U][VL];
    // ASYNC ring: slot d, row u, vector k of thread t at uint4 index ((d * DG_U + u) * VL + k) * blockDim.x + t
    const int      NT       = blockDim.x;
    uint4 *        ring     = reinterpret_cast<uint4 *>(part + ((p.ns > 1 ? p.rpc * p.ns : 0) + 3) / 4 * 4);
    const uint32_t ring_s   = (uint32_t) __cvta_generic_to_shared(ring);
    auto           load_rows = [&](int rbase, int slot) {
#pragma unroll
        for (int u = 0; u < DG_U; ++u) {
            const int row = min(rbase + u * p.nrw, r1 - 1);
            const unsigned char * rp = p.w + (size_t) row * row_bytes;
#pragma unroll
            for (int k = 0; k < VL; ++k) {
                const int v = min(v0 + lane + 32 * k, p.V - 1);
                if (ASYNC) {
                    dg_cp_async16(ring_s + (uint32_t) ((((slot * DG_U + u) * VL + k) * NT + (int) threadIdx.x) * 16),
                                  rp + (size_t) v * 16);
                } else {
                    wv[u][k] = ldg_stream_v4(rp + (size_t) v * 16);
                }
            }
        }
    };
    auto take_rows = [&](int slot) {  // ASYNC: my own bytes of the oldest batch -> registers
#pragma unroll
        for (int u = 0; u < DG_U; ++u) {
#pragma unroll
            for (int k = 0; k < VL; ++k) {
                wv[u][k] = ring[((slot * DG_U + u) * VL + k) * NT + threadIdx.x];
            }
        }
    };

    const bool have_work = (rl < p.nrw) && (r0 + rl < r1);
    const int  step      = DG_U * p.nrw;
    if (have_work) {
        if (ASYNC) {
#pragma unroll
            for (int d = 0; d < DG_DEPTH; ++d) {  // one commit group per batch, empty ones included: uniform counting
                if (r0 + rl + d * step < r1) {
                    load_rows(r0 + rl + d * step, d);
                }
                dg_cp_async_commit();
            }
        } else {
            load_rows(r0 + rl, 0);  // in flight across the PDL wait
        }
    }
    pdl_wait_prior_grid();  // x (h1) comes from the previous kernel
    pdl_launch_dependents();

    float xr[VL * 8];
    {
        const float * xt = p.x + (size_t) tok * p.K;
#pragma unroll
        for (int k = 0; k < VL; ++k) {
            const int v = v0 + lane + 32 * k;
            if (v < p.V) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(xt) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(xt) + 2 * v + 1);
                xr[k * 8 + 0] = a.x, xr[k * 8 + 1] = a.y, xr[k * 8 + 2] = a.z, xr[k * 8 + 3] = a.w;
                xr[k * 8 + 4] = b.x, xr[k * 8 + 5] = b.y, xr[k * 8 + 6] = b.z, xr[k * 8 + 7] = b.w;
                if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        xr[k * 8 + j] = round_through<WT>(xr[k * 8 + j]);
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xr[k * 8 + j] = 0.f;  // clamped (duplicate) vectors past the end contribute nothing
                }
            }
        }
    }

    auto finish = [&](int row, float z) {
        if (p.bias) {
            z += __ldg(p.bias + row);
        }
        if (EPI == DG_EPI_RELU) {
            p.out[(size_t) tok * p.rows + row] = (z > 0.f) ? z : 0.f;  // llama-graph.cpp:871-880
        } else {
            if (p.logits) {
                p.logits[(size_t) tok * p.rows + row] = z;
            }
            p.out[(size_t) tok * p.rows + row] = 1.f / (1.f + expf(-z));  // unary-ops.cpp:31-33 / unary.cu:48-50
        }
    };

    if (have_work) {
        int slot = 0;
        for (int rb = r0 + rl; rb < r1; rb += step) {
            if (ASYNC) {
                dg_cp_async_wait<DG_DEPTH - 1>();
                take_rows(slot);
            }
            float acc[DG_U][2];
#pragma unroll
            for (int u = 0; u < DG_U; ++u) {
                acc[u][0] = 0.f, acc[u][1] = 0.f;
#pragma unroll
                for (int k = 0; k < VL; ++k) {
                    float w[8];
                    unpack8<WT>(wv[u][k], w);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        acc[u][e & 1] = fmaf(w[e], xr[k * 8 + e], acc[u][e & 1]);
                    }
                }
            }
            if (ASYNC) {
                if (rb + DG_DEPTH * step < r1) {
                    load_rows(rb + DG_DEPTH * step, slot);  // the slot just read is free again
                }
                dg_cp_async_commit();
                slot = (slot + 1 == DG_DEPTH) ? 0 : slot + 1;
            } else if (rb + step < r1) {
                load_rows(rb + step, 0);  // next batch in flight while this one is reduced
            }
#pragma unroll
            for (int u = 0; u < DG_U; ++u) {
                const int   row = rb + u * p.nrw;
                const float z   = warp_sum(acc[u][0] + acc[u][1]);
                if (lane == 0 && row < r1) {
                    if (p.ns == 1) {
                        finish(row, z);
                    } else {
                        part[(row - r0) * p.ns + sl] = z;
                    }
                }
            }
        }
    }
    if (p.ns > 1) {
        __syncthreads();
        for (int i = threadIdx.x; i < r1 - r0; i += blockDim.x) {
            float z = 0.f;
            for (int s = 0; s < p.ns; ++s) {  // fixed order -> run-to-run bit-stable
                z += part[i * p.ns + s];
            }
            finish(r0 + i, z);
        }
    }
}

template <typename WT, int VL, int EPI, bool ASYNC>
static int launch_dg(const DenseGemvParams & p, dim3 grid, int threads, size_t smem, cudaStream_t stream) {
    if (smem > 48 * 1024) {
        static bool done = false;  // per instantiation
        if (!done) {
            SPIF_CUDA_TRY(cudaFuncSetAttribute(spif_dense_gemv_kernel<WT, VL, EPI, ASYNC>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            done = true;
        }
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
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_dense_gemv_kernel<WT, VL, EPI, ASYNC>, p));
    return 0;
}

template <typename WT, int EPI>
static int launch_dg_v(int vl, bool async, const DenseGemvParams & p, dim3 grid, int threads, size_t smem, cudaStream_t stream) {
    if (async) {
        switch (vl) {
            case 1:
                return launch_dg<WT, 1, EPI, true>(p, grid, threads, smem, stream);
            case 2:
                return launch_dg<WT, 2, EPI, true>(p, grid, threads, smem, stream);
            default:
                return launch_dg<WT, 4, EPI, true>(p, grid, threads, smem, stream);
        }
    }
    switch (vl) {
        case 1:
            return launch_dg<WT, 1, EPI, false>(p, grid, threads, smem, stream);
        case 2:
            return launch_dg<WT, 2, EPI, false>(p, grid, threads, smem, stream);
        default:
            return launch_dg<WT, 4, EPI, false>(p, grid, threads, smem, stream);
    }
}

// out[t][row] = epi(W[row] . x[t] + bias[row])
static int dense_gemv(int type, int epi, const void * w, const float * bias, const float * x, float * out, float * logits,
                      int rows, int K, int n_tok, int flags, int warp_cap, const spif_launch_ctx & ctx) {
    DenseGemvParams p = {};
    p.w      = (const unsigned char *) w;
    p.bias   = bias;
    p.x      = x;
    p.out    = out;
    p.logits = logits;
    p.rows   = rows;
    p.K      = K;
    p.V      = K / 8;
    p.flags  = flags;
    // slice width: VL*32 vectors of 8 elements per warp pass
    static const int k_vl = [] { const char * e = getenv("SPIF_DG_VL"); return e ? atoi(e) : 4; }();
    int              vl   = (p.V >= 128) ? 4 : (p.V >= 64 ? 2 : 1);
    vl                    = vl > k_vl ? k_vl : vl;
    p.ns = (p.V + vl * 32 - 1) / (vl * 32);
    SPIF_REQUIRE(p.ns <= DG_MAX_WARPS, SPIF_E_SHAPE);  // K <= 12 * 1024 elements
    p.nrw             = DG_MAX_WARPS / p.ns;
    // warp_cap: smaller CTAs that fit next to a resident FFN CTA (see spif_predictor); SPIF_DG_WARPS overrides, 0 = no cap
    static const int k_warps = [] { const char * e = getenv("SPIF_DG_WARPS"); return e ? atoi(e) : -1; }();
    const int        wcap    = k_warps >= 0 ? k_warps : warp_cap;
    if (wcap > 0) {
        const int cap = wcap / p.ns < 1 ? 1 : wcap / p.ns;
        p.nrw         = p.nrw > cap ? cap : p.nrw;
    }
    const int threads = p.ns * p.nrw * 32;
    int       per_sm  = (2 * DG_MAX_WARPS * 32) / threads;  // CTAs that fit the 2 x 384-thread register budget
    // cp.async ring variant (see the top of the file): SPIF_DG_ASYNC=1
    static const int k_async = [] { const char * e = getenv("SPIF_DG_ASYNC"); return e ? atoi(e) : 0; }();
    const bool   async      = k_async != 0;
    const size_t ring_bytes = async ? (size_t) DG_DEPTH * DG_U * vl * 16 * threads : 0;
    if (async) {
        const int by_smem = (int) ((200 * 1024) / (ring_bytes + 2048));
        per_sm            = per_sm > by_smem ? by_smem : per_sm;
    }
    const int G       = ctx.sm_count * (per_sm < 1 ? 1 : per_sm);
    p.rpc             = (rows + G - 1) / G;
    // keep every row lane busy: at least nrw rows per CTA when there are enough rows
    if (p.rpc < p.nrw) {
        p.rpc = p.nrw;
    }
    const int    grid_x    = (rows + p.rpc - 1) / p.rpc;
    const size_t part_b    = (p.ns > 1) ? sizeof(float) * (size_t) ((p.rpc * p.ns + 3) / 4 * 4) : 0;
    SPIF_REQUIRE(part_b <= 48 * 1024, SPIF_E_SHAPE);
    const size_t smem = part_b + ring_bytes;
    SPIF_REQUIRE(smem <= 200 * 1024, SPIF_E_SHAPE);
    const dim3 grid(grid_x, n_tok);
    if (type == SPIF_TYPE_F16) {
        return epi == DG_EPI_RELU ? launch_dg_v<spif_f16, DG_EPI_RELU>(vl, async, p, grid, threads, smem, ctx.stream) :
                                    launch_dg_v<spif_f16, DG_EPI_SIGMOID>(vl, async, p, grid, threads, smem, ctx.stream);
    }
    return epi == DG_EPI_RELU ? launch_dg_v<spif_bf16, DG_EPI_RELU>(vl, async, p, grid, threads, smem, ctx.stream) :
                                launch_dg_v<spif_bf16, DG_EPI_SIGMOID>(vl, async, p, grid, threads, smem, ctx.stream);
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
    const int fl = flags & SPIF_FLAG_CPU_ROUNDING;
    // The predictor of layer l+1 runs next to the FFN of layer l (look-ahead).  For 320 < n_embd/8 <= 640 the FFN kernel runs
    // 320-thread CTAs that leave 24 K registers per SM (spif_launch_ffn_direct): 6-warp predictor CTAs fit beside one, 12-warp
    // ones would wait for the FFN to leave.  profiles/r1d_coresidency.txt: 13B decode 1034 -> 1110 tok/s.
    const int V8   = n_embd / 8;
    const int wcap = (n_tok == 1 && V8 > 320 && V8 <= 640) ? 6 : 0;
    int       rc = dense_gemv(ptype, DG_EPI_RELU, wpu, bpu, x, h1, nullptr, r, n_embd, n_tok, fl, wcap, ctx);
    if (rc) {
        return rc;
    }
    return dense_gemv(ptype, DG_EPI_SIGMOID, wpd, bpd, h1, sparse_idx_out, logits_out, n_ff, r, n_tok, fl, wcap, ctx);
}
