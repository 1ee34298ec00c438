// spif_api.cu -- the extern "C" boundary declared in include/spif_b200.h: argument checks (the conditions the
// reference asserts, returned as codes instead of aborting), workspace carving, dispatch to the kernels.
// No CPU fallback anywhere: without a device every compute entry point fails.
#include "spif_common.cuh"

#include <cstring>

int spif_launch_ffn_q(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                      size_t ws_bytes, const spif_launch_ctx & ctx);
int    spif_launch_ffn_batch(const spif_ffn_args & a, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx);
size_t spif_batch_workspace_bytes(int n_embd, int n_ff, int n_tok);
int    spif_launch_shard_allreduce(float * y, int64_t n, const spif_shard_comm * comm, cudaStream_t stream);
size_t spif_active_list_bytes_impl(int m);
int    spif_launch_mask_to_list(const float * mask, int m, float thr, void * list, const spif_launch_ctx & ctx);
int    spif_launch_layer(const spif_ffn_args * ffn, const spif_pred_args * pred, const spif_layer_io * io, const void * list_in, void * list_out,
                         const spif_shard_comm * comm, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx);

static int make_ctx(spif_launch_ctx & ctx, spif_stream_t stream) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        return SPIF_E_NO_DEVICE;
    }
    // device attributes are immutable: cache them per device (benign race: every writer stores the same values)
    static int cached_sm[64], cached_smem[64];
    const int  slot = dev & 63;
    if (cached_sm[slot] == 0) {
        int sm = 0, smem = 0;
        SPIF_CUDA_TRY(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
        SPIF_CUDA_TRY(cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        cached_smem[slot] = smem;
        cached_sm[slot]   = sm;
    }
    ctx.device         = dev;
    ctx.sm_count       = cached_sm[slot];
    ctx.max_smem_optin = cached_smem[slot];
    ctx.stream         = (cudaStream_t) stream;
    ctx.comm           = nullptr;
    return 0;
}

static bool is_float_type(int t) {
    return t == SPIF_TYPE_F16 || t == SPIF_TYPE_BF16;
}
static bool is_quant_type(int t) {
    return t == SPIF_TYPE_Q8_0 || t == SPIF_TYPE_Q4_0;
}
static bool aligned16(const void * p) {
    return ((uintptr_t) p & 15) == 0;
}

extern "C" {

int spif_abi_version(void) {
    return SPIF_B200_ABI_VERSION;
}

const char * spif_error_string(int code) {
    switch (code) {
        case SPIF_OK:
            return "ok";
        case SPIF_E_TYPE:
            return "unsupported weight type for this op";
        case SPIF_E_SHAPE:
            return "shape / alignment precondition violated";
        case SPIF_E_NULL:
            return "required pointer is NULL";
        case SPIF_E_WORKSPACE:
            return "workspace missing, misaligned or too small";
        case SPIF_E_UNSUPPORTED:
            return "configuration not supported";
        case SPIF_E_NO_DEVICE:
            return "no CUDA device";
        case SPIF_E_STATIC_LIMIT:
            return "shape needs more scratch than the library's static array (SPIF_NOALLOC_WS_BYTES): use spif_sparse_ffn with a workspace";
        default:
            return code > 0 ? cudaGetErrorString((cudaError_t) code) : "unknown spif error";
    }
}

int spif_device_sm_count(int * sm_count) {
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, nullptr);
    if (rc) {
        return rc;
    }
    if (sm_count) {
        *sm_count = ctx.sm_count;
    }
    return 0;
}

size_t spif_row_size(int type, int64_t n) {
    switch (type) {
        case SPIF_TYPE_F32:
            return (size_t) n * 4;
        case SPIF_TYPE_F16:
        case SPIF_TYPE_BF16:
            return (size_t) n * 2;
        case SPIF_TYPE_Q8_0:
            return (size_t) (n / SPIF_QK) * 34;
        case SPIF_TYPE_Q4_0:
            return (size_t) (n / SPIF_QK) * 18;
        default:
            return 0;
    }
}

// partials for up to 320 CTAs (2 per SM) per token + per-CTA stats + predictor hidden + Q8_1-style activations; rounded up
size_t spif_workspace_size(int n_embd, int n_ff, int n_tok) {
    if (n_embd <= 0 || n_ff <= 0 || n_tok <= 0) {
        return 0;
    }
    const size_t P    = 320;
    size_t       need = sizeof(float) * (size_t) n_tok * P * n_embd;  // partial y per CTA
    need += sizeof(int32_t) * (size_t) n_tok * P * 2;                 // stats
    need += sizeof(float) * (size_t) n_tok * 2048;                    // predictor hidden (r <= 2048)
    need += (size_t) n_tok * ((size_t) n_embd / 32 * 40 + 64);        // quantised activations
    need += (size_t) n_tok * (size_t) n_ff * 4;                       // spare dense vector
    const size_t batch = spif_batch_workspace_bytes(n_embd, n_ff, n_tok);  // union-of-masks GEMM path (n_tok >= 2)
    need = need > batch ? need : batch;
    return (need + 4095) & ~(size_t) 4095;
}

int spif_mask_compact(const float * sparse_idx, const int32_t * neuron_idx, int m, int n_ff, int n_tok, float thr,
                      int flags, int32_t * idx_out, int32_t * count_out, int32_t * union_idx, int32_t * union_count,
                      spif_stream_t stream) {
    SPIF_REQUIRE(sparse_idx && idx_out && count_out, SPIF_E_NULL);
    SPIF_REQUIRE(m > 0 && n_ff > 0 && n_tok > 0 && (neuron_idx || m == n_ff), SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_mask_compact(sparse_idx, neuron_idx, m, n_ff, n_tok, thr, flags, idx_out, count_out, union_idx,
                                    union_count, ctx.stream);
}

int spif_topk_mask(const float * score, int n_ff, int n_tok, int k, float * sparse_idx_out, spif_stream_t stream) {
    SPIF_REQUIRE(score && sparse_idx_out, SPIF_E_NULL);
    SPIF_REQUIRE(n_ff > 0 && n_tok > 0, SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_topk_mask(score, n_ff, n_tok, k, sparse_idx_out, ctx.stream);
}

static int check_common(int type, const void * w, const float * mask, int n_embd, int n_ff, int m, int n_tok,
                        const int32_t * neuron_idx) {
    SPIF_REQUIRE(is_float_type(type) || is_quant_type(type), SPIF_E_TYPE);  // reference: GGML_ABORT unsupported type
    SPIF_REQUIRE(w != nullptr, SPIF_E_NULL);
    SPIF_REQUIRE(mask != nullptr, SPIF_E_NULL);                             // reference: "missing sparse_idx"
    SPIF_REQUIRE(n_embd > 0 && n_ff > 0 && m > 0 && n_tok > 0, SPIF_E_SHAPE);
    SPIF_REQUIRE(n_embd % 32 == 0, SPIF_E_SHAPE);                           // mm-sparse.cu:228, mmq-sparse.cu:177
    SPIF_REQUIRE(neuron_idx != nullptr || m == n_ff, SPIF_E_SHAPE);
    SPIF_REQUIRE(m <= n_ff, SPIF_E_SHAPE);
    SPIF_REQUIRE(aligned16(w), SPIF_E_SHAPE);
    return 0;
}

int spif_mul_mat_sparse(int type, const void * w, const float * x, const float * sparse_idx,
                        const int32_t * neuron_idx, float * dst, int n_embd, int n_ff, int m, int n_tok, float thr,
                        int flags, void * ws, size_t ws_bytes, spif_stream_t stream) {
    int rc = check_common(type, w, sparse_idx, n_embd, n_ff, m, n_tok, neuron_idx);
    if (rc) {
        return rc;
    }
    SPIF_REQUIRE(x && dst, SPIF_E_NULL);
    SPIF_REQUIRE(aligned16(x), SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    spif_ffn_args a = {};
    a.arch          = SPIF_ARCH_RELU_UP;
    a.type          = type;
    a.n_embd        = n_embd;
    a.n_ff          = n_ff;
    a.m             = m;
    a.n_tok         = n_tok;
    a.flags         = flags & (SPIF_FLAG_CPU_ROUNDING | SPIF_FLAG_MASK_BY_ROW);
    a.thr           = thr;
    a.w_up          = w;
    a.x             = x;
    a.sparse_idx    = sparse_idx;
    a.neuron_idx    = neuron_idx;
    return is_float_type(type) ? spif_launch_ffn(a, SPIF_MODE_GEMV, nullptr, dst, ws, ws_bytes, ctx) :
                                 spif_launch_ffn_q(a, SPIF_MODE_GEMV, nullptr, dst, ws, ws_bytes, ctx);
}

int spif_axpy_sparse(int type, const void * wt, const float * hid, const float * sparse_idx,
                     const int32_t * neuron_idx, float * dst, int n_embd, int n_ff, int m, int n_tok, float thr,
                     int flags, void * ws, size_t ws_bytes, spif_stream_t stream) {
    int rc = check_common(type, wt, sparse_idx, n_embd, n_ff, m, n_tok, neuron_idx);
    if (rc) {
        return rc;
    }
    SPIF_REQUIRE(hid && dst, SPIF_E_NULL);
    SPIF_REQUIRE(aligned16(dst), SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    spif_ffn_args a = {};
    a.arch          = SPIF_ARCH_RELU_UP;
    a.type          = type;
    a.n_embd        = n_embd;
    a.n_ff          = n_ff;
    a.m             = m;
    a.n_tok         = n_tok;
    a.flags         = flags & (SPIF_FLAG_CPU_ROUNDING | SPIF_FLAG_MASK_BY_ROW);
    a.thr           = thr;
    a.w_down_t      = wt;
    a.sparse_idx    = sparse_idx;
    a.neuron_idx    = neuron_idx;
    return is_float_type(type) ? spif_launch_ffn(a, SPIF_MODE_AXPY, hid, dst, ws, ws_bytes, ctx) :
                                 spif_launch_ffn_q(a, SPIF_MODE_AXPY, hid, dst, ws, ws_bytes, ctx);
}

int spif_fatrelu(const float * x, float * y, int64_t n, float threshold, spif_stream_t stream) {
    SPIF_REQUIRE(n == 0 || (x && y), SPIF_E_NULL);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_unary(0, x, y, n, threshold, ctx.stream);
}

int spif_shifted_step(const float * x, float * y, int64_t n, float threshold, spif_stream_t stream) {
    SPIF_REQUIRE(n == 0 || (x && y), SPIF_E_NULL);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_unary(1, x, y, n, threshold, ctx.stream);
}

int spif_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                   const float * x, float * sparse_idx_out, float * logits_out, int n_embd, int r, int n_ff, int n_tok,
                   int flags, void * ws, size_t ws_bytes, spif_stream_t stream) {
    SPIF_REQUIRE(wpu && wpd && x && sparse_idx_out, SPIF_E_NULL);
    SPIF_REQUIRE(aligned16(wpu) && aligned16(wpd) && aligned16(x), SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_predictor(ptype, wpu, bpu, wpd, bpd, x, sparse_idx_out, logits_out, n_embd, r, n_ff, n_tok, flags,
                                 ws, ws_bytes, ctx);
}

static int sparse_ffn_impl(const spif_ffn_args * args, const spif_shard_comm * comm, void * ws, size_t ws_bytes,
                           spif_stream_t stream) {
    SPIF_REQUIRE(args != nullptr, SPIF_E_NULL);
    const spif_ffn_args & a = *args;
    int rc = check_common(a.type, a.w_up, a.sparse_idx, a.n_embd, a.n_ff, a.m, a.n_tok, a.neuron_idx);
    if (rc) {
        return rc;
    }
    SPIF_REQUIRE(a.arch >= SPIF_ARCH_PROSPARSE && a.arch <= SPIF_ARCH_SILU_GATED, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(a.w_down_t && a.x && a.y, SPIF_E_NULL);
    SPIF_REQUIRE(a.arch == SPIF_ARCH_RELU_UP || a.w_gate != nullptr, SPIF_E_NULL);  // "FFN_GATE is not found"
    SPIF_REQUIRE(aligned16(a.w_down_t) && aligned16(a.x) && aligned16(a.y) && (a.w_gate == nullptr || aligned16(a.w_gate)),
                 SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    // quantised weights at small batches: the warp-per-neuron kernel with one CTA row per token re-streams a row per token that
    // needs it, but at <= 12 tokens that is still less time than dequantising the union of the rows into F16 operand tiles
    // (13B Q4_0, T = 8: 51 us against 92 us; T = 32: 194 against 124 -- profiles/r2_bench_n1_default.json `batched`)
    const bool quant_small = !is_float_type(a.type) && a.n_tok <= 12;
    if (a.n_tok >= 2 && !(a.flags & SPIF_FLAG_PER_TOKEN) && !quant_small) {
        // batched / prefill: union of the per-token masks -> gathered GEMM on tensor cores, each row read once
        rc = spif_launch_ffn_batch(a, ws, ws_bytes, ctx);
        if (rc == 0 && comm) {
            rc = spif_launch_shard_allreduce(a.y, (int64_t) a.n_embd * a.n_tok, comm, ctx.stream);
        }
        if (rc != SPIF_E_UNSUPPORTED) {
            return rc;
        }
    }
    ctx.comm = comm;  // the kernel that sums the per-CTA partials also sums over the shards
    return is_float_type(a.type) ? spif_launch_ffn(a, SPIF_MODE_FUSED, nullptr, nullptr, ws, ws_bytes, ctx) :
                                   spif_launch_ffn_q(a, SPIF_MODE_FUSED, nullptr, nullptr, ws, ws_bytes, ctx);
}

int spif_sparse_ffn(const spif_ffn_args * args, void * ws, size_t ws_bytes, spif_stream_t stream) {
    return sparse_ffn_impl(args, nullptr, ws, ws_bytes, stream);
}

int spif_sparse_ffn_sharded(const spif_ffn_args * args, const spif_shard_comm * comm, void * ws, size_t ws_bytes,
                            spif_stream_t stream) {
    SPIF_REQUIRE(comm != nullptr, SPIF_E_NULL);
    return sparse_ffn_impl(args, comm, ws, ws_bytes, stream);
}

size_t spif_shard_comm_data_bytes(int world, int cap_floats) {
    return (world < 1 || cap_floats < 1) ? 0 : 8 * 2 * (size_t) world * (size_t) cap_floats;
}

int spif_shard_allreduce(float * y, int64_t n, const spif_shard_comm * comm, spif_stream_t stream) {
    SPIF_REQUIRE(comm != nullptr, SPIF_E_NULL);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_shard_allreduce(y, n, comm, ctx.stream);
}

int spif_shard_comm_status(const spif_shard_comm * comm, int * error_out, spif_stream_t stream) {
    SPIF_REQUIRE(comm != nullptr && comm->seq != nullptr && error_out != nullptr, SPIF_E_NULL);
    SPIF_CUDA_TRY(cudaStreamSynchronize((cudaStream_t) stream));
    uint32_t w = 0;
    SPIF_CUDA_TRY(cudaMemcpy(&w, (const uint32_t *) comm->seq + 2, sizeof(w), cudaMemcpyDeviceToHost));
    *error_out = (int) w;
    return 0;
}

size_t spif_active_list_bytes(int m) {
    return spif_active_list_bytes_impl(m);
}

int spif_mask_to_list(const float * sparse_idx, int m, float thr, void * list_out, spif_stream_t stream) {
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_mask_to_list(sparse_idx, m, thr, list_out, ctx);
}

int spif_sparse_ffn_lookahead(const spif_ffn_args * ffn, const spif_pred_args * pred, const void * list_in, void * list_out,
                              const spif_shard_comm * comm, void * ws, size_t ws_bytes, spif_stream_t stream) {
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_layer(ffn, pred, nullptr, list_in, list_out, comm, ws, ws_bytes, ctx);
}

int spif_sparse_ffn_lookahead_ex(const spif_ffn_args * ffn, const spif_pred_args * pred, const spif_layer_io * io,
                                 const void * list_in, void * list_out, const spif_shard_comm * comm, void * ws, size_t ws_bytes,
                                 spif_stream_t stream) {
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_layer(ffn, pred, io, list_in, list_out, comm, ws, ws_bytes, ctx);
}

int spif_dfr_update(const float * sparse_idx, int n_ff, int n_tok, int g, float thr, float lambda, int ema,
                    float * dfr_scores, spif_stream_t stream) {
    SPIF_REQUIRE(sparse_idx && dfr_scores, SPIF_E_NULL);
    SPIF_REQUIRE(n_ff > 0 && n_tok > 0 && g > 0 && n_ff % g == 0, SPIF_E_SHAPE);
    spif_launch_ctx ctx;
    int             rc = make_ctx(ctx, stream);
    if (rc) {
        return rc;
    }
    return spif_launch_dfr_update(sparse_idx, n_ff, n_tok, g, thr, lambda, ema, dfr_scores, ctx.stream);
}

}  // extern "C"
