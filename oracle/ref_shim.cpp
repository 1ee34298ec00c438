// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
//
// Drives the UNMODIFIED reference CPU implementation of the sparse-FFN path through the
// reference's own public graph API (ggml_mul_mat_sparse / ggml_axpy_sparse / ggml_fatrelu / ...,
// ggml/include/ggml.h:964-980,1159-1165,1443-1455) exactly the way
// llm_graph_context::build_predictor / build_sparse_ffn do (src/llama-graph.cpp:865-1142, GPU-less
// variant: neuron_mask == 0 everywhere => "all neurons on CPU").
//
// This file is ours; every ggml symbol it calls is compiled by oracle/Makefile from the sources
// where they lie under /root/reference.  It is linked into oracle/_ref/libspif_ref.so and used
// (a) to pin oracle/spif_oracle.c, (b) to generate tests/golden/, (c) as the "reference" arm of
// bench.py (the reference's CPU path timed on the box's host cores).  The product never loads it.
#include "ggml.h"
#include "ggml-cpu.h"
#include "ggml-sparkinfer.hpp"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

// ggml-backend.cpp:1511 references these two libllama symbols (layering cycle in the reference);
// they are only reached in hybrid GPU<->CPU scheduling, never from a plain CPU graph compute.
const bool sparkinfer_layer_cache::k_enable_spif_reload = false;
void       sparkinfer_layer_cache::sparkinfer_reload_plan() {}

namespace {

struct shim_ctx {
    ggml_context * ctx = nullptr;
    explicit shim_ctx(size_t bytes) {
        ggml_init_params ip = { bytes, nullptr, /*no_alloc=*/true };
        ctx                 = ggml_init(ip);
    }
    ~shim_ctx() {
        if (ctx) {
            ggml_free(ctx);
        }
    }
    // leaf tensor that borrows caller memory
    ggml_tensor * leaf2(int type, int64_t ne0, int64_t ne1, const void * data) {
        ggml_set_no_alloc(ctx, true);
        ggml_tensor * t = ggml_new_tensor_2d(ctx, (ggml_type) type, ne0, ne1);
        t->data         = const_cast<void *>(data);
        ggml_set_no_alloc(ctx, false);
        return t;
    }
    ggml_tensor * leaf1(int type, int64_t ne0, const void * data) {
        ggml_set_no_alloc(ctx, true);
        ggml_tensor * t = ggml_new_tensor_1d(ctx, (ggml_type) type, ne0);
        t->data         = const_cast<void *>(data);
        ggml_set_no_alloc(ctx, false);
        return t;
    }
};

size_t scratch_bytes(int64_t n_embd, int64_t n_ff, int64_t r, int64_t n_tok) {
    // result tensors (f32) + work buffers + graph/tensor overhead; generous
    size_t per_tok = (size_t) (8 * n_ff + 8 * n_embd + 4 * r) * sizeof(float);
    return per_tok * (size_t) n_tok + ((size_t) 64 << 20);
}

ggml_tensor * predictor(shim_ctx & s, ggml_tensor * x, ggml_tensor * wpu, ggml_tensor * bpu, ggml_tensor * wpd,
                        ggml_tensor * bpd) {
    // mirrors llm_graph_context::build_predictor, src/llama-graph.cpp:865-894
    ggml_tensor * cur = ggml_mul_mat(s.ctx, wpu, x);
    if (bpu) {
        cur = ggml_add(s.ctx, cur, bpu);
    }
    cur = ggml_relu(s.ctx, cur);
    cur = ggml_mul_mat(s.ctx, wpd, cur);
    if (bpd) {
        cur = ggml_add(s.ctx, cur, bpd);
    }
    return ggml_sigmoid(s.ctx, cur);
}

// arch: 0 prosparse-llama, 1 bamboo/qwen2, 2 opt/falcon     (src/llama-graph.cpp:1063-1094)
ggml_tensor * sparse_ffn(shim_ctx & s, int arch, ggml_tensor * x, ggml_tensor * sparse_idx, ggml_tensor * mask,
                         ggml_tensor * up, ggml_tensor * gate, ggml_tensor * down, ggml_tensor * up_b,
                         ggml_tensor * gate_b, ggml_tensor * down_b, ggml_tensor ** o_up, ggml_tensor ** o_gate,
                         ggml_tensor ** o_hid) {
    ggml_tensor * cur_up   = ggml_mul_mat_sparse(s.ctx, up, x, sparse_idx, mask);
    ggml_tensor * cur_gate = nullptr;
    if (arch != 2) {
        cur_gate = ggml_mul_mat_sparse(s.ctx, gate, x, sparse_idx, mask);
    }
    if (o_up) {
        *o_up = cur_up;
    }
    if (o_gate) {
        *o_gate = cur_gate;
    }
    if (up_b) {
        cur_up = ggml_add(s.ctx, cur_up, up_b);
    }
    if (cur_gate && gate_b) {
        cur_gate = ggml_add(s.ctx, cur_gate, gate_b);
    }
    ggml_tensor * hid = nullptr;
    switch (arch) {
        case 0:
            cur_gate = ggml_fatrelu(s.ctx, cur_gate, 0.01f, false);
            hid      = ggml_mul(s.ctx, cur_gate, cur_up);
            break;
        case 1:
            cur_up   = ggml_relu(s.ctx, cur_up);
            cur_gate = ggml_relu(s.ctx, cur_gate);
            hid      = ggml_mul(s.ctx, cur_gate, cur_up);
            break;
        default:
            hid = ggml_relu(s.ctx, cur_up);
            break;
    }
    if (o_hid) {
        *o_hid = hid;
    }
    ggml_tensor * y = ggml_axpy_sparse(s.ctx, down, hid, sparse_idx, mask);
    if (down_b) {
        y = ggml_add(s.ctx, y, down_b);
    }
    return y;
}

int run(shim_ctx & s, ggml_cgraph * gf, int n_threads) {
    return ggml_graph_compute_with_ctx(s.ctx, gf, n_threads) == GGML_STATUS_SUCCESS ? 0 : 1;
}

}  // namespace

extern "C" {

const char * spif_ref_describe(void) {
    return "SparkInfer reference CPU ops (ggml 0.9.4 fork), compiled from /root/reference by oracle/Makefile, "
           "x86-64-v3 (AVX2/FMA/F16C), OpenMP, SPIF_SPARSE_THRESHOLD=0.5";
}

size_t spif_ref_row_size(int type, int64_t n) {
    return ggml_row_size((ggml_type) type, n);
}

// ggml_quantize_chunk: F16/BF16 conversion or quantize_row_q{8,4}_0_ref (ggml-quants.c:199,36)
size_t spif_ref_quantize(int type, const float * src, void * dst, int64_t nrows, int64_t n_per_row) {
    return ggml_quantize_chunk((ggml_type) type, src, dst, 0, nrows, n_per_row, nullptr);
}

int spif_ref_dequantize(int type, const void * src, float * dst, int64_t n) {
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->to_float) {
        return 1;
    }
    tt->to_float(src, dst, n);
    return 0;
}

// GGML_OP_MUL_MAT_SPARSE on the CPU backend (ggml-cpu.c:1785-1925). neuron_mask: I32{n_ff}, 1 = "lives on GPU, skip".
int spif_ref_mul_mat_sparse(int wtype, const void * w, const float * x, const float * sparse_idx,
                            const int32_t * neuron_mask, float * dst, int n_embd, int n_ff, int n_tok, int n_threads) {
    shim_ctx      s(scratch_bytes(n_embd, n_ff, 0, n_tok));
    ggml_tensor * W = s.leaf2(wtype, n_embd, n_ff, w);
    ggml_tensor * X = s.leaf2(GGML_TYPE_F32, n_embd, n_tok, x);
    ggml_tensor * P = s.leaf2(GGML_TYPE_F32, n_ff, n_tok, sparse_idx);
    ggml_tensor * M = s.leaf1(GGML_TYPE_I32, n_ff, neuron_mask);
    ggml_tensor * R = ggml_mul_mat_sparse(s.ctx, W, X, P, M);
    ggml_cgraph * gf = ggml_new_graph(s.ctx);
    ggml_build_forward_expand(gf, R);
    if (run(s, gf, n_threads)) {
        return 1;
    }
    memcpy(dst, R->data, sizeof(float) * (size_t) n_ff * n_tok);
    return 0;
}

// GGML_OP_AXPY_SPARSE on the CPU backend (ggml-cpu.c:2230-2337). wt = W_down transposed {n_embd, n_ff}.
int spif_ref_axpy_sparse(int wtype, const void * wt, const float * hid, const float * sparse_idx,
                         const int32_t * neuron_mask, float * dst, int n_embd, int n_ff, int n_tok, int n_threads) {
    shim_ctx      s(scratch_bytes(n_embd, n_ff, 0, n_tok));
    ggml_tensor * W = s.leaf2(wtype, n_embd, n_ff, wt);
    ggml_tensor * H = s.leaf2(GGML_TYPE_F32, n_ff, n_tok, hid);
    ggml_tensor * P = s.leaf2(GGML_TYPE_F32, n_ff, n_tok, sparse_idx);
    ggml_tensor * M = s.leaf1(GGML_TYPE_I32, n_ff, neuron_mask);
    ggml_tensor * R = ggml_axpy_sparse(s.ctx, W, H, P, M);
    ggml_cgraph * gf = ggml_new_graph(s.ctx);
    ggml_build_forward_expand(gf, R);
    if (run(s, gf, n_threads)) {
        return 1;
    }
    memcpy(dst, R->data, sizeof(float) * (size_t) n_embd * n_tok);
    return 0;
}

// build_norm(x, w, LLM_NORM_RMS) = ggml_mul(ggml_rms_norm(x, eps), w)  (src/llama-graph.cpp build_norm; src/models/llama.cpp:98-101)
// and, with add != NULL, the residual ggml_add(y, add) of :144 applied to y_in instead of the FFN output: out2 = y_in + add.
int spif_ref_rms_norm_mul(const float * x, const float * w, float eps, int n_embd, int n_tok, float * out, int n_threads) {
    shim_ctx      s(scratch_bytes(n_embd, 0, 0, n_tok));
    ggml_tensor * X = s.leaf2(GGML_TYPE_F32, n_embd, n_tok, x);
    ggml_tensor * W = s.leaf1(GGML_TYPE_F32, n_embd, w);
    ggml_tensor * R = ggml_mul(s.ctx, ggml_rms_norm(s.ctx, X, eps), W);
    ggml_cgraph * gf = ggml_new_graph(s.ctx);
    ggml_build_forward_expand(gf, R);
    if (run(s, gf, n_threads)) {
        return 1;
    }
    memcpy(out, R->data, sizeof(float) * (size_t) n_embd * n_tok);
    return 0;
}

// build_predictor (src/llama-graph.cpp:865-894): sigmoid(Wpd . relu(Wpu . x + bpu) + bpd). Biases nullable.
int spif_ref_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                       const float * x, float * sparse_idx_out, int n_embd, int r, int n_ff, int n_tok,
                       int n_threads) {
    shim_ctx      s(scratch_bytes(n_embd, n_ff, r, n_tok));
    ggml_tensor * X   = s.leaf2(GGML_TYPE_F32, n_embd, n_tok, x);
    ggml_tensor * WPU = s.leaf2(ptype, n_embd, r, wpu);
    ggml_tensor * WPD = s.leaf2(ptype, r, n_ff, wpd);
    ggml_tensor * BPU = bpu ? s.leaf1(GGML_TYPE_F32, r, bpu) : nullptr;
    ggml_tensor * BPD = bpd ? s.leaf1(GGML_TYPE_F32, n_ff, bpd) : nullptr;
    ggml_tensor * P   = predictor(s, X, WPU, BPU, WPD, BPD);
    ggml_cgraph * gf  = ggml_new_graph(s.ctx);
    ggml_build_forward_expand(gf, P);
    if (run(s, gf, n_threads)) {
        return 1;
    }
    memcpy(sparse_idx_out, P->data, sizeof(float) * (size_t) n_ff * n_tok);
    return 0;
}

// build_sparse_ffn (src/llama-graph.cpp:896-1142), all neurons CPU-side. Optional outputs nullable.
int spif_ref_sparse_ffn(int arch, int wtype, const void * w_up, const void * w_gate, const void * w_down_t,
                        const float * b_up, const float * b_gate, const float * b_down, const float * x,
                        const float * sparse_idx, float * y, float * o_up, float * o_gate, float * o_hid,
                        int n_embd, int n_ff, int n_tok, int n_threads) {
    shim_ctx             s(scratch_bytes(n_embd, n_ff, 0, n_tok));
    std::vector<int32_t> zero_mask((size_t) n_ff, 0);
    ggml_tensor *        X  = s.leaf2(GGML_TYPE_F32, n_embd, n_tok, x);
    ggml_tensor *        P  = s.leaf2(GGML_TYPE_F32, n_ff, n_tok, sparse_idx);
    ggml_tensor *        M  = s.leaf1(GGML_TYPE_I32, n_ff, zero_mask.data());
    ggml_tensor *        U  = s.leaf2(wtype, n_embd, n_ff, w_up);
    ggml_tensor *        G  = arch != 2 ? s.leaf2(wtype, n_embd, n_ff, w_gate) : nullptr;
    ggml_tensor *        D  = s.leaf2(wtype, n_embd, n_ff, w_down_t);
    ggml_tensor *        BU = b_up ? s.leaf1(GGML_TYPE_F32, n_ff, b_up) : nullptr;
    ggml_tensor *        BG = (b_gate && arch != 2) ? s.leaf1(GGML_TYPE_F32, n_ff, b_gate) : nullptr;
    ggml_tensor *        BD = b_down ? s.leaf1(GGML_TYPE_F32, n_embd, b_down) : nullptr;
    ggml_tensor *        tu = nullptr, *tg = nullptr, *th = nullptr;
    ggml_tensor *        Y  = sparse_ffn(s, arch, X, P, M, U, G, D, BU, BG, BD, &tu, &tg, &th);
    ggml_cgraph *        gf = ggml_new_graph(s.ctx);
    ggml_build_forward_expand(gf, Y);
    if (run(s, gf, n_threads)) {
        return 1;
    }
    memcpy(y, Y->data, sizeof(float) * (size_t) n_embd * n_tok);
    if (o_up && tu) {
        memcpy(o_up, tu->data, sizeof(float) * (size_t) n_ff * n_tok);
    }
    if (o_gate && tg) {
        memcpy(o_gate, tg->data, sizeof(float) * (size_t) n_ff * n_tok);
    }
    if (o_hid && th) {
        memcpy(o_hid, th->data, sizeof(float) * (size_t) n_ff * n_tok);
    }
    return 0;
}

// Times the reference CPU path for one decode step over a stack of layers, the way build_sparse_ffn chains
// them: layer l's mask is predicted from layer l-1's FFN input (look-ahead, llama-graph.cpp:939-946; layer 0
// predicts itself :933-938).  x_l are given (the attention blocks between FFNs are out of scope).
// Pointers are per-layer arrays.  Returns the median microseconds per step over `iters` timed steps.
double spif_ref_ffn_stack_time(int arch, int wtype, int ptype, int n_layers, const void * const * w_up,
                               const void * const * w_gate, const void * const * w_down_t,
                               const void * const * wpu, const void * const * wpd, const float * const * bpd,
                               const float * x /* [n_layers][n_tok][n_embd] */, float * y /* same shape */,
                               int n_embd, int n_ff, int r, int n_tok, int n_threads, int warmup, int iters) {
    shim_ctx             s(scratch_bytes(n_embd, n_ff, r, n_tok) * (size_t) n_layers);
    std::vector<int32_t> zero_mask((size_t) n_ff, 0);
    ggml_tensor *        M  = s.leaf1(GGML_TYPE_I32, n_ff, zero_mask.data());
    ggml_cgraph *        gf = ggml_new_graph_custom(s.ctx, 64 * n_layers + 64, false);
    std::vector<ggml_tensor *> Y((size_t) n_layers);
    ggml_tensor *              next_p = nullptr;
    for (int l = 0; l < n_layers; ++l) {
        ggml_tensor * X = s.leaf2(GGML_TYPE_F32, n_embd, n_tok, x + (size_t) l * n_tok * n_embd);
        ggml_tensor * P = next_p;
        if (l == 0) {
            P = predictor(s, X, s.leaf2(ptype, n_embd, r, wpu[0]), nullptr, s.leaf2(ptype, r, n_ff, wpd[0]),
                          bpd ? s.leaf1(GGML_TYPE_F32, n_ff, bpd[0]) : nullptr);
            ggml_build_forward_expand(gf, P);
        }
        if (l + 1 < n_layers) {
            next_p = predictor(s, X, s.leaf2(ptype, n_embd, r, wpu[l + 1]), nullptr,
                               s.leaf2(ptype, r, n_ff, wpd[l + 1]),
                               bpd ? s.leaf1(GGML_TYPE_F32, n_ff, bpd[l + 1]) : nullptr);
            ggml_build_forward_expand(gf, next_p);
        }
        Y[l] = sparse_ffn(s, arch, X, P, M, s.leaf2(wtype, n_embd, n_ff, w_up[l]),
                          arch != 2 ? s.leaf2(wtype, n_embd, n_ff, w_gate[l]) : nullptr,
                          s.leaf2(wtype, n_embd, n_ff, w_down_t[l]), nullptr, nullptr, nullptr, nullptr, nullptr,
                          nullptr);
        ggml_build_forward_expand(gf, Y[l]);
    }
    ggml_cplan           plan = ggml_graph_plan(gf, n_threads, nullptr);
    std::vector<uint8_t> work(plan.work_size + 64);
    plan.work_data = work.data();
    std::vector<double> us;
    for (int it = 0; it < warmup + iters; ++it) {
        auto t0 = std::chrono::steady_clock::now();
        if (ggml_graph_compute(gf, &plan) != GGML_STATUS_SUCCESS) {
            return -1.0;
        }
        auto t1 = std::chrono::steady_clock::now();
        if (it >= warmup) {
            us.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
        }
    }
    if (y) {
        for (int l = 0; l < n_layers; ++l) {
            memcpy(y + (size_t) l * n_tok * n_embd, Y[l]->data, sizeof(float) * (size_t) n_embd * n_tok);
        }
    }
    std::sort(us.begin(), us.end());
    return us.empty() ? -1.0 : us[us.size() / 2];
}

}  // extern "C"
