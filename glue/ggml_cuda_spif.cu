// glue/ggml_cuda_spif.cu -- the binding a SparkInfer maintainer adds to ggml/src/ggml-cuda/ so that the reference's own
// graph (GGML_OP_MUL_MAT_SPARSE / AXPY_SPARSE / FATRELU / SHIFTED_STEP nodes built by src/llama-graph.cpp:865-1142) runs on
// libspif_b200.so.  It is compiled against the reference's headers by tests/test_glue.py (nvcc -c, no link), so every
// field it reads from ggml_tensor / ggml_backend_cuda_context exists with that name and type in the reference tree.
//
// How to use it: `patch -p1 < glue/reference_tree.patch` in the reference tree and configure with
// -DGGML_CUDA_SPIF_B200_DIR=<this checkout> (tests/test_glue.py applies the patch to the reference's files and compiles the
// patched dispatch).  What the patch does: add this file to GGML_SOURCES_CUDA, link -lspif_b200, and in ggml-cuda.cu
//   * replace the bodies of ggml_cuda_mul_mat_sparse / ggml_cuda_axpy_sparse (ggml-cuda.cu:2463-2497) by calls to
//     spif_glue_mul_mat_sparse / spif_glue_axpy_sparse;
//   * route GGML_OP_FATRELU / GGML_OP_SHIFTED_STEP (ggml-cuda.cu:2802-2807) to spif_glue_fatrelu / spif_glue_shifted_step;
//   * in the fusion block of the node loop (ggml_cuda_can_fuse, ggml-cuda.cu:3279; fusion section :3578-3870) call
//     spif_glue_try_fuse_sparse_ffn(ctx, cgraph, i): it returns the number of nodes it consumed (0 = pattern not matched).
// Scratch comes from ctx.pool() exactly like the reference's own Q8_0 path (mmq-sparse.cu:339): a caching pool, so nothing
// is cudaMalloc'ed on the hot path after the first token and the calls stay CUDA-graph capturable.  No global state.
#include "common.cuh"
#include "ggml.h"

#include "spif_b200.h"

#ifndef SPIF_SPARSE_THRESHOLD
#define SPIF_SPARSE_THRESHOLD 0.5f /* mm-sparse.cu:5-7 */
#endif

static void spif_glue_check(int rc, const char * what) {
    if (rc != 0) {
        GGML_ABORT("%s: %s", what, spif_error_string(rc));  // same failure mode as GGML_ASSERT / CUDA_CHECK
    }
}

// ---- GGML_OP_MUL_MAT_SPARSE (ggml.c:3310-3331): src0 = weights {n_embd, m}, src1 = x {n_embd, n_tok}, dst->src[2] = sparse_idx
// {n_ff, n_tok} F32, dst->src[3] = neuron_idx I32 {m} or NULL (mm-sparse.cu:389-391)
void spif_glue_mul_mat_sparse(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, const ggml_tensor * src1, ggml_tensor * dst) {
    GGML_ASSERT(dst->src[2] != NULL && "dst->src[2] must be present for sparse matrix multiplication");
    const ggml_tensor * mask = dst->src[2];
    const ggml_tensor * nidx = dst->src[3];
    GGML_ASSERT(src1->type == GGML_TYPE_F32 && dst->type == GGML_TYPE_F32 && mask->type == GGML_TYPE_F32);
    GGML_ASSERT(ggml_is_contiguous(src0) && ggml_is_contiguous(src1) && ggml_is_contiguous(mask) && ggml_is_contiguous(dst));
    const int n_embd = (int) src0->ne[0], m = (int) src0->ne[1];
    const int n_ff = (int) mask->ne[0], n_tok = (int) src1->ne[1];
    const size_t               ws_bytes = spif_workspace_size(n_embd, n_ff, n_tok);
    ggml_cuda_pool_alloc<char> ws(ctx.pool(), ws_bytes + 256);
    char * ws_ptr = (char *) (((uintptr_t) ws.get() + 255) & ~(uintptr_t) 255);
    spif_glue_check(spif_mul_mat_sparse((int) src0->type, src0->data, (const float *) src1->data, (const float *) mask->data,
                                        nidx ? (const int32_t *) nidx->data : NULL, (float *) dst->data, n_embd, n_ff, m, n_tok,
                                        SPIF_SPARSE_THRESHOLD, /*flags=*/0, ws_ptr, ws_bytes, (spif_stream_t) ctx.stream()),
                    "spif_mul_mat_sparse");
}

// ---- GGML_OP_AXPY_SPARSE (ggml.c:3333-3356): src0 = W_down TRANSPOSED {n_embd, m}, src1 = hidden {n_ff, n_tok}
void spif_glue_axpy_sparse(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, const ggml_tensor * src1, ggml_tensor * dst) {
    GGML_ASSERT(dst->src[2] != NULL && "dst->src[2] must be present for sparse matrix multiplication");
    const ggml_tensor * mask = dst->src[2];
    const ggml_tensor * nidx = dst->src[3];
    GGML_ASSERT(src1->type == GGML_TYPE_F32 && dst->type == GGML_TYPE_F32);
    const int n_embd = (int) src0->ne[0], m = (int) src0->ne[1];
    const int n_ff = (int) src1->ne[0], n_tok = (int) src1->ne[1];
    const size_t               ws_bytes = spif_workspace_size(n_embd, n_ff, n_tok);
    ggml_cuda_pool_alloc<char> ws(ctx.pool(), ws_bytes + 256);
    char * ws_ptr = (char *) (((uintptr_t) ws.get() + 255) & ~(uintptr_t) 255);
    spif_glue_check(spif_axpy_sparse((int) src0->type, src0->data, (const float *) src1->data, (const float *) mask->data,
                                     nidx ? (const int32_t *) nidx->data : NULL, (float *) dst->data, n_embd, n_ff, m, n_tok,
                                     SPIF_SPARSE_THRESHOLD, 0, ws_ptr, ws_bytes, (spif_stream_t) ctx.stream()),
                    "spif_axpy_sparse");
}

// ---- GGML_OP_FATRELU / GGML_OP_SHIFTED_STEP (ggml.c:2748-2761; unary.cu:564-652): op_params[0] = threshold
void spif_glue_fatrelu(ggml_backend_cuda_context & ctx, ggml_tensor * dst) {
    GGML_ASSERT(dst->src[0]->type == GGML_TYPE_F32 && dst->type == GGML_TYPE_F32 && ggml_is_contiguous(dst->src[0]));
    float t;
    memcpy(&t, dst->op_params, sizeof(float));
    spif_glue_check(spif_fatrelu((const float *) dst->src[0]->data, (float *) dst->data, ggml_nelements(dst), t,
                                 (spif_stream_t) ctx.stream()), "spif_fatrelu");
}
void spif_glue_shifted_step(ggml_backend_cuda_context & ctx, ggml_tensor * dst) {
    GGML_ASSERT(dst->src[0]->type == GGML_TYPE_F32 && dst->type == GGML_TYPE_F32 && ggml_is_contiguous(dst->src[0]));
    float t;
    memcpy(&t, dst->op_params, sizeof(float));
    spif_glue_check(spif_shifted_step((const float *) dst->src[0]->data, (float *) dst->data, ggml_nelements(dst), t,
                                      (spif_stream_t) ctx.stream()), "spif_shifted_step");
}

// ---- whole-layer fusion.  build_sparse_ffn (gpu_only, no biases) emits, for the ProSparse recipe (llama-graph.cpp:978-1139):
//   n[i]   MUL_MAT_SPARSE(up)      n[i+1] MUL_MAT_SPARSE(gate)    n[i+2] FATRELU(gate)    n[i+3] MUL(fatrelu, up)
//   n[i+4] AXPY_SPARSE(down_t, hidden)
// and for Bamboo / Qwen2 RELU(gate), RELU(up), MUL; for OPT / Falcon RELU(up).  All sparse ops of a layer share sparse_idx
// (src[2]) and neuron_idx (src[3]).  Returns the number of nodes consumed; the caller advances its loop by that many.
int spif_glue_try_fuse_sparse_ffn(ggml_backend_cuda_context & ctx, ggml_cgraph * cgraph, int i) {
    const int     n_nodes = ggml_graph_n_nodes(cgraph);
    ggml_tensor * up      = ggml_graph_node(cgraph, i);
    if (up->op != GGML_OP_MUL_MAT_SPARSE || i + 2 >= n_nodes) {
        return 0;
    }
    int           arch = -1, used = 0;
    ggml_tensor * gate = NULL;
    ggml_tensor * down = NULL;
    ggml_tensor * n1   = ggml_graph_node(cgraph, i + 1);
    if (n1->op == GGML_OP_MUL_MAT_SPARSE && n1->src[1] == up->src[1] && n1->src[2] == up->src[2] && i + 4 < n_nodes) {
        gate               = n1;
        ggml_tensor * a    = ggml_graph_node(cgraph, i + 2);
        ggml_tensor * b    = ggml_graph_node(cgraph, i + 3);
        ggml_tensor * c    = ggml_graph_node(cgraph, i + 4);
        if (a->op == GGML_OP_FATRELU && a->src[0] == gate && b->op == GGML_OP_MUL && c->op == GGML_OP_AXPY_SPARSE && c->src[1] == b) {
            float t;
            memcpy(&t, a->op_params, sizeof(float));
            if (t != 0.01f) {
                return 0;  // the fused recipe hard-codes the ProSparse threshold (llama-graph.cpp:1067)
            }
            arch = SPIF_ARCH_PROSPARSE, down = c, used = 5;
        } else if (a->op == GGML_OP_UNARY && ggml_get_unary_op(a) == GGML_UNARY_OP_RELU && b->op == GGML_OP_UNARY &&
                   ggml_get_unary_op(b) == GGML_UNARY_OP_RELU && c->op == GGML_OP_MUL && i + 5 < n_nodes &&
                   ggml_graph_node(cgraph, i + 5)->op == GGML_OP_AXPY_SPARSE) {
            arch = SPIF_ARCH_RELU_GATED, down = ggml_graph_node(cgraph, i + 5), used = 6;
        }
    } else if (n1->op == GGML_OP_UNARY && ggml_get_unary_op(n1) == GGML_UNARY_OP_RELU && n1->src[0] == up &&
               ggml_graph_node(cgraph, i + 2)->op == GGML_OP_AXPY_SPARSE) {
        arch = SPIF_ARCH_RELU_UP, down = ggml_graph_node(cgraph, i + 2), used = 3;
    }
    if (arch < 0 || down->src[2] != up->src[2] || down->src[3] != up->src[3] || down->src[0]->type != up->src[0]->type) {
        return 0;
    }
    const ggml_tensor * mask = up->src[2];
    const ggml_tensor * nidx = up->src[3];
    spif_ffn_args a = {};
    a.arch       = arch;
    a.type       = (int32_t) up->src[0]->type;
    a.n_embd     = (int32_t) up->src[0]->ne[0];
    a.n_ff       = (int32_t) mask->ne[0];
    a.m          = (int32_t) up->src[0]->ne[1];
    a.n_tok      = (int32_t) up->src[1]->ne[1];
    a.flags      = 0;
    a.thr        = SPIF_SPARSE_THRESHOLD;
    a.w_up       = up->src[0]->data;
    a.w_gate     = gate ? gate->src[0]->data : NULL;
    a.w_down_t   = down->src[0]->data;
    a.x          = (const float *) up->src[1]->data;
    a.sparse_idx = (const float *) mask->data;
    a.neuron_idx = nidx ? (const int32_t *) nidx->data : NULL;
    a.y          = (float *) down->data;
    const size_t               ws_bytes = spif_workspace_size(a.n_embd, a.n_ff, a.n_tok);
    ggml_cuda_pool_alloc<char> ws(ctx.pool(), ws_bytes + 256);
    char * ws_ptr = (char *) (((uintptr_t) ws.get() + 255) & ~(uintptr_t) 255);
    spif_glue_check(spif_sparse_ffn(&a, ws_ptr, ws_bytes, (spif_stream_t) ctx.stream()), "spif_sparse_ffn");
    return used;
}

// ---- look-ahead layer: FFN(l) + predictor(l+1) in one launch.  The persistent objects (one spif_shard_comm with its zeroed
// buffers, two active-row lists) live next to the reference's per-layer cache state (sparkinfer_layer_cache,
// ggml/include/ggml-sparkinfer.hpp:26-81) and are created once by the cache manager (src/llama-sparkinfer.cpp:147-388).
struct spif_glue_lookahead_state {
    spif_shard_comm comm;      // world 1 on a single GPU; data[0] / seq are zero-initialised device buffers
    void *          list[2];   // spif_active_list_bytes(m) each
};

// up / gate / down: the layer's sparse nodes as above; x_pred: the tensor build_predictor is called on for layer l+1 (the FFN's
// own input, llama-graph.cpp:939-946; for the predictor-only call of layer 0 its own input); pred_*: the weights build_predictor
// reads (llama-graph.cpp:865-894); sparse_idx_next: the tensor the predictor's SIGMOID node writes.  il selects the list parity.
void spif_glue_lookahead(ggml_backend_cuda_context & ctx, spif_glue_lookahead_state & st, int il, int arch, const ggml_tensor * up,
                         const ggml_tensor * gate, ggml_tensor * down, const ggml_tensor * x_pred, const ggml_tensor * pred_up,
                         const ggml_tensor * pred_up_b, const ggml_tensor * pred_down, const ggml_tensor * pred_down_b,
                         ggml_tensor * sparse_idx_next) {
    spif_ffn_args  a = {};
    spif_pred_args q = {};
    const bool     has_ffn = up != NULL, has_pred = pred_up != NULL;
    int            n_embd = 0, n_ff = 0;
    if (has_ffn) {
        a.arch     = arch;
        a.type     = (int32_t) up->src[0]->type;
        a.n_embd   = n_embd = (int32_t) up->src[0]->ne[0];
        a.n_ff     = n_ff = (int32_t) up->src[2]->ne[0];
        a.m        = (int32_t) up->src[0]->ne[1];
        a.n_tok    = 1;
        a.thr      = SPIF_SPARSE_THRESHOLD;
        a.w_up     = up->src[0]->data;
        a.w_gate   = gate ? gate->src[0]->data : NULL;
        a.w_down_t = down->src[0]->data;
        a.x        = (const float *) up->src[1]->data;
        a.neuron_idx = up->src[3] ? (const int32_t *) up->src[3]->data : NULL;
        a.y        = (float *) down->data;
    }
    if (has_pred) {
        q.ptype  = (int32_t) pred_up->type;
        q.n_embd = (int32_t) pred_up->ne[0];
        q.r      = (int32_t) pred_up->ne[1];      // pred_lora
        q.m      = (int32_t) pred_down->ne[1];
        q.thr    = SPIF_SPARSE_THRESHOLD;
        q.wpu    = pred_up->data;
        q.bpu    = pred_up_b ? (const float *) pred_up_b->data : NULL;
        q.wpd    = pred_down->data;
        q.bpd    = pred_down_b ? (const float *) pred_down_b->data : NULL;
        GGML_ASSERT(x_pred != NULL && (!has_ffn || x_pred->data == up->src[1]->data));  // look-ahead: the FFN's own input
        q.x      = (const float *) x_pred->data;
        q.sparse_idx_out = (float *) sparse_idx_next->data;
        n_embd = q.n_embd, n_ff = n_ff ? n_ff : q.m;
    }
    const size_t               ws_bytes = spif_workspace_size(n_embd, n_ff, 1);
    ggml_cuda_pool_alloc<char> ws(ctx.pool(), ws_bytes + 256);
    char * ws_ptr = (char *) (((uintptr_t) ws.get() + 255) & ~(uintptr_t) 255);
    spif_glue_check(spif_sparse_ffn_lookahead(has_ffn ? &a : NULL, has_pred ? &q : NULL, has_ffn ? st.list[il & 1] : NULL,
                                              has_pred ? st.list[(il + 1) & 1] : NULL, &st.comm, ws_ptr, ws_bytes,
                                              (spif_stream_t) ctx.stream()),
                    "spif_sparse_ffn_lookahead");
}
