// glue_harness.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Runs glue/ggml_cuda_spif.cu -- the binding a SparkInfer maintainer adds to ggml-cuda -- for real: on the reference's OWN
// structures (ggml_context, ggml_tensor, ggml_cgraph, ggml_backend_cuda_context with its pool), on a graph built with the
// reference's OWN constructors in the order build_sparse_ffn emits it (src/llama-graph.cpp:969-971, 979-981, 1063-1098),
// next to the reference's OWN CUDA kernels on the same graph.  Three passes over one layer, n_tok == 1, gpu_only:
//   A  reference kernels node by node (dispatch as ggml-cuda.cu:2463-2497,2802-2807; kernels from libspif_refcuda.so)
//   B  the same node loop with the sparse ops routed to spif_glue_mul_mat_sparse / _axpy_sparse / _fatrelu
//   C  the node loop with the fusion hook: spif_glue_try_fuse_sparse_ffn(ctx, cgraph, 0) -- one spif_sparse_ffn call
// Built by oracle/Makefile into _ref/libspif_gluetest.so (links libspif_refcuda.so, libspif_ref.so, libspif_b200.so);
// driven by tests/test_glue.py (-m gpu).  The full llama-cli run needs the whole reference build and a
// model file and is not attempted; this is the op-level drop-in check on real hardware.
#include "common.cuh"
#include "ggml.h"
#include "mm-sparse.cuh"
#include "axpy-sparse.cuh"
#include "mmq-sparse.cuh"
#include "axpyq-sparse.cuh"
#include "unary.cuh"
#include "binbcast.cuh"

#include <cstring>
#include <vector>

// glue/ggml_cuda_spif.cu
void spif_glue_mul_mat_sparse(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, const ggml_tensor * src1, ggml_tensor * dst);
void spif_glue_axpy_sparse(ggml_backend_cuda_context & ctx, const ggml_tensor * src0, const ggml_tensor * src1, ggml_tensor * dst);
void spif_glue_fatrelu(ggml_backend_cuda_context & ctx, ggml_tensor * dst);
int  spif_glue_try_fuse_sparse_ffn(ggml_backend_cuda_context & ctx, ggml_cgraph * cgraph, int i);

namespace {

ggml_tensor * leaf(ggml_context * c, ggml_type type, int64_t ne0, int64_t ne1, const void * dev_ptr) {
    ggml_tensor * t = ggml_new_tensor_2d(c, type, ne0, ne1);
    t->data         = const_cast<void *>(dev_ptr);
    return t;
}

// one node through the reference's kernels (n_tok == 1: the arguments ggml_cuda_op_mul_mat would forward on one device)
void ref_node(ggml_backend_cuda_context & ctx, ggml_tensor * n) {
    cudaStream_t st = ctx.stream();
    switch (n->op) {
        case GGML_OP_MUL_MAT_SPARSE:
            if (n->src[0]->type == GGML_TYPE_Q8_0) {  // ggml-cuda.cu:2474-2476
                ggml_cuda_mul_mat_sparse_q(ctx, n->src[0], n->src[1], n);
                break;
            }
            ggml_cuda_op_mul_mat_sparse(ctx, n->src[0], n->src[1], n, (const char *) n->src[0]->data, (const float *) n->src[1]->data,
                                        nullptr, (float *) n->data, 0, n->src[0]->ne[1], 1, n->src[0]->ne[0], st);
            break;
        case GGML_OP_AXPY_SPARSE:
            if (n->src[0]->type == GGML_TYPE_Q8_0) {  // ggml-cuda.cu:2491-2493
                ggml_cuda_op_axpy_sparse_q(ctx, n->src[0], n->src[1], n, (const char *) n->src[0]->data, (const float *) n->src[1]->data,
                                           nullptr, (float *) n->data, 0, n->src[0]->ne[1], 1, n->src[0]->ne[1], st);
                break;
            }
            ggml_cuda_op_axpy_sparse(ctx, n->src[0], n->src[1], n, (const char *) n->src[0]->data, (const float *) n->src[1]->data,
                                     nullptr, (float *) n->data, 0, n->src[0]->ne[1], 1, n->src[0]->ne[1], st);
            break;
        case GGML_OP_FATRELU:
            ggml_cuda_op_fatrelu(ctx, n);
            break;
        case GGML_OP_UNARY:
            GGML_ASSERT(ggml_get_unary_op(n) == GGML_UNARY_OP_RELU);
            ggml_cuda_op_relu(ctx, n);
            break;
        case GGML_OP_MUL:
            ggml_cuda_op_mul(ctx, n);
            break;
        default:
            GGML_ABORT("glue harness: unexpected op %s", ggml_op_name(n->op));
    }
}

}  // namespace

extern "C" int spif_gluetest_layer(int arch, int wtype, const void * w_up, const void * w_gate, const void * w_down, const float * x,
                                   const float * sparse_idx, int n_embd, int n_ff, float * up_ref, float * up_glue, float * y_ref,
                                   float * y_glue_ops, float * y_glue_fused, int * nodes_total, int * nodes_fused) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        return -1;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    ggml_backend_cuda_context ctx(dev);
    cudaStream_t              st = ctx.stream();

    ggml_init_params ip = { ggml_tensor_overhead() * 64 + ggml_graph_overhead(), nullptr, /*no_alloc=*/true };
    ggml_context *   c  = ggml_init(ip);
    ggml_cgraph *    gf = ggml_new_graph(c);
    ggml_tensor *    Wu = leaf(c, (ggml_type) wtype, n_embd, n_ff, w_up);
    ggml_tensor *    Wg = arch != 2 ? leaf(c, (ggml_type) wtype, n_embd, n_ff, w_gate) : nullptr;
    ggml_tensor *    Wd = leaf(c, (ggml_type) wtype, n_embd, n_ff, w_down);
    ggml_tensor *    X  = leaf(c, GGML_TYPE_F32, n_embd, 1, x);
    ggml_tensor *    P  = leaf(c, GGML_TYPE_F32, n_ff, 1, sparse_idx);
    // build_sparse_ffn, gpu_only (src/llama-graph.cpp:969-971, 979-981, 1063-1098): each sparse op is expanded where it is built
    ggml_tensor * cur_up = ggml_mul_mat_sparse(c, Wu, X, P, nullptr);
    ggml_build_forward_expand(gf, cur_up);
    ggml_tensor * first_up = cur_up;
    ggml_tensor * cur_gate = nullptr;
    if (Wg) {
        cur_gate = ggml_mul_mat_sparse(c, Wg, X, P, nullptr);
        ggml_build_forward_expand(gf, cur_gate);
    }
    ggml_tensor * hid = nullptr;
    switch (arch) {
        case 0:
            cur_gate = ggml_fatrelu(c, cur_gate, 0.01f, false);
            hid      = ggml_mul(c, cur_gate, cur_up);
            break;
        case 1:
            cur_up   = ggml_relu(c, cur_up);
            cur_gate = ggml_relu(c, cur_gate);
            hid      = ggml_mul(c, cur_gate, cur_up);
            break;
        default:
            hid = ggml_relu(c, cur_up);
            break;
    }
    ggml_tensor * down = ggml_axpy_sparse(c, Wd, hid, P, nullptr);
    ggml_build_forward_expand(gf, down);

    const int          n_nodes = ggml_graph_n_nodes(gf);
    std::vector<void *> bufs;
    for (int i = 0; i < n_nodes; ++i) {
        ggml_tensor * n = ggml_graph_node(gf, i);
        void *        d = nullptr;
        CUDA_CHECK(cudaMalloc(&d, ggml_nbytes(n)));
        n->data = d;
        bufs.push_back(d);
    }
    auto poison = [&]() {
        for (int i = 0; i < n_nodes; ++i) {
            ggml_tensor * n = ggml_graph_node(gf, i);
            CUDA_CHECK(cudaMemsetAsync(n->data, 0xFF, ggml_nbytes(n), st));  // NaN: a node nobody wrote shows up
        }
    };
    auto fetch = [&](ggml_tensor * n, float * dst) {
        CUDA_CHECK(cudaMemcpyAsync(dst, n->data, ggml_nbytes(n), cudaMemcpyDeviceToDevice, st));
    };
    // A: the reference's kernels
    poison();
    for (int i = 0; i < n_nodes; ++i) {
        ref_node(ctx, ggml_graph_node(gf, i));
    }
    fetch(first_up, up_ref);
    fetch(down, y_ref);
    // B: the sparse ops through the glue, everything else through the reference's stock kernels
    poison();
    for (int i = 0; i < n_nodes; ++i) {
        ggml_tensor * n = ggml_graph_node(gf, i);
        if (n->op == GGML_OP_MUL_MAT_SPARSE) {
            spif_glue_mul_mat_sparse(ctx, n->src[0], n->src[1], n);
        } else if (n->op == GGML_OP_AXPY_SPARSE) {
            spif_glue_axpy_sparse(ctx, n->src[0], n->src[1], n);
        } else if (n->op == GGML_OP_FATRELU) {
            spif_glue_fatrelu(ctx, n);
        } else {
            ref_node(ctx, n);
        }
    }
    fetch(first_up, up_glue);
    fetch(down, y_glue_ops);
    // C: the node loop with the fusion hook
    poison();
    int fused = 0;
    for (int i = 0; i < n_nodes;) {
        const int used = spif_glue_try_fuse_sparse_ffn(ctx, gf, i);
        if (used > 0) {
            fused += used;
            i += used;
        } else {
            ref_node(ctx, ggml_graph_node(gf, i));
            ++i;
        }
    }
    fetch(down, y_glue_fused);
    CUDA_CHECK(cudaStreamSynchronize(st));
    *nodes_total = n_nodes;
    *nodes_fused = fused;
    for (void * d : bufs) {
        cudaFree(d);
    }
    ggml_free(c);
    return 0;
}

// ---- the look-ahead glue on a two-layer chain ---------------------------------------------------------------------------
// Reference order (src/llama-graph.cpp:932-946): layer 0 predicts its own mask from x_0; while layer 0 runs, the mask of
// layer 1 is predicted from x_0 too (the FFN's input); then layer 1 runs on x_1.  Pass A: build_predictor's six nodes and
// build_sparse_ffn's nodes through the reference's kernels.  Pass B: three spif_glue_lookahead calls (predictor only;
// FFN 0 + predictor 1; FFN 1 only) on the same tensors.
#include "mmvf.cuh"
#include "spif_b200.h"

struct spif_glue_lookahead_state {
    spif_shard_comm comm;
    void *          list[2];
};
void spif_glue_lookahead(ggml_backend_cuda_context & ctx, spif_glue_lookahead_state & st, int il, int arch, const ggml_tensor * up,
                         const ggml_tensor * gate, ggml_tensor * down, const ggml_tensor * x_pred, const ggml_tensor * pred_up,
                         const ggml_tensor * pred_up_b, const ggml_tensor * pred_down, const ggml_tensor * pred_down_b,
                         ggml_tensor * sparse_idx_next);

namespace {
void ref_pred_node(ggml_backend_cuda_context & ctx, ggml_tensor * n) {
    switch (n->op) {
        case GGML_OP_MUL_MAT:
            ggml_cuda_mul_mat_vec_f(ctx, n->src[0], n->src[1], nullptr, n, nullptr);  // n_tok == 1 (ggml-cuda.cu mul_mat dispatch)
            break;
        case GGML_OP_ADD:
            ggml_cuda_op_add(ctx, n);
            break;
        case GGML_OP_UNARY:
            if (ggml_get_unary_op(n) == GGML_UNARY_OP_RELU) {
                ggml_cuda_op_relu(ctx, n);
            } else {
                GGML_ASSERT(ggml_get_unary_op(n) == GGML_UNARY_OP_SIGMOID);
                ggml_cuda_op_sigmoid(ctx, n);
            }
            break;
        default:
            GGML_ABORT("glue harness: unexpected predictor op %s", ggml_op_name(n->op));
    }
}
}  // namespace

// weights: [2] device pointers per matrix (layer 0, layer 1); x: [2][n_embd]; outputs: p_* [2][n_ff], y_* [2][n_embd] (device)
extern "C" int spif_gluetest_lookahead(int wtype, void ** w_up, void ** w_gate, void ** w_down, void ** wpu, void ** wpd, float ** bpd,
                                       const float * x, int n_embd, int n_ff, int r, float * p_ref, float * p_glue, float * y_ref,
                                       float * y_glue) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        return -1;
    }
    ggml_backend_cuda_context ctx(dev);
    cudaStream_t              st = ctx.stream();
    ggml_init_params ip = { ggml_tensor_overhead() * 128 + ggml_graph_overhead(), nullptr, /*no_alloc=*/true };
    ggml_context *   c  = ggml_init(ip);
    ggml_cgraph *    gf = ggml_new_graph(c);
    std::vector<void *> bufs;
    auto out_buf = [&](ggml_tensor * n) {
        void * d = nullptr;
        CUDA_CHECK(cudaMalloc(&d, ggml_nbytes(n)));
        n->data = d;
        bufs.push_back(d);
    };
    ggml_tensor * X[2], * PU[2], * PD[2], * BPD[2], * Wu[2], * Wg[2], * Wd[2], * sig[2], * up[2], * gate[2], * down[2];
    for (int l = 0; l < 2; ++l) {
        X[l]   = leaf(c, GGML_TYPE_F32, n_embd, 1, x + (size_t) l * n_embd);
        PU[l]  = leaf(c, (ggml_type) wtype, n_embd, r, wpu[l]);
        PD[l]  = leaf(c, (ggml_type) wtype, r, n_ff, wpd[l]);
        BPD[l] = leaf(c, GGML_TYPE_F32, n_ff, 1, bpd[l]);
        Wu[l]  = leaf(c, (ggml_type) wtype, n_embd, n_ff, w_up[l]);
        Wg[l]  = leaf(c, (ggml_type) wtype, n_embd, n_ff, w_gate[l]);
        Wd[l]  = leaf(c, (ggml_type) wtype, n_embd, n_ff, w_down[l]);
    }
    // build_predictor (llama-graph.cpp:865-894) for both layers, both from x_0 (layer 0: its own input; layer 1: look-ahead)
    std::vector<ggml_tensor *> pred_nodes[2];
    for (int l = 0; l < 2; ++l) {
        ggml_tensor * cur = ggml_mul_mat(c, PU[l], X[0]);
        pred_nodes[l].push_back(cur);
        cur = ggml_relu(c, cur);
        pred_nodes[l].push_back(cur);
        cur = ggml_mul_mat(c, PD[l], cur);
        pred_nodes[l].push_back(cur);
        cur = ggml_add(c, cur, BPD[l]);
        pred_nodes[l].push_back(cur);
        cur = ggml_sigmoid(c, cur);
        pred_nodes[l].push_back(cur);
        sig[l] = cur;
        for (ggml_tensor * n : pred_nodes[l]) {
            out_buf(n);
        }
    }
    // build_sparse_ffn (ProSparse recipe) for both layers on x_l with the predicted sparse_idx
    std::vector<ggml_tensor *> ffn_nodes[2];
    for (int l = 0; l < 2; ++l) {
        up[l]   = ggml_mul_mat_sparse(c, Wu[l], X[l], sig[l], nullptr);
        gate[l] = ggml_mul_mat_sparse(c, Wg[l], X[l], sig[l], nullptr);
        ggml_tensor * act = ggml_fatrelu(c, gate[l], 0.01f, false);
        ggml_tensor * hid = ggml_mul(c, act, up[l]);
        down[l]           = ggml_axpy_sparse(c, Wd[l], hid, sig[l], nullptr);
        ffn_nodes[l]      = { up[l], gate[l], act, hid, down[l] };
        for (ggml_tensor * n : ffn_nodes[l]) {
            out_buf(n);
        }
    }
    (void) gf;
    const size_t pb = sizeof(float) * (size_t) n_ff, yb = sizeof(float) * (size_t) n_embd;
    // A: reference kernels, reference order
    for (ggml_tensor * n : pred_nodes[0]) ref_pred_node(ctx, n);
    for (ggml_tensor * n : pred_nodes[1]) ref_pred_node(ctx, n);
    for (int l = 0; l < 2; ++l) {
        for (ggml_tensor * n : ffn_nodes[l]) ref_node(ctx, n);
        CUDA_CHECK(cudaMemcpyAsync(p_ref + (size_t) l * n_ff, sig[l]->data, pb, cudaMemcpyDeviceToDevice, st));
        CUDA_CHECK(cudaMemcpyAsync(y_ref + (size_t) l * n_embd, down[l]->data, yb, cudaMemcpyDeviceToDevice, st));
    }
    // B: the look-ahead glue.  Persistent state as the cache manager would create it once: zeroed comm (world 1), two lists.
    spif_glue_lookahead_state ls = {};
    ls.comm.rank = 0, ls.comm.world = 1, ls.comm.cap_floats = (n_embd + r + 3) / 4 * 4;
    const size_t cb = spif_shard_comm_data_bytes(1, ls.comm.cap_floats);
    CUDA_CHECK(cudaMalloc(&ls.comm.data[0], cb));
    CUDA_CHECK(cudaMalloc(&ls.comm.seq, sizeof(uint32_t) * SPIF_COMM_SEQ_WORDS));
    CUDA_CHECK(cudaMemsetAsync(ls.comm.data[0], 0, cb, st));
    CUDA_CHECK(cudaMemsetAsync(ls.comm.seq, 0, sizeof(uint32_t) * SPIF_COMM_SEQ_WORDS, st));
    for (int k = 0; k < 2; ++k) {
        CUDA_CHECK(cudaMalloc(&ls.list[k], spif_active_list_bytes(n_ff)));
    }
    for (int l = 0; l < 2; ++l) {  // poison what pass B must write
        CUDA_CHECK(cudaMemsetAsync(sig[l]->data, 0xFF, pb, st));
        CUDA_CHECK(cudaMemsetAsync(down[l]->data, 0xFF, yb, st));
    }
    spif_glue_lookahead(ctx, ls, 0, 0, nullptr, nullptr, nullptr, X[0], PU[0], nullptr, PD[0], BPD[0], sig[0]);   // predictor(0) from x_0
    spif_glue_lookahead(ctx, ls, 0, 0, up[0], gate[0], down[0], X[0], PU[1], nullptr, PD[1], BPD[1], sig[1]);     // FFN(0) + predictor(1)
    spif_glue_lookahead(ctx, ls, 1, 0, up[1], gate[1], down[1], nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);  // FFN(1)
    for (int l = 0; l < 2; ++l) {
        CUDA_CHECK(cudaMemcpyAsync(p_glue + (size_t) l * n_ff, sig[l]->data, pb, cudaMemcpyDeviceToDevice, st));
        CUDA_CHECK(cudaMemcpyAsync(y_glue + (size_t) l * n_embd, down[l]->data, yb, cudaMemcpyDeviceToDevice, st));
    }
    int err = 0;
    const int rc = spif_shard_comm_status(&ls.comm, &err, (spif_stream_t) st);
    CUDA_CHECK(cudaStreamSynchronize(st));
    for (void * d : bufs) {
        cudaFree(d);
    }
    cudaFree(ls.comm.data[0]), cudaFree(ls.comm.seq), cudaFree(ls.list[0]), cudaFree(ls.list[1]);
    ggml_free(c);
    return rc ? rc : err;
}
