// ref_cuda_shim.cu -- TEST / BASELINE INFRASTRUCTURE ONLY (never linked into the product).
//
// GPU baseline (last row): the reference's OWN CUDA kernels for the sparse-FFN path, compiled unmodified for
// sm_100 from the sources where they lie under /root/reference (oracle/Makefile target `refcuda`):
//     ggml/src/ggml-cuda/mm-sparse.cu     mul_mat_vec_sparse / mul_mat_batch_sparse        (GGML_OP_MUL_MAT_SPARSE, F16/BF16)
//     ggml/src/ggml-cuda/axpy-sparse.cu   mul_mat_axpy_sparse_rowwise                     (GGML_OP_AXPY_SPARSE, F16/BF16)
//     ggml/src/ggml-cuda/mmq-sparse.cu, axpyq-sparse.cu, quantize.cu                      (the Q8_0 variants)
//     ggml/src/ggml-cuda/mmvf.cu          mul_mat_vec_f                                   (the predictor's two stock MUL_MATs)
//     ggml/src/ggml-cuda/unary.cu, binbcast.cu                                            (relu, sigmoid, fatrelu, mul, add)
// This file supplies what those translation units expect from ggml-cuda.cu (which is the whole backend and is NOT built):
// ggml_cuda_info / set_device / error, the context destructor and a trivial pool -- and drives the kernels in the order
// build_predictor / build_sparse_ffn emit them (src/llama-graph.cpp:865-894, 896-1142; dispatch ggml-cuda.cu:2463-2497),
// on ggml_tensor descriptors that point at the caller's device buffers.  The generic driver ggml_cuda_op_mul_mat
// (ggml-cuda.cu:1499-1800) only forwards pointers on a single device, so its callbacks are called directly.
// The predictor's bias add is folded into mul_mat_vec_f through the reference's own fusion argument (x_bias), as its graph
// fusion pass would (ggml-cuda.cu:3578-3870): the baseline gets the benefit of the doubt.
#include "common.cuh"
#include "ggml.h"
#include "mm-sparse.cuh"
#include "axpy-sparse.cuh"
#include "mmq-sparse.cuh"
#include "axpyq-sparse.cuh"
#include "mmvf.cuh"
#include "unary.cuh"
#include "binbcast.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

// ---- what ggml-cuda.cu would provide ---------------------------------------------------------------------------------
[[noreturn]] void ggml_cuda_error(const char * stmt, const char * func, const char * file, int line, const char * msg) {
    fprintf(stderr, "refcuda: CUDA error: %s\n  in %s at %s:%d: %s\n", msg, func, file, line, stmt);
    abort();
}
void ggml_cuda_set_device(int device) {
    cudaSetDevice(device);
}
int ggml_cuda_get_device() {
    int id = 0;
    cudaGetDevice(&id);
    return id;
}
const ggml_cuda_device_info & ggml_cuda_info() {
    static ggml_cuda_device_info info = [] {
        ggml_cuda_device_info i = {};
        cudaGetDeviceCount(&i.device_count);
        i.device_count = std::min(i.device_count, (int) GGML_CUDA_MAX_DEVICES);
        for (int id = 0; id < i.device_count; ++id) {
            cudaDeviceProp prop;
            cudaGetDeviceProperties(&prop, id);
            i.devices[id].cc         = 100 * prop.major + 10 * prop.minor;
            i.devices[id].nsm        = prop.multiProcessorCount;
            i.devices[id].smpb       = prop.sharedMemPerBlock;
            i.devices[id].smpbo      = prop.sharedMemPerBlockOptin;
            i.devices[id].total_vram = prop.totalGlobalMem;
            i.devices[id].warp_size  = prop.warpSize;
        }
        return i;
    }();
    return info;
}
// a caching pool like the reference's ggml_cuda_pool_leg (ggml-cuda.cu): freed buffers are kept and handed out again, so the
// Q8_0 path's scratch (quantised activations, mmq-sparse.cu:339) is allocated in the eager warm-up and only REUSED while the
// stream is being captured into a CUDA graph
struct refcuda_pool : public ggml_cuda_pool {
    std::vector<std::pair<void *, size_t>> cache;
    void * alloc(size_t size, size_t * actual_size) override {
        int best = -1;
        for (int i = 0; i < (int) cache.size(); ++i) {
            if (cache[i].second >= size && (best < 0 || cache[i].second < cache[best].second)) {
                best = i;
            }
        }
        if (best >= 0) {
            void * p     = cache[best].first;
            *actual_size = cache[best].second;
            cache.erase(cache.begin() + best);
            return p;
        }
        void * p = nullptr;
        CUDA_CHECK(cudaMalloc(&p, size));
        *actual_size = size;
        return p;
    }
    void free(void * ptr, size_t size) override { cache.emplace_back(ptr, size); }
    ~refcuda_pool() {
        for (auto & c : cache) {
            cudaFree(c.first);
        }
    }
};
std::unique_ptr<ggml_cuda_pool> ggml_backend_cuda_context::new_pool_for_device(int, int) {
    return std::unique_ptr<ggml_cuda_pool>(new refcuda_pool());
}
ggml_backend_cuda_context::~ggml_backend_cuda_context() {}

// ---- tensor descriptors over caller memory ---------------------------------------------------------------------------
static ggml_tensor mk(ggml_type type, int64_t ne0, int64_t ne1, void * data) {
    ggml_tensor t;
    memset(&t, 0, sizeof(t));
    t.type  = type;
    t.ne[0] = ne0, t.ne[1] = ne1, t.ne[2] = 1, t.ne[3] = 1;
    t.nb[0] = ggml_type_size(type);
    t.nb[1] = ggml_row_size(type, ne0);
    t.nb[2] = t.nb[1] * ne1;
    t.nb[3] = t.nb[2];
    t.data  = data;
    return t;
}

struct Step {
    ggml_backend_cuda_context * ctx;
    int arch, wtype, ptype, L, n_embd, n_ff, r;
    void ** w_up; void ** w_gate; void ** w_down; void ** wpu; void ** wpd; float ** bpd;
    float * x;      // [L][n_embd]
    float * y;      // [L][n_embd]
    float * p;      // [L][n_ff]   sparse_idx per layer
    float * h1;     // [r]
    float * up; float * gate; float * hid;  // [n_ff] each
};

// build_predictor (llama-graph.cpp:865-894): MUL_MAT, [ADD], RELU, MUL_MAT, [ADD], SIGMOID
static void predictor(Step & s, int l, const float * x) {
    ggml_tensor wpu = mk((ggml_type) s.ptype, s.n_embd, s.r, s.wpu[l]);
    ggml_tensor wpd = mk((ggml_type) s.ptype, s.r, s.n_ff, s.wpd[l]);
    ggml_tensor xt  = mk(GGML_TYPE_F32, s.n_embd, 1, (void *) x);
    ggml_tensor h   = mk(GGML_TYPE_F32, s.r, 1, s.h1);
    ggml_tensor z   = mk(GGML_TYPE_F32, s.n_ff, 1, s.p + (size_t) l * s.n_ff);
    ggml_cuda_mul_mat_vec_f(*s.ctx, &wpu, &xt, nullptr, &h, nullptr);
    ggml_tensor hr = h;
    hr.src[0]      = &h;
    ggml_cuda_op_relu(*s.ctx, &hr);  // in place
    ggml_tensor                    bias = mk(GGML_TYPE_F32, s.n_ff, 1, s.bpd ? s.bpd[l] : nullptr);
    ggml_cuda_mm_fusion_args_host fus  = {};
    fus.x_bias                          = &bias;
    ggml_cuda_mul_mat_vec_f(*s.ctx, &wpd, &h, nullptr, &z, (s.bpd && s.bpd[l]) ? &fus : nullptr);
    ggml_tensor zs = z;
    zs.src[0]      = &z;
    ggml_cuda_op_sigmoid(*s.ctx, &zs);  // in place
}

// build_sparse_ffn, gpu_only (llama-graph.cpp:896-1142): MUL_MAT_SPARSE(up), MUL_MAT_SPARSE(gate), activation, AXPY_SPARSE
static void sparse_ffn(Step & s, int l) {
    cudaStream_t st   = s.ctx->stream();
    float *      x    = s.x + (size_t) l * s.n_embd;
    ggml_tensor  wu   = mk((ggml_type) s.wtype, s.n_embd, s.n_ff, s.w_up[l]);
    ggml_tensor  wg   = mk((ggml_type) s.wtype, s.n_embd, s.n_ff, s.w_gate ? s.w_gate[l] : nullptr);
    ggml_tensor  wd   = mk((ggml_type) s.wtype, s.n_embd, s.n_ff, s.w_down[l]);
    ggml_tensor  xt   = mk(GGML_TYPE_F32, s.n_embd, 1, x);
    ggml_tensor  idx  = mk(GGML_TYPE_F32, s.n_ff, 1, s.p + (size_t) l * s.n_ff);
    ggml_tensor  up   = mk(GGML_TYPE_F32, s.n_ff, 1, s.up);
    ggml_tensor  gate = mk(GGML_TYPE_F32, s.n_ff, 1, s.gate);
    ggml_tensor  hid  = mk(GGML_TYPE_F32, s.n_ff, 1, s.hid);
    ggml_tensor  yt   = mk(GGML_TYPE_F32, s.n_embd, 1, s.y + (size_t) l * s.n_embd);
    const bool   q8   = s.wtype == GGML_TYPE_Q8_0;
    auto mms = [&](ggml_tensor & w, ggml_tensor & dst) {
        dst.src[0] = &w, dst.src[1] = &xt, dst.src[2] = &idx, dst.src[3] = nullptr;
        if (q8) {
            ggml_cuda_mul_mat_sparse_q(*s.ctx, &w, &xt, &dst);
        } else {
            ggml_cuda_op_mul_mat_sparse(*s.ctx, &w, &xt, &dst, (const char *) w.data, (const float *) xt.data, nullptr,
                                        (float *) dst.data, 0, s.n_ff, 1, s.n_embd, st);
        }
    };
    mms(wu, up);
    const bool gated = s.arch != 2;
    if (gated) {
        mms(wg, gate);
    }
    // activation recipe per arch (llama-graph.cpp:1063-1094)
    if (s.arch == 0) {          // ProSparse: fatrelu(gate, 0.01) * up
        ggml_tensor g2 = gate;
        g2.src[0]      = &gate;
        const float t  = 0.01f;
        memcpy(g2.op_params, &t, sizeof(float));
        ggml_cuda_op_fatrelu(*s.ctx, &g2);
        hid.src[0] = &gate, hid.src[1] = &up;
        ggml_cuda_op_mul(*s.ctx, &hid);
    } else if (s.arch == 1) {   // Bamboo / Qwen2: relu(gate) * relu(up)
        ggml_tensor g2 = gate, u2 = up;
        g2.src[0] = &gate, u2.src[0] = &up;
        ggml_cuda_op_relu(*s.ctx, &g2);
        ggml_cuda_op_relu(*s.ctx, &u2);
        hid.src[0] = &gate, hid.src[1] = &up;
        ggml_cuda_op_mul(*s.ctx, &hid);
    } else {                    // OPT / Falcon: relu(up)
        hid.src[0] = &up;
        ggml_cuda_op_relu(*s.ctx, &hid);
    }
    yt.src[0] = &wd, yt.src[1] = &hid, yt.src[2] = &idx, yt.src[3] = nullptr;
    if (q8) {
        ggml_cuda_op_axpy_sparse_q(*s.ctx, &wd, &hid, &yt, (const char *) wd.data, (const float *) hid.data, nullptr,
                                   (float *) yt.data, 0, s.n_ff, 1, s.n_ff, st);
    } else {
        ggml_cuda_op_axpy_sparse(*s.ctx, &wd, &hid, &yt, (const char *) wd.data, (const float *) hid.data, nullptr,
                                 (float *) yt.data, 0, s.n_ff, 1, s.n_ff, st);
    }
}

static void step(Step & s, bool with_pred) {
    if (with_pred) {
        predictor(s, 0, s.x);
    }
    for (int l = 0; l < s.L; ++l) {
        if (with_pred && l + 1 < s.L) {
            predictor(s, l + 1, s.x + (size_t) l * s.n_embd);  // look-ahead (llama-graph.cpp:939-946)
        }
        sparse_ffn(s, l);
    }
}

extern "C" {

const char * spif_refcuda_describe(void) {
    return "reference CUDA kernels (mm-sparse.cu, axpy-sparse.cu, mmq-sparse.cu, axpyq-sparse.cu, mmvf.cu, unary.cu, "
           "binbcast.cu) compiled unmodified for sm_100";
}

// One decode step = predictor + sparse FFN of L layers in the reference's node order on one stream.
// All pointers are DEVICE pointers except the arrays of pointers and out_us.  p_io: [L][n_ff] masks -- written by the
// predictor when with_pred != 0, otherwise read as given.  out_us[0] = CUDA-graph replay, us per step (median of iters);
// out_us[1] = eager launches on the stream, us per step.  Returns 0, or -1 without a device.
int spif_refcuda_stack(int arch, int wtype, int ptype, int L, void ** w_up, void ** w_gate, void ** w_down, void ** wpu,
                       void ** wpd, float ** bpd, float * x, float * y, float * p_io, int n_embd, int n_ff, int r,
                       int with_pred, int warmup, int iters, double * out_us) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        return -1;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    ggml_backend_cuda_context ctx(dev);
    Step s = {};
    s.ctx = &ctx, s.arch = arch, s.wtype = wtype, s.ptype = ptype, s.L = L, s.n_embd = n_embd, s.n_ff = n_ff, s.r = r;
    s.w_up = w_up, s.w_gate = w_gate, s.w_down = w_down, s.wpu = wpu, s.wpd = wpd, s.bpd = bpd;
    s.x = x, s.y = y, s.p = p_io;
    float * scratch = nullptr;
    CUDA_CHECK(cudaMalloc(&scratch, sizeof(float) * ((size_t) 3 * n_ff + r)));
    s.up = scratch, s.gate = scratch + n_ff, s.hid = scratch + 2 * (size_t) n_ff, s.h1 = scratch + 3 * (size_t) n_ff;
    cudaStream_t st = ctx.stream();
    cudaEvent_t  e0, e1;
    CUDA_CHECK(cudaEventCreate(&e0));
    CUDA_CHECK(cudaEventCreate(&e1));
    auto median = [](std::vector<float> & v) {
        std::sort(v.begin(), v.end());
        return (double) v[v.size() / 2] * 1e3;
    };
    // eager
    for (int i = 0; i < std::max(warmup, 1); ++i) {
        step(s, with_pred != 0);
    }
    CUDA_CHECK(cudaStreamSynchronize(st));
    std::vector<float> te, tg;
    for (int i = 0; i < iters; ++i) {
        CUDA_CHECK(cudaEventRecord(e0, st));
        step(s, with_pred != 0);
        CUDA_CHECK(cudaEventRecord(e1, st));
        CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        te.push_back(ms);
    }
    // CUDA graph of the same launches (the reference decodes through CUDA graphs when built with GGML_CUDA_GRAPHS)
    cudaGraph_t     g  = nullptr;
    cudaGraphExec_t ge = nullptr;
    CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    step(s, with_pred != 0);
    CUDA_CHECK(cudaStreamEndCapture(st, &g));
    CUDA_CHECK(cudaGraphInstantiate(&ge, g, 0));
    for (int i = 0; i < std::max(warmup, 1); ++i) {
        CUDA_CHECK(cudaGraphLaunch(ge, st));
    }
    CUDA_CHECK(cudaStreamSynchronize(st));
    for (int i = 0; i < iters; ++i) {
        CUDA_CHECK(cudaEventRecord(e0, st));
        CUDA_CHECK(cudaGraphLaunch(ge, st));
        CUDA_CHECK(cudaEventRecord(e1, st));
        CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        tg.push_back(ms);
    }
    out_us[0] = median(tg);
    out_us[1] = median(te);
    CUDA_CHECK(cudaGraphExecDestroy(ge));
    CUDA_CHECK(cudaGraphDestroy(g));
    CUDA_CHECK(cudaEventDestroy(e0));
    CUDA_CHECK(cudaEventDestroy(e1));
    CUDA_CHECK(cudaFree(scratch));
    return 0;
}

}  // extern "C"
