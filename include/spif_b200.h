/* spif_b200.h -- C-ABI of the B200-native (sm_100a) sparse-FFN kernels.
 *
 * This is the drop-in boundary for SparkInfer's activation-sparse FFN path: every entry point below is what the
 * reference's ggml CUDA backend would bind in place of its own kernels.  Citations are file:line in the reference
 * tree (ganminghao/SparkInfer).  INTEGRATION.md shows the glue a maintainer adds to ggml-cuda.cu.
 *
 * Conventions (identical to the reference's op contract):
 *   - all pointers are DEVICE pointers unless a name ends in _host; the callee borrows, never frees or retains;
 *   - weights are contiguous row-major, one row per neuron, nb[1] = ggml_row_size(type, n_embd), for ALL THREE
 *     matrices (W_down is stored transposed, src/llama-model.cpp:2763); rows must be 16-byte aligned;
 *   - activations / mask / outputs are F32, contiguous, token-major (x + n_embd*tok, dst + n_ff*tok);
 *   - `type` uses ggml's type ids (ggml/include/ggml.h:385-415);
 *   - a neuron is active for token t  <=>  !(sparse_idx[t][neu] < thr)   (mm-sparse.cu:22; NaN counts as active);
 *   - `neuron_idx` (nullable, I32[m]) maps weight row -> original neuron id exactly like the reference's GPU cache
 *     (mm-sparse.cu:20): NULL means identity and m == n_ff.  The same mechanism expresses a GPU *shard*;
 *   - work is enqueued on `stream`; no allocation, no implicit synchronisation, no default-stream use, re-entrant;
 *     scratch comes from the caller (`ws`, sized by spif_workspace_size) so CUDA-graph capture is legal;
 *   - every function returns 0 on success, a negative SPIF_E* code for argument errors, or a positive cudaError_t.
 *     The glue turns non-zero into GGML_ABORT (the reference aborts on the same conditions).
 * There is NO CPU fallback: without a CUDA device every compute entry point returns an error.
 *
 * Environment: the library reads a handful of SPIF_* variables, once per process, as EXPERIMENT switches between kernel
 * variants that all compute the same results (SPIF_LAYER_IMPL, SPIF_LAYER_L2_HINT, SPIF_Q_RING, SPIF_QW_HALF, SPIF_DIRECT,
 * SPIF_DIRECT_SMALL, SPIF_RED_FIRST, SPIF_FFN_CTAS, SPIF_PDL_LATE, SPIF_DG_WARPS, SPIF_DG_VL, SPIF_DG_ASYNC, SPIF_CTAS_PER_SM,
 * SPIF_STAGES, SPIF_SMEM_SPLIT, SPIF_GU_PCT, SPIF_DSLOTS, SPIF_TOPK_CLUSTER) and two debug switches that do NOT
 * (SPIF_LAYER_DEBUG, SPIF_BATCH_DEBUG: timing floors with the arithmetic or the loads skipped).  Defaults are the measured
 * best; DESIGN.md section 4 lists meaning and default of each.  None is needed in production.  The Python host reads two
 * more: SPIF_COMM_MULTICAST=0 (ShardComm: leave spif_shard_comm.mc NULL, unicast exchange stores) and SPIF_BENCH_VERBOSE.
 */
#ifndef SPIF_B200_H
#define SPIF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* bumped when a struct layout or a signature changes; additions of entry points do not bump it.
 * 3 -> 4: spif_shard_comm gained `mc` (multicast mapping of the exchange buffers) at its end */
#define SPIF_B200_ABI_VERSION 4

typedef struct CUstream_st * spif_stream_t; /* == cudaStream_t */

enum spif_type { /* ggml_type values */
    SPIF_TYPE_F32  = 0,
    SPIF_TYPE_F16  = 1,
    SPIF_TYPE_Q4_0 = 2,
    SPIF_TYPE_Q8_0 = 8,
    SPIF_TYPE_BF16 = 30,
};

enum spif_arch { /* activation recipe between gate/up and down, src/llama-graph.cpp:1063-1094 */
    SPIF_ARCH_PROSPARSE  = 0, /* h = fatrelu(gate, 0.01) * up      (LLM_ARCH_PROSPARSE_LLAMA) */
    SPIF_ARCH_RELU_GATED = 1, /* h = relu(gate) * relu(up)         (BAMBOO, QWEN2)            */
    SPIF_ARCH_RELU_UP    = 2, /* h = relu(up), no gate matrix      (OPT, FALCON)              */
    SPIF_ARCH_SILU_GATED = 3, /* h = silu(gate) * up -- Llama-3 top-k config; no reference code path */
};

enum spif_flags {
    /* Round x to the weight float type before the dot and the hidden value before the axpy, like the reference CPU
     * backend does (ggml-cpu.c:1832-1856, 2196-2207).  Default (0) keeps both in F32 like the reference CUDA kernels
     * (mm-sparse.cu:55-57, axpy-sparse.cu:44).  Only meaningful for F16/BF16 weights. */
    SPIF_FLAG_CPU_ROUNDING = 1 << 0,
    /* sparse_idx is indexed by weight row (shard-local mask of length m) instead of by neuron id */
    SPIF_FLAG_MASK_BY_ROW  = 1 << 1,
    /* (bit 2 is reserved: round 1 declared a float-atomics down projection here that no kernel implemented; every
     * down projection is the deterministic fixed-order sum, unlike axpy-sparse.cu:84) */
    /* Caller's promise (spif_sparse_ffn only): sparse_idx was produced at least TWO kernel launches before this call on
     * `stream` (or is otherwise known to be complete), as in the reference's look-ahead order where the mask of layer l
     * is predicted one whole layer earlier (src/llama-graph.cpp:939-946).  The kernel then scans the mask and starts
     * streaming weight rows BEFORE its programmatic-dependent-launch wait, i.e. under the tail of the previous kernel;
     * only x, y and the workspace wait.  Without the flag everything waits (always safe). */
    SPIF_FLAG_MASK_STABLE  = 1 << 3,
    /* spif_sparse_ffn with n_tok >= 2: keep the reference's per-(row, token) schedule (mm-sparse.cu:107-210: every token
     * re-streams its active rows) instead of the default union-of-masks gathered GEMM on tensor cores. */
    SPIF_FLAG_PER_TOKEN    = 1 << 4,
};

enum spif_error {
    SPIF_OK            = 0,
    SPIF_E_TYPE        = -1, /* unsupported weight type for this op (reference: GGML_ABORT "unsupported type") */
    SPIF_E_SHAPE       = -2, /* shape/alignment precondition violated (reference: GGML_ASSERT) */
    SPIF_E_NULL        = -3, /* required pointer missing (reference: "missing sparse_idx") */
    SPIF_E_WORKSPACE   = -4, /* ws too small / misaligned */
    SPIF_E_UNSUPPORTED = -5,
    SPIF_E_NO_DEVICE   = -6,
    SPIF_E_STATIC_LIMIT = -7, /* spif_sparse_ffn_noalloc: the shape needs more scratch than the library's static array holds */
};

int          spif_abi_version(void);
const char * spif_error_string(int code);
/* number of SMs etc. of the current device; returns an error without a GPU */
int          spif_device_sm_count(int * sm_count);

size_t spif_row_size(int type, int64_t n_embd); /* == ggml_row_size */

/* Scratch needed by any entry point below for these sizes (bytes; 256-byte aligned pointer required). */
size_t spif_workspace_size(int n_embd, int n_ff, int n_tok);

/* ---- mask -> compact active-index list -------------------------------------------------------------------------
 * The reference never materialises the list (every kernel re-tests the float mask, mm-sparse.cu:22, axpy-sparse.cu:53);
 * this is the north-star's "threshold mask + warp-ballot/prefix-scan compaction".  For each token t:
 *   idx_out[t*m .. t*m+count_out[t])  = ascending weight rows r whose neuron is active for t.
 * If union_idx/union_count are non-NULL they receive the same for "active for ANY token" (batched path).       */
int spif_mask_compact(const float * sparse_idx, const int32_t * neuron_idx, int m, int n_ff, int n_tok, float thr,
                      int flags, int32_t * idx_out, int32_t * count_out, int32_t * union_idx, int32_t * union_count,
                      spif_stream_t stream);

/* ---- top-k mask (Llama-3 / SiLU config; no reference code) ----------------------------------------
 * sparse_idx_out[t][i] = 1.0f for the k largest score[t][i] (ties -> lower index), else 0.0f.                    */
int spif_topk_mask(const float * score, int n_ff, int n_tok, int k, float * sparse_idx_out, spif_stream_t stream);

/* ---- GGML_OP_MUL_MAT_SPARSE (ggml.c:3310-3331; CUDA mm-sparse.cu:366-431, mmq-sparse.cu:293-394) ------------------
 * dst[t][neu] = dot(W[row], x[t]) for active rows, 0 elsewhere (dst {n_ff, n_tok} is zero-filled by the callee).
 * Types: F16, BF16, Q8_0 (reference) and Q4_0 (claimed by supports_op, ggml-cuda.cu:4594-4663, never implemented). */
int spif_mul_mat_sparse(int type, const void * w, const float * x, const float * sparse_idx,
                        const int32_t * neuron_idx, float * dst, int n_embd, int n_ff, int m, int n_tok, float thr,
                        int flags, void * ws, size_t ws_bytes, spif_stream_t stream);

/* ---- GGML_OP_AXPY_SPARSE (ggml.c:3333-3356; CUDA axpy-sparse.cu:139-204, axpyq-sparse.cu:135-185) -----------------
 * dst[t][:] = sum over active rows with hid[t][neu] != 0 of hid[t][neu] * Wt[row][:]   (dst {n_embd, n_tok}).
 * Deterministic (fixed summation order), unlike the reference's float atomics (axpy-sparse.cu:84).               */
int spif_axpy_sparse(int type, const void * wt, const float * hid, const float * sparse_idx,
                     const int32_t * neuron_idx, float * dst, int n_embd, int n_ff, int m, int n_tok, float thr,
                     int flags, void * ws, size_t ws_bytes, spif_stream_t stream);

/* ---- GGML_OP_FATRELU / GGML_UNARY-style helpers next to the path (unary.cu:564-652) ----------------------------- */
int spif_fatrelu(const float * x, float * y, int64_t n, float threshold, spif_stream_t stream);
int spif_shifted_step(const float * x, float * y, int64_t n, float threshold, spif_stream_t stream);

/* ---- predictor (build_predictor, src/llama-graph.cpp:865-894: six stock nodes in the reference) ------------------
 * sparse_idx_out[t][i] = sigmoid(Wpd[i] . relu(Wpu . x[t] + bpu) + bpd[i]).  Wpu {n_embd, r}, Wpd {r, n_ff} row-major
 * F16/BF16; biases nullable.  logits_out (nullable) receives the pre-sigmoid values (flip accounting).
 * flags: SPIF_FLAG_CPU_ROUNDING rounds x and the hidden vector to the weight type like the CPU backend's mul_mat.   */
int spif_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                   const float * x, float * sparse_idx_out, float * logits_out, int n_embd, int r, int n_ff, int n_tok,
                   int flags, void * ws, size_t ws_bytes, spif_stream_t stream);

/* ---- whole layer: build_sparse_ffn (src/llama-graph.cpp:896-1142), GPU-resident rows ----------------------------
 * n_tok == 1: one persistent kernel (+ fixed-order partial sum).  2 <= n_tok <= 32 with F16/BF16 weights and
 * n_embd % 512 == 0: the union of the per-token masks becomes a gathered GEMM on the tensor cores (tcgen05, accumulators
 * in tensor memory) -- every union row is read once instead of once per token as in mm-sparse.cu:107-210; x and the
 * hidden values are rounded to the weight float type there (the reference CPU backend's semantics; F16 hidden values beyond
 * +-65504 saturate instead of becoming inf -- the reference CUDA kernels keep them in F32 and stay finite).  Q8_0 / Q4_0 take the
 * same union GEMM (rows dequantised into F16 tiles) from 13 tokens on; up to 12 tokens, other shapes and
 * SPIF_FLAG_PER_TOKEN use the per-(row, token) schedule (quantised: one CTA row per token, 8-bit activations like mmq-sparse.cu).  stats_out[t] = {active, non-zero hidden} per token; the
 * union path also reports {.., .., union rows, union rows with a non-zero hidden value}.
 * Equivalent to MUL_MAT_SPARSE(up), MUL_MAT_SPARSE(gate), [+bias], activation, AXPY_SPARSE(down), [+bias] on the rows
 * this device holds; the place to call it from is the fusion block of the node loop (ggml-cuda.cu:3578-3870).       */
typedef struct spif_ffn_args {
    int32_t arch;   /* enum spif_arch */
    int32_t type;   /* enum spif_type of the three FFN matrices */
    int32_t n_embd;
    int32_t n_ff;   /* length of the mask / of a dense hidden vector */
    int32_t m;      /* rows held here (== n_ff when neuron_idx == NULL) */
    int32_t n_tok;
    int32_t flags;  /* enum spif_flags */
    float   thr;    /* SPIF_SPARSE_THRESHOLD, default 0.5f */
    const void *    w_up;       /* {n_embd, m} */
    const void *    w_gate;     /* {n_embd, m}; NULL for SPIF_ARCH_RELU_UP */
    const void *    w_down_t;   /* {n_embd, m} (transposed down) */
    const float *   b_up;       /* nullable, indexed by neuron id */
    const float *   b_gate;     /* nullable */
    const float *   b_down;     /* nullable {n_embd}; added once (pass NULL on all but one shard) */
    const float *   x;          /* {n_embd, n_tok} */
    const float *   sparse_idx; /* {n_ff, n_tok} (or {m, n_tok} with SPIF_FLAG_MASK_BY_ROW) */
    const int32_t * neuron_idx; /* nullable I32[m] */
    float *         y;          /* out {n_embd, n_tok} */
    float *         hidden_out; /* nullable out, dense {n_ff, n_tok}: hid for active rows held here, 0 elsewhere */
    int32_t *       stats_out;  /* nullable out I32[n_tok][4]: {n_active, n_nonzero_hidden, 0, 0} */
} spif_ffn_args;

int spif_sparse_ffn(const spif_ffn_args * args, void * ws, size_t ws_bytes, spif_stream_t stream);

/* ---- neuron-sharded layers: the per-layer exchange (the reference is single-GPU, src/llama-sparkinfer.cpp:129)
 * Each GPU holds the rows of a subset of neurons and produces a partial y; the sum over GPUs is a one-shot exchange over
 * NVLink peer memory: every rank stores its chunk, as 8-byte {value, sequence number} words, into its slot of every peer's
 * exchange buffer; the receiver polls the words themselves (no flag, no fence, one NVLink hop) and adds the slots in RANK
 * ORDER (bit-identical on all ranks, run-to-run stable).  In spif_sparse_ffn_sharded / spif_sparse_ffn_lookahead this is the
 * epilogue of the kernel that sums the per-CTA partials (one kernel does both); spif_shard_allreduce is the stand-alone form.
 * Buffers are owned by the caller: data[r] is rank r's buffer as mapped into THIS process (CUDA IPC / VMM / torch symmetric
 * memory), zero-initialised before first use; seq is LOCAL device memory, zero-initialised before first use.
 * The sequence number is one counter per comm object and per call (seq[0]), advanced by the last block of every call, so
 * the stand-alone, the partial-sum-fused and the layer-fused forms can be mixed freely on one comm as long as every rank
 * issues the same sequence of calls, stream-ordered.  A peer that does not answer within the spin bound (several seconds)
 * makes the affected outputs NaN and sets a sticky error word that spif_shard_comm_status() returns: a wrong sum is never
 * returned silently.  world == 1 is legal (no peers; the look-ahead layer kernel needs a comm for its own control words).   */
#define SPIF_COMM_MAX_WORLD  8
#define SPIF_COMM_SEQ_WORDS  1024
typedef struct spif_shard_comm {
    int32_t rank, world;     /* world <= SPIF_COMM_MAX_WORLD */
    int32_t cap_floats;      /* capacity of one slot in floats: >= n_tok * n_embd (+ pred rank r for the look-ahead layer), multiple of 4 */
    int32_t reserved;
    void *  data[SPIF_COMM_MAX_WORLD];  /* 8-byte words [2][world][cap_floats]  (spif_shard_comm_data_bytes), 16-byte aligned */
    void *  seq;                        /* local uint32_t [SPIF_COMM_SEQ_WORDS], zero-initialised: [0] calls completed, [2] error */
    void *  mc;                         /* nullable: a MULTICAST mapping of the same buffers (NVLS: cuMulticastBindMem / torch symmetric
                                           memory's multicast_ptr).  One store through it lands in every rank's copy: the kernels then
                                           issue one store per word instead of one per peer.  NULL: unicast stores to data[r]. */
} spif_shard_comm;
size_t spif_shard_comm_data_bytes(int world, int cap_floats);
/* y[0..n) = sum over ranks of y, in place; n % 4 == 0, n <= cap_floats */
int spif_shard_allreduce(float * y, int64_t n, const spif_shard_comm * comm, spif_stream_t stream);
/* spif_sparse_ffn on this rank's rows + the exchange: args->y receives the FULL layer output on every rank */
int spif_sparse_ffn_sharded(const spif_ffn_args * args, const spif_shard_comm * comm, void * ws, size_t ws_bytes,
                            spif_stream_t stream);
/* synchronises `stream`, then *error_out = the comm's sticky error word (0 = every exchange completed; 1 = a peer timed out) */
int spif_shard_comm_status(const spif_shard_comm * comm, int * error_out, spif_stream_t stream);

/* ---- look-ahead layer: sparse FFN of layer l + predictor of layer l+1 in ONE launch ------------------------------------
 * The reference builds the predictor of layer l+1 from layer l's FFN input while layer l runs (src/llama-graph.cpp:932-946):
 * both read the same x.  spif_sparse_ffn_lookahead runs, in one persistent kernel, build_sparse_ffn of layer l on the rows
 * this device holds (`ffn`, n_tok == 1, F16/BF16) and build_predictor of layer l+1 (`pred`), and hands the next launch the
 * ACTIVE-ROW LIST of layer l+1 (`list_out`, warp-ballot compaction per block of rows) so that no kernel ever scans the float
 * mask again.  ffn == NULL: predictor only (layer 0's own mask).  pred == NULL: FFN only (last layer).
 *   - `list_in` / `list_out`: opaque device buffers of spif_active_list_bytes(m) bytes, 16-byte aligned; produced by this
 *     call, or from a float mask by spif_mask_to_list (mask indexed by ROW, like SPIF_FLAG_MASK_BY_ROW);
 *   - ffn->sparse_idx is ignored (the list replaces it); ffn->flags: SPIF_FLAG_CPU_ROUNDING;
 *   - pred->sparse_idx_out / logits_out (nullable) receive sigmoid / logit per Wpd row held here; list_out is nullable
 *     (top-k configs build the mask from the logits with spif_topk_mask + spif_mask_to_list);
 *   - `comm` is REQUIRED (world 1 on a single GPU): its buffers carry the predictor's hidden vector (every shard computes
 *     r / world rows of it) and the sum of y over shards, its seq words the call counter; cap_floats >= n_embd + r;
 *   - ws as for spif_sparse_ffn.  y is the full layer output on every rank (b_down added once per rank that passes it).
 * Limits: n_embd <= 8192, r % 256 == 0, r <= n_embd, r % world == 0, pred->ptype == ffn->type; otherwise
 * SPIF_E_UNSUPPORTED and the caller uses spif_predictor + spif_sparse_ffn.                                              */
typedef struct spif_pred_args {
    int32_t ptype;   /* SPIF_TYPE_F16 / SPIF_TYPE_BF16 */
    int32_t n_embd;
    int32_t r;       /* pred_lora */
    int32_t m;       /* rows of Wpd held here == rows of the next layer's FFN matrices on this device */
    int32_t flags;   /* SPIF_FLAG_CPU_ROUNDING */
    float   thr;     /* threshold of the list: active <=> !(sigmoid < thr) */
    const void *  wpu;  /* {n_embd, r}, replicated on every shard */
    const float * bpu;  /* nullable {r} */
    const void *  wpd;  /* {r, m} */
    const float * bpd;  /* nullable {m} */
    const float * x;    /* {n_embd}; must equal ffn->x when both are given */
    float * sparse_idx_out; /* nullable {m} */
    float * logits_out;     /* nullable {m} */
} spif_pred_args;
/* Optional fusion of what stands either side of the sparse FFN in the reference's layer graph (src/models/llama.cpp:98-101,
 * 144): build_norm(ffn_inp, ffn_norm, LLM_NORM_RMS) before it -- x_n = (x * (1 / sqrtf(mean(x^2) + eps))) * w,
 * the arithmetic of ggml_compute_forward_rms_norm_f32 (ggml-cpu/ops.cpp:3730-3779: double sum, float mean) followed by
 * ggml_mul -- and ggml_add(cur, ffn_inp) after it.  With norm_w the kernel normalises ffn->x / pred->x (the RAW ffn_inp)
 * itself: the FFN of layer l and the look-ahead predictor of layer l + 1 both read the normalised vector, as in the reference;
 * with residual, y = (ffn + b_down) + residual.  Replaces three graph nodes (RMS_NORM, MUL, ADD) and their launches. */
typedef struct spif_layer_io {
    const float * norm_w;    /* nullable {n_embd}: ffn_norm weight */
    float         norm_eps;  /* f_norm_rms_eps */
    int32_t       reserved;
    const float * residual;  /* nullable {n_embd}: ffn_inp; may alias x */
} spif_layer_io;
size_t spif_active_list_bytes(int m);
int spif_mask_to_list(const float * sparse_idx, int m, float thr, void * list_out, spif_stream_t stream);
int spif_sparse_ffn_lookahead(const spif_ffn_args * ffn, const spif_pred_args * pred, const void * list_in, void * list_out,
                              const spif_shard_comm * comm, void * ws, size_t ws_bytes, spif_stream_t stream);
/* the same with the layer's RMSNorm / residual add fused in (io may be NULL = spif_sparse_ffn_lookahead) */
int spif_sparse_ffn_lookahead_ex(const spif_ffn_args * ffn, const spif_pred_args * pred, const spif_layer_io * io,
                                 const void * list_in, void * list_out, const spif_shard_comm * comm, void * ws, size_t ws_bytes,
                                 spif_stream_t stream);

/* ---- balancer helpers (build_dfr, src/llama-graph.cpp:910-930; binbcast.cu:26-42; sumcols.cu; argsort.cu) ---------
 * One fused kernel: per group of `g` consecutive neurons, delta = #(active, token) pairs, then
 *   ema != 0 : dfr = lambda*dfr + (1-lambda)*delta/(n_tok*g)        (SPIF_DFR_EMA variant, binbcast.cu:28-34)
 *   ema == 0 : dfr = lambda*dfr + delta/(n_tok*g)
 * replacing shifted_step + sum_cols + reshape + sum_rows + scale_add.  In place on dfr_scores {n_ff/g}.            */
int spif_dfr_update(const float * sparse_idx, int n_ff, int n_tok, int g, float thr, float lambda, int ema,
                    float * dfr_scores, spif_stream_t stream);

/* ---- shard placement and swap planning (host code, no device work) ------------------------------------------------
 * Replaces the planning half of the reference's cache manager (sparkinfer_reload_plan, src/llama-sparkinfer.cpp:45-91:
 * scores -> copy_pair list, ggml-sparkinfer.hpp:16-24), re-targeted from "which groups live on the GPU" to "which GPU
 * shard holds which group" (SURVEY.md 8e).  Every group is on exactly one of n_shards shards, equal group count per shard.
 *
 * spif_initial_placement: groups sorted by hotness (descending, ties by index), dealt to the shards in snake order ->
 *   shard_of_group[n_g].  n_g % n_shards == 0.
 * spif_rebalance_plan: up to max_swaps times, take the most- and the least-loaded shard (load = sum of dfr over its groups)
 *   and swap the pair of groups (one from each) that lowers the larger of the two loads most; stop when the gain is
 *   <= min_gain.  shard_of_group is updated in place; plan[4*i .. 4*i+3] = {group leaving the hot shard, group leaving the
 *   cold shard, hot shard, cold shard}; *n_swaps = number of swaps planned.  The executor moves the rows. */
int spif_initial_placement(const float * group_hotness, int n_g, int n_shards, int32_t * shard_of_group);
/* Load-time half of the cache manager (sparkinfer_cache_manager ctor, src/llama-sparkinfer.cpp:179-202, 283-323):
 * spif_cache_group_budget: groups per layer that fit `vram_budget` bytes -- budget = vram_budget / (group_bytes * (has_gate ?
 *   3 : 2)); layer l starts with min((int)(budget * pattern[l]), n_group) (pattern = the model-split file's
 *   `ffn_normalized_pattern`), the remainder is dealt round-robin, one group per layer per turn, to layers that are not full.
 *   On a B200 every layer is normally full (m == n_ff, the reference's gpu_only); smaller budgets give the GPU-cache subset
 *   that spif_mul_mat_sparse / spif_axpy_sparse address through neuron_idx.
 * spif_reorder_rows: dst row new = src row perm[new] (`blk.N.ffn_reorder_perms`, new_row -> old_row, :283-304); host
 *   memory, dst != src, row_bytes per row (a 1-D tensor is n rows of its element size).                                   */
int spif_cache_group_budget(int64_t vram_budget, int64_t group_bytes, int has_gate, const float * pattern, int n_layer,
                            int n_group, int32_t * groups_per_layer);
int spif_reorder_rows(const void * src, void * dst, const int32_t * perm, int64_t n_rows, int64_t row_bytes);
int spif_rebalance_plan(const float * dfr, int32_t * shard_of_group, int n_g, int n_shards, int max_swaps, float min_gain,
                        int32_t * plan, int * n_swaps);

/* ---- model-split file (`-ms`): consumer and producer, host code ------------------------------------------------------
 * The file the reference's cache manager loads (sparkinfer_cache_manager ctor, src/llama-sparkinfer.cpp:150-158, 257,
 * 269-276): a GGUF v3 container with KV `ffn_group_size` (i32), `ffn_normalized_pattern` (f32[n_layer]) and one tensor
 * `blk.N.ffn_reorder_perms` (I32 {n_ff}, new_row -> old_row) per layer.  spif_model_split_open parses it (unknown keys and
 * tensors are skipped); _pattern feeds spif_cache_group_budget, _perm feeds spif_reorder_rows.  The reference has no
 * producer for this file; spif_model_split_write emits one the unmodified reference loads (tests/test_model_split.py reads
 * it back with the reference's gguf_init_from_file).  SPIF_E_SHAPE = malformed / truncated file or not a permutation.  */
typedef struct spif_model_split spif_model_split;
int  spif_model_split_open(const char * path_host, spif_model_split ** out);
void spif_model_split_close(spif_model_split * ms);
int  spif_model_split_info(const spif_model_split * ms, int32_t * group_size, int32_t * n_layer, int32_t * n_ff);
int  spif_model_split_pattern(const spif_model_split * ms, float * pattern_out_host, int n_layer);
int  spif_model_split_perm(const spif_model_split * ms, int layer, int32_t * perm_out_host, int n_ff);
int  spif_model_split_write(const char * path_host, int group_size, const float * pattern_host, int n_layer,
                            const int32_t * perms_host /* [n_layer][n_ff] */, int n_ff);

/* ---- model file (GGUF): the path's hyper-parameters and tensors, host code ------------------------------------------------
 * Consumer of what the reference's loader reads for this path from the model GGUF (src/llama-model.cpp:661, 2730-2768;
 * written by convert_hf_to_gguf.py:4597-4648 -- sparkinfer_b200/convert.py is the producer here):
 *   KV      general.architecture, {arch}.block_count / embedding_length / feed_forward_length /
 *           attention.layer_norm_rms_epsilon, {arch}.pred_lora (predictor rank per layer, 0 = dense layer);
 *   tensors blk.N.ffn_up.weight, blk.N.ffn_gate.weight {n_embd, n_ff}; blk.N.ffn_down.weight TRANSPOSED {n_embd, n_ff} (one
 *           row per neuron, the w_down_t of spif_ffn_args); blk.N.ffn_pred_up.weight {n_embd, r}, blk.N.ffn_pred_down.weight
 *           {r, n_ff}, their optional .bias; blk.N.ffn_norm.weight {n_embd}.
 * GGUF v2/v3, mapped read-only; spif_gguf_read copies rows [row0, row0 + n_rows) of one tensor (a row = ne[0] elements; a 1-D
 * tensor = ne[0] rows of one element) into caller memory, so a GPU shard stages only the neuron rows it holds.  `type` is the
 * ggml type id, equal to SPIF_TYPE_* for the types the library computes in (I32 = 26 is readable too); nbytes = 0 marks a type
 * the path does not read (spif_gguf_read: SPIF_E_TYPE).  SPIF_E_SHAPE: malformed / truncated file, missing key or tensor.  */
typedef struct spif_gguf spif_gguf;
typedef struct spif_gguf_tensor {
    int32_t  type;
    int32_t  n_dims;
    int64_t  ne[4];   /* ggml order: ne[0] = row length in elements, ne[1] = rows */
    uint64_t offset;  /* of the tensor's first byte from the start of the file */
    uint64_t nbytes;
} spif_gguf_tensor;
int  spif_gguf_open(const char * path_host, spif_gguf ** out);
void spif_gguf_close(spif_gguf * g);
int  spif_gguf_hparams(const spif_gguf * g, char * arch_out, int arch_cap, int32_t * n_layer, int32_t * n_embd, int32_t * n_ff,
                       float * rms_eps);
int  spif_gguf_pred_rank(const spif_gguf * g, int32_t * ranks_out, int n_layer);
int  spif_gguf_n_tensors(const spif_gguf * g);
int  spif_gguf_tensor_at(const spif_gguf * g, int i, char * name_out, int name_cap, spif_gguf_tensor * out);
int  spif_gguf_find(const spif_gguf * g, const char * name, spif_gguf_tensor * out);
int  spif_gguf_read(const spif_gguf * g, const spif_gguf_tensor * t, int64_t row0, int64_t n_rows, void * dst_host,
                    size_t dst_bytes);
int  spif_gguf_gather(const spif_gguf * g, const spif_gguf_tensor * t, const int32_t * rows_host, int64_t n_rows, void * dst_host,
                      size_t dst_bytes); /* row rows_host[i] -> row i of dst: a model-split permutation or a shard's neuron_idx */

/* ---- online shard balancer: host-side state (plan -> exec -> anchor) ----------------------------------------------------
 * Counterpart of the reference's per-layer cache state (cache_meta, group_maps, dfr_decay_pack: src/llama-sparkinfer.h,
 * ggml/include/ggml-sparkinfer.hpp:16-24) and of the plan / exec / anchor split of ggml_sparkinfer_reload
 * (ggml/src/ggml-cuda/ggml-cuda.cu:2499-2606, lambda adaptation ggml-sparkinfer.hpp:143-175), re-targeted to GPU shards:
 * every rank holds an identical spif_balancer (the plan depends on all-gathered scores only, so no coordinator is needed).
 *   create   slot s of shard r initially holds the s-th smallest group id with shard_of_group[layer][g] == r;
 *   plan     slot_scores_host = the all-gathered decayed firing rates [n_shards][n_layer][n_slot] (spif_dfr_update output of
 *            every shard).  Per layer: spif_rebalance_plan, then the moves THIS rank executes, ordered by (round, layer) --
 *            both ends of a move see the same order.  The new tables are pending; lambda adapts: raised (x (1 + step)) when the
 *            balance the previous plan promised did not hold (mean imbalance now > promised x (1 + drift_tol)), else lowered,
 *            clamped to [0.05, 0.95].  SPIF_E_WORKSPACE: moves_cap too small, *n_moves = entries needed, nothing pending;
 *   exec     the caller moves, per move, the g rows of up / gate / down^T / predictor-down (+ bias, + score) of `slot`
 *            to `peer` and receives the peer's into the same slot (device work: NVLink send/recv or peer copies);
 *   commit   (anchor) the pending tables become current; abort drops them.                                              */
typedef struct spif_balancer spif_balancer;
typedef struct spif_move {
    int32_t layer, round;         /* swap `round` of `layer` */
    int32_t slot, peer;           /* my slot whose group leaves / arrives; the other shard */
    int32_t group_out, group_in;  /* group ids (for logging / checks) */
} spif_move;
int  spif_balancer_create(int n_layer, int n_g, int n_shards, int rank, const int32_t * shard_of_group_host /* [n_layer][n_g] */,
                          float lambda0, float lambda_step, float drift_tol, spif_balancer ** out);
void spif_balancer_destroy(spif_balancer * b);
int  spif_balancer_slots(const spif_balancer * b, int layer, int32_t * slots_out_host /* [n_shards][n_slot] -> group id */);
int  spif_balancer_state(const spif_balancer * b, float * lambda, int64_t * swaps_done, int * has_pending);
int  spif_balancer_plan(spif_balancer * b, const float * slot_scores_host, int max_swaps, float min_gain, spif_move * moves_out,
                        int moves_cap, int * n_moves, float * imbalance_before, float * imbalance_after, int * n_swaps_total);
int  spif_balancer_commit(spif_balancer * b);
int  spif_balancer_abort(spif_balancer * b);

/* ---- allocation-free single-token layer (NO counterpart in the reference) ----
 * spif_sparse_ffn for callers that can neither call cudaMalloc nor own a workspace (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid allocation): the scratch is a file-scope static __device__ array inside the library,
 * SPIF_NOALLOC_WS_BYTES (12 MiB) on every GPU that loads the library, held until the process exits.  The call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine); inputs and outputs are still the caller's (args as for
 * spif_sparse_ffn, n_tok == 1 only -- otherwise SPIF_E_UNSUPPORTED).  A shape whose spif_workspace_size(n_embd, n_ff, 1)
 * exceeds the array returns SPIF_E_STATIC_LIMIT (every BASELINE.json architecture fits: ReluLLaMA-70B needs 10.2 MiB).
 * ALL calls on a device share the one array, which keeps the previous call's contents (nothing in it has to start at zero:
 * every CTA writes its partial row and counters before they are summed).  Calls on one stream are ordered and safe; calls
 * that OVERLAP on the GPU (other streams, concurrent host threads on one device) RACE on it -- use spif_sparse_ffn with
 * separate workspaces there.  Results are bit-identical to spif_sparse_ffn on the same inputs. */
#define SPIF_NOALLOC_WS_BYTES (12u << 20)
size_t spif_noalloc_max_bytes(void);
int    spif_sparse_ffn_noalloc(const spif_ffn_args * args, spif_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SPIF_B200_H */
