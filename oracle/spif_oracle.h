/* oracle/spif_oracle.h -- TEST INFRASTRUCTURE ONLY (see spif_oracle.c). Never linked into the product. */
#ifndef SPIF_ORACLE_H
#define SPIF_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ggml type ids (ggml/include/ggml.h:385-415) */
enum { SPIFO_F32 = 0, SPIFO_F16 = 1, SPIFO_Q4_0 = 2, SPIFO_Q8_0 = 8, SPIFO_BF16 = 30 };
/* arch activation recipes (src/llama-graph.cpp:1063-1094) */
enum { SPIFO_ARCH_PROSPARSE = 0, SPIFO_ARCH_RELU_GATED = 1, SPIFO_ARCH_RELU_UP = 2,
       SPIFO_ARCH_SILU_GATED = 3 /* Llama-3 / top-k config: no reference code path, parity unpinned */ };

const char * spifo_describe(void);
size_t       spifo_row_size(int type, int64_t n);

/* storage conversions: how synthetic weights are made (ggml-quants.c:36,199; ggml-impl.h:401,550) */
void spifo_fp32_to_fp16_row(const float * x, uint16_t * y, int64_t n);
void spifo_fp16_to_fp32_row(const uint16_t * x, float * y, int64_t n);
void spifo_fp32_to_bf16_row(const float * x, uint16_t * y, int64_t n);
int  spifo_quantize_rows(int type, const float * src, void * dst, int64_t nrows, int64_t n_per_row);
int  spifo_dequantize_row(int type, const void * src, float * dst, int64_t n);

/* activation quantiser used by the CPU backend for Q8_0/Q4_0 weights (arch/x86/quants.c:290, AVX2 branch) */
void spifo_quantize_row_q8_0_act(const float * x, void * y, int64_t n);

/* mask -> ascending index list {i : !(sparse_idx[i] < thr)}  (mm-sparse.cu:22, ggml-cpu.c:1775). returns count */
int spifo_mask_compact(const float * sparse_idx, int n_ff, float thr, int32_t * idx_out);

/* top-k mask (config 4; NO reference implementation -- parity unpinned): the k largest scores, ties to the
 * lower index; writes sparse_idx_out[i] = 1.0f for selected, 0.0f otherwise. returns k' = min(k, n) */
int spifo_topk_mask(const float * score, int n_ff, int k, float * sparse_idx_out);

/* GGML_OP_MUL_MAT_SPARSE. W has m rows; row -> neuron via neuron_idx (NULL: identity, m == n_ff).
 * neuron_mask (nullable) is the CPU-side "lives on GPU, skip" mask (ggml-cpu.c:1775). dst {n_ff, n_tok} zero-filled. */
int spifo_mul_mat_sparse(int type, const void * w, int m, const int32_t * neuron_idx, const int32_t * neuron_mask,
                         const float * x, const float * sparse_idx, float * dst, int n_embd, int n_ff, int n_tok,
                         float thr);

/* GGML_OP_AXPY_SPARSE with n_threads = 1 summation order (ggml-cpu.c:2178-2337). dst {n_embd, n_tok}. */
int spifo_axpy_sparse(int type, const void * wt, int m, const int32_t * neuron_idx, const int32_t * neuron_mask,
                      const float * hid, const float * sparse_idx, float * dst, int n_embd, int n_ff, int n_tok,
                      float thr);

/* build_predictor (src/llama-graph.cpp:865-894). logits_out (nullable) = pre-sigmoid values. */
int spifo_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                    const float * x, float * sparse_idx_out, float * logits_out, int n_embd, int r, int n_ff,
                    int n_tok);

/* elementwise recipe between the two sparse ops (src/llama-graph.cpp:1049-1094) */
void spifo_activation(int arch, const float * up, const float * gate, const float * b_up, const float * b_gate,
                      float * hid, int n_ff, int n_tok);

/* the whole layer, O-cpu semantics (x->F16/Q8_0 rounding, alpha->F16 rounding, 1-thread order) */
int spifo_sparse_ffn(int arch, int type, const void * w_up, const void * w_gate, const void * w_down_t,
                     const float * b_up, const float * b_gate, const float * b_down, const float * x,
                     const float * sparse_idx, float * y, float * o_up, float * o_gate, float * o_hid, int n_embd,
                     int n_ff, int n_tok, float thr);

/* O-f64: same math on the same stored weights in double, F32 x and alpha un-rounded ("truth" for error budgets) */
int spifo_sparse_ffn_f64(int arch, int type, const void * w_up, const void * w_gate, const void * w_down_t,
                         const float * b_up, const float * b_gate, const float * b_down, const float * x,
                         const float * sparse_idx, double * y, double * o_hid, int n_embd, int n_ff, int n_tok,
                         float thr);

/* multi-threaded (OpenMP) port of the per-token FFN stack for the timed CPU baseline when _ref is absent.
 * returns median microseconds per step. Same layout as spif_ref_ffn_stack_time. */
double spifo_ffn_stack_time(int arch, int wtype, int ptype, int n_layers, const void * const * w_up,
                            const void * const * w_gate, const void * const * w_down_t, const void * const * wpu,
                            const void * const * wpd, const float * const * bpd, const float * x, float * y,
                            int n_embd, int n_ff, int r, int n_tok, int n_threads, int warmup, int iters);

#ifdef __cplusplus
}
#endif
#endif
