// spif_ffn_q.cu -- Q8_0 / Q4_0 variants of the sparse-FFN hot path (dequantisation fused into the dot products).
#include "spif_common.cuh"

int spif_launch_ffn_q(const spif_ffn_args & a, int mode, const float * hid_in, float * dst_dense, void * ws,
                      size_t ws_bytes, const spif_launch_ctx & ctx) {
    (void) a; (void) mode; (void) hid_in; (void) dst_dense; (void) ws; (void) ws_bytes; (void) ctx;
    return SPIF_E_UNSUPPORTED;
}
