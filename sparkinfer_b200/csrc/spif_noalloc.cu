// spif_noalloc.cu -- spif_sparse_ffn_noalloc: the single-token fused sparse FFN with its scratch in a STATIC device array.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA-graph capture, tight serving loops, harnesses that forbid allocation) and do not want to own a workspace either.
// The scratch of the kernels behind spif_sparse_ffn at n_tok == 1 (one partial y row + two counters per CTA, summed in a fixed
// order by spif_reduce_partials_kernel) lives in g_spif_static_ws below: SPIF_NOALLOC_WS_BYTES = 12 MiB per GPU that loads the
// library, held until the process exits.  That covers every architecture of BASELINE.json (ReluLLaMA-70B needs 10.2 MiB).
//
// In host code the NAME of a __device__ array is a host-side stand-in, not a device address: a kernel given it faults or, where
// the GPU can reach host memory, silently uses the wrong memory.  So the address is taken here, in the file that defines the
// array, once per device with that device current, by cudaGetSymbolAddress (return code checked), and that pointer is what
// the launchers receive.  This file calls no cudaMalloc / cudaFree / allocating library routine, and neither do the launchers
// it reaches (tests/test_zz_no_alloc.py runs it twice in a row against spif_sparse_ffn).
//
// Sharing: every call on a device uses the SAME array, and it keeps the previous call's contents.  Nothing in it has to start
// at zero -- every CTA writes its whole partial row and both of its counters before the summing kernel reads them -- so nothing
// is cleared; but two calls that OVERLAP on the GPU (different streams, or concurrent host threads on one device) race on it.
// Calls on one stream are ordered and safe.  Callers that need concurrency pass their own workspaces to spif_sparse_ffn.
#include "spif_common.cuh"

#include <atomic>

static __device__ __align__(256) unsigned char g_spif_static_ws[SPIF_NOALLOC_WS_BYTES];

// device address of the array per device ordinal; benign race: every writer stores the same value
static std::atomic<void *> g_ws_addr[64];

static int static_ws_ptr(void ** out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        return SPIF_E_NO_DEVICE;
    }
    SPIF_REQUIRE(dev >= 0 && dev < 64, SPIF_E_UNSUPPORTED);
    void * p = g_ws_addr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        // set-up, once per device, with that device current (it is: cudaGetDevice above)
        const cudaError_t e = cudaGetSymbolAddress(&p, g_spif_static_ws);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? SPIF_E_NO_DEVICE : (int) e;
        }
        SPIF_REQUIRE(p != nullptr && ((uintptr_t) p & 255) == 0, SPIF_E_WORKSPACE);
        g_ws_addr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

extern "C" size_t spif_noalloc_max_bytes(void) {
    return (size_t) SPIF_NOALLOC_WS_BYTES;
}

extern "C" int spif_sparse_ffn_noalloc(const spif_ffn_args * args, spif_stream_t stream) {
    SPIF_REQUIRE(args != nullptr, SPIF_E_NULL);
    SPIF_REQUIRE(args->n_tok == 1, SPIF_E_UNSUPPORTED);  // single-token decode only: the batched path's scratch is not "modest"
    SPIF_REQUIRE(args->n_embd > 0 && args->n_ff > 0, SPIF_E_SHAPE);
    // sizes beyond the static array: a clear error, never a silent overrun
    const size_t need = spif_workspace_size(args->n_embd, args->n_ff, 1);
    SPIF_REQUIRE(need != 0 && need <= (size_t) SPIF_NOALLOC_WS_BYTES, SPIF_E_STATIC_LIMIT);
    void * ws = nullptr;
    int    rc = static_ws_ptr(&ws);
    if (rc) {
        return rc;
    }
    return spif_sparse_ffn(args, ws, (size_t) SPIF_NOALLOC_WS_BYTES, stream);
}
