// spif_layer.cu -- ONE persistent kernel per decode layer: the sparse FFN of layer l and the look-ahead predictor of layer l+1.
//
// Reference graph order (src/llama-graph.cpp:932-946, 896-1142): while layer l runs, the mask of layer l+1 is predicted from
// layer l's FFN input x_l.  The reference issues 6 predictor nodes + 2 MUL_MAT_SPARSE + activation + AXPY_SPARSE (+ memsets)
// per layer; round 1 of this repo used 4 launches on two streams (predictor up, predictor down, fused FFN, partial sum).
// Here everything that shares x_l is one launch, and every grid-wide dependency is hidden behind other streaming work:
//
//   per CTA (one per SM, all co-resident), in queue order:
//     A  its share of the predictor up-projection rows  Wpu[r][n_embd]      -> h1 rows, published as {value, seq} words
//     B  its share of layer l's ACTIVE neurons: gate/up rows -> hidden value -> W_down^T row (only if hidden != 0)
//        -> partial y in registers -> written once, counted
//     C  its contiguous block of predictor down-projection rows Wpd[m][r]   (needs all of h1: long complete by now)
//        -> sigmoid -> sparse_idx of layer l+1, and the block's ACTIVE-ROW LIST for the next launch (warp ballot + popc)
//     D  column-sliced fixed-order sum of the partial y of all CTAs (long complete by now) [+ the shard exchange]
//
//   * The mask is never scanned again: launch l+1 reads 148 block counts + its own ~9 list entries instead of 55 KB of floats
//     per CTA (round 1: 2.5 us on every layer's critical path, 148 times the same scan).
//   * No second kernel for the partial sum (round 1: 3-4 us per layer, 10.7 % of the launch list).
//   * One register budget: round 1's FFN and predictor grids fought for the SMs and overlapped by 2.7 us of 25.
//
// Data movement: ONE elected thread (the issuer warp) streams everything with 1-D TMA bulk copies (cp.async.bulk -> UBLKCP)
// into a ring of slots of RPS = 4 (2 for rows > 12 KB) weight rows each (13B: 5 x 40 KB = 200 KB in flight per SM, Little's
// law needs ~64 KB); completion by mbarrier complete_tx.  Consumer warps own fixed 16-byte column slices of a row (x in
// registers), multiply whatever item arrives (RPS rows -> RPS partial dots, reduced with ONE packed shuffle tree), publish
// per-warp partials in shared memory and arrive on the slot's `done` barrier; they never wait for each other.  A third
// role, the epilogue warp, retires items in order (cross-warp sum in a fixed order, bias, activation, sigmoid, list
// compaction), pushes the W_down^T rows of neurons whose hidden value is non-zero to the issuer's queue and frees the slot.
//
// Determinism: every sum has a fixed shape (packed shuffle tree per warp, fixed warp order, W_down^T rows in ascending
// neuron order per CTA, partials in CTA order) -> run-to-run bit-stable, unlike axpy-sparse.cu:84 (float atomics).
#include "spif_comm.cuh"

#include <cstdlib>
#include <type_traits>

#define SPIF_LIST_MAX_BLOCKS 1024
#define SPIF_LIST_HDR        4 /* int32: {nblk, m, 0, 0} */

enum { IT_WPU = 0, IT_GU = 1, IT_DN = 2, IT_WPD = 3, IT_FLUSH = 4, IT_END = 5 };

struct LayerParams {
    // ---- sparse FFN of this layer (has_ffn)
    const unsigned char * w_up;
    const unsigned char * w_gate;
    const unsigned char * w_down;
    const float *         b_up;
    const float *         b_gate;
    const float *         b_down;
    const float *         x;           // [n_embd] (also the predictor's input: look-ahead)
    float *               y;           // [n_embd]
    const int32_t *       neuron_idx;  // nullable [m]
    float *               hidden_out;  // nullable dense [n_ff]
    int32_t *             stats_out;   // nullable [4]
    const int32_t *       lin_cnt;     // active-row list of this layer: counts per block
    const int32_t *       lin_rows;    //   rows, block b at offset floor(b * m / nblk)
    const float *         norm_w;      // nullable [n_embd]: RMSNorm weight, x is normalised in the kernel (spif_layer_io)
    const float *         residual;    // nullable [n_embd]: added to y after b_down
    float                 norm_eps;
    float *               partials;    // [P][n_embd]
    int32_t *             cta_stats;   // [P][2]
    // ---- predictor of the next layer (has_pred)
    const unsigned char * wpu;         // [r][n_embd]
    const unsigned char * wpd;         // [pm][r]
    const float *         bpu;
    const float *         bpd;
    float *               p_out;       // nullable [pm]
    float *               logits_out;  // nullable [pm]
    int32_t *             lout_cnt;    // nullable: active-row list of the next layer
    int32_t *             lout_rows;
    int32_t *             lout_hdr;
    int                   has_ffn, has_pred;
    int                   n_embd, V, row_bytes;
    int                   m, n_ff, lin_nblk;
    int                   r, pm, Vr, RP, WR;  // Vr = r/8 vectors per Wpd row, RP = Wpd rows per slot row, WR = 32-vector groups per Wpd row
    int                   nw;                 // consumer warps
    int                   nslots;
    int                   cap_ffn, cap_pd, cap_pu;
    float                 thr_next;           // threshold for the list of the next layer
    int                   arch, flags, pflags;
    int                   l2_hint;            // 1: TMA copies carry an L2 evict_first policy
    int                   dbg;                // SPIF_LAYER_DEBUG bit 0: consumers skip the arithmetic (streaming floor; wrong results)
    spif_comm_dev         comm;
    unsigned long long *  trace;
};

struct LayerSmem {
    uint32_t ring, full, done, freeb, desc, dalpha, part, ctl, list_row, list_neu, list_bu, list_bg, pend_row, pend_alpha, bpu_s,
        bpd_s, lout_s, h1_s, blk_pre, hred, total;
};

__host__ __device__ static inline LayerSmem layer_smem_layout(int nslots, int slot_bytes, int cap_ffn, int cap_pd, int cap_pu,
                                                              int r, int nblk) {
    LayerSmem L;
    uint32_t  o = 0;
    auto      take = [&](uint32_t bytes, uint32_t align) {
        o               = (o + align - 1) / align * align;
        const uint32_t r = o;
        o += bytes;
        return r;
    };
    L.ring       = take((uint32_t) nslots * slot_bytes, 128);
    L.full       = take(8u * nslots, 8);
    L.done       = take(8u * nslots, 8);
    L.freeb      = take(8u * nslots, 8);
    L.desc       = take(16u * nslots, 16);
    L.dalpha     = take(16u * nslots, 16);
    L.part       = take(4u * 128 * nslots, 16);
    L.ctl        = take(64u, 16);  // int[4] {pend_tail, gu items retired, -, -}, mbarrier h1 ready, mbarrier partial y stored
    L.list_row   = take(4u * cap_ffn, 4);
    L.list_neu   = take(4u * cap_ffn, 4);
    L.list_bu    = take(4u * cap_ffn, 4);
    L.list_bg    = take(4u * cap_ffn, 4);
    L.pend_row   = take(4u * cap_ffn, 4);
    L.pend_alpha = take(4u * cap_ffn, 4);
    L.bpu_s      = take(4u * cap_pu, 4);
    L.bpd_s      = take(4u * cap_pd, 4);
    L.lout_s     = take(4u * cap_pd, 4);
    L.h1_s       = take(4u * r, 16);
    L.blk_pre    = take(4u * (nblk + 1), 4);
    L.hred       = take(16u * 4 * 16, 16);  // helper warps' scratch of the column-sliced sum
    L.total      = (o + 127) / 128 * 128;
    return L;
}

#define LAYER_STAMP(k)                                                   \
    do {                                                                 \
        if (p.trace && lane == 0) {                                      \
            unsigned long long _t;                                       \
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(_t));        \
            p.trace[(size_t) blockIdx.x * 512 + (k)] = _t;               \
        }                                                                \
    } while (0)

// per-item stamps (debug): [16 + 4*q + {0: type, 1: issued, 2: landed (consumer warp 0), 3: retired}] for q < 120, in SM
// clock cycles (clock64: cheap; %globaltimer ticks every 256 ns).  [15] = the CTA's clock at kernel entry.
#define ITEM_STAMP(q, j, val)                                                        \
    do {                                                                             \
        if (p.trace && lane == 0 && (q) < 120) {                                     \
            unsigned long long _t = (val);                                           \
            if (_t == 0) {                                                           \
                _t = (unsigned long long) clock64();                                 \
            }                                                                        \
            p.trace[(size_t) blockIdx.x * 512 + 16 + 4 * (q) + (j)] = _t;            \
        }                                                                            \
    } while (0)

// lanes 0..15 return the warp sum of a, lanes 16..31 the warp sum of b: two reductions for the price of one tree
__device__ __forceinline__ float reduce2(float a, float b, int lane) {
    const bool hi   = (lane & 16) != 0;
    float      keep = hi ? b : a;
    float      send = hi ? a : b;
    keep += __shfl_xor_sync(0xffffffffu, send, 16);
    keep += __shfl_xor_sync(0xffffffffu, keep, 8);
    keep += __shfl_xor_sync(0xffffffffu, keep, 4);
    keep += __shfl_xor_sync(0xffffffffu, keep, 2);
    keep += __shfl_xor_sync(0xffffffffu, keep, 1);
    return keep;
}
// four reductions with six shuffles: lanes [8j, 8j+8) return the warp sum of v[j]
__device__ __forceinline__ float reduce4(float v0, float v1, float v2, float v3, int lane) {
    const bool hi = (lane & 16) != 0, h8 = (lane & 8) != 0;
    float k0 = (hi ? v2 : v0) + __shfl_xor_sync(0xffffffffu, hi ? v0 : v2, 16);
    float k1 = (hi ? v3 : v1) + __shfl_xor_sync(0xffffffffu, hi ? v1 : v3, 16);
    float m  = (h8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, h8 ? k0 : k1, 8);
    m += __shfl_xor_sync(0xffffffffu, m, 4);
    m += __shfl_xor_sync(0xffffffffu, m, 2);
    m += __shfl_xor_sync(0xffffffffu, m, 1);
    return m;
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ int ld_volatile_s32(const int * p) {
    return *reinterpret_cast<const volatile int *>(p);
}
__device__ __forceinline__ void st_volatile_s32(int * p, int v) {
    *reinterpret_cast<volatile int *>(p) = v;
}

template <typename WT> __device__ __forceinline__ float dot8(const uint4 & wv, const float * xv) {
    float w[8];
    unpack8<WT>(wv, w);
    float s0 = w[0] * xv[0], s1 = w[1] * xv[1];
    s0 = fmaf(w[2], xv[2], s0), s1 = fmaf(w[3], xv[3], s1);
    s0 = fmaf(w[4], xv[4], s0), s1 = fmaf(w[5], xv[5], s1);
    s0 = fmaf(w[6], xv[6], s0), s1 = fmaf(w[7], xv[7], s1);
    return s0 + s1;
}

// floor(a * b / d) for a * b < 2^32 (checked by the launcher).  A 64-bit division is a ~100-instruction dependent chain: six of
// them at kernel entry cost 1.4 us before the first TMA copy was issued (profiles/r2_layer_trace.txt, v2).
__device__ __forceinline__ int share(int a, int b, int d) {
    return (int) ((unsigned) a * (unsigned) b / (unsigned) d);
}

// D: y[columns of this CTA] = sum over all CTAs of their partial y (+ b_down) [+ sum over the shards].  The SHAPE of the sum does
// not depend on who executes it: G virtual groups, group g adds the partials of CTAs g, g + G, g + 2G, ... in ascending order;
// quarter j adds its groups [j*G/4, (j+1)*G/4) in ascending order; the four quarters are added in order.  WIDE: 16 * G threads,
// one group each (one L2 round trip; scratch [(G + 4)][16] float4).  !WIDE: 64 threads, one quarter each in registers (scratch
// [4][16] float4) -- slower, but it runs in the shadow of the predictor stream.  Both give the same bits.
__device__ __forceinline__ float4 layer_group_sum(const float4 * base, int g, int G, int P, int n4) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int pp0 = g; pp0 < P; pp0 += 4 * G) {
        float4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int pp = pp0 + q * G;
            v[q]         = (pp < P) ? __ldcg(base + (size_t) pp * n4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            s.x += v[q].x, s.y += v[q].y, s.z += v[q].z, s.w += v[q].w;
        }
    }
    return s;
}

// out of line: its 8 x 4 words in flight must not set the register budget of the streaming loops
__device__ __noinline__ float4 layer_allreduce_f4(const spif_comm_dev & c, uint32_t seq, float4 v, int idx4, bool has_value) {
    return spif_peer_allreduce_f4(c, seq, v, idx4, has_value);
}

template <bool WIDE>
__device__ __forceinline__ void layer_sum_partials(const LayerParams & p, int t, int nthr, int bar_id, float4 * scratch,
                                                   uint32_t seq, int G) {
    const int P = gridDim.x, c = blockIdx.x, lane = t & 31;
    // b_down of my column: requested before the wait below instead of on the path between the sum and the store of y
    float4 bb_pre = make_float4(0.f, 0.f, 0.f, 0.f), res_pre = bb_pre;
    if (WIDE && t < 64 && (t & 3) == 0 && (p.b_down || p.residual)) {
        const int n4p = p.n_embd >> 2, c0p = share(c, n4p, P), c1p = share(c + 1, n4p, P);
        if ((t >> 2) < c1p - c0p) {
            if (p.b_down) {
                bb_pre = __ldg(reinterpret_cast<const float4 *>(p.b_down) + c0p + (t >> 2));
            }
            if (p.residual) {  // ggml_add(cur, ffn_inp), src/models/llama.cpp:144 (plain load: may be this launch's own x)
                res_pre = reinterpret_cast<const float4 *>(p.residual)[c0p + (t >> 2)];
            }
        }
    }
    if (t == 0) {  // every CTA has published its partial y
        unsigned spins = 0;
        while (ld_u32_acquire_gpu(p.comm.seq + 3) < (unsigned) P) {
            if (++spins > SPIF_COMM_SPIN_LIMIT) {
                atomicOr(p.comm.seq + 2, SPIF_COMM_ERR_TIMEOUT);
                break;
            }
            if (!WIDE) {
                __nanosleep(200);
            }
        }
    }
    named_bar_sync(bar_id, nthr);
    if (t < 32) {
        LAYER_STAMP(11);
    }
    const int      n4 = p.n_embd >> 2;
    const int      c0 = share(c, n4, P), c1 = share(c + 1, n4, P);
    const int      ncol = c1 - c0;  // <= 16 (checked by the launcher)
    const int      col  = t & 15;
    const float4 * base = reinterpret_cast<const float4 *>(p.partials) + c0 + col;
    float4 *       red2;
    if (WIDE) {
        const int pg   = t >> 4;
        float4 *  red  = scratch;                 // [G][16]
        red2           = red + (size_t) G * 16;   // [4][16]
        red[pg * 16 + col] = (col < ncol) ? layer_group_sum(base, pg, G, P, n4) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (t < 32) {
            LAYER_STAMP(3);
        }
        named_bar_sync(bar_id, nthr);
        if (t < 32) {
            LAYER_STAMP(5);
        }
        // second level: the four quarters of a column sit in four adjacent lanes, combined as (q0 + q1) + (q2 + q3) by two
        // shuffles -- no third pass through shared memory
        if (t < 64) {
            const int c2 = t >> 2, j = t & 3, lo = j * G / 4, hi = (j + 1) * G / 4;
            float4    a = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int g = lo; g < hi; ++g) {
                const float4 v = red[g * 16 + c2];
                a.x += v.x, a.y += v.y, a.z += v.z, a.w += v.w;
            }
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                a.x += __shfl_xor_sync(0xffffffffu, a.x, o), a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
                a.z += __shfl_xor_sync(0xffffffffu, a.z, o), a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
            }
            if (j == 0) {
                const bool own = c2 < ncol;
                if (own && p.b_down) {
                    a.x += bb_pre.x, a.y += bb_pre.y, a.z += bb_pre.z, a.w += bb_pre.w;
                }
                a = layer_allreduce_f4(p.comm, seq, a, c0 + c2, own);
                if (own && p.residual) {  // after the sum over the shards: added once
                    a.x += res_pre.x, a.y += res_pre.y, a.z += res_pre.z, a.w += res_pre.w;
                }
                if (own) {
                    reinterpret_cast<float4 *>(p.y)[c0 + c2] = a;
                }
            }
        }
        if (c == 0 && (t >> 5) == (nthr > 64 ? 2 : 1) && p.stats_out) {
            int nz = 0;
            for (int q = lane; q < P; q += 32) {
                nz += __ldcg(p.cta_stats + q * 2 + 1);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                nz += __shfl_xor_sync(0xffffffffu, nz, o);
            }
            if (lane == 0) {
                p.stats_out[0] = __ldcg(p.cta_stats + 0);
                p.stats_out[1] = nz;
                p.stats_out[2] = 0;
                p.stats_out[3] = 0;
            }
        }
        if (t < 32) {
            LAYER_STAMP(12);
        }
        return;
    } else {
        const int j = t >> 4, lo = j * G / 4, hi = (j + 1) * G / 4;
        red2        = scratch;
        float4 a    = make_float4(0.f, 0.f, 0.f, 0.f);
        if (col < ncol && 4 * G >= P) {  // two groups = 8 loads in flight per thread (same order of additions as below)
            for (int g = lo; g < hi; g += 2) {
                const bool two = g + 1 < hi;
                float4     v[8];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int pp = g + q * G;
                    v[q]         = (pp < P) ? __ldcg(base + (size_t) pp * n4) : make_float4(0.f, 0.f, 0.f, 0.f);
                    v[4 + q]     = (two && pp + 1 < P) ? __ldcg(base + (size_t) (pp + 1) * n4) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    s0.x += v[q].x, s0.y += v[q].y, s0.z += v[q].z, s0.w += v[q].w;
                    s1.x += v[4 + q].x, s1.y += v[4 + q].y, s1.z += v[4 + q].z, s1.w += v[4 + q].w;
                }
                a.x += s0.x, a.y += s0.y, a.z += s0.z, a.w += s0.w;
                if (two) {
                    a.x += s1.x, a.y += s1.y, a.z += s1.z, a.w += s1.w;
                }
            }
        } else if (col < ncol) {
            for (int g = lo; g < hi; ++g) {
                const float4 v = layer_group_sum(base, g, G, P, n4);
                a.x += v.x, a.y += v.y, a.z += v.z, a.w += v.w;
            }
        }
        red2[j * 16 + col] = a;
    }
    named_bar_sync(bar_id, nthr);
    if (t < 16) {
        const bool own = col < ncol;
        float4     a   = red2[col];
#pragma unroll
        for (int j = 1; j < 4; ++j) {
            const float4 v = red2[j * 16 + col];
            a.x += v.x, a.y += v.y, a.z += v.z, a.w += v.w;
        }
        if (own && p.b_down) {
            const float4 bb = __ldg(reinterpret_cast<const float4 *>(p.b_down) + c0 + col);
            a.x += bb.x, a.y += bb.y, a.z += bb.z, a.w += bb.w;
        }
        a = layer_allreduce_f4(p.comm, seq, a, c0 + col, own);
        if (own) {
            reinterpret_cast<float4 *>(p.y)[c0 + col] = a;
        }
    }
    if (c == 0 && (t >> 5) == 1 && p.stats_out) {
        int nz = 0;
        for (int q = lane; q < P; q += 32) {
            nz += __ldcg(p.cta_stats + q * 2 + 1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nz += __shfl_xor_sync(0xffffffffu, nz, o);
        }
        if (lane == 0) {
            p.stats_out[0] = __ldcg(p.cta_stats + 0);
            p.stats_out[1] = nz;
            p.stats_out[2] = 0;
            p.stats_out[3] = 0;
        }
    }
    if (t < 32) {
        LAYER_STAMP(12);
    }
}

#define SPIF_LAYER_HELPER_WARPS 2

// Warp roles: [0, NW) consumers, NW = TMA issuer, NW + 1 = epilogue, NW + 2 .. NW + 3 = helpers.  RPS = weight rows per ring slot (one queue item).
//   issuer    waits for a free slot, decides what goes in (pending W_down^T rows first, then the static stream Wpu -> gate/up ->
//             Wpd), issues the bulk copies.  Nothing else: a single warp executes ~1 dependent instruction per 5 cycles, so the
//             first version (issue + retire in one warp, ~1 us per item) capped the SM at 20 GB/s (profiles/r2_layer_trace.txt).
//   consumers multiply the RPS rows of an item with x (or the predictor's hidden vector), RPS partial dots -> one packed
//             shuffle tree -> per-warp partials in shared memory -> arrive on done[slot].
//   epilogue  retires items in order: cross-warp sums (fixed order), bias / activation / sigmoid / list compaction, pushes
//             W_down^T rows of non-zero hidden values to the issuer's queue, frees the slot.
//   helpers   everything that waits on OTHER CTAs, off the streaming path: poll the predictor's hidden vector (words from every
//             CTA of every shard) into shared memory; publish this CTA's partial y (fence + counter); when a predictor stream
//             is still running, the column-sliced sum over all CTAs' partial y [+ the shard exchange] in its shadow.
template <typename WT, int VPT, int RPS, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) spif_layer_kernel(const __grid_constant__ LayerParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NW = p.nw, NC = NW * 32;
    const int P = gridDim.x, c = blockIdx.x;
    const int N = p.nslots;
    const int slot_bytes = RPS * p.row_bytes;

    const LayerSmem L = layer_smem_layout(N, slot_bytes, p.cap_ffn, p.cap_pd, p.cap_pu, p.has_pred ? p.r : 0, p.lin_nblk);
    unsigned char * ring       = smem + L.ring;
    uint64_t *      full       = reinterpret_cast<uint64_t *>(smem + L.full);
    uint64_t *      done       = reinterpret_cast<uint64_t *>(smem + L.done);
    uint64_t *      freeb      = reinterpret_cast<uint64_t *>(smem + L.freeb);
    int4 *          desc       = reinterpret_cast<int4 *>(smem + L.desc);
    float *         dalpha     = reinterpret_cast<float *>(smem + L.dalpha);  // [N][4]
    float *         part       = reinterpret_cast<float *>(smem + L.part);    // [N][4][32]
    int *           ctl        = reinterpret_cast<int *>(smem + L.ctl);       // [0] pend_tail, [1] gu items retired
    uint64_t *      h1bar      = reinterpret_cast<uint64_t *>(smem + L.ctl + 16);
    uint64_t *      flushbar   = reinterpret_cast<uint64_t *>(smem + L.ctl + 24);
    uint64_t *      listbar    = reinterpret_cast<uint64_t *>(smem + L.ctl + 32);
    int *           dbgflag    = reinterpret_cast<int *>(smem + L.ctl + 40);
    float4 *        hred       = reinterpret_cast<float4 *>(smem + L.hred);
    int *           list_row   = reinterpret_cast<int *>(smem + L.list_row);
    int *           list_neu   = reinterpret_cast<int *>(smem + L.list_neu);
    float *         list_bu    = reinterpret_cast<float *>(smem + L.list_bu);
    float *         list_bg    = reinterpret_cast<float *>(smem + L.list_bg);
    int *           pend_row   = reinterpret_cast<int *>(smem + L.pend_row);
    float *         pend_alpha = reinterpret_cast<float *>(smem + L.pend_alpha);
    float *         bpu_s      = reinterpret_cast<float *>(smem + L.bpu_s);
    float *         bpd_s      = reinterpret_cast<float *>(smem + L.bpd_s);
    int *           lout_s     = reinterpret_cast<int *>(smem + L.lout_s);
    float *         h1_s       = reinterpret_cast<float *>(smem + L.h1_s);
    int *           blk_pre    = reinterpret_cast<int *>(smem + L.blk_pre);

    if (tid == 0) {
        if (p.trace) {
            p.trace[(size_t) blockIdx.x * 512 + 15] = (unsigned long long) clock64();
            LAYER_STAMP(7);
        }
        for (int s = 0; s < N; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], NW);
            mbar_init(&freeb[s], 1);
        }
        ctl[0] = 0, ctl[1] = 0;
        dbgflag[0] = 0;
        mbar_init(h1bar, 32);
        mbar_init(flushbar, 1);
        mbar_init(listbar, 1);
        mbar_fence_init();
    }
    __syncthreads();

    const bool     gated    = (p.arch != SPIF_ARCH_RELU_UP);
    const bool     rounding = (p.flags & SPIF_FLAG_CPU_ROUNDING) != 0;
    const int      world    = p.comm.world;
    const int      h1_off   = p.n_embd;  // h1 words live behind the y words of a slot
    uint32_t       seq      = 0;

    // my shares of the dense predictor matrices (weights: independent of the previous kernel)
    int ra = 0, rb = 0, pa = 0, pb = 0;
    if (p.has_pred) {
        const int rw = p.r / world, R0 = p.comm.rank * rw;  // r % world == 0 (launcher)
        ra = R0 + share(c, rw, P);
        rb = R0 + share(c + 1, rw, P);
        pa = share(c, p.pm, P);
        pb = share(c + 1, p.pm, P);
    }
    const int neu_per_gu = gated ? RPS / 2 : RPS;  // neurons per gate/up item

    if (warp == NW) {
        // ================================================= TMA issuer warp =================================================
        const uint64_t pol  = l2_policy_evict_first();
        const bool     hint = p.l2_hint != 0;
        auto tma = [&](void * dst, const void * src, uint32_t bytes, uint64_t * bar) {
            if (hint) {
                tma_load_1d(dst, src, bytes, bar, pol);
            } else {
                tma_load_1d_nohint(dst, src, bytes, bar);
            }
        };
        int      q = 0, s = 0;
        unsigned fph = 0;          // phase of free[s] to wait for once q >= N
        int      a_next = ra, b_next = 0, b_end = 0, c_next = pa;
        int      pend_head = 0, gu_issued = 0, total_active = 0;
        bool     flushed = false, list_ready = !p.has_ffn;
        const uint32_t wrow = (uint32_t) p.r * 2;

        // returns 1 if an item was issued, 0 if nothing can be issued right now, -1 after END
        auto issue_next = [&]() -> int {
            int gu_ret = 0, np = 0;
            if (!list_ready && mbar_test(listbar, 0)) {  // the helper warp has built this CTA's row list
                list_ready   = true;
                b_end        = ctl[2];
                total_active = ctl[3];
                LAYER_STAMP(2);
            }
            if (p.has_ffn) {
                gu_ret = ld_volatile_s32(ctl + 1);
                __threadfence_block();
                np = ld_volatile_s32(ctl + 0) - pend_head;  // read AFTER gu_ret: covers every push of the retired items
            }
            const bool gu_all_issued = list_ready && b_next >= b_end;
            const bool gu_all_done   = gu_all_issued && gu_ret == gu_issued;
            int        type          = -1;
            if (np >= RPS || (np >= 2 && !gu_all_done) || (np >= 1 && gu_all_done)) {
                type = IT_DN;
            } else if (a_next < rb) {
                type = IT_WPU;
            } else if (list_ready && b_next < b_end) {
                type = IT_GU;  // the head of the longest dependency chain (gate/up -> hidden -> W_down^T -> partial y -> sum)
            } else if (p.has_ffn && list_ready && !flushed && gu_all_done && np == 0) {
                type = IT_FLUSH;
            } else if (c_next < pb && list_ready) {
                type = IT_WPD;  // dense, known at entry, needs only h1 (a helper warp collects it meanwhile)
            } else if (flushed || !p.has_ffn) {
                type = IT_END;
            } else {
                return 0;  // gate/up items still in flight: their W_down^T rows may follow
            }
            if (q >= N) {
                mbar_wait_spin(&freeb[s], fph);
            }
            unsigned char * dst = ring + (size_t) s * slot_bytes;
            if (lane == 0) {
                switch (type) {
                    case IT_WPU: {
                        const int n = min(RPS, rb - a_next);
                        desc[s]     = make_int4(IT_WPU, a_next, n, 0);
                        mbar_arrive_expect_tx(&full[s], (uint32_t) n * p.row_bytes);
                        tma(dst, p.wpu + (size_t) a_next * p.row_bytes, (uint32_t) n * p.row_bytes, &full[s]);  // rows are adjacent
                    } break;
                    case IT_GU: {
                        const int n = min(neu_per_gu, b_end - b_next);
                        desc[s]     = make_int4(IT_GU, b_next, n, 0);
                        mbar_arrive_expect_tx(&full[s], (uint32_t) (gated ? 2 * n : n) * p.row_bytes);
                        for (int j = 0; j < n; ++j) {
                            const size_t off = (size_t) list_row[b_next + j] * p.row_bytes;
                            if (gated) {
                                tma(dst + (size_t) (2 * j) * p.row_bytes, p.w_up + off, p.row_bytes, &full[s]);
                                tma(dst + (size_t) (2 * j + 1) * p.row_bytes, p.w_gate + off, p.row_bytes, &full[s]);
                            } else {
                                tma(dst + (size_t) j * p.row_bytes, p.w_up + off, p.row_bytes, &full[s]);
                            }
                        }
                    } break;
                    case IT_DN: {
                        const int n = min(RPS, np);
                        desc[s]     = make_int4(IT_DN, n, 0, 0);
                        mbar_arrive_expect_tx(&full[s], (uint32_t) n * p.row_bytes);
                        for (int j = 0; j < n; ++j) {
                            dalpha[s * 4 + j] = pend_alpha[pend_head + j];
                            tma(dst + (size_t) j * p.row_bytes, p.w_down + (size_t) pend_row[pend_head + j] * p.row_bytes, p.row_bytes,
                                &full[s]);
                        }
                    } break;
                    case IT_WPD: {
                        const int n = min(RPS * p.RP, pb - c_next);
                        desc[s]     = make_int4(IT_WPD, c_next, n, 0);
                        mbar_arrive_expect_tx(&full[s], (uint32_t) n * wrow);
                        for (int j = 0, left = n; left > 0; ++j, left -= p.RP) {  // RP rows of Wpd per slot row
                            tma(dst + (size_t) j * p.row_bytes, p.wpd + (size_t) (c_next + j * p.RP) * wrow,
                                (uint32_t) min(left, p.RP) * wrow, &full[s]);
                        }
                    } break;
                    case IT_FLUSH:
                        desc[s] = make_int4(IT_FLUSH, total_active, 0, 0);
                        mbar_arrive(&full[s]);
                        break;
                    default:
                        desc[s] = make_int4(IT_END, 0, 0, 0);
                        mbar_arrive(&full[s]);
                        break;
                }
            }
            switch (type) {
                case IT_WPU: a_next += min(RPS, rb - a_next); break;
                case IT_GU: b_next += min(neu_per_gu, b_end - b_next); ++gu_issued; break;
                case IT_DN: pend_head += min(RPS, np); break;
                case IT_WPD: c_next += min(RPS * p.RP, pb - c_next); break;
                case IT_FLUSH: flushed = true; break;
                default: break;
            }
            ITEM_STAMP(q, 0, (unsigned long long) (type + 1));
            ITEM_STAMP(q, 1, 0);
            ++q;
            if (++s == N) {
                s = 0;
                fph ^= (q > N) ? 1u : 0u;
            }
            return type == IT_END ? -1 : 1;
        };

        // ---- before the programmatic-dependent-launch wait: the predictor up-projection rows are already on their way.
        // (Not the Wpd rows: consumers take items in order, a Wpd item ahead of the gate/up items makes those wait for h1; and
        // the TMA unit accepts ~80 KB of outstanding copies, then the issuing thread blocks -- v3 trace, profiles/.)
        while (q < N && a_next < rb) {
            issue_next();
        }
        LAYER_STAMP(0);
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        LAYER_STAMP(1);

        for (;;) {
            const int rc = issue_next();
            if (rc < 0) {
                break;
            }
            if (rc == 0) {
                __nanosleep(64);
            }
        }
        LAYER_STAMP(5);
    } else if (warp == NW + 1) {
        // ================================================== epilogue warp ==================================================
        for (int i = lane; i < rb - ra; i += 32) {
            bpu_s[i] = p.bpu ? __ldg(p.bpu + ra + i) : 0.f;
        }
        for (int i = lane; i < pb - pa; i += 32) {
            bpd_s[i] = p.bpd ? __ldg(p.bpd + pa + i) : 0.f;
        }
        __syncwarp();
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq                = spif_comm_call_seq(p.comm);
        const uint32_t par = seq & 1u;
        if (p.has_pred && rb == ra && lane == 0) {
            atomicAdd(p.comm.seq + 4, 1u);  // no Wpu rows here: nothing to wait for from this CTA
        }
        int            pend_tail = 0, gu_retired = 0, n_nonzero = 0, cnt_out = 0;
        int            s = 0, q = 0;
        unsigned       ph = 0;
        for (;; ++q) {
            mbar_wait_spin(&done[s], ph);
            ITEM_STAMP(q, 3, 0);
            const int4    d  = desc[s];
            const float * ps = part + s * 128;
            if (d.x == IT_END) {
                break;
            }
            if (d.x == IT_WPU || d.x == IT_GU) {
                float k;
                if (RPS == 4) {
                    const float a0 = (lane < NW) ? ps[lane] : 0.f, a1 = (lane < NW) ? ps[32 + lane] : 0.f;
                    const float a2 = (lane < NW) ? ps[64 + lane] : 0.f, a3 = (lane < NW) ? ps[96 + lane] : 0.f;
                    k              = reduce4(a0, a1, a2, a3, lane);  // lanes [8j, 8j+8): row j
                } else {
                    const float a0 = (lane < NW) ? ps[lane] : 0.f, a1 = (lane < NW) ? ps[32 + lane] : 0.f;
                    k              = reduce2(a0, a1, lane);          // lanes [16j, 16j+16): row j
                }
                constexpr int LPR = 32 / RPS;  // lanes per row sum
                if (d.x == IT_WPU) {
                    // h1[row] = relu(Wpu[row] . x + bpu[row])  (llama-graph.cpp:871-880), published to every shard
                    const int j = lane / LPR;
                    if ((lane % LPR) == 0 && j < d.z) {
                        const int row = d.y + j;
                        float     z   = k + bpu_s[row - ra];
                        z             = (z > 0.f) ? z : 0.f;
                        const size_t off = ((size_t) par * world + p.comm.rank) * p.comm.cap + h1_off + row;
                        if (p.comm.mc) {
                            st_word_sys(p.comm.mc + off, __float_as_uint(z), seq);
                        } else {
                            for (int r = 0; r < world; ++r) {
                                st_word_sys(p.comm.data[r] + off, __float_as_uint(z), seq);
                            }
                        }
                    }
                    if (d.y + d.z >= rb) {  // hint for the pollers: this CTA's words are on their way (not a release: every
                        __syncwarp();       // word carries its own sequence number)
                        if (lane == 0) {
                            atomicAdd(p.comm.seq + 4, 1u);
                        }
                        LAYER_STAMP(13);
                    }
                } else {
                    // gated: slot rows (up_0, gate_0, up_1, gate_1); ungated: (up_0 .. up_{RPS-1})
                    for (int j = 0; j < d.z; ++j) {
                        const int   i = d.y + j;
                        const float u = __shfl_sync(0xffffffffu, k, (gated ? 2 * j : j) * LPR) + list_bu[i];
                        const float g = gated ? __shfl_sync(0xffffffffu, k, (2 * j + 1) * LPR) + list_bg[i] : 0.f;
                        float       h = spif_activation(p.arch, g, u);
                        if (lane == 0 && p.hidden_out) {
                            p.hidden_out[list_neu[i]] = h;
                        }
                        if (rounding) {
                            h = round_through<WT>(h);
                        }
                        if (!(h == 0.0f)) {  // axpy-sparse.cu:53: rows with hid == 0 are never read
                            if (lane == 0) {
                                pend_row[pend_tail]   = list_row[i];
                                pend_alpha[pend_tail] = h;
                            }
                            ++pend_tail;
                            ++n_nonzero;
                        }
                    }
                    ++gu_retired;
                    if (lane == 0) {
                        __threadfence_block();
                        st_volatile_s32(ctl + 0, pend_tail);
                        __threadfence_block();
                        st_volatile_s32(ctl + 1, gu_retired);
                    }
                }
            } else if (d.x == IT_WPD) {
                const int row0 = d.y, n = d.z;
                for (int base = 0; base < n; base += 32) {
                    const int  idx   = base + lane;
                    const bool valid = idx < n;
                    float      z     = 0.f;
                    if (valid) {
                        // Wpd row idx of the item lives in slot row j = idx / RP; its 32-vector groups are g0 .. g0 + WR - 1
                        // of that slot row; group g was reduced by consumer warp g % NW for its vector k = g / NW
                        const int j = idx / p.RP, g0 = (idx - j * p.RP) * p.WR;
                        for (int t = 0; t < p.WR; ++t) {  // fixed order
                            const int g = g0 + t, k = g / NW, w = g - k * NW;
                            z += ps[(j * VPT + k) * 32 + w];
                        }
                        z += bpd_s[row0 - pa + idx];
                    }
                    const float pr = 1.f / (1.f + expf(-z));  // unary-ops.cpp:31-33 / unary.cu:48-50
                    if (valid) {
                        if (p.p_out) {
                            p.p_out[row0 + idx] = pr;
                        }
                        if (p.logits_out) {
                            p.logits_out[row0 + idx] = z;
                        }
                    }
                    const bool     act  = valid && spif_is_active(pr, p.thr_next);
                    const unsigned ball = __ballot_sync(0xffffffffu, act);
                    if (act) {
                        lout_s[cnt_out + __popc(ball & ((1u << lane) - 1u))] = row0 + idx;
                    }
                    cnt_out += __popc(ball);
                }
            } else if (d.x == IT_FLUSH) {
                // every consumer thread has stored its slice of the partial y (their arrivals on done[s] ordered the stores
                // before this point); a helper thread makes them visible to the other CTAs (fence + counter) off this path
                if (lane == 0) {
                    p.cta_stats[c * 2 + 0] = d.y;
                    p.cta_stats[c * 2 + 1] = n_nonzero;
                    mbar_arrive(flushbar);
                }
                LAYER_STAMP(4);
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&freeb[s]);
            }
            if (++s == N) {
                s = 0;
                ph ^= 1u;
            }
        }
        // ---- the next layer's active-row list: this block's rows in ascending order + its count
        if (p.has_pred && p.lout_cnt) {
            __syncwarp();
            for (int i = lane; i < cnt_out; i += 32) {
                p.lout_rows[pa + i] = lout_s[i];
            }
            if (lane == 0) {
                p.lout_cnt[c] = cnt_out;
                if (c == 0) {
                    p.lout_hdr[0] = P;
                    p.lout_hdr[1] = p.pm;
                }
            }
        }
        LAYER_STAMP(6);
    } else if (warp < NW) {
        // ================================================ consumer warps ================================================
        float xr[VPT * 8], acc[VPT * 8], hr[VPT * 8];
#pragma unroll
        for (int e = 0; e < VPT * 8; ++e) {
            xr[e] = 0.f, acc[e] = 0.f, hr[e] = 0.f;
        }
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq = spif_comm_call_seq(p.comm);
        bool vok[VPT];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = tid + k * NC;
            vok[k]      = v < p.V;
            if (vok[k]) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(p.x) + 2 * v + 1);
                xr[k * 8 + 0] = a.x, xr[k * 8 + 1] = a.y, xr[k * 8 + 2] = a.z, xr[k * 8 + 3] = a.w;
                xr[k * 8 + 4] = b.x, xr[k * 8 + 5] = b.y, xr[k * 8 + 6] = b.z, xr[k * 8 + 7] = b.w;
                if (rounding) {  // ggml-cpu.c:1832-1856: x is converted to the weight type
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        xr[k * 8 + e] = round_through<WT>(xr[k * 8 + e]);
                    }
                }
            }
        }
        const uint32_t ring_s  = smem_u32(ring) + (uint32_t) tid * 16u;  // my first vector of slot 0, row 0
        const uint32_t koff    = (uint32_t) NC * 16u;                    // my next vector (VPT == 2)
        const uint32_t rbytes  = (uint32_t) p.row_bytes;
        bool           h1_loaded = false;
        int            s = 0, qc = 0;
        unsigned       ph = 0;
        uint32_t       slot_s = ring_s;
        for (;; ++qc) {
            mbar_wait(&full[s], ph);
            if (warp == 0) {
                if (qc == 0) {
                    LAYER_STAMP(8);
                }
                ITEM_STAMP(qc, 2, 0);
            }
            const int4 d = desc[s];
            float *    pw = part + s * 128 + warp;
            if (p.dbg & 1) {
            } else if (d.x == IT_WPU || d.x == IT_GU) {
                float v[RPS];
#pragma unroll
                for (int r = 0; r < RPS; ++r) {
                    v[r] = 0.f;
                }
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    if (vok[k]) {
                        uint4 w[RPS];
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            w[r] = lds128(slot_s + r * rbytes + k * koff);
                        }
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            v[r] += dot8<WT>(w[r], xr + k * 8);
                        }
                    }
                }
                // rows that were not loaded hold stale bytes: their sums are never read by the epilogue (NaN bits stay in
                // their own lane group of the packed tree)
                float k2;
                if (RPS == 4) {
                    k2 = reduce4(v[0], v[1], v[2], v[3], lane);
                    if ((lane & 7) == 0) {
                        pw[(lane >> 3) * 32] = k2;
                    }
                } else {
                    k2 = reduce2(v[0], v[1], lane);
                    if ((lane & 15) == 0) {
                        pw[(lane >> 4) * 32] = k2;
                    }
                }
            } else if (d.x == IT_DN) {
                const float4 al = *reinterpret_cast<const float4 *>(dalpha + s * 4);
                const float  alv[4] = { al.x, al.y, al.z, al.w };
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    if (vok[k]) {
                        uint4 w[RPS];
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            w[r] = lds128(slot_s + r * rbytes + k * koff);
                        }
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            if (r < d.y) {  // uniform: ascending neuron order, a fixed FMA chain per element
                                float f[8];
                                unpack8<WT>(w[r], f);
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    acc[k * 8 + e] = fmaf(f[e], alv[r], acc[k * 8 + e]);
                                }
                            }
                        }
                    }
                }
            } else if (d.x == IT_WPD) {
                if (!h1_loaded) {
                    // the hidden vector of the predictor, collected by the helper warps while the first items streamed
                    h1_loaded = true;
                    mbar_wait(h1bar, 0);
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        if (vok[k]) {
                            const int col = (tid + k * NC) % p.Vr;
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                float hv = h1_s[col * 8 + e];
                                if (p.pflags & SPIF_FLAG_CPU_ROUNDING) {
                                    hv = round_through<WT>(hv);
                                }
                                hr[k * 8 + e] = hv;
                            }
                        }
                    }
                }
                // slot row r holds RP rows of Wpd back to back; my vector k of that slot row belongs to one of them (the same
                // one for the whole warp): RPS * VPT partial dots per item, one per (slot row, vector)
                float v[RPS * VPT];
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    uint4 w[RPS];
#pragma unroll
                    for (int r = 0; r < RPS; ++r) {
                        w[r] = vok[k] ? lds128(slot_s + r * rbytes + k * koff) : make_uint4(0u, 0u, 0u, 0u);
                    }
#pragma unroll
                    for (int r = 0; r < RPS; ++r) {
                        v[r * VPT + k] = dot8<WT>(w[r], hr + k * 8);
                    }
                }
                if (RPS * VPT == 4) {
                    const float k2 = reduce4(v[0], v[1], v[2], v[3], lane);
                    if ((lane & 7) == 0) {
                        pw[(lane >> 3) * 32] = k2;
                    }
                } else {
                    const float k2 = reduce2(v[0], v[1], lane);
                    if ((lane & 15) == 0) {
                        pw[(lane >> 4) * 32] = k2;
                    }
                }
            } else if (d.x == IT_FLUSH) { /* this CTA's partial y */
                float * prt = p.partials + (size_t) c * p.n_embd;
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    if (vok[k]) {
                        float4 * o = reinterpret_cast<float4 *>(prt) + 2 * (tid + k * NC);
                        __stcg(o, make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]));
                        __stcg(o + 1, make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]));
                    }
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&done[s]);
            }
            if (d.x == IT_END) {
                break;
            }
            slot_s += (uint32_t) slot_bytes;
            if (++s == N) {
                s = 0;
                ph ^= 1u;
                slot_s = ring_s;
            }
        }
        if (warp == 0) {
            LAYER_STAMP(10);
        }

        // ---- D (launches without a predictor stream to hide it behind: all consumer threads, one L2 round trip)
        if (p.has_ffn && (!p.has_pred || !(p.dbg & 4))) {
            named_bar_sync(1, NC);  // every consumer is done with the ring: reuse it as scratch
            layer_sum_partials<true>(p, tid, NC, 1, reinterpret_cast<float4 *>(ring), seq, 2 * NW);
        }
    } else {
        // ================================================== helper warps ==================================================
        const int ht = tid - (NW + 2) * 32, HT = 32 * SPIF_LAYER_HELPER_WARPS;
        pdl_wait_prior_grid();
        pdl_launch_dependents();
        seq = spif_comm_call_seq(p.comm);
        // ---- my share of this layer's active rows: ranks [c*A/P, (c+1)*A/P) of the concatenated block lists
        if (p.has_ffn && ht >= 32) {
            // lane L owns blocks [L*kb, (L+1)*kb): all counts are requested at once (one L2 round trip), one warp scan, the
            // exclusive prefix of every block goes to shared memory; then one lane per list entry finds its block by binary
            // search and fetches the row (second round trip, all entries in parallel)
            const int nb = p.lin_nblk, kb = (nb + 31) / 32;  // kb <= 8 (checked by the launcher: nblk <= 256)
            const int b0 = min(lane * kb, nb), b1 = min(b0 + kb, nb);
            int       cb[8];
            int       local = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                cb[j] = (b0 + j < b1) ? __ldcg(p.lin_cnt + b0 + j) : 0;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                local += cb[j];
            }
            int incl = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                incl += (lane >= o) ? t : 0;
            }
            const int total_active = __shfl_sync(0xffffffffu, incl, 31);
            int pbk      = incl - local;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (b0 + j < b1) {
                    blk_pre[b0 + j] = pbk;
                }
                pbk += cb[j];
            }
            if (lane == 31) {
                blk_pre[nb] = total_active;
            }
            __syncwarp();
            const int lo = share(c, total_active, P), hi = share(c + 1, total_active, P);
            const int b_end = hi - lo;
            for (int j = lane; j < b_end; j += 32) {
                const int rank = lo + j;
                int       bl = 0, bh = nb;  // blk_pre[bl] <= rank < blk_pre[bh]
                while (bh - bl > 1) {
                    const int mid = (bl + bh) >> 1;
                    if (blk_pre[mid] <= rank) {
                        bl = mid;
                    } else {
                        bh = mid;
                    }
                }
                const int row = __ldcg(p.lin_rows + share(bl, p.m, nb) + (rank - blk_pre[bl]));
                const int neu = p.neuron_idx ? __ldg(p.neuron_idx + row) : row;
                list_row[j]   = row;
                list_neu[j]   = neu;
                list_bu[j]    = p.b_up ? __ldg(p.b_up + neu) : 0.f;
                list_bg[j]    = p.b_gate ? __ldg(p.b_gate + neu) : 0.f;
            }
            __syncwarp();
            if (lane == 0) {
                ctl[2] = b_end;
                ctl[3] = total_active;
                mbar_arrive(listbar);  // release: the issuer's test_wait acquires the list
            }
        }
        if (p.has_pred && ht < 32) {
            // the hidden vector of the predictor: {value, seq} words, complete once every CTA (of every shard) has retired its
            // Wpu rows -- first in everybody's queue.  Few threads poll (all threads polling floods the L2 with spinning loads
            // and starves the producers' stores: 100 us per launch in the first version), 8 words per thread per pass.
            const uint32_t par = seq & 1u;
            const int      rw  = p.r / world;  // rows per owner rank
            unsigned       spins = 0;
            // one word is polled until every CTA of THIS shard has sent its rows (148 pollers x 8 KB of words per round used a
            // quarter of the L2's request rate and cost 2.7 us per layer); the words themselves are then read, normally once
            while (ld_u32_relaxed_gpu(p.comm.seq + 4) < (unsigned) P && ++spins < SPIF_COMM_SPIN_LIMIT) {
                __nanosleep(100);
            }
            LAYER_STAMP(14);
            for (int row = ht * 16; row < p.r; row += 512) {  // 16 words per thread per pass (r % 256 == 0: whole threads)
                const int     owner = row / rw;
                const uint2 * src   = p.comm.data[p.comm.rank] + ((size_t) par * world + owner) * p.comm.cap + h1_off + row;
                uint2         w16[16];
                bool          ok = false;
                while (true) {
#pragma unroll
                    for (int e = 0; e < 16; e += 2) {
                        ld_word2_sys(src + e, w16[e], w16[e + 1]);
                    }
                    ok = true;
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        ok = ok && (w16[e].y == seq);
                    }
                    if (ok || ++spins > SPIF_COMM_SPIN_LIMIT) {
                        break;
                    }
                    __nanosleep(100);
                }
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    h1_s[row + e] = ok ? __uint_as_float(w16[e].x) : __uint_as_float(0x7fc00000u);
                }
                if (!ok) {
                    atomicOr(p.comm.seq + 2, SPIF_COMM_ERR_TIMEOUT);
                }
            }
            mbar_arrive(h1bar);  // release: the consumers' wait acquires h1_s
            if (ht < 32) {
                LAYER_STAMP(9);
            }
        }
        if (p.has_ffn) {
            if (ht == 0) {
                mbar_wait(flushbar, 0);  // the epilogue warp has seen every consumer's arrival after its partial-y stores
                __threadfence();
                atomicAdd(p.comm.seq + 3, 1u);
            }
            if (p.has_pred && (p.dbg & 4)) {
                layer_sum_partials<false>(p, ht, HT, 3, hred, seq, 2 * NW);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        spif_comm_block_done(p.comm, seq, (unsigned) P);
    }
}

#include "spif_layer_tma.cuh"

// ---------------------------------------------------------------------------------------------------------------
// float mask -> block lists (the form the layer kernel consumes).  One warp per block of rows, ballot + popc, ascending.
__global__ void __launch_bounds__(32) spif_mask_to_list_kernel(const float * __restrict__ mask, int m, float thr, int nblk,
                                                               int32_t * __restrict__ hdr, int32_t * __restrict__ cnt,
                                                               int32_t * __restrict__ rows) {
    const int b = blockIdx.x, lane = threadIdx.x;
    const int r0 = (int) ((long long) b * m / nblk), r1 = (int) ((long long) (b + 1) * m / nblk);
    int       n = 0;
    for (int base = r0; base < r1; base += 32) {
        const int      row  = base + lane;
        const bool     a    = (row < r1) && spif_is_active(__ldg(mask + row), thr);
        const unsigned ball = __ballot_sync(0xffffffffu, a);
        if (a) {
            rows[r0 + n + __popc(ball & ((1u << lane) - 1u))] = row;
        }
        n += __popc(ball);
    }
    if (lane == 0) {
        cnt[b] = n;
        if (b == 0) {
            hdr[0] = nblk;
            hdr[1] = m;
        }
    }
}

size_t spif_active_list_bytes_impl(int m) {
    return sizeof(int32_t) * ((size_t) SPIF_LIST_HDR + SPIF_LIST_MAX_BLOCKS + (size_t) (m > 0 ? m : 0));
}

int spif_launch_mask_to_list(const float * mask, int m, float thr, void * list, const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(mask && list, SPIF_E_NULL);
    SPIF_REQUIRE(m > 0 && ((uintptr_t) list & 15) == 0, SPIF_E_SHAPE);
    const int nblk = ctx.sm_count;
    SPIF_REQUIRE(nblk <= SPIF_LIST_MAX_BLOCKS, SPIF_E_UNSUPPORTED);
    int32_t * hdr = (int32_t *) list;
    spif_mask_to_list_kernel<<<nblk, 32, 0, ctx.stream>>>(mask, m, thr, nblk, hdr, hdr + SPIF_LIST_HDR,
                                                          hdr + SPIF_LIST_HDR + SPIF_LIST_MAX_BLOCKS);
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
extern unsigned long long * g_spif_trace;

template <typename WT, int VPT, int RPS, int MAXT>
static int launch_layer_t(const LayerParams & kp, int P, int threads, size_t smem, int device, cudaStream_t stream) {
    auto kern = spif_layer_kernel<WT, VPT, RPS, MAXT>;
    static unsigned long long done = 0;  // per instantiation: bit per device
    const unsigned long long  bit  = 1ull << (device & 63);
    if (!(done & bit)) {
        SPIF_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        done |= bit;
    }
    // every CTA waits for words written by every other CTA: the whole grid must be resident (one CTA per SM)
    int per_sm = 0;
    SPIF_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
    SPIF_REQUIRE(per_sm >= 1, SPIF_E_UNSUPPORTED);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(P);
    cfg.blockDim           = dim3(threads);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, kp));
    return 0;
}

template <typename WT, int VPT, int MAXT, bool RND, bool TRACE = false>
static int launch_layer4_t(const LayerParams & kp, int P, int threads, size_t smem, int device, cudaStream_t stream) {
    constexpr bool can_trace = VPT == 1 && !RND && std::is_same<WT, spif_f16>::value;
    if (!TRACE && can_trace && kp.trace) {  // tools/trace_layer.py: per-item time stamps
        return launch_layer4_t<WT, VPT, MAXT, RND, can_trace>(kp, P, threads, smem, device, stream);
    }
    auto kern = spif_layer4_kernel<WT, VPT, MAXT, TRACE, RND>;
    static unsigned long long done = 0;  // per instantiation: bit per device
    const unsigned long long  bit  = 1ull << (device & 63);
    if (!(done & bit)) {
        SPIF_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        // every CTA waits for words written by every other CTA: the whole grid must be resident (one CTA per SM)
        int per_sm = 0;
        SPIF_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
        SPIF_REQUIRE(per_sm >= 1, SPIF_E_UNSUPPORTED);
        done |= bit;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(P);
    cfg.blockDim           = dim3(threads);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, kp));
    return 0;
}

int spif_launch_layer(const spif_ffn_args * ffn, const spif_pred_args * pred, const spif_layer_io * io, const void * list_in, void * list_out,
                      const spif_shard_comm * comm, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(ffn != nullptr || pred != nullptr, SPIF_E_NULL);
    SPIF_REQUIRE(comm != nullptr, SPIF_E_NULL);
    LayerParams kp = {};
    int         rc = spif_comm_to_dev(comm, kp.comm);
    if (rc) {
        return rc;
    }
    const int P = ctx.sm_count;
    SPIF_REQUIRE(P >= 1 && P <= 256, SPIF_E_UNSUPPORTED);  // block counts of a list: 8 per lane of the producer warp
    int type = 0, n_embd = 0;
    const float * x = nullptr;
    if (ffn) {
        SPIF_REQUIRE(ffn->n_tok == 1, SPIF_E_UNSUPPORTED);
        SPIF_REQUIRE(ffn->type == SPIF_TYPE_F16 || ffn->type == SPIF_TYPE_BF16, SPIF_E_TYPE);
        SPIF_REQUIRE(ffn->w_up && ffn->w_down_t && ffn->x && ffn->y && list_in, SPIF_E_NULL);
        SPIF_REQUIRE(ffn->arch == SPIF_ARCH_RELU_UP || ffn->w_gate, SPIF_E_NULL);
        SPIF_REQUIRE(ffn->arch >= SPIF_ARCH_PROSPARSE && ffn->arch <= SPIF_ARCH_SILU_GATED, SPIF_E_UNSUPPORTED);
        SPIF_REQUIRE(ffn->m > 0 && ffn->n_ff >= ffn->m && (ffn->neuron_idx || ffn->m == ffn->n_ff), SPIF_E_SHAPE);
        SPIF_REQUIRE((((uintptr_t) ffn->w_up | (uintptr_t) ffn->w_gate | (uintptr_t) ffn->w_down_t | (uintptr_t) ffn->x |
                       (uintptr_t) ffn->y | (uintptr_t) ffn->b_down | (uintptr_t) list_in) & 15) == 0, SPIF_E_SHAPE);
        type   = ffn->type;
        n_embd = ffn->n_embd;
        x      = ffn->x;
    }
    const spif_pred_args * pv = pred;
    if (pv) {
        SPIF_REQUIRE(pv->ptype == SPIF_TYPE_F16 || pv->ptype == SPIF_TYPE_BF16, SPIF_E_TYPE);
        SPIF_REQUIRE(!ffn || pv->ptype == type, SPIF_E_UNSUPPORTED);  // one float codec per launch
        SPIF_REQUIRE(pv->wpu && pv->wpd && pv->x, SPIF_E_NULL);
        SPIF_REQUIRE(!ffn || pv->x == ffn->x, SPIF_E_UNSUPPORTED);    // look-ahead: the predictor reads the FFN's own input
        SPIF_REQUIRE(!ffn || pv->n_embd == n_embd, SPIF_E_SHAPE);
        SPIF_REQUIRE(pv->r > 0 && pv->r % 256 == 0 && pv->r <= pv->n_embd && pv->r % kp.comm.world == 0 && pv->m > 0, SPIF_E_SHAPE);
        SPIF_REQUIRE((((uintptr_t) pv->wpu | (uintptr_t) pv->wpd | (uintptr_t) pv->x | (uintptr_t) list_out) & 15) == 0, SPIF_E_SHAPE);
        type   = pv->ptype;
        n_embd = pv->n_embd;
        x      = pv->x;
    }
    SPIF_REQUIRE(n_embd > 0 && n_embd % 8 == 0, SPIF_E_SHAPE);
    const int V = n_embd / 8;
    SPIF_REQUIRE(V <= 1024, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE((n_embd / 4 + P - 1) / P <= 16, SPIF_E_UNSUPPORTED);  // column slice of the final sum
    SPIF_REQUIRE(kp.comm.cap >= n_embd + (pv ? pv->r : 0), SPIF_E_WORKSPACE);

    const int vpt = (V <= 896) ? 1 : 2;
    kp.nw         = (V + 32 * vpt - 1) / (32 * vpt);
    SPIF_REQUIRE(kp.nw >= 2, SPIF_E_UNSUPPORTED);  // n_embd >= 512: the final column-sliced sum uses 64 threads
    const int threads = (kp.nw + 2 + SPIF_LAYER_HELPER_WARPS) * 32;  // consumers + TMA issuer + epilogue + helpers
    // share(): every a * b / d on the device is done in 32 bits
    const long long max_dim = std::max<long long>(std::max<long long>(ffn ? ffn->m : 0, pv ? pv->m : 0), std::max(n_embd, pv ? pv->r : 0));
    SPIF_REQUIRE((long long) (P + 1) * max_dim < (1ll << 31), SPIF_E_UNSUPPORTED);
    kp.n_embd     = n_embd;
    kp.V          = V;
    kp.row_bytes  = n_embd * 2;
    kp.x          = x;
    kp.trace      = g_spif_trace;
    static const int k_hint = [] { const char * e = getenv("SPIF_LAYER_L2_HINT"); return e ? atoi(e) : 0; }();
    kp.l2_hint    = k_hint;
    static const int k_dbg = [] { const char * e = getenv("SPIF_LAYER_DEBUG"); return e ? atoi(e) : 0; }();
    kp.dbg        = k_dbg;
    if (ffn) {
        kp.has_ffn    = 1;
        kp.w_up       = (const unsigned char *) ffn->w_up;
        kp.w_gate     = (const unsigned char *) ffn->w_gate;
        kp.w_down     = (const unsigned char *) ffn->w_down_t;
        kp.b_up       = ffn->b_up;
        kp.b_gate     = ffn->b_gate;
        kp.b_down     = ffn->b_down;
        kp.y          = ffn->y;
        kp.neuron_idx = ffn->neuron_idx;
        kp.hidden_out = ffn->hidden_out;
        kp.stats_out  = ffn->stats_out;
        kp.m          = ffn->m;
        kp.n_ff       = ffn->n_ff;
        kp.arch       = ffn->arch;
        kp.flags      = ffn->flags;
        const int32_t * lh = (const int32_t *) list_in;
        kp.lin_cnt    = lh + SPIF_LIST_HDR;
        kp.lin_rows   = lh + SPIF_LIST_HDR + SPIF_LIST_MAX_BLOCKS;
        const size_t part_bytes = sizeof(float) * (size_t) P * n_embd;
        const size_t stat_bytes = sizeof(int32_t) * (size_t) P * 2;
        SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0 && ws_bytes >= part_bytes + stat_bytes, SPIF_E_WORKSPACE);
        kp.partials  = (float *) ws;
        kp.cta_stats = (int32_t *) ((char *) ws + part_bytes);
        kp.cap_ffn   = (ffn->m + P - 1) / P + 2;
    } else {
        kp.cap_ffn = 1;
    }
    if (pv) {
        kp.has_pred   = 1;
        kp.wpu        = (const unsigned char *) pv->wpu;
        kp.wpd        = (const unsigned char *) pv->wpd;
        kp.bpu        = pv->bpu;
        kp.bpd        = pv->bpd;
        kp.p_out      = pv->sparse_idx_out;
        kp.logits_out = pv->logits_out;
        kp.r          = pv->r;
        kp.pm         = pv->m;
        kp.Vr         = pv->r / 8;
        kp.RP         = V / kp.Vr;
        kp.WR         = kp.Vr / 32;
        kp.thr_next   = pv->thr;
        kp.pflags     = pv->flags;
        if (!ffn) {
            kp.flags = pv->flags & SPIF_FLAG_CPU_ROUNDING;  // x rounding of a predictor-only launch
        }
        if (list_out) {
            int32_t * lh = (int32_t *) list_out;
            kp.lout_hdr  = lh;
            kp.lout_cnt  = lh + SPIF_LIST_HDR;
            kp.lout_rows = lh + SPIF_LIST_HDR + SPIF_LIST_MAX_BLOCKS;
        }
        kp.cap_pd = (pv->m + P - 1) / P + 2;
        kp.cap_pu = (pv->r + P - 1) / P + 2;
    } else {
        kp.cap_pd = 1, kp.cap_pu = 1;
        kp.Vr = 32, kp.RP = 1, kp.WR = 1, kp.r = 256;
    }
    kp.lin_nblk          = P;  // lists on a device are always built with one block per SM
    // SPIF_LAYER_IMPL=0 selects the first version (issuer thread + per-slot descriptors, below): kept as the A/B reference of
    // profiles/r2_bench_n1_layer_tmaring.json; the register-direct and cp.async versions in between live in the git history only
    static const int k_impl = [] { const char * e = getenv("SPIF_LAYER_IMPL"); return e ? atoi(e) : 3; }();
    if (io && (io->norm_w || io->residual)) {
        SPIF_REQUIRE(k_impl != 0, SPIF_E_UNSUPPORTED);
        SPIF_REQUIRE(((uintptr_t) io->norm_w & 15) == 0 && ((uintptr_t) io->residual & 15) == 0, SPIF_E_SHAPE);
        SPIF_REQUIRE(!io->norm_w || io->norm_eps >= 0.0f, SPIF_E_SHAPE);
        SPIF_REQUIRE(!io->residual || ffn, SPIF_E_NULL);
        kp.norm_w   = io->norm_w;
        kp.norm_eps = io->norm_eps;
        kp.residual = io->residual;
    }
    if (k_impl != 0) {  // version 6: row tables + a producer warp + TMA bulk copies into a ring of 4-row slots (spif_layer_tma.cuh)
        const int vpt4 = (V <= 640) ? 1 : 2;
        kp.nw          = (V + 32 * vpt4 - 1) / (32 * vpt4);
        SPIF_REQUIRE(kp.nw >= 2, SPIF_E_UNSUPPORTED);
        const int threads4   = (kp.nw + 4) * 32;  // consumers + epilogue + 2 helpers + producer
        const int slot_bytes = (4 / vpt4) * kp.row_bytes;
        int       nslots     = 8;
        Layer4Smem L4        = layer4_smem_layout(nslots, slot_bytes, kp.cap_ffn, kp.cap_pd, kp.cap_pu, pv ? pv->r : 0, P);
        while (nslots > 2 && L4.total > (uint32_t) ctx.max_smem_optin) {
            --nslots;
            L4 = layer4_smem_layout(nslots, slot_bytes, kp.cap_ffn, kp.cap_pd, kp.cap_pu, pv ? pv->r : 0, P);
        }
        SPIF_REQUIRE(L4.total <= (uint32_t) ctx.max_smem_optin, SPIF_E_SHAPE);
        // the final column-sliced sum reuses the ring as scratch: [(2 nw + 4) x 16] float4
        SPIF_REQUIRE((size_t) nslots * slot_bytes >= (size_t) (kp.nw * 2 + 4) * 16 * 16, SPIF_E_SHAPE);
        kp.nslots = nslots;
        if (ffn && ffn->hidden_out) {
            SPIF_CUDA_TRY(cudaMemsetAsync(ffn->hidden_out, 0, sizeof(float) * (size_t) ffn->n_ff, ctx.stream));
        }
        const bool h16 = type == SPIF_TYPE_F16;
        const bool rnd = (!ffn || (ffn->flags & SPIF_FLAG_CPU_ROUNDING)) && (!pv || (pv->flags & SPIF_FLAG_CPU_ROUNDING));
#define SPIF_L4_GO(WT_, VPT_, MAXT_)                                                                                       \
    (rnd ? launch_layer4_t<WT_, VPT_, MAXT_, true>(kp, P, threads4, L4.total, ctx.device, ctx.stream) :                    \
           launch_layer4_t<WT_, VPT_, MAXT_, false>(kp, P, threads4, L4.total, ctx.device, ctx.stream))
        if (vpt4 == 2) {
            return h16 ? SPIF_L4_GO(spif_f16, 2, 640) : SPIF_L4_GO(spif_bf16, 2, 640);
        }
        return h16 ? SPIF_L4_GO(spif_f16, 1, 768) : SPIF_L4_GO(spif_bf16, 1, 768);
#undef SPIF_L4_GO
    }
    const int rps        = (vpt == 1 && kp.row_bytes <= 12288) ? 4 : 2;  // weight rows per ring slot
    const int slot_bytes = rps * kp.row_bytes;
    int       nslots     = 12;
    LayerSmem L          = layer_smem_layout(nslots, slot_bytes, kp.cap_ffn, kp.cap_pd, kp.cap_pu, pv ? pv->r : 0, P);
    while (nslots > 3 && L.total > (uint32_t) ctx.max_smem_optin) {
        --nslots;
        L = layer_smem_layout(nslots, slot_bytes, kp.cap_ffn, kp.cap_pd, kp.cap_pu, pv ? pv->r : 0, P);
    }
    SPIF_REQUIRE(L.total <= (uint32_t) ctx.max_smem_optin, SPIF_E_SHAPE);
    // the final column-sliced sum reuses the ring as scratch: [(G + 4) x 16] float4
    SPIF_REQUIRE((size_t) nslots * slot_bytes >= (size_t) (kp.nw * 2 + 4) * 16 * 16, SPIF_E_SHAPE);
    SPIF_REQUIRE(!pv || rps * vpt * 32 >= 1, SPIF_E_SHAPE);
    kp.nslots = nslots;
    if (ffn && ffn->hidden_out) {
        SPIF_CUDA_TRY(cudaMemsetAsync(ffn->hidden_out, 0, sizeof(float) * (size_t) ffn->n_ff, ctx.stream));
    }
    const bool f16 = type == SPIF_TYPE_F16;
    if (vpt == 2) {
        return f16 ? launch_layer_t<spif_f16, 2, 2, 640>(kp, P, threads, L.total, ctx.device, ctx.stream) :
                     launch_layer_t<spif_bf16, 2, 2, 640>(kp, P, threads, L.total, ctx.device, ctx.stream);
    }
    if (rps == 2) {
        return f16 ? launch_layer_t<spif_f16, 1, 2, 1024>(kp, P, threads, L.total, ctx.device, ctx.stream) :
                     launch_layer_t<spif_bf16, 1, 2, 1024>(kp, P, threads, L.total, ctx.device, ctx.stream);
    }
    if (threads <= 768) {
        return f16 ? launch_layer_t<spif_f16, 1, 4, 768>(kp, P, threads, L.total, ctx.device, ctx.stream) :
                     launch_layer_t<spif_bf16, 1, 4, 768>(kp, P, threads, L.total, ctx.device, ctx.stream);
    }
    return f16 ? launch_layer_t<spif_f16, 1, 4, 1024>(kp, P, threads, L.total, ctx.device, ctx.stream) :
                 launch_layer_t<spif_bf16, 1, 4, 1024>(kp, P, threads, L.total, ctx.device, ctx.stream);
}
