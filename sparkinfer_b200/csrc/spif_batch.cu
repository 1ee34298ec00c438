// spif_batch.cu -- batched (2 <= n_tok <= 32) sparse FFN for F16/BF16 weights on sm_100a tensor cores.
//
// Reference: mul_mat_batch_sparse (mm-sparse.cu:107-210) and mul_mat_axpy_sparse_rowwise with T > 1 (axpy-sparse.cu:16-86)
// launch one CTA per (row, token): every token re-streams every active row -> T x the weight traffic.  Here the UNION of
// the per-token masks becomes a dense gathered GEMM, every union row is read from HBM exactly once per matrix:
//   prep   : bits[row] = OR_t active(row, t) << t                 (spif_batch_bits_kernel, also x F32 -> F16/BF16)
//            union_idx = ascending rows with bits != 0            (spif_batch_union_kernel, ballot + prefix scan)
//   gate/up: CTA c owns a contiguous slice of the union; D[64 tokens x (up rows | gate rows)] = X . W[rows,:]^T with
//            tcgen05.mma (A = x tile, B = gathered weight rows, both K-major, cp.async into SWIZZLE_128B tiles;
//            accumulators in TMEM), epilogue applies the per-token mask bits + bias + activation and writes H[t][slot]
//   down   : Y[T x n_embd] = H . Wdown^T[union rows, :] as a stream-K GEMM over (512-column tile, 64-row chunk) items;
//            A = gathered 1 KB row segments of W_down^T (MN-major operand), B = H (K-major); each CTA accumulates a
//            contiguous item range in TMEM and writes <= 2 partial tiles, summed in a fixed order by the last kernel.
// x and the hidden values are rounded to the weight float type (what the reference CPU backend does: ggml-cpu.c:1832-1856,
// 2196-2207); accumulation is FP32 in tensor memory.
#include "spif_batch.cuh"

#include <cstdlib>

#define SPIF_B_ROWS_MAX 1024 /* union rows one CTA may own (m <= 1024 * SMs) */
// producer threads of the two GEMM kernels: copying F16 rows needs few, dequantising Q8_0/Q4_0 blocks is ALU work
#define BATCH_NP(QT) ((QT) ? 384 : 128)

struct BatchParams {
    const unsigned char * w_up;
    const unsigned char * w_gate;
    const unsigned char * w_down;
    const float *         b_up;
    const float *         b_gate;
    const float *         b_down;
    const float *         x;           // [n_tok][n_embd]
    const float *         mask;        // [n_tok][mask_len]
    const int32_t *       neuron_idx;  // nullable [m]
    float *               y;           // [n_tok][n_embd]
    float *               hidden_out;  // nullable [n_tok][n_ff]
    int32_t *             stats_out;   // nullable [n_tok][4]
    // workspace
    uint32_t * bits;         // [m]
    int32_t *  union_idx;    // [m]
    int32_t *  union_count;  // [4]: {U, rows with a non-zero hidden value, 0, 0}
    uint16_t * xh;           // [TP][n_embd]
    uint16_t * hh;           // [TP][m_pad]
    float *    part;         // [P][2][TP][512]
    int        n_embd, n_ff, m, m_pad, mask_len, n_tok;
    int        row_bytes;
    int        arch, gated, by_row;
    float      thr;
    int        debug;  // timing experiments only (SPIF_BATCH_DEBUG): 1 = no MMAs, 2 = no weight loads
};

// ---------------------------------------------------------------------------------------------------------------
// prep 1: per-row token bits; x -> weight float type (rows n_tok..TP-1 are zero)
template <typename WT>
__global__ void __launch_bounds__(256) spif_batch_bits_kernel(const BatchParams p, int nb_bits, int TP) {
    const int tid = threadIdx.x;
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    if ((int) blockIdx.x < nb_bits) {
        const int row = blockIdx.x * 256 + tid;
        if (row < p.m) {
            const int neu = p.neuron_idx ? __ldg(p.neuron_idx + row) : row;
            const int mi  = p.by_row ? row : neu;
            uint32_t  b   = 0;
            for (int t = 0; t < p.n_tok; ++t) {
                b |= (spif_is_active(__ldg(p.mask + (size_t) t * p.mask_len + mi), p.thr) ? 1u : 0u) << t;
            }
            p.bits[row] = b;
        }
        return;
    }
    const int nvec = TP * (p.n_embd >> 3);
    for (int v = ((int) blockIdx.x - nb_bits) * 256 + tid; v < nvec; v += (gridDim.x - nb_bits) * 256) {
        const int t  = v / (p.n_embd >> 3);
        const int c8 = v - t * (p.n_embd >> 3);
        uint4     o  = make_uint4(0u, 0u, 0u, 0u);
        if (t < p.n_tok) {
            const float4 a = __ldg(reinterpret_cast<const float4 *>(p.x + (size_t) t * p.n_embd) + 2 * c8);
            const float4 b = __ldg(reinterpret_cast<const float4 *>(p.x + (size_t) t * p.n_embd) + 2 * c8 + 1);
            o.x = (uint32_t) f32_to_wt_bits<WT>(a.x) | ((uint32_t) f32_to_wt_bits<WT>(a.y) << 16);
            o.y = (uint32_t) f32_to_wt_bits<WT>(a.z) | ((uint32_t) f32_to_wt_bits<WT>(a.w) << 16);
            o.z = (uint32_t) f32_to_wt_bits<WT>(b.x) | ((uint32_t) f32_to_wt_bits<WT>(b.y) << 16);
            o.w = (uint32_t) f32_to_wt_bits<WT>(b.z) | ((uint32_t) f32_to_wt_bits<WT>(b.w) << 16);
        }
        reinterpret_cast<uint4 *>(p.xh)[v] = o;
    }
}

// prep 2: union list (one CTA).  Warp w owns a contiguous row chunk; a lane keeps the ballots of up to 32 iterations in
// registers (all loads in flight at once), so the list comes out ascending after ONE pass + a cross-warp prefix.
__global__ void __launch_bounds__(1024) spif_batch_union_kernel(const BatchParams p) {
    __shared__ int warp_cnt[32];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    const int      chunk = ((p.m + nwarp * 32 - 1) / (nwarp * 32)) * 32;
    const int      r0 = warp * chunk, r1 = min(r0 + chunk, p.m);
    const int      n_it = chunk >> 5;
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    unsigned ball[32];
    int      cnt = 0;
    if (n_it <= 32) {
        uint32_t v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int row = r0 + i * 32 + lane;
            v[i]          = (i < n_it && row < r1) ? p.bits[row] : 0u;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            ball[i] = __ballot_sync(0xffffffffu, v[i] != 0u);
            cnt += __popc(ball[i]);
        }
    } else {
        for (int base = r0; base < r1; base += 32) {
            const int row = base + lane;
            cnt += __popc(__ballot_sync(0xffffffffu, (row < r1) && (p.bits[row] != 0u)));
        }
    }
    if (lane == 0) {
        warp_cnt[warp] = cnt;
    }
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < nwarp; ++w) {
        const int v = warp_cnt[w];
        before += (w < warp) ? v : 0;
        total += v;
    }
    int run = before;
    if (n_it <= 32) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if (ball[i] & (1u << lane)) {
                p.union_idx[run + __popc(ball[i] & ((1u << lane) - 1u))] = r0 + i * 32 + lane;
            }
            run += __popc(ball[i]);
        }
    } else {
        for (int base = r0; base < r1; base += 32) {
            const int      row = base + lane;
            const bool     a   = (row < r1) && (p.bits[row] != 0u);
            const unsigned b   = __ballot_sync(0xffffffffu, a);
            if (a) {
                p.union_idx[run + __popc(b & ((1u << lane) - 1u))] = row;
            }
            run += __popc(b);
        }
    }
    if (tid == 0) {
        p.union_count[0] = total;
        p.union_count[1] = 0;
    }
    if (p.stats_out && tid < p.n_tok * 4) {
        p.stats_out[tid] = 0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// gate/up gathered GEMM.  160 threads: warps 0-3 = cp.async producers (warps 0-1 also run the epilogue); warp 4 = TMEM
// owner + MMA issuer.  K block = 64 elements (one 128-byte swizzle row).  A stage holds
//   [up tile][gate tile][x tile TP x 128 B]      (K-major SWIZZLE_128B, 8-row atoms of 1 KB)
// for one K block; a weight tile only occupies ceil(rows / 8) atoms, so a CTA with few rows gets a deep ring (up to 32
// K blocks in flight).
// MMA orientation: D[64 tokens x N] = X[64 x 64] . W[N x 64]^T with N = up rows followed by gate rows (the two tiles are
// adjacent in shared memory, so ONE instruction per K = 16 step covers both matrices).  Measured on B200: a tcgen05.mma
// with both operands in shared memory costs ~1 cycle per A-operand row whatever N is (M = 128 weight rows x N = 16 tokens:
// ~135 cycles per instruction, 8 per K block -> the kernel was MMA-issue bound at 45 us); with the tokens as the (M = 64)
// A operand the weight rows ride in the cheap N dimension.  Rows 32..63 of the x tile are whatever follows it in shared
// memory: they only produce accumulator lanes nobody reads (K1_SLACK keeps those reads inside the allocation).
#define K1_MAXST 32
#define K1_RING  (190 * 1024)
#define K1_SLACK (16 * 1024)
#define K1_TAIL  (1024 + 4 * SPIF_B_ROWS_MAX + 2 * 128 * 16 + 256)
#define K1_SMEM  (K1_RING + K1_SLACK + K1_TAIL + 1024)

// QT = 0: F16/BF16 rows copied as they are; QT = Q8_0 / Q4_0: raw blocks are staged in shared memory (cp.async, groups of
// 4 K blocks = 8 quant blocks per row, double-buffered) and dequantised to F16 operand tiles by the producer warps.
template <typename WT, int TP, int QT>
__global__ void __launch_bounds__(BATCH_NP(QT) + 32, 1) spif_batch_gateup_kernel(const BatchParams p) {
    constexpr int NP = BATCH_NP(QT);  // producer threads; warp NP / 32 owns TMEM and issues the MMAs
    constexpr int TR = (QT == SPIF_TYPE_Q8_0) ? 64 : ((QT == SPIF_TYPE_Q4_0) ? 96 : 128);  // rows per tile
    extern __shared__ unsigned char smem_raw[];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int      P = gridDim.x, c = blockIdx.x;
    const uint32_t raw0  = smem_u32(smem_raw);
    const uint32_t sbase = (raw0 + 1023u) & ~1023u;
    unsigned char * sm   = smem_raw + (sbase - raw0);
    unsigned char * tail = sm + K1_RING + K1_SLACK;
    uint64_t *      full       = reinterpret_cast<uint64_t *>(tail);
    uint64_t *      empty      = full + K1_MAXST;
    uint64_t *      accum_full = empty + K1_MAXST;
    uint64_t *      tmem_empty = accum_full + 1;
    uint32_t *      tmem_slot  = reinterpret_cast<uint32_t *>(tmem_empty + 1);
    int *           s_rows     = reinterpret_cast<int *>(tail + 1024);
    // per-tile row metadata for the epilogue, double-buffered by tile parity: [2][128] x {bits, neuron, b_up, b_gate}
    uint32_t *      s_bits     = reinterpret_cast<uint32_t *>(tail + 1024 + 4 * SPIF_B_ROWS_MAX);
    int *           s_neu      = reinterpret_cast<int *>(s_bits + 256);
    float *         s_bu       = reinterpret_cast<float *>(s_neu + 256);
    float *         s_bg       = s_bu + 256;
    uint32_t *      s_nz       = reinterpret_cast<uint32_t *>(s_bg + 256);  // [2 warps][8 chunks] rows with a non-zero hidden

    if (tid == 0) {
        for (int s = 0; s < K1_MAXST; ++s) {
            mbar_init(&full[s], NP);
            mbar_init(&empty[s], 1);
        }
        mbar_init(accum_full, 1);
        mbar_init(tmem_empty, 64);
        mbar_fence_init();
    }
    if (warp == NP / 32) {
        tmem_alloc(tmem_slot, 256);
    }
    pdl_wait_prior_grid();  // union list, bits, xh
    pdl_launch_dependents();

    const int U   = p.union_count[0];
    const int q   = (U + P - 1) / P;
    const int s0  = c * q;
    const int cnt = max(0, min(U - s0, q));
    if (c == 0 && tid < p.n_tok && p.stats_out) {
        p.stats_out[tid * 4 + 2] = U;
    }
    for (int i = tid; i < cnt; i += blockDim.x) {
        s_rows[i] = p.union_idx[s0 + i];
    }
    if (cnt > 0 && s0 + cnt == U) {
        // owner of the last slice: slots U .. roundup(U, 8) feed the down GEMM's 16-byte H chunks and must be finite
        const int pad = ((U + 7) & ~7) - U;
        for (int i = tid; i < pad * TP; i += blockDim.x) {
            p.hh[(size_t) (i / pad) * p.m_pad + U + (i % pad)] = 0;
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    const int      KB      = p.n_embd >> 6;
    const int      n_tiles = (cnt + TR - 1) / TR;
    const bool     gated   = p.gated != 0;
    const int      R8      = (min(cnt, TR) + 7) >> 3;          // 8-row atoms per weight tile
    const uint32_t a_bytes = (uint32_t) R8 * 1024u;
    const uint32_t x_off   = a_bytes * (gated ? 2u : 1u);
    const uint32_t stage_b = x_off + TP * 128;
    // quantised weights: [raw ring: 2 buffers x {up, gate} x TR rows x GB bytes][operand ring]
    constexpr int  GB      = (QT == SPIF_TYPE_Q8_0) ? 272 : ((QT == SPIF_TYPE_Q4_0) ? 144 : 0);  // 8 quant blocks
    constexpr uint32_t RAWB = (uint32_t) (2 * TR * GB);                                         // one raw buffer
    constexpr uint32_t RAW_ALL = (2 * RAWB + 1023u) & ~1023u;
    const uint32_t ring0   = sbase + RAW_ALL;
    uint32_t       NST     = min((uint32_t) K1_MAXST, ((uint32_t) K1_RING - RAW_ALL) / stage_b);
    if (QT != 0) {
        NST &= ~1u;  // the dequantiser fills stages in pairs
    }
    const int      ncol_up = R8 * 8;                            // accumulator columns: [up rows][gate rows]

    if (cnt > 0 && warp < NP / 32) {
        const int cc = tid & 7, rsub = tid >> 3;
        uint32_t  st = 0, lap = 0;
        for (int tile = 0; tile < n_tiles; ++tile) {
            const int cnt_t = min(TR, cnt - tile * TR);
            const int buf   = (tile & 1) * 128;
            if (tid < cnt_t) {
                const int row    = s_rows[tile * TR + tid];
                const int neu    = p.neuron_idx ? __ldg(p.neuron_idx + row) : row;
                s_bits[buf + tid] = p.bits[row];
                s_neu[buf + tid]  = neu;
                s_bu[buf + tid]   = p.b_up ? __ldg(p.b_up + neu) : 0.f;
                s_bg[buf + tid]   = (gated && p.b_gate) ? __ldg(p.b_gate + neu) : 0.f;
            }
            named_bar_sync(1, NP);
            int rows[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int r = rsub + 16 * j;
                rows[j]     = (r < cnt_t) ? s_rows[tile * TR + r] : -1;
            }
            if constexpr (QT == 0) {
            // CTAs walk K in rotated order so that they do not all read the same slice of x at the same moment
                int kb = (int) (((unsigned) c * 29u) % (unsigned) KB);
                for (int ki = 0; ki < KB; ++ki, kb = (kb + 1 == KB) ? 0 : kb + 1) {
                    if (lap > 0) {
                        mbar_wait(&empty[st], (lap - 1u) & 1u);
                    }
                    const uint32_t sA = ring0 + st * stage_b;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (rows[j] >= 0 && p.debug != 2) {
                            const uint32_t off = swz_kmajor(rsub + 16 * j, cc);
                            const size_t   go  = (size_t) rows[j] * p.row_bytes + (size_t) kb * 128 + cc * 16;
                            cp_async16(sA + off, p.w_up + go);
                            if (gated) {
                                cp_async16(sA + a_bytes + off, p.w_gate + go);
                            }
                        }
                    }
#pragma unroll
                    for (int h = 0; h < TP * 8 / 128; ++h) {
                        const int idx = tid + 128 * h, t = idx >> 3, c8 = idx & 7;
                        cp_async16(sA + x_off + swz_kmajor(t, c8), p.xh + (size_t) t * p.n_embd + kb * 64 + c8 * 8);
                    }
                    cp_async_mbar_arrive_noinc(&full[st]);
                    if (++st == NST) {
                        st = 0;
                        ++lap;
                    }
                }
            } else {
                using QP = qpair<QT>;
                const int      nmat = gated ? 2 : 1;
                const int      NG   = KB >> 2;
                const int      RP   = (cnt_t + 31) & ~31;
                constexpr int  NCH  = GB / 16;
                const uint32_t raw0s = sbase;
                auto issue_raw = [&](int g, int buf) {
                    const int total = nmat * cnt_t * NCH;
                    for (int idx = tid; idx < total; idx += NP) {
                        const int unit = idx / NCH, ch = idx - unit * NCH;
                        const int mt = unit >= cnt_t ? 1 : 0, r = unit - mt * cnt_t;
                        const unsigned char * src = (mt ? p.w_gate : p.w_up) + (size_t) s_rows[tile * TR + r] * p.row_bytes +
                                                    (size_t) g * GB + ch * 16;
                        cp_async16(raw0s + buf * RAWB + (uint32_t) ((mt * TR + r) * GB + ch * 16), src);
                    }
                    cp_async_commit();
                };
                int g = (int) (((unsigned) c * 7u) % (unsigned) NG);
                issue_raw(g, 0);
                for (int gi = 0; gi < NG; ++gi) {
                    const int buf = gi & 1, gn = (g + 1 == NG) ? 0 : g + 1;
                    if (gi + 1 < NG) {
                        issue_raw(gn, buf ^ 1);
                    } else {
                        cp_async_commit();
                    }
                    cp_async_wait<1>();
                    named_bar_sync(1, NP);  // everybody's share of group gi has landed
                    const unsigned char * rawp = sm + buf * RAWB;
                    for (int h = 0; h < 2; ++h) {
                        if (lap > 0) {
                            mbar_wait(&empty[st], (lap - 1u) & 1u);
                            mbar_wait(&empty[st + 1], (lap - 1u) & 1u);
                        }
                        const uint32_t sA0 = ring0 + st * stage_b;
                        // x tiles of the two K blocks
                        for (int idx = tid; idx < 2 * TP * 8; idx += NP) {
                            const int s = idx / (TP * 8), rem = idx - s * (TP * 8), t = rem >> 3, c8 = rem & 7;
                            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p.xh + (size_t) t * p.n_embd + (g * 4 + h * 2 + s) * 64) + c8);
                            *reinterpret_cast<uint4 *>(sm + (sA0 - sbase) + s * stage_b + x_off + swz_kmajor(t, c8)) = v;
                        }
                        // weights: task = (row, matrix, K block of the pair): 2 quant blocks -> 64 F16 = one 128-byte tile row
                        for (int id = tid; id < RP * nmat * 2; id += NP) {
                            const int r = id % RP, rest = id / RP, mt = rest % nmat, s = rest / nmat;
                            if (r < cnt_t) {
                                const uint32_t * wp = reinterpret_cast<const uint32_t *>(rawp + (size_t) (mt * TR + r) * GB + (h * 2 + s) * (2 * QP::BS));
                                uint32_t w[QP::NW];
#pragma unroll
                                for (int i = 0; i < QP::NW; ++i) {
                                    w[i] = wp[i];
                                }
                                uint4 out[8];
                                QP::dequant(w, out);
                                unsigned char * dst = sm + (sA0 - sbase) + s * stage_b + mt * a_bytes;
#pragma unroll
                                for (int c8 = 0; c8 < 8; ++c8) {
                                    *reinterpret_cast<uint4 *>(dst + swz_kmajor(r, c8)) = out[c8];
                                }
                            }
                        }
                        fence_proxy_async_smem();
                        mbar_arrive(&full[st]);
                        mbar_arrive(&full[st + 1]);
                        st += 2;
                        if (st == NST) {
                            st = 0;
                            ++lap;
                        }
                    }
                    named_bar_sync(1, NP);  // raw[buf] may be overwritten by the copy issued in the next iteration
                    g = gn;
                }
            }
            if (warp >= 2) {
                continue;
            }
            // ---- epilogue (warps 0-1): token t lives in TMEM lane (t % 16) + 32 * (t / 16); thread = token, loop over rows
            mbar_wait(accum_full, tile & 1);
            tc_fence_after_sync();
            const int      t       = warp * 16 + lane;
            const bool     tok_ok  = lane < 16 && t < p.n_tok;
            const uint32_t taddr   = tmem_base + ((uint32_t) (warp * 32) << 16);
            const int      slot0   = s0 + tile * TR;
            int            n_act = 0, n_nz = 0;
            for (int r0 = 0; r0 < cnt_t; r0 += 16) {
                float up[16], gt[16];
                tmem_ld16(taddr + r0, up);
                if (gated) {
                    tmem_ld16(taddr + ncol_up + r0, gt);
                }
                uint32_t nzmask = 0;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int r = r0 + j;
                    if (r < cnt_t) {
                        const bool act = tok_ok && ((s_bits[buf + r] >> t) & 1u);
                        float      h   = 0.f;
                        if (act) {
                            h = spif_activation(p.arch, gated ? gt[j] + s_bg[buf + r] : 0.f, up[j] + s_bu[buf + r]);
                            if (p.hidden_out) {
                                p.hidden_out[(size_t) t * p.n_ff + s_neu[buf + r]] = h;
                            }
                        }
                        const uint16_t hb = f32_to_wt_bits<WT>(saturate_to_wt<WT>(h));
                        const bool     nz = (hb & 0x7fffu) != 0;
                        if (lane < 16 && t < TP) {
                            p.hh[(size_t) t * p.m_pad + slot0 + r] = hb;
                        }
                        n_act += act ? 1 : 0;
                        n_nz += (act && nz) ? 1 : 0;
                        nzmask |= nz ? (1u << j) : 0u;
                    }
                }
                nzmask = __reduce_or_sync(0xffffffffu, nzmask);
                if (lane == 0) {
                    s_nz[warp * 8 + (r0 >> 4)] = nzmask;
                }
            }
            tc_fence_before_sync();
            mbar_arrive(tmem_empty);
            if (p.stats_out && tok_ok) {
                atomicAdd(p.stats_out + t * 4 + 0, n_act);
                atomicAdd(p.stats_out + t * 4 + 1, n_nz);
            }
            named_bar_sync(2, 64);
            if (tid < ((cnt_t + 15) >> 4)) {
                const int z = __popc(s_nz[tid] | s_nz[8 + tid]);
                if (z) {
                    atomicAdd(p.union_count + 1, z);
                }
            }
            named_bar_sync(2, 64);
        }
    } else if (cnt > 0) {
        const uint32_t idesc = umma_idesc_f16(wt_fmt<WT>::id, 0, 0, 64, gated ? 2 * ncol_up : ncol_up);
        uint32_t       st = 0, lap = 0;
        for (int tile = 0; tile < n_tiles; ++tile) {
            if (tile > 0) {
                mbar_wait(tmem_empty, (tile - 1) & 1);
                tc_fence_after_sync();
            }
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait(&full[st], lap & 1u);
                tc_fence_after_sync();
                if (lane == 0) {
                    const uint32_t sA = ring0 + st * stage_b;
#pragma unroll
                    for (int k = 0; k < (p.debug == 1 ? 0 : 4); ++k) {
                        umma_f16(tmem_base, umma_smem_desc(sA + x_off + k * 32, 16, 1024), umma_smem_desc(sA + k * 32, 16, 1024),
                                 idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[st]);
                    if (kb == KB - 1) {
                        umma_commit(accum_full);
                    }
                }
                __syncwarp();
                if (++st == NST) {
                    st = 0;
                    ++lap;
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == NP / 32) {
        tmem_dealloc(tmem_base, 256);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// down projection: stream-K over items (ct = 512-column tile of n_embd, kc = chunk of 64 union rows), item id = ct * n_kc + kc.
// CTA c owns items [c * ipc, (c + 1) * ipc): at most two column tiles ("segments"), four M = 128 accumulators each.
//   A tile  : W_down^T[64 union rows][1 KB column segment], MN-major SWIZZLE_128B.  Column block b (128 columns) = two
//             64-column halves (LBO 8 KB apart), each half = 8 atoms of 8 k-rows x 128 B (SBO 1 KB)  -> 16 KB per block
//   B tile  : H[t][64 slots], K-major SWIZZLE_128B, TP rows of 128 B
#define K2_NST   3
#define K2_ATILE 65536
#define K2_STAGE (K2_ATILE + 4096)
#define K2_SMEM  (K2_NST * K2_STAGE + 256 + 1024)

// QT = Q8_0 / Q4_0: the 64 x 1 KB-of-columns item is 64 x 16 quant blocks; raw blocks are staged (cp.async, two buffers in
// place of the third operand stage) and dequantised by the producer warps into the same MN-major operand tile.
template <typename WT, int TP, int QT>
__global__ void __launch_bounds__(BATCH_NP(QT) + 32, 1) spif_batch_down_kernel(const BatchParams p) {
    constexpr int NP = BATCH_NP(QT);
    constexpr uint32_t NSTK = QT ? 2 : K2_NST;
    constexpr int      SEGB = (QT == SPIF_TYPE_Q8_0) ? 544 : ((QT == SPIF_TYPE_Q4_0) ? 288 : 0);  // 16 quant blocks
    constexpr uint32_t RAWB = 64u * SEGB;
    constexpr uint32_t RAW_ALL = QT ? (uint32_t) K2_STAGE : 0u;  // raw ring lives where the third stage would be
    extern __shared__ unsigned char smem_raw[];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int      P = gridDim.x, c = blockIdx.x;
    const uint32_t raw0  = smem_u32(smem_raw);
    const uint32_t sbase = (raw0 + 1023u) & ~1023u;
    unsigned char * sm   = smem_raw + (sbase - raw0);
    unsigned char * tail = sm + K2_NST * K2_STAGE;
    uint64_t *      full       = reinterpret_cast<uint64_t *>(tail);
    uint64_t *      empty      = full + K2_NST;
    uint64_t *      accum_full = empty + K2_NST;  // [2]
    uint32_t *      tmem_slot  = reinterpret_cast<uint32_t *>(accum_full + 2);

    if (tid == 0) {
        for (int s = 0; s < K2_NST; ++s) {
            mbar_init(&full[s], NP);
            mbar_init(&empty[s], 1);
        }
        mbar_init(&accum_full[0], 1);
        mbar_init(&accum_full[1], 1);
        mbar_fence_init();
    }
    if (warp == NP / 32) {
        tmem_alloc(tmem_slot, 256);
    }
    pdl_wait_prior_grid();  // H, union list
    pdl_launch_dependents();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t ring0     = sbase + RAW_ALL;
    static_assert(2 * RAWB <= K2_STAGE, "raw ring does not fit");

    const int  U    = p.union_count[0];
    const int  U8   = (U + 7) & ~7;
    const int  n_kc = (U + 63) >> 6;
    const int  n_ct = p.n_embd >> 9;
    const int  I    = n_ct * n_kc;
    const int  ipc  = (I + P - 1) / P;
    const int  i0   = c * ipc;
    const int  i1   = min(I, i0 + ipc);
    const bool work = i0 < i1;
    const int  ct0  = work ? i0 / n_kc : 0;
    const int  nseg = work ? (i1 - 1) / n_kc - ct0 + 1 : 0;  // 1 or 2 (launcher guarantees n_ct <= P)

    if (work && warp < NP / 32) {
        const int ch = tid & 63, ksub = tid >> 6;
        // byte offset of this thread's 16-byte column chunk inside a k-row's swizzle atom set
        const uint32_t ch_off = (uint32_t) ((ch >> 4) * 16384 + ((ch >> 3) & 1) * 8192);
        uint32_t       st = 0, lap = 0;
        if constexpr (QT == 0) {
        for (int i = i0; i < i1; ++i) {
                const int ct = i / n_kc, kc = i - ct * n_kc;
                if (lap > 0) {
                    mbar_wait(&empty[st], (lap - 1u) & 1u);
                }
                const uint32_t        sA   = ring0 + st * K2_STAGE;
                const unsigned char * wcol = p.w_down + (size_t) ct * 1024 + ch * 16;
#pragma unroll 8
                for (int j = 0; j < 32; ++j) {
                    const int      k     = ksub + 2 * j;
                    const int      slot  = kc * 64 + k;
                    const bool     valid = slot < U;
                    const int      row   = valid ? __ldg(p.union_idx + slot) : 0;
                    const uint32_t off   = ch_off + (uint32_t) ((k >> 3) * 1024 + (k & 7) * 128 + (((ch & 7) ^ (k & 7)) << 4));
                    cp_async16_zfill(sA + off, wcol + (size_t) row * p.row_bytes, valid ? 16u : 0u);
                }
#pragma unroll
                for (int h = 0; h < TP * 8 / 128; ++h) {
                    const int  idx = tid + 128 * h, t = idx >> 3, c8 = idx & 7;
                    const int  s8  = kc * 64 + c8 * 8;
                    cp_async16_zfill(sA + K2_ATILE + swz_kmajor(t, c8), p.hh + (size_t) t * p.m_pad + s8, s8 < U8 ? 16u : 0u);
                }
                cp_async_mbar_arrive_noinc(&full[st]);
                if (++st == NSTK) {
                    st = 0;
                    ++lap;
                }
            }
        } else {
            using QP = qpair<QT>;
            constexpr int NCH = SEGB / 16;
            auto issue_raw = [&](int i, int buf) {
                const int ct = i / n_kc, kc = i - ct * n_kc;
                for (int idx = tid; idx < 64 * NCH; idx += NP) {
                    const int kr = idx / NCH, chk = idx - kr * NCH;
                    const int slot = kc * 64 + kr;
                    if (slot < U) {
                        cp_async16(sbase + buf * RAWB + (uint32_t) (kr * SEGB + chk * 16),
                                   p.w_down + (size_t) __ldg(p.union_idx + slot) * p.row_bytes + (size_t) ct * SEGB + chk * 16);
                    }
                }
                cp_async_commit();
            };
            issue_raw(i0, 0);
            for (int i = i0; i < i1; ++i) {
                const int ct = i / n_kc, kc = i - ct * n_kc;
                const int buf = (i - i0) & 1;
                if (i + 1 < i1) {
                    issue_raw(i + 1, buf ^ 1);
                } else {
                    cp_async_commit();
                }
                cp_async_wait<1>();
                named_bar_sync(1, NP);
                if (lap > 0) {
                    mbar_wait(&empty[st], (lap - 1u) & 1u);
                }
                unsigned char *       dstA = sm + RAW_ALL + st * K2_STAGE;
                const unsigned char * rawp = sm + buf * RAWB;
                // H tile
                for (int idx = tid; idx < TP * 8; idx += NP) {
                    const int t = idx >> 3, c8 = idx & 7, s8 = kc * 64 + c8 * 8;
                    uint4     v = make_uint4(0u, 0u, 0u, 0u);
                    if (s8 < U8) {
                        v = *reinterpret_cast<const uint4 *>(p.hh + (size_t) t * p.m_pad + s8);
                    }
                    *reinterpret_cast<uint4 *>(dstA + K2_ATILE + swz_kmajor(t, c8)) = v;
                }
                // weights: task = (k-row, pair of quant blocks = 64 columns = one 128-byte MN-major row of one half)
                for (int id = tid; id < 64 * 8; id += NP) {
                    const int kr = id & 63, pp = id >> 6;
                    uint4     out[8];
                    if (kc * 64 + kr < U) {
                        const uint32_t * wp = reinterpret_cast<const uint32_t *>(rawp + kr * SEGB + pp * (2 * QP::BS));
                        uint32_t         w[QP::NW];
#pragma unroll
                        for (int q = 0; q < QP::NW; ++q) {
                            w[q] = wp[q];
                        }
                        QP::dequant(w, out);
                    } else {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            out[q] = make_uint4(0u, 0u, 0u, 0u);
                        }
                    }
                    unsigned char * d = dstA + (pp >> 1) * 16384 + (pp & 1) * 8192 + (kr >> 3) * 1024 + (kr & 7) * 128;
#pragma unroll
                    for (int c8 = 0; c8 < 8; ++c8) {
                        *reinterpret_cast<uint4 *>(d + ((c8 ^ (kr & 7)) << 4)) = out[c8];
                    }
                }
                fence_proxy_async_smem();
                mbar_arrive(&full[st]);
                if (++st == NSTK) {
                    st = 0;
                    ++lap;
                }
                named_bar_sync(1, NP);
            }
        }
        // ---- epilogue: one partial tile per segment, part[c][seg][t][512 columns]
        for (int seg = 0; seg < (warp < 4 ? nseg : 0); ++seg) {
            mbar_wait(&accum_full[seg], 0);
            tc_fence_after_sync();
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                float          v[TP];
                const uint32_t taddr = tmem_base + ((uint32_t) (warp * 32) << 16) + (seg * 4 + b) * TP;
#pragma unroll
                for (int h = 0; h < TP / 16; ++h) {
                    tmem_ld16(taddr + h * 16, *reinterpret_cast<float(*)[16]>(v + h * 16));
                }
                float * o = p.part + ((size_t) (c * 2 + seg) * TP) * 512 + b * 128 + tid;
#pragma unroll
                for (int t = 0; t < TP; ++t) {
                    if (t < p.n_tok) {
                        o[t * 512] = v[t];
                    }
                }
            }
        }
    } else if (work) {
        constexpr uint32_t idesc = umma_idesc_f16(wt_fmt<WT>::id, 1, 0, 128, TP);
        uint32_t           st = 0, lap = 0;
        for (int i = i0; i < i1; ++i) {
            const int  ct = i / n_kc, kc = i - ct * n_kc;
            const int  seg   = ct - ct0;
            const bool first = (i == i0) || (kc == 0);
            const bool last  = (i == i1 - 1) || (kc == n_kc - 1);
            mbar_wait(&full[st], lap & 1u);
            fence_proxy_async_smem();
            tc_fence_after_sync();
            if (lane == 0) {
                const uint32_t sA = ring0 + st * K2_STAGE;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint64_t db = umma_smem_desc(sA + K2_ATILE + k * 32, 16, 1024);
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const uint64_t da = umma_smem_desc(sA + b * 16384 + k * 2048, 8192, 1024);
                        umma_f16(tmem_base + (seg * 4 + b) * TP, da, db, idesc, (first && k == 0) ? 0u : 1u);
                    }
                }
                umma_commit(&empty[st]);
                if (last) {
                    umma_commit(&accum_full[seg]);
                }
            }
            __syncwarp();
            if (++st == NSTK) {
                st = 0;
                ++lap;
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == NP / 32) {
        tmem_dealloc(tmem_base, 256);
    }
}

// y[t][col] = sum over the CTAs whose item range touched the column's 512-wide tile, ascending CTA order (fixed) (+ b_down)
__global__ void __launch_bounds__(128) spif_batch_reduce_kernel(const BatchParams p, int P, int TP) {
    const int ct = blockIdx.x >> 2, t = blockIdx.y, col = (blockIdx.x & 3) * 128 + threadIdx.x;
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int U    = p.union_count[0];
    const int n_kc = (U + 63) >> 6;
    const int n_ct = p.n_embd >> 9;
    const int I    = n_ct * n_kc;
    float     s    = 0.f;
    if (I > 0) {
        const int ipc = (I + P - 1) / P;
        const int lo = ct * n_kc, hi = lo + n_kc - 1;
        for (int c = lo / ipc; c <= hi / ipc; ++c) {
            const int seg = ct - (c * ipc) / n_kc;
            s += p.part[((size_t) (c * 2 + seg) * TP + t) * 512 + col];
        }
    }
    if (p.b_down) {
        s += __ldg(p.b_down + ct * 512 + col);
    }
    p.y[(size_t) t * p.n_embd + ct * 512 + col] = s;
    if (p.stats_out && blockIdx.x == 0 && threadIdx.x == 0) {
        p.stats_out[t * 4 + 3] = p.union_count[1];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host side
static inline size_t al256(size_t v) {
    return (v + 255) & ~(size_t) 255;
}

size_t spif_batch_workspace_bytes(int n_embd, int n_ff, int n_tok) {
    if (n_tok < 2) {
        return 0;
    }
    const size_t TP = 32, P = 320, m_pad = ((size_t) n_ff + 127) / 128 * 128;
    return al256(4 * (size_t) n_ff) * 2 + 256 + al256(TP * (size_t) n_embd * 2) + al256(TP * m_pad * 2) + al256(P * 2 * TP * 512 * 4);
}

// every kernel of the chain opts into programmatic dependent launch and starts with griddepcontrol.wait
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = grid;
    cfg.blockDim           = block;
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    return cudaLaunchKernelEx(&cfg, kern, args...);
}

template <typename WT, int TP, int QT>
static int launch_batch_t(const BatchParams & bp, int P, const spif_launch_ctx & ctx) {
    static unsigned long long done1 = 0, done2 = 0;
    auto k1 = spif_batch_gateup_kernel<WT, TP, QT>;
    auto k2 = spif_batch_down_kernel<WT, TP, QT>;
    int  rc = spif_set_max_smem_once(k1, K1_SMEM, ctx.device, done1);
    if (rc) {
        return rc;
    }
    rc = spif_set_max_smem_once(k2, K2_SMEM, ctx.device, done2);
    if (rc) {
        return rc;
    }
    const int nb_bits = (bp.m + 255) / 256;
    const int nvec    = TP * (bp.n_embd >> 3);
    const int nb_cvt  = min(64, (nvec + 255) / 256);
    SPIF_CUDA_TRY(launch_pdl(spif_batch_bits_kernel<WT>, dim3(nb_bits + nb_cvt), dim3(256), 0, ctx.stream, bp, nb_bits, (int) TP));
    SPIF_CUDA_TRY(launch_pdl(spif_batch_union_kernel, dim3(1), dim3(1024), 0, ctx.stream, bp));
    SPIF_CUDA_TRY(launch_pdl(k1, dim3(P), dim3(BATCH_NP(QT) + 32), K1_SMEM, ctx.stream, bp));
    SPIF_CUDA_TRY(launch_pdl(k2, dim3(P), dim3(BATCH_NP(QT) + 32), K2_SMEM, ctx.stream, bp));
    SPIF_CUDA_TRY(launch_pdl(spif_batch_reduce_kernel, dim3(bp.n_embd >> 7, bp.n_tok), dim3(128), 0, ctx.stream, bp, P, (int) TP));
    return 0;
}

// returns SPIF_E_UNSUPPORTED when the shape does not qualify: the caller then uses the per-token kernel
int spif_launch_ffn_batch(const spif_ffn_args & a, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx) {
    const bool quant = a.type == SPIF_TYPE_Q8_0 || a.type == SPIF_TYPE_Q4_0;
    SPIF_REQUIRE(a.type == SPIF_TYPE_F16 || a.type == SPIF_TYPE_BF16 || quant, SPIF_E_UNSUPPORTED);
    // Q8_0/Q4_0 with SPIF_FLAG_CPU_ROUNDING asks for the reference CPU arithmetic (x quantised to 8-bit blocks, integer
    // dots): only the per-token kernel reproduces that; the union GEMM multiplies dequantised weights by F16 activations
    SPIF_REQUIRE(!(quant && (a.flags & SPIF_FLAG_CPU_ROUNDING)), SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(a.n_tok >= 2 && a.n_tok <= 32, SPIF_E_UNSUPPORTED);
    const int P = ctx.sm_count;
    SPIF_REQUIRE(a.n_embd % 512 == 0 && (a.n_embd >> 9) <= P && a.m <= SPIF_B_ROWS_MAX * P, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(ctx.max_smem_optin >= K1_SMEM && ctx.max_smem_optin >= K2_SMEM, SPIF_E_UNSUPPORTED);
    const int    TP    = a.n_tok <= 16 ? 16 : 32;
    const size_t m_pad = ((size_t) a.m + 127) / 128 * 128;

    BatchParams bp = {};
    bp.w_up       = (const unsigned char *) a.w_up;
    bp.w_gate     = (const unsigned char *) a.w_gate;
    bp.w_down     = (const unsigned char *) a.w_down_t;
    bp.b_up       = a.b_up;
    bp.b_gate     = a.b_gate;
    bp.b_down     = a.b_down;
    bp.x          = a.x;
    bp.mask       = a.sparse_idx;
    bp.neuron_idx = a.neuron_idx;
    bp.y          = a.y;
    bp.hidden_out = a.hidden_out;
    bp.stats_out  = a.stats_out;
    bp.n_embd     = a.n_embd;
    bp.n_ff       = a.n_ff;
    bp.m          = a.m;
    bp.m_pad      = (int) m_pad;
    bp.by_row     = (a.flags & SPIF_FLAG_MASK_BY_ROW) ? 1 : 0;
    bp.mask_len   = bp.by_row ? a.m : a.n_ff;
    bp.n_tok      = a.n_tok;
    bp.row_bytes  = quant ? (a.n_embd / 32) * (a.type == SPIF_TYPE_Q8_0 ? 34 : 18) : a.n_embd * 2;
    bp.arch       = a.arch;
    bp.gated      = (a.arch != SPIF_ARCH_RELU_UP) ? 1 : 0;
    bp.thr        = a.thr;
    static const int k_debug = [] { const char * e = getenv("SPIF_BATCH_DEBUG"); return e ? atoi(e) : 0; }();
    bp.debug      = k_debug;

    SPIF_REQUIRE(ws != nullptr && ((uintptr_t) ws & 255) == 0, SPIF_E_WORKSPACE);
    char * w = (char *) ws;
    size_t o = 0;
    bp.bits = (uint32_t *) (w + o);
    o += al256(4 * (size_t) a.m);
    bp.union_idx = (int32_t *) (w + o);
    o += al256(4 * (size_t) a.m);
    bp.union_count = (int32_t *) (w + o);
    o += 256;
    bp.xh = (uint16_t *) (w + o);
    o += al256((size_t) TP * a.n_embd * 2);
    bp.hh = (uint16_t *) (w + o);
    o += al256((size_t) TP * m_pad * 2);
    bp.part = (float *) (w + o);
    o += al256((size_t) P * 2 * TP * 512 * 4);
    SPIF_REQUIRE(ws_bytes >= o, SPIF_E_WORKSPACE);

    if (a.hidden_out) {
        SPIF_CUDA_TRY(cudaMemsetAsync(a.hidden_out, 0, sizeof(float) * (size_t) a.n_ff * a.n_tok, ctx.stream));
    }
    // quantised weights are dequantised to F16 operand tiles, so x and H are F16 as well
    if (a.type == SPIF_TYPE_Q8_0) {
        return TP == 16 ? launch_batch_t<spif_f16, 16, SPIF_TYPE_Q8_0>(bp, P, ctx) : launch_batch_t<spif_f16, 32, SPIF_TYPE_Q8_0>(bp, P, ctx);
    }
    if (a.type == SPIF_TYPE_Q4_0) {
        return TP == 16 ? launch_batch_t<spif_f16, 16, SPIF_TYPE_Q4_0>(bp, P, ctx) : launch_batch_t<spif_f16, 32, SPIF_TYPE_Q4_0>(bp, P, ctx);
    }
    const bool f16 = a.type == SPIF_TYPE_F16;
    if (TP == 16) {
        return f16 ? launch_batch_t<spif_f16, 16, 0>(bp, P, ctx) : launch_batch_t<spif_bf16, 16, 0>(bp, P, ctx);
    }
    return f16 ? launch_batch_t<spif_f16, 32, 0>(bp, P, ctx) : launch_batch_t<spif_bf16, 32, 0>(bp, P, ctx);
}
