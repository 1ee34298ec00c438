// spif_batch.cu -- batched (2 <= n_tok <= 32) sparse FFN for F16/BF16 weights on sm_100a tensor cores.
//
// Reference: mul_mat_batch_sparse (mm-sparse.cu:107-210) and mul_mat_axpy_sparse_rowwise with T > 1 (axpy-sparse.cu:16-86)
// launch one CTA per (row, token): every token re-streams every active row -> T x the weight traffic.  Here the UNION of
// the per-token masks becomes a dense gathered GEMM, every union row is read from HBM exactly once per matrix:
//   prep   : bits[row] = OR_t active(row, t) << t                 (spif_batch_bits_kernel, also x F32 -> F16/BF16)
//            union_idx = ascending rows with bits != 0            (spif_batch_union_kernel, ballot + prefix scan)
//   gate/up: CTA c owns a contiguous slice of the union; D[128 rows x T] = W[rows,:] . X^T with tcgen05.mma
//            (A = gathered weight rows, K-major, cp.async into SWIZZLE_128B tiles; B = X; accumulators in TMEM),
//            epilogue applies the per-token mask bits + bias + activation and writes H[t][slot] (weight float type)
//   down   : Y[T x n_embd] = H . Wdown^T[union rows, :] as a stream-K GEMM over (128-column tile, 128-row chunk) items;
//            A = gathered 256-byte row segments of W_down^T (MN-major operand), B = H (K-major); each CTA accumulates a
//            contiguous item range in TMEM and writes <= 2 partial tiles, summed in a fixed order by the last kernel.
// x and the hidden values are rounded to the weight float type (what the reference CPU backend does: ggml-cpu.c:1832-1856,
// 2196-2207); accumulation is FP32 in tensor memory.
#include "spif_batch.cuh"

#define SPIF_B_ROWS_MAX 1024 /* union rows one CTA may own (m <= 1024 * SMs) */

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
    float *    part;         // [P][2][TP][128]
    int        n_embd, n_ff, m, m_pad, mask_len, n_tok;
    int        row_bytes;
    int        arch, gated, by_row;
    float      thr;
};

// ---------------------------------------------------------------------------------------------------------------
// prep 1: per-row token bits; x -> weight float type (rows n_tok..TP-1 are zero)
template <typename WT>
__global__ void __launch_bounds__(256) spif_batch_bits_kernel(const BatchParams p, int nb_bits, int TP) {
    const int tid = threadIdx.x;
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

// prep 2: union list (one CTA; warps own contiguous row chunks, two passes of ballot + popc -> ascending order)
__global__ void __launch_bounds__(1024) spif_batch_union_kernel(const BatchParams p) {
    __shared__ int warp_cnt[32];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    const int      chunk = ((p.m + nwarp * 32 - 1) / (nwarp * 32)) * 32;
    const int      r0 = warp * chunk, r1 = min(r0 + chunk, p.m);
    int            cnt = 0;
    for (int base = r0; base < r1; base += 32) {
        const int  row = base + lane;
        const bool a   = (row < r1) && (p.bits[row] != 0u);
        cnt += __popc(__ballot_sync(0xffffffffu, a));
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
    for (int base = r0; base < r1; base += 32) {
        const int      row  = base + lane;
        const bool     a    = (row < r1) && (p.bits[row] != 0u);
        const unsigned ball = __ballot_sync(0xffffffffu, a);
        if (a) {
            p.union_idx[run + __popc(ball & ((1u << lane) - 1u))] = row;
        }
        run += __popc(ball);
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
// gate/up gathered GEMM.  160 threads: warps 0-3 = cp.async producers, then epilogue; warp 4 = TMEM owner + MMA issuer.
// Stage = [up tile 128 rows x 128 B][gate tile][x tile TP rows x 128 B], K block = 64 elements, SWIZZLE_128B K-major.
#define K1_NST   5
#define K1_LAG   3
#define K1_ATILE 16384
#define K1_STAGE (2 * K1_ATILE + 4096)
#define K1_TAIL  (256 + 4 * SPIF_B_ROWS_MAX)
#define K1_SMEM  (K1_NST * K1_STAGE + K1_TAIL + 1024)

template <typename WT, int TP>
__global__ void __launch_bounds__(160, 1) spif_batch_gateup_kernel(const BatchParams p) {
    extern __shared__ unsigned char smem_raw[];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int      P = gridDim.x, c = blockIdx.x;
    const uint32_t raw0  = smem_u32(smem_raw);
    const uint32_t sbase = (raw0 + 1023u) & ~1023u;
    unsigned char * sm   = smem_raw + (sbase - raw0);
    unsigned char * tail = sm + K1_NST * K1_STAGE;
    uint64_t *      full       = reinterpret_cast<uint64_t *>(tail);
    uint64_t *      empty      = full + K1_NST;
    uint64_t *      accum_full = empty + K1_NST;
    uint64_t *      tmem_empty = accum_full + 1;
    uint32_t *      tmem_slot  = reinterpret_cast<uint32_t *>(tmem_empty + 1);
    int *           s_rows     = reinterpret_cast<int *>(tail + 256);

    const int U   = p.union_count[0];
    const int q   = (U + P - 1) / P;
    const int s0  = c * q;
    const int cnt = max(0, min(U - s0, q));
    if (c == 0 && tid < p.n_tok && p.stats_out) {
        p.stats_out[tid * 4 + 2] = U;
    }
    if (cnt == 0) {
        return;
    }
    for (int i = tid; i < cnt; i += blockDim.x) {
        s_rows[i] = p.union_idx[s0 + i];
    }
    if (tid == 0) {
        for (int s = 0; s < K1_NST; ++s) {
            mbar_init(&full[s], 128);
            mbar_init(&empty[s], 1);
        }
        mbar_init(accum_full, 1);
        mbar_init(tmem_empty, 128);
        mbar_fence_init();
    }
    if (warp == 4) {
        tmem_alloc(tmem_slot, 64);
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    const int  KB      = p.n_embd >> 6;
    const int  n_tiles = (cnt + 127) >> 7;
    const bool gated   = p.gated != 0;

    if (warp < 4) {
        const int cc = tid & 7, rsub = tid >> 3;
        uint32_t  it = 0;
        for (int tile = 0; tile < n_tiles; ++tile) {
            const int cnt_t = min(128, cnt - tile * 128);
            int       rows[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int r = rsub + 16 * j;
                rows[j]     = (r < cnt_t) ? s_rows[tile * 128 + r] : -1;
            }
            for (int kb = 0; kb < KB; ++kb, ++it) {
                const uint32_t st = it % K1_NST;
                if (it >= K1_NST) {
                    mbar_wait(&empty[st], ((it / K1_NST) - 1u) & 1u);
                }
                const uint32_t sA = sbase + st * K1_STAGE;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (rows[j] >= 0) {
                        const uint32_t off = swz_kmajor(rsub + 16 * j, cc);
                        const size_t   go  = (size_t) rows[j] * p.row_bytes + (size_t) kb * 128 + cc * 16;
                        cp_async16(sA + off, p.w_up + go);
                        if (gated) {
                            cp_async16(sA + K1_ATILE + off, p.w_gate + go);
                        }
                    }
                }
#pragma unroll
                for (int h = 0; h < TP * 8 / 128; ++h) {
                    const int idx = tid + 128 * h, t = idx >> 3, c8 = idx & 7;
                    cp_async16(sA + 2 * K1_ATILE + swz_kmajor(t, c8), p.xh + (size_t) t * p.n_embd + kb * 64 + c8 * 8);
                }
                cp_async_commit();
                if (kb >= K1_LAG) {
                    cp_async_wait<K1_LAG>();
                    fence_proxy_async_smem();
                    mbar_arrive(&full[(it - K1_LAG) % K1_NST]);
                }
            }
            cp_async_wait<0>();
            fence_proxy_async_smem();
            for (int d = min(K1_LAG, KB); d >= 1; --d) {
                mbar_arrive(&full[(it - d) % K1_NST]);
            }

            // ---- epilogue of this tile: TMEM -> registers -> mask bits, bias, activation -> H
            mbar_wait(accum_full, tile & 1);
            tc_fence_after_sync();
            float          up[TP], gt[TP];
            const uint32_t taddr = tmem_base + ((uint32_t) (warp * 32) << 16);
#pragma unroll
            for (int h = 0; h < TP / 16; ++h) {
                tmem_ld16(taddr + h * 16, *reinterpret_cast<float(*)[16]>(up + h * 16));
                if (gated) {
                    tmem_ld16(taddr + TP + h * 16, *reinterpret_cast<float(*)[16]>(gt + h * 16));
                }
            }
            tc_fence_before_sync();
            mbar_arrive(tmem_empty);
            const int  r     = tid;
            const bool valid = r < cnt_t;
            const int  row   = valid ? s_rows[tile * 128 + r] : 0;
            const int  neu   = p.neuron_idx ? __ldg(p.neuron_idx + row) : row;
            const uint32_t bits = valid ? p.bits[row] : 0u;
            const float    bu   = (valid && p.b_up) ? __ldg(p.b_up + neu) : 0.f;
            const float    bg   = (valid && gated && p.b_gate) ? __ldg(p.b_gate + neu) : 0.f;
            const int      slot = s0 + tile * 128 + r;
            bool           any_nz = false;
#pragma unroll
            for (int t = 0; t < TP; ++t) {
                float      h   = 0.f;
                const bool act = (t < p.n_tok) && ((bits >> t) & 1u);
                if (act) {
                    h = spif_activation(p.arch, gated ? gt[t] + bg : 0.f, up[t] + bu);
                    if (p.hidden_out) {
                        p.hidden_out[(size_t) t * p.n_ff + neu] = h;
                    }
                }
                const uint16_t hb = f32_to_wt_bits<WT>(h);
                const bool     nz = (hb & 0x7fffu) != 0;
                if (valid) {
                    p.hh[(size_t) t * p.m_pad + slot] = hb;
                }
                any_nz = any_nz || nz;
                if (p.stats_out && t < p.n_tok) {
                    const unsigned ba = __ballot_sync(0xffffffffu, act);
                    const unsigned bz = __ballot_sync(0xffffffffu, act && nz);
                    if (lane == 0) {
                        if (ba) {
                            atomicAdd(p.stats_out + t * 4 + 0, __popc(ba));
                        }
                        if (bz) {
                            atomicAdd(p.stats_out + t * 4 + 1, __popc(bz));
                        }
                    }
                }
            }
            const unsigned bn = __ballot_sync(0xffffffffu, any_nz);
            if (lane == 0 && bn) {
                atomicAdd(p.union_count + 1, __popc(bn));
            }
        }
    } else {
        constexpr uint32_t idesc = umma_idesc_f16(wt_fmt<WT>::id, 0, 0, 128, TP);
        uint32_t           it    = 0;
        for (int tile = 0; tile < n_tiles; ++tile) {
            if (tile > 0) {
                mbar_wait(tmem_empty, (tile - 1) & 1);
                tc_fence_after_sync();
            }
            for (int kb = 0; kb < KB; ++kb, ++it) {
                const uint32_t st = it % K1_NST;
                mbar_wait(&full[st], (it / K1_NST) & 1u);
                tc_fence_after_sync();
                if (lane == 0) {
                    const uint32_t sA = sbase + st * K1_STAGE;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t db  = umma_smem_desc(sA + 2 * K1_ATILE + k * 32, 16, 1024);
                        const uint32_t acc = (kb | k) != 0 ? 1u : 0u;
                        umma_f16(tmem_base, umma_smem_desc(sA + k * 32, 16, 1024), db, idesc, acc);
                        if (gated) {
                            umma_f16(tmem_base + TP, umma_smem_desc(sA + K1_ATILE + k * 32, 16, 1024), db, idesc, acc);
                        }
                    }
                    umma_commit(&empty[st]);
                    if (kb == KB - 1) {
                        umma_commit(accum_full);
                    }
                }
                __syncwarp();
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4) {
        tmem_dealloc(tmem_base, 64);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// down projection: stream-K over items (ct = 128-column tile of n_embd, kc = chunk of 128 union rows), item id = ct * n_kc + kc.
// CTA c owns items [c * ipc, (c + 1) * ipc): at most two column tiles ("segments"), one TMEM accumulator each.
//   A tile  : W_down^T[union rows of the chunk][256-byte column segment], MN-major SWIZZLE_128B: k-row = 128 B of 64 columns,
//             8 k-rows per 1024-byte atom, atoms of one 64-column half contiguous along k (SBO 1024), halves LBO = 16 KB apart
//   B tile  : H[t][128 slots], K-major SWIZZLE_128B, two 64-slot atoms of TP rows
#define K2_NST   5
#define K2_LAG   3
#define K2_ATILE 32768
#define K2_STAGE (K2_ATILE + 8192)
#define K2_SMEM  (K2_NST * K2_STAGE + 256 + 1024)

template <typename WT, int TP>
__global__ void __launch_bounds__(160, 1) spif_batch_down_kernel(const BatchParams p) {
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

    const int U    = p.union_count[0];
    const int n_kc = (U + 127) >> 7;
    const int n_ct = p.n_embd >> 7;
    const int I    = n_ct * n_kc;
    const int ipc  = (I + P - 1) / P;
    const int i0   = c * ipc;
    const int i1   = min(I, i0 + ipc);
    if (i0 >= i1) {
        return;
    }
    const int ct0  = i0 / n_kc;
    const int nseg = (i1 - 1) / n_kc - ct0 + 1;  // 1 or 2 (launcher guarantees n_ct <= P)

    if (tid == 0) {
        for (int s = 0; s < K2_NST; ++s) {
            mbar_init(&full[s], 128);
            mbar_init(&empty[s], 1);
        }
        mbar_init(&accum_full[0], 1);
        mbar_init(&accum_full[1], 1);
        mbar_fence_init();
    }
    if (warp == 4) {
        tmem_alloc(tmem_slot, 64);
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
        const int ch = tid & 15, ksub = tid >> 4;
        uint32_t  it = 0;
        for (int i = i0; i < i1; ++i, ++it) {
            const int      ct = i / n_kc, kc = i - ct * n_kc;
            const uint32_t st = it % K2_NST;
            if (it >= K2_NST) {
                mbar_wait(&empty[st], ((it / K2_NST) - 1u) & 1u);
            }
            const uint32_t sA = sbase + st * K2_STAGE;
#pragma unroll 4
            for (int j = 0; j < 16; ++j) {
                const int      k     = ksub + 8 * j;
                const int      slot  = kc * 128 + k;
                const bool     valid = slot < U;
                const int      row   = valid ? __ldg(p.union_idx + slot) : 0;
                const uint32_t off   = (uint32_t) ((ch >> 3) * 16384 + (k >> 3) * 1024 + (k & 7) * 128 + (((ch & 7) ^ (k & 7)) << 4));
                cp_async16_zfill(sA + off, p.w_down + (size_t) row * p.row_bytes + (size_t) ct * 256 + ch * 16, valid ? 16u : 0u);
            }
#pragma unroll
            for (int h = 0; h < TP * 16 / 128; ++h) {
                const int idx = tid + 128 * h, t = idx >> 4, c16 = idx & 15;
                cp_async16(sA + K2_ATILE + (c16 >> 3) * (TP * 128) + swz_kmajor(t, c16 & 7),
                           p.hh + (size_t) t * p.m_pad + kc * 128 + c16 * 8);
            }
            cp_async_commit();
            if (it >= K2_LAG) {
                cp_async_wait<K2_LAG>();
                fence_proxy_async_smem();
                mbar_arrive(&full[(it - K2_LAG) % K2_NST]);
            }
        }
        cp_async_wait<0>();
        fence_proxy_async_smem();
        for (int d = min((uint32_t) K2_LAG, it); d >= 1; --d) {
            mbar_arrive(&full[(it - d) % K2_NST]);
        }
        // ---- epilogue: one partial tile per segment, part[c][seg][t][col]
        for (int seg = 0; seg < nseg; ++seg) {
            mbar_wait(&accum_full[seg], 0);
            tc_fence_after_sync();
            float          v[TP];
            const uint32_t taddr = tmem_base + ((uint32_t) (warp * 32) << 16) + seg * TP;
#pragma unroll
            for (int h = 0; h < TP / 16; ++h) {
                tmem_ld16(taddr + h * 16, *reinterpret_cast<float(*)[16]>(v + h * 16));
            }
            float * o = p.part + ((size_t) (c * 2 + seg) * TP) * 128 + tid;
#pragma unroll
            for (int t = 0; t < TP; ++t) {
                if (t < p.n_tok) {
                    o[t * 128] = v[t];
                }
            }
        }
    } else {
        constexpr uint32_t idesc = umma_idesc_f16(wt_fmt<WT>::id, 1, 0, 128, TP);
        uint32_t           it    = 0;
        for (int i = i0; i < i1; ++i, ++it) {
            const int      ct = i / n_kc, kc = i - ct * n_kc;
            const int      seg = ct - ct0;
            const uint32_t st  = it % K2_NST;
            const bool     first = (i == i0) || (kc == 0);
            const bool     last  = (i == i1 - 1) || (kc == n_kc - 1);
            mbar_wait(&full[st], (it / K2_NST) & 1u);
            tc_fence_after_sync();
            if (lane == 0) {
                const uint32_t sA = sbase + st * K2_STAGE;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint64_t da = umma_smem_desc(sA + k * 2048, 16384, 1024);
                    const uint64_t db = umma_smem_desc(sA + K2_ATILE + (k >> 2) * (TP * 128) + (k & 3) * 32, 16, 1024);
                    umma_f16(tmem_base + seg * TP, da, db, idesc, (first && k == 0) ? 0u : 1u);
                }
                umma_commit(&empty[st]);
                if (last) {
                    umma_commit(&accum_full[seg]);
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4) {
        tmem_dealloc(tmem_base, 64);
    }
}

// y[t][col] = sum over the CTAs whose item range touched column tile ct, ascending CTA order (fixed) (+ b_down)
__global__ void __launch_bounds__(128) spif_batch_reduce_kernel(const BatchParams p, int P, int TP) {
    const int ct = blockIdx.x, t = blockIdx.y, col = threadIdx.x;
    const int U    = p.union_count[0];
    const int n_kc = (U + 127) >> 7;
    const int n_ct = p.n_embd >> 7;
    const int I    = n_ct * n_kc;
    float     s    = 0.f;
    if (I > 0) {
        const int ipc = (I + P - 1) / P;
        const int lo = ct * n_kc, hi = lo + n_kc - 1;
        for (int c = lo / ipc; c <= hi / ipc; ++c) {
            const int seg = ct - (c * ipc) / n_kc;
            s += p.part[((size_t) (c * 2 + seg) * TP + t) * 128 + col];
        }
    }
    if (p.b_down) {
        s += __ldg(p.b_down + ct * 128 + col);
    }
    p.y[(size_t) t * p.n_embd + ct * 128 + col] = s;
    if (p.stats_out && ct == 0 && col == 0) {
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
    return al256(4 * (size_t) n_ff) * 2 + 256 + al256(TP * (size_t) n_embd * 2) + al256(TP * m_pad * 2) + al256(P * 2 * TP * 128 * 4);
}

template <typename WT, int TP>
static int launch_batch_t(const BatchParams & bp, int P, const spif_launch_ctx & ctx) {
    static unsigned long long done1 = 0, done2 = 0;
    auto k1 = spif_batch_gateup_kernel<WT, TP>;
    auto k2 = spif_batch_down_kernel<WT, TP>;
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
    spif_batch_bits_kernel<WT><<<nb_bits + nb_cvt, 256, 0, ctx.stream>>>(bp, nb_bits, TP);
    spif_batch_union_kernel<<<1, 1024, 0, ctx.stream>>>(bp);
    k1<<<P, 160, K1_SMEM, ctx.stream>>>(bp);
    k2<<<P, 160, K2_SMEM, ctx.stream>>>(bp);
    spif_batch_reduce_kernel<<<dim3(bp.n_embd >> 7, bp.n_tok), 128, 0, ctx.stream>>>(bp, P, TP);
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}

// returns SPIF_E_UNSUPPORTED when the shape does not qualify: the caller then uses the per-token kernel
int spif_launch_ffn_batch(const spif_ffn_args & a, void * ws, size_t ws_bytes, const spif_launch_ctx & ctx) {
    SPIF_REQUIRE(a.type == SPIF_TYPE_F16 || a.type == SPIF_TYPE_BF16, SPIF_E_UNSUPPORTED);
    SPIF_REQUIRE(a.n_tok >= 2 && a.n_tok <= 32, SPIF_E_UNSUPPORTED);
    const int P = ctx.sm_count;
    SPIF_REQUIRE(a.n_embd % 128 == 0 && (a.n_embd >> 7) <= P && a.m <= SPIF_B_ROWS_MAX * P, SPIF_E_UNSUPPORTED);
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
    bp.row_bytes  = a.n_embd * 2;
    bp.arch       = a.arch;
    bp.gated      = (a.arch != SPIF_ARCH_RELU_UP) ? 1 : 0;
    bp.thr        = a.thr;

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
    const size_t hh_bytes = (size_t) TP * m_pad * 2;
    o += al256(hh_bytes);
    bp.part = (float *) (w + o);
    o += al256((size_t) P * 2 * TP * 128 * 4);
    SPIF_REQUIRE(ws_bytes >= o, SPIF_E_WORKSPACE);

    SPIF_CUDA_TRY(cudaMemsetAsync(bp.hh, 0, hh_bytes, ctx.stream));
    if (a.hidden_out) {
        SPIF_CUDA_TRY(cudaMemsetAsync(a.hidden_out, 0, sizeof(float) * (size_t) a.n_ff * a.n_tok, ctx.stream));
    }
    const bool f16 = a.type == SPIF_TYPE_F16;
    if (TP == 16) {
        return f16 ? launch_batch_t<spif_f16, 16>(bp, P, ctx) : launch_batch_t<spif_bf16, 16>(bp, P, ctx);
    }
    return f16 ? launch_batch_t<spif_f16, 32>(bp, P, ctx) : launch_batch_t<spif_bf16, 32>(bp, P, ctx);
}
