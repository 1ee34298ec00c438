// spif_mask.cu -- mask compaction (warp ballot + prefix scan), top-k mask, the two SparkInfer unary ops and the fused
// DFR (decayed firing rate) update used by the balancer.
//
// Reference: the mask is the float tensor `sparse_idx`; no index list is ever built (mm-sparse.cu:22, axpy-sparse.cu:53).
// fatrelu / shifted_step: ggml/src/ggml-cuda/unary.cu:564-652.  build_dfr: src/llama-graph.cpp:910-930 with
// op_scale_add(_ema) ggml/src/ggml-cuda/binbcast.cu:26-42.
#include <cooperative_groups.h>

#include "spif_common.cuh"

// ---------------------------------------------------------------------------------------------------------------
// One CTA per token (+ one for the union over tokens).  Warps own contiguous row chunks; two passes (count, then
// rank = warp base + running ballot popcount), output ascending == the order the reference CPU loops visit rows
// (ggml-cpu.c:1773-1779, 2195-2212).
__global__ void __launch_bounds__(1024) spif_mask_compact_kernel(const float * __restrict__ sparse_idx,
                                                                  const int32_t * __restrict__ neuron_idx, int m, int n_ff,
                                                                  int n_tok, float thr, int by_row,
                                                                  int32_t * __restrict__ idx_out,
                                                                  int32_t * __restrict__ count_out,
                                                                  int32_t * __restrict__ union_idx,
                                                                  int32_t * __restrict__ union_count) {
    __shared__ int warp_cnt[32];
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    const bool     is_union = (blockIdx.x == n_tok);
    const int      mask_len = by_row ? m : n_ff;
    const int      chunk    = ((m + nwarp * 32 - 1) / (nwarp * 32)) * 32;
    const int      r0 = warp * chunk, r1 = min(r0 + chunk, m);

    auto test = [&](int row) -> bool {
        const int neu = neuron_idx ? __ldg(neuron_idx + row) : row;
        const int mi  = by_row ? row : neu;
        if (!is_union) {
            return spif_is_active(__ldg(sparse_idx + (size_t) blockIdx.x * mask_len + mi), thr);
        }
        bool a = false;
        for (int t = 0; t < n_tok; ++t) {
            a = a || spif_is_active(__ldg(sparse_idx + (size_t) t * mask_len + mi), thr);
        }
        return a;
    };

    int cnt = 0;
    for (int base = r0; base < r1; base += 32) {
        const int  row = base + lane;
        const bool a   = (row < r1) && test(row);
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
    int32_t * out = is_union ? union_idx : idx_out + (size_t) blockIdx.x * m;
    int       run = before;
    for (int base = r0; base < r1; base += 32) {
        const int      row  = base + lane;
        const bool     a    = (row < r1) && test(row);
        const unsigned ball = __ballot_sync(0xffffffffu, a);
        if (a) {
            out[run + __popc(ball & ((1u << lane) - 1u))] = row;
        }
        run += __popc(ball);
    }
    if (tid == 0) {
        if (is_union) {
            *union_count = total;
        } else {
            count_out[blockIdx.x] = total;
        }
    }
}

int spif_launch_mask_compact(const float * sparse_idx, const int32_t * neuron_idx, int m, int n_ff, int n_tok, float thr,
                             int flags, int32_t * idx_out, int32_t * count_out, int32_t * union_idx, int32_t * union_count,
                             cudaStream_t stream) {
    const bool want_union = union_idx != nullptr && union_count != nullptr;
    spif_mask_compact_kernel<<<n_tok + (want_union ? 1 : 0), 1024, 0, stream>>>(
        sparse_idx, neuron_idx, m, n_ff, n_tok, thr, (flags & SPIF_FLAG_MASK_BY_ROW) ? 1 : 0, idx_out, count_out,
        union_idx, union_count);
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// top-k mask: one CTA per token.  4-pass MSB radix select on order-preserving uint keys finds the k-th largest key;
// everything above it is selected, ties at the threshold are taken in ascending index order (ballot prefix scan).
__device__ __forceinline__ uint32_t float_key(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // larger float -> larger key (NaN sorts by bit pattern)
}

__global__ void __launch_bounds__(1024) spif_topk_mask_kernel(const float * __restrict__ score, int n_ff, int k,
                                                               float * __restrict__ out) {
    __shared__ int      hist[256];
    __shared__ uint32_t s_prefix, s_prefix_mask;
    __shared__ int      s_k;
    __shared__ int      warp_cnt[32];
    const float *       sc = score + (size_t) blockIdx.x * n_ff;
    float *             o  = out + (size_t) blockIdx.x * n_ff;
    const int           tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;

    if (k <= 0) {
        for (int i = tid; i < n_ff; i += blockDim.x) {
            o[i] = 0.f;
        }
        return;
    }
    if (k >= n_ff) {
        for (int i = tid; i < n_ff; i += blockDim.x) {
            o[i] = 1.f;
        }
        return;
    }
    if (tid == 0) {
        s_prefix      = 0;
        s_prefix_mask = 0;
        s_k           = k;
    }
    for (int pass = 3; pass >= 0; --pass) {
        for (int i = tid; i < 256; i += blockDim.x) {
            hist[i] = 0;
        }
        __syncthreads();
        const uint32_t pre = s_prefix, pmask = s_prefix_mask;
        const int      sh  = pass * 8;
        // histogram of the current digit among the keys that still match the prefix.  The top digit (sign + exponent) has
        // a handful of distinct values, so lanes with the same digit are merged first (match_any) and ONE lane adds the
        // group's size: without it this pass serialises ~n_ff shared-memory atomics on 3-4 addresses.
        for (int base = 0; base < n_ff; base += blockDim.x) {
            const int      i   = base + tid;
            const uint32_t key = (i < n_ff) ? float_key(sc[i]) : 0u;
            const bool     in  = (i < n_ff) && ((key & pmask) == pre);
            const unsigned dig = in ? ((key >> sh) & 255u) : (0x1000u + lane);
            if (pass == 3) {
                const unsigned grp = __match_any_sync(0xffffffffu, dig);
                if (in && (__ffs(grp) - 1) == lane) {
                    atomicAdd(&hist[dig], __popc(grp));
                }
            } else if (in) {  // lower digits are well spread and few keys still match the prefix
                atomicAdd(&hist[dig], 1);
            }
        }
        __syncthreads();
        if (warp == 0) {
            // lane l owns digits 8l .. 8l+7; suffix sums over the lanes above, then the lane holding the k-th key scans its 8
            int own[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                own[j] = hist[lane * 8 + j];
                sum += own[j];
            }
            int above = sum;  // inclusive suffix sum over lanes >= lane
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_down_sync(0xffffffffu, above, o);
                above += (lane + o < 32) ? t : 0;
            }
            above -= sum;  // keys in strictly higher lanes
            const int need = s_k;
            __syncwarp();  // every lane has read s_k before the owning lane replaces it
            if (above < need && need <= above + sum) {
                int rem = need - above, bdig = lane * 8 + 7;
#pragma unroll
                for (int j = 7; j >= 0; --j) {
                    if (rem > 0 && own[j] >= rem) {
                        bdig = lane * 8 + j;
                        s_k  = rem;  // how many to take among keys with this prefix digit
                        rem  = -1;
                    } else if (rem > 0) {
                        rem -= own[j];
                    }
                }
                s_prefix      = pre | ((uint32_t) bdig << sh);
                s_prefix_mask = pmask | (255u << sh);
            }
        }
        __syncthreads();
    }
    const uint32_t kth      = s_prefix;  // key of the k-th largest
    const int      need_eq  = s_k;       // how many keys == kth to take (lowest indices first)
    // ordered pass over indices: warps own contiguous chunks
    const int chunk = ((n_ff + nwarp * 32 - 1) / (nwarp * 32)) * 32;
    const int r0 = warp * chunk, r1 = min(r0 + chunk, n_ff);
    int       cnt = 0;
    for (int base = r0; base < r1; base += 32) {
        const int  i  = base + lane;
        const bool eq = (i < r1) && float_key(sc[i]) == kth;
        cnt += __popc(__ballot_sync(0xffffffffu, eq));
    }
    if (lane == 0) {
        warp_cnt[warp] = cnt;
    }
    __syncthreads();
    int run = 0;
    for (int w = 0; w < warp; ++w) {
        run += warp_cnt[w];
    }
    for (int base = r0; base < r1; base += 32) {
        const int      i    = base + lane;
        const uint32_t key  = (i < r1) ? float_key(sc[i]) : 0u;
        const bool     eq   = (i < r1) && key == kth;
        const unsigned ball = __ballot_sync(0xffffffffu, eq);
        if (i < r1) {
            bool sel = key > kth;
            if (eq) {
                sel = (run + __popc(ball & ((1u << lane) - 1u))) < need_eq;
            }
            o[i] = sel ? 1.0f : 0.0f;
        }
        run += __popc(ball);
    }
}

// Register-resident variant for n_ff <= KPT * 1024: every thread loads its KPT keys ONCE (warp w owns the contiguous index
// chunk [w * chunk, (w+1) * chunk), key j of a lane is index chunk_base + 32 j + lane -> coalesced, and already in the order
// the tie-break needs), the four radix passes and the ordered output pass then run on registers only.
template <int KPT>
__global__ void __launch_bounds__(1024) spif_topk_mask_reg_kernel(const float * __restrict__ score, int n_ff, int k,
                                                                   float * __restrict__ out) {
    __shared__ int      hist[256];
    __shared__ uint32_t s_prefix, s_prefix_mask;
    __shared__ int      s_k;
    __shared__ int      warp_cnt[32];
    const float *       sc = score + (size_t) blockIdx.x * n_ff;
    float *             o  = out + (size_t) blockIdx.x * n_ff;
    const int           tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int           chunk = ((n_ff + 1023) / 1024) * 32;
    const int           r0 = warp * chunk, r1 = min(r0 + chunk, n_ff);
    if (k <= 0 || k >= n_ff) {
        for (int i = tid; i < n_ff; i += blockDim.x) {
            o[i] = (k <= 0) ? 0.f : 1.f;
        }
        return;
    }
    uint32_t key[KPT];
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        const int i = r0 + j * 32 + lane;
        key[j]      = (i < r1) ? float_key(sc[i]) : 0u;
    }
    if (tid == 0) {
        s_prefix      = 0;
        s_prefix_mask = 0;
        s_k           = k;
    }
    for (int pass = 3; pass >= 0; --pass) {
        for (int i = tid; i < 256; i += blockDim.x) {
            hist[i] = 0;
        }
        __syncthreads();
        const uint32_t pre = s_prefix, pmask = s_prefix_mask;
        const int      sh  = pass * 8;
#pragma unroll
        for (int j = 0; j < KPT; ++j) {
            if (r0 + j * 32 >= r1) {
                continue;  // warp-uniform: nothing of this warp's chunk in this slot
            }
            const bool     in  = (r0 + j * 32 + lane < r1) && ((key[j] & pmask) == pre);
            const unsigned dig = in ? ((key[j] >> sh) & 255u) : 0x1000u;
            if (pass == 3) {  // sign + exponent: a handful of distinct values -> merge equal lanes, one atomic per group
                const unsigned grp = __match_any_sync(0xffffffffu, dig);
                if (in && (__ffs(grp) - 1) == lane) {
                    atomicAdd(&hist[dig], __popc(grp));
                }
            } else if (in) {
                atomicAdd(&hist[dig], 1);
            }
        }
        __syncthreads();
        if (warp == 0) {
            int own[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                own[j] = hist[lane * 8 + j];
                sum += own[j];
            }
            int above = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_down_sync(0xffffffffu, above, d);
                above += (lane + d < 32) ? t : 0;
            }
            above -= sum;
            const int need = s_k;
            __syncwarp();
            if (above < need && need <= above + sum) {
                int rem = need - above, bdig = lane * 8 + 7;
#pragma unroll
                for (int j = 7; j >= 0; --j) {
                    if (rem > 0 && own[j] >= rem) {
                        bdig = lane * 8 + j;
                        s_k  = rem;
                        rem  = -1;
                    } else if (rem > 0) {
                        rem -= own[j];
                    }
                }
                s_prefix      = pre | ((uint32_t) bdig << sh);
                s_prefix_mask = pmask | (255u << sh);
            }
        }
        __syncthreads();
    }
    const uint32_t kth     = s_prefix;
    const int      need_eq = s_k;
    int            cnt = 0;
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        cnt += __popc(__ballot_sync(0xffffffffu, (r0 + j * 32 + lane < r1) && key[j] == kth));
    }
    if (lane == 0) {
        warp_cnt[warp] = cnt;
    }
    __syncthreads();
    int run = 0;
    for (int w = 0; w < warp; ++w) {
        run += warp_cnt[w];
    }
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        const int      i  = r0 + j * 32 + lane;
        const unsigned bj = __ballot_sync(0xffffffffu, (i < r1) && key[j] == kth);
        if (i < r1) {
            bool sel = key[j] > kth;
            if (key[j] == kth) {
                sel = (run + __popc(bj & ((1u << lane) - 1u))) < need_eq;
            }
            o[i] = sel ? 1.0f : 0.0f;
        }
        run += __popc(bj);
    }
}

// Cluster variant (default for n_ff <= 16384 since round 2: parity-tested on hardware, Llama-3-8B top-k 775 -> 855 tok/s;
// SPIF_TOPK_CLUSTER=0 selects the single-CTA kernels): one cluster of TK_CL CTAs per token.  Every CTA keeps 1/TK_CL of the keys in registers and a private histogram; after one cluster
// barrier per pass each CTA sums the TK_CL histograms through distributed shared memory and takes the SAME decision, so
// no result has to be broadcast.  Index ranges ascend with (cluster rank, warp, slot, lane): the tie-break order is the
// same as in the single-CTA kernels.  Meant for the Llama-3-8B top-k chain where the 1-CTA kernel (16 us) is the longest
// link (open items).
#define TK_CL      8
#define TK_THREADS 256
template <int KPT>
__global__ void __launch_bounds__(TK_THREADS) spif_topk_mask_cluster_kernel(const float * __restrict__ score, int n_ff, int k,
                                                                            float * __restrict__ out) {
    namespace cg = cooperative_groups;
    cg::cluster_group   cluster = cg::this_cluster();
    __shared__ int      hist[2][256];  // double-buffered: a peer may still read pass p while pass p+1 is being zeroed
    __shared__ int      ghist[256];
    __shared__ uint32_t s_prefix, s_prefix_mask;
    __shared__ int      s_k;
    __shared__ int      warp_cnt[TK_THREADS / 32];
    __shared__ int      s_cta_eq;
    const float *       sc = score + (size_t) blockIdx.y * n_ff;
    float *             o  = out + (size_t) blockIdx.y * n_ff;
    const int           tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int           rank   = (int) cluster.block_rank();
    const int           nwarp  = TK_THREADS / 32;
    const int           wchunk = ((n_ff + TK_CL * nwarp - 1) / (TK_CL * nwarp) + 31) / 32 * 32;  // <= 32 * KPT (launcher)
    const int           r0 = min((rank * nwarp + warp) * wchunk, n_ff), r1 = min(r0 + wchunk, n_ff);
    if (k <= 0 || k >= n_ff) {  // uniform over the cluster: nobody reaches a barrier
        for (int i = r0 + lane; i < r1; i += 32) {
            o[i] = (k <= 0) ? 0.f : 1.f;
        }
        return;
    }
    uint32_t key[KPT];
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        const int i = r0 + j * 32 + lane;
        key[j]      = (i < r1) ? float_key(sc[i]) : 0u;
    }
    if (tid == 0) {
        s_prefix      = 0;
        s_prefix_mask = 0;
        s_k           = k;
    }
    for (int pass = 3; pass >= 0; --pass) {
        int * h = hist[pass & 1];
        for (int i = tid; i < 256; i += TK_THREADS) {
            h[i] = 0;
        }
        __syncthreads();
        const uint32_t pre = s_prefix, pmask = s_prefix_mask;
        const int      sh  = pass * 8;
#pragma unroll
        for (int j = 0; j < KPT; ++j) {
            if (r0 + j * 32 >= r1) {
                continue;  // warp-uniform
            }
            const bool     in  = (r0 + j * 32 + lane < r1) && ((key[j] & pmask) == pre);
            const unsigned dig = in ? ((key[j] >> sh) & 255u) : 0x1000u;
            if (pass == 3) {
                const unsigned grp = __match_any_sync(0xffffffffu, dig);
                if (in && (__ffs(grp) - 1) == lane) {
                    atomicAdd(&h[dig], __popc(grp));
                }
            } else if (in) {
                atomicAdd(&h[dig], 1);
            }
        }
        cluster.sync();  // every CTA's histogram of this pass is complete (and visible cluster-wide)
        {
            int sum = 0;
#pragma unroll
            for (int r = 0; r < TK_CL; ++r) {
                sum += cluster.map_shared_rank(h, r)[tid];  // TK_THREADS == 256 bins
            }
            ghist[tid] = sum;
        }
        __syncthreads();
        if (warp == 0) {
            int own[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                own[j] = ghist[lane * 8 + j];
                sum += own[j];
            }
            int above = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_down_sync(0xffffffffu, above, d);
                above += (lane + d < 32) ? t : 0;
            }
            above -= sum;
            const int need = s_k;
            __syncwarp();
            if (above < need && need <= above + sum) {
                int rem = need - above, bdig = lane * 8 + 7;
#pragma unroll
                for (int j = 7; j >= 0; --j) {
                    if (rem > 0 && own[j] >= rem) {
                        bdig = lane * 8 + j;
                        s_k  = rem;
                        rem  = -1;
                    } else if (rem > 0) {
                        rem -= own[j];
                    }
                }
                s_prefix      = pre | ((uint32_t) bdig << sh);
                s_prefix_mask = pmask | (255u << sh);
            }
        }
        __syncthreads();
    }
    const uint32_t kth     = s_prefix;
    const int      need_eq = s_k;
    int            cnt = 0;
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        cnt += __popc(__ballot_sync(0xffffffffu, (r0 + j * 32 + lane < r1) && key[j] == kth));
    }
    if (lane == 0) {
        warp_cnt[warp] = cnt;
    }
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int w = 0; w < nwarp; ++w) {
            t += warp_cnt[w];
        }
        s_cta_eq = t;
    }
    cluster.sync();
    int run = 0;
    for (int r = 0; r < rank; ++r) {
        run += *cluster.map_shared_rank(&s_cta_eq, r);
    }
    for (int w = 0; w < warp; ++w) {
        run += warp_cnt[w];
    }
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
        const int      i  = r0 + j * 32 + lane;
        const unsigned bj = __ballot_sync(0xffffffffu, (i < r1) && key[j] == kth);
        if (i < r1) {
            bool sel = key[j] > kth;
            if (key[j] == kth) {
                sel = (run + __popc(bj & ((1u << lane) - 1u))) < need_eq;
            }
            o[i] = sel ? 1.0f : 0.0f;
        }
        run += __popc(bj);
    }
    cluster.sync();  // no CTA may exit while a peer can still read its shared memory
}

static int launch_topk_cluster(const float * score, int n_ff, int n_tok, int k, float * out, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(TK_CL, n_tok);
    cfg.blockDim           = dim3(TK_THREADS);
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id               = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = TK_CL;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs              = at;
    cfg.numAttrs           = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_topk_mask_cluster_kernel<8>, score, n_ff, k, out));
    return 0;
}

int spif_launch_topk_mask(const float * score, int n_ff, int n_tok, int k, float * out, cudaStream_t stream) {
    // 8-CTA cluster kernel by default; SPIF_TOPK_CLUSTER=0 keeps one CTA per token (read per call: the tests cover both)
    const char * e_cl = getenv("SPIF_TOPK_CLUSTER");
    if ((!e_cl || atoi(e_cl) != 0) && n_ff <= TK_CL * TK_THREADS * 8) {
        return launch_topk_cluster(score, n_ff, n_tok, k, out, stream);
    }
    if (n_ff <= 16 * 1024) {
        spif_topk_mask_reg_kernel<16><<<n_tok, 1024, 0, stream>>>(score, n_ff, k, out);
    } else if (n_ff <= 32 * 1024) {
        spif_topk_mask_reg_kernel<32><<<n_tok, 1024, 0, stream>>>(score, n_ff, k, out);
    } else {
        spif_topk_mask_kernel<<<n_tok, 1024, 0, stream>>>(score, n_ff, k, out);
    }
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// GGML_OP_FATRELU: x > t ? x : 0 ; GGML_OP_SHIFTED_STEP: (x + t) > 0 ? 1 : 0     (unary.cu:571,616)
template <int OP> __global__ void spif_unary_kernel(const float * __restrict__ x, float * __restrict__ y, int64_t n, float t) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const float v = x[i];
        y[i]          = (OP == 0) ? ((v > t) ? v : 0.0f) : (((v + t) > 0.0f) ? 1.0f : 0.0f);
    }
}

int spif_launch_unary(int op, const float * x, float * y, int64_t n, float t, cudaStream_t stream) {
    if (n <= 0) {
        return 0;
    }
    const int     threads = 256;
    const int64_t blocks  = (n + threads - 1) / threads;
    if (op == 0) {
        spif_unary_kernel<0><<<(unsigned) blocks, threads, 0, stream>>>(x, y, n, t);
    } else {
        spif_unary_kernel<1><<<(unsigned) blocks, threads, 0, stream>>>(x, y, n, t);
    }
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// DFR update: one warp per neuron group.  delta = number of (neuron in group, token) pairs that are active.
__global__ void spif_dfr_update_kernel(const float * __restrict__ sparse_idx, int n_ff, int n_tok, int g, float thr,
                                       float lambda, int ema, float * __restrict__ dfr) {
    const int grp  = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int n_g  = n_ff / g;
    if (grp >= n_g) {
        return;
    }
    int cnt = 0;
    for (int e = lane; e < g * n_tok; e += 32) {
        const int t = e / g, j = e - t * g;
        // build_dfr tests with ggml_shifted_step(sparse_idx, -0.5): (p + (-thr)) > 0, i.e. strictly above the threshold
        // (llama-graph.cpp:911, unary.cu:611-613) -- not the kernels' !(p < thr)
        cnt += ((sparse_idx[(size_t) t * n_ff + (size_t) grp * g + j] + (-thr)) > 0.0f) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if (lane == 0) {
        const float norm  = (float) n_tok * (float) g;
        const float delta = (float) cnt / norm;
        const float old   = dfr[grp];
        dfr[grp]          = ema ? (lambda * old + (1.0f - lambda) * delta) : (lambda * old + delta);
    }
}

int spif_launch_dfr_update(const float * sparse_idx, int n_ff, int n_tok, int g, float thr, float lambda, int ema,
                           float * dfr, cudaStream_t stream) {
    const int n_g     = n_ff / g;
    const int threads = 256;
    const int blocks  = (n_g * 32 + threads - 1) / threads;
    spif_dfr_update_kernel<<<blocks, threads, 0, stream>>>(sparse_idx, n_ff, n_tok, g, thr, lambda, ema, dfr);
    SPIF_CUDA_TRY(cudaGetLastError());
    return 0;
}
