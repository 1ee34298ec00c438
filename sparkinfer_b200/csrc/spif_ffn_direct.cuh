// spif_ffn_direct.cuh -- register-streaming variant of the fused sparse-FFN layer (F16/BF16, n_tok == 1 per CTA row).
//
// Same work split as spif_ffn_kernel (every CTA scans the mask, owns the active rows rank % P == cta, one partial y per
// CTA, fixed-order sum afterwards) but the weight rows never touch shared memory: thread v owns 16-byte vector v of every
// row, x lives in registers, and the rows of the next batches are already in flight as plain 128-bit global loads
// (ld.global.nc.L1::no_allocate) while the current batch is reduced.  Why: profiles/r1_ubench_stream_rows.txt -- for one
// layer's worth of rows per launch (30 rows per SM) register-direct loads reach 4.9 TB/s where the TMA / cp.async rings
// stop at 3.9-4.0 TB/s, and in steady state 6.5 TB/s against 5.7 TB/s: with 512-640 threads x 10 loads the SM keeps
// ~100 KB in flight with no ring-slot granularity, no producer warp and no mbarrier round trips.
//
// Software pipeline per CTA (NB neurons per batch, all indices static after unrolling by 2):
//   iteration b:  dot(batch b)                     -- registers GU[b & 1], loaded two iterations ago
//                 issue gate/up loads of batch b+2  -- into the registers just freed
//                 cross-warp reduction, activation  -- hidden values of batch b
//                 axpy(down rows of batch b-2)      -- registers DN[b & 1], loaded two iterations ago
//                 issue down loads of batch b       -- only for hidden != 0 (axpy-sparse.cu:53)
#pragma once
#include "spif_ffn_kernel.cuh"

template <typename WT, int VPT, int NB, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) spif_ffn_direct_kernel(const FfnKParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarp = blockDim.x >> 5;
    const int P = gridDim.x, c = blockIdx.x, tok = blockIdx.y;
    const int NT = blockDim.x;

    float * red        = reinterpret_cast<float *>(smem);                 // [2][NB * 2][32]
    int *   warp_cnt   = reinterpret_cast<int *>(red + 2 * NB * 2 * 32);  // [32]
    int *   s_counts   = warp_cnt + 32;                                   // [4]
    int *   list_row   = s_counts + 4;
    int *   list_neu   = list_row + p.list_cap;
    float * list_alpha = reinterpret_cast<float *>(list_neu + p.list_cap);

    const bool gated = (p.arch != SPIF_ARCH_RELU_UP);
    const bool early = (p.flags & SPIF_FLAG_MASK_STABLE) != 0;
    SPIF_STAMP(0);
    if (!early) {
        pdl_wait_prior_grid();
    }
    SPIF_STAMP(1);
    const float * mask  = p.mask + (size_t) tok * p.mask_len;
    const float * xg    = p.x + (size_t) tok * p.n_embd;
    float *       dense = p.dense_out ? p.dense_out + (size_t) tok * p.n_ff : nullptr;
    build_my_list<SPIF_MODE_FUSED, FloatCodec<WT, VPT>>(p, mask, nullptr, dense, warp_cnt, list_row, list_neu, list_alpha,
                                                        s_counts);
    const int n  = s_counts[0];
    const int nb = (n + NB - 1) / NB;
    SPIF_STAMP(2);
    const int V  = p.n_embd >> 3;

    // this thread's vectors: v = tid + k * NT
    uint4 GU[2][NB][2][VPT];
    uint4 DN[2][NB][VPT];
    float alpha[2][NB];
    auto  load_gu = [&](int b, uint4 (&dst)[NB][2][VPT]) {
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const int    i   = b * NB + j;
            const bool   ok  = i < n;
            const size_t off = ok ? (size_t) list_row[i] * p.row_bytes : 0;
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                const int v = tid + k * NT;
                dst[j][0][k] = make_uint4(0u, 0u, 0u, 0u);
                dst[j][1][k] = make_uint4(0u, 0u, 0u, 0u);
                if (ok && v < V) {
                    dst[j][0][k] = ldg_stream_v4(p.w_up + off + (size_t) v * 16);
                    if (gated) {
                        dst[j][1][k] = ldg_stream_v4(p.w_gate + off + (size_t) v * 16);
                    }
                }
            }
        }
    };
    load_gu(0, GU[0]);
    load_gu(1, GU[1]);
    if (early) {
        pdl_wait_prior_grid();  // x, y and the partials workspace belong to the previous kernels until here
    }
    const bool pdl_late = (p.flags & SPIF_KFLAG_PDL_LATE) != 0;
    if (!pdl_late) {
        pdl_launch_dependents();
    }

    float xr[VPT * 8], acc[VPT * 8];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = tid + k * NT;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            xr[k * 8 + e] = 0.f, acc[k * 8 + e] = 0.f;
        }
        if (v < V) {
            const float4 a = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v);
            const float4 b = __ldg(reinterpret_cast<const float4 *>(xg) + 2 * v + 1);
            xr[k * 8 + 0] = a.x, xr[k * 8 + 1] = a.y, xr[k * 8 + 2] = a.z, xr[k * 8 + 3] = a.w;
            xr[k * 8 + 4] = b.x, xr[k * 8 + 5] = b.y, xr[k * 8 + 6] = b.z, xr[k * 8 + 7] = b.w;
            if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    xr[k * 8 + e] = round_through<WT>(xr[k * 8 + e]);
                }
            }
        }
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            alpha[s][j] = 0.f;
        }
    }
    int n_nonzero = 0;

    auto axpy_prev = [&](const uint4 (&dn)[NB][VPT], const float (&al)[NB]) {
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            if (al[j] != 0.0f) {  // uniform over the CTA
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    float w[8];
                    unpack8<WT>(dn[j][k], w);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        acc[k * 8 + e] = fmaf(w[e], al[j], acc[k * 8 + e]);
                    }
                }
            }
        }
    };

    auto iteration = [&](int b, uint4 (&gu)[NB][2][VPT], uint4 (&dn)[NB][VPT], float (&al)[NB], float * rb) {
        // ---- partial dots of batch b
        float pu[NB], pg[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            pu[j] = 0.f, pg[j] = 0.f;
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                float w[8];
                unpack8<WT>(gu[j][0][k], w);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    pu[j] = fmaf(w[e], xr[k * 8 + e], pu[j]);
                }
                if (gated) {
                    unpack8<WT>(gu[j][1][k], w);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        pg[j] = fmaf(w[e], xr[k * 8 + e], pg[j]);
                    }
                }
            }
        }
        load_gu(b + 2, gu);  // the registers are free again: rows of batch b+2 go in flight now
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            pu[j] = warp_sum(pu[j]);
            pg[j] = warp_sum(pg[j]);
        }
        if (lane == 0) {
#pragma unroll
            for (int j = 0; j < NB; ++j) {
                rb[(j * 2 + 0) * 32 + warp] = pu[j];
                rb[(j * 2 + 1) * 32 + warp] = pg[j];
            }
        }
        __syncthreads();
        // every warp re-sums the per-warp partials with the same fixed shuffle tree
        float tot[NB * 2];
#pragma unroll
        for (int q = 0; q < NB * 2; ++q) {
            tot[q] = warp_sum(lane < nwarp ? rb[q * 32 + lane] : 0.f);
        }
        // ---- down rows issued two iterations ago
        axpy_prev(dn, al);
        // ---- hidden values of batch b, then their down rows go in flight
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const int i = b * NB + j;
            al[j]       = 0.f;
            if (i < n) {
                const int neu = list_neu[i];
                float     u = tot[j * 2], g = tot[j * 2 + 1];
                if (p.b_up) {
                    u += __ldg(p.b_up + neu);
                }
                if (gated && p.b_gate) {
                    g += __ldg(p.b_gate + neu);
                }
                float h = spif_activation(p.arch, g, u);
                if (tid == 0 && dense) {
                    dense[neu] = h;
                }
                if (p.flags & SPIF_FLAG_CPU_ROUNDING) {
                    h = round_through<WT>(h);
                }
                al[j] = h;
                if (!(h == 0.0f)) {
                    ++n_nonzero;
                    const size_t off = (size_t) list_row[i] * p.row_bytes;
#pragma unroll
                    for (int k = 0; k < VPT; ++k) {
                        const int v = tid + k * NT;
                        dn[j][k]    = (v < V) ? ldg_stream_v4(p.w_down + off + (size_t) v * 16) : make_uint4(0u, 0u, 0u, 0u);
                    }
                }
            }
        }
    };

    for (int b = 0; b < nb; b += 2) {
        iteration(b, GU[0], DN[0], alpha[0], red);
        if (b == 0) {
            SPIF_STAMP(3);
        }
        if (b + 1 < nb) {
            iteration(b + 1, GU[1], DN[1], alpha[1], red + NB * 2 * 32);
        }
    }
    SPIF_STAMP(4);
    if (pdl_late) {
        pdl_launch_dependents();  // the partial-sum CTAs take their registers only now (room for the predictor until here)
    }
    axpy_prev(DN[0], alpha[0]);
    axpy_prev(DN[1], alpha[1]);

    float * part = p.partials + ((size_t) tok * P + c) * p.n_embd;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = tid + k * NT;
        if (v < V) {
            float4 * o = reinterpret_cast<float4 *>(part) + 2 * v;
            o[0]       = make_float4(acc[k * 8 + 0], acc[k * 8 + 1], acc[k * 8 + 2], acc[k * 8 + 3]);
            o[1]       = make_float4(acc[k * 8 + 4], acc[k * 8 + 5], acc[k * 8 + 6], acc[k * 8 + 7]);
        }
    }
    if (tid == 0 && p.cta_stats) {
        int32_t * st = p.cta_stats + ((size_t) tok * P + c) * 2;
        st[0]        = s_counts[1];
        st[1]        = n_nonzero;
    }
    SPIF_STAMP(5);
    if (p.trace && tid == 0 && blockIdx.y == 0) {
        p.trace[(size_t) blockIdx.x * 8 + 6] = (unsigned long long) n;
    }
}
