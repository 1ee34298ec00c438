// Microbenchmark behind the layer kernel's load path (round 2): 20 consumer warps per CTA, one CTA per SM, every warp owns the
// 16-byte column slices [32w, 32w+32) of every 10 KB row (x in registers, real f16 dot + packed shuffle tree + smem deposit),
// rows gathered at random from a 4 GB pool.  How should a warp keep its loads in flight?
//   ldg8     : 8 rows per batch straight into registers, wait, compute (warps overlap each other)
//   ldg4x2   : two 4-row batches, the next one issued before the current one is used (register double buffer)
//   cpasync  : cp.async 16 B per lane into the thread's own staging slots, D batches of 4 rows in flight, wait_group
//   wtma     : ONE lane per warp issues cp.async.bulk copies of the warp's 512-byte slice of each row into the warp's own
//              staging ring (mbarrier complete_tx), D batches of 4 rows in flight
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o stream_items stream_items.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint4 ldg_na(const void* p) {
    uint4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r;
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ float dot8(const uint4& v, const float* x) {
    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i])); s0 = fmaf(t.x, x[2 * i], s0); s1 = fmaf(t.y, x[2 * i + 1], s1); }
    return s0 + s1;
}
__device__ __forceinline__ float reduce4(float v0, float v1, float v2, float v3, int lane) {
    const bool hi = (lane & 16) != 0, h8 = (lane & 8) != 0;
    float k0 = (hi ? v2 : v0) + __shfl_xor_sync(0xffffffffu, hi ? v0 : v2, 16);
    float k1 = (hi ? v3 : v1) + __shfl_xor_sync(0xffffffffu, hi ? v1 : v3, 16);
    float m  = (h8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, h8 ? k0 : k1, 8);
    m += __shfl_xor_sync(0xffffffffu, m, 4); m += __shfl_xor_sync(0xffffffffu, m, 2); m += __shfl_xor_sync(0xffffffffu, m, 1);
    return m;
}
__device__ __forceinline__ void deposit4(const uint4 (&w)[4], const float* x, float* part, int lane, int warp) {
    const float s = reduce4(dot8(w[0], x), dot8(w[1], x), dot8(w[2], x), dot8(w[3], x), lane);
    if ((lane & 7) == 0) part[(lane >> 3) * 32 + warp] = s;
}

enum { M_LDG8 = 0, M_LDG4X2 = 1, M_CPASYNC = 2, M_WTMA = 3 };

template <int MODE, int D>
__global__ void __launch_bounds__(640, 1) k_items(const uint8_t* __restrict__ buf, const int* __restrict__ rows, int n_rows, const float* __restrict__ xg, float* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ int   s_rows[512];
    __shared__ float s_part[16 * 128];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int* my = rows + (size_t)blockIdx.x * n_rows;
    for (int i = tid; i < n_rows; i += 640) s_rows[i] = my[i];
    float x[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) x[e] = xg[tid * 8 + e];
    __syncthreads();
    const uint32_t toff = tid * 16u;
    const size_t RB = 10240;
    int k = 0;
    if (MODE == M_LDG8) {
        for (int i = 0; i < n_rows; i += 8) {
            uint4 w[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = ldg_na(buf + (size_t)s_rows[i + j] * RB + toff);
            const uint4 (&a)[4] = *reinterpret_cast<const uint4 (*)[4]>(&w[0]);
            const uint4 (&b)[4] = *reinterpret_cast<const uint4 (*)[4]>(&w[4]);
            deposit4(a, x, s_part + (k++ & 15) * 128, lane, warp);
            deposit4(b, x, s_part + (k++ & 15) * 128, lane, warp);
        }
    } else if (MODE == M_LDG4X2) {
        uint4 A[4], B[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) A[j] = ldg_na(buf + (size_t)s_rows[j] * RB + toff);
        for (int i = 4; i < n_rows; i += 8) {
#pragma unroll
            for (int j = 0; j < 4; ++j) B[j] = ldg_na(buf + (size_t)s_rows[i + j] * RB + toff);
            deposit4(A, x, s_part + (k++ & 15) * 128, lane, warp);
            if (i + 4 < n_rows) {
#pragma unroll
                for (int j = 0; j < 4; ++j) A[j] = ldg_na(buf + (size_t)s_rows[i + 4 + j] * RB + toff);
            }
            deposit4(B, x, s_part + (k++ & 15) * 128, lane, warp);
        }
    } else if (MODE == M_CPASYNC) {
        // staging [D][4][640] x 16 B: a thread reads back only what it copied itself -> no barrier at all
        const uint32_t st = smem_u32(smem) + toff;
        auto issue = [&](int it) {
            if (it * 4 < n_rows) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(st + (uint32_t)(((it % D) * 4 + j) * 10240)), "l"(buf + (size_t)s_rows[it * 4 + j] * RB + toff) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
#pragma unroll
        for (int it = 0; it < D - 1; ++it) issue(it);
        const int n_it = n_rows / 4;
        for (int it = 0; it < n_it; ++it) {
            issue(it + D - 1);
            asm volatile("cp.async.wait_group %0;" :: "n"(D - 1) : "memory");
            uint4 w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = lds128(st + (uint32_t)(((it % D) * 4 + j) * 10240));
            deposit4(w, x, s_part + (k++ & 15) * 128, lane, warp);
        }
    } else {
        // per-warp ring [D][4][512 B] + D mbarriers per warp
        uint8_t*  ring = smem + (size_t)warp * D * 2048;
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)20 * D * 2048) + warp * D;
        if (lane == 0) {
            for (int s = 0; s < D; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bars[s])) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        const uint32_t woff = warp * 512u;
        auto issue = [&](int it) {
            if (it * 4 < n_rows && lane == 0) {
                const uint32_t bar = smem_u32(&bars[it % D]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 2048;" :: "r"(bar) : "memory");
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];"
                                 :: "r"(smem_u32(ring + ((it % D) * 4 + j) * 512)), "l"(buf + (size_t)s_rows[it * 4 + j] * RB + woff), "r"(bar) : "memory");
            }
        };
        for (int it = 0; it < D - 1; ++it) issue(it);
        const int n_it = n_rows / 4;
        for (int it = 0; it < n_it; ++it) {
            issue(it + D - 1);
            const uint32_t bar = smem_u32(&bars[it % D]), ph = (it / D) & 1;
            asm volatile("{\n.reg .pred P1;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n@P1 bra WD;\nbra WL;\nWD:\n}" :: "r"(bar), "r"(ph) : "memory");
            uint4 w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = lds128(smem_u32(ring + ((it % D) * 4 + j) * 512) + lane * 16u);
            deposit4(w, x, s_part + (k++ & 15) * 128, lane, warp);
            __syncwarp();  // every lane has its values in registers before lane 0 reuses the slot D-1 items later
        }
    }
    __syncthreads();
    if (tid < 128) out[blockIdx.x * 128 + tid] = s_part[tid];
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int P = prop.multiProcessorCount;
    const size_t n_pool = 400000, RB = 10240;
    uint8_t* buf; CK(cudaMalloc(&buf, n_pool * RB)); CK(cudaMemset(buf, 1, n_pool * RB));
    float *out, *xg; CK(cudaMalloc(&out, P * 128 * 4)); CK(cudaMalloc(&xg, 5120 * 4)); CK(cudaMemset(xg, 0, 5120 * 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    printf("SMs %d, 640 consumer threads per CTA, 1 CTA per SM, f16 dot + shuffle tree + smem deposit per 4 rows\n", P);
    for (int n_rows : {48, 480}) {
        const int REP = 20;
        std::vector<int> h((size_t)P * n_rows * REP);
        uint64_t sd = 88172645463325252ull;
        for (auto& r : h) { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; r = (int)(sd % n_pool); }
        int* rows; CK(cudaMalloc(&rows, h.size() * 4)); CK(cudaMemcpy(rows, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
        const double bytes = (double)P * n_rows * RB;
        auto timeit = [&](const char* name, auto launch) {
            cudaGraph_t g; cudaGraphExec_t ge;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
            for (int r = 0; r < REP; ++r) launch(st, rows + (size_t)r * P * n_rows);
            CK(cudaStreamEndCapture(st, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
            CK(cudaGraphLaunch(ge, st)); CK(cudaStreamSynchronize(st));
            float best = 1e9f;
            for (int it = 0; it < 5; ++it) { CK(cudaEventRecord(e0, st)); CK(cudaGraphLaunch(ge, st)); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = ms < best ? ms : best; }
            CK(cudaGetLastError());
            printf("rows/SM %4d %-28s %8.2f us/launch  %7.1f GB/s\n", n_rows, name, best * 1e3 / REP, bytes * REP / best / 1e6);
            CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
        };
#define RUN(MODE, D, SM, NAME) do { CK(cudaFuncSetAttribute(k_items<MODE, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SM))); \
        timeit(NAME, [&](cudaStream_t s_, const int* r_) { k_items<MODE, D><<<P, 640, (SM), s_>>>(buf, r_, n_rows, xg, out); }); } while (0)
        RUN(M_LDG8, 1, 0, "ldg8 (batch, no pipeline)");
        RUN(M_LDG4X2, 1, 0, "ldg4x2 (register dbl buf)");
        RUN(M_CPASYNC, 2, 2 * 40960, "cpasync D=2");
        RUN(M_CPASYNC, 3, 3 * 40960, "cpasync D=3");
        RUN(M_CPASYNC, 4, 4 * 40960, "cpasync D=4");
        RUN(M_CPASYNC, 5, 5 * 40960, "cpasync D=5");
        RUN(M_WTMA, 2, 20 * 2 * 2048 + 20 * 2 * 8, "wtma D=2");
        RUN(M_WTMA, 3, 20 * 3 * 2048 + 20 * 3 * 8, "wtma D=3");
        RUN(M_WTMA, 4, 20 * 4 * 2048 + 20 * 4 * 8, "wtma D=4");
        RUN(M_WTMA, 5, 20 * 5 * 2048 + 20 * 5 * 8, "wtma D=5");
        CK(cudaFree(rows));
    }
    return 0;
}
