// Microbenchmark: how fast can one CTA per SM gather random 10 KB rows from HBM, by load path?
//   tma   : cp.async.bulk (1-D TMA) into a smem ring, one producer thread, chunk = row/split bytes
//   ldg   : ld.global.nc.v4 straight into registers, U rows in flight per thread
//   ldgsts: cp.async 16 B into a private smem ring, S rows in flight per thread
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o stream_rows stream_rows.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
    asm volatile("{\n.reg .pred P1;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra WD;\nbra WL;\nWD:\n}" :: "r"(smem_u32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---------------- TMA ring
template <int PRODUCERS>
__global__ void __launch_bounds__(512, 1) k_tma(const uint8_t* __restrict__ buf, const int* __restrict__ rows, int n_per_cta, int row_bytes, int split, int S, float* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = (uint64_t*)(smem + (size_t)S * row_bytes);
    uint64_t* empty = full + S;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nwc = (blockDim.x >> 5) - 1;
    if (tid == 0) { for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nwc); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const int* my = rows + (size_t)blockIdx.x * n_per_cta;
    if (warp == nwc) {
        if (lane < PRODUCERS) {
            const int chunk = row_bytes / split;
            for (int i = 0; i < n_per_cta; ++i) {
                const int s = i % S;
                if (i >= S) mbar_wait(&empty[s], ((i / S) - 1) & 1);
                if (lane == 0) mbar_expect(&full[s], row_bytes);
                __syncwarp((1u << PRODUCERS) - 1);
                const uint8_t* src = buf + (size_t)my[i] * row_bytes;
                for (int c = lane; c < split; c += PRODUCERS) tma1d(smem + (size_t)s * row_bytes + c * chunk, src + c * chunk, chunk, &full[s]);
            }
        }
        return;
    }
    float acc = 0.f;
    const int nc = nwc * 32;
    for (int i = 0; i < n_per_cta; ++i) {
        const int s = i % S;
        mbar_wait(&full[s], (i / S) & 1);
        const uint4* r = (const uint4*)(smem + (size_t)s * row_bytes);
        for (int v = tid; v < row_bytes / 16; v += nc) { uint4 w = r[v]; acc += __uint_as_float(w.x) + __uint_as_float(w.w); }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (acc == 12345.f) out[0] = acc;
}

// ---------------- TMA ring, NP producer warps (each issues rows i == w mod NP)
__global__ void __launch_bounds__(512, 1) k_tma_np(const uint8_t* __restrict__ buf, const int* __restrict__ rows, int n_per_cta, int row_bytes, int S, int NP, float* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = (uint64_t*)(smem + (size_t)S * row_bytes);
    uint64_t* empty = full + S;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nwc = (blockDim.x >> 5) - NP;
    if (tid == 0) { for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nwc); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const int* my = rows + (size_t)blockIdx.x * n_per_cta;
    if (warp >= nwc) {
        const int w = warp - nwc;
        if (lane == 0) {
            for (int i = w; i < n_per_cta; i += NP) {
                const int s = i % S;
                if (i >= S) mbar_wait(&empty[s], ((i / S) - 1) & 1);
                mbar_expect(&full[s], row_bytes);
                tma1d(smem + (size_t)s * row_bytes, buf + (size_t)my[i] * row_bytes, row_bytes, &full[s]);
            }
        }
        return;
    }
    float acc = 0.f;
    const int nc = nwc * 32;
    for (int i = 0; i < n_per_cta; ++i) {
        const int s = i % S;
        mbar_wait(&full[s], (i / S) & 1);
        const uint4* r = (const uint4*)(smem + (size_t)s * row_bytes);
        for (int v = tid; v < row_bytes / 16; v += nc) { uint4 w = r[v]; acc += __uint_as_float(w.x) + __uint_as_float(w.w); }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (acc == 12345.f) out[0] = acc;
}

// ---------------- LDG to registers, U rows in flight
template <int U, int VPT>
__global__ void __launch_bounds__(640) k_ldg(const uint8_t* __restrict__ buf, const int* __restrict__ rows, int n_per_cta, int row_bytes, float* out) {
    const int tid = threadIdx.x, nc = blockDim.x;
    const int* my = rows + (size_t)blockIdx.x * n_per_cta;
    const int V = row_bytes / 16;
    float acc = 0.f;
    for (int i0 = 0; i0 < n_per_cta; i0 += U) {
        uint4 w[U][VPT];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = min(i0 + u, n_per_cta - 1);
            const uint4* r = (const uint4*)(buf + (size_t)my[i] * row_bytes);
#pragma unroll
            for (int k = 0; k < VPT; ++k) { const int v = min(tid + k * nc, V - 1); asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[u][k].x), "=r"(w[u][k].y), "=r"(w[u][k].z), "=r"(w[u][k].w) : "l"(r + v)); }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < VPT; ++k) acc += __uint_as_float(w[u][k].x) + __uint_as_float(w[u][k].w);
    }
    if (acc == 12345.f) out[0] = acc;
}

// ---------------- cp.async (LDGSTS) private ring, S rows in flight
template <int VPT>
__global__ void __launch_bounds__(640) k_ldgsts(const uint8_t* __restrict__ buf, const int* __restrict__ rows, int n_per_cta, int row_bytes, int S, float* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, nc = blockDim.x;
    const int* my = rows + (size_t)blockIdx.x * n_per_cta;
    const int V = row_bytes / 16;
    auto issue = [&](int i) {
        if (i < n_per_cta) {
            const uint8_t* r = buf + (size_t)my[i] * row_bytes;
            uint8_t* d = smem + (size_t)(i % S) * row_bytes;
#pragma unroll
            for (int k = 0; k < VPT; ++k) { const int v = tid + k * nc; if (v < V) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(d + v * 16)), "l"(r + v * 16) : "memory"); }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int i = 0; i < S - 1; ++i) issue(i);
    float acc = 0.f;
    for (int i = 0; i < n_per_cta; ++i) {
        issue(i + S - 1);
        // groups committed so far: i + S; need group i complete -> allow S-1 pending
        switch (S - 1) {
            case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
            case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
            case 7: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
            case 15: asm volatile("cp.async.wait_group 15;" ::: "memory"); break;
            default: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        }
        const uint4* r = (const uint4*)(smem + (size_t)(i % S) * row_bytes);
#pragma unroll
        for (int k = 0; k < VPT; ++k) { const int v = tid + k * nc; if (v < V) { uint4 w = r[v]; acc += __uint_as_float(w.x) + __uint_as_float(w.w); } }
    }
    if (acc == 12345.f) out[0] = acc;
}

int main(int argc, char** argv) {
    int dev = 0; cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, dev));
    const int P = prop.multiProcessorCount;
    const int row_bytes = 10240;
    const size_t n_rows_total = 400000;  // 4 GB pool
    uint8_t* buf; CK(cudaMalloc(&buf, n_rows_total * row_bytes)); CK(cudaMemset(buf, 1, n_rows_total * row_bytes));
    float* out; CK(cudaMalloc(&out, 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    printf("SMs %d\n", P);
    cudaStream_t st; CK(cudaStreamCreate(&st));
    for (int n_per_cta : {30, 400}) {
        for (int ctas_per_sm : {1, 2}) {
            const int G = P * ctas_per_sm;
            const int npc = n_per_cta / ctas_per_sm;
            const int REP = 20;
            std::vector<int> h((size_t)G * npc * REP);
            uint64_t sd = 88172645463325252ull;
            for (auto& r : h) { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; r = (int)(sd % n_rows_total); }
            int* rows; CK(cudaMalloc(&rows, h.size() * 4)); CK(cudaMemcpy(rows, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
            const double bytes = (double)G * npc * row_bytes;
            // time REP back-to-back launches (different rows each) inside one CUDA graph: steady-state per-launch time
            auto timeit = [&](const char* name, auto launch) {
                cudaGraph_t g; cudaGraphExec_t ge;
                CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
                for (int r = 0; r < REP; ++r) launch(st, rows + (size_t)r * G * npc);
                CK(cudaStreamEndCapture(st, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
                CK(cudaGraphLaunch(ge, st)); CK(cudaStreamSynchronize(st));
                float best = 1e9f;
                for (int it = 0; it < 5; ++it) { CK(cudaEventRecord(e0, st)); CK(cudaGraphLaunch(ge, st)); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = ms < best ? ms : best; }
                CK(cudaGetLastError());
                printf("rows/SM %4d ctas/SM %d %-34s %8.2f us/launch  %7.1f GB/s\n", n_per_cta, ctas_per_sm, name, best * 1e3 / REP, bytes * REP / best / 1e6);
                CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
            };
            for (int S : {4, 8}) for (int NP : {1, 2, 4}) {
                if (ctas_per_sm * ((size_t)S * row_bytes + 1024) > 220 * 1024) continue;
                size_t sm = (size_t)S * row_bytes + 16 * S + 64;
                char nm[64];
                CK(cudaFuncSetAttribute(k_tma_np, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                snprintf(nm, 64, "tma S=%d producers=%d", S, NP);
                timeit(nm, [&](cudaStream_t s_, const int* r_) { k_tma_np<<<G, 320 + 32 * NP, sm, s_>>>(buf, r_, npc, row_bytes, S, NP, out); });
            }
            timeit("ldg T=320 U=2 VPT=2", [&](cudaStream_t s_, const int* r_) { k_ldg<2, 2><<<G, 320, 0, s_>>>(buf, r_, npc, row_bytes, out); });
            timeit("ldg T=320 U=4 VPT=2", [&](cudaStream_t s_, const int* r_) { k_ldg<4, 2><<<G, 320, 0, s_>>>(buf, r_, npc, row_bytes, out); });
            timeit("ldg T=640 U=4 VPT=1", [&](cudaStream_t s_, const int* r_) { k_ldg<4, 1><<<G, 640, 0, s_>>>(buf, r_, npc, row_bytes, out); });
            timeit("ldg T=640 U=8 VPT=1", [&](cudaStream_t s_, const int* r_) { k_ldg<8, 1><<<G, 640, 0, s_>>>(buf, r_, npc, row_bytes, out); });
            for (int S : {4, 8}) {
                if (ctas_per_sm * ((size_t)S * row_bytes) > 220 * 1024) continue;
                size_t sm = (size_t)S * row_bytes;
                CK(cudaFuncSetAttribute(k_ldgsts<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                CK(cudaFuncSetAttribute(k_ldgsts<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                char nm[64]; snprintf(nm, 64, "ldgsts T=320 S=%d", S);
                timeit(nm, [&](cudaStream_t s_, const int* r_) { k_ldgsts<2><<<G, 320, sm, s_>>>(buf, r_, npc, row_bytes, S, out); });
                snprintf(nm, 64, "ldgsts T=640 S=%d", S);
                timeit(nm, [&](cudaStream_t s_, const int* r_) { k_ldgsts<1><<<G, 640, sm, s_>>>(buf, r_, npc, row_bytes, S, out); });
            }
            CK(cudaFree(rows));
        }
    }
    return 0;
}
