// Microbenchmark: latency of a small dependent global load (L2 hit) issued from an SM whose TMA unit has D bulk copies of
// B bytes outstanding (all SMs streaming from HBM at once).  Question: do small loads queue behind bulk traffic?
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o lat_under_load lat_under_load.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
    asm volatile("{\n.reg .pred P1;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n@P1 bra WD;\nbra WL;\nWD:\n}" :: "r"(smem_u32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// warp 0: keeps D copies of B bytes in flight (n_items total).  warp 1: pointer chase through `chain` (L2 resident), n_lat hops,
// records min/avg/max latency in cycles.  warp 2: st + poll ping (store a word, another ... skipped)
__global__ void __launch_bounds__(64, 1) k(const uint8_t* __restrict__ big, size_t big_bytes, int D, int B, int n_items,
                                           const int* __restrict__ chain, int n_lat, int mode, unsigned long long* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = (uint64_t*)(smem + (size_t)D * B);
    __shared__ volatile int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { for (int s = 0; s < D; ++s) mbar_init(&full[s], 1); stop = 0; asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (warp == 0) {
        if (lane == 0 && D > 0) {
            unsigned long long t_issue = 0;
            size_t off = ((size_t)blockIdx.x * 7919 * B) % (big_bytes - B);
            for (int i = 0; i < n_items; ++i) {
                const int s = i % D;
                if (i >= D) mbar_wait(&full[s], ((i / D) - 1) & 1);
                mbar_expect(&full[s], B);
                const long long t0 = clock64();
                tma1d(smem + (size_t)s * B, big + off, B, &full[s]);
                t_issue += clock64() - t0;
                off += (size_t)gridDim.x * B; if (off + B > big_bytes) off -= (big_bytes - B) / B * B;
                if (stop) { n_items = i + 1; break; }
            }
            for (int i = max(0, n_items - D); i < n_items; ++i) mbar_wait(&full[i % D], (i / D) & 1);  // drain
            out[blockIdx.x * 8 + 4] = t_issue / n_items;
        }
    } else {
        if (lane == 0) {
            // let the stream ramp up
            long long t = clock64(); while (clock64() - t < 6000) { }
            int idx = blockIdx.x; unsigned long long mn = ~0ull, mx = 0, sum = 0;
            for (int rep = 0; rep < 10; ++rep) {   // 10 chains of n_lat dependent hops: per-chain average latency
                const long long t0 = clock64();
                for (int i = 0; i < n_lat; ++i) {
                    int v;
                    if (mode == 0) asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(chain + idx) : "memory");
                    else asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(chain + idx) : "memory");
                    idx = v;
                }
                const unsigned long long dt = (unsigned long long)(clock64() - t0 + (idx < 0)) / n_lat;
                mn = min(mn, dt); mx = max(mx, dt); sum += dt;
            }
            sum = sum / 10 * n_lat;
            out[blockIdx.x * 8 + 0] = mn; out[blockIdx.x * 8 + 1] = sum / n_lat; out[blockIdx.x * 8 + 2] = mx; out[blockIdx.x * 8 + 3] = idx;
            stop = 1;
        }
    }
}
int main() {
    int dev = 0; cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, dev));
    const int P = pr.multiProcessorCount;
    const size_t big_bytes = 4ull << 30;
    uint8_t* big; CK(cudaMalloc(&big, big_bytes)); CK(cudaMemset(big, 1, big_bytes));
    const int NCH = 1 << 14; std::vector<int> h(NCH); for (int i = 0; i < NCH; ++i) h[i] = (i * 977 + 331) % NCH;
    int* chain; CK(cudaMalloc(&chain, NCH * 4)); CK(cudaMemcpy(chain, h.data(), NCH * 4, cudaMemcpyHostToDevice));
    unsigned long long* out; CK(cudaMalloc(&out, P * 8 * 8));
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    printf("SMs %d, clock %d kHz\n", P, pr.clockRate);
    const int Bs[] = {10240, 40960};
    for (int mode = 0; mode < 2; ++mode)
    for (int B : Bs) for (int D : {0, 1, 2, 3, 4, 5}) {
        if ((size_t)D * B > 210 * 1024) continue;
        if (D == 0 && B != Bs[0]) continue;
        CK(cudaMemset(out, 0, P * 64));
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        const int n_items = 40000 * 10240 / B, n_lat = 20;
        cudaEventRecord(e0);
        k<<<P, 64, (size_t)D * B + 64>>>(big, big_bytes, D, B, n_items, chain, n_lat, mode, out);
        cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        std::vector<unsigned long long> o(P * 8); CK(cudaMemcpy(o.data(), out, P * 64, cudaMemcpyDeviceToHost));
        std::vector<unsigned long long> avg, mx, mn, iss; for (int c = 0; c < P; ++c) { mn.push_back(o[c*8]); avg.push_back(o[c*8+1]); mx.push_back(o[c*8+2]); iss.push_back(o[c*8+4]); }
        std::sort(avg.begin(), avg.end()); std::sort(mx.begin(), mx.end()); std::sort(mn.begin(), mn.end()); std::sort(iss.begin(), iss.end());
        printf("%s B=%6d D=%d (%3d KB in flight): load latency cycles min %llu  avg(med over SMs) %llu  avg(max SM) %llu  max %llu | TMA issue cycles med %llu | %.1f us\n",
               mode ? "ld.relaxed.sys" : "ld.cg         ", B, D, D * B / 1024, mn[0], avg[P / 2], avg[P - 1], mx[P - 1], iss[P / 2], ms * 1e3);
    }
    return 0;
}
