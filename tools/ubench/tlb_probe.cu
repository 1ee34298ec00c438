// Microbenchmark: does a cold TLB explain the ~22 GB/s-per-SM ceiling of the row-gather kernels?
// One CTA per SM gathers R rows of 10 KB (640 threads x 16 B per row, 8 rows in flight per thread) from a big pool.
//   warm : the same rows every launch, L2 flushed in between by streaming a 512 MB buffer (TLB entries of the rows survive? the
//          flush touches 256 pages)
//   cold : new random rows every launch, spread over POOL GB (every row in a different 2 MB page, never seen before)
//   local: new rows every launch, but all inside one 256 MB window (128 pages) that moves every launch
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tlb_probe tlb_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__global__ void __launch_bounds__(640, 1) gather(const uint8_t* __restrict__ buf, const long long* __restrict__ rows, int R, float* out) {
    const long long* my = rows + (size_t)blockIdx.x * R;
    float acc = 0.f;
    for (int i = 0; i < R; i += 8) {
        uint4 w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint4* p = (const uint4*)(buf + my[i + j] * 10240) + threadIdx.x;
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[j].x), "=r"(w[j].y), "=r"(w[j].z), "=r"(w[j].w) : "l"(p));
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += __uint_as_float(w[j].x) + __uint_as_float(w[j].w);
    }
    if (acc == 12345.f) out[0] = acc;
}
__global__ void flush(const uint4* __restrict__ b, size_t n, float* out) {
    float acc = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) { uint4 v = b[i]; acc += __uint_as_float(v.x); }
    if (acc == 12345.f) out[0] = acc;
}
int main(int argc, char** argv) {
    cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
    const int P = pr.multiProcessorCount;
    const size_t POOL = (size_t)(argc > 1 ? atoi(argv[1]) : 32) << 30;
    const long long n_rows_total = POOL / 10240;
    uint8_t* buf; CK(cudaMalloc(&buf, POOL)); CK(cudaMemset(buf, 1, POOL));
    const size_t FL = 512ull << 20; uint4* fl; CK(cudaMalloc(&fl, FL)); CK(cudaMemset(fl, 1, FL));
    float* out; CK(cudaMalloc(&out, 4));
    const int REP = 20;
    uint64_t sd = 88172645463325252ull;
    auto rnd = [&]() { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; return sd; };
    printf("SMs %d pool %zu GB\n", P, POOL >> 30);
    for (int R : {8, 16, 32, 48}) {
        for (int mode = 0; mode < 3; ++mode) {
            std::vector<long long> h((size_t)REP * P * R);
            for (int rep = 0; rep < REP; ++rep) for (int i = 0; i < P * R; ++i) {
                long long r;
                if (mode == 0) r = (rep == 0) ? (long long)(rnd() % n_rows_total) : h[i];
                else if (mode == 1) r = (long long)(rnd() % n_rows_total);
                else { const long long win = (256ll << 20) / 10240; const long long base = (long long)((rnd() % 1) + rep) * win * 3 % (n_rows_total - win); r = base + (long long)(rnd() % win); }
                h[(size_t)rep * P * R + i] = r;
            }
            long long* rows; CK(cudaMalloc(&rows, h.size() * 8)); CK(cudaMemcpy(rows, h.data(), h.size() * 8, cudaMemcpyHostToDevice));
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            float tot = 0, best = 1e9;
            for (int rep = 0; rep < REP; ++rep) {
                flush<<<P * 4, 512>>>(fl, FL / 16, out);
                cudaEventRecord(e0);
                gather<<<P, 640>>>(buf, rows + (size_t)rep * P * R, R, out);
                cudaEventRecord(e1); CK(cudaDeviceSynchronize());
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (rep > 0) { tot += ms; best = std::min(best, ms); }
            }
            const double bytes = (double)P * R * 10240;
            printf("rows/SM %3d %-6s avg %7.2f us  best %7.2f us  %7.1f GB/s (avg)  %6.1f GB/s/SM\n", R, mode == 0 ? "warm" : mode == 1 ? "cold" : "local",
                   tot / (REP - 1) * 1e3, best * 1e3, bytes * (REP - 1) / tot / 1e6, bytes * (REP - 1) / tot / 1e6 / P);
            cudaFree(rows);
        }
    }
    return 0;
}
