// spif_comm.cu -- stand-alone form of the shard exchange (spif_comm.cuh): y = sum over GPU shards of y, in place.
// Used after the batched / quantised layers; the single-token F16/BF16 layer has it fused into its partial-sum kernel.
#include "spif_comm.cuh"

__global__ void __launch_bounds__(256) spif_shard_allreduce_kernel(float * __restrict__ y, int n4, const spif_comm_dev comm) {
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const uint32_t seq = spif_comm_call_seq(comm);
    const int      i   = blockIdx.x * blockDim.x + threadIdx.x;
    const bool     own = i < n4;
    float4         v   = own ? reinterpret_cast<const float4 *>(y)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    v                  = spif_peer_allreduce_f4(comm, seq, v, i, own);
    if (own) {
        reinterpret_cast<float4 *>(y)[i] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        spif_comm_block_done(comm, seq, gridDim.x);
    }
}

int spif_launch_shard_allreduce(float * y, int64_t n, const spif_shard_comm * comm, cudaStream_t stream) {
    spif_comm_dev cd = {};
    int           rc = spif_comm_to_dev(comm, cd);
    if (rc) {
        return rc;
    }
    SPIF_REQUIRE(y != nullptr, SPIF_E_NULL);
    SPIF_REQUIRE(n > 0 && n % 4 == 0 && n <= comm->cap_floats && ((uintptr_t) y & 15) == 0, SPIF_E_SHAPE);
    const int n4     = (int) (n / 4);
    // few, fat blocks keep the flag traffic low; at least 64 threads per block so that a block's stores fill NVLink packets
    int       blocks = (n4 + 255) / 256;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(blocks);
    cfg.blockDim           = dim3(256);
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    cfg.numAttrs                                     = 1;
    SPIF_CUDA_TRY(cudaLaunchKernelEx(&cfg, spif_shard_allreduce_kernel, y, n4, cd));
    return 0;
}
