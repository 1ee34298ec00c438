// spif_comm.cuh -- one-shot sum over the GPU shards through NVLink peer memory (device side).
//
// Low-latency protocol (the idea of NCCL's LL): every 4-byte value travels in an 8-byte word {value, sequence number} written
// with ONE naturally aligned 8-byte store, which is single-copy atomic -- the receiver polls the word itself, so there is no
// separate flag, no fence and no extra NVLink round trip.  Per block and call:
//   seq = local counter + 1 (never 0, the buffers start zeroed), parity = seq & 1
//   1. a thread holding a float4 stores 4 words into slot [parity][my rank] of every PEER's buffer (NVLink peer stores)
//   2. it polls the 4 words of every peer's slot in its OWN buffer until they carry seq (bounded spin), and adds the values
//      in rank order, its own contribution at its rank's position -> bit-identical on all ranks
// Slot reuse is safe with two parities: a rank can only be two calls ahead of a peer after that peer has sent the words of the
// call in between, which it does after it finished reading the older slot (calls are stream-ordered on every rank).
#pragma once
#include "spif_common.cuh"

struct spif_comm_dev {
    uint2 *    data[SPIF_COMM_MAX_WORLD];  // [2][world][cap] words {float bits, seq}
    uint32_t * seq;
    int        rank, world, cap;
};

__device__ __forceinline__ void st_word_sys(uint2 * p, uint32_t v, uint32_t seq) {
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v), "r"(seq) : "memory");
}
__device__ __forceinline__ uint2 ld_word_sys(const uint2 * p) {
    uint2 v;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

// idx4: float4 index of this thread's value inside y (only read when has_value); block: counter index, unique per block.
// Must be called by all threads of the block.  Returns the sum over ranks for threads with has_value.
__device__ __forceinline__ float4 spif_peer_allreduce_f4(const spif_comm_dev & c, float4 v, int idx4, bool has_value, int block) {
    const int      tid = threadIdx.x + threadIdx.y * blockDim.x;
    const uint32_t seq = c.seq[block] + 1u;
    const uint32_t par = seq & 1u;
    __syncthreads();  // everybody has read the counter before thread 0 advances it
    float4 s = v;
    if (has_value) {
        const float  mine[4] = { v.x, v.y, v.z, v.w };
        const size_t out_off = ((size_t) par * c.world + c.rank) * c.cap + (size_t) idx4 * 4;
        for (int r = 0; r < c.world; ++r) {
            if (r != c.rank) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    st_word_sys(c.data[r] + out_off + e, __float_as_uint(mine[e]), seq);
                }
            }
        }
        float acc[4] = { 0.f, 0.f, 0.f, 0.f };
        for (int r = 0; r < c.world; ++r) {
            if (r == c.rank) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    acc[e] += mine[e];
                }
                continue;
            }
            const uint2 * in = c.data[c.rank] + ((size_t) par * c.world + r) * c.cap + (size_t) idx4 * 4;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                uint2 w = ld_word_sys(in + e);
                // bounded: a peer that died must not hang this GPU forever (the sum is then wrong, which the tests catch)
                for (int spin = 0; w.y != seq && spin < (1 << 22); ++spin) {
                    w = ld_word_sys(in + e);
                }
                acc[e] += __uint_as_float(w.x);
            }
        }
        s = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
    if (tid == 0) {
        c.seq[block] = seq;
    }
    return s;
}

static inline int spif_comm_to_dev(const spif_shard_comm * h, spif_comm_dev & d) {
    SPIF_REQUIRE(h != nullptr && h->seq != nullptr, SPIF_E_NULL);
    SPIF_REQUIRE(h->world >= 1 && h->world <= SPIF_COMM_MAX_WORLD && h->rank >= 0 && h->rank < h->world, SPIF_E_SHAPE);
    SPIF_REQUIRE(h->cap_floats > 0 && h->cap_floats % 4 == 0, SPIF_E_SHAPE);
    for (int r = 0; r < h->world; ++r) {
        SPIF_REQUIRE(h->data[r] != nullptr && ((uintptr_t) h->data[r] & 15) == 0, SPIF_E_NULL);
        d.data[r] = (uint2 *) h->data[r];
    }
    d.seq   = (uint32_t *) h->seq;
    d.rank  = h->rank;
    d.world = h->world;
    d.cap   = h->cap_floats;
    return 0;
}
