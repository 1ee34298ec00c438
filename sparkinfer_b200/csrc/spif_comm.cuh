// spif_comm.cuh -- one-shot sum over the GPU shards through NVLink peer memory (device side).
//
// Low-latency protocol (the idea of NCCL's LL): every 4-byte value travels in an 8-byte word {value, sequence number} written
// with ONE naturally aligned 8-byte store, which is single-copy atomic -- the receiver polls the word itself, so there is no
// separate flag, no fence and no extra NVLink round trip.
//
// Sequence numbers are PER COMM OBJECT AND PER CALL, never per block: comm.seq[0] holds the number of calls completed on this
// rank ("epoch"); every kernel that uses the buffers reads it after its programmatic-dependent-launch wait (the previous call
// on the stream has completed then), uses seq = epoch + 1 for every word of the call, and its last block to finish stores
// seq back.  All ranks issue the same sequence of calls on a comm, so they agree on seq whatever the launch shapes of the
// individual calls are (fused layer kernel, partial-sum kernel, stand-alone allreduce can be mixed freely on one comm).
// A stale word can never be mistaken for a fresh one: it carries an older seq.  parity = seq & 1 selects one of two slot
// sets: a rank can only be two calls ahead of a peer after that peer has sent the words of the call in between, which it
// does after it finished reading the older slot set (calls are stream-ordered on every rank).
//
// Local control words (comm.seq, uint32, zero-initialised by the caller):
//   [0] epoch   [1] blocks that have finished the current call   [2] sticky error (SPIF_COMM_ERR_*)   [3] layer kernel:
//   CTAs whose partial y is written   [4] layer kernel: CTAs whose predictor rows are sent (a hint)   [5..7] reserved
// A peer that does not answer within the spin bound sets the error word, and the affected outputs are NaN: a wrong sum is
// never returned silently.  spif_shard_comm_status() reads the error word.
#pragma once
#include "spif_common.cuh"

#define SPIF_COMM_ERR_TIMEOUT 1u
#define SPIF_COMM_SPIN_LIMIT  (1u << 24) /* polls of local memory: several seconds */

struct spif_comm_dev {
    uint2 *    data[SPIF_COMM_MAX_WORLD];  // [2][world][cap] words {float bits, seq}
    uint32_t * seq;                        // local control words, see above
    uint2 *    mc;                         // nullable: multicast mapping of the buffers (one store reaches every rank's copy)
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
// two adjacent words with one 16-byte load whose ELEMENTS are the 8-byte words (each element is single-copy atomic)
__device__ __forceinline__ void ld_word2_sys(const uint2 * p, uint2 & a, uint2 & b) {
    unsigned long long x, y;
    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(p) : "memory");
    a = make_uint2((uint32_t) x, (uint32_t) (x >> 32));
    b = make_uint2((uint32_t) y, (uint32_t) (y >> 32));
}
__device__ __forceinline__ uint32_t ld_u32_relaxed_gpu(const uint32_t * p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_u32_acquire_gpu(const uint32_t * p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// seq of the call this kernel belongs to.  Must be read after griddepcontrol.wait (or in a normally ordered launch).
__device__ __forceinline__ uint32_t spif_comm_call_seq(const spif_comm_dev & c) {
    return ld_u32_relaxed_gpu(c.seq + 0) + 1u;
}

// Called by ONE thread per block after all threads of the block are done with the comm (i.e. after a __syncthreads()).
// The last block of the call publishes the new epoch and clears the per-call counters.
__device__ __forceinline__ void spif_comm_block_done(const spif_comm_dev & c, uint32_t seq, unsigned n_blocks) {
    __threadfence();
    const unsigned old = atomicAdd(c.seq + 1, 1u);
    if (old == n_blocks - 1u) {
        c.seq[1] = 0u;
        c.seq[3] = 0u;
        c.seq[4] = 0u;
        __threadfence();
        c.seq[0] = seq;
    }
}

// idx4: float4 index of this thread's value inside the slot (only read when has_value).  Returns the sum over ranks, added
// in RANK ORDER (bit-identical on all ranks) for threads with has_value.  All peer words are requested at once and only the
// missing ones are polled again (the first version polled 4 x (world-1) words one after the other: 12 us at 8 GPUs).
__device__ __forceinline__ float4 spif_peer_allreduce_f4(const spif_comm_dev & c, uint32_t seq, float4 v, int idx4,
                                                         bool has_value) {
    if (!has_value || c.world == 1) {
        return v;
    }
    const uint32_t par     = seq & 1u;
    const size_t   out_off = ((size_t) par * c.world + c.rank) * c.cap + (size_t) idx4 * 4;
    const uint32_t mine[4] = { __float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w) };
    if (c.mc) {  // NVLS multicast: the switch replicates the word into every rank's buffer (our own copy is never read)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            st_word_sys(c.mc + out_off + e, mine[e], seq);
        }
    } else {
#pragma unroll 1
        for (int r = 1; r < c.world; ++r) {  // start with the next rank: spreads the stores over the links
            const int dst = (c.rank + r) % c.world;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                st_word_sys(c.data[dst] + out_off + e, mine[e], seq);
            }
        }
    }
    uint2 w[SPIF_COMM_MAX_WORLD][4];
    const uint2 * in0 = c.data[c.rank] + (size_t) par * c.world * c.cap + (size_t) idx4 * 4;
    unsigned      pending = 0;
#pragma unroll
    for (int r = 0; r < SPIF_COMM_MAX_WORLD; ++r) {
        if (r < c.world && r != c.rank) {
            ld_word2_sys(in0 + (size_t) r * c.cap, w[r][0], w[r][1]);
            ld_word2_sys(in0 + (size_t) r * c.cap + 2, w[r][2], w[r][3]);
            pending |= 1u << r;
        }
    }
    unsigned spins = 0;
    while (pending) {
#pragma unroll
        for (int r = 0; r < SPIF_COMM_MAX_WORLD; ++r) {
            if (pending & (1u << r)) {
                if (w[r][0].y == seq && w[r][1].y == seq && w[r][2].y == seq && w[r][3].y == seq) {
                    pending &= ~(1u << r);
                } else {
                    ld_word2_sys(in0 + (size_t) r * c.cap, w[r][0], w[r][1]);
                    ld_word2_sys(in0 + (size_t) r * c.cap + 2, w[r][2], w[r][3]);
                }
            }
        }
        if (++spins > SPIF_COMM_SPIN_LIMIT) {
            break;
        }
    }
    if (pending) {  // a peer never answered: poison the result and raise the sticky error word
        atomicOr(c.seq + 2, SPIF_COMM_ERR_TIMEOUT);
        const float qnan = __uint_as_float(0x7fc00000u);
        return make_float4(qnan, qnan, qnan, qnan);
    }
    float acc[4] = { 0.f, 0.f, 0.f, 0.f };
#pragma unroll
    for (int r = 0; r < SPIF_COMM_MAX_WORLD; ++r) {
        if (r < c.world) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                acc[e] += (r == c.rank) ? __uint_as_float(mine[e]) : __uint_as_float(w[r][e].x);
            }
        }
    }
    return make_float4(acc[0], acc[1], acc[2], acc[3]);
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
    d.mc    = (h->world > 1) ? (uint2 *) h->mc : nullptr;
    d.rank  = h->rank;
    d.world = h->world;
    d.cap   = h->cap_floats;
    return 0;
}
