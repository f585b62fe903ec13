// xxh32_cta.cuh — XXH32_update by one CTA of 256 threads (xxhash.c:452-526): ONE hashing warp (a lane per accumulator, the
// recurrence cannot be split) fed from double-buffered shared-memory tiles that the other seven warps prefetch.  Used for the
// per-frame content checksum (xxh32.cu) and — one CTA per block — for the block checksums of decode.cu / encode.cu.
#pragma once
#include "common.cuh"

namespace b2 {

constexpr uint32_t XXH_TILE = 16384;      // bytes per staged tile (two tiles in shared memory)
struct XxhCtaShared { uint8_t tile[2 * XXH_TILE]; XxhState s; uint64_t pos, stripes; uint32_t out; uint32_t pad[3]; };

__device__ __forceinline__ void xxh32_state_reset(XxhState& s) {
    s = XxhState{};
    s.v[0] = XP1 + XP2; s.v[1] = XP2; s.v[2] = 0; s.v[3] = 0u - XP1;     // xxhash.c:437-449
}
__device__ __forceinline__ uint32_t xxh32_state_digest(const XxhState& s) {   // xxhash.c:529-554
    uint32_t h = s.large ? xxh32_merge(s.v[0], s.v[1], s.v[2], s.v[3]) : s.v[2] + XP5;
    h += s.total_len;
    return xxh32_tail(h, (const uint8_t*)s.mem, s.memsize);
}

#ifndef B2_XXH_CHAIN
#define B2_XXH_CHAIN 1
#endif
// XXH32_update over data[0, n) (xxhash.c:452-526).  Launch with ONE CTA of 256 threads.
// S.s holds the state (loaded by the caller, visible to the CTA: a __syncthreads() lies between); every thread of the CTA
// calls this; on return S.s is the updated state (and a __syncthreads() has been passed).
__device__ __forceinline__ void xxh32_cta_update(XxhCtaShared& S, const uint8_t* __restrict__ data, uint64_t n)
{
    uint8_t* const tile = S.tile; XxhState& s = S.s; uint64_t& s_pos = S.pos; uint64_t& s_stripes = S.stripes;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        s.total_len += (uint32_t)n;                                           // wraps mod 2^32 by design
        s.large |= (uint32_t)((n >= 16) | (s.total_len >= 16));
        uint64_t pos = 0;
        if (s.memsize + n < 16) {                                             // not a full stripe yet
            for (uint64_t i = 0; i < n; ++i) ((uint8_t*)s.mem)[s.memsize + i] = data[i];
            s.memsize += (uint32_t)n;
            pos = n;
        } else if (s.memsize) {                                               // complete the pending stripe
            uint32_t take = 16 - s.memsize;
            for (uint32_t i = 0; i < take; ++i) ((uint8_t*)s.mem)[s.memsize + i] = data[i];
            for (int k = 0; k < 4; ++k) s.v[k] = xxh32_round(s.v[k], s.mem[k]);
            s.memsize = 0;
            pos = take;
        }
        s_pos = pos;
        s_stripes = (n - pos) >> 4;
    }
    __syncthreads();
    const uint64_t pos = s_pos, stripes = s_stripes;
    const uint64_t ntiles = (stripes * 16 + XXH_TILE - 1) / XXH_TILE;
    // lanes 0..3 of warp 0 own one accumulator each: a stripe then costs the warp one load and three dependent
    // arithmetic instructions instead of sixteen issue slots of a single thread (the chain latency is what remains)
    uint32_t acc = s.v[tid & 3u];
    if (ntiles) {
        uint64_t len0 = stripes * 16 < XXH_TILE ? stripes * 16 : XXH_TILE;
        group_copy(tile, data + pos, len0, tid, 256);
    }
    __syncthreads();
    for (uint64_t t = 0; t < ntiles; ++t) {
        const uint64_t off = t * XXH_TILE;
        const uint64_t len = (stripes * 16 - off) < XXH_TILE ? (stripes * 16 - off) : XXH_TILE;
        if (t + 1 < ntiles) {                                                 // prefetch the next tile
            const uint64_t off2 = off + XXH_TILE;
            const uint64_t len2 = (stripes * 16 - off2) < XXH_TILE ? (stripes * 16 - off2) : XXH_TILE;
            if (tid >= 32) group_copy(tile + ((t + 1) & 1) * XXH_TILE, data + pos + off2, len2, tid - 32, 224);
        }
        if (tid < 4u) {
            const uint32_t* w = (const uint32_t*)(tile + (t & 1) * XXH_TILE) + tid;
            const uint32_t ns = (uint32_t)(len >> 4);
            uint32_t i = 0;
            for (; i + 8 <= ns; i += 8) {
                uint32_t x[8];
                #pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = w[4u * (i + k)];
#if B2_XXH_CHAIN == 0 || !B2_PTX
                #pragma unroll
                for (int k = 0; k < 8; ++k) acc = xxh32_round(acc, x[k]);
#else
                // round(acc, w) = rotl13(acc + w*P2) * P1.  With X = acc + w*P2:  rotl13(X) = (X << 13) + (X >> 19) (disjoint
                // bits), so round = X*(P1 << 13) + (X >> 19)*P1, and the NEXT X is (X >> 19)*P1 + [X*(P1 << 13) + w'*P2]:
                // the bracket does not wait for the shift, which leaves two dependent instructions per stripe instead of three
                // (measured on B200: 1.97 -> 2.44 GB/s for one stream).
                constexpr uint32_t P1 = 2654435761u, P2 = 2246822519u, K1 = P1 << 13;
                #pragma unroll
                for (int k = 0; k < 8; ++k) x[k] *= P2;
                uint32_t X = acc + x[0];
                #pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t t = X >> 19;                               // (__umulhi(X, 1 << 13) on the multiplier pipe: 1.36 GB/s)
                    uint32_t u;                                                // inline PTX: nvcc re-associates the C form into a 3-deep chain
                    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(u) : "r"(X), "r"(K1), "r"(k < 7 ? x[k < 7 ? k + 1 : 7] : 0u));
                    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(X) : "r"(t), "r"(P1), "r"(u));
                }
                acc = X;
#endif
            }
            for (; i < ns; ++i) acc = xxh32_round(acc, w[4u * i]);
        }
        __syncthreads();
    }
    if (tid < 4u) s.v[tid] = acc;
    __syncthreads();
    if (tid == 0) {
        uint64_t done = pos + stripes * 16;
        if (done < n) {                                                       // keep the remainder for later
            uint32_t rem = (uint32_t)(n - done);
            for (uint32_t i = 0; i < rem; ++i) ((uint8_t*)s.mem)[i] = data[done + i];
            s.memsize = rem;
        }
    }
    __syncthreads();
}

// whole hash of [data, data + n) by the CTA; the result is returned to every thread
__device__ __forceinline__ uint32_t xxh32_cta(XxhCtaShared& S, const uint8_t* __restrict__ data, uint64_t n) {
    if (threadIdx.x == 0) xxh32_state_reset(S.s);
    __syncthreads();
    xxh32_cta_update(S, data, n);
    if (threadIdx.x == 0) S.out = xxh32_state_digest(S.s);
    __syncthreads();
    const uint32_t h = S.out;
    __syncthreads();
    return h;
}

}  // namespace b2
