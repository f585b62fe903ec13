// common.cuh — device helpers shared by the sm_100a kernels (byte access, warp copies, XXH32).
#pragma once
#include "b2_cuda.h"
#include <stdint.h>
#include "device_types.h"

namespace b2 {

constexpr unsigned FULL_MASK = 0xffffffffu;

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

#if B2_PTX
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
#else
__device__ __forceinline__ void prefetch_l1(const void*) {}
__device__ __forceinline__ void prefetch_l2(const void*) {}
#endif

// 4 bytes at an arbitrary address, little endian.  Touches only the aligned words that contain them.
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    uintptr_t a = (uintptr_t)p;
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3u) * 8u;
    uint32_t lo = w[0];
    if (sh == 0) return lo;
    return __funnelshift_r(lo, w[1], sh);
}

// ... of memory the kernel only reads (the encoder's input): explicit global loads through the read-only path — a pointer
// rebuilt from an integer is a GENERIC pointer to the compiler (LD instead of LDG)
__device__ __forceinline__ uint32_t ldg32(const uint32_t* p) { return __ldg(p); }
__device__ __forceinline__ uint32_t ld32u_ro(const uint8_t* p) {
    uintptr_t a = (uintptr_t)p;
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3u) * 8u;
    uint32_t lo = __ldg(w);
    if (sh == 0) return lo;
    return __funnelshift_r(lo, __ldg(w + 1), sh);
}

// Copy n bytes with every thread of a group (nthreads threads, this one is `tid`).  dst/src may have
// any alignment and must not overlap.  Bulk part: 16-byte stores to the aligned destination; a misaligned
// source is read as the two aligned 16-byte chunks that hold the bytes and re-aligned with funnel shifts
// (every chunk that is read contains at least one byte of [src, src + n)).
template <int WS>
__device__ __forceinline__ void copy_shifted(uint4* __restrict__ d4, const uint4* __restrict__ s4, uint64_t nq,
                                             uint32_t bs, uint32_t tid, uint32_t nthreads) {
    #pragma unroll 2
    for (uint64_t i = tid; i < nq; i += nthreads) {
        const uint4 a = s4[i], b = s4[i + 1];
        const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        uint4 o;
        o.x = __funnelshift_r(w[WS], w[WS + 1], bs); o.y = __funnelshift_r(w[WS + 1], w[WS + 2], bs);
        o.z = __funnelshift_r(w[WS + 2], w[WS + 3], bs); o.w = __funnelshift_r(w[WS + 3], w[WS + 4], bs);
        d4[i] = o;
    }
}
__device__ __forceinline__ void group_copy(uint8_t* dst, const uint8_t* src, uint64_t n,
                                           uint32_t tid, uint32_t nthreads) {
    if (n < 64) { for (uint64_t i = tid; i < n; i += nthreads) dst[i] = src[i]; return; }
    const uint32_t head = (uint32_t)((16u - ((uintptr_t)dst & 15u)) & 15u);
    if (tid < head) dst[tid] = src[tid];
    dst += head; src += head; n -= head;
    const uint64_t nq = n >> 4;
    const uint32_t sh = (uint32_t)((uintptr_t)src & 15u);
    const uint4* s4 = (const uint4*)(src - sh);
    uint4* d4 = (uint4*)dst;
    if (sh == 0) {
        #pragma unroll 4
        for (uint64_t i = tid; i < nq; i += nthreads) d4[i] = s4[i];
    } else {
        const uint32_t bs = (sh & 3u) * 8u;
        switch (sh >> 2) {
            case 0: copy_shifted<0>(d4, s4, nq, bs, tid, nthreads); break;
            case 1: copy_shifted<1>(d4, s4, nq, bs, tid, nthreads); break;
            case 2: copy_shifted<2>(d4, s4, nq, bs, tid, nthreads); break;
            default: copy_shifted<3>(d4, s4, nq, bs, tid, nthreads); break;
        }
    }
    const uint64_t done = nq << 4;
    if (tid < (uint32_t)(n - done)) dst[done + tid] = src[done + tid];
}

// Small-footprint variant for code that sits inside a hot loop (few registers): 4-byte stores, source re-aligned
// with a funnel shift.
__device__ __forceinline__ void group_copy_small(uint8_t* dst, const uint8_t* src, uint64_t n,
                                                 uint32_t tid, uint32_t nthreads) {
    if (n < 64) { for (uint64_t i = tid; i < n; i += nthreads) dst[i] = src[i]; return; }
    uint32_t head = (uint32_t)((4u - ((uintptr_t)dst & 3u)) & 3u);
    if (tid < head) dst[tid] = src[tid];
    dst += head; src += head; n -= head;
    uint64_t nw = n >> 2;
    uint32_t sh = (uint32_t)((uintptr_t)src & 3u) * 8u;
    const uint32_t* sw = (const uint32_t*)((uintptr_t)src & ~(uintptr_t)3);
    uint32_t* dw = (uint32_t*)dst;
    if (sh == 0) { for (uint64_t i = tid; i < nw; i += nthreads) dw[i] = sw[i]; }
    else { for (uint64_t i = tid; i < nw; i += nthreads) dw[i] = __funnelshift_r(sw[i], sw[i + 1], sh); }
    uint64_t done = nw << 2;
    if (tid < (uint32_t)(n - done)) dst[done + tid] = src[done + tid];
}

// ---- XXH32 over [p, p+n) by ONE thread, seed 0 (xxhash.c:350-416).  The four accumulators are
// independent, so the loop body has 4-way ILP.
__device__ __forceinline__ uint32_t xxh32_thread(const uint8_t* p, uint64_t n) {
    uint32_t h;
    uint64_t i = 0;
    if (n >= 16) {
        uint32_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0u - XP1;
        if (((uintptr_t)p & 3u) == 0) {
            const uint32_t* w = (const uint32_t*)p;
            for (; i + 16 <= n; i += 16) {
                uint32_t a = w[0], b = w[1], c = w[2], d = w[3]; w += 4;
                v1 = xxh32_round(v1, a); v2 = xxh32_round(v2, b); v3 = xxh32_round(v3, c); v4 = xxh32_round(v4, d);
            }
        } else {
            for (; i + 16 <= n; i += 16) {
                v1 = xxh32_round(v1, ld32u(p + i)); v2 = xxh32_round(v2, ld32u(p + i + 4));
                v3 = xxh32_round(v3, ld32u(p + i + 8)); v4 = xxh32_round(v4, ld32u(p + i + 12));
            }
        }
        h = xxh32_merge(v1, v2, v3, v4);
    } else {
        h = XP5;
    }
    h += (uint32_t)n;
    return xxh32_tail(h, p + i, (uint32_t)(n - i));
}

// ---- XXH32 over [p, p+n) by ONE WARP, seed 0; every lane returns the hash.  The recurrence is serial per accumulator
// (xxhash.c:269-275), so a warp cannot split the stream — what it can do is load it coalesced: 32 consecutive words (eight
// 16-byte stripes) per load, every lane multiplying its own word by PRIME2 on the way in, then lanes 0..3 (one accumulator
// each) pull their eight words over with shuffles.  ~3 x a lone thread on long blocks (block checksums of 4 MB blocks).
__device__ __forceinline__ uint32_t xxh32_warp(const uint8_t* p, uint64_t n, uint32_t lane) {
    uint32_t h;
    uint64_t done = 0;
    if (n >= 16) {
        const uint32_t q = lane & 3u;
        uint32_t acc = q == 0u ? XP1 + XP2 : q == 1u ? XP2 : q == 2u ? 0u : 0u - XP1;
        const uint64_t nstripes = n >> 4;
        uint32_t w = (uint64_t)(lane >> 2) < nstripes ? ld32u(p + 4u * lane) * XP2 : 0u;
        for (uint64_t s0 = 0; s0 < nstripes; s0 += 8) {
            const uint32_t ns = nstripes - s0 < 8 ? (uint32_t)(nstripes - s0) : 8u;
            const uint32_t cur = w;
            const uint64_t s1 = s0 + 8;                                       // the next group is requested before this one is hashed
            w = (s1 + (lane >> 2) < nstripes) ? ld32u(p + s1 * 16 + 4u * lane) * XP2 : 0u;
            #pragma unroll
            for (uint32_t k = 0; k < 8; ++k) {
                const uint32_t m = __shfl_sync(FULL_MASK, cur, 4u * k + q);
                if (k < ns) { acc += m; acc = (acc << 13) | (acc >> 19); acc *= XP1; }
            }
        }
        const uint32_t v1 = __shfl_sync(FULL_MASK, acc, 0), v2 = __shfl_sync(FULL_MASK, acc, 1);
        const uint32_t v3 = __shfl_sync(FULL_MASK, acc, 2), v4 = __shfl_sync(FULL_MASK, acc, 3);
        h = xxh32_merge(v1, v2, v3, v4);
        done = nstripes << 4;
    } else {
        h = XP5;
    }
    h += (uint32_t)n;
    return xxh32_tail(h, p + done, (uint32_t)(n - done));
}

}  // namespace b2
