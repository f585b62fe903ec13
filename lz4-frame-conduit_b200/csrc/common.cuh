// common.cuh — device helpers shared by the sm_100a kernels (byte access, warp copies, XXH32).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "device_types.h"

namespace b2 {

constexpr unsigned FULL_MASK = 0xffffffffu;

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

// 4 bytes at an arbitrary address, little endian.  Touches only the aligned words that contain them.
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    uintptr_t a = (uintptr_t)p;
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3u) * 8u;
    uint32_t lo = w[0];
    if (sh == 0) return lo;
    return __funnelshift_r(lo, w[1], sh);
}

// Copy n bytes with every thread of a group (nthreads threads, this one is `tid`).  dst/src may have
// any alignment and must not overlap.  Bulk part: 4-byte stores, source re-aligned with a funnel shift,
// so a warp writes 128 contiguous bytes per instruction.
__device__ __forceinline__ void group_copy(uint8_t* dst, const uint8_t* src, uint64_t n,
                                           uint32_t tid, uint32_t nthreads) {
    if (n < 64) { for (uint64_t i = tid; i < n; i += nthreads) dst[i] = src[i]; return; }
    uint32_t head = (uint32_t)((4u - ((uintptr_t)dst & 3u)) & 3u);
    if (tid < head) dst[tid] = src[tid];
    dst += head; src += head; n -= head;
    uint64_t nw = n >> 2;
    uint32_t sh = (uint32_t)((uintptr_t)src & 3u) * 8u;
    const uint32_t* sw = (const uint32_t*)((uintptr_t)src & ~(uintptr_t)3);
    uint32_t* dw = (uint32_t*)dst;
    if (sh == 0) {
        if ((((uintptr_t)dw | (uintptr_t)sw) & 15u) == 0) {
            uint64_t nq = nw >> 2;
            const uint4* s4 = (const uint4*)sw; uint4* d4 = (uint4*)dw;
            for (uint64_t i = tid; i < nq; i += nthreads) d4[i] = s4[i];
            for (uint64_t i = (nq << 2) + tid; i < nw; i += nthreads) dw[i] = sw[i];
        } else {
            for (uint64_t i = tid; i < nw; i += nthreads) dw[i] = sw[i];
        }
    } else {
        for (uint64_t i = tid; i < nw; i += nthreads) dw[i] = __funnelshift_r(sw[i], sw[i + 1], sh);
    }
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

}  // namespace b2
