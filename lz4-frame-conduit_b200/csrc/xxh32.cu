// xxh32.cu — XXH32 on sm_100a (seed 0 everywhere on this path).
//
// Replaces XXH32 / XXH32_reset / XXH32_update / XXH32_digest (lz4/lib/xxhash.c:392-416, 437-554) where
// the frame layer uses them: per-block checksums (many independent hashes, one thread each — see
// xxh32_thread in common.cuh) and the per-frame content checksum.  The content checksum is a single
// serial recurrence (4 accumulators, xxhash.c:269-275): it cannot be split across threads, so the
// stream kernel below uses ONE hashing warp (a lane per accumulator) fed from shared memory while the rest of the CTA
// prefetches the next tile; it is meant to run on its own stream, overlapped with the codec kernels.
#include "xxh32_cta.cuh"
#include "kernels.h"

namespace b2 {

__global__ void xxh32_reset_kernel(XxhState* st) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        XxhState s{};
        s.v[0] = XP1 + XP2; s.v[1] = XP2; s.v[2] = 0; s.v[3] = 0u - XP1;     // xxhash.c:437-449
        *st = s;
    }
}

// XXH32_update over data[0, n) (xxhash.c:452-526).  Launch with ONE CTA of 256 threads.
__global__ void __launch_bounds__(256)
xxh32_stream_kernel(const uint8_t* __restrict__ data, uint64_t n, const uint64_t* __restrict__ n_ptr,
                    const uint32_t* __restrict__ enable_ptr, XxhState* __restrict__ st)
{
    if (enable_ptr && !*enable_ptr) return;      // e.g. &FrameIndex::desc.content_checksum
    if (n_ptr) n = *n_ptr;                        // e.g. &DecodeResult::total_out
    B2_SHARED_VAR(XxhCtaShared, S);
    if (threadIdx.x == 0) S.s = *st;
    __syncthreads();
    xxh32_cta_update(S, data, n);
    if (threadIdx.x == 0) *st = S.s;
}

// XXH32_digest (xxhash.c:529-554); optionally stores the value little-endian at base + *off_ptr
__global__ void xxh32_digest_kernel(const XxhState* __restrict__ st, uint32_t* __restrict__ out,
                                    uint8_t* base, const uint64_t* off_ptr)
{
    if (threadIdx.x || blockIdx.x) return;
    XxhState s = *st;
    uint32_t h = s.large ? xxh32_merge(s.v[0], s.v[1], s.v[2], s.v[3]) : s.v[2] + XP5;
    h += s.total_len;
    h = xxh32_tail(h, (const uint8_t*)s.mem, s.memsize);
    if (out) *out = h;
    if (base && off_ptr && *off_ptr) st_le32(base + *off_ptr, h);
}

// decode side: compare the digest with the checksum stored in the frame (lz4frame.c:1734-1745)
__global__ void content_check_kernel(const FrameIndex* __restrict__ fi, const XxhState* __restrict__ st,
                                     DecodeResult* __restrict__ res)
{
    if (threadIdx.x || blockIdx.x) return;
    if (fi->err != OK || !fi->desc.content_checksum) return;
    XxhState s = *st;
    uint32_t h = s.large ? xxh32_merge(s.v[0], s.v[1], s.v[2], s.v[3]) : s.v[2] + XP5;
    h += s.total_len;
    h = xxh32_tail(h, (const uint8_t*)s.mem, s.memsize);
    res->content_checksum = h;
    if (res->err == OK && res->layout_ok && h != fi->stored_content_checksum) res->err = E_contentChecksum_invalid;
}

// one hash per (offset, length) range, one thread each
__global__ void __launch_bounds__(128)
xxh32_many_kernel(const uint8_t* __restrict__ data, const uint64_t* __restrict__ offs,
                  const uint32_t* __restrict__ lens, uint32_t n, uint32_t* __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (; i < n; i += gridDim.x * blockDim.x) out[i] = xxh32_thread(data + offs[i], lens[i]);
}

}  // namespace b2
