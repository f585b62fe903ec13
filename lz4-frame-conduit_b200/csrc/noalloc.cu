// noalloc.cu — b2lz4_xxh32_ranges_static: block checksums of many byte ranges of a device buffer by an entry point that
// ALLOCATES NOTHING (no cudaMalloc / cudaFree / cudaMallocHost, no context, no stream or event creation).  It has no
// counterpart in the reference; it was added for callers that cannot allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The arithmetic is the frame layer's
// per-block XXH32 (lz4frame.c:969-972 on the write side, 1590-1600 on the read side; xxhash.c:350-416), one thread per
// range like xxh32_many_kernel (xxh32.cu) and verify_block_checksums_kernel (decode.cu).
//
// Those kernels take their (offset, length) descriptors and their results from cudaMalloc'ed workspaces of a context.
// Here the workspace is a file-scope __device__ array of fixed size: B2LZ4_STATIC_MAX_RANGES descriptors + results + one
// counter, 1 MiB + 256 B, which every GPU that loads the library holds until the process exits.  In HOST code the name
// of that array is a stand-in, not a device address; the address comes from cudaGetSymbolAddress, asked once per device
// with that device current.
//
// Every call on a device shares the one array (it keeps the last call's contents; the counter is cleared by every call).
// Calls of one process are serialised by a mutex and each call waits for its own work, so they cannot overlap; work that
// a caller queues by other means against the same array would race.
#include "common.cuh"
#include "kernels.h"
#include "../../include/b200lz4.h"
#include <mutex>
#include <cstdio>

namespace b2 {

constexpr uint32_t NA_MAX = B2LZ4_STATIC_MAX_RANGES;
// layout of the scratch array
constexpr size_t NA_OFFS = 0;                                          // uint64_t[NA_MAX]   range offsets (from the caller)
constexpr size_t NA_LENS = NA_OFFS + (size_t)NA_MAX * 8;               // uint32_t[NA_MAX]   range lengths (from the caller)
constexpr size_t NA_OUT  = NA_LENS + (size_t)NA_MAX * 4;               // uint32_t[NA_MAX]   hashes
constexpr size_t NA_BAD  = NA_OUT + (size_t)NA_MAX * 4;                // uint32_t           ranges that leave the buffer (must start at 0)
constexpr size_t NA_BYTES = NA_BAD + 256;

static __device__ __align__(256) unsigned char g_static_scratch[NA_BYTES];

// variant of xxh32_many_kernel: descriptors and results live in the static scratch, ranges are checked against the buffer
__global__ void __launch_bounds__(128)
xxh32_ranges_static_kernel(const uint8_t* __restrict__ data, uint64_t n_src, const uint64_t* __restrict__ offs,
                           const uint32_t* __restrict__ lens, uint32_t n, uint32_t* __restrict__ out, uint32_t* __restrict__ n_bad)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t o = offs[i], l = lens[i];
        if (o > n_src || l > n_src - o) { atomicAdd(n_bad, 1u); out[i] = 0u; continue; }
        out[i] = xxh32_thread(data + o, l);
    }
}

namespace {
constexpr int NA_MAX_DEVICES = 64;
struct NaDevice { std::mutex mu; unsigned char* scratch = nullptr; int sm_count = 0; };
NaDevice g_na[NA_MAX_DEVICES];
thread_local char g_na_msg[256] = "";

size_t na_fail(uint32_t code, const char* fmt, unsigned long long a = 0, unsigned long long b = 0) {
    snprintf(g_na_msg, sizeof g_na_msg, fmt, a, b);
    return (size_t)0 - (size_t)code;
}
size_t na_cuda(cudaError_t e, const char* what) {
    snprintf(g_na_msg, sizeof g_na_msg, "%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return (size_t)0 - (size_t)B2LZ4_ERROR_cudaFailure;
}
struct DeviceGuard {             // the call leaves the caller's current device as it found it
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
}  // namespace
}  // namespace b2

using namespace b2;

extern "C" {

const char* b2lz4_static_last_error(void) { return g_na_msg; }

size_t b2lz4_xxh32_ranges_static(int device, void* stream, const void* d_src, size_t n_src,
                                 const uint64_t* offs, const uint32_t* lens, uint32_t n_ranges, uint32_t* out)
{
    g_na_msg[0] = 0;
    if (n_ranges > NA_MAX)
        return na_fail(E_srcSize_tooLarge, "b2lz4_xxh32_ranges_static: %llu ranges exceed B2LZ4_STATIC_MAX_RANGES = %llu",
                       n_ranges, NA_MAX);
    if (device < 0 || device >= NA_MAX_DEVICES)
        return na_fail(E_GENERIC, "b2lz4_xxh32_ranges_static: device %llu is outside 0..%llu", (unsigned)device, NA_MAX_DEVICES - 1);
    if (n_ranges && (!offs || !lens || !out)) return na_fail(E_srcPtr_wrong, "b2lz4_xxh32_ranges_static: NULL offs / lens / out");
    if (!d_src && n_src) return na_fail(E_srcPtr_wrong, "b2lz4_xxh32_ranges_static: NULL d_src");
    for (uint32_t i = 0; i < n_ranges; ++i)
        if (offs[i] > n_src || lens[i] > n_src - offs[i])
            return na_fail(E_srcSize_tooLarge, "b2lz4_xxh32_ranges_static: range %llu leaves the buffer of %llu bytes", i, n_src);
    if (!n_ranges) return 0;

    DeviceGuard guard;
    cudaError_t e = cudaGetDevice(&guard.prev);
    if (e != cudaSuccess) { guard.prev = -1; return na_cuda(e, "cudaGetDevice"); }
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_cuda(e, "cudaSetDevice");

    NaDevice& D = g_na[device];
    std::lock_guard<std::mutex> lock(D.mu);
    if (!D.scratch) {            // set-up, once per device, with that device current: the DEVICE address of the array
        unsigned char* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
        if (e != cudaSuccess || !p) return na_cuda(e, "cudaGetSymbolAddress(g_static_scratch)");
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) return na_cuda(e, "cudaDeviceGetAttribute");
        D.sm_count = sms > 0 ? sms : 1;
        D.scratch = p;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t* d_offs = (uint64_t*)(D.scratch + NA_OFFS);
    uint32_t* d_lens = (uint32_t*)(D.scratch + NA_LENS);
    uint32_t* d_out  = (uint32_t*)(D.scratch + NA_OUT);
    uint32_t* d_bad  = (uint32_t*)(D.scratch + NA_BAD);
    if ((e = cudaMemsetAsync(d_bad, 0, sizeof(uint32_t), st)) != cudaSuccess) return na_cuda(e, "cudaMemsetAsync");
    if ((e = cudaMemcpyAsync(d_offs, offs, (size_t)n_ranges * 8, cudaMemcpyHostToDevice, st)) != cudaSuccess) return na_cuda(e, "cudaMemcpyAsync(offs)");
    if ((e = cudaMemcpyAsync(d_lens, lens, (size_t)n_ranges * 4, cudaMemcpyHostToDevice, st)) != cudaSuccess) return na_cuda(e, "cudaMemcpyAsync(lens)");
    const uint32_t grid = std::max(1u, std::min((n_ranges + 127u) / 128u, (uint32_t)D.sm_count * 16u));
    B2_LAUNCH(xxh32_ranges_static_kernel, grid, 128, 0, st, (const uint8_t*)d_src, (uint64_t)n_src,
              (const uint64_t*)d_offs, (const uint32_t*)d_lens, n_ranges, d_out, d_bad);
    if ((e = cudaGetLastError()) != cudaSuccess) return na_cuda(e, "xxh32_ranges_static_kernel launch");
    uint32_t bad = 0;
    if ((e = cudaMemcpyAsync(out, d_out, (size_t)n_ranges * 4, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return na_cuda(e, "cudaMemcpyAsync(out)");
    if ((e = cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return na_cuda(e, "cudaMemcpyAsync(bad)");
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return na_cuda(e, "cudaStreamSynchronize");
    if (bad) return na_fail(E_srcSize_tooLarge, "b2lz4_xxh32_ranges_static: %llu ranges left the buffer (device check)", bad);
    return 0;
}

}  // extern "C"
