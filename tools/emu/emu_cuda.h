// emu_cuda.h — TEST INFRASTRUCTURE ONLY.  A SIMT emulator that lets the .cu sources of lz4-frame-conduit_b200/csrc
// compile with plain g++ (-DB2_EMU -x c++) and run on the host, so that kernel LOGIC (warp collectives, shared-memory
// protocols, inter-warp dataflow waits, forward progress) can be checked in this GPU-less container before GPU time is
// spent.  It is not a CPU codec and not a fallback: the product library (csrc/Makefile) never defines B2_EMU, nothing
// under lz4-frame-conduit_b200/ or bench.py loads the emulated build, and only tests/ (-m "not gpu") build and call it.
//
// Model: every CUDA thread is a fiber (own stack, hand-written context switch); the 32 fibers of a warp run one after the
// other until each blocks at a warp collective (__shfl_sync, __ballot_sync, __syncwarp, ...), at __syncthreads or in
// __nanosleep; when all live lanes of a warp wait at a collective it is resolved and they continue.  Warps and CTAs are
// scheduled round robin on the calling OS thread; at most EMU_RESIDENT_THREADS threads are resident, further CTAs are
// dispatched in blockIdx order as earlier ones retire (like the hardware).  A kernel that waits for work no resident warp
// will ever do hangs the scheduler's watchdog ("no progress") instead of the GPU.
// What it cannot show: memory-model races between SMs, performance.  Lanes of a warp do not run in lockstep between
// collectives (B2_EMU_LANE_ORDER=reverse|random changes their order to shake out missing __syncwarp()s).
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#include <stdio.h>
#include <new>
#include <algorithm>
#include <vector>
#include <functional>
#include <chrono>
#include <future>
#include <mutex>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct __attribute__((aligned(16))) uint4 { uint32_t x, y, z, w; };
struct __attribute__((aligned(8))) uint2 { uint32_t x, y; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { uint4 v; v.x = x; v.y = y; v.z = z; v.w = w; return v; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 v; v.x = x; v.y = y; return v; }
struct dim3 { unsigned x, y, z; dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {} };

namespace emu {
struct Idx { unsigned x, y, z; };
struct Warp;
struct Lane {
    void* sp; int state; int lane; Warp* warp;
    Idx tidx, bidx, bdim, gdim;
    int op; uint32_t a, b, res; const void* site;     // pending collective
    unsigned shared_cursor;
    int stack_slot;
};
extern Lane* cur;
enum Op { OP_SYNCWARP = 1, OP_SHFL, OP_SHFL_UP, OP_SHFL_DOWN, OP_SHFL_XOR, OP_BALLOT, OP_ANY, OP_ALL, OP_RED_OR, OP_RED_ADD,
          OP_RED_MIN, OP_RED_MAX, OP_RED_AND };
uint32_t collective(int op, uint32_t a, uint32_t b);
void syncthreads();
void nanosleep_yield();
void note_progress();
void* shared_alloc(size_t bytes);
void* dyn_shared();
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body);
int sm_count();
int device_count();
}  // namespace emu

#define threadIdx (emu::cur->tidx)
#define blockIdx  (emu::cur->bidx)
#define blockDim  (emu::cur->bdim)
#define gridDim   (emu::cur->gdim)

// ---- shared memory: arrays through macros (per-CTA storage); see csrc/b2_cuda.h
#define B2_SHARED_ARRAY(type, name, count) type* const name = (type*)emu::shared_alloc(sizeof(type) * (size_t)(count))
#define B2_SHARED_VAR(type, name) type& name = *(type*)emu::shared_alloc(sizeof(type))
#define B2_DYN_SHARED(type, name) type* const name = (type*)emu::dyn_shared()
#define B2_LAUNCH(kern, grid, block, smem, stream, ...) emu::launch(dim3(grid), dim3(block), (size_t)(smem), [=]() { kern(__VA_ARGS__); })

// ---- warp collectives
template <typename T> static inline T __shfl_sync(unsigned, T v, int src, int = 32) {
    static_assert(sizeof(T) <= 8, "shfl: at most 64 bits");
    if (sizeof(T) <= 4) { uint32_t x = 0; memcpy(&x, &v, sizeof(T)); x = emu::collective(emu::OP_SHFL, x, (uint32_t)src); T r; memcpy(&r, &x, sizeof(T)); return r; }
    uint64_t x = 0; memcpy(&x, &v, sizeof(T));
    uint32_t lo = emu::collective(emu::OP_SHFL, (uint32_t)x, (uint32_t)src), hi = emu::collective(emu::OP_SHFL, (uint32_t)(x >> 32), (uint32_t)src);
    x = (uint64_t)hi << 32 | lo; T r; memcpy(&r, &x, sizeof(T)); return r;
}
#define EMU_SHFL2(NAME, OP) \
template <typename T> static inline T NAME(unsigned, T v, unsigned d, int = 32) { \
    static_assert(sizeof(T) <= 8, "shfl: at most 64 bits"); \
    if (sizeof(T) <= 4) { uint32_t x = 0; memcpy(&x, &v, sizeof(T)); x = emu::collective(OP, x, d); T r; memcpy(&r, &x, sizeof(T)); return r; } \
    uint64_t x = 0; memcpy(&x, &v, sizeof(T)); \
    uint32_t lo = emu::collective(OP, (uint32_t)x, d), hi = emu::collective(OP, (uint32_t)(x >> 32), d); \
    x = (uint64_t)hi << 32 | lo; T r; memcpy(&r, &x, sizeof(T)); return r; }
EMU_SHFL2(__shfl_up_sync, emu::OP_SHFL_UP)
EMU_SHFL2(__shfl_down_sync, emu::OP_SHFL_DOWN)
EMU_SHFL2(__shfl_xor_sync, emu::OP_SHFL_XOR)
static inline unsigned __ballot_sync(unsigned, int p) { return emu::collective(emu::OP_BALLOT, p ? 1u : 0u, 0); }
static inline int __any_sync(unsigned, int p) { return (int)emu::collective(emu::OP_ANY, p ? 1u : 0u, 0); }
static inline int __all_sync(unsigned, int p) { return (int)emu::collective(emu::OP_ALL, p ? 1u : 0u, 0); }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return emu::collective(emu::OP_RED_OR, v, 0); }
static inline unsigned __reduce_and_sync(unsigned, unsigned v) { return emu::collective(emu::OP_RED_AND, v, 0); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return emu::collective(emu::OP_RED_ADD, v, 0); }
static inline unsigned __reduce_min_sync(unsigned, unsigned v) { return emu::collective(emu::OP_RED_MIN, v, 0); }
static inline unsigned __reduce_max_sync(unsigned, unsigned v) { return emu::collective(emu::OP_RED_MAX, v, 0); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::collective(emu::OP_SYNCWARP, 0, 0); }
static inline void __syncthreads() { emu::syncthreads(); }
static inline void __nanosleep(unsigned) { emu::nanosleep_yield(); }
static inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline void __threadfence_block() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }

// ---- atomics (the scheduler is single-threaded; builtins keep them honest against host threads)
template <typename T> static inline T atomicAdd(T* p, T v) { emu::note_progress(); return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicSub(T* p, T v) { emu::note_progress(); return __atomic_fetch_sub(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicOr(T* p, T v) { emu::note_progress(); return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicAnd(T* p, T v) { emu::note_progress(); return __atomic_fetch_and(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicExch(T* p, T v) { emu::note_progress(); return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicCAS(T* p, T cmp, T v) { emu::note_progress(); __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST); return cmp; }
template <typename T> static inline T atomicMin(T* p, T v) { emu::note_progress(); T o = __atomic_load_n(p, __ATOMIC_SEQ_CST); while (v < o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {} return o; }
template <typename T> static inline T atomicMax(T* p, T v) { emu::note_progress(); T o = __atomic_load_n(p, __ATOMIC_SEQ_CST); while (v > o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {} return o; }

template <typename T> static inline T __ldcg(const T* p) { return *(const volatile T*)p; }
template <typename T> static inline T __ldcv(const T* p) { return *(const volatile T*)p; }
template <typename T> static inline T __ldg(const T* p) { return *p; }
static inline uint4 __ldcg(const uint4* p) { uint4 v; memcpy(&v, p, 16); return v; }
template <typename T> static inline void __stcg(T* p, T v) { *(volatile T*)p = v; }

// ---- bit intrinsics
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; ++i) r |= ((x >> i) & 1u) << (31 - i); return r; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { return (unsigned)((((uint64_t)hi << 32) | lo) >> (sh & 31u)); }
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned sh) { return (unsigned)(((((uint64_t)hi << 32) | lo) << (sh & 31u)) >> 32); }
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned long long min(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
static inline unsigned long long max(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
static inline unsigned long min(unsigned long a, unsigned long b) { return a < b ? a : b; }
static inline unsigned long max(unsigned long a, unsigned long b) { return a > b ? a : b; }

// ---- the slice of the CUDA runtime API engine.cu / lz4f_compat.cu use; everything executes at once, in program order
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2, cudaErrorInvalidDevice = 101 };
typedef struct EmuStream_* cudaStream_t;
struct EmuEvent_ { double t; };
typedef EmuEvent_* cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaMemcpyDefault = 4 };
enum { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaHostAllocPortable = 1, cudaHostAllocMapped = 2, cudaHostAllocWriteCombined = 4, cudaHostRegisterDefault = 0 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
namespace emu { int& current_device(); double now_ms(); void* dev_alloc(size_t n); }
static inline const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : e == cudaErrorInvalidDevice ? "invalid device ordinal (emulator)" : "error (emulator)"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = emu::device_count(); return *n ? cudaSuccess : 100; }
static inline cudaError_t cudaSetDevice(int d) { if (d < 0 || d >= emu::device_count()) return cudaErrorInvalidDevice; emu::current_device() = d; return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int* d) { *d = emu::current_device(); return cudaSuccess; }
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return cudaSuccess; }   // one "device" image
static inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr, int) { *v = emu::sm_count(); return cudaSuccess; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { if (!emu::device_count()) return 100; *s = (cudaStream_t)malloc(8); return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new EmuEvent_{0}; return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = new EmuEvent_{0}; return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->t = emu::now_ms(); return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return cudaSuccess; }
static inline cudaError_t cudaMalloc(void** p, size_t n) { if (!emu::device_count()) return 100; *p = emu::dev_alloc(n); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
template <typename T> static inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc((void**)p, n); }
static inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { if (!emu::device_count()) return 100; *p = malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
template <typename T> static inline cudaError_t cudaMallocHost(T** p, size_t n) { return cudaMallocHost((void**)p, n); }
static inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { return cudaMallocHost(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
static inline cudaError_t cudaHostRegister(void*, size_t, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaHostUnregister(void*) { return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { if (n) memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { if (n) memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { if (n) memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { if (n) memset(d, v, n); return cudaSuccess; }
