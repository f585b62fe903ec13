// b2_cuda.h — the one place the sources meet the CUDA toolkit.  The product build (csrc/Makefile, nvcc, sm_100a) takes
// the first branch.  -DB2_EMU (tests only, tools/emu/) compiles the same sources with g++ against a SIMT emulator so
// that kernel logic can be exercised in a container without a GPU; the product never defines it.
#pragma once
#ifdef B2_EMU
#include "../../tools/emu/emu_cuda.h"
#define B2_NOINLINE __attribute__((noinline))
#define B2_PTX 0
struct B2Range { explicit B2Range(const char*) {} };
#else
#include <cuda_runtime.h>
// shared memory is declared through these so that the emulator can give every CTA its own copy
#define B2_SHARED_ARRAY(type, name, count) __shared__ __align__(16) type name[count]
#define B2_SHARED_VAR(type, name) __shared__ type name
#define B2_DYN_SHARED(type, name) extern __shared__ __align__(16) type name[]
#define B2_LAUNCH(kern, grid, block, smem, stream, ...) kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define B2_NOINLINE __noinline__
#define B2_PTX 1
// NVTX ranges around the host-side phases (batch submit = H2D + kernels, batch completion = wait + D2H): visible in Nsight
// Systems / ncu timelines, a no-op without a profiler attached (header-only NVTX v3)
#include <nvtx3/nvToolsExt.h>
struct B2Range { explicit B2Range(const char* name) { nvtxRangePushA(name); } ~B2Range() { nvtxRangePop(); } };
#endif
