// mg_platform.h -- CUDA qualifiers and the handful of intrinsics the simulator uses.
//
// The product is compiled by nvcc for sm_100a only. Defining MG_SIMT_EMU swaps in
// tests/simt_emu/simt_emu.h, a fiber-based SIMT emulator used by the CPU test-suite to execute
// these same kernel sources without a GPU (debugging aid; it is never built into, loaded by or
// reachable from the metta_b200 package).
#ifndef MG_PLATFORM_H_
#define MG_PLATFORM_H_

#ifdef MG_SIMT_EMU
#include "simt_emu.h"
#else
#include <cuda_runtime.h>
#define MG_DEVICE __device__
#define MG_GLOBAL __global__
#define MG_NOINLINE __noinline__
#define MG_SHARED __shared__
#define MG_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#define MG_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define MG_FADD(a, b) __fadd_rn((a), (b))
#define MG_FMUL(a, b) __fmul_rn((a), (b))
#define MG_INT_AS_FLOAT(x) __int_as_float(x)
#define MG_LAUNCH_BOUNDS(t, b) __launch_bounds__(t, b)
#define MG_PREFETCH_L1(p) asm volatile("prefetch.global.L1 [%0];" ::"l"(p))
#define MG_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#define MG_LDG(p) __ldg(p)
// Counters in HBM that nobody needs back within the step are bumped with a reduction (no load, no
// stall); code that reads such a slot later in the same kernel reads it at L2, past any stale L1 line.
#define MG_RED_ADD(p, v) ((void)atomicAdd((p), (v)))
#define MG_LDCG(p) __ldcg(p)
#define MG_NANOSLEEP(ns) __nanosleep(ns)
#define MG_GRID_CONSTANT __grid_constant__ /* the parameter's address may be taken without a per-thread copy */
#endif

#define MG_FULL_MASK 0xFFFFFFFFu

#endif  // MG_PLATFORM_H_
