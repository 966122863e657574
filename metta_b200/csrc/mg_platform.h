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
// Stat counters in HBM. Measured both ways on B200 (arena, 4096 envs): bumping them with reductions
// (no load, no stall) and reading the few the rewards need at L2 (__ldcg, past any stale L1 line) is
// SLOWER than plain read-modify-write on lines the load phase prefetched into L1 (step kernel 123.5 us
// vs 119 us): the reward reads are on the critical path, the bumps are not. -DMG_STAT_RED selects the
// reduction variant.
#ifdef MG_STAT_RED
#define MG_RED_ADD(p, v) ((void)atomicAdd((p), (v)))
#define MG_LDCG(p) __ldcg(p)
#else
#define MG_RED_ADD(p, v) ((void)(*(p) += (v)))
#define MG_LDCG(p) (*(p))
#undef MG_PREFETCH_L2
#define MG_PREFETCH_L2(p) MG_PREFETCH_L1(p)
#endif
#define MG_LD_L2(p) __ldcg(p) /* read at L2: for words other lanes may have bumped with atomics */
#define MG_NANOSLEEP(ns) __nanosleep(ns)
// A counter in HBM that nothing on the step's critical path reads back through L1: bumped with a reduction
// (no load, no stall on a cold line). Readers use MG_LD_L2.
#define MG_STAT_BUMP(p, v) ((void)atomicAdd((p), (v)))
#define MG_PREFETCH_TO_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#define MG_WARP_SUM(v) ((int)__reduce_add_sync(0xFFFFFFFFu, (int)(v))) /* one redux.sync instead of five shuffle rounds */
#define MG_ROLLED _Pragma("unroll 1")
#define MG_GRID_CONSTANT __grid_constant__ /* the parameter's address may be taken without a per-thread copy */
#endif

#define MG_FULL_MASK 0xFFFFFFFFu

#endif  // MG_PLATFORM_H_
