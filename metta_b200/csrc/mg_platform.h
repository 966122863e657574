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

#if !defined(MG_SIMT_EMU) && !defined(MG_NO_BULK_COPY)
// ---- 1-D bulk copies (TMA, cp.async.bulk) between this GPU's HBM and shared memory, completion on an mbarrier.
// One thread issues a copy of any multiple of 16 bytes; the data never passes through registers.
#define MG_HAS_BULK_COPY 1
#ifndef MG_NO_BULK_LOAD
#define MG_HAS_BULK_LOAD 1 /* the staging of records at the start of a step; the write-back uses bulk copies either way */
#endif
__device__ __forceinline__ uint32_t mg_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mg_mbar_init(uint64_t* bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mg_smem_addr(bar)), "r"(arrivals) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // the async proxy must see the initialised barrier
}
__device__ __forceinline__ void mg_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mg_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mg_bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(mg_smem_addr(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(mg_smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mg_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MG_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MG_DONE_%=;\n"
      "bra MG_WAIT_%=;\n"
      "MG_DONE_%=:\n"
      "}\n" ::"r"(mg_smem_addr(bar)),
      "r"(parity)
      : "memory");
}
// shared -> global; the caller has made its shared-memory writes visible to the async proxy (mg_bulk_store_fence)
__device__ __forceinline__ void mg_bulk_store_fence() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mg_bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(mg_smem_addr(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mg_bulk_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void mg_bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }  // shared memory may be reused
__device__ __forceinline__ void mg_bulk_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }        // the writes have been performed
#endif

#ifndef MG_SIMT_EMU
// A 4-byte global -> shared copy that completes in the background (LDGSTS); MG_ASYNC_COPY_WAIT before the first read.
#define MG_ASYNC_COPY4(smem_dst, gmem_src) \
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory")
#define MG_ASYNC_COPY_WAIT() asm volatile("cp.async.wait_all;" ::: "memory")
#endif

#define MG_FULL_MASK 0xFFFFFFFFu

#endif  // MG_PLATFORM_H_
