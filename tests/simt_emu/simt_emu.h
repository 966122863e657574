// simt_emu.h -- TEST INFRASTRUCTURE: a tiny SIMT emulator so the CPU test-suite can execute the
// simulator's kernel sources (metta_b200/csrc/*.cuh, compiled with -DMG_SIMT_EMU) without a GPU.
//
// Every CUDA thread of a block is a fiber (ucontext). Fibers run round-robin and only switch at
// __syncwarp / __syncthreads / warp collectives, which are implemented as barriers with an
// exchange buffer, so warp-synchronous code behaves as on hardware as long as collectives are
// reached by all 32 lanes (the kernels are written that way). Blocks run one after another.
// The handful of CUDA runtime calls the C-ABI makes are mapped onto malloc/memcpy.
//
// This is a debugging aid for kernel logic. It is never compiled into libmettagrid_b200.so and
// the metta_b200 package cannot load it.
#ifndef SIMT_EMU_H_
#define SIMT_EMU_H_

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <functional>
#include <vector>

#define MG_DEVICE
#define MG_GLOBAL
#define MG_NOINLINE __attribute__((noinline))
#define MG_SHARED static
#define MG_DYN_SMEM(name) unsigned char* name = simt_emu::dyn_smem()
#define MG_LAUNCH_BOUNDS(t, b)
#define MG_PREFETCH_L1(p) ((void)(p))
#define MG_PREFETCH_L2(p) ((void)(p))
#define MG_RED_ADD(p, v) ((void)(*(p) += (v)))
#define MG_LDCG(p) (*(p))
#define MG_ASYNC_COPY4(smem_dst, gmem_src) (*(smem_dst) = *(gmem_src))
#define MG_ASYNC_COPY_WAIT() ((void)0)
#define MG_STAT_BUMP(p, v) ((void)(*(p) += (v)))
#define MG_PREFETCH_TO_L2(p) ((void)(p))
#define MG_LD_L2(p) (*(p))
#define MG_NANOSLEEP(ns) ((void)0)
#define MG_ROLLED
#define MG_LDG(p) (*(p))
#define MG_GRID_CONSTANT
#define MG_FADD(a, b) ((float)(a) + (float)(b))
#define MG_FMUL(a, b) ((float)(a) * (float)(b))
#define MG_LAUNCH(kernel, grid, block, smem, stream, ...) \
  simt_emu::launch((grid), (block), (smem), [=]() { kernel(__VA_ARGS__); })

static inline float MG_INT_AS_FLOAT(int32_t x) { float f; memcpy(&f, &x, 4); return f; }

struct uint4 { uint32_t x, y, z, w; } __attribute__((aligned(16)));
struct uint2 { uint32_t x, y; } __attribute__((aligned(8)));
struct int4 { int32_t x, y, z, w; } __attribute__((aligned(16)));
struct emu_dim3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};

namespace simt_emu {
struct Fiber {
  ucontext_t ctx;
  char* stack;
  bool done;
  int site;  // call site of the collective this lane is in (see MG_EMU_SITE)
};
struct WarpState {
  int arrived, alive;
  unsigned gen;
  uint64_t xchg[32];
  int site, phase;  // call site of the first arrival at the barrier being filled; sub-barrier count within one collective
};
struct State {
  ucontext_t sched;
  std::vector<Fiber> fibers;
  std::vector<WarpState> warps;
  int cur;
  int block_arrived, block_alive;
  unsigned block_gen;
  unsigned char* dyn;
  std::function<void()> body;
  bool progressed;
};
State& st();
void launch(dim3 grid, int block, size_t smem, std::function<void()> body);
void yield();
unsigned char* dyn_smem();
void warp_barrier();
void block_barrier();
}  // namespace simt_emu

extern emu_dim3 threadIdx, blockIdx, blockDim, gridDim;

static inline void emu__syncwarp(unsigned = 0xFFFFFFFFu) { simt_emu::warp_barrier(); }
static inline void __syncthreads() { simt_emu::block_barrier(); }

template <typename T>
static inline T emu_exchange(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "emu_exchange: type too wide");
  simt_emu::State& s = simt_emu::st();
  simt_emu::WarpState& w = s.warps[s.cur / 32];
  uint64_t raw = 0;
  memcpy(&raw, &v, sizeof(T));
  w.xchg[s.cur % 32] = raw;
  simt_emu::warp_barrier();
  T out = v;
  if (src_lane >= 0 && src_lane < 32) memcpy(&out, &w.xchg[src_lane], sizeof(T));
  simt_emu::warp_barrier();
  return out;
}
// width: the warp is split into segments of `width` lanes (a power of two), as the hardware does
template <typename T>
static inline T emu__shfl_sync(unsigned, T v, int src, int width = 32) {
  int lane = (int)(threadIdx.x & 31);
  return emu_exchange(v, (lane & ~(width - 1)) | (src & (width - 1)));
}
template <typename T>
static inline T emu__shfl_up_sync(unsigned, T v, unsigned d, int width = 32) {
  int lane = (int)(threadIdx.x & 31), in_seg = lane & (width - 1);
  return emu_exchange(v, in_seg - (int)d >= 0 ? lane - (int)d : lane);
}
template <typename T>
static inline T emu__shfl_down_sync(unsigned, T v, unsigned d, int width = 32) {
  int lane = (int)(threadIdx.x & 31), in_seg = lane & (width - 1);
  return emu_exchange(v, in_seg + (int)d < width ? lane + (int)d : lane);
}
template <typename T>
static inline T emu__shfl_xor_sync(unsigned, T v, int m, int width = 32) {
  int lane = (int)(threadIdx.x & 31), target = lane ^ m;
  return emu_exchange(v, (target & ~(width - 1)) == (lane & ~(width - 1)) ? target : lane);
}
static inline unsigned emu__ballot_sync(unsigned, int pred) {
  simt_emu::State& s = simt_emu::st();
  simt_emu::WarpState& w = s.warps[s.cur / 32];
  w.xchg[s.cur % 32] = pred ? 1 : 0;
  simt_emu::warp_barrier();
  unsigned m = 0;
  int nl = (int)blockDim.x - (s.cur / 32) * 32;
  if (nl > 32) nl = 32;
  for (int i = 0; i < nl; i++)
    if (w.xchg[i]) m |= 1u << i;
  simt_emu::warp_barrier();
  return m;
}
static inline unsigned emu__match_any_sync(unsigned, int v) {
  simt_emu::State& s = simt_emu::st();
  simt_emu::WarpState& w = s.warps[s.cur / 32];
  w.xchg[s.cur % 32] = (uint64_t)(uint32_t)v;
  simt_emu::warp_barrier();
  unsigned m = 0;
  int nl = (int)blockDim.x - (s.cur / 32) * 32;
  if (nl > 32) nl = 32;
  for (int i = 0; i < nl; i++)
    if (w.xchg[i] == (uint64_t)(uint32_t)v) m |= 1u << i;
  simt_emu::warp_barrier();
  return m;
}
static inline unsigned atomicOr(unsigned* p, unsigned v) { unsigned o = *p; *p |= v; return o; }
static inline int emu_warp_sum(int v) { for (int o = 16; o > 0; o >>= 1) v += emu__shfl_xor_sync(0xFFFFFFFFu, v, o); return v; }
#define MG_WARP_SUM(v) emu_warp_sum((int)(v))
static inline int atomicMax(int* p, int v) { int o = *p; if (v > o) *p = v; return o; }
static inline int emu__any_sync(unsigned m, int p) { return emu__ballot_sync(m, p) != 0; }
static inline long long clock64() { return 0; }
static inline void __threadfence() {}
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline double atomicAdd(double* p, double v) { double o = *p; *p += v; return o; }
static inline float atomicAdd(float* p, float v) { float o = *p; *p += v; return o; }
static inline int atomicAdd(int* p, int v) { int o = *p; *p += v; return o; }
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p += v; return o; }

// ---- the few CUDA runtime entry points used by mg_abi.cu
typedef int cudaError_t;
typedef void* cudaStream_t;
enum { cudaSuccess = 0 };
enum { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
enum { cudaStreamNonBlocking = 1 };
static inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = aligned_alloc(256, (n + 255) & ~(size_t)255); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { free(p); return 0; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
static inline cudaError_t cudaMemset(void* p, int v, size_t n) { memset(p, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, void*) { memset(p, v, n); return 0; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, int) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, int, cudaStream_t) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, int) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaGetLastError() { return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
#define __device__
#define __align__(n) __attribute__((aligned(n)))
template <typename T>
static inline cudaError_t cudaGetSymbolAddress(void** p, const T& symbol) { *p = (void*)&symbol; return 0; }
typedef void* cudaEvent_t;
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = nullptr; return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.0f; return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return 0; }
enum { cudaLimitStackSize = 0 };
enum { cudaEventDisableTiming = 2 };
typedef void* cudaGraph_t;
typedef void* cudaGraphExec_t;
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, int) { *e = nullptr; return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, int) { return 0; }
static inline cudaError_t cudaGraphExecDestroy(cudaGraphExec_t) { return 0; }
static inline cudaError_t cudaGraphDestroy(cudaGraph_t) { return 0; }
static inline cudaError_t cudaDeviceSetLimit(int, size_t) { return 0; }

// Every collective records where it was called from; a barrier that completes with lanes arriving from
// different call sites means the kernel's lanes took different numbers of collectives (a bug that real
// hardware would turn into a hang or garbage): the emulator aborts with both sites.
namespace simt_emu { void set_site(int site); }
#define MG_EMU_SITE(call) (simt_emu::set_site(__COUNTER__ * 100000 + __LINE__), call)
#define __syncwarp(...) MG_EMU_SITE(emu__syncwarp(__VA_ARGS__))
#define __shfl_sync(...) MG_EMU_SITE(emu__shfl_sync(__VA_ARGS__))
#define __shfl_up_sync(...) MG_EMU_SITE(emu__shfl_up_sync(__VA_ARGS__))
#define __shfl_down_sync(...) MG_EMU_SITE(emu__shfl_down_sync(__VA_ARGS__))
#define __shfl_xor_sync(...) MG_EMU_SITE(emu__shfl_xor_sync(__VA_ARGS__))
#define __ballot_sync(...) MG_EMU_SITE(emu__ballot_sync(__VA_ARGS__))
#define __match_any_sync(...) MG_EMU_SITE(emu__match_any_sync(__VA_ARGS__))
#define __any_sync(...) MG_EMU_SITE(emu__any_sync(__VA_ARGS__))
#undef MG_WARP_SUM
#define MG_WARP_SUM(v) MG_EMU_SITE(emu_warp_sum((int)(v)))

#endif  // SIMT_EMU_H_
