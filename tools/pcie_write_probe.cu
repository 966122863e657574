// How fast can SMs write scattered pieces into page-locked host memory? (tuning probe for the host export
// of mg_step_host; build: nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/pcie_write_probe tools/pcie_write_probe.cu)
// Every variant moves the same `pieces` x `piece_bytes` from a device buffer to a mapped host buffer of
// 60 MB at piece-aligned pseudo-random offsets (or contiguously), timed with CUDA events.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// lanes_per_piece lanes write one piece of PIECE bytes with VEC-byte stores each
template <int PIECE, int VEC>
__global__ void scatter_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t n_pieces, uint32_t n_slots, int contiguous) {
  constexpr int LPP = PIECE / VEC;  // lanes per piece
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t piece = t / LPP, sub = t % LPP;
  if (piece >= n_pieces) return;
  const uint32_t slot = contiguous ? piece : mix(piece) % n_slots;
  const size_t off = (size_t)slot * PIECE + (size_t)sub * VEC;
  if (VEC == 8) *reinterpret_cast<uint2*>(dst + off) = *reinterpret_cast<const uint2*>(src + off);
  else if (VEC == 16) *reinterpret_cast<uint4*>(dst + off) = *reinterpret_cast<const uint4*>(src + off);
  else *reinterpret_cast<uint32_t*>(dst + off) = *reinterpret_cast<const uint32_t*>(src + off);
}

// one warp stages 32 pieces' worth... simpler: one warp per piece group: lane 0 issues a bulk store of PIECE bytes
// from shared memory (filled by the warp) to the host address
template <int PIECE>
__global__ void bulk_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t n_pieces, uint32_t n_slots, int contiguous) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t piece = blockIdx.x * (blockDim.x >> 5) + warp;
  if (piece >= n_pieces) return;
  const uint32_t slot = contiguous ? piece : mix(piece) % n_slots;
  const size_t off = (size_t)slot * PIECE;
  uint8_t* stage = smem + (size_t)warp * PIECE;
  for (int i = lane * 16; i < PIECE; i += 32 * 16) *reinterpret_cast<uint4*>(stage + i) = *reinterpret_cast<const uint4*>(src + off + i);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(stage);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + off), "r"(s), "r"((uint32_t)PIECE) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

template <typename F>
static float time_it(F launch, int reps) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  launch(); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; i++) launch();
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms; CK(cudaEventElapsedTime(&ms, a, b));
  return ms / reps;
}

int main() {
  const size_t buf = 60u << 20, total = 12u << 20;
  uint8_t *d_src, *h_dst, *h_dev;
  CK(cudaMalloc(&d_src, buf));
  CK(cudaMemset(d_src, 1, buf));
  CK(cudaHostAlloc(&h_dst, buf, cudaHostAllocMapped));
  CK(cudaHostGetDevicePointer((void**)&h_dev, h_dst, 0));
  const int reps = 20;
  {
    float ms = time_it([&] { CK(cudaMemcpyAsync(h_dst, d_src, total, cudaMemcpyDeviceToHost, 0)); }, reps);
    printf("%-44s %7.3f ms %6.1f GB/s\n", "copy engine, 12 MB contiguous", ms, total / ms / 1e6);
  }
#define RUN_SCATTER(PIECE, VEC, CONTIG) { \
    const uint32_t n_pieces = (uint32_t)(total / PIECE), n_slots = (uint32_t)(buf / PIECE); \
    const uint32_t threads = n_pieces * (PIECE / VEC); \
    float ms = time_it([&] { scatter_kernel<PIECE, VEC><<<(threads + 255) / 256, 256>>>(d_src, h_dev, n_pieces, n_slots, CONTIG); }, reps); \
    char name[96]; snprintf(name, sizeof name, "SM stores, %4d-byte pieces, %2d B/lane, %s", PIECE, VEC, CONTIG ? "contiguous" : "scattered"); \
    printf("%-44s %7.3f ms %6.1f GB/s\n", name, ms, total / ms / 1e6); }
  RUN_SCATTER(32, 8, 0) RUN_SCATTER(64, 8, 0) RUN_SCATTER(64, 16, 0) RUN_SCATTER(128, 8, 0) RUN_SCATTER(128, 16, 0)
  RUN_SCATTER(256, 16, 0) RUN_SCATTER(512, 16, 0) RUN_SCATTER(64, 8, 1) RUN_SCATTER(128, 16, 1) RUN_SCATTER(256, 16, 1)
#define RUN_BULK(PIECE, CONTIG) { \
    const uint32_t n_pieces = (uint32_t)(total / PIECE), n_slots = (uint32_t)(buf / PIECE); \
    float ms = time_it([&] { bulk_kernel<PIECE><<<(n_pieces + 7) / 8, 256, 8 * PIECE>>>(d_src, h_dev, n_pieces, n_slots, CONTIG); }, reps); \
    char name[96]; snprintf(name, sizeof name, "bulk store (TMA), %4d-byte pieces, %s", PIECE, CONTIG ? "contiguous" : "scattered"); \
    printf("%-44s %7.3f ms %6.1f GB/s\n", name, ms, total / ms / 1e6); }
  RUN_BULK(64, 0) RUN_BULK(128, 0) RUN_BULK(256, 0) RUN_BULK(512, 0) RUN_BULK(1024, 0) RUN_BULK(256, 1)
  return 0;
}
