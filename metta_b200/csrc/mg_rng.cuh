// mg_rng.cuh -- std::mt19937 and the libstdc++ distributions the reference draws from it,
// restated for one warp per environment.
//
// The 624-word state lives in global memory exactly as std::mt19937 holds it (_M_x, _M_p), so
// the stream is the reference's stream word for word. Draws are made by all 32 lanes together
// (same address, broadcast load) so that the twist, which is needed every 624 words, can be
// done cooperatively in 20 chunks instead of 624 serial steps.
#ifndef MG_RNG_CUH_
#define MG_RNG_CUH_

#include "mg_platform.h"

#define MT_N 624
#define MT_M 397

// std::mersenne_twister_engine::seed (bits/random.tcc:323-337); serial recurrence, one lane.
MG_DEVICE inline void mt_seed_serial(uint32_t* mt, uint32_t seed) {
  uint32_t x = seed;
  mt[0] = x;
  for (uint32_t i = 1; i < MT_N; i++) {
    x = 1812433253u * (x ^ (x >> 30)) + i;
    mt[i] = x;
  }
}

// _M_gen_rand (bits/random.tcc:395-433) in chunks of 32 consecutive words: within a chunk every
// lane reads old[i], old[i+1] and the word 397 ahead before anyone writes; across chunks the
// order of the serial algorithm is kept, so word i+397 (mod 624) is old for i < 227 and already
// regenerated for i >= 227, and word 623 sees the new word 0.
MG_DEVICE inline void mt_twist_warp(uint32_t* mt, int lane) {
  for (int s = 0; s < MT_N; s += 32) {
    int i = s + lane;
    uint32_t v = 0;
    if (i < MT_N) {
      int i1 = i + 1 == MT_N ? 0 : i + 1;
      int im = i + MT_M >= MT_N ? i + MT_M - MT_N : i + MT_M;
      uint32_t y = (mt[i] & 0x80000000u) | (mt[i1] & 0x7fffffffu);
      v = mt[im] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    __syncwarp();
    if (i < MT_N) mt[i] = v;
    __syncwarp();
  }
}

// operator() (bits/random.tcc:435-455). Called by all lanes; every lane gets the same word.
MG_DEVICE inline uint32_t mt_next_warp(uint32_t* mt, uint32_t& idx, int lane) {
  if (idx >= MT_N) {
    mt_twist_warp(mt, lane);
    idx = 0;
  }
  uint32_t z = mt[idx++];
  z ^= (z >> 11);
  z ^= (z << 7) & 0x9d2c5680u;
  z ^= (z << 15) & 0xefc60000u;
  z ^= (z >> 18);
  return z;
}

// uniform_int_distribution<T>(0, range-1), range < 2^32, on a 32-bit engine: Lemire's method
// (bits/uniform_int_dist.h:257-330).
MG_DEVICE inline uint32_t uniform_below_warp(uint32_t* mt, uint32_t& idx, int lane, uint32_t range) {
  uint64_t product = (uint64_t)mt_next_warp(mt, idx, lane) * (uint64_t)range;
  uint32_t low = (uint32_t)product;
  if (low < range) {
    uint32_t threshold = (0u - range) % range;
    while (low < threshold) {
      product = (uint64_t)mt_next_warp(mt, idx, lane) * (uint64_t)range;
      low = (uint32_t)product;
    }
  }
  return (uint32_t)(product >> 32);
}

// std::shuffle over n indices (bits/stl_algo.h:3742-3800, two swap positions per draw).
// Every lane runs the same loop so the draws stay warp-uniform; lane 0 owns `order`.
MG_DEVICE inline void std_shuffle_warp(uint32_t* mt, uint32_t& idx, int lane, uint8_t* order, int n) {
  if (lane == 0)
    for (int i = 0; i < n; i++) order[i] = (uint8_t)i;
  if (n <= 1) return;
  int i = 1;
  if ((n & 1) == 0) {
    uint32_t j = uniform_below_warp(mt, idx, lane, 2u);
    if (lane == 0) { uint8_t t = order[1]; order[1] = order[j]; order[j] = t; }
    i = 2;
  }
  while (i < n) {
    uint32_t s = (uint32_t)i + 1u;
    uint32_t x = uniform_below_warp(mt, idx, lane, s * (s + 1u));
    uint32_t p0 = x / (s + 1u), p1 = x % (s + 1u);
    if (lane == 0) {
      uint8_t t = order[i]; order[i] = order[p0]; order[p0] = t;
      t = order[i + 1]; order[i + 1] = order[p1]; order[p1] = t;
    }
    i += 2;
  }
}

// std::discrete_distribution<size_t>(weights)(g) (bits/random.tcc:2657-2730); generate_canonical
// draws two words (:3349-3382). A single weight returns 0 without drawing. Doubles, as the reference.
MG_DEVICE inline int discrete_pick_warp(uint32_t* mt, uint32_t& idx, int lane, const int* weights, int n) {
  if (n < 2) return 0;
  double sum = 0.0;
  for (int i = 0; i < n; i++) sum += (double)weights[i];
  double s = (double)mt_next_warp(mt, idx, lane);
  s += (double)mt_next_warp(mt, idx, lane) * 4294967296.0;
  double u = s / 18446744073709551616.0;
  if (u >= 1.0) u = 0.99999999999999988897769753748;  // nextafter(1.0, 0.0)
  double acc = 0.0;
  int pos = 0;
  for (; pos < n; pos++) {
    acc += (double)weights[pos] / sum;
    double cp = pos == n - 1 ? 1.0 : acc;
    if (!(cp < u)) break;
  }
  return pos < n ? pos : n - 1;
}

#endif  // MG_RNG_CUH_
