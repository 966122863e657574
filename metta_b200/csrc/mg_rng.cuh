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

// _M_gen_rand (bits/random.tcc:395-433), cooperatively: every lane reads old[i], old[i+1] and the word 397
// ahead before anyone writes; the order of the serial algorithm is kept where it matters, so word i+397
// (mod 624) is old for i < 227 and already regenerated for i >= 227, and word 623 sees the new word 0.
MG_DEVICE MG_NOINLINE void mt_twist_warp(uint32_t* mt, int lane) {  // once per 624 words: kept out of line, out of the hot code
  // The whole state in registers: word i lives in x[i / 32] of lane i % 32 (twenty independent loads, one
  // memory round trip). Word i needs word i+1 (the next lane) and word i+397 (old, thirteen lanes on and twelve
  // registers up) for i < 227, or the regenerated word i-227 (three lanes back, seven registers down) after
  // that; registers are regenerated in ascending order, so both are in the state the serial loop would see.
  uint32_t x[20];
#pragma unroll
  for (int k = 0; k < 20; k++) {
    const int i = lane + 32 * k;
    x[k] = i < MT_N ? mt[i] : 0u;
  }
#pragma unroll
  for (int k = 0; k < 20; k++) {
    uint32_t next = __shfl_down_sync(MG_FULL_MASK, x[k], 1);
    if (k < 19) {
      const uint32_t wrap = __shfl_sync(MG_FULL_MASK, x[k + 1], 0);
      if (lane == 31) next = wrap;
    } else {
      const uint32_t first = __shfl_sync(MG_FULL_MASK, x[0], 0);  // word 623 pairs with the NEW word 0
      if (lane == 15) next = first;
    }
    uint32_t far = 0u;
    if (k <= 7) {  // old word i + 397
      const uint32_t a = __shfl_sync(MG_FULL_MASK, x[k + 12], (lane + 13) & 31);
      uint32_t b = 0u;
      if (k + 13 < 20) b = __shfl_sync(MG_FULL_MASK, x[k + 13], (lane + 13) & 31);
      far = lane + 13 < 32 ? a : b;
    }
    if (k >= 7) {  // new word i - 227
      const uint32_t c = __shfl_sync(MG_FULL_MASK, x[k - 7], (lane - 3) & 31);
      uint32_t d = 0u;
      if (k >= 8) d = __shfl_sync(MG_FULL_MASK, x[k - 8], (lane + 29) & 31);
      const uint32_t regenerated = lane >= 3 ? c : d;
      if (k > 7 || lane >= 3) far = regenerated;
    }
    const uint32_t y = (x[k] & 0x80000000u) | (next & 0x7fffffffu);
    x[k] = far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
  }
#pragma unroll
  for (int k = 0; k < 20; k++) {
    const int i = lane + 32 * k;
    if (i < MT_N) mt[i] = x[k];
  }
  __syncwarp();
}

// A warp's view of one env's generator. `buf` optionally holds the next untempered words
// (state[buf_base .. buf_base + buf_n)), fetched by all lanes in one go so that a run of draws
// does not pay one HBM round trip each.
struct MtStream {
  uint32_t* mt;
  uint32_t idx;  // std::mt19937::_M_p
  const uint32_t* buf;
  uint32_t buf_base, buf_n;
};

MG_DEVICE inline MtStream mt_stream(uint32_t* mt, uint32_t idx) {
  MtStream s;
  s.mt = mt; s.idx = idx; s.buf = nullptr; s.buf_base = 0; s.buf_n = 0;
  return s;
}

// Regenerate if the state is exhausted, then stage the next (up to 32) words in `buf` (shared memory).
MG_DEVICE inline void mt_prefetch_warp(MtStream& s, uint32_t* buf, int lane) {
  if (s.idx >= MT_N) {
    mt_twist_warp(s.mt, lane);
    s.idx = 0;
  }
  const uint32_t n = MT_N - s.idx < 32u ? MT_N - s.idx : 32u;
  if ((uint32_t)lane < n) buf[lane] = s.mt[s.idx + lane];
  __syncwarp();
  s.buf = buf; s.buf_base = s.idx; s.buf_n = n;
}

// operator() (bits/random.tcc:435-455). Called by all lanes; every lane gets the same word.
MG_DEVICE inline uint32_t mt_next_warp(MtStream& s, int lane) {
  if (s.idx >= MT_N) {
    mt_twist_warp(s.mt, lane);
    s.idx = 0;
    s.buf_n = 0;
  }
  const uint32_t k = s.idx - s.buf_base;
  uint32_t z = (s.buf_n != 0u && k < s.buf_n) ? s.buf[k] : s.mt[s.idx];
  s.idx++;
  z ^= (z >> 11);
  z ^= (z << 7) & 0x9d2c5680u;
  z ^= (z << 15) & 0xefc60000u;
  z ^= (z >> 18);
  return z;
}

// uniform_int_distribution<T>(0, range-1), range < 2^32, on a 32-bit engine: Lemire's method
// (bits/uniform_int_dist.h:257-330).
MG_DEVICE inline uint32_t uniform_below_warp(MtStream& s, int lane, uint32_t range) {
  uint64_t product = (uint64_t)mt_next_warp(s, lane) * (uint64_t)range;
  uint32_t low = (uint32_t)product;
  if (low < range) {
    uint32_t threshold = (0u - range) % range;
    while (low < threshold) {
      product = (uint64_t)mt_next_warp(s, lane) * (uint64_t)range;
      low = (uint32_t)product;
    }
  }
  return (uint32_t)(product >> 32);
}

// std::shuffle over n indices (bits/stl_algo.h:3742-3800, two swap positions per draw).
// Every lane runs the same loop so the draws stay warp-uniform; lane 0 owns `order`.
MG_DEVICE MG_NOINLINE void std_shuffle_warp(MtStream& s, int lane, uint8_t* order, int n) {  // the fallback path
  if (lane == 0)
    for (int i = 0; i < n; i++) order[i] = (uint8_t)i;
  if (n <= 1) return;
  int i = 1;
  if ((n & 1) == 0) {
    uint32_t j = uniform_below_warp(s, lane, 2u);
    if (lane == 0) { uint8_t t = order[1]; order[1] = order[j]; order[j] = t; }
    i = 2;
  }
  while (i < n) {
    uint32_t sr = (uint32_t)i + 1u;
    uint32_t x = uniform_below_warp(s, lane, sr * (sr + 1u));
    uint32_t p0 = x / (sr + 1u), p1 = x % (sr + 1u);
    if (lane == 0) {
      uint8_t t = order[i]; order[i] = order[p0]; order[p0] = t;
      t = order[i + 1]; order[i + 1] = order[p1]; order[p1] = t;
    }
    i += 2;
  }
}

// The same shuffle in two halves, because the order only matters to the agents that go through the
// ordered pass, while the draws always have to be taken. Draw d only depends on word d of the stream
// (unless Lemire's method rejects a word, about once in 10^7 draws), so lane d tempers word d and
// derives its pair of swap positions into `scratch` (shared memory, 32 words). Returns false when the
// staged words do not cover the draws or a word was rejected: then nothing has been consumed and the
// caller runs std_shuffle_warp.
MG_DEVICE inline bool std_shuffle_draws_warp(MtStream& s, int lane, int n, uint32_t* scratch) {
  const int nd = n >> 1;  // draws: n even: 1 + (n - 2) / 2, n odd: (n - 1) / 2
  if (n <= 1) return true;
  const uint32_t first = s.idx - s.buf_base;
  if (nd > 32 || s.buf_n == 0u || s.idx < s.buf_base || first + (uint32_t)nd > s.buf_n) return false;
  bool reject = false;
  if (lane < nd) {
    const bool even = (n & 1) == 0;
    const uint32_t i = even ? 2u * (uint32_t)lane : 2u * (uint32_t)lane + 1u;  // index of the first element this draw places
    const uint32_t sr = i + 1u;
    const uint32_t range = (even && lane == 0) ? 2u : sr * (sr + 1u);
    uint32_t z = s.buf[first + lane];
    z ^= (z >> 11);
    z ^= (z << 7) & 0x9d2c5680u;
    z ^= (z << 15) & 0xefc60000u;
    z ^= (z >> 18);
    const uint64_t product = (uint64_t)z * (uint64_t)range;
    const uint32_t low = (uint32_t)product;
    if (low < range && low < (0u - range) % range) reject = true;
    const uint32_t x = (uint32_t)(product >> 32);
    scratch[lane] = (even && lane == 0) ? x : ((x / (sr + 1u)) | ((x % (sr + 1u)) << 16));
  }
  if (__ballot_sync(MG_FULL_MASK, reject) != 0u) return false;
  s.idx += (uint32_t)nd;
  __syncwarp();
  return true;
}

// Second half, one lane: apply the swaps std_shuffle_draws_warp left in `scratch`.
MG_DEVICE inline void std_shuffle_apply(uint8_t* order, int n, const uint32_t* scratch) {
  for (int k = 0; k < n; k++) order[k] = (uint8_t)k;
  if (n <= 1) return;
  int i = 1, d = 0;
  if ((n & 1) == 0) {
    const uint32_t j = scratch[0];
    uint8_t t = order[1]; order[1] = order[j]; order[j] = t;
    i = 2; d = 1;
  }
  for (; i < n; i += 2, d++) {
    const uint32_t p0 = scratch[d] & 0xFFFFu, p1 = scratch[d] >> 16;
    uint8_t t = order[i]; order[i] = order[p0]; order[p0] = t;
    t = order[i + 1]; order[i + 1] = order[p1]; order[p1] = t;
  }
}

// std::discrete_distribution<size_t>(weights)(g) (bits/random.tcc:2657-2730); generate_canonical
// draws two words (:3349-3382). A single weight returns 0 without drawing. Doubles, as the reference.
MG_DEVICE inline int discrete_pick_warp(MtStream& st, int lane, const int* weights, int n) {
  if (n < 2) return 0;
  double sum = 0.0;
  for (int i = 0; i < n; i++) sum += (double)weights[i];
  double s = (double)mt_next_warp(st, lane);
  s += (double)mt_next_warp(st, lane) * 4294967296.0;
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
