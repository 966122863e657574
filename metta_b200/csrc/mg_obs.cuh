// mg_obs.cuh -- the ObservationEncoder: one warp encodes one agent's egocentric window.
//
// Reference: MettaGrid::_compute_observation (bindings/mettagrid_c.cpp:397-555) walks the window
// in Manhattan order (systems/packed_coordinate.hpp:74-143), asks each object for its feature
// list and appends (location, feature, value) byte triples until num_tokens are written
// (systems/observation_encoder.hpp:107-117), counting the overflow.
//
// Here nothing is rebuilt per observer: every object, static or not, has a token record
// (mg_tokens.cuh) and every agent a record of its global tokens. Each lane owns a run of the
// Manhattan-ordered cells: it reads the 8-byte header of its cells' records (token count,
// cooldown end), a warp prefix sum turns counts into token offsets, and the lane copies its cells'
// tokens (2 bytes each) behind the location byte into a shared-memory image of the agent's row,
// which is then stored to HBM with 16-byte stores.
#ifndef MG_OBS_CUH_
#define MG_OBS_CUH_

#include "mg_tokens.cuh"

MG_DEVICE inline int warp_exclusive_scan(int v, int lane, int* total) {
  int x = v;
  for (int d = 1; d < 32; d <<= 1) {
    int y = __shfl_up_sync(MG_FULL_MASK, x, d);
    if (lane >= d) x += y;
  }
  *total = __shfl_sync(MG_FULL_MASK, x, 31);
  return x - v;
}

MG_DEVICE inline void row_put(uint8_t* row, int& pos, int cap, uint32_t loc, uint32_t tok) {
  if (pos < cap) {
    uint8_t* p = row + pos * 3;
    p[0] = (uint8_t)loc;
    p[1] = (uint8_t)tok;
    p[2] = (uint8_t)(tok >> 8);
  }
  pos++;
}

// Encodes agent `ai` of env `env` into `row` (shared-memory image of its T*3 bytes, pre-filled with
// 0xFF) when EMIT, or only counts tokens. Must be called by all 32 lanes. CPL = window cells per
// lane (CPL * 32 >= obs_h * obs_w). Returns the number of tokens attempted.
template <int CPL, bool EMIT>
MG_DEVICE inline int obs_encode_agent_warp(const MgBatch& b, int env, int ai, int lane, uint8_t* row, uint32_t current_step) {
  const int32_t* cfg = b.cfg;
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const uint32_t pos_rc = *reinterpret_cast<const uint32_t*>(b.agents + (size_t)env * b.A + ai);  // r | c << 16
  const int my_r = (int)(pos_rc & 0xFFFFu), my_c = (int)(pos_rc >> 16);
  const int hr = cfg[MGC_OBS_H] >> 1, wr = cfg[MGC_OBS_W] >> 1;
  const int T = b.T;
  const int cooldown_feature = cfg[MGC_FEAT_COOLDOWN];

  // ---- global tokens: prepared by the action kernel, copied byte by byte
  const uint8_t* gt = b.gtok + ((size_t)env * b.A + ai) * MG_GTOK_STRIDE;
  const int n_global = gt[0];
  if (EMIT) {
    const int nb = (n_global < T ? n_global : T) * 3;
    for (int i = lane; i < nb; i += 32) row[i] = gt[1 + i];
  }

  // ---- pass 1: my cells' records and token counts
  const uint8_t* ebase = b.etok + (((size_t)env * b.rec_slots) << b.rec_shift);
  const uint64_t kind_base = (uint64_t)b.rec_kind_base_lo | ((uint64_t)b.rec_kind_base_hi << 32);
  const int ncells = b.obs_cells;
  const uint8_t* rec[CPL];
  uint32_t cnt[CPL];  // tokens | cooldown_remaining << 8
  int mine = 0;
#pragma unroll
  for (int j = 0; j < CPL; j++) {
    const int p = lane * CPL + j;
    uint32_t g = 0;
    if (p < ncells) {
      const uint32_t loc = b.pattern[p];
      const int r = my_r + (int)(loc >> 4) - hr, c = my_c + (int)(loc & 0xF) - wr;
      if (r >= 0 && r < b.H && c >= 0 && c < b.W) g = grid[r * b.W + c];
    }
    uint32_t n = 0;
    rec[j] = nullptr;
    if (g != 0) {
      const uint32_t slot = (uint32_t)(kind_base >> (MG_GRID_KIND(g) * 16u)) & 0xFFFFu;
      rec[j] = ebase + ((size_t)(slot + MG_GRID_INDEX(g)) << b.rec_shift);
      const uint2 h = *reinterpret_cast<const uint2*>(rec[j]);
      uint32_t rem = h.y > current_step ? h.y - current_step : 0u;  // Assembler::cooldown_remaining, capped for the token
      rem = rem > 255u ? 255u : rem;
      n = (h.x & 0xFFu) | (rem << 8);
      mine += (int)(h.x & 0xFFu) + (rem != 0u ? 1 : 0);
    }
    cnt[j] = n;
  }
  int total;
  const int offset = warp_exclusive_scan(mine, lane, &total);
  if (EMIT) {
    // ---- pass 2: copy my cells' tokens behind their location byte
    int pos = n_global + offset;
#pragma unroll
    for (int j = 0; j < CPL; j++) {
      if (cnt[j] == 0 || pos >= T) continue;
      const uint32_t loc = b.pattern[lane * CPL + j];
      const uint32_t rem = cnt[j] >> 8, n = cnt[j] & 0xFFu;
      if (rem != 0u) row_put(row, pos, T, loc, (uint32_t)cooldown_feature | (rem << 8));
      const uint16_t* src = reinterpret_cast<const uint16_t*>(rec[j] + MG_REC_TOKENS);
      for (uint32_t k = 0; k < n; k++) row_put(row, pos, T, loc, src[k]);
    }
  }
  return n_global + total;
}

// Count-only pass used by the reset path (any window size).
MG_DEVICE inline int obs_count_agent_warp(const MgBatch& b, int env, int ai, int lane, uint32_t current_step) {
  const int cpl = (b.obs_cells + 31) >> 5;
  if (cpl <= 4) return obs_encode_agent_warp<4, false>(b, env, ai, lane, nullptr, current_step);
  return obs_encode_agent_warp<8, false>(b, env, ai, lane, nullptr, current_step);
}

#endif  // MG_OBS_CUH_
