// mg_obs.cuh -- the ObservationEncoder: one warp encodes one agent's egocentric window.
//
// Reference: MettaGrid::_compute_observation (bindings/mettagrid_c.cpp:397-555) walks the window
// in Manhattan order (systems/packed_coordinate.hpp:74-143), asks each object for its feature
// list and appends (location, feature, value) byte triples until num_tokens are written
// (systems/observation_encoder.hpp:107-117), counting the overflow.
//
// Here nothing is rebuilt per observer: every object, static or not, has a token record
// (mg_tokens.cuh) and every agent a record of its global tokens. Each lane gathers a run of the
// Manhattan-ordered window cells; the warp compacts the non-empty ones (most of a window is empty),
// reads one record header per lane (token count, cooldown end), turns the counts into token offsets
// with a prefix sum, and then emits the row one token per lane (2 bytes fetched from the object's
// record, behind the location byte) into a shared-memory image of the agent's row, which is stored
// to HBM with 16-byte stores.
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

// Window cells of one lane: CPL consecutive pattern bytes in one load (the table is padded with 0xFF).
template <int CPL>
MG_DEVICE inline uint64_t load_pattern_run(const uint8_t* pattern, int lane) {
  const uint8_t* p = pattern + lane * CPL;
  if (CPL == 8) { const uint2 v = *reinterpret_cast<const uint2*>(p); return (uint64_t)v.x | ((uint64_t)v.y << 32); }
  if (CPL == 4) return *reinterpret_cast<const uint32_t*>(p);
  if (CPL == 2) return *reinterpret_cast<const uint16_t*>(p);
  return *p;
}

// Bytes of shared memory one warp needs besides its row image: one 8-byte descriptor per window cell.
#define MG_OBS_DESC_BYTES(cpl) ((cpl) * 32 * 8)

// Token count of agent `ai`'s observation (what the encoder below would attempt to write), without
// writing anything: the reset path counts the observation passes the reference makes in its
// constructor (tokens_written / dropped / free_space stats). Must be called by all 32 lanes; CPL =
// window cells per lane (CPL * 32 >= obs_h * obs_w).
template <int CPL>
MG_DEVICE inline int obs_count_cells_warp(const MgBatch& b, int env, int ai, int lane, uint32_t current_step) {
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const uint32_t pos_rc = *reinterpret_cast<const uint32_t*>(b.agents + (size_t)env * b.A + ai);  // r | c << 16
  const int my_r = (int)(pos_rc & 0xFFFFu), my_c = (int)(pos_rc >> 16);
  const int hr = b.obs_hr, wr = b.obs_wr;
  const int n_global = b.gtok[((size_t)env * b.A + ai) * MG_GTOK_STRIDE];
  const uint8_t* ebase = b.etok + (((size_t)env * b.rec_slots) << b.rec_shift);
  const uint64_t kind_base = (uint64_t)b.rec_kind_base_lo | ((uint64_t)b.rec_kind_base_hi << 32);
  const uint64_t locs = load_pattern_run<CPL>(b.pattern, lane);
  int mine = 0;
#pragma unroll
  for (int j = 0; j < CPL; j++) {
    const uint32_t loc = (uint32_t)(locs >> (8 * j)) & 0xFFu;
    const int r = my_r + (int)(loc >> 4) - hr, c = my_c + (int)(loc & 0xF) - wr;
    uint32_t g = 0;
    if (loc != 0xFFu && (unsigned)r < (unsigned)b.H && (unsigned)c < (unsigned)b.W) g = grid[r * b.W + c];
    if (g != 0) {
      const uint32_t slot = ((uint32_t)(kind_base >> (MG_GRID_KIND(g) * 16u)) & 0xFFFFu) + MG_GRID_INDEX(g);
      const uint2 h = *reinterpret_cast<const uint2*>(ebase + ((size_t)slot << b.rec_shift));
      mine += (int)(h.x & 0xFFu) + (h.y > current_step ? 1 : 0);  // + the cooldown_remaining token while it is > 0
    }
  }
  int total;
  warp_exclusive_scan(mine, lane, &total);
  return n_global + total;
}

// The encoder the observation kernels use, arranged for few issued instructions: most window cells are empty, so after the grid gather the
// warp compacts its non-empty cells (ballots + popcounts, window order kept) into shared memory and
// then looks at one OBJECT per lane (record header, cooldown, token count, prefix sum, descriptor)
// instead of CPL cells per lane. The cell list shares `desc`'s second half: an entry is read before the
// descriptors written in the same round can reach it (descriptors are 8 bytes, at most one per object).
// `fill_words`: > 0 = first set that many 16-byte words at `row` (the whole image) to 0xFF. The caller's fill is
// done here so that the loads everything else waits for -- the agent's position, its global tokens, the
// lane's pattern entries -- are already in flight while the image is filled.
template <int CPL, int TT = 0>
MG_DEVICE inline int obs_emit_agent_warp(const MgBatch& b, int env, int ai, int lane, uint8_t* row, uint2* desc, uint32_t current_step,
                                         int fill_words = 0) {
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const uint32_t pos_rc = *reinterpret_cast<const uint32_t*>(b.agents + (size_t)env * b.A + ai);  // r | c << 16
  // ---- global tokens: prepared by the step kernel (a count, then up to 21 triples in a 64-byte record)
  const uint8_t* gt = b.gtok + ((size_t)env * b.A + ai) * MG_GTOK_STRIDE;
  const int n_global = gt[0];
  const uint8_t gt_lo = gt[1 + lane], gt_hi = lane < 31 ? gt[33 + lane] : (uint8_t)0xFF;
  const uint64_t locs = load_pattern_run<CPL>(b.pattern, lane);
  if (fill_words > 0) {
    uint4 ff;
    ff.x = ff.y = ff.z = ff.w = 0xFFFFFFFFu;
    uint4* p = reinterpret_cast<uint4*>(row);
    if (lane < fill_words) p[lane] = ff;
    if (lane + 32 < fill_words) p[lane + 32] = ff;
#pragma unroll 1
    for (int i = lane + 64; i < fill_words; i += 32) p[i] = ff;
    __syncwarp();
  }
  const int my_r = (int)(pos_rc & 0xFFFFu), my_c = (int)(pos_rc >> 16);
  const int hr = b.obs_hr, wr = b.obs_wr;
  const int T = TT ? TT : b.T;
  uint32_t* cells = reinterpret_cast<uint32_t*>(desc) + CPL * 32;
  {
    const int nb = (n_global < T ? n_global : T) * 3;  // <= 63 bytes (MG_GTOK_STRIDE)
    if (lane < nb) row[lane] = gt_lo;
    if (lane + 32 < nb) row[lane + 32] = gt_hi;
  }

  // ---- pass 1a: grid values of my CPL cells; non-empty ones go to the cell list in window order
  uint32_t cell[CPL];  // grid value | location << 16, 0 = nothing there
  const uint32_t lower = (1u << lane) - 1u;
  int rank = 0, n_obj = 0;
#ifdef MG_OBS_NO_FASTGATHER
  if (false) {
#else
  if (my_r >= hr && my_c >= wr && my_r + hr < b.H && my_c + wr < b.W) {
#endif
    // the whole window lies inside the map (always, where the map's border is at least as wide as the window's
    // radius, as in the arena): one add per cell, no range tests
    const uint16_t* corner = grid + (my_r - hr) * b.W + (my_c - wr);
    int32_t off[CPL];  // the lane's cells as offsets from the corner, precomputed per config (-1: no cell)
    if (CPL == 4) {
      const int4 o4 = *reinterpret_cast<const int4*>(b.pattern_off + lane * 4);
      off[0] = o4.x; off[1 % CPL] = o4.y; off[2 % CPL] = o4.z; off[3 % CPL] = o4.w;
    } else {
#pragma unroll
      for (int j = 0; j < CPL; j++) off[j] = b.pattern_off[lane * CPL + j];
    }
    uint32_t g[CPL];
#pragma unroll
    for (int j = 0; j < CPL; j++) g[j] = off[j] >= 0 ? (uint32_t)corner[off[j]] : 0u;
#pragma unroll
    for (int j = 0; j < CPL; j++) cell[j] = g[j] != 0u ? (g[j] | (((uint32_t)(locs >> (8 * j)) & 0xFFu) << 16)) : 0u;
  } else {
#pragma unroll
    for (int j = 0; j < CPL; j++) {
      const uint32_t loc = (uint32_t)(locs >> (8 * j)) & 0xFFu;
      const uint32_t r = (uint32_t)(my_r - hr) + (loc >> 4), c = (uint32_t)(my_c - wr) + (loc & 0xFu);  // wraps below zero: fails the range test
      uint32_t g = 0;
      if (r < (uint32_t)b.H && c < (uint32_t)b.W && loc != 0xFFu) g = grid[r * (uint32_t)b.W + c];
      cell[j] = g != 0u ? (g | (loc << 16)) : 0u;
    }
  }
#pragma unroll
  for (int j = 0; j < CPL; j++) {
    const uint32_t m = __ballot_sync(MG_FULL_MASK, cell[j] != 0u);
    rank += __popc(m & lower);
    n_obj += __popc(m);
  }
#pragma unroll
  for (int j = 0; j < CPL; j++)
    if (cell[j] != 0u) cells[rank++] = cell[j];
  __syncwarp();

  // ---- pass 1b: one object per lane: header of its token record, token offset, descriptor
  const uint8_t* ebase = b.etok + (((size_t)env * b.rec_slots) << b.rec_shift);
  const uint64_t kind_base = (uint64_t)b.rec_kind_base_lo | ((uint64_t)b.rec_kind_base_hi << 32);
  int pos_base = n_global, di_base = 0;
#pragma unroll 1
  for (int o0 = 0; o0 < n_obj; o0 += 32) {
    const int o = o0 + lane;
    const uint32_t cv = o < n_obj ? cells[o] : 0u;
    uint32_t slot = 0, rem = 0;
    int tokens = 0;
    if (cv != 0u) {
      const uint32_t g = cv & 0xFFFFu;
      slot = ((uint32_t)(kind_base >> (MG_GRID_KIND(g) * 16u)) & 0xFFFFu) + MG_GRID_INDEX(g);
      const uint2 h = *reinterpret_cast<const uint2*>(ebase + ((size_t)slot << b.rec_shift));
      rem = h.y > current_step ? h.y - current_step : 0u;  // Assembler::cooldown_remaining, capped for the token
      rem = rem > 255u ? 255u : rem;
      tokens = (int)(h.x & 0xFFu) + (rem != 0u ? 1 : 0);
    }
    int total;
    const int offset = warp_exclusive_scan(tokens != 0 ? tokens | 0x10000 : 0, lane, &total);
    __syncwarp();  // every lane has read its cell before descriptors overwrite the list's neighbourhood
    if (tokens != 0) {
      const int pos = pos_base + (offset & 0xFFFF);
      const uint32_t loc = cv >> 16;
      uint2 dv;
      dv.x = slot | ((uint32_t)pos << 16);
      dv.y = loc | (rem << 8);
      desc[di_base + (offset >> 16)] = dv;
      if (pos < T) row[pos * 3] = (uint8_t)loc;
    }
    pos_base += total & 0xFFFF;
    di_base += total >> 16;
  }
  __syncwarp();

  // ---- pass 2: one token per lane
  const int end = pos_base < T ? pos_base : T;
  const uint32_t cooldown_feature = (uint32_t)b.feat_cooldown;
  int before = 0;  // descriptors that start before this group of 32 tokens
  for (int t0 = n_global; t0 < end; t0 += 32) {
    const int t = t0 + lane;
    uint8_t* out = row + t * 3;
    const bool live = t < end;
    const uint32_t starts = __ballot_sync(MG_FULL_MASK, live && out[0] != 0xFFu);
    if (live) {
      const uint2 d = desc[before + __popc(starts & (0xFFFFFFFFu >> (31 - lane))) - 1];
      const int k = t - (int)(d.x >> 16);
      const uint32_t rem = (d.y >> 8) & 0xFFu;
      uint32_t tok;
      if (rem != 0u && k == 0) {
        tok = cooldown_feature | (rem << 8);
      } else {
        const uint8_t* rec = ebase + ((size_t)(d.x & 0xFFFFu) << b.rec_shift) + MG_REC_TOKENS;
        tok = reinterpret_cast<const uint16_t*>(rec)[k - (rem != 0u ? 1 : 0)];
      }
      out[0] = (uint8_t)d.y;
      out[1] = (uint8_t)tok;
      out[2] = (uint8_t)(tok >> 8);
    }
    before += __popc(starts);
  }
  return pos_base;
}

// Count-only pass for U consecutive agents at once: the loads of all of them (position, grid cells, record
// headers) are issued together, so that the pass costs about two memory round trips per group of agents
// instead of per agent (an env's reset runs it for every agent of the env on one warp).
template <int CPL, int U>
MG_DEVICE inline void obs_count_group_warp(const MgBatch& b, int env, int ai0, int n, int lane, uint32_t current_step, int* totals) {
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const int hr = b.obs_hr, wr = b.obs_wr;
  const uint8_t* ebase = b.etok + (((size_t)env * b.rec_slots) << b.rec_shift);
  const uint64_t kind_base = (uint64_t)b.rec_kind_base_lo | ((uint64_t)b.rec_kind_base_hi << 32);
  const uint64_t locs = load_pattern_run<CPL>(b.pattern, lane);
  uint32_t pos[U];
  int n_global[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int ai = ai0 + (u < n ? u : 0);
    pos[u] = *reinterpret_cast<const uint32_t*>(b.agents + (size_t)env * b.A + ai);  // r | c << 16
    n_global[u] = b.gtok[((size_t)env * b.A + ai) * MG_GTOK_STRIDE];
  }
  uint32_t g[U][CPL];
#pragma unroll
  for (int u = 0; u < U; u++)
#pragma unroll
    for (int j = 0; j < CPL; j++) {
      const uint32_t loc = (uint32_t)(locs >> (8 * j)) & 0xFFu;
      const uint32_t r = (pos[u] & 0xFFFFu) - (uint32_t)hr + (loc >> 4), c = (pos[u] >> 16) - (uint32_t)wr + (loc & 0xFu);
      g[u][j] = (u < n && loc != 0xFFu && r < (uint32_t)b.H && c < (uint32_t)b.W) ? grid[r * (uint32_t)b.W + c] : 0u;
    }
  uint2 h[U][CPL];
#pragma unroll
  for (int u = 0; u < U; u++)
#pragma unroll
    for (int j = 0; j < CPL; j++) {
      h[u][j].x = h[u][j].y = 0u;
      if (g[u][j] != 0u) {
        const uint32_t slot = ((uint32_t)(kind_base >> (MG_GRID_KIND(g[u][j]) * 16u)) & 0xFFFFu) + MG_GRID_INDEX(g[u][j]);
        h[u][j] = *reinterpret_cast<const uint2*>(ebase + ((size_t)slot << b.rec_shift));
      }
    }
#pragma unroll
  for (int u = 0; u < U; u++) {
    int mine = 0;
#pragma unroll
    for (int j = 0; j < CPL; j++)
      if (g[u][j] != 0u) mine += (int)(h[u][j].x & 0xFFu) + (h[u][j].y > current_step ? 1 : 0);  // + the cooldown_remaining token while it is > 0
    totals[u] = n_global[u] + MG_WARP_SUM(mine);
  }
}

// Count-only pass used by the reset path (any window size).
MG_DEVICE inline int obs_count_agent_warp(const MgBatch& b, int env, int ai, int lane, uint32_t current_step) {
  const int cpl = (b.obs_cells + 31) >> 5;
  if (cpl <= 4) return obs_count_cells_warp<4>(b, env, ai, lane, current_step);
  return obs_count_cells_warp<8>(b, env, ai, lane, current_step);
}

#endif  // MG_OBS_CUH_
