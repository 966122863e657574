// mg_obs.cuh -- the ObservationEncoder: one warp encodes one agent's egocentric window.
//
// Reference: MettaGrid::_compute_observation (bindings/mettagrid_c.cpp:397-555) walks the window
// in Manhattan order (systems/packed_coordinate.hpp:74-143), asks each object for its feature
// list (agent.cpp:187-213, assembler.hpp:528-578, chest.hpp:124-146, wall.hpp:26-38) and appends
// (location, feature, value) byte triples until num_tokens are written
// (systems/observation_encoder.hpp:107-117), counting the overflow.
//
// Here the per-object feature lists are not rebuilt per observer:
//   * walls, protocol details, tag/vibe tails come from token templates built once per config;
//   * what an agent or a chest shows (group, frozen, vibe, inventory digits, tags) is written
//     once per step by the action kernel into a small token record per entity;
//   * only an assembler's cooldown / clipped / remaining-uses tokens are computed here.
// Each lane owns a run of the Manhattan-ordered cells: it looks up the token count of its cells,
// a warp prefix sum turns counts into token offsets, and the lane copies its cells' tokens
// (2 bytes each) behind the location byte into a shared-memory image of the agent's row, which
// is then stored to HBM with 16-byte stores.
#ifndef MG_OBS_CUH_
#define MG_OBS_CUH_

#include "mg_platform.h"
#include "mg_state.h"

// ---------------------------------------------------------------- entity token records (written by the action kernel)
struct TokenWriter {
  uint16_t* out;
  int n, cap;
};
MG_DEVICE inline void tw_put(TokenWriter& w, int feature, int value) {
  if (w.n < w.cap) w.out[w.n] = (uint16_t)((feature & 0xFF) | ((value & 0xFF) << 8));
  w.n++;
}
// ObservationEncoder::append_inventory_tokens, observation_encoder.hpp:160-173
MG_DEVICE inline void tw_inventory(TokenWriter& w, const int32_t* cfg, uint64_t order, const uint16_t* amt) {
  const uint32_t base = (uint32_t)cfg[MGC_TOKEN_BASE];
  const int ntok = cfg[MGC_NUM_INV_TOKENS];
  const int32_t* feat = cfg + cfg[MGC_OFF_INV_FEATURES];
  for (int k = 0; k < 16; k++) {
    const uint32_t item = (uint32_t)(order & 0xFull);
    if (item == 0xFu) break;
    order >>= 4;
    const uint32_t amount = amt[item];
    tw_put(w, feat[item * MG_MAX_INV_TOKENS], (int)(amount % base));
    uint32_t rem = amount / base;
    for (int p = 1; p < ntok && rem > 0; p++) {
      tw_put(w, feat[item * MG_MAX_INV_TOKENS + p], (int)(rem % base));
      rem /= base;
    }
  }
}
MG_DEVICE inline void tw_template(TokenWriter& w, const MgBatch& b, int ix) {
  const uint32_t e = MG_LDG(&b.tmpl_ix[ix]);
  const uint16_t* src = b.tmpl_tok + (e >> 8);
  for (uint32_t k = 0; k < (e & 0xFFu); k++) {
    if (w.n < w.cap) w.out[w.n] = MG_LDG(&src[k]);
    w.n++;
  }
}
// Agent::obs_features (agent.cpp:187-213): group, frozen, vibe if set, inventory in map order, tags
MG_DEVICE inline void write_agent_tokens(const MgBatch& b, const AgentRec& a, uint8_t* rec) {
  const int32_t* cfg = b.cfg;
  TokenWriter w;
  w.out = reinterpret_cast<uint16_t*>(rec + 2);
  w.n = 0;
  w.cap = (b.atok_stride - 2) >> 1;
  const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + a.group * MG_GRP_WORDS;
  tw_put(w, cfg[MGC_FEAT_GROUP], grp[MG_GRP_GROUP_ID]);
  tw_put(w, cfg[MGC_FEAT_FROZEN], a.frozen != 0 ? 1 : 0);
  if (a.vibe != 0) tw_put(w, cfg[MGC_FEAT_VIBE], a.vibe);
  tw_inventory(w, cfg, a.order, a.amt);
  tw_template(w, b, b.ix_group + a.group);
  rec[0] = (uint8_t)w.n;
}
// Chest::obs_features (chest.hpp:124-146): vibe if set, inventory, tags
MG_DEVICE inline void write_chest_tokens(const MgBatch& b, const ChestRec& c, uint8_t* rec) {
  const int32_t* cfg = b.cfg;
  TokenWriter w;
  w.out = reinterpret_cast<uint16_t*>(rec + 2);
  w.n = 0;
  w.cap = (b.ctok_stride - 2) >> 1;
  const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c.cell * MG_CELL_WORDS;
  if (cc[MG_CELL_VIBE] != 0) tw_put(w, cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]);
  tw_inventory(w, cfg, c.order, c.amt);
  tw_template(w, b, c.cell);
  rec[0] = (uint8_t)w.n;
}

// Assembler::obs_features (assembler.hpp:528-578): cooldown_remaining, clipped, remaining_uses, the
// current protocol's inputs and outputs, tags, vibe. Rewritten every step by the action kernel
// because cooldown_remaining changes with time.
struct AsmView;
MG_DEVICE inline void asm_view(const MgBatch& b, const int32_t* cfg, const AsmRec* a, uint32_t current_step, AsmView& v);
MG_DEVICE inline void write_asm_tokens(const MgBatch& b, const AsmRec* a, uint32_t current_step, uint8_t* rec);

// ---------------------------------------------------------------- per-observer encoding
// What an assembler shows (assembler.hpp:528-578): [cooldown_remaining][clipped][remaining_uses]
// computed here, then the current protocol's inputs/outputs (template), then tags and vibe (template).
struct AsmView {
  uint32_t dyn[3];
  int ndyn;
  uint32_t proto_ix, tail_ix;  // template index entries (offset << 8 | count); count 0 = nothing
};
MG_DEVICE inline void asm_view(const MgBatch& b, const int32_t* cfg, const AsmRec* a, uint32_t current_step, AsmView& v) {
  const uint4 lo = *reinterpret_cast<const uint4*>(a);        // r,c | cell,flags,unclip | cooldown_end | cooldown_duration
  const uint4 hi = *(reinterpret_cast<const uint4*>(a) + 1);  // uses_count | infection_weight | cache_protocol | cache_step
  const uint32_t cell = lo.y & 0xFFu, flags = (lo.y >> 8) & 0xFFu;
  const int unclip = (int)(int16_t)(lo.y >> 16);
  const uint32_t cooldown_end = lo.z, uses = hi.x;
  const int32_t* ty = cfg + cfg[MGC_OFF_ASM_TYPES] + cfg[cfg[MGC_OFF_CELLTYPES] + cell * MG_CELL_WORDS + MG_CELL_PARAM] * MG_ASM_WORDS;
  v.ndyn = 0;
  uint32_t rem = cooldown_end <= current_step ? 0u : cooldown_end - current_step;
  if (rem > 255u) rem = 255u;
  if (rem > 0u) v.dyn[v.ndyn++] = (uint32_t)cfg[MGC_FEAT_COOLDOWN] | (rem << 8);
  const bool clipped = (flags & MG_ASMF_CLIPPED) != 0;
  if (clipped) v.dyn[v.ndyn++] = (uint32_t)cfg[MGC_FEAT_CLIPPED] | (1u << 8);
  const uint32_t max_uses = (uint32_t)ty[MG_ASM_MAX_USES];
  if (max_uses > 0u) {
    uint32_t left = uses < max_uses ? max_uses - uses : 0u;
    if (left > 255u) left = 255u;
    v.dyn[v.ndyn++] = (uint32_t)cfg[MGC_FEAT_REMAINING_USES] | (left << 8);
  }
  v.proto_ix = 0;
  if (cfg[MGC_PROTOCOL_DETAILS_OBS]) {
    // The protocol an observer sees depends on the agents around the assembler *now*
    // (get_current_protocol, assembler.hpp:378-408). The action kernel caches it on the record for
    // every assembler that has an agent next to it this step; all others are idle.
    int pi;
    if (hi.w == current_step) {
      pi = (int)hi.z;
    } else if (clipped) {
      pi = -1;
      if (unclip >= 0) {
        const int32_t* p = cfg + cfg[MGC_OFF_PROTOCOLS] + unclip * MG_PROTO_WORDS;
        if (p[MG_PROTO_KEY_LO] == 0 && p[MG_PROTO_KEY_HI] == 0 && p[MG_PROTO_MIN_AGENTS] <= 0) pi = unclip;
      }
    } else {
      pi = ty[MG_ASM_IDLE_PROTOCOL];
    }
    if (pi >= 0) v.proto_ix = MG_LDG(&b.tmpl_ix[b.ix_proto + pi]);
  }
  v.tail_ix = MG_LDG(&b.tmpl_ix[cell]);
}

MG_DEVICE inline void write_asm_tokens(const MgBatch& b, const AsmRec* a, uint32_t current_step, uint8_t* rec) {
  AsmView v;
  asm_view(b, b.cfg, a, current_step, v);
  uint16_t* out = reinterpret_cast<uint16_t*>(rec + 2);
  const int cap = (b.mtok_stride - 2) >> 1;
  int n = 0;
  for (int k = 0; k < v.ndyn; k++) { if (n < cap) out[n] = (uint16_t)v.dyn[k]; n++; }
  const uint16_t* ps = b.tmpl_tok + (v.proto_ix >> 8);
  for (uint32_t k = 0; k < (v.proto_ix & 0xFFu); k++) { if (n < cap) out[n] = MG_LDG(&ps[k]); n++; }
  const uint16_t* ts = b.tmpl_tok + (v.tail_ix >> 8);
  for (uint32_t k = 0; k < (v.tail_ix & 0xFFu); k++) { if (n < cap) out[n] = MG_LDG(&ts[k]); n++; }
  rec[0] = (uint8_t)n;
}

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
MG_DEVICE inline int obs_encode_agent_warp(const MgBatch& b, int env, int ai, int lane, uint8_t* row,
                                           uint32_t current_step, int last_action) {
  const int32_t* cfg = b.cfg;
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const AgentRec* me = b.agents + (size_t)env * b.A + ai;
  const uint32_t pos_rc = *reinterpret_cast<const uint32_t*>(me);  // r | c << 16
  const int my_r = (int)(pos_rc & 0xFFFFu), my_c = (int)(pos_rc >> 16);
  const int oh = cfg[MGC_OBS_H], ow = cfg[MGC_OBS_W];
  const int hr = oh >> 1, wr = ow >> 1;
  const int T = b.T;

  // ---- global tokens at the window centre, then the compass token (mettagrid_c.cpp:425-517).
  // Uniform across lanes; lane 0 stores.
  int gpos = 0;
  uint8_t* grow = (EMIT && lane == 0) ? row : nullptr;
  const int gcap = grow != nullptr ? T : 0;
  {
    const uint32_t centre = (uint32_t)((hr << 4) | (wr & 0xF));
    if (cfg[MGC_GOBS_ECP]) {
      const uint32_t max_steps = (uint32_t)cfg[MGC_MAX_STEPS];
      uint32_t v = 0;
      if (max_steps > 0u) v = current_step >= max_steps ? 255u : ((256u * current_step / max_steps) & 0xFFu);
      row_put(grow, gpos, gcap, centre, (uint32_t)cfg[MGC_FEAT_ECP] | (v << 8));
    }
    if (cfg[MGC_GOBS_LAST_ACTION]) row_put(grow, gpos, gcap, centre, (uint32_t)cfg[MGC_FEAT_LAST_ACTION] | ((uint32_t)(last_action & 0xFF) << 8));
    // rewards are zeroed before and written after the observations of a step (mettagrid_c.cpp:578,656):
    // the last_reward token is always round(0 * 100) = 0.
    if (cfg[MGC_GOBS_LAST_REWARD]) row_put(grow, gpos, gcap, centre, (uint32_t)cfg[MGC_FEAT_LAST_REWARD]);
    if (cfg[MGC_GOBS_GOAL]) {
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + me->group * MG_GRP_WORDS;
      for (int i = 0; i < grp[MG_GRP_NUM_GOALS]; i++)
        row_put(grow, gpos, gcap, centre, (uint32_t)cfg[MGC_FEAT_GOAL] | ((uint32_t)(grp[MG_GRP_GOALS + i] & 0xFF) << 8));
    }
    if (cfg[MGC_GOBS_COMPASS]) {
      const int dr = b.H / 2 - my_r, dc = b.W / 2 - my_c;
      const int sr = dr > 0 ? 1 : (dr < 0 ? -1 : 0), sc = dc > 0 ? 1 : (dc < 0 ? -1 : 0);
      if (sr != 0 || sc != 0) {
        const int orr = hr + sr, occ = wr + sc;
        if (orr >= 0 && orr < oh && occ >= 0 && occ < ow)
          row_put(grow, gpos, gcap, (uint32_t)((orr << 4) | (occ & 0xF)), (uint32_t)cfg[MGC_FEAT_COMPASS] | (1u << 8));
      }
    }
  }
  const int n_global = gpos;

  // ---- pass 1: my cells' grid values and token counts. Every object, static or not, shows a token
  // record: byte 0 = count, u16 tokens from byte 2.
  const uint8_t* rec_base[4];
  int rec_stride[4];
  rec_base[MG_GK_STATIC] = b.static_tok;                                            rec_stride[MG_GK_STATIC] = b.stok_stride;
  rec_base[MG_GK_AGENT] = b.agent_tok + (size_t)env * b.A * b.atok_stride;          rec_stride[MG_GK_AGENT] = b.atok_stride;
  rec_base[MG_GK_ASM] = b.asm_tok + (size_t)env * b.max_asm * b.mtok_stride;        rec_stride[MG_GK_ASM] = b.mtok_stride;
  rec_base[MG_GK_CHEST] = b.chest_tok + (size_t)env * b.max_chest * b.ctok_stride;  rec_stride[MG_GK_CHEST] = b.ctok_stride;
  const int ncells = b.obs_cells;
  const uint8_t* rec[CPL];
  uint32_t cnt[CPL];
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
      const uint32_t kind = MG_GRID_KIND(g), idx = MG_GRID_INDEX(g);
      const uint8_t* base = kind == MG_GK_STATIC ? rec_base[0] : (kind == MG_GK_AGENT ? rec_base[1] : (kind == MG_GK_ASM ? rec_base[2] : rec_base[3]));
      const int stride = kind == MG_GK_STATIC ? rec_stride[0] : (kind == MG_GK_AGENT ? rec_stride[1] : (kind == MG_GK_ASM ? rec_stride[2] : rec_stride[3]));
      rec[j] = base + (size_t)idx * stride;
      n = rec[j][0];
    }
    cnt[j] = n;
    mine += (int)n;
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
      const uint16_t* src = reinterpret_cast<const uint16_t*>(rec[j] + 2);
      for (uint32_t k = 0; k < cnt[j]; k++) row_put(row, pos, T, loc, src[k]);
    }
  }
  return n_global + total;
}

// Count-only pass used by the reset path (any window size).
MG_DEVICE inline int obs_count_agent_warp(const MgBatch& b, int env, int ai, int lane, uint32_t current_step) {
  const int cpl = (b.obs_cells + 31) >> 5;
  if (cpl <= 4) return obs_encode_agent_warp<4, false>(b, env, ai, lane, nullptr, current_step, 0);
  return obs_encode_agent_warp<8, false>(b, env, ai, lane, nullptr, current_step, 0);
}

#endif  // MG_OBS_CUH_
