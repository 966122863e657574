// mg_obs.cuh -- the ObservationEncoder: one warp encodes one agent's egocentric window.
//
// Reference: MettaGrid::_compute_observation (bindings/mettagrid_c.cpp:397-555) walks the window
// in Manhattan order (systems/packed_coordinate.hpp:74-143), asks each object for its feature
// list (agent.cpp:187-213, assembler.hpp:528-578, chest.hpp:124-146, wall.hpp:26-38) and appends
// (location, feature, value) byte triples until num_tokens are written
// (systems/observation_encoder.hpp:107-117), counting the overflow.
//
// Here each lane owns a contiguous run of the Manhattan-ordered cells. Pass 1 counts the tokens
// of the lane's cells, a warp prefix sum turns counts into token offsets, pass 2 writes the
// tokens into a shared-memory image of the agent's row (pre-filled with 0xFF), and the row is
// stored to HBM with 16-byte stores.
#ifndef MG_OBS_CUH_
#define MG_OBS_CUH_

#include "mg_platform.h"
#include "mg_state.h"

struct TokenSink {
  uint8_t* out;  // row image (may be null: count only)
  int pos;       // next token index
  int cap;       // num_tokens
  uint32_t loc;  // packed location byte
};

MG_DEVICE inline void sink_put(TokenSink& s, int feature, int value) {
  if (s.out != nullptr && s.pos < s.cap) {
    uint8_t* p = s.out + s.pos * 3;
    p[0] = (uint8_t)s.loc;
    p[1] = (uint8_t)feature;
    p[2] = (uint8_t)value;
  }
  s.pos++;
}

// ObservationEncoder::append_inventory_tokens, observation_encoder.hpp:160-173
MG_DEVICE inline void sink_inventory(TokenSink& s, const int32_t* cfg, uint64_t order, const uint16_t* amt) {
  const uint32_t base = (uint32_t)cfg[MGC_TOKEN_BASE];
  const int ntok = cfg[MGC_NUM_INV_TOKENS];
  const int32_t* feat = cfg + cfg[MGC_OFF_INV_FEATURES];
  for (int k = 0; k < 16; k++) {
    uint32_t item = (uint32_t)(order & 0xFull);
    if (item == 0xFu) break;
    order >>= 4;
    uint32_t amount = amt[item];
    sink_put(s, feat[item * MG_MAX_INV_TOKENS], (int)(amount % base));
    uint32_t rem = amount / base;
    for (int p = 1; p < ntok && rem > 0; p++) {
      sink_put(s, feat[item * MG_MAX_INV_TOKENS + p], (int)(rem % base));
      rem /= base;
    }
  }
}

MG_DEVICE inline void sink_tags(TokenSink& s, const int32_t* cfg, const int32_t* tags, int n) {
  const int f = cfg[MGC_FEAT_TAG];
  for (int i = 0; i < n; i++) sink_put(s, f, tags[i]);
}

// Tokens of whatever occupies grid value g. `agents/asms/chests` are this env's global records.
MG_DEVICE inline void sink_object(TokenSink& s, const MgBatch& b, const int32_t* cfg, uint32_t g, uint32_t current_step,
                                  const AgentRec* agents, const AsmRec* asms, const ChestRec* chests) {
  const uint32_t kind = MG_GRID_KIND(g), idx = MG_GRID_INDEX(g);
  if (kind == MG_GK_STATIC) {  // Wall::obs_features: tags, then vibe
    const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + idx * MG_CELL_WORDS;
    sink_tags(s, cfg, cc + MG_CELL_TAGS, cc[MG_CELL_NUM_TAGS]);
    if (cc[MG_CELL_VIBE] != 0) sink_put(s, cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]);
  } else if (kind == MG_GK_AGENT) {  // Agent::obs_features: group, frozen, vibe, inventory, tags
    const AgentRec* a = &agents[idx];
    const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + a->group * MG_GRP_WORDS;
    sink_put(s, cfg[MGC_FEAT_GROUP], grp[MG_GRP_GROUP_ID]);
    sink_put(s, cfg[MGC_FEAT_FROZEN], a->frozen != 0 ? 1 : 0);
    if (a->vibe != 0) sink_put(s, cfg[MGC_FEAT_VIBE], a->vibe);
    sink_inventory(s, cfg, a->order, a->amt);
    sink_tags(s, cfg, grp + MG_GRP_TAGS, grp[MG_GRP_NUM_TAGS]);
  } else if (kind == MG_GK_ASM) {  // Assembler::obs_features
    const AsmRec* a = &asms[idx];
    const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + a->cell * MG_CELL_WORDS;
    const int32_t* ty = cfg + cfg[MGC_OFF_ASM_TYPES] + cc[MG_CELL_PARAM] * MG_ASM_WORDS;
    uint32_t rem = a->cooldown_end <= current_step ? 0u : a->cooldown_end - current_step;
    if (rem > 255u) rem = 255u;
    if (rem > 0u) sink_put(s, cfg[MGC_FEAT_COOLDOWN], (int)rem);
    const bool clipped = (a->flags & MG_ASMF_CLIPPED) != 0;
    if (clipped) sink_put(s, cfg[MGC_FEAT_CLIPPED], 1);
    const uint32_t max_uses = (uint32_t)ty[MG_ASM_MAX_USES];
    if (max_uses > 0u) {
      uint32_t left = a->uses_count < max_uses ? max_uses - a->uses_count : 0u;
      sink_put(s, cfg[MGC_FEAT_REMAINING_USES], (int)(left > 255u ? 255u : left));
    }
    if (cfg[MGC_PROTOCOL_DETAILS_OBS]) {
      // The protocol an observer sees depends on the agents around the assembler *now*
      // (get_current_protocol, assembler.hpp:378-408). The action kernel caches it on the record
      // for every assembler that has an agent next to it this step; all others are idle.
      int pi;
      if (a->cache_step == current_step) {
        pi = a->cache_protocol;
      } else if (clipped) {
        pi = -1;
        if (a->unclip_protocol >= 0) {
          const int32_t* p = cfg + cfg[MGC_OFF_PROTOCOLS] + a->unclip_protocol * MG_PROTO_WORDS;
          if (p[MG_PROTO_KEY_LO] == 0 && p[MG_PROTO_KEY_HI] == 0 && p[MG_PROTO_MIN_AGENTS] <= 0) pi = a->unclip_protocol;
        }
      } else {
        pi = ty[MG_ASM_IDLE_PROTOCOL];
      }
      if (pi >= 0) {
        const int32_t* p = cfg + cfg[MGC_OFF_PROTOCOLS] + pi * MG_PROTO_WORDS;
        const int32_t* pf = cfg + cfg[MGC_OFF_PROTO_FEATURES];
        const int32_t* l = p + MG_PROTO_INPUTS;
        for (int i = 0; i < l[MG_LIST_N]; i++)
          if (l[MG_LIST_VAL(i)] > 0) sink_put(s, pf[l[MG_LIST_ITEM(i)]], l[MG_LIST_VAL(i)]);
        l = p + MG_PROTO_OUTPUTS;
        for (int i = 0; i < l[MG_LIST_N]; i++)
          if (l[MG_LIST_VAL(i)] > 0) sink_put(s, pf[MG_RES_STRIDE + l[MG_LIST_ITEM(i)]], l[MG_LIST_VAL(i)]);
      }
    }
    sink_tags(s, cfg, cc + MG_CELL_TAGS, cc[MG_CELL_NUM_TAGS]);
    if (cc[MG_CELL_VIBE] != 0) sink_put(s, cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]);
  } else {  // Chest::obs_features: vibe, inventory, tags
    const ChestRec* c = &chests[idx];
    const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c->cell * MG_CELL_WORDS;
    if (cc[MG_CELL_VIBE] != 0) sink_put(s, cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]);
    sink_inventory(s, cfg, c->order, c->amt);
    sink_tags(s, cfg, cc + MG_CELL_TAGS, cc[MG_CELL_NUM_TAGS]);
  }
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

// Encodes agent `ai` of env `env` into `row` (shared memory image of its T*3 bytes, already filled
// with 0xFF) or only counts when row == nullptr. Must be called by all 32 lanes.
// Returns the number of tokens attempted (written = min(T, attempted)).
MG_DEVICE inline int obs_encode_agent_warp(const MgBatch& b, int env, int ai, int lane, uint8_t* row,
                                           uint32_t current_step, int last_action) {
  const int32_t* cfg = b.cfg;
  const uint16_t* grid = b.grid + (size_t)env * b.grid_stride;
  const AgentRec* agents = b.agents + (size_t)env * b.A;
  const AsmRec* asms = b.asms + (size_t)env * b.max_asm;
  const ChestRec* chests = b.chests + (size_t)env * b.max_chest;
  const AgentRec* me = &agents[ai];
  const int my_r = me->r, my_c = me->c;
  const int oh = cfg[MGC_OBS_H], ow = cfg[MGC_OBS_W];
  const int hr = oh >> 1, wr = ow >> 1;

  // ---- global tokens at the window centre, then the compass token (mettagrid_c.cpp:425-517)
  TokenSink gs;
  gs.out = lane == 0 ? row : nullptr;
  gs.pos = 0;
  gs.cap = b.T;
  gs.loc = (uint32_t)((hr << 4) | (wr & 0xF));
  if (cfg[MGC_GOBS_ECP]) {
    const uint32_t max_steps = (uint32_t)cfg[MGC_MAX_STEPS];
    int v = 0;
    if (max_steps > 0u) v = current_step >= max_steps ? 255 : (int)((256u * current_step / max_steps) & 0xFFu);
    sink_put(gs, cfg[MGC_FEAT_ECP], v);
  }
  if (cfg[MGC_GOBS_LAST_ACTION]) sink_put(gs, cfg[MGC_FEAT_LAST_ACTION], last_action & 0xFF);
  // rewards are zeroed before and written after the observations of a step (mettagrid_c.cpp:578,656):
  // the last_reward token is always round(0 * 100) = 0.
  if (cfg[MGC_GOBS_LAST_REWARD]) sink_put(gs, cfg[MGC_FEAT_LAST_REWARD], 0);
  if (cfg[MGC_GOBS_GOAL]) {
    const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + me->group * MG_GRP_WORDS;
    for (int i = 0; i < grp[MG_GRP_NUM_GOALS]; i++) sink_put(gs, cfg[MGC_FEAT_GOAL], grp[MG_GRP_GOALS + i]);
  }
  if (cfg[MGC_GOBS_COMPASS]) {
    const int dr = b.H / 2 - my_r, dc = b.W / 2 - my_c;
    const int sr = dr > 0 ? 1 : (dr < 0 ? -1 : 0), sc = dc > 0 ? 1 : (dc < 0 ? -1 : 0);
    if (sr != 0 || sc != 0) {
      const int orr = hr + sr, occ = wr + sc;
      if (orr >= 0 && orr < oh && occ >= 0 && occ < ow) {
        gs.loc = (uint32_t)((orr << 4) | (occ & 0xF));
        sink_put(gs, cfg[MGC_FEAT_COMPASS], 1);
      }
    }
  }
  const int n_global = gs.pos;

  // ---- window cells: lane owns cells [lane*cpl, lane*cpl + cpl) of the Manhattan order
  const int ncells = b.obs_cells;
  const int cpl = (ncells + 31) >> 5;
  const int p0 = lane * cpl;
  TokenSink cs;
  cs.out = nullptr;
  cs.pos = 0;
  cs.cap = b.T;
  uint32_t gv[8];  // grid values of my cells (cpl <= 8: windows up to 15x15)
  for (int j = 0; j < cpl; j++) {
    gv[j] = 0;
    const int p = p0 + j;
    if (p < ncells) {
      const uint32_t loc = b.pattern[p];
      const int r = my_r + (int)(loc >> 4) - hr, c = my_c + (int)(loc & 0xF) - wr;
      if (r >= 0 && r < b.H && c >= 0 && c < b.W) gv[j] = grid[r * b.W + c];
    }
    if (gv[j] != 0) sink_object(cs, b, cfg, gv[j], current_step, agents, asms, chests);
  }
  int total;
  const int offset = warp_exclusive_scan(cs.pos, lane, &total);
  if (row != nullptr) {
    cs.out = row;
    cs.pos = n_global + offset;
    for (int j = 0; j < cpl; j++) {
      if (gv[j] == 0) continue;
      if (cs.pos >= cs.cap) break;
      cs.loc = b.pattern[p0 + j];
      sink_object(cs, b, cfg, gv[j], current_step, agents, asms, chests);
    }
  }
  return n_global + total;
}

#endif  // MG_OBS_CUH_
