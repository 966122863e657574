// mg_tokens.cuh -- token records: what each object shows to observers, written where the object
// changes (action kernel, reset) instead of being rebuilt by every observer.
//
// Reference feature lists: Agent::obs_features (agent.cpp:187-213), Assembler::obs_features
// (assembler.hpp:528-578), Chest::obs_features (chest.hpp:124-146), Wall::obs_features
// (wall.hpp:26-38), and the global tokens of _compute_observation (mettagrid_c.cpp:425-517).
// Record format: see MgBatch in mg_state.h.
#ifndef MG_TOKENS_CUH_
#define MG_TOKENS_CUH_

#include "mg_platform.h"
#include "mg_state.h"

MG_DEVICE inline uint32_t rec_kind_base(const MgBatch& b, uint32_t kind) {
  const uint32_t w = kind < 2u ? b.rec_kind_base_lo : b.rec_kind_base_hi;
  return (w >> ((kind & 1u) * 16u)) & 0xFFFFu;
}
MG_DEVICE inline uint8_t* token_record(const MgBatch& b, int env, uint32_t kind, uint32_t idx) {
  return b.etok + (((size_t)env * b.rec_slots + rec_kind_base(b, kind) + idx) << b.rec_shift);
}

MG_DEVICE inline uint2 make_rec_header(uint32_t n, uint32_t cooldown_end) {
  uint2 h;
  h.x = n & 0xFFu;
  h.y = cooldown_end;
  return h;
}

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
  w.out = reinterpret_cast<uint16_t*>(rec + MG_REC_TOKENS);
  w.n = 0;
  w.cap = ((1 << b.rec_shift) - MG_REC_TOKENS) >> 1;
  const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + a.group * MG_GRP_WORDS;
  tw_put(w, cfg[MGC_FEAT_GROUP], grp[MG_GRP_GROUP_ID]);
  tw_put(w, cfg[MGC_FEAT_FROZEN], a.frozen != 0 ? 1 : 0);
  if (a.vibe != 0) tw_put(w, cfg[MGC_FEAT_VIBE], a.vibe);
  tw_inventory(w, cfg, a.order, a.amt);
  tw_template(w, b, b.ix_group + a.group);
  *reinterpret_cast<uint2*>(rec) = make_rec_header((uint32_t)w.n, 0u);
}
// Chest::obs_features (chest.hpp:124-146): vibe if set, inventory, tags
MG_DEVICE inline void write_chest_tokens(const MgBatch& b, const ChestRec& c, uint8_t* rec) {
  const int32_t* cfg = b.cfg;
  TokenWriter w;
  w.out = reinterpret_cast<uint16_t*>(rec + MG_REC_TOKENS);
  w.n = 0;
  w.cap = ((1 << b.rec_shift) - MG_REC_TOKENS) >> 1;
  const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c.cell * MG_CELL_WORDS;
  if (cc[MG_CELL_VIBE] != 0) tw_put(w, cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]);
  tw_inventory(w, cfg, c.order, c.amt);
  tw_template(w, b, c.cell);
  *reinterpret_cast<uint2*>(rec) = make_rec_header((uint32_t)w.n, 0u);
}

// Assembler::obs_features (assembler.hpp:528-578): cooldown_remaining, clipped, remaining_uses, the
// current protocol's inputs and outputs, tags, vibe. Rewritten every step by the action kernel
// because cooldown_remaining changes with time.
struct AsmView;
MG_DEVICE inline void asm_view(const MgBatch& b, const int32_t* cfg, const AsmRec* a, uint32_t current_step, AsmView& v);
MG_DEVICE inline void write_asm_tokens(const MgBatch& b, const AsmRec* a, uint32_t current_step, uint8_t* rec);

// What an assembler shows (assembler.hpp:528-578): [cooldown_remaining, added by the observer][clipped]
// [remaining_uses], then the current protocol's inputs/outputs (template), then tags and vibe (template).
struct AsmView {
  uint32_t dyn[2];
  int ndyn;
  uint32_t cooldown_end;
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
  v.cooldown_end = cooldown_end;  // the cooldown_remaining token depends on the time of observation: observers add it
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
  uint16_t* out = reinterpret_cast<uint16_t*>(rec + MG_REC_TOKENS);
  const int cap = ((1 << b.rec_shift) - MG_REC_TOKENS) >> 1;
  int n = 0;
  for (int k = 0; k < v.ndyn; k++) { if (n < cap) out[n] = (uint16_t)v.dyn[k]; n++; }
  const uint16_t* ps = b.tmpl_tok + (v.proto_ix >> 8);
  for (uint32_t k = 0; k < (v.proto_ix & 0xFFu); k++) { if (n < cap) out[n] = MG_LDG(&ps[k]); n++; }
  const uint16_t* ts = b.tmpl_tok + (v.tail_ix >> 8);
  for (uint32_t k = 0; k < (v.tail_ix & 0xFFu); k++) { if (n < cap) out[n] = MG_LDG(&ts[k]); n++; }
  *reinterpret_cast<uint2*>(rec) = make_rec_header((uint32_t)n, v.cooldown_end);
}

// The same record written by the whole warp (all lanes with the same arguments): a token per lane.
MG_DEVICE inline void write_asm_tokens_warp(const MgBatch& b, const AsmRec* a, uint32_t current_step, uint8_t* rec, int lane) {
  AsmView v;
  asm_view(b, b.cfg, a, current_step, v);
  uint16_t* out = reinterpret_cast<uint16_t*>(rec + MG_REC_TOKENS);
  const int cap = ((1 << b.rec_shift) - MG_REC_TOKENS) >> 1;
  const int np = (int)(v.proto_ix & 0xFFu), nt = (int)(v.tail_ix & 0xFFu);
  const int n = v.ndyn + np + nt;
  const uint16_t* ps = b.tmpl_tok + (v.proto_ix >> 8);
  const uint16_t* ts = b.tmpl_tok + (v.tail_ix >> 8);
  for (int k = lane; k < n && k < cap; k += 32) {
    uint16_t tok;
    if (k < v.ndyn) tok = (uint16_t)(k == 0 ? v.dyn[0] : v.dyn[1]);
    else if (k < v.ndyn + np) tok = MG_LDG(&ps[k - v.ndyn]);
    else tok = MG_LDG(&ts[k - v.ndyn - np]);
    out[k] = tok;
  }
  if (lane == 0) *reinterpret_cast<uint2*>(rec) = make_rec_header((uint32_t)n, v.cooldown_end);
  __syncwarp();
}

// The tokens every agent gets at the centre of its window, then the compass token
// (mettagrid_c.cpp:425-517), as (location, feature, value) triples ready to be copied.
MG_DEVICE inline void write_global_tokens(const MgBatch& b, const AgentRec& me, uint32_t current_step, int last_action, uint8_t* rec) {
  const int32_t* cfg = b.cfg;
  const int oh = cfg[MGC_OBS_H], ow = cfg[MGC_OBS_W];
  const int hr = oh >> 1, wr = ow >> 1;
  const uint8_t centre = (uint8_t)((hr << 4) | (wr & 0xF));
  int n = 0;
  uint8_t* out = rec + 1;
  const int cap = (MG_GTOK_STRIDE - 1) / 3;
#define MG_GPUT(loc, f, v) do { if (n < cap) { out[n * 3] = (uint8_t)(loc); out[n * 3 + 1] = (uint8_t)(f); out[n * 3 + 2] = (uint8_t)(v); } n++; } while (0)
  if (cfg[MGC_GOBS_ECP]) {
    const uint32_t max_steps = (uint32_t)cfg[MGC_MAX_STEPS];
    uint32_t v = 0;
    if (max_steps > 0u) v = current_step >= max_steps ? 255u : ((256u * current_step / max_steps) & 0xFFu);
    MG_GPUT(centre, cfg[MGC_FEAT_ECP], v);
  }
  if (cfg[MGC_GOBS_LAST_ACTION]) MG_GPUT(centre, cfg[MGC_FEAT_LAST_ACTION], last_action & 0xFF);
  // rewards are zeroed before and written after the observations of a step (mettagrid_c.cpp:578,656):
  // the last_reward token is always round(0 * 100) = 0.
  if (cfg[MGC_GOBS_LAST_REWARD]) MG_GPUT(centre, cfg[MGC_FEAT_LAST_REWARD], 0);
  if (cfg[MGC_GOBS_GOAL]) {
    const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + me.group * MG_GRP_WORDS;
    for (int i = 0; i < grp[MG_GRP_NUM_GOALS]; i++) MG_GPUT(centre, cfg[MGC_FEAT_GOAL], grp[MG_GRP_GOALS + i]);
  }
  if (cfg[MGC_GOBS_COMPASS]) {
    const int dr = b.H / 2 - (int)me.r, dc = b.W / 2 - (int)me.c;
    const int sr = dr > 0 ? 1 : (dr < 0 ? -1 : 0), sc = dc > 0 ? 1 : (dc < 0 ? -1 : 0);
    if (sr != 0 || sc != 0) {
      const int orr = hr + sr, occ = wr + sc;
      if (orr >= 0 && orr < oh && occ >= 0 && occ < ow) MG_GPUT((orr << 4) | (occ & 0xF), cfg[MGC_FEAT_COMPASS], 1);
    }
  }
#undef MG_GPUT
  rec[0] = (uint8_t)n;
}

#endif  // MG_TOKENS_CUH_
