// mg_device.cuh -- per-environment game rules as device functions.
//
// These run inside mg_step_actions_kernel (one warp per environment). The ordered action pass
// is executed by one lane on agent records staged in shared memory; everything that touches a
// single agent (regen, rewards) is done one agent per lane. Each function names the reference
// code whose behaviour it reproduces.
#ifndef MG_DEVICE_CUH_
#define MG_DEVICE_CUH_

#include "mg_platform.h"
#include "mg_state.h"
#include "mg_tokens.cuh"

#define RS MG_RES_STRIDE

// The rules below run once per agent and step on single lanes; their loops are short and data dependent.
// Left to itself the compiler unrolls them, and the step kernel's code outgrows the instruction cache
// (24 warps per SM, each somewhere else in the rules): keeping every loop here rolled (MG_ROLLED) took the
// kernel's code from 429 KB to 343 KB and its time from 106 us to 101 us (B200, arena, 4096 envs).
// MG_INLINE_ACTIONS: bit 0 inlines action_begin / action_end into their callers, bit 1 assembler_use_warp (the
// context then is a known shared-memory block instead of a generic pointer). Measured on B200, arena steady
// state: 910 M agent-steps/s with both out of line, 954 M with both inlined.
#ifdef MG_SIMT_EMU
#undef MG_INLINE_ACTIONS
#define MG_INLINE_ACTIONS 0
#endif
#ifndef MG_INLINE_ACTIONS
#define MG_INLINE_ACTIONS 3
#endif
#if MG_INLINE_ACTIONS & 1
#define MG_ACTION_FN __forceinline__
#else
#define MG_ACTION_FN MG_NOINLINE
#endif
#if MG_INLINE_ACTIONS & 2
#define MG_USE_FN __forceinline__
#else
#define MG_USE_FN MG_NOINLINE
#endif

struct EnvCtx {
  const int32_t* cfg;
  int env, H, W, A, R, S;
  uint32_t current_step;
  uint16_t* grid;     // global, this env
  AgentRec* ag;       // shared memory copy of this env's agents
  AsmRec* asms;       // global
  ChestRec* chests;   // global
  float* astats;      // global [A][S]
  float* hot;         // shared copy of the first MG_HOT_STATS slots of every agent (null: use hotg)
  float* hotg;        // global [A][MG_HOT_STATS]: where those slots live in HBM
  float* gstats;      // global [GS]
  int n_asm;
  int clip_enabled, clip_length_scale;
  uint32_t req_mask[3], con_mask[3];  // per action kind: which resources are required / consumed
  bool concurrent;    // several lanes of the warp are running handle_action for different agents
  const MgBatch* batch;
  bool asm_records_on_use;  // assembler token records are rewritten where an assembler changes (else: all, every step)
  int inv_masks_off;  // word offset of the limit member/modifier masks behind the config (mg_create)
};

// The per-step counters of an agent (action success/failure, frozen ticks, motion) live in the first
// MG_HOT_STATS slots; the step kernel stages them in shared memory so that the ordered action pass
// does not wait on HBM for every agent.
MG_DEVICE inline float* agent_stat(const EnvCtx& e, int ai, int slot) {
  if (slot < MG_HOT_STATS) return (e.hot != nullptr ? e.hot : e.hotg) + ai * MG_HOT_STATS + slot;
  return e.astats + (size_t)ai * e.S + slot;
}

// ---------------------------------------------------------------- inventory
// The reference keeps each inventory as unordered_map<item, amount> plus shared limit records
// (objects/inventory.hpp:13-43). Here: amounts in a dense u16 array, the map's iteration order as
// a list of 4-bit item ids in one u64, and limit totals recomputed from the amounts.
struct InvView {
  uint16_t* amt;
  uint64_t* order;
  const int32_t* invcfg;  // INVCFG record
  const int32_t* masks;   // per limit of that record: [MG_MAX_LIMITS] items sharing it, [MG_MAX_LIMITS] items that modify it (bit per item)
  int owner_kind;         // 0 agent, 1 chest
  int owner_idx;
};

MG_DEVICE inline InvView agent_inv(const EnvCtx& e, int ai) {
  InvView v;
  AgentRec* a = &e.ag[ai];
  v.amt = a->amt;
  v.order = &a->order;
  const int32_t* grp = e.cfg + e.cfg[MGC_OFF_GROUPS] + a->group * MG_GRP_WORDS;
  v.invcfg = e.cfg + e.cfg[MGC_OFF_INVCFG] + grp[MG_GRP_INVCFG] * MG_INVCFG_WORDS;
  v.masks = e.cfg + e.inv_masks_off + grp[MG_GRP_INVCFG] * (2 * MG_MAX_LIMITS);
  v.owner_kind = 0;
  v.owner_idx = ai;
  return v;
}

MG_DEVICE inline InvView chest_inv(const EnvCtx& e, int ci) {
  InvView v;
  ChestRec* c = &e.chests[ci];
  v.amt = c->amt;
  v.order = &c->order;
  const int32_t* cell = e.cfg + e.cfg[MGC_OFF_CELLTYPES] + c->cell * MG_CELL_WORDS;
  const int32_t* ct = e.cfg + e.cfg[MGC_OFF_CHEST_TYPES] + cell[MG_CELL_PARAM] * MG_CHEST_WORDS;
  v.invcfg = e.cfg + e.cfg[MGC_OFF_INVCFG] + ct[MG_CHEST_INVCFG] * MG_INVCFG_WORDS;
  v.masks = e.cfg + e.inv_masks_off + ct[MG_CHEST_INVCFG] * (2 * MG_MAX_LIMITS);
  v.owner_kind = 1;
  v.owner_idx = ci;
  return v;
}

// SharedInventoryLimit::amount: kept by the reference as a running u16 sum (inventory.cpp:66-69).
MG_DEVICE inline uint32_t inv_limit_amount(const InvView& v, int limit) {
  uint32_t sum = 0;
  MG_ROLLED
  for (uint32_t m = (uint32_t)v.masks[limit]; m != 0u; m &= m - 1u) sum += v.amt[__ffs((int)m) - 1];
  return sum & 0xFFFFu;
}

// SharedInventoryLimit::effective_limit, objects/inventory.hpp:21-34
MG_DEVICE inline int inv_effective_limit(const InvView& v, int limit) {
  int effective = v.invcfg[MG_INVCFG_LIMIT_BASE + limit];
  const int32_t* mod = v.invcfg + MG_INVCFG_LIMIT_MOD + limit * RS;
  MG_ROLLED
  for (uint32_t m = (uint32_t)v.masks[MG_MAX_LIMITS + limit]; m != 0u; m &= m - 1u) {
    const int i = __ffs((int)m) - 1;
    effective += (int)v.amt[i] * mod[i];
  }
  effective = effective < 0 ? 0 : effective;
  return effective > 65535 ? 65535 : effective;
}

// Inventory::free_space, inventory.cpp:95-109
MG_DEVICE inline int inv_free_space(const InvView& v, int item) {
  int limit = v.invcfg[MG_INVCFG_ITEM_LIMIT + item];
  if (limit < 0) return 65535 - (int)v.amt[item];
  int used = (int)inv_limit_amount(v, limit);
  int effective = inv_effective_limit(v, limit);
  return effective > used ? effective - used : 0;
}

MG_DEVICE inline void order_push_front(uint64_t* order, int item) { *order = (*order << 4) | (uint64_t)item; }

MG_DEVICE inline void order_erase(uint64_t* order, int item) {
  // find the nibble equal to `item` (ids are 0..12, 0xF ends the list): xor with the id in every nibble,
  // then the classic zero-nibble test; the lowest hit is the entry (an id occurs at most once)
  const uint64_t o = *order;
  const uint64_t x = o ^ ((uint64_t)item * 0x1111111111111111ull);
  const uint64_t hit = (x - 0x1111111111111111ull) & ~x & 0x8888888888888888ull;
  if (hit == 0ull) return;
  const int k = (__ffsll((long long)hit) - 1) >> 2;  // nibble index
  const uint64_t low = k == 0 ? 0ull : (o & ((1ull << (4 * k)) - 1ull));
  const uint64_t high = k == 15 ? 0ull : ((o >> (4 * (k + 1))) << (4 * k));
  *order = low | high | (0xFull << 60);
}

MG_DEVICE int inv_update(EnvCtx& e, InvView& v, int item, int attempted_delta, bool ignore_limits, int depth);

// Agent::on_inventory_change (agent.cpp:74-101), Chest::on_inventory_change (chest.hpp:54-63)
MG_DEVICE inline void on_inventory_change(EnvCtx& e, const InvView& v, int item, int delta) {
  if (v.owner_kind == 0) {
    float* st = e.astats + (size_t)v.owner_idx * e.S;
    AgentRec* a = &e.ag[v.owner_idx];
    a->pad0 = 1;  // token record is stale
    int amount = a->amt[item];
    if (delta > 0) MG_STAT_BUMP(&st[MG_AS_GAINED + item], (float)delta);
    else MG_STAT_BUMP(&st[MG_AS_LOST + item], (float)(-delta));
    st[MG_AS_AMOUNT + item] = (float)amount;
    const int32_t* grp = e.cfg + e.cfg[MGC_OFF_GROUPS] + a->group * MG_GRP_WORDS;
    if (grp[MG_GRP_DIVERSITY + item]) {
      // presence flips exactly when the amount crosses zero
      bool has = amount > 0, had = (amount - delta) > 0;
      if (had != has) {
        int n = 0;
        MG_ROLLED
        for (int i = 0; i < MG_MAX_RESOURCES; i++) n += (grp[MG_GRP_DIVERSITY + i] && a->amt[i] > 0) ? 1 : 0;
        *agent_stat(e, v.owner_idx, MG_AS_DIVERSITY) = (float)n;
      }
    }
  } else {
    if (delta > 0) MG_RED_ADD(&e.gstats[MG_GS_CHEST_DEPOSITED + item], (float)delta);
    else MG_RED_ADD(&e.gstats[MG_GS_CHEST_WITHDRAWN + item], (float)(-delta));
    e.gstats[MG_GS_CHEST_AMOUNT + item] = (float)v.amt[item];
  }
}

// Inventory::enforce_all_limits, inventory.cpp:128-158 (limit_defs order; the reference's
// pointer-ordered std::set is not reproducible)
MG_DEVICE MG_NOINLINE void inv_enforce_all_limits(EnvCtx& e, InvView& v, int depth) {
  int nlim = v.invcfg[MG_INVCFG_NUM_LIMITS];
  int norder = v.invcfg[MG_INVCFG_NUM_ORDER];
  MG_ROLLED
  for (int limit = 0; limit < nlim; limit++) {
    bool reachable = false;
    MG_ROLLED
    for (int k = 0; k < norder; k++)
      if (v.invcfg[MG_INVCFG_ITEM_LIMIT + v.invcfg[MG_INVCFG_ORDER + k]] == limit) reachable = true;
    if (!reachable) continue;
    int excess = (int)inv_limit_amount(v, limit) - inv_effective_limit(v, limit);
    if (excess <= 0) continue;
    MG_ROLLED
    for (int k = 0; k < norder; k++) {
      int item = v.invcfg[MG_INVCFG_ORDER + k];
      if (v.invcfg[MG_INVCFG_ITEM_LIMIT + item] != limit) continue;
      int current = v.amt[item];
      int to_drop = current < excess ? current : excess;
      if (to_drop > 0) {
        inv_update(e, v, item, -to_drop, false, depth + 1);
        excess = (int)inv_limit_amount(v, limit) - inv_effective_limit(v, limit);
      }
      if (excess <= 0) break;
    }
  }
}

// Inventory::update, inventory.cpp:37-85
MG_DEVICE MG_NOINLINE int inv_update(EnvCtx& e, InvView& v, int item, int attempted_delta, bool ignore_limits, int depth) {
  int initial = v.amt[item];
  int new_amount = initial + attempted_delta;
  int max = 65535;
  int limit = v.invcfg[MG_INVCFG_ITEM_LIMIT + item];
  if (!ignore_limits && limit >= 0) {
    int effective = inv_effective_limit(v, limit);
    int used_by_others = (int)inv_limit_amount(v, limit) - initial;
    if (used_by_others < 0) used_by_others = 0;
    int max_int = effective - used_by_others;
    max = max_int < 0 ? 0 : max_int;
  }
  int clamped = new_amount < 0 ? 0 : (new_amount > max ? max : new_amount);
  if (clamped == 0) {
    if (initial > 0) order_erase(v.order, item);
  } else if (initial == 0) {
    order_push_front(v.order, item);  // new hash node: head of the list (ids < 13 never share a bucket)
  }
  v.amt[item] = (uint16_t)clamped;
  int clamped_delta = clamped - initial;
  if (clamped_delta != 0) on_inventory_change(e, v, item, clamped_delta);
  if (clamped_delta < 0 && v.invcfg[MG_INVCFG_IS_MODIFIER + item] && depth < 8) inv_enforce_all_limits(e, v, depth);
  return clamped_delta;
}

// The inventories that take part in an assembler use, without an array of views (arrays indexed at
// run time live in per-thread local memory, whose lane-interleaved layout makes a single lane's
// accesses miss L1 on every word): up to 8 agents as 8-bit indices packed in a u64, then chests.
struct InvList {
  uint64_t agents;
  int n_agents, n_chests;
  const uint16_t* chests;  // only with chest_search_distance > 0
};
MG_DEVICE inline int inv_list_size(const InvList& l) { return l.n_agents + l.n_chests; }
MG_DEVICE inline InvView inv_list_view(const EnvCtx& e, const InvList& l, int i) {
  if (i < l.n_agents) return agent_inv(e, (int)((l.agents >> (8 * i)) & 0xFFull));
  return chest_inv(e, (int)l.chests[i - l.n_agents]);
}

// HasInventory::shared_update, has_inventory.cpp:7-74
#define MG_MAX_SHARED 40
MG_DEVICE MG_NOINLINE int shared_update(EnvCtx& e, const InvList& l, int item, int delta) {
  const int n = inv_list_size(l);
  if (n == 0) return 0;
  if (n == 1) {  // one inventory: both branches below come down to this single update with the whole delta
    InvView v = inv_list_view(e, l, 0);
    return inv_update(e, v, item, delta, false, 0);
  }
  int delta_remaining = delta;
  uint64_t active = n >= 64 ? ~0ull : ((1ull << n) - 1ull);  // inventories still to consider, in order
  int num_remaining = n;
  int per = delta_remaining / num_remaining;
  MG_ROLLED
  for (;;) {
    uint64_t next = 0;
    int ncur = 0, nnext = 0;
    MG_ROLLED
    for (int i = 0; i < n; i++) {
      if (!((active >> i) & 1ull)) continue;
      ncur++;
      InvView v = inv_list_view(e, l, i);
      bool now;
      if (delta_remaining > 0) now = inv_free_space(v, item) <= per;
      else now = (int)v.amt[item] <= -per;
      if (now) {
        delta_remaining -= inv_update(e, v, item, per, false, 0);
        num_remaining--;
        if (num_remaining > 0) per = delta_remaining / num_remaining;
      } else {
        next |= 1ull << i;
        nnext++;
      }
    }
    active = next;
    if (ncur == nnext) break;
  }
  if (num_remaining == 0) return delta - delta_remaining;
  int cnt = 0;
  MG_ROLLED
  for (int i = 0; i < n; i++) cnt += (int)((active >> i) & 1ull);
  MG_ROLLED
  for (int i = n - 1; i >= 0; i--) {  // reverse order: earlier inventories get the remainder
    if (!((active >> i) & 1ull)) continue;
    int d = delta_remaining / cnt;
    InvView v = inv_list_view(e, l, i);
    delta_remaining -= inv_update(e, v, item, d, false, 0);
    cnt--;
  }
  return delta - delta_remaining;
}

// HasInventory::transfer_resources, has_inventory.cpp:77-108
MG_DEVICE inline int transfer_resources(EnvCtx& e, InvView& src, InvView& dst, int item, int delta, bool destroy) {
  if (delta <= 0) return 0;
  int avail = src.amt[item];
  int can_give = avail < delta ? avail : delta;
  int can_recv = inv_free_space(dst, item);
  int amount = can_give < can_recv ? can_give : can_recv;
  int loss = destroy ? can_give : amount;
  inv_update(e, src, item, -loss, false, 0);
  inv_update(e, dst, item, amount, false, 0);
  return amount;
}

// ---------------------------------------------------------------- grid helpers
MG_DEVICE inline uint16_t grid_at(const EnvCtx& e, int r, int c) { return e.grid[r * e.W + c]; }
MG_DEVICE inline const int32_t* proto_cfg(const int32_t* cfg, int p) { return cfg + cfg[MGC_OFF_PROTOCOLS] + p * MG_PROTO_WORDS; }
MG_DEVICE inline const int32_t* cell_cfg(const int32_t* cfg, int cell) { return cfg + cfg[MGC_OFF_CELLTYPES] + cell * MG_CELL_WORDS; }
MG_DEVICE inline const int32_t* asm_type_cfg(const int32_t* cfg, int cell) {
  return cfg + cfg[MGC_OFF_ASM_TYPES] + cell_cfg(cfg, cell)[MG_CELL_PARAM] * MG_ASM_WORDS;
}

// Assembler::get_local_vibe + surrounding agent count (assembler.hpp:30-45,353-375)
MG_DEVICE inline uint64_t asm_local_vibe(const EnvCtx& e, int r, int c, int* num_agents) {
  // the eight cells first (eight loads in flight), then what stands there; row-major order as in the reference
  uint16_t cell[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int rr = r + (int)((0xA940u >> (2 * k)) & 3u) - 1, cc = c + (int)((0x9224u >> (2 * k)) & 3u) - 1;
    cell[k] = (rr < 0 || cc < 0 || rr >= e.H || cc >= e.W) ? (uint16_t)0 : grid_at(e, rr, cc);
  }
  uint64_t key = 0;  // non-zero vibes, sorted, smallest in the highest used byte (inserted one by one)
  int nv = 0, na = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    if (MG_GRID_KIND(cell[k]) != MG_GK_AGENT) continue;
    na++;
    const uint32_t vb = e.ag[MG_GRID_INDEX(cell[k])].vibe;
    if (vb == 0u) continue;
    int gt = 0;  // entries already in the key that are larger: they sit at the low end
    while (gt < nv && ((key >> (8 * gt)) & 0xFFull) > vb) gt++;
    const uint64_t low = gt == 0 ? 0ull : (key & ((1ull << (8 * gt)) - 1ull));
    const uint64_t high = gt >= 7 ? 0ull : ((key >> (8 * gt)) << (8 * (gt + 1)));
    key = high | ((uint64_t)vb << (8 * gt)) | low;
    nv++;
  }
  *num_agents = na;
  return key;
}

MG_DEVICE inline int find_protocol(const int32_t* cfg, int first, int n, uint64_t key, int num_agents) {
  MG_ROLLED
  for (int i = first; i < first + n; i++) {
    const int32_t* p = proto_cfg(cfg, i);
    uint64_t k = ((uint64_t)(uint32_t)p[MG_PROTO_KEY_HI] << 32) | (uint32_t)p[MG_PROTO_KEY_LO];
    if (k == key && p[MG_PROTO_MIN_AGENTS] <= num_agents) return i;
  }
  return -1;
}

// Assembler::get_current_protocol, assembler.hpp:378-408
MG_DEVICE inline int asm_current_protocol(const EnvCtx& e, const AsmRec* a) {
  int num_agents;
  uint64_t key = asm_local_vibe(e, a->r, a->c, &num_agents);
  int first, n;
  if (a->flags & MG_ASMF_CLIPPED) {
    if (a->unclip_protocol < 0) return -1;
    first = a->unclip_protocol;
    n = 1;
  } else {
    const int32_t* t = asm_type_cfg(e.cfg, a->cell);
    first = t[MG_ASM_FIRST_PROTOCOL];
    n = t[MG_ASM_NUM_PROTOCOLS];
  }
  int p = find_protocol(e.cfg, first, n, key, num_agents);
  if (p < 0 && key != 0) p = find_protocol(e.cfg, first, n, 0, num_agents);
  return p;
}

// The same from a neighbourhood that has already been scanned (key of sorted vibes, agent count).
MG_DEVICE inline int asm_protocol_for(const EnvCtx& e, const AsmRec* a, uint64_t key, int num_agents) {
  int first, n;
  if (a->flags & MG_ASMF_CLIPPED) {
    if (a->unclip_protocol < 0) return -1;
    first = a->unclip_protocol;
    n = 1;
  } else {
    const int32_t* t = asm_type_cfg(e.cfg, a->cell);
    first = t[MG_ASM_FIRST_PROTOCOL];
    n = t[MG_ASM_NUM_PROTOCOLS];
  }
  int p = find_protocol(e.cfg, first, n, key, num_agents);
  if (p < 0 && key != 0) p = find_protocol(e.cfg, first, n, 0, num_agents);
  return p;
}

// ---------------------------------------------------------------- clipper bookkeeping (systems/clipper.hpp)
MG_DEVICE inline int asm_cheb(const AsmRec* a, const AsmRec* b) {
  int dr = (int)a->r - (int)b->r, dc = (int)a->c - (int)b->c;
  dr = dr < 0 ? -dr : dr;
  dc = dc < 0 ? -dc : dc;
  return dr > dc ? dr : dc;
}
MG_DEVICE inline uint32_t clip_infection_weight(const EnvCtx& e, const AsmRec* from, const AsmRec* to) {  // :93-97
  const int32_t* cc = e.cfg + e.cfg[MGC_OFF_CLIPPER];
  uint32_t cutoff = (uint32_t)cc[MG_CLIP_CUTOFF];
  uint32_t scaled = (uint32_t)asm_cheb(from, to) / (uint32_t)e.clip_length_scale;
  return scaled > cutoff ? 0u : (1u << (cutoff - scaled));
}
MG_DEVICE inline bool clip_adjacent(const EnvCtx& e, const AsmRec* a, const AsmRec* b) {  // compute_adjacencies :106-138
  const int32_t* cc = e.cfg + e.cfg[MGC_OFF_CLIPPER];
  if ((uint32_t)cc[MG_CLIP_PERIOD] == 0u) return false;
  uint32_t lim = (uint32_t)cc[MG_CLIP_CUTOFF] * (uint32_t)e.clip_length_scale;
  return a != b && (a->flags & MG_ASMF_TRACKED) && (b->flags & MG_ASMF_TRACKED) && (uint32_t)asm_cheb(a, b) <= lim;
}
// Clipper::clip_assembler :140-158 + Assembler::become_clipped (assembler.hpp:411-421); `sel` drawn by the caller
MG_DEVICE inline void clip_assembler(EnvCtx& e, AsmRec* t, uint32_t sel) {
  const int32_t* cc = e.cfg + e.cfg[MGC_OFF_CLIPPER];
  MG_ROLLED
  for (int i = 0; i < e.n_asm; i++) {
    AsmRec* adj = &e.asms[i];
    if (!clip_adjacent(e, t, adj)) continue;
    adj->infection_weight += clip_infection_weight(e, t, adj);
    if (adj->flags & MG_ASMF_CLIPPED) continue;
    adj->flags |= MG_ASMF_BORDER;
  }
  t->flags = (uint8_t)((t->flags & ~(MG_ASMF_BORDER | MG_ASMF_UNCLIPPED)) | MG_ASMF_CLIPPED);
  t->unclip_protocol = (int16_t)(cc[MG_CLIP_FIRST_PROTOCOL] + (int)sel);
  t->cooldown_end = e.current_step;
  t->cooldown_duration = 0;
}
MG_DEVICE inline void clip_on_unclip(EnvCtx& e, AsmRec* t) {  // :160-168
  MG_ROLLED
  for (int i = 0; i < e.n_asm; i++) {
    AsmRec* adj = &e.asms[i];
    if (clip_adjacent(e, t, adj)) adj->infection_weight -= clip_infection_weight(e, t, adj);
  }
  t->flags |= MG_ASMF_UNCLIPPED;
  if (t->infection_weight > 0) t->flags |= MG_ASMF_BORDER;
}

// ---------------------------------------------------------------- assembler use (assembler.hpp:455-526)
struct ItemList {
  int n;
  int item[RS];
  int val[RS];
};
MG_DEVICE inline void list_load(ItemList& l, const int32_t* w) {
  l.n = w[MG_LIST_N];
  MG_ROLLED
  for (int i = 0; i < l.n; i++) {
    l.item[i] = w[MG_LIST_ITEM(i)];
    l.val[i] = w[MG_LIST_VAL(i)];
  }
}
MG_DEVICE inline bool list_has_positive(const ItemList& l) {
  MG_ROLLED
  for (int i = 0; i < l.n; i++)
    if (l.val[i] > 0) return true;
  return false;
}

MG_DEVICE MG_NOINLINE bool assembler_on_use(EnvCtx& e, int asm_idx, int actor) {
#ifdef MG_TIMING
  long long tk0 = clock64(), tk1 = 0, tk2 = 0, tk3 = 0;
#endif
  AsmRec* a = &e.asms[asm_idx];
  const int32_t* t = asm_type_cfg(e.cfg, a->cell);
  uint32_t max_uses = (uint32_t)t[MG_ASM_MAX_USES];
  if (max_uses > 0 && a->uses_count >= max_uses) return false;
  uint32_t remaining = a->cooldown_end <= e.current_step ? 0u : a->cooldown_end - e.current_step;  // :318-323
  bool partial = t[MG_ASM_ALLOW_PARTIAL] != 0;
  if (remaining > 0 && !partial) return false;
  // One pass over the eight cells around the assembler, in the ring order of get_surrounding_agents
  // (:48-92: NW,N,NE,W,E,SW,S,SE): who stands there (agent index per ring position, 0xFF = nobody), how
  // many, and their non-zero vibes sorted into the protocol key (get_local_vibe :353-375). The key is
  // kept sorted in a register (smallest vibe in the highest used byte), not in an array.
  // In the parallel pass none of this is needed: an agent only gets there when no other agent stands in
  // or moves into the block (step_phase_decide, rule 3), so the actor is the whole neighbourhood.
  const int ar = (int)a->r, ac = (int)a->c;
  uint64_t ring = ~0ull, key = 0ull;
  int num_agents = 0, nv = 0;
  if (e.concurrent) {
    ring = (ring & ~0xFFull) | (uint64_t)actor;
    num_agents = 1;
    key = e.ag[actor].vibe;
  }
  MG_ROLLED
  for (int k = e.concurrent ? 8 : 0; k < 8; k++) {
    const int rr = ar + (int)((0xA940u >> (2 * k)) & 3u) - 1, cc = ac + (int)((0x9224u >> (2 * k)) & 3u) - 1;
    if (rr < 0 || cc < 0 || rr >= e.H || cc >= e.W) continue;
    const uint16_t g = grid_at(e, rr, cc);
    if (MG_GRID_KIND(g) != MG_GK_AGENT) continue;
    const uint32_t idx = MG_GRID_INDEX(g);
    ring = (ring & ~(0xFFull << (8 * k))) | ((uint64_t)idx << (8 * k));
    num_agents++;
    const uint32_t vb = e.ag[idx].vibe;
    if (vb != 0u) {
      int gt = 0;  // entries already in the key that are larger: they sit at the low end
      while (gt < nv && ((key >> (8 * gt)) & 0xFFull) > vb) gt++;
      const uint64_t low = gt == 0 ? 0ull : (key & ((1ull << (8 * gt)) - 1ull));
      const uint64_t high = gt >= 7 ? 0ull : ((key >> (8 * gt)) << (8 * (gt + 1)));
      key = high | ((uint64_t)vb << (8 * gt)) | low;
      nv++;
    }
  }
  int pi = asm_protocol_for(e, a, key, num_agents);
  if (pi < 0) return false;
  const int32_t* p = proto_cfg(e.cfg, pi);
  bool clipped = (a->flags & MG_ASMF_CLIPPED) != 0;
  // scale_protocol_for_partial_usage :427-453. The scaled protocol is a rebuilt map: its lists are
  // stored in that map's iteration order; amounts are scaled on the fly (inputs up, outputs down).
  const bool scaled = remaining > 0 && partial;
  const uint32_t duration = a->cooldown_duration, elapsed = duration - remaining;
  const int32_t* lin = p + (scaled ? MG_PROTO_SCALED_INPUTS : MG_PROTO_INPUTS);
  const int32_t* lout = p + (scaled ? MG_PROTO_SCALED_OUTPUTS : MG_PROTO_OUTPUTS);
  const int n_in = lin[MG_LIST_N], n_out = lout[MG_LIST_N];
#define MG_IN_AMOUNT(i) (scaled ? (int)(uint16_t)(((uint32_t)(uint16_t)lin[MG_LIST_VAL(i)] * elapsed + duration - 1) / duration) : (int)(uint16_t)lin[MG_LIST_VAL(i)])
#define MG_OUT_AMOUNT(i) (scaled ? (int)(uint16_t)(((uint32_t)(uint16_t)lout[MG_LIST_VAL(i)] * elapsed) / duration) : (int)(uint16_t)lout[MG_LIST_VAL(i)])
  bool has_output = false;
  MG_ROLLED
  for (int i = 0; i < n_out; i++) has_output = has_output || MG_OUT_AMOUNT(i) > 0;
  if (scaled && !has_output && p[MG_PROTO_HAS_OUTPUT] && !clipped) return false;
  // the inputs: the ring again, starting at the actor's position
  int start = 0;
  MG_ROLLED
  for (int k = 0; k < 8; k++)
    if (((ring >> (8 * k)) & 0xFFull) == (uint64_t)actor) start = k;
  InvList inputs;
  inputs.agents = 0; inputs.n_agents = 0; inputs.n_chests = 0; inputs.chests = nullptr;
  MG_ROLLED
  for (int j = 0; j < 8; j++) {
    const uint64_t idx = (ring >> (8 * ((start + j) & 7))) & 0xFFull;
    if (idx != 0xFFull) inputs.agents |= idx << (8 * inputs.n_agents++);
  }
  uint16_t chest_idx[MG_MAX_SHARED - 8];
  const int dist = t[MG_ASM_CHEST_DISTANCE];
  if (dist > 0) {  // get_surrounding_chests :96-121
    inputs.chests = chest_idx;
    MG_ROLLED
    for (int dr = -dist; dr <= dist; dr++)
      MG_ROLLED
      for (int dc = -dist; dc <= dist; dc++) {
        if (dr == 0 && dc == 0) continue;
        const int rr = ar + dr, cc = ac + dc;
        if (rr < 0 || cc < 0 || rr >= e.H || cc >= e.W) continue;
        const uint16_t g = grid_at(e, rr, cc);
        if (MG_GRID_KIND(g) == MG_GK_CHEST && inputs.n_chests < MG_MAX_SHARED - 8) chest_idx[inputs.n_chests++] = (uint16_t)MG_GRID_INDEX(g);
      }
  }
  const int nin = inv_list_size(inputs);
#ifdef MG_TIMING
  tk1 = clock64();
#endif
  MG_ROLLED
  for (int i = 0; i < n_in; i++) {  // can_afford_protocol :124-137 (u16 accumulation)
    const int item = lin[MG_LIST_ITEM(i)];
    uint32_t total = 0;
    MG_ROLLED
    for (int k = 0; k < nin; k++) total = (total + inv_list_view(e, inputs, k).amt[item]) & 0xFFFFu;
    if (total < (uint32_t)MG_IN_AMOUNT(i)) return false;
  }
  InvList outputs;  // get_output_inventories :198-223
  outputs.agents = 0; outputs.n_agents = 0; outputs.n_chests = 0; outputs.chests = nullptr;
  const int nvibes = p[MG_PROTO_NUM_VIBES];
  if (nvibes > 1) {
    uint32_t used = 0;  // which of the protocol's vibe slots are taken
    MG_ROLLED
    for (int i = 0; i < inputs.n_agents && outputs.n_agents < nvibes; i++) {
      const int ai = (int)((inputs.agents >> (8 * i)) & 0xFFull);
      const int v = e.ag[ai].vibe;
      if (v == 0) continue;
      MG_ROLLED
      for (int sidx = 0; sidx < nvibes; sidx++) {
        if (!((used >> sidx) & 1u) && p[MG_PROTO_VIBES + sidx] == v) {
          used |= 1u << sidx;
          outputs.agents |= (uint64_t)ai << (8 * outputs.n_agents++);
          break;
        }
      }
    }
  }
  if (outputs.n_agents == 0) { outputs.agents = (uint64_t)actor; outputs.n_agents = 1; }
  bool can_receive = true;  // can_receive_output :143-171
  if (has_output) {
    can_receive = false;
    MG_ROLLED
    for (int i = 0; i < n_out && !can_receive; i++) {
      if (MG_OUT_AMOUNT(i) <= 0) continue;
      uint32_t free_total = 0;
      MG_ROLLED
      for (int k = 0; k < outputs.n_agents; k++)
        free_total = (free_total + (uint32_t)inv_free_space(inv_list_view(e, outputs, k), lout[MG_LIST_ITEM(i)])) & 0xFFFFu;
      if (free_total >= 1) can_receive = true;
    }
  }
  if (!can_receive && !clipped) return false;
#ifdef MG_TIMING
  tk2 = clock64();
#endif
  MG_ROLLED
  for (int i = 0; i < n_in; i++) shared_update(e, inputs, lin[MG_LIST_ITEM(i)], -MG_IN_AMOUNT(i));
  MG_ROLLED
  for (int i = 0; i < n_out; i++) {  // give_output_for_protocol :174-184
    const int item = lout[MG_LIST_ITEM(i)];
    const int distributed = shared_update(e, outputs, item, MG_OUT_AMOUNT(i));
    if (distributed > 0) {
      // integer-valued float counter: the sum does not depend on the order of the additions
      if (e.concurrent) atomicAdd(&e.gstats[MG_GS_ASM_CREATED + item], (float)distributed);  // several lanes may create the same item
      else MG_RED_ADD(&e.gstats[MG_GS_ASM_CREATED + item], (float)distributed);
    }
  }
#undef MG_IN_AMOUNT
#undef MG_OUT_AMOUNT
  a->cooldown_duration = (uint32_t)(uint16_t)p[MG_PROTO_COOLDOWN];
  a->cooldown_end = e.current_step + a->cooldown_duration;
  if (clipped) {  // become_unclipped :581-590
    a->flags &= (uint8_t)~MG_ASMF_CLIPPED;
    a->unclip_protocol = -1;
    if (e.clip_enabled && (a->flags & MG_ASMF_TRACKED)) clip_on_unclip(e, a);
  } else {
    a->uses_count++;
  }
#ifdef MG_TIMING
  tk3 = clock64();
#endif
  if (e.asm_records_on_use) write_asm_tokens(*e.batch, a, e.current_step, token_record(*e.batch, e.env, MG_GK_ASM, (uint32_t)asm_idx));
#ifdef MG_TIMING
  if (e.batch->dbg != nullptr) {  // successful uses only: set-up and scans | checks | consume and give | record
    uint32_t* d = e.batch->dbg + (size_t)e.env * 16 + 12;
    const long long tk4 = clock64();
    atomicAdd(d + 0, (uint32_t)(tk1 - tk0));
    atomicAdd(d + 1, (uint32_t)(tk2 - tk1));
    atomicAdd(d + 2, (uint32_t)(tk3 - tk2));
    atomicAdd(d + 3, (uint32_t)(tk4 - tk3));
  }
#endif
  return true;
}

// Chest::onUse / transfer_resources, chest.hpp:31-52,103-122
MG_DEVICE MG_NOINLINE bool chest_on_use(EnvCtx& e, int chest_idx, int actor) {
  ChestRec* ch = &e.chests[chest_idx];
  const int32_t* ct = e.cfg + e.cfg[MGC_OFF_CHEST_TYPES] + cell_cfg(e.cfg, ch->cell)[MG_CELL_PARAM] * MG_CHEST_WORDS;
  int nv = ct[MG_CHEST_NUM_VIBES];
  if (nv == 0) return false;
  int vibe = e.ag[actor].vibe, which = -1;
  MG_ROLLED
  for (int i = 0; i < nv; i++)
    if (ct[MG_CHEST_VIBE_IDS + i] == vibe) which = i;
  if (which < 0) return false;
  ItemList deltas;
  list_load(deltas, ct + MG_CHEST_VIBE_LISTS + which * MG_LIST_WORDS);
  InvView ai = agent_inv(e, actor), ci = chest_inv(e, chest_idx);
  float* st = e.astats + (size_t)actor * e.S;
  bool any = false;
  MG_ROLLED
  for (int i = 0; i < deltas.n; i++) {
    int item = deltas.item[i], d = deltas.val[i];
    if (d > 0) {
      int moved = transfer_resources(e, ai, ci, item, d, true);
      if (moved > 0) { any = true; MG_RED_ADD(&st[MG_AS_CHEST_DEPOSITED + item], (float)moved); }
    } else if (d < 0) {
      int moved = transfer_resources(e, ci, ai, item, -d, true);
      if (moved > 0) { any = true; MG_RED_ADD(&st[MG_AS_CHEST_WITHDRAWN + item], (float)moved); }
    }
  }
  return any;
}

// Attack::try_attack and helpers, attack.hpp:93-253
MG_DEVICE MG_NOINLINE bool try_attack(EnvCtx& e, int actor, uint16_t target_cell) {
  if (MG_GRID_KIND(target_cell) != MG_GK_AGENT) return false;
  int target = MG_GRID_INDEX(target_cell);
  AgentRec* ac = &e.ag[actor];
  AgentRec* tg = &e.ag[target];
  if (tg->frozen > 0) return false;
  const int32_t* k = e.cfg + e.cfg[MGC_OFF_ATTACK];
  MG_ROLLED
  for (int i = 0; i < MG_MAX_RESOURCES; i++)
    if ((int)ac->amt[i] < (k[MG_ATK_CONSUMED + i] & 0xFFFF)) return false;
  InvView ai = agent_inv(e, actor), ti = agent_inv(e, target);
  bool blocked = false;
  if (k[MG_ATK_HAS_DEFENSE]) {
    const int32_t* vb = e.cfg + e.cfg[MGC_OFF_VIBES] + tg->vibe * MG_VIBE_WORDS;
    int weapon = 0, armor = 0;
    MG_ROLLED
    for (int i = 0; i < MG_MAX_RESOURCES; i++) weapon += (int)ac->amt[i] * k[MG_ATK_WEAPON + i];
    MG_ROLLED
    for (int i = 0; i < MG_MAX_RESOURCES; i++) {
      int w = k[MG_ATK_ARMOR + i];
      if (w == 0) continue;
      int amount = tg->amt[i];
      if (vb[MG_VIBE_RESOURCE] == i) amount += vb[MG_VIBE_BONUS];
      armor += amount * w;
    }
    int bonus = weapon - armor > 0 ? weapon - armor : 0;
    ItemList defense;
    list_load(defense, k + MG_ATK_DEFENSE);
    bool can_defend = true;
    MG_ROLLED
    for (int i = 0; i < defense.n; i++) {
      int required = defense.val[i] + bonus;
      if ((int)tg->amt[defense.item[i]] < (required & 0xFFFF)) { can_defend = false; break; }
    }
    if (can_defend) {
      MG_ROLLED
      for (int i = 0; i < defense.n; i++) inv_update(e, ti, defense.item[i], -(defense.val[i] + bonus), false, 0);
      *agent_stat(e, actor, MG_AS_ATTACK_BLOCKED_BY + tg->group) += 1.0f;
      blocked = true;
    }
  }
  if (!blocked) {
    if (k[MG_ATK_FREEZE] > 0) { tg->frozen = (int16_t)k[MG_ATK_FREEZE]; tg->pad0 = 1; }
    ItemList l;
    list_load(l, k + MG_ATK_ACTOR_DELTA);
    MG_ROLLED
    for (int i = 0; i < l.n; i++) inv_update(e, ai, l.item[i], l.val[i], false, 0);
    list_load(l, k + MG_ATK_TARGET_DELTA);
    MG_ROLLED
    for (int i = 0; i < l.n; i++) inv_update(e, ti, l.item[i], l.val[i], false, 0);
    list_load(l, k + MG_ATK_LOOT);
    MG_ROLLED
    for (int i = 0; i < l.n; i++) {
      int amount = tg->amt[l.item[i]];
      if (amount > 0) {
        int stolen = inv_update(e, ai, l.item[i], amount, false, 0);
        inv_update(e, ti, l.item[i], -stolen, false, 0);
      }
    }
    if (ac->group == tg->group) *agent_stat(e, actor, MG_AS_ATTACK_FRIENDLY_FIRE) += 1.0f;
    else {
      *agent_stat(e, actor, MG_AS_ATTACK_HIT + tg->group) += 1.0f;
      *agent_stat(e, target, MG_AS_ATTACK_HIT_BY + ac->group) += 1.0f;
    }
  }
  MG_ROLLED
  for (int i = 0; i < MG_MAX_RESOURCES; i++) {
    int amt = k[MG_ATK_CONSUMED + i] & 0xFFFF;
    if (amt > 0) inv_update(e, ai, i, -amt, false, 0);
  }
  return true;
}

// Transfer::try_transfer, transfer.hpp:72-150
MG_DEVICE MG_NOINLINE bool try_transfer(EnvCtx& e, int actor, uint16_t target_cell) {
  AgentRec* ac = &e.ag[actor];
  const int32_t* x = e.cfg + e.cfg[MGC_OFF_TRANSFER];
  MG_ROLLED
  for (int i = 0; i < MG_MAX_RESOURCES; i++)
    if ((int)ac->amt[i] < (x[MG_XFER_REQUIRED + i] & 0xFFFF)) return false;
  if (MG_GRID_KIND(target_cell) != MG_GK_AGENT) return false;
  int target = MG_GRID_INDEX(target_cell);
  AgentRec* tg = &e.ag[target];
  if (tg->frozen > 0) return false;
  int eidx = e.cfg[e.cfg[MGC_OFF_VIBES] + ac->vibe * MG_VIBE_WORDS + MG_VIBE_TRANSFER];
  if (eidx < 0) return false;
  const int32_t* eff = x + MG_XFER_EFFECTS + eidx * MG_XFER_EFFECT_WORDS;
  ItemList ad, td;
  list_load(ad, eff + MG_XFER_ACTOR);
  list_load(td, eff + MG_XFER_TARGET);
  InvView ai = agent_inv(e, actor), ti = agent_inv(e, target);
  MG_ROLLED
  for (int i = 0; i < ad.n; i++)
    if (ad.val[i] < 0 && (int)ac->amt[ad.item[i]] < -ad.val[i]) return false;
  MG_ROLLED
  for (int i = 0; i < td.n; i++)
    if (td.val[i] < 0 && (int)tg->amt[td.item[i]] < -td.val[i]) return false;
  MG_ROLLED
  for (int i = 0; i < ad.n; i++)
    if (ad.val[i] > 0 && ad.val[i] > inv_free_space(ai, ad.item[i])) return false;
  MG_ROLLED
  for (int i = 0; i < td.n; i++)
    if (td.val[i] > 0 && td.val[i] > inv_free_space(ti, td.item[i])) return false;
  float* ast = e.astats + (size_t)actor * e.S;
  MG_ROLLED
  for (int i = 0; i < ad.n; i++)
    if (ad.val[i] != 0) {
      int actual = inv_update(e, ai, ad.item[i], ad.val[i], false, 0);
      if (actual != 0) ast[MG_AS_TRANSFER_SELF + ad.item[i]] += (float)actual;
    }
  MG_ROLLED
  for (int i = 0; i < td.n; i++)
    if (td.val[i] != 0) {
      int actual = inv_update(e, ti, td.item[i], td.val[i], false, 0);
      if (actual != 0) ast[MG_AS_TRANSFER_TO + tg->group * RS + td.item[i]] += (float)actual;
    }
  *agent_stat(e, actor, MG_AS_TRANSFER_COUNT) += 1.0f;
  return true;
}

// Results of action_begin. MG_ACT_USE_PENDING: the action came down to using assembler *asm_out and the caller
// asked to run that use itself (with the whole warp, assembler_use_warp); nothing of the use has happened yet.
#define MG_ACT_FAIL 0
#define MG_ACT_OK 1
#define MG_ACT_USE_PENDING 2
#define MG_ACT_FROZEN 3 /* counted and done: no motion bookkeeping, no success/failure stats (action_handler.hpp:100-110) */

// Move::_handle_action, move.hpp:74-141
MG_DEVICE inline int handle_move(EnvCtx& e, int actor, int arg, bool defer_use, int* asm_out) {
  AgentRec* ac = &e.ag[actor];
  // orientation.hpp:28-48: N,S,W,E,NW,NE,SW,SE packed as 2-bit (delta+1) fields
  int dr = (int)((0xA058u >> (2 * arg)) & 3u) - 1;   // -1,+1,0,0,-1,-1,+1,+1
  int dc = (int)((0x8885u >> (2 * arg)) & 3u) - 1;   //  0, 0,-1,+1,-1,+1,-1,+1
  int tr = (int)ac->r + dr, tc = (int)ac->c + dc;
  if (tr < 0 || tc < 0 || tr >= e.H || tc >= e.W) return MG_ACT_FAIL;  // u16 wrap + is_valid_location
  uint16_t target = grid_at(e, tr, tc);
  int handler = e.cfg[e.cfg[MGC_OFF_VIBES] + ac->vibe * MG_VIBE_WORDS + MG_VIBE_HANDLER];
  if (handler == MG_HANDLER_ATTACK && try_attack(e, actor, target)) return MG_ACT_OK;
  if (handler == MG_HANDLER_TRANSFER && try_transfer(e, actor, target)) return MG_ACT_OK;
  if (target == 0) {  // Grid::move_object, core/grid.hpp:70-83
    e.grid[tr * e.W + tc] = (uint16_t)(0x4000 | actor);
    e.grid[(int)ac->r * e.W + (int)ac->c] = 0;
    ac->r = (uint16_t)tr;
    ac->c = (uint16_t)tc;
    return MG_ACT_OK;
  }
  int kind = MG_GRID_KIND(target), idx = MG_GRID_INDEX(target);
  if (kind == MG_GK_AGENT) {
    AgentRec* tg = &e.ag[idx];
    if (tg->frozen > 0) {  // Grid::swap_objects :85-94
      e.grid[(int)ac->r * e.W + (int)ac->c] = target;
      e.grid[tr * e.W + tc] = (uint16_t)(0x4000 | actor);
      tg->r = ac->r;
      tg->c = ac->c;
      ac->r = (uint16_t)tr;
      ac->c = (uint16_t)tc;
      *agent_stat(e, actor, MG_AS_SWAP) += 1.0f;
      return MG_ACT_OK;
    }
    return MG_ACT_FAIL;  // Agent::onUse, agent.cpp:179-185
  }
  if (kind == MG_GK_ASM) {
    if (defer_use) { *asm_out = idx; return MG_ACT_USE_PENDING; }
    return assembler_on_use(e, idx, actor) ? MG_ACT_OK : MG_ACT_FAIL;
  }
  if (kind == MG_GK_CHEST) return chest_on_use(e, idx, actor) ? MG_ACT_OK : MG_ACT_FAIL;
  return MG_ACT_FAIL;  // walls are not Usable
}

// ActionHandler::handle_action, action_handler.hpp:99-153, in two halves around the handler's own work, so
// that an assembler use can be run by the whole warp in between (assembler_use_warp).
MG_DEVICE MG_ACTION_FN int action_begin(EnvCtx& e, int actor, int kind, int arg, bool defer_use, int* asm_out) {
  AgentRec* ac = &e.ag[actor];
  if (ac->frozen != 0) {
    *agent_stat(e, actor, MG_AS_FROZEN_TICKS) += 1.0f;
    if (ac->frozen > 0) { ac->frozen -= 1; ac->pad0 = 1; }
    return MG_ACT_FROZEN;
  }
  const int32_t* res = e.cfg + e.cfg[MGC_OFF_ACTION_RES] + kind * MG_ACTRES_WORDS;
  MG_ROLLED
  for (uint32_t m = e.req_mask[kind]; m != 0u; m &= m - 1u) {
    const int i = __ffs((int)m) - 1;
    if ((int)ac->amt[i] < (res[MG_ACTRES_REQUIRED + i] & 0xFFFF)) return MG_ACT_FAIL;
  }
  if (kind == MG_ACTION_NOOP) return MG_ACT_OK;
  if (kind == MG_ACTION_MOVE) return handle_move(e, actor, arg, defer_use, asm_out);
  ac->vibe = (uint8_t)arg;  // change_vibe.hpp:48-51
  ac->pad0 = 1;
  return MG_ACT_OK;
}

MG_DEVICE MG_ACTION_FN void action_end(EnvCtx& e, int actor, int kind, bool success) {
  AgentRec* ac = &e.ag[actor];
  if (ac->r == ac->prev_r && ac->c == ac->prev_c) {
    ac->steps_without_motion += 1;
    float* mx = agent_stat(e, actor, MG_AS_MAX_STEPS_WITHOUT_MOTION);
    if ((float)ac->steps_without_motion > *mx) *mx = (float)ac->steps_without_motion;
  } else {
    ac->steps_without_motion = 0;
  }
  ac->prev_r = ac->r;
  ac->prev_c = ac->c;
  if (success) {
    *agent_stat(e, actor, MG_AS_ACTION_SUCCESS + kind) += 1.0f;
    if (e.con_mask[kind] != 0u) {
      const int32_t* res = e.cfg + e.cfg[MGC_OFF_ACTION_RES] + kind * MG_ACTRES_WORDS;
      InvView ai = agent_inv(e, actor);
      MG_ROLLED
      for (uint32_t m = e.con_mask[kind]; m != 0u; m &= m - 1u) {
        const int i = __ffs((int)m) - 1;
        inv_update(e, ai, i, -(res[MG_ACTRES_CONSUMED + i] & 0xFFFF), false, 0);
      }
    }
  } else {
    *agent_stat(e, actor, MG_AS_ACTION_FAILED + kind) += 1.0f;
    *agent_stat(e, actor, MG_AS_ACTION_FAILED_ANY) += 1.0f;
  }
}

// ---------------------------------------------------------------- assembler use by the whole warp
// Inventory::update (inventory.cpp:37-85) for an agent, all 32 lanes with the same arguments: the limit's
// member sum and its modifier bonus are one warp sum each (a lane per resource) instead of two loops.
MG_DEVICE inline int inv_update_warp(EnvCtx& e, InvView& v, int item, int attempted_delta, int lane) {
  const int initial = v.amt[item];
  int max = 65535;
  const int limit = v.invcfg[MG_INVCFG_ITEM_LIMIT + item];
  if (limit >= 0) {
    const uint32_t members = (uint32_t)v.masks[limit], mods = (uint32_t)v.masks[MG_MAX_LIMITS + limit];
    const int mine = lane < MG_MAX_RESOURCES ? (int)v.amt[lane] : 0;
    const int used = MG_WARP_SUM(((members >> lane) & 1u) ? mine : 0) & 0xFFFF;
    int effective = v.invcfg[MG_INVCFG_LIMIT_BASE + limit];
    if (mods != 0u) effective += MG_WARP_SUM(((mods >> lane) & 1u) ? mine * v.invcfg[MG_INVCFG_LIMIT_MOD + limit * RS + lane] : 0);
    effective = effective < 0 ? 0 : (effective > 65535 ? 65535 : effective);
    int others = used - initial;
    others = others < 0 ? 0 : others;
    max = effective - others;
    max = max < 0 ? 0 : max;
  }
  const int want = initial + attempted_delta;
  const int clamped = want < 0 ? 0 : (want > max ? max : want);
  const int delta = clamped - initial;
  __syncwarp();  // every lane has read the amounts before lane 0 changes them
  if (lane == 0) {
    if (clamped == 0) {
      if (initial > 0) order_erase(v.order, item);
    } else if (initial == 0) {
      order_push_front(v.order, item);
    }
    v.amt[item] = (uint16_t)clamped;
    if (delta != 0) on_inventory_change(e, v, item, delta);
    if (delta < 0 && v.invcfg[MG_INVCFG_IS_MODIFIER + item]) inv_enforce_all_limits(e, v, 0);
  }
  __syncwarp();
  return delta;
}

// HasInventory::shared_update (has_inventory.cpp:7-74) for a negative delta over 2..8 agent inventories, all 32
// lanes with the same arguments: the reference's rounds (whoever holds no more than an equal share gives all
// of it, the share of the others grows, until nobody drops out; the rest is split from the back) walked by
// every lane alike, each single update by inv_update_warp.
MG_DEVICE inline int shared_consume_warp(EnvCtx& e, uint64_t agents, int n, int item, int delta, int lane) {
  int delta_remaining = delta;
  uint32_t active = (1u << n) - 1u;
  int num_remaining = n;
  int per = delta_remaining / num_remaining;
  MG_ROLLED
  for (;;) {
    uint32_t next = 0u;
    int ncur = 0, nnext = 0;
    MG_ROLLED
    for (int i = 0; i < n; i++) {
      if (!((active >> i) & 1u)) continue;
      ncur++;
      InvView v = agent_inv(e, (int)((agents >> (8 * i)) & 0xFFull));
      if ((int)v.amt[item] <= -per) {
        delta_remaining -= inv_update_warp(e, v, item, per, lane);
        num_remaining--;
        if (num_remaining > 0) per = delta_remaining / num_remaining;
      } else {
        next |= 1u << i;
        nnext++;
      }
    }
    active = next;
    if (ncur == nnext) break;
  }
  if (num_remaining == 0) return delta - delta_remaining;
  int cnt = __popc(active);
  MG_ROLLED
  for (int i = n - 1; i >= 0; i--) {  // reverse order: earlier inventories get the remainder
    if (!((active >> i) & 1u)) continue;
    const int d = delta_remaining / cnt;
    InvView v = agent_inv(e, (int)((agents >> (8 * i)) & 0xFFull));
    delta_remaining -= inv_update_warp(e, v, item, d, lane);
    cnt--;
  }
  return delta - delta_remaining;
}

// Assembler::onUse (assembler.hpp:455-526) called by all 32 lanes with the same arguments. The common case
// -- a one-protocol assembler that is neither clipped, nor cooling down, nor searching chests -- runs here
// with a lane per ring cell / per protocol item, whether the actor is alone at the assembler or other agents
// stand around it and pay their share (shared_consume_warp); everything else is handed to assembler_on_use
// on lane 0, unchanged. `alone`: the caller knows that no other agent stands next to the assembler
// (step_phase_decide, rule 3).
MG_DEVICE MG_USE_FN bool assembler_use_warp(EnvCtx& e, int asm_idx, int actor, bool alone, int lane) {
#ifdef MG_TIMING
  const long long tk0 = clock64();
#endif
  AsmRec* a = &e.asms[asm_idx];
  // everything the type contributes, in one line: lane i holds word i of the cell code's record
  const int32_t fw = e.batch->asm_fast[(int)a->cell * MG_AF_WORDS + lane];
#define MG_AF(i) __shfl_sync(MG_FULL_MASK, fw, (i))
  const uint32_t max_uses = (uint32_t)MG_AF(MG_AF_MAX_USES);
  if (max_uses > 0 && a->uses_count >= max_uses) return false;
  const uint32_t remaining = a->cooldown_end <= e.current_step ? 0u : a->cooldown_end - e.current_step;
  const bool partial = MG_AF(MG_AF_PARTIAL) != 0;
  if (remaining > 0 && !partial) return false;
  bool simple = MG_AF(MG_AF_VALID) != 0 && remaining == 0 && (a->flags & MG_ASMF_CLIPPED) == 0 && MG_AF(MG_AF_CHEST_DISTANCE) <= 0;
  uint64_t payers = (uint64_t)actor;  // the agents around the assembler, from the actor on in ring order (8 bits each)
  int n_payers = 1;
  if (simple && !alone) {  // who stands in the ring? One cell per lane (get_surrounding_agents :48-92).
    int here = 0xFF;
    if (lane < 8) {
      const int rr = (int)a->r + (int)((0xA940u >> (2 * lane)) & 3u) - 1, cc = (int)a->c + (int)((0x9224u >> (2 * lane)) & 3u) - 1;
      if (rr >= 0 && cc >= 0 && rr < e.H && cc < e.W) {
        const uint16_t g = grid_at(e, rr, cc);
        if (MG_GRID_KIND(g) == MG_GK_AGENT) here = (int)MG_GRID_INDEX(g);
      }
    }
    const uint32_t present = __ballot_sync(MG_FULL_MASK, here != 0xFF);
    if (__popc(present) > 1) {
      const uint32_t at_actor = __ballot_sync(MG_FULL_MASK, here == actor);
      const int start = at_actor != 0u ? __ffs((int)at_actor) - 1 : 0;
      payers = 0ull;
      n_payers = 0;
      MG_ROLLED
      for (int j = 0; j < 8; j++) {
        const int idx = __shfl_sync(MG_FULL_MASK, here, (start + j) & 7);
        if (idx != 0xFF) payers |= (uint64_t)idx << (8 * n_payers++);
      }
    }
  }
  if (!simple) {  // the general case, as before
    int ok = 0;
    __syncwarp();  // every lane has read the assembler's record before lane 0 changes it
    if (lane == 0) ok = assembler_on_use(e, asm_idx, actor) ? 1 : 0;
    return __shfl_sync(MG_FULL_MASK, ok, 0) != 0;
  }
  // One protocol (key 0, min_agents <= 1: what get_current_protocol :378-408 finds whatever the vibes around
  // are, directly or through its key-0 fallback); every agent around pays, from the actor on in ring order;
  // the protocol names no vibes, so the actor alone receives (get_output_inventories :198-223)
  const int n_in = MG_AF(MG_AF_N_IN), n_out = MG_AF(MG_AF_N_OUT);
  const AgentRec* ac = &e.ag[actor];
  // lanes [0, n_in): inputs; lanes [16, 16 + n_out): outputs
  const int my_item = __shfl_sync(MG_FULL_MASK, fw, lane < 16 ? MG_AF_IN_ITEM + (lane & 3) : MG_AF_OUT_ITEM + (lane & 3));
  const int my_amount = __shfl_sync(MG_FULL_MASK, fw, lane < 16 ? MG_AF_IN_AMOUNT + (lane & 3) : MG_AF_OUT_AMOUNT + (lane & 3));
  if (n_payers == 1) {
    const bool short_of = lane < n_in && (int)ac->amt[my_item] < my_amount;  // can_afford_protocol :124-137
    if (__ballot_sync(MG_FULL_MASK, short_of) != 0u) return false;
  } else {
    // everybody around pays: the u16 sum over the payers per input, input i on lanes [8 i, 8 i + 8), payer k on lane 8 i + k
    const int i = lane >> 3, k = lane & 7;
    const int item = __shfl_sync(MG_FULL_MASK, my_item, i), amount = __shfl_sync(MG_FULL_MASK, my_amount, i);
    uint32_t total = (i < n_in && k < n_payers) ? (uint32_t)e.ag[(int)((payers >> (8 * k)) & 0xFFull)].amt[item] : 0u;
    total += __shfl_xor_sync(MG_FULL_MASK, total, 1);
    total += __shfl_xor_sync(MG_FULL_MASK, total, 2);
    total += __shfl_xor_sync(MG_FULL_MASK, total, 4);
    const bool short_of = i < n_in && (total & 0xFFFFu) < (uint32_t)amount;
    if (__ballot_sync(MG_FULL_MASK, short_of) != 0u) return false;
  }
  const bool gives = lane >= 16 && lane - 16 < n_out && my_amount > 0;
  const uint32_t giving = __ballot_sync(MG_FULL_MASK, gives);
  InvView v = agent_inv(e, actor);
  if (giving != 0u) {  // can_receive_output :143-171: room for at least one unit of some output
    const bool room = gives && inv_free_space(v, my_item) >= 1;
    if (__ballot_sync(MG_FULL_MASK, room) == 0u) return false;
  }
#ifdef MG_TIMING
  const long long tk1 = clock64();
#endif
  for (int i = 0; i < n_in; i++) {  // consume_resources_for_protocol :228-234
    const int item = __shfl_sync(MG_FULL_MASK, my_item, i), amount = __shfl_sync(MG_FULL_MASK, my_amount, i);
    if (n_payers == 1) inv_update_warp(e, v, item, -amount, lane);
    else shared_consume_warp(e, payers, n_payers, item, -amount, lane);
  }
  for (int i = 0; i < n_out; i++) {  // give_output_for_protocol :174-184
    const int item = __shfl_sync(MG_FULL_MASK, my_item, 16 + i), amount = __shfl_sync(MG_FULL_MASK, my_amount, 16 + i);
    const int distributed = inv_update_warp(e, v, item, amount, lane);
    if (distributed > 0 && lane == 0) MG_RED_ADD(&e.gstats[MG_GS_ASM_CREATED + item], (float)distributed);
  }
#ifdef MG_TIMING
  const long long tk2 = clock64();
#endif
  const uint32_t cooldown = (uint32_t)MG_AF(MG_AF_COOLDOWN);
#undef MG_AF
  if (lane == 0) {
    a->cooldown_duration = cooldown;
    a->cooldown_end = e.current_step + cooldown;
    a->uses_count++;
    if (e.asm_records_on_use) {
      // What observers see of this assembler only changes in two places (it is neither clipped nor does its
      // protocol depend on neighbours here): the cooldown end in the record's header and, where uses are
      // limited, the remaining_uses token, which is the record's first (write_asm_tokens: [clipped][remaining_uses]...)
      uint8_t* rec = token_record(*e.batch, e.env, MG_GK_ASM, (uint32_t)asm_idx);
      *reinterpret_cast<uint32_t*>(rec + 4) = a->cooldown_end;
      if (max_uses > 0u) {
        uint32_t left = a->uses_count < max_uses ? max_uses - a->uses_count : 0u;
        left = left > 255u ? 255u : left;
        rec[MG_REC_TOKENS + 1] = (uint8_t)left;
      }
    }
  }
  __syncwarp();
#ifdef MG_TIMING
  if (lane == 0 && e.batch->dbg != nullptr) {  // set-up and checks | (unused) | consume and give | record
    uint32_t* d = e.batch->dbg + (size_t)e.env * 16 + 12;
    const long long tk3 = clock64();
    atomicAdd(d + 0, (uint32_t)(tk1 - tk0));
    atomicAdd(d + 2, (uint32_t)(tk2 - tk1));
    atomicAdd(d + 3, (uint32_t)(tk3 - tk2));
  }
#endif
  return true;
}

// The whole of handle_action on one lane (paths that never hand an assembler use to the warp).
MG_DEVICE inline bool handle_action(EnvCtx& e, int actor, int kind, int arg) {
  int unused = -1;
  const int code = action_begin(e, actor, kind, arg, false, &unused);
  if (code == MG_ACT_FROZEN) return false;
  action_end(e, actor, kind, code == MG_ACT_OK);
  return code == MG_ACT_OK;
}

// inventory regeneration for one agent, mettagrid_c.cpp:625-640
MG_DEVICE inline void regen_agent(EnvCtx& e, int ai) {
  AgentRec* a = &e.ag[ai];
  const int32_t* grp = e.cfg + e.cfg[MGC_OFF_GROUPS] + a->group * MG_GRP_WORDS;
  if (!grp[MG_GRP_HAS_REGEN]) return;
  int which = -1, fallback = -1;
  MG_ROLLED
  for (int i = 0; i < grp[MG_GRP_NUM_REGEN]; i++) {
    if (grp[MG_GRP_REGEN_VIBE + i] == (int)a->vibe) which = i;
    if (grp[MG_GRP_REGEN_VIBE + i] == 0) fallback = i;
  }
  if (which < 0) which = fallback;
  if (which < 0) return;
  const int32_t* l = grp + MG_GRP_REGEN_LISTS + which * MG_LIST_WORDS;
  InvView iv = agent_inv(e, ai);
  MG_ROLLED
  for (int i = 0; i < l[MG_LIST_N]; i++) inv_update(e, iv, l[MG_LIST_ITEM(i)], l[MG_LIST_VAL(i)], false, 0);
}

// Agent::compute_stat_rewards, agent.cpp:103-132. Returns the reward delta of this step.
MG_DEVICE inline float stat_reward_delta(EnvCtx& e, int ai) {
  AgentRec* a = &e.ag[ai];
  const int32_t* grp = e.cfg + e.cfg[MGC_OFF_GROUPS] + a->group * MG_GRP_WORDS;
  int n = grp[MG_GRP_NUM_REWARDS];
  if (n == 0) return 0.0f;
  float total = 0.0f;
  MG_ROLLED
  for (int i = 0; i < n; i++) {
    const int32_t* rw = grp + MG_GRP_REWARDS + i * MG_RWD_WORDS;
    float v = 0.0f;
    const int as = rw[MG_RWD_AGENT_STAT];
    if (as >= MG_AS_AMOUNT && as < MG_AS_AMOUNT + MG_MAX_RESOURCES) {
      // "<resource>.amount" is by construction the agent's current amount (on_inventory_change writes
      // it at every change, the initial inventory included): read it from the record, not from HBM
      v = (float)a->amt[as - MG_AS_AMOUNT];
    } else if (as >= 0) {
      const float* sp = agent_stat(e, ai, as);
      v = as < MG_HOT_STATS ? *sp : MG_LD_L2(sp);  // some of these slots are bumped with reductions (at L2)
    }
    v = MG_FADD(v, rw[MG_RWD_GAME_STAT] >= 0 ? MG_LD_L2(&e.gstats[rw[MG_RWD_GAME_STAT]]) : 0.0f);
    v = MG_FADD(v, MG_INT_AS_FLOAT(rw[MG_RWD_COLLECTIVE]));  // the agent's collective's value of the stat: a constant (mg_config.h)
    float r = MG_FMUL(v, MG_INT_AS_FLOAT(rw[MG_RWD_WEIGHT]));
    if (rw[MG_RWD_HAS_MAX]) {
      float cap = MG_INT_AS_FLOAT(rw[MG_RWD_MAX]);
      r = cap < r ? cap : r;  // std::min(stats_reward, cap)
    }
    total = MG_FADD(total, r);
  }
  float delta = MG_FADD(total, -a->current_stat_reward);
  if (delta != 0.0f) a->current_stat_reward = total;
  return delta;
}

#endif  // MG_DEVICE_CUH_
