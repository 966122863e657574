// mg_state.h -- how a batch of environments lives in HBM.
//
// One record array per entity kind, indexed [env][instance]; records are small fixed-size
// structs so that a warp loads an env's agents with a few 16-byte loads per lane, and the
// observation kernel reaches any visible object with one or two sector reads.
//
// What the reference keeps per env -> where it lives here:
//   Grid::grid (core/grid.hpp:23)                     -> grid[env][H*W] u16 (encoding below)
//   Agent (objects/agent.hpp:23-79) + Inventory        -> AgentRec, 64 B
//   Assembler (objects/assembler.hpp:239-279)          -> AsmRec, 32 B
//   Chest (objects/chest.hpp) + Inventory              -> ChestRec, 48 B
//   MettaGrid scalars (bindings/mettagrid_c.hpp:60-143)-> EnvRec, 64 B
//   std::mt19937 _rng (+ the Clipper's private copy)   -> rng[env][624] u32
//   StatsTracker maps (systems/stats_tracker.hpp)      -> dense float slots (mg_config.h MG_AS_*/MG_GS_*)
#ifndef MG_STATE_H_
#define MG_STATE_H_

#include <stdint.h>

#include "../../include/mg_config.h"

// grid cell encoding (u16): 0 = empty, 1..0x3FFF = cell code of a static object (walls),
// 0x4000|i agent i, 0x8000|i assembler instance i, 0xC000|i chest instance i.
#define MG_GRID_KIND(v) ((v) >> 14)
#define MG_GRID_INDEX(v) ((v) & 0x3FFF)
#define MG_GK_STATIC 0
#define MG_GK_AGENT 1
#define MG_GK_ASM 2
#define MG_GK_CHEST 3

#define MG_HOT_STATS 16 /* agent stat slots 0..15: touched by nearly every action, kept in an array of their own */
#define MG_GTOK_STRIDE 64
#define MG_REC_TOKENS 8 /* byte offset of the tokens inside a token record */

/* asm_fast record (32 words): valid, max_uses, allow_partial, chest_distance, cooldown, n_in, n_out, protocol, then
 * input items [8..12), input amounts [12..16), output items [16..20), output amounts [20..24) */
#define MG_AF_VALID 0
#define MG_AF_MAX_USES 1
#define MG_AF_PARTIAL 2
#define MG_AF_CHEST_DISTANCE 3
#define MG_AF_COOLDOWN 4
#define MG_AF_N_IN 5
#define MG_AF_N_OUT 6
#define MG_AF_PROTOCOL 7
#define MG_AF_IN_ITEM 8
#define MG_AF_IN_AMOUNT 12
#define MG_AF_OUT_ITEM 16
#define MG_AF_OUT_AMOUNT 20
#define MG_AF_WORDS 32

#define MG_ORDER_EMPTY 0xFFFFFFFFFFFFFFFFull /* inventory order list: 4-bit item ids, 0xF ends the list */

struct alignas(16) AgentRec {
  uint16_t r, c, prev_r, prev_c;
  uint8_t vibe;
  uint8_t group;  // index of the agent config (GROUPS section)
  int16_t frozen;
  uint32_t steps_without_motion;
  uint64_t order;  // iteration order of the reference's unordered_map<item, amount>
  uint16_t amt[MG_MAX_RESOURCES];
  uint16_t pad0;  // step kernel scratch: token record needs a refresh
  float current_stat_reward;
  float episode_reward;
  uint32_t pad1;
};
static_assert(sizeof(AgentRec) == 64, "AgentRec must be 64 bytes");

#define MG_ASMF_CLIPPED 1
#define MG_ASMF_TRACKED 2
#define MG_ASMF_BORDER 4
#define MG_ASMF_UNCLIPPED 8 /* member of Clipper::unclipped_assemblers */

struct alignas(16) AsmRec {
  uint16_t r, c;
  uint8_t cell;   // cell code -> tags, type id, assembler type
  uint8_t flags;  // MG_ASMF_*
  int16_t unclip_protocol;  // protocol index while clipped, -1 none
  uint32_t cooldown_end, cooldown_duration, uses_count;
  uint32_t infection_weight;
  int32_t cache_protocol;  // current protocol as seen by observers, valid when cache_step == current_step
  uint32_t cache_step;
};
static_assert(sizeof(AsmRec) == 32, "AsmRec must be 32 bytes");

struct alignas(16) ChestRec {
  uint16_t r, c;
  uint8_t cell;
  uint8_t pad[3];
  uint64_t order;
  uint16_t amt[MG_MAX_RESOURCES];
  uint16_t pad1[3];
};
static_assert(sizeof(ChestRec) == 48, "ChestRec must be 48 bytes");

struct alignas(16) EnvRec {
  uint32_t current_step;
  uint32_t rng_idx;       // std::mt19937::_M_p
  uint32_t clip_rng_idx;
  uint32_t seed;
  int32_t map_idx;
  int32_t n_asm, n_chest;
  int32_t clip_length_scale;
  uint32_t early_reset_step;  // EarlyResetHandler (envs/early_reset_handler.py:12-20); 0 = none
  uint32_t episodes;          // episodes completed so far
  uint8_t done;               // bit 0: terminals set, bit 1: truncations set (for all agents of the env)
  uint8_t pad[3];
  uint32_t spare[5];
};
static_assert(sizeof(EnvRec) == 64, "EnvRec must be 64 bytes");

// Hand-over of finished envs from the step kernel to the observation kernel running beside it
// (mg_step_queue_kernel / mg_obs_queue_kernel): an entry is epoch << 32 | env, valid for the launch
// whose epoch it carries, so nothing has to be cleared between steps.
struct MgQueueCtl {
  uint32_t epoch;  // bumped by mg_queue_begin_kernel before every step
  uint32_t tail;   // entries published so far
  uint32_t head;   // observation tasks handed out so far
  uint32_t error;  // set if a consumer gave up waiting (never in a healthy run)
  uint32_t env_next;  // fused kernel: envs handed out so far
  uint32_t pad[3];
};

// Everything a kernel needs, passed by value.
struct MgBatch {
  int32_t num_envs, A, H, W, T, R, S, GS, max_asm, max_chest, num_maps, has_clipper;
  int32_t obs_cells;           // obs_h * obs_w
  int32_t obs_hr, obs_wr;      // obs_h / 2, obs_w / 2 (the window's radius), and
  int32_t feat_cooldown;       // the feature id of cooldown_remaining: launch parameters, so that the observation kernel does not fetch them from the config
  int32_t grid_stride;         // cells per env grid, padded to a multiple of 8 (16-byte rows)
  int32_t auto_reset, map_stride;
  int32_t env_begin, env_end;  // the slice of envs this launch works on (the step is launched in chunks)
  int32_t env_flags;           // flag mode: the step kernel marks every finished env in queue[env], the observation kernel waits per env
  int32_t after_step;          // this observation pass follows a step: envs whose episode just ended are summed into episode_accum here
  int32_t cfg_words;           // length of cfg in int32 words, including what mg_create appends to the blob
  // derived from the config once, on the host (mg_create):
  uint32_t req_mask[3], con_mask[3];  // per action kind: resources it requires / consumes (bit per item)
  int32_t any_damage;          // some agent group has damage enabled
  int32_t any_chest_search;    // some assembler type searches chests (chest_search_distance > 0): its users are ordered
  int32_t reward_stat_prefetch;  // some reward term reads a stat slot that lives in HBM (not a hot slot, not <res>.amount)
  int32_t reward_game_stats;   // some reward term reads a game stat
  int32_t inv_masks_off;       // appended: per INVCFG record [MG_MAX_LIMITS] member masks, [MG_MAX_LIMITS] modifier masks
  const int32_t* asm_fast;     // [MG_MAX_CELLTYPES][32] per map cell code: what a use of a one-protocol assembler needs, in one 128-byte line (mg_create)
  const int32_t* cfg;          // lowered config
  const uint8_t* pattern;      // Manhattan-order packed window locations (packed_coordinate.hpp:74-143)
  const int32_t* pattern_off;  // the same cells as offsets (row * W + col) from the window's top-left cell, -1 behind the last; padded to 8 per lane
  int32_t obs_slot, obs_warp_bytes;  // shared memory per observation warp: row image bytes (16-byte multiple), and with the descriptors behind it
  // static token templates, built once per config on the host (mg_abi.cu build_templates):
  // a token is feature | value << 8; an index entry is offset << 8 | count
  const uint16_t* tmpl_tok;
  const uint32_t* tmpl_ix;     // [cells: full list for walls, tags+vibe tail for assemblers, tags for chests][protocols][groups: tags]
  int32_t ix_proto, ix_group;  // section starts inside tmpl_ix
  // Token records: what an observer copies for one object. Uniform format for every kind of object:
  // byte 0 = token count n, bytes 4..7 = cooldown_end (assemblers; 0 otherwise: a cooldown_remaining
  // token is prepended while cooldown_end > current_step), n u16 tokens from byte 8.
  // One array per env, slots [static cell codes | agents | assemblers | chests], stride 1 << rec_shift.
  int32_t rec_shift, rec_slots;       // bytes per record = 1 << rec_shift; records per env
  uint32_t rec_kind_base_lo, rec_kind_base_hi;  // first slot of grid kinds 0..3, 16 bits each
  const uint8_t* static_rec;   // [cell codes][1 << rec_shift]: records of static objects, copied into every env at reset
  // map pool
  const uint16_t* pool_grid;   // [M][H*W]
  const int32_t* pool_agents;  // [M][A][3]
  const int32_t* pool_n_asm;   // [M]
  const int32_t* pool_asm;     // [M][max_asm][3]
  const int32_t* pool_n_chest; // [M]
  const int32_t* pool_chest;   // [M][max_chest][3]
  const float* pool_obj_counts; // [M][MG_MAX_CELLTYPES] walls, assemblers and chests per cell code: the objects.<name> game stats
  uint16_t* pool_tok0;         // [M][A] token count of every agent's first observation on this map (filled on first use; without a clipper it depends on the map alone)
  uint32_t* pool_tok0_valid;   // [M] 1 once pool_tok0[m] is complete
  // state
  EnvRec* env;
  uint16_t* grid;              // [N][H*W]
  AgentRec* agents;            // [N][A]
  AsmRec* asms;                // [N][max_asm]
  ChestRec* chests;            // [N][max_chest]
  float* agent_stats;          // [N][A][S]; slots below MG_HOT_STATS live in hot_stats instead
  float* hot_stats;            // [N][A][MG_HOT_STATS]: the per-step counters, dense so that a warp brings its env's in with 16-byte loads
  float* game_stats;           // [N][GS]
  uint32_t* rng;               // [N][624]
  uint32_t* clip_rng;          // [N][624] (only with a clipper)
  uint32_t* rng_seeded;        // [N][624] the generator right after std::mt19937(seed): an auto-reset (same seed) copies it instead of re-seeding
  uint8_t* etok;               // [N][rec_slots][1 << rec_shift] token records (see above)
  uint8_t* gtok;               // [N][A][MG_GTOK_STRIDE]: byte 0 = n, then n (location, feature, value) triples: the global tokens of an agent
  uint16_t* tok_counts;        // [N][A][2] written, attempted: produced by the observation kernel
  int32_t* last_action;        // [N][A] executed action (0 when it failed), mettagrid_c.cpp:595-620
  uint8_t* action_success;     // [N][A]
  int32_t* invalid_ids;        // [N][A][MG_INVALID_WORDS] out-of-range action ids per agent (action.invalid_index.<id>)
  long long* episode_accum;    // [2 + S + GS] sums over finished episodes (for infos / NCCL reduce), fixed point with 16 fraction bits:
                               // integer atomics only (double arithmetic measured ~1000 cycles per instruction on this part); mg_read converts
  MgQueueCtl* qctl;            // queue control block and entries [N] (see above)
  unsigned long long* queue;
  uint32_t* dbg;               // [N][16] cycle counters of the step kernel's phases (builds with -DMG_TIMING only, else null)
  // caller-owned buffers (set_buffers)
  uint8_t* obs;                // [N*A][T][3]
  uint8_t* terminals;          // [N*A]
  uint8_t* truncations;        // [N*A]
  float* rewards;              // [N*A]
  const int32_t* actions;      // [N*A]
  // optional, caller-owned (mg_set_obs_aux): what the policy's first op would otherwise recompute from
  // the observations (agent/src/metta/agent/components/obs_shim.py:181-196)
  int32_t* row_len;            // [N*A] tokens written per row = index of the first 0xFF location byte (T if none)
  int32_t* max_flip;           // [1]   max of row_len over the rows of the last full observation pass
  uint8_t* obs_mask;           // [N*A][T] 1 where the token slot is padding (location byte 0xFF)
};

#endif  // MG_STATE_H_
