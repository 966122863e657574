// mg_kernels.cuh -- the three kernels of a simulator step.
//
//   mg_reset_kernel   one warp per env: (re)build an env from its pool map     (episode start)
//   mg_step_kernel    one warp per env: MettaGrid::_step up to the observations (mettagrid_c.cpp:572-650)
//                     plus rewards / truncation (:655-677) and the auto-reset of
//                     MettaGridPufferEnv.step (envs/mettagrid_puffer_env.py:158-164)
//   mg_obs_kernel     one warp per agent: _compute_observations (:557-563)
//
// The last_reward token is identically 0 in the reference (see mg_obs.cuh), so computing rewards
// before the observation kernel does not change any output byte.
#ifndef MG_KERNELS_CUH_
#define MG_KERNELS_CUH_

#include "mg_device.cuh"
#include "mg_obs.cuh"
#include "mg_rng.cuh"

#include <math.h>

// -DMG_TIMING: per-env cycle counts of the step kernel's phases in EnvRec::spare (tuning builds only)
#ifdef MG_TIMING
#define MG_TICK(i) do { const long long now__ = clock64(); tick_acc[i] += (uint32_t)(now__ - tick_last); tick_last = now__; } while (0)
#define MG_SM_TICK(sm, i) do { __syncwarp(); if (lane == 0) { const long long now__ = clock64(); (sm).ticks[i] += (uint32_t)(now__ - (sm).tick_last); (sm).tick_last = now__; } } while (0)
#else
#define MG_TICK(i) do { } while (0)
#define MG_SM_TICK(sm, i) do { } while (0)
#endif

// The five phases of a step are inlined into the kernel (a bit per phase: load, decide, act, post, finish;
// -DMG_INLINE_PHASES=0 makes them out-of-line functions): inlined code reads the launch parameters from the
// constant bank and the warp's block with shared-memory loads, an out-of-line phase gets both as generic
// pointers. Measured on B200 (arena steady state): step kernel 70.4 us out of line, 61.1 us inlined.
#ifdef MG_SIMT_EMU
#undef MG_INLINE_PHASES
#define MG_INLINE_PHASES 0
#endif
#ifndef MG_INLINE_PHASES
#define MG_INLINE_PHASES 31
#endif
#if MG_INLINE_PHASES & 1
#define MG_PHASE_LOAD_FN __forceinline__
#else
#define MG_PHASE_LOAD_FN MG_NOINLINE
#endif
#if MG_INLINE_PHASES & 2
#define MG_PHASE_DECIDE_FN __forceinline__
#else
#define MG_PHASE_DECIDE_FN MG_NOINLINE
#endif
#if MG_INLINE_PHASES & 4
#define MG_PHASE_ACT_FN __forceinline__
#else
#define MG_PHASE_ACT_FN MG_NOINLINE
#endif
#if MG_INLINE_PHASES & 8
#define MG_PHASE_POST_FN __forceinline__
#else
#define MG_PHASE_POST_FN MG_NOINLINE
#endif
#if MG_INLINE_PHASES & 16
#define MG_PHASE_FINISH_FN __forceinline__
#else
#define MG_PHASE_FINISH_FN MG_NOINLINE
#endif
// The config copy in shared memory per CTA (step kernel) is off: measured slower on B200 (step kernel 128.6 us
// with it, 120.7 us without); -DMG_CFG_SMEM_LIMIT=98304 turns it on.
#ifndef MG_CFG_SMEM_LIMIT
#define MG_CFG_SMEM_LIMIT 0
#endif
#ifndef MG_STEP_WARPS
#define MG_STEP_WARPS 8    /* envs per CTA in the reset/step kernels */
#endif
#ifndef MG_OBS_WARPS
#define MG_OBS_WARPS 8     /* agents per CTA in the observation kernel */
#endif
#define MG_MAX_AGENTS 64   /* agents per env that fit the shared-memory staging */
#ifndef MG_OBS_MIN_CTAS
#define MG_OBS_MIN_CTAS 8 /* CTAs of the observation kernel resident per SM: 64 warps of 32 registers (6 CTAs: 68 us against 60 us) */
#endif
#ifndef MG_STEP_MIN_CTAS
#define MG_STEP_MIN_CTAS 3 /* CTAs of the step kernel resident per SM (80 registers): 3 552 of 4096 envs at once, the rest as a second wave */
#endif

template <int MAXA>
struct StepSmem {
  AgentRec ag[MAXA];
  float hot[MAXA * MG_HOT_STATS];  // (16-byte aligned: it follows whole 64-byte records)
  unsigned long long stage_bar;    // mbarrier of the bulk copies that stage ag and hot (step_smem_init)
  uint32_t stage_phase;            // its current phase bit
  uint32_t publish;                // the env's writes must be complete when the step returns (the fused kernel hands envs to other SMs)
  int32_t action[MAXA];
  int32_t executed[MAXA];
  uint8_t success[MAXA];
  uint8_t order[MAXA];
  int32_t scratch[2 * MG_RES_STRIDE + 8];
  uint32_t rngbuf[32];
  uint32_t rng_base, rng_n;  // rngbuf holds state words [rng_base, rng_base + rng_n) of the env's generator
  uint32_t intent[MAXA];  // per agent: bit 0 = act in the parallel pass, bits 8..15 action kind, 16..31 argument
  uint32_t ordered_mask;
  long long tick_last;
#ifdef MG_TIMING
  uint32_t ticks[16];
#endif
  EnvRec er;
  EnvCtx ctx;  // warp-uniform; lives here so that passing it to out-of-line functions costs no stack traffic
};

// n 16-byte words from src to dst by one warp, eight loads in flight per lane (a plain copy loop waits for
// every load before it issues the next).
MG_DEVICE inline void copy16_warp(uint4* dst, const uint4* src, int n, int lane) {
  for (int i0 = 0; i0 < n; i0 += 32 * 8) {
    uint4 w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int i = i0 + lane + 32 * k;
      if (i < n) w[k] = src[i];
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int i = i0 + lane + 32 * k;
      if (i < n) dst[i] = w[k];
    }
  }
}

// Ask for a whole range at once (one request per 128-byte line, spread over the lanes): what follows reads it
// in dependent steps, and would otherwise pay an HBM round trip for each of them.
MG_DEVICE inline void prefetch_range_warp(const void* p, size_t bytes, int lane) {
  const char* c = reinterpret_cast<const char*>(p);
  for (size_t off = (size_t)lane * 128; off < bytes; off += 32 * 128) MG_PREFETCH_TO_L2(c + off);
}

// An EnvRec moves as four 16-byte words, not field by field.
MG_DEVICE inline void copy_env_rec(EnvRec* dst, const EnvRec* src) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  const uint4 a = s4[0], b = s4[1], c = s4[2], d = s4[3];
  d4[0] = a; d4[1] = b; d4[2] = c; d4[3] = d;
}

// Fills `e` where it lives (the step kernel keeps it in shared memory: no copy through registers).
MG_DEVICE inline void init_ctx(EnvCtx& e, const MgBatch& b, int env, AgentRec* ag, const EnvRec& er) {
  e.cfg = b.cfg;
  e.env = env;
  e.H = b.H; e.W = b.W; e.A = b.A; e.R = b.R; e.S = b.S;
  e.current_step = er.current_step;
  e.grid = b.grid + (size_t)env * b.grid_stride;
  e.ag = ag;
  e.asms = b.asms + (size_t)env * b.max_asm;
  e.chests = b.chests + (size_t)env * b.max_chest;
  e.astats = b.agent_stats + (size_t)env * b.A * b.S;
  e.hot = nullptr;
  e.hotg = b.hot_stats + (size_t)env * b.A * MG_HOT_STATS;
  e.concurrent = false;
  e.batch = &b;
  e.asm_records_on_use = false;
  e.inv_masks_off = b.inv_masks_off;
  for (int k = 0; k < 3; k++) e.req_mask[k] = e.con_mask[k] = 0x1FFFu;
  e.gstats = b.game_stats + (size_t)env * b.GS;
  e.n_asm = er.n_asm;
  e.clip_enabled = b.has_clipper;
  e.clip_length_scale = er.clip_length_scale;
}

MG_DEVICE inline EnvCtx make_ctx(const MgBatch& b, int env, AgentRec* ag, const EnvRec& er) {
  EnvCtx e;
  init_ctx(e, b, env, ag, er);
  return e;
}

// tokens_written / tokens_dropped / tokens_free_space (mettagrid_c.cpp:552-554) are produced by the
// observation kernel as integer counts per agent and folded into the float game stats here, agent
// by agent, so that the float additions happen in the reference's order.
MG_DEVICE MG_NOINLINE void fold_token_stats_warp(const MgBatch& b, int env, int lane) {
  uint16_t* tc = b.tok_counts + (size_t)env * b.A * 2;
  float* gs = b.game_stats + (size_t)env * b.GS;
  // All counts are small integers. While a total stays below 2^24 every float addition is exact, so
  // the sum of the counts can be added in one step; only beyond that does the reference's
  // agent-by-agent rounding matter, and then the additions are replayed in its order.
  int w = 0, d = 0, f = 0, any = 0;
  for (int ai = lane; ai < b.A; ai += 32) {
    const uint32_t pair = *reinterpret_cast<const uint32_t*>(tc + ai * 2);
    const int written = (int)(pair & 0xFFFFu), attempted = (int)(pair >> 16);
    if (attempted != 0xFFFF) { any = 1; w += written; d += attempted - written; f += b.T - written; }  // 0xFFFF: nothing pending
  }
  for (int off = 16; off > 0; off >>= 1) {
    w += __shfl_xor_sync(MG_FULL_MASK, w, off);
    d += __shfl_xor_sync(MG_FULL_MASK, d, off);
    f += __shfl_xor_sync(MG_FULL_MASK, f, off);
    any |= __shfl_xor_sync(MG_FULL_MASK, any, off);
  }
  if (!any) return;
  if (lane < 3) {
    const int total = lane == 0 ? w : (lane == 1 ? d : f);
    float acc = gs[MG_GS_TOKENS_WRITTEN + lane];
    if (acc + (float)total < 16777216.0f) {
      acc = MG_FADD(acc, (float)total);
    } else {
      for (int ai = 0; ai < b.A; ai++) {
        const int written = tc[ai * 2], attempted = tc[ai * 2 + 1];
        if (attempted == 0xFFFF) continue;
        const int v = lane == 0 ? written : (lane == 1 ? attempted - written : b.T - written);
        acc = MG_FADD(acc, (float)v);
      }
    }
    gs[MG_GS_TOKENS_WRITTEN + lane] = acc;
  }
  __syncwarp();
  for (int i = lane; i < b.A; i += 32) tc[i * 2 + 1] = 0xFFFF;
  __syncwarp();
}

// Cache, on every assembler that has an agent next to it, the protocol observers will see this
// step (see sink_object). One agent per lane; duplicates write identical values.
MG_DEVICE inline void update_protocol_cache_warp(const MgBatch& b, EnvCtx& e, int lane) {
  if (!b.cfg[MGC_HAS_VIBE_ASM] || !b.cfg[MGC_PROTOCOL_DETAILS_OBS]) return;
  // one (agent, neighbouring cell) pair per lane: with a handful of agents per env a lane per agent would
  // leave most of the warp idle while each agent walks its eight cells and every assembler it finds walks its
  // own eight. An assembler next to two agents is computed twice, with the same result.
  for (int pair = lane; pair < e.A * 8; pair += 32) {
    const int ai = pair >> 3, k = pair & 7;
    const int rr = (int)e.ag[ai].r + (int)((0xA940u >> (2 * k)) & 3u) - 1, cc = (int)e.ag[ai].c + (int)((0x9224u >> (2 * k)) & 3u) - 1;
    if (rr < 0 || cc < 0 || rr >= e.H || cc >= e.W) continue;
    const uint16_t g = grid_at(e, rr, cc);
    if (MG_GRID_KIND(g) != MG_GK_ASM) continue;
    AsmRec* a = &e.asms[MG_GRID_INDEX(g)];
    const int pi = asm_current_protocol(e, a);
    a->cache_protocol = pi;
    a->cache_step = e.current_step;
  }
  __syncwarp();
}

// What observers will see of every assembler this step (see write_asm_tokens).
MG_DEVICE MG_NOINLINE void write_all_asm_tokens_warp(const MgBatch& b, int env, int lane, int n_asm, uint32_t current_step) {
  const AsmRec* asms = b.asms + (size_t)env * b.max_asm;
  for (int i = lane; i < n_asm; i += 32) write_asm_tokens(b, &asms[i], current_step, token_record(b, env, MG_GK_ASM, (uint32_t)i));
  __syncwarp();
}

// One observation pass that only counts tokens (nothing is written): the reference runs such
// passes on buffers nobody reads (see reset_env_warp), but their token counts stay in the stats.
// `map`: >= 0 when the pass looks at the untouched start of an episode on that pool map and nothing but the
// map decides what the agents see (no clipper): the counts are then kept per map and reused.
MG_DEVICE MG_NOINLINE void count_observation_pass_warp(const MgBatch& b, int env, int lane, uint32_t current_step, int folds, int map) {
  __syncwarp();
  uint16_t* tc = b.tok_counts + (size_t)env * b.A * 2;
  const bool cacheable = map >= 0 && b.pool_tok0 != nullptr;
  if (cacheable && *reinterpret_cast<volatile uint32_t*>(&b.pool_tok0_valid[map]) != 0u) {
    __threadfence();
    for (int ai = lane; ai < b.A; ai += 32) {
      const int attempted = MG_LD_L2(&b.pool_tok0[(size_t)map * b.A + ai]);
      tc[ai * 2] = (uint16_t)(attempted < b.T ? attempted : b.T);
      tc[ai * 2 + 1] = (uint16_t)attempted;
    }
  } else if (b.obs_cells <= 128) {  // four agents' loads in flight at a time
    for (int ai = 0; ai < b.A; ai += 4) {
      int totals[4];
      const int n = b.A - ai < 4 ? b.A - ai : 4;
      obs_count_group_warp<4, 4>(b, env, ai, n, lane, current_step, totals);
      if (lane < n) {
        const int attempted = lane == 0 ? totals[0] : (lane == 1 ? totals[1] : (lane == 2 ? totals[2] : totals[3]));
        tc[(ai + lane) * 2] = (uint16_t)(attempted < b.T ? attempted : b.T);
        tc[(ai + lane) * 2 + 1] = (uint16_t)attempted;
      }
    }
  } else {
    for (int ai = 0; ai < b.A; ai++) {
      const int attempted = obs_count_agent_warp(b, env, ai, lane, current_step);
      if (lane == 0) {
        tc[ai * 2] = (uint16_t)(attempted < b.T ? attempted : b.T);
        tc[ai * 2 + 1] = (uint16_t)attempted;
      }
    }
  }
  __syncwarp();
  if (cacheable && *reinterpret_cast<volatile uint32_t*>(&b.pool_tok0_valid[map]) == 0u) {
    // several envs may get here for the same map at once: they all write the same values
    for (int ai = lane; ai < b.A; ai += 32) b.pool_tok0[(size_t)map * b.A + ai] = tc[ai * 2 + 1];
    __threadfence();
    __syncwarp();
    if (lane == 0) *reinterpret_cast<volatile uint32_t*>(&b.pool_tok0_valid[map]) = 1u;
  }
  // `folds` identical passes (the reference runs the pass again on an unchanged state): counted that many times
  for (int k = 0; k < folds; k++) {
    uint32_t keep[(MG_MAX_AGENTS + 31) / 32];
    for (int i = lane, q = 0; i < b.A; i += 32, q++) keep[q] = *reinterpret_cast<const uint32_t*>(tc + i * 2);
    fold_token_stats_warp(b, env, lane);
    if (k + 1 < folds) {
      for (int i = lane, q = 0; i < b.A; i += 32, q++) *reinterpret_cast<uint32_t*>(tc + i * 2) = keep[q];
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------- reset
// MettaGrid::MettaGrid + _init_grid + Clipper ctor (mettagrid_c.cpp:46-245, clipper.hpp:35-91),
// from a map that the host already split into per-kind object tables (metta_b200/maps.py).
// `seeded_cached`: the env's seed is the one its cached generator state was made from (auto-reset).
// `folds`: how many times the closing observation count is added to the token stats (see the end).
MG_DEVICE MG_NOINLINE void reset_env_warp(const MgBatch& b, int env, int lane, EnvRec& er, bool seeded_cached, int folds) {
#ifdef MG_TIMING
  const long long rt_begin = clock64();
#endif
  const int32_t* cfg = b.cfg;
  const int map = er.map_idx;
  // grid: 16-byte copies (grid_stride is a multiple of 8 cells)
  copy16_warp(reinterpret_cast<uint4*>(b.grid + (size_t)env * b.grid_stride),
              reinterpret_cast<const uint4*>(b.pool_grid + (size_t)map * b.grid_stride), b.grid_stride / 8, lane);
  er.current_step = 0;
  er.done = 0;
  er.n_asm = b.pool_n_asm[map];
  er.n_chest = b.pool_n_chest[map];
  er.rng_idx = MT_N;
  er.clip_rng_idx = MT_N;
  float* astats = b.agent_stats + (size_t)env * b.A * b.S;
  {
    const uint4 zero = {0u, 0u, 0u, 0u};
    const size_t n = (size_t)b.A * b.S;  // the env's slice starts 16-byte aligned when A * S is a multiple of 4
    if ((n & 3) == 0 && (reinterpret_cast<uintptr_t>(astats) & 15) == 0) {
      for (int i = lane; i < (int)(n >> 2); i += 32) reinterpret_cast<uint4*>(astats)[i] = zero;
    } else {
      for (int i = lane; i < (int)n; i += 32) astats[i] = 0.0f;
    }
    uint4* hot = reinterpret_cast<uint4*>(b.hot_stats + (size_t)env * b.A * MG_HOT_STATS);
    for (int i = lane; i < b.A * MG_HOT_STATS / 4; i += 32) hot[i] = zero;
  }
  float* gstats = b.game_stats + (size_t)env * b.GS;
  for (int i = lane; i < b.GS; i += 32)  // objects.<cell name> are counted at construction: constants of the map
    gstats[i] = (i >= MG_GS_OBJECTS && i < MG_GS_OBJECTS + MG_MAX_CELLTYPES && b.pool_obj_counts != nullptr)
                    ? b.pool_obj_counts[(size_t)map * MG_MAX_CELLTYPES + (i - MG_GS_OBJECTS)] : 0.0f;
  for (int i = lane; i < b.A * MG_INVALID_WORDS; i += 32) b.invalid_ids[(size_t)env * b.A * MG_INVALID_WORDS + i] = 0;
  uint16_t* tc = b.tok_counts + (size_t)env * b.A * 2;
  for (int i = lane; i < b.A; i += 32) { tc[i * 2] = 0; tc[i * 2 + 1] = 0xFFFF; }
  AgentRec* agents = b.agents + (size_t)env * b.A;
  AsmRec* asms = b.asms + (size_t)env * b.max_asm;
  ChestRec* chests = b.chests + (size_t)env * b.max_chest;
  const int32_t* pa = b.pool_agents + (size_t)map * b.A * 3;
  for (int ai = lane; ai < b.A; ai += 32) {
    AgentRec a;
    const int32_t* cc = cell_cfg(cfg, pa[ai * 3 + 2]);
    const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + cc[MG_CELL_PARAM] * MG_GRP_WORDS;
    a.r = a.prev_r = (uint16_t)pa[ai * 3 + 0];
    a.c = a.prev_c = (uint16_t)pa[ai * 3 + 1];
    a.vibe = (uint8_t)grp[MG_GRP_INITIAL_VIBE];
    a.group = (uint8_t)cc[MG_CELL_PARAM];
    a.frozen = 0;
    a.steps_without_motion = 0;
    a.order = MG_ORDER_EMPTY;
    for (int i = 0; i < MG_MAX_RESOURCES; i++) a.amt[i] = 0;
    a.pad0 = 0;
    a.current_stat_reward = 0.0f;
    a.episode_reward = 0.0f;
    a.pad1 = 0;
    agents[ai] = a;
    b.last_action[(size_t)env * b.A + ai] = 0;
    b.action_success[(size_t)env * b.A + ai] = 0;
  }
  const int32_t* ps = b.pool_asm + (size_t)map * b.max_asm * 3;
  for (int i = lane; i < er.n_asm; i += 32) {
    AsmRec a;
    a.r = (uint16_t)ps[i * 3 + 0];
    a.c = (uint16_t)ps[i * 3 + 1];
    a.cell = (uint8_t)ps[i * 3 + 2];
    const int32_t* ty = asm_type_cfg(cfg, a.cell);
    a.flags = (uint8_t)(ty[MG_ASM_START_CLIPPED] ? MG_ASMF_CLIPPED : 0);
    if (b.has_clipper && !ty[MG_ASM_CLIP_IMMUNE]) a.flags |= MG_ASMF_TRACKED | MG_ASMF_UNCLIPPED;
    a.unclip_protocol = -1;
    a.cooldown_end = a.cooldown_duration = a.uses_count = 0;
    a.infection_weight = 0;
    a.cache_protocol = -1;
    a.cache_step = 0xFFFFFFFFu;
    asms[i] = a;
  }
  const int32_t* pc = b.pool_chest + (size_t)map * b.max_chest * 3;
  for (int i = lane; i < er.n_chest; i += 32) {
    ChestRec c;
    c.r = (uint16_t)pc[i * 3 + 0];
    c.c = (uint16_t)pc[i * 3 + 1];
    c.cell = (uint8_t)pc[i * 3 + 2];
    c.pad[0] = c.pad[1] = c.pad[2] = 0;
    c.order = MG_ORDER_EMPTY;
    for (int k = 0; k < MG_MAX_RESOURCES; k++) c.amt[k] = 0;
    c.pad1[0] = c.pad1[1] = c.pad1[2] = 0;
    chests[i] = c;
  }
  uint32_t* mt = b.rng + (size_t)env * MT_N;
  {  // std::mt19937(seed), mettagrid_c.cpp:57: 623 dependent steps on one lane, or a copy of the state made last time
    uint4* cache = reinterpret_cast<uint4*>(b.rng_seeded + (size_t)env * MT_N);
    uint4* st = reinterpret_cast<uint4*>(mt);
    if (seeded_cached) {
      copy16_warp(st, cache, MT_N / 4, lane);
    } else {
      if (lane == 0) mt_seed_serial(mt, er.seed);
      __syncwarp();
      copy16_warp(cache, st, MT_N / 4, lane);
    }
  }
  __syncwarp();
  EnvCtx e = make_ctx(b, env, agents, er);
  // initial inventories bypass the limits but still report through on_inventory_change
  // (agent.cpp:48-52, chest.hpp:78-82)
  for (int ai = lane; ai < b.A; ai += 32) {
    const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + agents[ai].group * MG_GRP_WORDS;
    const int32_t* l = grp + MG_GRP_INITIAL_INV;
    InvView iv = agent_inv(e, ai);
    for (int k = 0; k < l[MG_LIST_N]; k++) inv_update(e, iv, l[MG_LIST_ITEM(k)], l[MG_LIST_VAL(k)] & 0xFFFF, true, 0);
  }
  if (lane == 0) {  // chests share the game stat slots: one lane, creation order
    for (int i = 0; i < er.n_chest; i++) {
      const int32_t* ct = cfg + cfg[MGC_OFF_CHEST_TYPES] + cell_cfg(cfg, chests[i].cell)[MG_CELL_PARAM] * MG_CHEST_WORDS;
      const int32_t* l = ct + MG_CHEST_INITIAL_INV;
      InvView iv = chest_inv(e, i);
      for (int k = 0; k < l[MG_LIST_N]; k++)
        if (l[MG_LIST_VAL(k)] > 0) inv_update(e, iv, l[MG_LIST_ITEM(k)], l[MG_LIST_VAL(k)], true, 0);
    }
  }
  __syncwarp();
  {  // token records: static objects from the per-config templates, then agents, chests, global tokens
    const uint4* src = reinterpret_cast<const uint4*>(b.static_rec);
    uint4* dst = reinterpret_cast<uint4*>(token_record(b, env, MG_GK_STATIC, 0u));
    copy16_warp(dst, src, (cfg[MGC_NUM_CELLTYPES] << b.rec_shift) >> 4, lane);
  }
  for (int ai = lane; ai < b.A; ai += 32) {
    write_agent_tokens(b, agents[ai], token_record(b, env, MG_GK_AGENT, (uint32_t)ai));
    write_global_tokens(b, agents[ai], 0u, 0, b.gtok + ((size_t)env * b.A + ai) * MG_GTOK_STRIDE);
  }
  for (int i = lane; i < er.n_chest; i += 32) write_chest_tokens(b, chests[i], token_record(b, env, MG_GK_CHEST, (uint32_t)i));
  __syncwarp();
  // The reference's constructor ends with an observation pass on its internal buffers
  // (_make_buffers -> set_buffers -> _init_buffers, mettagrid_c.cpp:139-140,247-262), BEFORE the
  // clipper exists; the caller's set_buffers (simulator.py:93-100) then runs the pass whose output
  // is kept. Both passes add to tokens_written/dropped/free_space, so the first is counted here.
#ifdef MG_TIMING
  const long long rt1 = clock64();
#endif
  update_protocol_cache_warp(b, e, lane);
  write_all_asm_tokens_warp(b, env, lane, er.n_asm, 0u);
#ifdef MG_TIMING
  const long long rt2 = clock64();
#endif
  count_observation_pass_warp(b, env, lane, 0u, folds, b.has_clipper ? -1 : er.map_idx);
#if defined(MG_TIMING) && !defined(MG_TIMING_ACC)
  if (lane == 0 && b.dbg != nullptr) {  // reset: [13] up to the assembler records | [14] assembler records | [15] count pass
    b.dbg[(size_t)env * 16 + 13] = (uint32_t)(rt1 - rt_begin);
    b.dbg[(size_t)env * 16 + 14] = (uint32_t)(rt2 - rt1);
    b.dbg[(size_t)env * 16 + 15] = (uint32_t)(clock64() - rt2);
  }
#endif
  if (!b.has_clipper) return;  // nothing below changes what observers see
  {
    uint32_t* cmt = b.clip_rng + (size_t)env * MT_N;
    for (int i = lane; i < MT_N; i += 32) cmt[i] = mt[i];  // the clipper keeps its own copy of the rng
    const int32_t* cc = cfg + cfg[MGC_OFF_CLIPPER];
    int scale = cc[MG_CLIP_LENGTH_SCALE];
    int ntracked = 0;
    for (int i = 0; i < er.n_asm; i++) ntracked += (asms[i].flags & MG_ASMF_TRACKED) ? 1 : 0;
    if (scale <= 0 && ntracked > 0) {  // clipper.hpp:62-78 integer Newton square root
      const uint32_t sparsity = (uint32_t)(b.W * b.H) / (uint32_t)ntracked;
      int32_t root = 1, next = 10;
      for (int i = 0; i < 10; i++) {
        if (next == root || root == 0) break;
        root = next;
        next = (int32_t)((sparsity + (uint32_t)(root * root) + (uint32_t)(2 * root - 2)) / (uint32_t)(2 * root));
      }
      scale = root / 2 > 1 ? root / 2 : 1;
    }
    if (scale <= 0) scale = 1;
    er.clip_length_scale = scale;
    e.clip_length_scale = scale;
    __syncwarp();
    for (int i = 0; i < er.n_asm; i++) {  // starting_clipped_assemblers, object id order
      const bool start = (asms[i].flags & MG_ASMF_TRACKED) && asm_type_cfg(cfg, asms[i].cell)[MG_ASM_START_CLIPPED];
      if (!start) continue;
      MtStream cs = mt_stream(cmt, er.clip_rng_idx);
      const uint32_t sel = uniform_below_warp(cs, lane, (uint32_t)cc[MG_CLIP_NUM_PROTOCOLS]);
      er.clip_rng_idx = cs.idx;
      if (lane == 0) {
        asms[i].flags &= (uint8_t)~MG_ASMF_CLIPPED;
        clip_assembler(e, &asms[i], sel);
      }
      __syncwarp();
    }
  }
  update_protocol_cache_warp(b, e, lane);
  write_all_asm_tokens_warp(b, env, lane, er.n_asm, 0u);
}

// v as fixed point with 16 fraction bits, from 32-bit conversions only (|v| < 2^31)
MG_DEVICE inline long long to_fixed16(float v) {
  const float whole = floorf(v);
  return (long long)(int)whole * 65536ll + (long long)(int)((v - whole) * 65536.0f + 0.5f);
}
MG_DEVICE inline void accum_add(long long* slot, float v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(slot), (unsigned long long)to_fixed16(v));
}

// An episode that has just ended is summed into the batch accumulators by the observation pass that follows
// the step: the warp that encodes agent ai's row adds that agent's stat row and episode reward (and agent
// 0's warp the env's game stats). Twenty-four warps share the work there, off the step's critical chain; the
// state itself stays as it is until the next step starts the new episode.
MG_DEVICE MG_NOINLINE void accumulate_agent_warp(const MgBatch& b, int env, int ai, int lane) {
  const size_t g = (size_t)env * b.A + ai;
  const float* row = b.agent_stats + g * b.S;
  const float* hot = b.hot_stats + g * MG_HOT_STATS;
  if (lane == 0) {
    const float r = b.agents[g].episode_reward;
    if (r != 0.0f) accum_add(&b.episode_accum[1], r);
    if (ai == 0) accum_add(&b.episode_accum[0], 1.0f);
  }
  if (lane < MG_HOT_STATS) {
    const float v = hot[lane];
    if (v != 0.0f) accum_add(&b.episode_accum[2 + lane], v);
  }
  for (int s0 = MG_HOT_STATS; s0 < b.S; s0 += 32 * 4) {  // four loads in flight per lane
    float x[4];
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = s0 + lane + 32 * k < b.S ? row[s0 + lane + 32 * k] : 0.0f;
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (x[k] != 0.0f) accum_add(&b.episode_accum[2 + s0 + lane + 32 * k], x[k]);
  }
  if (ai == 0) {
    const float* gstats = b.game_stats + (size_t)env * b.GS;
    for (int s = lane; s < b.GS; s += 32) {
      // the token counts of this very pass are still pending (folded at the next step's start): add them here
      float v = gstats[s];
      if (v != 0.0f) accum_add(&b.episode_accum[2 + b.S + s], v);
    }
  }
}

MG_GLOBAL void mg_reset_kernel(const MG_GRID_CONSTANT MgBatch b, const uint8_t* mask, const uint32_t* seeds, const int32_t* map_idx) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * MG_STEP_WARPS + (threadIdx.x >> 5);
  if (env >= b.num_envs) return;
  if (mask != nullptr && !mask[env]) return;
  EnvRec er = b.env[env];
  er.seed = seeds[env];
  er.map_idx = map_idx[env];
  er.early_reset_step = 0u;
  reset_env_warp(b, env, lane, er, false, 1);
  for (int ai = lane; ai < b.A; ai += 32) {  // _init_buffers, mettagrid_c.cpp:264-289
    const size_t g = (size_t)env * b.A + ai;
    b.rewards[g] = 0.0f;
    b.terminals[g] = 0;
    b.truncations[g] = 0;
  }
  if (lane == 0) b.env[env] = er;
}

// Agent::check_and_apply_damage for every agent of the env (agent.cpp:134-177), in agent order,
// drawing from the env's generator.
MG_DEVICE MG_NOINLINE void damage_pass_warp(const MgBatch& b, EnvCtx& e, int32_t* scratch, MtStream& rs, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  for (int base = 0; base < A; base += 32) {
    const int ai = base + lane;
    bool met = false;
    if (ai < A) {
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + e.ag[ai].group * MG_GRP_WORDS;
      if (grp[MG_GRP_DAMAGE_ENABLED]) {
        met = true;
        const int32_t* th = grp + MG_GRP_DAMAGE_THRESHOLD;
        for (int i = 0; i < th[MG_LIST_N]; i++)
          if ((int)e.ag[ai].amt[th[MG_LIST_ITEM(i)]] < (th[MG_LIST_VAL(i)] & 0xFFFF)) { met = false; break; }
      }
    }
    uint32_t todo = __ballot_sync(MG_FULL_MASK, met);
    while (todo) {
      const int who = base + (__ffs((int)todo) - 1);
      todo &= todo - 1u;
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + e.ag[who].group * MG_GRP_WORDS;
      if (lane == 0) {
        InvView iv = agent_inv(e, who);
        const int32_t* th = grp + MG_GRP_DAMAGE_THRESHOLD;
        for (int i = 0; i < th[MG_LIST_N]; i++) inv_update(e, iv, th[MG_LIST_ITEM(i)], -th[MG_LIST_VAL(i)], false, 0);
        const int32_t* rs = grp + MG_GRP_DAMAGE_RESOURCES;
        int n = 0;
        for (int i = 0; i < rs[MG_LIST_N]; i++) {
          const int removable = (int)e.ag[who].amt[rs[MG_LIST_ITEM(i)]] - rs[MG_LIST_VAL(i)];
          if (removable > 0) { scratch[1 + n] = rs[MG_LIST_ITEM(i)]; scratch[1 + MG_RES_STRIDE + n] = removable; n++; }
        }
        scratch[0] = n;
      }
      __syncwarp();
      const int n = scratch[0];
      if (n > 0) {
        int weights[MG_RES_STRIDE];
        for (int i = 0; i < n; i++) weights[i] = scratch[1 + MG_RES_STRIDE + i];
        const int sel = discrete_pick_warp(rs, lane, weights, n);
        if (lane == 0) {
          InvView iv = agent_inv(e, who);
          const int item = scratch[1 + sel];
          inv_update(e, iv, item, -1, false, 0);
          *agent_stat(e, who, MG_AS_DAMAGE_ITEMS_LOST) += 1.0f;
          *agent_stat(e, who, MG_AS_DAMAGED + item) += 1.0f;
        }
      }
      if (lane == 0) *agent_stat(e, who, MG_AS_DAMAGE_TRIGGERED) += 1.0f;
      __syncwarp();
    }
  }

}

// Clipper::maybe_clip_new_assembler (clipper.hpp:203-238) with the clipper's private generator.
MG_DEVICE MG_NOINLINE void clipper_step_warp(const MgBatch& b, EnvCtx& e, int32_t* scratch, uint32_t& clip_rng_idx, int lane) {
  const int32_t* cfg = b.cfg;
  const int env = e.env;
  {
    const int32_t* cc = cfg + cfg[MGC_OFF_CLIPPER];
    uint32_t* cmt = b.clip_rng + (size_t)env * MT_N;
    const uint32_t period = (uint32_t)cc[MG_CLIP_PERIOD];
    MtStream cs = mt_stream(cmt, clip_rng_idx);
    bool fire;
    if (period == 0u) {
      // uniform_int_distribution<uint32_t>(1, 0): the 64-bit range wraps, libstdc++ composes two draws
      // (uniform_int_dist.h:331-356) and truncates: the result is 1 only when the second word is 0.
      (void)mt_next_warp(cs, lane);
      fire = mt_next_warp(cs, lane) == 0u;
    } else {
      fire = uniform_below_warp(cs, lane, period) == 0u;
    }
    if (fire) {
      if (lane == 0) {
        int nun = 0;
        uint32_t total = 0;
        for (int i = 0; i < e.n_asm; i++) {
          nun += (e.asms[i].flags & MG_ASMF_UNCLIPPED) ? 1 : 0;
          if (e.asms[i].flags & MG_ASMF_BORDER) total += e.asms[i].infection_weight;
        }
        scratch[0] = nun;
        scratch[1] = (int32_t)total;
      }
      __syncwarp();
      const int nun = scratch[0];
      const uint32_t total = (uint32_t)scratch[1];
      __syncwarp();
      if (nun > 0) {
        int pick = -1;
        if (total == 0u) {
          uint32_t k = uniform_below_warp(cs, lane, (uint32_t)nun);
          for (int i = 0; i < e.n_asm && pick < 0; i++)
            if (e.asms[i].flags & MG_ASMF_UNCLIPPED) { if (k == 0u) pick = i; else k--; }
        } else {
          uint32_t rw = uniform_below_warp(cs, lane, total) + 1u;
          for (int i = 0; i < e.n_asm && pick < 0; i++) {
            if (!(e.asms[i].flags & MG_ASMF_BORDER)) continue;
            if (rw <= e.asms[i].infection_weight) pick = i; else rw -= e.asms[i].infection_weight;
          }
        }
        if (pick >= 0) {
          const uint32_t sel = uniform_below_warp(cs, lane, (uint32_t)cc[MG_CLIP_NUM_PROTOCOLS]);
          __syncwarp();  // every lane has walked the records before lane 0 changes them
          if (lane == 0) clip_assembler(e, &e.asms[pick], sel);
        }
      }
      __syncwarp();
    }
    clip_rng_idx = cs.idx;
  }

}

// ---------------------------------------------------------------- step
// The step kernel is written as a short sequence of out-of-line phases that communicate through
// the warp's shared-memory block (StepSmem): each phase gets its own register allocation, so the
// kernel needs no spills (as one function it wanted 210 registers per thread; spill traffic to local
// memory was three times the real data traffic, see profiles/r01e). All phases are called by the
// full warp.

// Once per warp, before its first env.
template <int MAXA>
MG_DEVICE inline void step_smem_init(StepSmem<MAXA>& sm, int lane, bool publish) {
  if (lane == 0) {
#ifdef MG_HAS_BULK_COPY
    mg_mbar_init(reinterpret_cast<uint64_t*>(&sm.stage_bar), 1u);
#endif
    sm.stage_phase = 0u;
    sm.publish = publish ? 1u : 0u;
  }
  __syncwarp();
}

// Phase 1: bring the env in (and start a new episode first if the last one ended).
template <int MAXA>
MG_DEVICE MG_PHASE_LOAD_FN void step_phase_load(const MgBatch& b, StepSmem<MAXA>& sm, int env, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  EnvRec er;
  copy_env_rec(&er, &b.env[env]);
  // ---- bring the env in. Every load that does not depend on another is issued before the first
  // use, so the whole stage costs about one HBM round trip instead of one per array: agent records
  // and their 16 per-step stat slots (16-byte loads), actions, the pending token counts of the last
  // observation pass and the three float totals they are folded into.
  constexpr int NV = MAXA * 4 / 32;  // 16-byte loads per lane for records (and for hot stats)
  (void)NV;
  constexpr int NA = (MAXA + 31) / 32;
  const size_t abase = (size_t)env * A;
#ifdef MG_PREFETCH_ENV
  // The env's grid and assembler records are read later in dependent steps (target cells, then what stands
  // there, then its neighbourhood), each of which would be an HBM round trip after the policy's forward pass
  // has swept L2. Ask for all of it now (a few KB per env): the later reads, and the observation pass
  // behind this kernel, find it in L2.
  prefetch_range_warp(b.grid + (size_t)env * b.grid_stride, (size_t)b.H * b.W * sizeof(uint16_t), lane);
#if MG_PREFETCH_ENV > 1
  prefetch_range_warp(b.asms + (size_t)env * b.max_asm, (size_t)b.max_asm * sizeof(AsmRec), lane);
#endif
#if MG_PREFETCH_ENV > 2
  prefetch_range_warp(b.etok + (((size_t)env * b.rec_slots) << b.rec_shift), (size_t)b.rec_slots << b.rec_shift, lane);
#endif
#endif
  for (bool fresh = false;;) {
#ifndef MG_HAS_BULK_LOAD
    uint4 ag_v[NV], hot_v[NV];
#endif
    int32_t act_v[NA];
    uint32_t tc_v[NA];
    const uint4* ag_src = reinterpret_cast<const uint4*>(b.agents + abase);
    const uint4* hot_src = reinterpret_cast<const uint4*>(b.hot_stats + abase * MG_HOT_STATS);
    uint16_t* tc = b.tok_counts + abase * 2;
    float* gs = b.game_stats + (size_t)env * b.GS;
#ifdef MG_HAS_BULK_LOAD
    // agent records and hot stat slots: two bulk copies (TMA) straight into the warp's shared block, issued by
    // one lane and waited for on the block's mbarrier further down; no registers, no per-lane address math
    if (fresh) {
      // the records were just written by this warp's own (generic) stores: read them back the same way
      for (int i = lane; i < A * 4; i += 32) {
        reinterpret_cast<uint4*>(sm.ag)[i] = ag_src[i];
        reinterpret_cast<uint4*>(sm.hot)[i] = hot_src[i];
      }
    } else if (lane == 0) {
      uint64_t* bar = reinterpret_cast<uint64_t*>(&sm.stage_bar);
      const uint32_t bytes = (uint32_t)A * (uint32_t)sizeof(AgentRec);
      mg_mbar_expect_tx(bar, 2u * bytes);
      mg_bulk_load(sm.ag, ag_src, bytes, bar);
      mg_bulk_load(sm.hot, hot_src, bytes, bar);
    }
#else
#pragma unroll
    for (int k = 0; k < NV; k++) {
      const int i = lane + 32 * k;
      if (i < A * 4) {
        ag_v[k] = ag_src[i];
        hot_v[k] = hot_src[i];
      }
    }
#endif
#pragma unroll
    for (int k = 0; k < NA; k++) {
      const int ai = lane + 32 * k;
      act_v[k] = ai < A ? b.actions[abase + ai] : 0;
      tc_v[k] = ai < A ? *reinterpret_cast<const uint32_t*>(tc + ai * 2) : 0xFFFF0000u;
    }
    float gs3 = lane < 3 ? gs[MG_GS_TOKENS_WRITTEN + lane] : 0.0f;
    // the generator words the shuffle will draw from: one more load in this batch instead of an HBM
    // round trip of its own in the next phase
    const uint32_t rng_first = er.rng_idx;
    const uint32_t rng_n = rng_first < MT_N ? (MT_N - rng_first < 32u ? MT_N - rng_first : 32u) : 0u;
    // straight into the shared block, in the background: the decide phase waits for it where it draws (this load
    // depends on EnvRec, so it is a memory round trip of its own; nothing in this phase needs its result)
    if ((uint32_t)lane < rng_n) MG_ASYNC_COPY4(&sm.rngbuf[lane], &b.rng[(size_t)env * MT_N + rng_first + lane]);

    // fold the pending token counts (written | attempted << 16; attempted 0xFFFF = nothing pending)
    {
      int w = 0, d = 0, f = 0, any = 0;
#pragma unroll
      for (int k = 0; k < NA; k++) {
        const int written = (int)(tc_v[k] & 0xFFFFu), attempted = (int)(tc_v[k] >> 16);
        if (attempted != 0xFFFF) { any = 1; w += written; d += attempted - written; f += b.T - written; }
      }
      for (int off = 16; off > 0; off >>= 1) {
        w += __shfl_xor_sync(MG_FULL_MASK, w, off);
        d += __shfl_xor_sync(MG_FULL_MASK, d, off);
        f += __shfl_xor_sync(MG_FULL_MASK, f, off);
        any |= __shfl_xor_sync(MG_FULL_MASK, any, off);
      }
      if (any) {
        const int total = lane == 0 ? w : (lane == 1 ? d : f);
        const bool exact = gs3 + (float)total < 16777216.0f;  // see fold_token_stats_warp
        float folded = MG_FADD(gs3, (float)total);
        if (__ballot_sync(MG_FULL_MASK, lane < 3 && !exact) != 0u) {
          // a total beyond 2^24 (tokens_free_space gets there after ~4600 arena steps): the reference's float
          // additions round agent by agent, so they are replayed in its order. Every lane converts its own
          // agents' three addends once and parks them in shared memory (intent / executed / action are not in
          // use yet; an agent with nothing pending adds 0, which changes nothing); lanes 0..2 then run one
          // load-and-add chain each.
          float* parked[3] = {reinterpret_cast<float*>(sm.intent), reinterpret_cast<float*>(sm.executed), reinterpret_cast<float*>(sm.action)};
#pragma unroll
          for (int k = 0; k < NA; k++) {
            const int ai = lane + 32 * k;
            if (ai < A) {
              const int written = (int)(tc_v[k] & 0xFFFFu), attempted = (int)(tc_v[k] >> 16);
              const bool pending = attempted != 0xFFFF;
              parked[0][ai] = pending ? (float)written : 0.0f;
              parked[1][ai] = pending ? (float)(attempted - written) : 0.0f;
              parked[2][ai] = pending ? (float)(b.T - written) : 0.0f;
            }
          }
          __syncwarp();
          const float* mine = lane == 0 ? parked[0] : (lane == 1 ? parked[1] : parked[2]);
          folded = gs3;
#pragma unroll 4
          for (int ai = 0; ai < A; ai++) folded = MG_FADD(folded, mine[ai]);
          __syncwarp();
        }
        if (lane < 3) gs[MG_GS_TOKENS_WRITTEN + lane] = folded;
#pragma unroll
        for (int k = 0; k < NA; k++) {
          const int ai = lane + 32 * k;
          if (ai < A) tc[ai * 2 + 1] = 0xFFFF;
        }
      }
    }
    MG_SM_TICK(sm, 0);
    if (er.done && b.auto_reset && !fresh) {  // MettaGridPufferEnv.step: new Simulation, then apply the actions
#ifdef MG_HAS_BULK_LOAD
      mg_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_bar), sm.stage_phase);  // the old episode's records: let the copy land, then drop it
      __syncwarp();
      if (lane == 0) sm.stage_phase ^= 1u;
      __syncwarp();
#endif
      {  // everything the episode change will read, requested now
        const int next_map = b.num_maps > 0 ? (er.map_idx + b.map_stride) % b.num_maps : er.map_idx;
        prefetch_range_warp(b.pool_grid + (size_t)next_map * b.grid_stride, (size_t)b.grid_stride * sizeof(uint16_t), lane);
        prefetch_range_warp(b.rng_seeded + (size_t)env * MT_N, MT_N * sizeof(uint32_t), lane);
        prefetch_range_warp(b.pool_asm + (size_t)next_map * b.max_asm * 3, (size_t)b.max_asm * 3 * sizeof(int32_t), lane);
        prefetch_range_warp(b.pool_agents + (size_t)next_map * A * 3, (size_t)A * 3 * sizeof(int32_t), lane);
        if (lane == 0) { MG_PREFETCH_TO_L2(b.pool_n_asm + next_map); MG_PREFETCH_TO_L2(b.pool_n_chest + next_map); }
        if (b.pool_tok0 != nullptr) prefetch_range_warp(b.pool_tok0 + (size_t)next_map * A, (size_t)A * sizeof(uint16_t), lane);
      }
      __syncwarp();
      if (b.num_maps > 0) er.map_idx = (er.map_idx + b.map_stride) % b.num_maps;
      er.episodes += 1;
      // set_buffers of the new Simulation: the initial observation nobody sees is still counted. Without a
      // clipper it equals the constructor's own pass, which is then simply counted twice.
      reset_env_warp(b, env, lane, er, true, b.has_clipper ? 1 : 2);
      __syncwarp();
      if (b.has_clipper) count_observation_pass_warp(b, env, lane, 0u, 1, -1);
      fresh = true;
      continue;  // load the new episode's state
    }
#ifdef MG_HAS_BULK_LOAD
    if (!fresh) {
      mg_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_bar), sm.stage_phase);
      __syncwarp();
      if (lane == 0) sm.stage_phase ^= 1u;
    }
#else
#pragma unroll
    for (int k = 0; k < NV; k++) {
      const int i = lane + 32 * k;
      if (i < A * 4) {
        reinterpret_cast<uint4*>(sm.ag)[i] = ag_v[k];
        reinterpret_cast<uint4*>(sm.hot)[i] = hot_v[k];
      }
    }
#endif
#pragma unroll
    for (int k = 0; k < NA; k++) {
      const int ai = lane + 32 * k;
      if (ai < A) {
        sm.action[ai] = act_v[k];
        sm.executed[ai] = 0;
        sm.success[ai] = 0;
      }
    }
    if (lane == 0) { sm.rng_base = rng_first; sm.rng_n = rng_n; }
    break;
  }
  __syncwarp();
  MG_SM_TICK(sm, 1);
  // the stat slots the rewards will read at the end of the step: start pulling them in now (configs whose
  // rewards only look at hot counters and <res>.amount, like the arena, have none)
  if (b.reward_stat_prefetch) {
    for (int ai = lane; ai < A; ai += 32) {
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + sm.ag[ai].group * MG_GRP_WORDS;
      const float* st = b.agent_stats + (abase + ai) * b.S;
      for (int i = 0; i < grp[MG_GRP_NUM_REWARDS]; i++) {
        const int slot = grp[MG_GRP_REWARDS + i * MG_RWD_WORDS + MG_RWD_AGENT_STAT];
        if (slot >= MG_HOT_STATS && !(slot >= MG_AS_AMOUNT && slot < MG_AS_AMOUNT + MG_MAX_RESOURCES)) MG_PREFETCH_TO_L2(st + slot);
      }
    }
  }
  if (b.reward_game_stats && lane < (b.GS * 4 + 127) / 128) MG_PREFETCH_L2(b.game_stats + (size_t)env * b.GS + lane * 32);
  er.current_step += 1;
  if (lane == 0) {
    copy_env_rec(&sm.er, &er);
    EnvCtx& e = sm.ctx;
    init_ctx(e, b, env, sm.ag, er);
    e.hot = sm.hot;
    e.asm_records_on_use = cfg[MGC_HAS_VIBE_ASM] == 0;
    for (int k = 0; k < 3; k++) { e.req_mask[k] = b.req_mask[k]; e.con_mask[k] = b.con_mask[k]; }
  }
  __syncwarp();
}

// Phase 2: processing order, every agent's intent, and who may act independently of the order.
// Leaves order[], intent[] and ordered_mask in shared memory.
template <int MAXA>
MG_DEVICE MG_PHASE_DECIDE_FN void step_phase_decide(const MgBatch& b, StepSmem<MAXA>& sm, int env, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  EnvCtx& e = sm.ctx;
  uint32_t* mt = b.rng + (size_t)env * MT_N;
  const int nact = cfg[MGC_NUM_ACTIONS];
  const int32_t* acts = cfg + cfg[MGC_OFF_ACTIONS];
  // ---- what does every agent want to do? One agent per lane: decode the action, find the target
  // cell of a move and what stands there. These loads (and the prefetch of the assembler or chest
  // record behind the cell) are issued for all agents at once; the passes below hit L1.
  int my_kind = 0, my_arg = 0, my_tr = -1, my_tc = -1;
  bool my_valid = false;
  uint32_t my_target = 0;  // grid value of the cell a move goes to (0: empty, or no in-bounds move)
  if (A <= 32 && lane < A) {
    const int a = sm.action[lane];
    my_valid = a >= 0 && a < nact;
    if (my_valid) {
      my_kind = acts[a * MG_ACT_WORDS + MG_ACT_KIND];
      my_arg = acts[a * MG_ACT_WORDS + MG_ACT_ARG];
      if (my_kind == MG_ACTION_MOVE && sm.ag[lane].frozen == 0) {
        const int tr = (int)sm.ag[lane].r + (int)((0xA058u >> (2 * my_arg)) & 3u) - 1;
        const int tc2 = (int)sm.ag[lane].c + (int)((0x8885u >> (2 * my_arg)) & 3u) - 1;
        if (tr >= 0 && tc2 >= 0 && tr < b.H && tc2 < b.W) {
          my_tr = tr; my_tc = tc2;
          my_target = e.grid[tr * b.W + tc2];
          if (MG_GRID_KIND(my_target) == MG_GK_ASM) MG_PREFETCH_L1(&e.asms[MG_GRID_INDEX(my_target)]);
          else if (MG_GRID_KIND(my_target) == MG_GK_CHEST) MG_PREFETCH_L1(&e.chests[MG_GRID_INDEX(my_target)]);
        }
      }
    }
  }
  MG_SM_TICK(sm, 3);
  // random processing order (mettagrid_c.cpp:591-593); the words it needs are fetched in one go
  MtStream rs = mt_stream(mt, sm.er.rng_idx);
  MG_ASYNC_COPY_WAIT();  // the generator words the load phase asked for
  __syncwarp();
  if (sm.rng_n != 0u && sm.rng_base == rs.idx) {  // staged by the load phase
    rs.buf = sm.rngbuf; rs.buf_base = sm.rng_base; rs.buf_n = sm.rng_n;
  } else {
    mt_prefetch_warp(rs, sm.rngbuf, lane);
  }
  // the draws are always taken; the swaps are only applied if somebody needs the order (below)
  uint32_t* draws = reinterpret_cast<uint32_t*>(sm.scratch);
  const bool fast_shuffle = std_shuffle_draws_warp(rs, lane, A, draws);
  if (!fast_shuffle) std_shuffle_warp(rs, lane, sm.order, A);

  if (lane == 0) sm.er.rng_idx = rs.idx;
  // out-of-range actions: counted once per priority pass, two passes (mettagrid_c.cpp:599-608)
  for (int ai = lane; ai < A; ai += 32) {
    const int a = sm.action[ai];
    if (a < 0 || a >= nact) {
      float* st = agent_stat(e, ai, MG_AS_INVALID_INDEX);
      *st = MG_FADD(MG_FADD(*st, 1.0f), 1.0f);
      // action.invalid_index.<id>: the first few distinct ids with their counts, then a count of the rest
      int32_t* t = b.invalid_ids + ((size_t)env * A + ai) * MG_INVALID_WORDS;
      const int n = t[2 * MG_INVALID_IDS + 1];
      int hit = -1;
      for (int q = 0; q < n; q++)
        if (t[2 * q] == a) hit = q;
      if (hit < 0 && n < MG_INVALID_IDS) { hit = n; t[2 * n] = a; t[2 * n + 1] = 0; t[2 * MG_INVALID_IDS + 1] = n + 1; }
      if (hit >= 0) t[2 * hit + 1] += 2; else t[2 * MG_INVALID_IDS] += 2;
    }
  }
  __syncwarp();

  MG_SM_TICK(sm, 4);
  // ---- which agents can act independently of the processing order?
  // An action touches the mover's cell and the cell it moves into; using an assembler touches the
  // agents standing in the 3x3 block around it (they pay the inputs). Agents stand on distinct
  // cells, so the reference's shuffled order can only matter for an agent when
  //   (1) another agent moves into the same cell,
  //   (2) the cell it moves into holds an agent (swap, attack, transfer, or that agent leaves first),
  //   (3) it uses an assembler and another agent stands in, or moves into, that assembler's block,
  //   (4) it touches a chest or a chest-searching assembler (chest.<res>.amount is last-writer-wins)
  //       or anything under a clipper (shared infection weights).
  // Everybody else is handled one agent per lane, all at once; the rest goes through the ordered
  // pass below in the shuffled order. Both sides of a conflict are ordered.
  uint32_t ordered_mask;
  if (A <= 32) {
    if (lane == 0) sm.ordered_mask = 0u;  // bits of agents that someone else's action can affect
    __syncwarp();
    bool ordered = false;
    const uint32_t tkind = MG_GRID_KIND(my_target);
    const bool moving = my_tr >= 0;
    const uint32_t same_target = __match_any_sync(MG_FULL_MASK, moving ? my_tr * b.W + my_tc : -1 - lane);
    if (moving && (same_target & (same_target - 1u)) != 0u) ordered = true;                 // (1)
    if (moving && my_target != 0u && tkind == MG_GK_AGENT) {                                // (2)
      ordered = true;
      atomicOr(&sm.ordered_mask, 1u << MG_GRID_INDEX(my_target));
    }
    bool uses_asm = moving && tkind == MG_GK_ASM;
    if (moving && tkind == MG_GK_CHEST) ordered = true;                                     // (4)
    if (uses_asm && (b.has_clipper || b.any_chest_search)) {  // (4); most configs have neither: no look at the record then
      const AsmRec* ar = &e.asms[MG_GRID_INDEX(my_target)];
      if (b.has_clipper || asm_type_cfg(cfg, ar->cell)[MG_ASM_CHEST_DISTANCE] > 0) ordered = true;
    }
    // (3) every agent compares its own cell and the cell it moves into with the block of each user: no
    // grid reads (agents stand where their records say)
    const int my_r = lane < A ? (int)sm.ag[lane].r : -4, my_c = lane < A ? (int)sm.ag[lane].c : -4;
    uint32_t users = __ballot_sync(MG_FULL_MASK, uses_asm);
    while (users != 0u) {
      const int u = __ffs((int)users) - 1;
      users &= users - 1u;
      const int br = __shfl_sync(MG_FULL_MASK, my_tr, u), bc = __shfl_sync(MG_FULL_MASK, my_tc, u);
      const bool stands = lane != u && my_r >= br - 1 && my_r <= br + 1 && my_c >= bc - 1 && my_c <= bc + 1;
      const bool enters = moving && lane != u && my_tr >= br - 1 && my_tr <= br + 1 && my_tc >= bc - 1 && my_tc <= bc + 1;
      const bool hit = stands || enters;
      if (hit) ordered = true;
      if (__ballot_sync(MG_FULL_MASK, hit) != 0u && lane == u) ordered = true;
    }
    __syncwarp();
    if ((sm.ordered_mask >> lane) & 1u) ordered = true;
    __syncwarp();
    ordered_mask = __ballot_sync(MG_FULL_MASK, ordered && lane < A);
    if (lane < A) sm.intent[lane] = (my_valid && !ordered) ? (1u | ((uint32_t)my_kind << 8) | ((uint32_t)(my_arg & 0xFFFF) << 16)) : 0u;
  } else {
    ordered_mask = 0xFFFFFFFFu;
  }
  if (lane == 0) {
    sm.ordered_mask = ordered_mask;
    sm.er.spare[0] += A <= 32 ? (uint32_t)__popc(ordered_mask) : (uint32_t)A;  // agents that went through the ordered pass, since mg_create (tests read it)
    if (fast_shuffle && ordered_mask != 0u) std_shuffle_apply(sm.order, A, draws);
  }
  __syncwarp();

}

// Phase 3: the actions. Order-independent agents one per lane, then the ordered pass on one lane.
template <int MAXA>
MG_DEVICE MG_PHASE_ACT_FN void step_phase_act(const MgBatch& b, StepSmem<MAXA>& sm, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  EnvCtx& e = sm.ctx;
  const int nact = cfg[MGC_NUM_ACTIONS];
  const int32_t* acts = cfg + cfg[MGC_OFF_ACTIONS];
  const uint32_t ordered_mask = sm.ordered_mask;
  if (A <= 32) {
    if (lane == 0) e.concurrent = true;
    __syncwarp();
    const uint32_t intent = lane < A ? sm.intent[lane] : 0u;
    const int kind = (int)((intent >> 8) & 0xFFu);
    int code = MG_ACT_FROZEN, my_asm = -1;
    if (intent & 1u) {
      code = action_begin(e, lane, kind, (int)(intent >> 16), true, &my_asm);
      if (code == MG_ACT_OK || code == MG_ACT_FAIL) {
        action_end(e, lane, kind, code == MG_ACT_OK);
        sm.success[lane] = code == MG_ACT_OK ? 1 : 0;
        if (code == MG_ACT_OK) sm.executed[lane] = sm.action[lane];
      }
    }
    __syncwarp();
    // assembler uses that came up in this pass: one after the other, each by the whole warp (they touch
    // disjoint neighbourhoods, so their order does not matter)
    uint32_t pending = __ballot_sync(MG_FULL_MASK, code == MG_ACT_USE_PENDING);
    while (pending != 0u) {
      const int who = __ffs((int)pending) - 1;
      pending &= pending - 1u;
      const int asm_idx = __shfl_sync(MG_FULL_MASK, my_asm, who);
      const bool ok = assembler_use_warp(e, asm_idx, who, true, lane);
      if (lane == who) {
        action_end(e, who, MG_ACTION_MOVE, ok);
        sm.success[who] = ok ? 1 : 0;
        if (ok) sm.executed[who] = sm.action[who];
      }
      __syncwarp();
    }
    if (lane == 0) e.concurrent = false;
    __syncwarp();
  }
  MG_SM_TICK(sm, 6);
  // the ordered action pass: later agents see what earlier agents did. The agents that need it are
  // first pulled out of the shuffled order (one ballot), so that the serial loop only visits them.
  int n_ordered = A;
  if (A <= 32 && ordered_mask != 0u) {
    const int ai = lane < A ? (int)sm.order[lane] : 0;
    const bool mine = lane < A && ((ordered_mask >> ai) & 1u) != 0u;
    const uint32_t picks = __ballot_sync(MG_FULL_MASK, mine);
    __syncwarp();
    if (mine) sm.order[__popc(picks & ((1u << lane) - 1u))] = (uint8_t)ai;  // slot <= lane: never ahead of an unread entry
    n_ordered = __popc(picks);
    __syncwarp();
  }
  if (ordered_mask != 0u) {
    // every lane walks the list; lane 0 acts, and an assembler use is done by the whole warp
    for (int k = 0; k < n_ordered; k++) {
      const int ai = sm.order[k];
      const int a = sm.action[ai];
      if (a < 0 || a >= nact) continue;
      const int kind = acts[a * MG_ACT_WORDS + MG_ACT_KIND];
      int code = MG_ACT_FROZEN, asm_idx = -1;
      if (lane == 0) code = action_begin(e, ai, kind, acts[a * MG_ACT_WORDS + MG_ACT_ARG], true, &asm_idx);
      code = __shfl_sync(MG_FULL_MASK, code, 0);
      bool ok = code == MG_ACT_OK;
      if (code == MG_ACT_USE_PENDING) {
        asm_idx = __shfl_sync(MG_FULL_MASK, asm_idx, 0);
        ok = assembler_use_warp(e, asm_idx, ai, false, lane);
      }
      if (lane == 0 && code != MG_ACT_FROZEN) {
        action_end(e, ai, kind, ok);
        sm.success[ai] = ok ? 1 : 0;
        if (ok) sm.executed[ai] = a;
      }
      __syncwarp();
    }
  }
  __syncwarp();
  MG_SM_TICK(sm, 7);
}

// Phase 4: what happens after the actions: regeneration, damage, clipper, assembler views.
template <int MAXA>
MG_DEVICE MG_PHASE_POST_FN void step_phase_post(const MgBatch& b, StepSmem<MAXA>& sm, int env, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  EnvCtx& e = sm.ctx;
  EnvRec& er = sm.er;
  uint32_t* mt = b.rng + (size_t)env * MT_N;
  MtStream rs = mt_stream(mt, er.rng_idx);
  // inventory regeneration (mettagrid_c.cpp:625-640): independent per agent
  const uint32_t regen = (uint32_t)cfg[MGC_REGEN_INTERVAL];
  if (regen > 0u && er.current_step % regen == 0u) {
    for (int ai = lane; ai < A; ai += 32) regen_agent(e, ai);
    __syncwarp();
  }

  // damage (agent.cpp:134-177): agent order, draws from the env rng
  if (b.any_damage) damage_pass_warp(b, e, sm.scratch, rs, lane);
  // clipper (clipper.hpp:203-238): draws from its private rng copy
  if (b.has_clipper) {
    uint32_t cidx = er.clip_rng_idx;
    clipper_step_warp(b, e, sm.scratch, cidx, lane);
    __syncwarp();  // every lane has read the old position
    if (lane == 0) er.clip_rng_idx = cidx;
  }

  update_protocol_cache_warp(b, e, lane);
  // where the visible protocol can change without the assembler being used (vibes of neighbours, the
  // clipper), all assembler records are refreshed; otherwise assembler_on_use already did it
  if (cfg[MGC_HAS_VIBE_ASM]) write_all_asm_tokens_warp(b, env, lane, er.n_asm, er.current_step);

  __syncwarp();  // every lane has read the old position
  if (lane == 0) er.rng_idx = rs.idx;
  __syncwarp();
}

// Phase 5: rewards, flags, what observers will see, write-back.
template <int MAXA>
MG_DEVICE MG_PHASE_FINISH_FN void step_phase_finish(const MgBatch& b, StepSmem<MAXA>& sm, int env, int lane) {
  const int32_t* cfg = b.cfg;
  const int A = b.A;
  EnvCtx& e = sm.ctx;
  EnvRec er;
  copy_env_rec(&er, &sm.er);
  // rewards, episode totals, truncation (mettagrid_c.cpp:655-677) and the write-back
  const uint32_t max_steps = (uint32_t)cfg[MGC_MAX_STEPS];
  bool term = false, trunc = false;
  if (max_steps > 0u && er.current_step >= max_steps) {
    if (cfg[MGC_EPISODE_TRUNCATES]) trunc = true; else term = true;
  }
  if (er.early_reset_step != 0u && er.current_step >= er.early_reset_step) {  // Simulation.end_episode
    trunc = true;
    er.early_reset_step = 0u;
  }
  // flags are sticky within an episode (the reference only ever sets them) and cleared by reset
  er.done = (uint8_t)(er.done | (term ? 1 : 0) | (trunc ? 2 : 0));
  for (int ai = lane; ai < A; ai += 32) {
    const size_t g = (size_t)env * A + ai;
    const float delta = stat_reward_delta(e, ai);
    sm.ag[ai].episode_reward = MG_FADD(sm.ag[ai].episode_reward, delta);
    b.rewards[g] = delta;
    b.terminals[g] = (uint8_t)(er.done & 1);
    b.truncations[g] = (uint8_t)((er.done >> 1) & 1);
    b.last_action[g] = sm.executed[ai];
    b.action_success[g] = sm.success[ai];
    if (sm.ag[ai].pad0 != 0) {  // vibe, frozen or inventory changed: refresh what observers see of this agent
      write_agent_tokens(b, sm.ag[ai], token_record(b, env, MG_GK_AGENT, (uint32_t)ai));
      sm.ag[ai].pad0 = 0;
    }
    write_global_tokens(b, sm.ag[ai], er.current_step, sm.executed[ai], b.gtok + g * MG_GTOK_STRIDE);
  }
  for (int i = lane; i < er.n_chest; i += 32) write_chest_tokens(b, e.chests[i], token_record(b, env, MG_GK_CHEST, (uint32_t)i));
  __syncwarp();
  MG_SM_TICK(sm, 9);
  {
    // records and hot stat slots back to HBM, 16 bytes per lane and round (both arrays are dense per env)
    static_assert(MG_HOT_STATS == 16, "four 16-byte chunks of hot stats per agent, like the record");
    constexpr int NV = MAXA * 4 / 32;
    uint4* dst = reinterpret_cast<uint4*>(b.agents + (size_t)env * A);
    const uint4* src = reinterpret_cast<const uint4*>(sm.ag);
    const uint4* hot = reinterpret_cast<const uint4*>(sm.hot);
    uint4* hot_dst = reinterpret_cast<uint4*>(b.hot_stats + (size_t)env * A * MG_HOT_STATS);
#ifdef MG_HAS_BULK_COPY
    // two bulk copies (TMA) from the shared block: every lane's writes to it are made visible to the async proxy
    // first; one lane issues and then only waits until the block has been read (until the writes have been
    // performed where the env is handed to another SM right away)
    (void)src; (void)hot; (void)NV;
    mg_bulk_store_fence();
    __syncwarp();
    if (lane == 0) {
      const uint32_t bytes = (uint32_t)A * (uint32_t)sizeof(AgentRec);
      mg_bulk_store(dst, sm.ag, bytes);
      mg_bulk_store(hot_dst, sm.hot, bytes);
      mg_bulk_store_commit();
      if (sm.publish) mg_bulk_store_wait_all(); else mg_bulk_store_wait_read();
    }
#else
#pragma unroll
    for (int k = 0; k < NV; k++) {
      const int i = lane + 32 * k;
      if (i < A * 4) {
        dst[i] = src[i];
        hot_dst[i] = hot[i];
      }
    }
#endif
  }
  if (lane == 0) copy_env_rec(&b.env[env], &er);
  __syncwarp();
}

// One env, one warp: the five phases.
template <int MAXA>
MG_DEVICE inline void step_env_warp(const MgBatch& b, StepSmem<MAXA>& sm, int env, int lane) {
#ifdef MG_TIMING
  if (lane == 0) {
    sm.tick_last = clock64();
    for (int i = 0; i < 16; i++) sm.ticks[i] = 0;
    unsigned long long wall;  // slot 10: when this env started (ns, low word), to tell waves from chains
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(wall));
    sm.ticks[10] = (uint32_t)wall;
  }
  if (lane >= 12 && lane < 16 && b.dbg != nullptr) b.dbg[(size_t)env * 16 + lane] = 0u;
  __syncwarp();
#endif
  step_phase_load<MAXA>(b, sm, env, lane);
  MG_SM_TICK(sm, 2);
  step_phase_decide<MAXA>(b, sm, env, lane);
  MG_SM_TICK(sm, 5);
  step_phase_act<MAXA>(b, sm, lane);
  step_phase_post<MAXA>(b, sm, env, lane);
  MG_SM_TICK(sm, 8);
  step_phase_finish<MAXA>(b, sm, env, lane);
  MG_SM_TICK(sm, 11);
#ifdef MG_TIMING
  if (lane < 12 && b.dbg != nullptr) b.dbg[(size_t)env * 16 + lane] = sm.ticks[lane];  // 12..15: assembler_on_use adds to them
#endif
}

// MAXA: agents per env the shared-memory staging is sized for (32 or MG_MAX_AGENTS)
template <int MAXA>
MG_GLOBAL void MG_LAUNCH_BOUNDS(MG_STEP_WARPS * 32, MG_STEP_MIN_CTAS) mg_step_kernel(const MG_GRID_CONSTANT MgBatch b_param, int cfg_bytes_in_smem) {
  // dynamic shared memory: [StepSmem per warp][config copy]
  MG_SHARED MgBatch batch_s;  // the launch parameters with cfg redirected to the shared-memory copy
  MG_DYN_SMEM(dyn_smem);
  StepSmem<MAXA>* smem = reinterpret_cast<StepSmem<MAXA>*>(dyn_smem);
  unsigned char* cfg_smem = dyn_smem + sizeof(StepSmem<MAXA>) * MG_STEP_WARPS;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  // Optionally (cfg_bytes_in_smem, off by default: measured slower on B200) each CTA first copies the
  // lowered config next to its envs, so that the rules' dependent lookups (section offset -> record ->
  // field) are shared-memory accesses; otherwise they go to HBM through L1 and the launch parameters are
  // used where they are (the constant bank).
#if MG_CFG_SMEM_LIMIT > 0
  const bool cfg_fits = cfg_bytes_in_smem > 0 && cfg_bytes_in_smem >= b_param.cfg_words * 4;
  if (cfg_fits) {
    const uint4* src = reinterpret_cast<const uint4*>(b_param.cfg);
    uint4* dst = reinterpret_cast<uint4*>(cfg_smem);
    for (int i = threadIdx.x; i < (b_param.cfg_words + 3) / 4; i += blockDim.x) dst[i] = src[i];
    if (threadIdx.x == 0) {
      batch_s = b_param;
      batch_s.cfg = reinterpret_cast<const int32_t*>(cfg_smem);
    }
    __syncthreads();
  }
  const MgBatch& b = cfg_fits ? batch_s : b_param;
#else
  (void)cfg_smem; (void)cfg_bytes_in_smem; (void)batch_s;
  const MgBatch& b = b_param;  // known to be the launch parameters: inlined code reads them from the constant bank
#endif
  const int env = b.env_begin + blockIdx.x * MG_STEP_WARPS + warp;
  if (env >= b.env_end) return;
  step_smem_init<MAXA>(smem[warp], lane, b.env_flags != 0);
  step_env_warp<MAXA>(b, smem[warp], env, lane);
#ifndef MG_SIMT_EMU
  if (b.env_flags) {  // hand-over mode: every lane's writes first, then the env's entry in the order of completion
    __threadfence();
    __syncwarp();
    if (lane == 0) {
      const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&b.qctl->epoch);
      const uint32_t pos = atomicAdd(&b.qctl->tail, 1u);
      const unsigned long long v = ((unsigned long long)epoch << 32) | (unsigned long long)(uint32_t)env;
      asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(b.queue + pos), "l"(v) : "memory");
    }
  }
#endif
}

#ifndef MG_SIMT_EMU
// ---- the same step with one env per CTA, handing every finished env to the observation kernel that
// runs beside it (mg_obs_queue_kernel) instead of making it wait for the slowest env of a slice.
// One-warp CTAs give their registers back as soon as their env is done, which is what lets the
// observation CTAs in.
#ifndef MG_QUEUE_STEP_MIN_CTAS
#define MG_QUEUE_STEP_MIN_CTAS 24
#endif
template <int MAXA>
MG_GLOBAL void MG_LAUNCH_BOUNDS(32, MG_QUEUE_STEP_MIN_CTAS) mg_step_queue_kernel(const MG_GRID_CONSTANT MgBatch b) {
  MG_DYN_SMEM(dyn_smem);
  StepSmem<MAXA>& sm = *reinterpret_cast<StepSmem<MAXA>*>(dyn_smem);
  const int lane = threadIdx.x & 31;
  const int env = b.env_begin + (int)blockIdx.x;
  if (env >= b.env_end) return;
  step_smem_init<MAXA>(sm, lane, true);
  step_env_warp<MAXA>(b, sm, env, lane);
  // publish: every lane's writes first, then the entry
  __threadfence();
  __syncwarp();
  if (lane == 0) {
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&b.qctl->epoch);
    const uint32_t pos = atomicAdd(&b.qctl->tail, 1u);
    *reinterpret_cast<volatile unsigned long long*>(b.queue + pos) = ((unsigned long long)epoch << 32) | (unsigned long long)(uint32_t)env;
  }
}

MG_GLOBAL void mg_queue_begin_kernel(MgQueueCtl* ctl) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    ctl->epoch += 1u;
    ctl->tail = 0u;
    ctl->head = 0u;
    ctl->env_next = 0u;
  }
}
#endif  // !MG_SIMT_EMU

// ---------------------------------------------------------------- observations
// What the observation pass leaves behind per row besides the row itself: token counts for the stats fold of
// the next step, the side outputs for the policy (row length, running maximum, padding mask), and the
// episode sums when this step ended the episode.
// AUX = false: the caller knows that none of the three optional outputs is bound (the tests are compiled out)
template <bool AUX = true, int TT = 0>
MG_DEVICE inline void obs_row_finish_warp(const MgBatch& b, int env, int ai, size_t g, int attempted, uint32_t done_word, int lane) {
  const int T = TT ? TT : b.T;
  const int written = attempted < T ? attempted : T;
  if (lane == 0) {
    uint16_t* tc = b.tok_counts + (size_t)g * 2;
    tc[0] = (uint16_t)written;
    tc[1] = (uint16_t)attempted;
    if (AUX && b.row_len != nullptr) b.row_len[g] = written;
    // the running maximum only moves a few times per pass: look before bumping it, so that nearly
    // all rows do without an atomic on the one shared word
    if (AUX && b.max_flip != nullptr && written > MG_LD_L2(b.max_flip)) atomicMax(b.max_flip, written);
  }
  {
    if ((done_word & 0xFFu) != 0u) {  // the episode ended in this step: add it to the batch accumulators
      accumulate_agent_warp(b, env, ai, lane);
      if (lane < 3)  // this pass's own token counts (folded into the game stats only when the next step starts)
        accum_add(&b.episode_accum[2 + b.S + MG_GS_TOKENS_WRITTEN + lane], (float)(lane == 0 ? written : (lane == 1 ? attempted - written : T - written)));
    }
  }
  if (AUX && b.obs_mask != nullptr) {  // 8 slots per lane and round; a row of T bytes starts 8-byte aligned when T % 8 == 0
    uint8_t* m = b.obs_mask + (size_t)g * b.T;
    if ((b.T & 7) == 0) {
      for (int j0 = lane * 8; j0 < b.T; j0 += 256) {
        const int k = written - j0;  // slots of this group that hold tokens
        const unsigned long long v = k <= 0 ? 0x0101010101010101ull : (k >= 8 ? 0ull : (0x0101010101010101ull << (8 * k)));
        *reinterpret_cast<unsigned long long*>(m + j0) = v;
      }
    } else {
      for (int j = lane; j < b.T; j += 32) m[j] = j >= written ? 1 : 0;
    }
  }
}

// One agent's row by one warp. `warp_smem`: this warp's (slot + MG_OBS_DESC_BYTES(CPL)) bytes of shared
// memory, 16-byte aligned: [row image | cell descriptors]. TT > 0: the caller knows the token count per row
// (TT, a multiple of 8) and that the observation buffer is 8-byte aligned, so sizes, loop bounds and the
// alignment test are compile-time constants.
template <int CPL, bool AUX = true, int TT = 0>
MG_DEVICE inline void obs_row_warp(const MgBatch& b, int env, int ai, int lane, uint8_t* warp_smem) {
  const size_t g = (size_t)env * b.A + ai;  // global agent index
  const int row_bytes = (TT ? TT : b.T) * 3;
  const int slot = TT ? ((TT * 3 + 16 + 15) & ~15) : b.obs_slot;
  uint8_t* out = b.obs + (size_t)g * row_bytes;
  uint8_t* row = warp_smem;
  uint2* desc = reinterpret_cast<uint2*>(warp_smem + slot);
  // read past L1 (two env records share a cache line, and the hand-over modes encode envs other SMs stepped):
  // the step counter, and whether this step ended the episode (EnvRec::done, byte 40)
  const uint32_t current_step = MG_LD_L2(&b.env[env].current_step);
  const uint32_t done_word = (b.after_step && b.auto_reset) ? MG_LD_L2(reinterpret_cast<const uint32_t*>(&b.env[env]) + 10) : 0u;
  // (the image is filled with 0xFF inside, behind the first loads)
  const int attempted = obs_emit_agent_warp<CPL, TT>(b, env, ai, lane, row, desc, current_step, slot >> 4);
  __syncwarp();
  // store: rows are multiples of 8 bytes for every even token count, and then 8-byte aligned whenever the
  // buffer is: 8-byte chunks, no head or tail
  if (TT || ((row_bytes & 7) == 0 && (reinterpret_cast<uintptr_t>(b.obs) & 7) == 0)) {
    const uint2* s2 = reinterpret_cast<const uint2*>(row);
    uint2* d2 = reinterpret_cast<uint2*>(out);
    const int n8 = row_bytes >> 3;  // 75 for 200 tokens: three rounds, spelled out (a compiler-unrolled loop cost 79 instructions per row)
    if (lane < n8) d2[lane] = s2[lane];
    if (lane + 32 < n8) d2[lane + 32] = s2[lane + 32];
    if (lane + 64 < n8) d2[lane + 64] = s2[lane + 64];
#pragma unroll 1
    for (int i = lane + 96; i < n8; i += 32) d2[i] = s2[i];
  } else {
    for (int i = lane; i < row_bytes; i += 32) out[i] = row[i];
  }
  obs_row_finish_warp<AUX, TT>(b, env, ai, g, attempted, done_word, lane);
}

// MASKED: only the envs with mask[env] != 0 (mg_reset / mg_compute_observations of some envs); the pass behind a
// step is the unmasked instantiation, which does not carry the test.
// FLAT: the slice's rows back to back, grid (ceil(rows / MG_OBS_WARPS)), for envs with few agents -- with the
// default grid (envs, ceil(A / MG_OBS_WARPS)) a 4-agent env would leave half of every CTA's warps without a row.
// WAIT (hand-over mode): the kernel is launched beside the step kernel; CTA x takes the x-th env to finish its
// step (the step kernel appends finished envs to a queue), so CTAs are dispatched in the order their work
// becomes available and only the ones at the frontier wait (bounded: a broken launch ends with the error flag
// of the queue block instead of a hung GPU).
// AUX = false: row lengths, their maximum and the padding mask are not bound (mg_set_obs_aux).
// TT: see obs_row_warp.
template <int CPL, bool MASKED = true, bool FLAT = false, bool WAIT = false, bool AUX = true, int TT = 0>
MG_GLOBAL void MG_LAUNCH_BOUNDS(MG_OBS_WARPS * 32, MG_OBS_MIN_CTAS) mg_obs_kernel(const MG_GRID_CONSTANT MgBatch b, const uint8_t* mask) {
  MG_DYN_SMEM(dyn);
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  int env, ai;
#ifndef MG_SIMT_EMU
  int handed = -1;
#ifdef MG_TIMING
  unsigned long long t_dispatch = 0;
  if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_dispatch));
#endif
  if (WAIT) {
    MG_SHARED int s_env;
    if (threadIdx.x == 0) {
      const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&b.qctl->epoch);
      const unsigned long long* entry = b.queue + blockIdx.y;  // grid (groups, envs): an env's CTAs are dispatched together
      const long long t0 = clock64();
      int got = -1;
      for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(entry) : "memory");
        if ((uint32_t)(v >> 32) == epoch) { got = (int)(uint32_t)v; break; }
        if (clock64() - t0 > 6000000000ll) { atomicExch(&b.qctl->error, 1u); break; }
        MG_NANOSLEEP(100);
      }
      s_env = got;
    }
    __syncthreads();
    handed = s_env;
    if (handed < 0) return;
  }
#endif
  if (FLAT) {
    const int row = (int)blockIdx.x * MG_OBS_WARPS + warp;
    const int e = row / b.A;
    env = b.env_begin + e;
    ai = row - e * b.A;
    if (env >= b.env_end) return;
  } else {
    env = b.env_begin + (int)blockIdx.x;
    ai = (int)blockIdx.y * MG_OBS_WARPS + warp;
#ifndef MG_SIMT_EMU
    if (WAIT) { env = handed; ai = (int)blockIdx.x * MG_OBS_WARPS + warp; }
#endif
    if (ai >= b.A) return;
  }
  if (MASKED && mask != nullptr && !mask[env]) return;
#if defined(MG_TIMING) && !defined(MG_SIMT_EMU)
  if (threadIdx.x == 0 && (WAIT ? blockIdx.x == 0 : blockIdx.y == 0) && b.dbg != nullptr && b.after_step) {  // probe: when this env's first CTA was dispatched, and when its env was there
    unsigned long long t_go;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_go));
    b.dbg[(size_t)env * 16 + 13] = (uint32_t)t_dispatch;
    b.dbg[(size_t)env * 16 + 14] = (uint32_t)t_go;
  }
#endif
  obs_row_warp<CPL, AUX, TT>(b, env, ai, lane, dyn + warp * b.obs_warp_bytes);
#if defined(MG_TIMING) && !defined(MG_SIMT_EMU)
  if (threadIdx.x == 0 && (WAIT ? blockIdx.x == 0 : blockIdx.y == 0) && b.dbg != nullptr && b.after_step) {
    unsigned long long t_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
    b.dbg[(size_t)env * 16 + 15] = (uint32_t)t_end;
  }
#endif
}

#ifndef MG_SIMT_EMU
// The consumer side of the queue: CTAs take (env, group of MG_OBS_WARPS agents) tasks in the order
// envs finish their step. A task whose env has not been published yet is waited for (the step kernel
// is resident and makes progress on its own); the wait is bounded so that a broken launch ends with
// an error flag instead of a hung GPU.
#define MG_QUEUE_WAIT_CYCLES 6000000000ll /* ~3 s */
template <int CPL>
MG_GLOBAL void MG_LAUNCH_BOUNDS(MG_OBS_WARPS * 32, MG_OBS_MIN_CTAS) mg_obs_queue_kernel(const MG_GRID_CONSTANT MgBatch b) {
  MG_DYN_SMEM(dyn);
  MG_SHARED int s_task[2];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int slot = (b.T * 3 + 16 + 15) & ~15;
  uint8_t* warp_smem = dyn + (size_t)warp * (slot + MG_OBS_DESC_BYTES(CPL));
  const int groups = (b.A + MG_OBS_WARPS - 1) / MG_OBS_WARPS;
  const uint32_t total = (uint32_t)(b.env_end - b.env_begin) * (uint32_t)groups;
  for (;;) {
    if (threadIdx.x == 0) {
      int env = -1, group = 0;
      const uint32_t t = atomicAdd(&b.qctl->head, 1u);
      if (t < total) {
        const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&b.qctl->epoch);
        const volatile unsigned long long* entry = b.queue + t / (uint32_t)groups;
        const long long t0 = clock64();
        for (;;) {
          const unsigned long long v = *entry;
          if ((uint32_t)(v >> 32) == epoch) { env = (int)(uint32_t)v; break; }
          if (clock64() - t0 > MG_QUEUE_WAIT_CYCLES) { atomicExch(&b.qctl->error, 1u); break; }
          MG_NANOSLEEP(200);
        }
        group = (int)(t % (uint32_t)groups);
        __threadfence();
      }
      s_task[0] = env;
      s_task[1] = group;
    }
    __syncthreads();
    const int env = s_task[0], ai = s_task[1] * MG_OBS_WARPS + warp;
    __syncthreads();
    if (env < 0) return;
    if (ai < b.A) obs_row_warp<CPL>(b, env, ai, lane, warp_smem);
  }
}
#endif  // !MG_SIMT_EMU

#ifndef MG_SIMT_EMU
// ---------------------------------------------------------------- one kernel for the whole step
// mg_fused_kernel: a persistent grid in which every warp first steps envs (claimed one at a time from a
// counter) and then, once no env is left to claim, encodes observation rows of envs whose step has
// finished (tasks of MG_FUSED_ROWS rows, handed out in the order the envs were published). Why: the step
// of an env is one dependent chain (tens of microseconds, with a long tail where assemblers are used) that
// leaves the issue slots idle, while the observation pass is issue-bound. As two kernels per slice, every
// slice's observation pass waited for the slowest env of the slice and the tail of every launch ran on a
// half-empty GPU; here the encoding of finished envs fills the issue slots under the chains of the slow
// ones, and only the last env's own rows are left over at the end.
// Hand-over: the stepping warp makes its writes visible (__threadfence) and publishes epoch << 32 | env
// into the queue; a consumer polls the entry, fences, and reads the env's state. State that a consumer
// reads lives in cache lines owned by one env (grid and token-record strides are padded to 128 bytes; the
// one exception, EnvRec::current_step, is read at L2), so a line cached while encoding one env never holds
// another env's pre-step bytes.
#ifndef MG_FUSED_WARPS
#define MG_FUSED_WARPS 4
#endif
#ifndef MG_FUSED_MIN_CTAS
#define MG_FUSED_MIN_CTAS 6
#endif
#ifndef MG_FUSED_ROWS
#define MG_FUSED_ROWS 4
#endif
template <int MAXA, int CPL>
MG_GLOBAL void MG_LAUNCH_BOUNDS(MG_FUSED_WARPS * 32, MG_FUSED_MIN_CTAS) mg_fused_kernel(const MG_GRID_CONSTANT MgBatch b, int warp_bytes) {
  MG_DYN_SMEM(dyn);
  const int lane = threadIdx.x & 31;
  unsigned char* mine = dyn + (size_t)(threadIdx.x >> 5) * warp_bytes;  // [StepSmem] while stepping, [row image | descriptors] while encoding
  StepSmem<MAXA>& sm = *reinterpret_cast<StepSmem<MAXA>*>(mine);
  MgQueueCtl* ctl = b.qctl;
  const uint32_t n_envs = (uint32_t)(b.env_end - b.env_begin);
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&ctl->epoch);
  step_smem_init<MAXA>(sm, lane, true);
  for (;;) {  // ---- stepping
    uint32_t k = 0;
    if (lane == 0) k = atomicAdd(&ctl->env_next, 1u);
    k = __shfl_sync(MG_FULL_MASK, k, 0);
    if (k >= n_envs) break;
    const int env = b.env_begin + (int)k;
    step_env_warp<MAXA>(b, sm, env, lane);
    __threadfence();  // every lane's writes, then the entry
    __syncwarp();
    if (lane == 0) {
      const uint32_t pos = atomicAdd(&ctl->tail, 1u);
      *reinterpret_cast<volatile unsigned long long*>(b.queue + pos) = ((unsigned long long)epoch << 32) | (unsigned long long)(uint32_t)env;
    }
    __syncwarp();
  }
  // ---- encoding
  const uint32_t per_env = (uint32_t)((b.A + MG_FUSED_ROWS - 1) / MG_FUSED_ROWS);
  const uint32_t total = n_envs * per_env;
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(&ctl->head, 1u);
    t = __shfl_sync(MG_FULL_MASK, t, 0);
    if (t >= total) break;
    int env = -1;
    if (lane == 0) {
      const volatile unsigned long long* entry = b.queue + t / per_env;
      const long long t0 = clock64();
      for (;;) {
        const unsigned long long v = *entry;
        if ((uint32_t)(v >> 32) == epoch) { env = (int)(uint32_t)v; break; }
        if (clock64() - t0 > MG_QUEUE_WAIT_CYCLES) { atomicExch(&ctl->error, 1u); break; }
        MG_NANOSLEEP(100);
      }
    }
    env = __shfl_sync(MG_FULL_MASK, env, 0);
    if (env < 0) break;
    __threadfence();
    const int a0 = (int)(t % per_env) * MG_FUSED_ROWS;
#pragma unroll 1
    for (int ai = a0; ai < a0 + MG_FUSED_ROWS && ai < b.A; ai++) {
      obs_row_warp<CPL>(b, env, ai, lane, mine);
      __syncwarp();
    }
  }
}
#endif  // !MG_SIMT_EMU

// Host export of observations for mg_step_host (page-locked caller buffer, written over PCIe by the
// SMs): a row only differs from what the host already holds in its first max(previous, current)
// tokens, because everything behind the tokens written is 0xFF in both. One warp copies that prefix
// (rounded up to 8 bytes: the device row is right for every byte, so copying too much is harmless)
// of MG_EXPORT_ROWS rows; the CTA's first warp also writes the rows' rewards and flags.
// host.obs == nullptr only records the counts (after a full copy).
#ifndef MG_EXPORT_ROWS
#define MG_EXPORT_ROWS 4
#endif
#ifndef MG_EXPORT_ALIGN
#define MG_EXPORT_ALIGN 32
#endif
#define MG_EXPORT_WARPS 8
struct MgHostTargets {  // device-visible addresses of the caller's page-locked buffers (all rows; nullptr = not exported here)
  uint8_t* obs;
  float* rewards;
  uint8_t* terminals;
  uint8_t* truncations;
};
MG_GLOBAL void MG_LAUNCH_BOUNDS(MG_EXPORT_WARPS * 32, 4)
    mg_export_rows_kernel(const uint8_t* obs, const float* rewards, const uint8_t* terminals, const uint8_t* truncations, MgHostTargets host,
                          const uint16_t* tok_counts, uint16_t* prev, uint8_t* shadow, int row_begin, int row_end, int row_bytes,
                          unsigned long long* bytes_out) {
  const int lane = threadIdx.x & 31;
  const int cta_first = row_begin + (int)blockIdx.x * (MG_EXPORT_WARPS * MG_EXPORT_ROWS);
  const int first = cta_first + (int)(threadIdx.x >> 5) * MG_EXPORT_ROWS;
  int cur = 0, pv = 0;
  if (lane < MG_EXPORT_ROWS && first + lane < row_end) {
    cur = tok_counts[(size_t)(first + lane) * 2];
    pv = prev[first + lane];
    prev[first + lane] = (uint16_t)cur;
  }
  if (host.rewards != nullptr && threadIdx.x < MG_EXPORT_WARPS * MG_EXPORT_ROWS && cta_first + (int)threadIdx.x < row_end) {
    const int r = cta_first + (int)threadIdx.x;  // the CTA's rows: one thread each for the per-agent scalars
    host.rewards[r] = rewards[r];
    host.terminals[r] = terminals[r];
    host.truncations[r] = truncations[r];
  }
  if (host.obs == nullptr) return;
  if (shadow != nullptr) {
    // With a device copy of what the host buffer holds (`shadow`), a row's prefix is compared 64 bytes
    // at a time (the unit the writes cross PCIe in: 8 lanes x 8 bytes, aligned) and only the pieces that
    // differ are written, to the host and to the shadow. An agent that stood still keeps most of its row.
    unsigned long long total = 0;
#pragma unroll 1
    for (int k = 0; k < MG_EXPORT_ROWS; k++) {
      const int c = __shfl_sync(MG_FULL_MASK, cur, k), p = __shfl_sync(MG_FULL_MASK, pv, k);
      if (first + k >= row_end) break;
      const int n = c > p ? c : p;
      if (n == 0) continue;
      int nb = (n * 3 + 7) & ~7;
      nb = nb < row_bytes ? nb : row_bytes;
      const size_t lo = (size_t)(first + k) * row_bytes, hi = lo + nb;
#pragma unroll 1
      for (size_t w = lo & ~(size_t)63; w < hi; w += 256) {
        const size_t piece = w + (size_t)lane * 8;
        const bool in_range = piece >= lo && piece < hi;
        uint2 nv, sv;
        nv.x = nv.y = sv.x = sv.y = 0u;
        if (in_range) {
          nv = *reinterpret_cast<const uint2*>(obs + piece);
          sv = *reinterpret_cast<const uint2*>(shadow + piece);
        }
        const uint32_t differ = __ballot_sync(MG_FULL_MASK, in_range && (nv.x != sv.x || nv.y != sv.y));
        const bool write = in_range && ((differ >> (lane & 24)) & 0xFFu) != 0u;
        if (write) {
          *reinterpret_cast<uint2*>(host.obs + piece) = nv;
          *reinterpret_cast<uint2*>(shadow + piece) = nv;
        }
        total += 8ull * (unsigned long long)__popc(__ballot_sync(MG_FULL_MASK, write));
      }
    }
    if (lane == 0 && total != 0) atomicAdd(bytes_out, total);
    return;
  }
  const bool sector_aligned = MG_EXPORT_ALIGN > 8 && ((reinterpret_cast<uintptr_t>(obs) | reinterpret_cast<uintptr_t>(host.obs)) & (MG_EXPORT_ALIGN - 1)) == 0;
  unsigned long long total = 0;
#pragma unroll 1
  for (int k = 0; k < MG_EXPORT_ROWS; k++) {
    const int c = __shfl_sync(MG_FULL_MASK, cur, k), p = __shfl_sync(MG_FULL_MASK, pv, k);
    if (first + k >= row_end) break;
    const int n = c > p ? c : p;
    if (n == 0) continue;
    int nb = (n * 3 + 7) & ~7;
    nb = nb < row_bytes ? nb : row_bytes;
    // widen to whole 32-byte sectors (neighbouring rows of this launch are final too, so their bytes
    // are the right ones): PCIe writes of whole sectors need no byte enables
    size_t lo = (size_t)(first + k) * row_bytes, hi = lo + nb;
    if (sector_aligned) {
      const size_t lim_lo = (size_t)row_begin * row_bytes, lim_hi = (size_t)row_end * row_bytes;
      lo &= ~(size_t)(MG_EXPORT_ALIGN - 1);
      hi = (hi + MG_EXPORT_ALIGN - 1) & ~(size_t)(MG_EXPORT_ALIGN - 1);
      lo = lo < lim_lo ? lim_lo : lo;
      hi = hi > lim_hi ? lim_hi : hi;
    }
    if (((lo | hi) & 15) == 0) {
      const uint4* src = reinterpret_cast<const uint4*>(obs + lo);
      uint4* dst = reinterpret_cast<uint4*>(host.obs + lo);
      const int nv = (int)((hi - lo) >> 4);
#pragma unroll 1
      for (int i = lane; i < nv; i += 32) dst[i] = src[i];
    } else {
      const uint2* src = reinterpret_cast<const uint2*>(obs + lo);
      uint2* dst = reinterpret_cast<uint2*>(host.obs + lo);
      const int nv = (int)((hi - lo) >> 3);
#pragma unroll 1
      for (int i = lane; i < nv; i += 32) dst[i] = src[i];
    }
    total += (unsigned long long)(hi - lo);
  }
  if (lane == 0 && total != 0) atomicAdd(bytes_out, total);
}

// Folds pending token counts so that game stats can be read out (mg_read).
MG_GLOBAL void mg_fold_kernel(const MG_GRID_CONSTANT MgBatch b) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * MG_STEP_WARPS + (threadIdx.x >> 5);
  if (env >= b.num_envs) return;
  fold_token_stats_warp(b, env, lane);
}

// Agent::set_inventory (agent.cpp:54-72) for one agent; a test/tooling entry point of the reference API.
MG_GLOBAL void mg_set_inventory_kernel(const MG_GRID_CONSTANT MgBatch b, int env, int ai, const int32_t* items, const int32_t* amounts, int n) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  EnvRec er = b.env[env];
  AgentRec* agents = b.agents + (size_t)env * b.A;
  EnvCtx e = make_ctx(b, env, agents, er);
  InvView iv = agent_inv(e, ai);
  int existing[16], ne = 0;
  uint64_t o = agents[ai].order;
  while ((o & 0xFull) != 0xFull && ne < 16) { existing[ne++] = (int)(o & 0xFull); o >>= 4; }
  for (int i = 0; i < ne; i++) {
    inv_update(e, iv, existing[i], -(int)agents[ai].amt[existing[i]], false, 0);
    e.astats[(size_t)ai * e.S + MG_AS_AMOUNT + existing[i]] = 0.0f;
  }
  for (int i = 0; i < n; i++) inv_update(e, iv, items[i], amounts[i] - (int)agents[ai].amt[items[i]], false, 0);
  write_agent_tokens(b, agents[ai], token_record(b, env, MG_GK_AGENT, (uint32_t)ai));
}

#endif  // MG_KERNELS_CUH_
